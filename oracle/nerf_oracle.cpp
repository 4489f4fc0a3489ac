// CPU oracle for the NeRF driver kernels. TEST INFRASTRUCTURE ONLY (see oracle.h).
// Restates, routine by routine (paths relative to the reference root):
//   constants, stepping, cascades, warps      include/neural-graphics-primitives/nerf_device.cuh:20-40, 265-458
//   activations, losses                       nerf_device.cuh:74-143, 203-263, 600-615
//   image_idx, random image position          nerf_device.cuh:552-598
//   lens undistortion, uv_to_ray, read_rgba   include/neural-graphics-primitives/common_device.cuh:248-460, 730-806
//   sRGB                                      common_device.cuh:34-64
//   ray/box                                   include/neural-graphics-primitives/bounding_box.cuh:172-230
//   Morton                                    dependencies/tiny-cuda-nn/include/tiny-cuda-nn/common_device.h:759-785
//   generate_training_samples_nerf            src/testbed_nerf.cu:679-838
//   compute_loss_kernel_train_nerf            src/testbed_nerf.cu:841-1160 (no envmap / depth / error map / exposure)
//   grid_to_bitfield, bitfield_max_pool, mean src/testbed_nerf.cu:284-331, 2362-2379
//   mark_untrained_density_grid, generate_grid_samples_nerf_nonuniform, splat_grid_samples_nerf_max_nearest_neighbor,
//   ema_grid_samples_nerf                     src/testbed_nerf.cu:74-282; pos_to_uv common_device.cuh:497-531
// PARITY UNPINNED by the reference (no golden vectors, GPU-only, built with --use_fast_math). log / exp in
// the exponential stepping are the single-precision Cephes algorithms so that host and device agree bit
// for bit; everything else that decides where a sample lies is IEEE add/mul/div/sqrt (-ffp-contract=off).
#include "oracle.h"
#include "pcg32_oracle.h"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

typedef _Float16 half_t;
static inline half_t nbits_to_half(uint16_t b) { half_t h; memcpy(&h, &b, 2); return h; }

namespace {

constexpr uint32_t GRIDSIZE = 128, GRID_CELLS = 128 * 128 * 128, NERF_STEPS = 1024, CASCADES = 8;
constexpr float MAX_DEPTH = 16384.0f; // common_device.cuh:32
constexpr float SQRT3 = 1.73205080757f;
constexpr float MIN_CONE_STEPSIZE = SQRT3 / NERF_STEPS;
constexpr float MAX_CONE_STEPSIZE = MIN_CONE_STEPSIZE * (1 << (CASCADES - 1)) * NERF_STEPS / GRIDSIZE;
constexpr uint32_t MAX_RANDOM_SAMPLES_PER_RAY = 16;
constexpr float MIN_OPTICAL_THICKNESS = 0.01f;

struct V { float x, y, z; };
inline V operator+(V a, V b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V operator-(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V operator*(V a, V b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V operator/(V a, V b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline V operator*(float s, V a) { return {s * a.x, s * a.y, s * a.z}; }
inline V operator*(V a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V operator/(V a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V operator+(V a, float s) { return {a.x + s, a.y + s, a.z + s}; }
inline V operator-(V a, float s) { return {a.x - s, a.y - s, a.z - s}; }
inline float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float length(V a) { return sqrtf(dot(a, a)); }
inline V normalize(V a) { return a / length(a); }

struct Box { V min, max; };
inline bool contains(const Box& b, V p) { return p.x >= b.min.x && p.x <= b.max.x && p.y >= b.min.y && p.y <= b.max.y && p.z >= b.min.z && p.z <= b.max.z; }
inline V diag(const Box& b) { return b.max - b.min; }
void ray_intersect(const Box& b, V pos, V dir, float& tmin_o, float& tmax_o) {
	float tmin = (b.min.x - pos.x) / dir.x, tmax = (b.max.x - pos.x) / dir.x;
	if (tmin > tmax) std::swap(tmin, tmax);
	float tymin = (b.min.y - pos.y) / dir.y, tymax = (b.max.y - pos.y) / dir.y;
	if (tymin > tymax) std::swap(tymin, tymax);
	if (tmin > tymax || tymin > tmax) { tmin_o = tmax_o = FLT_MAX; return; }
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = (b.min.z - pos.z) / dir.z, tzmax = (b.max.z - pos.z) / dir.z;
	if (tzmin > tzmax) std::swap(tzmin, tzmax);
	if (tmin > tzmax || tzmin > tmax) { tmin_o = tmax_o = FLT_MAX; return; }
	if (tzmin > tmin) tmin = tzmin;
	if (tzmax < tmax) tmax = tzmax;
	tmin_o = tmin; tmax_o = tmax;
}

inline uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
inline uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
inline uint32_t morton3D_invert(uint32_t x) {
	x = x & 0x49249249;
	x = (x | (x >> 2)) & 0xc30c30c3;
	x = (x | (x >> 4)) & 0x0f00f00f;
	x = (x | (x >> 8)) & 0xff0000ff;
	x = (x | (x >> 16)) & 0x0000ffff;
	return x;
}

// Cephes single-precision logf / expf (see file header).
float c_logf(float xx) {
	int e;
	float x = frexpf(xx, &e);
	if (x < 0.707106781186547524f) { e -= 1; x = x + x - 1.0f; } else { x = x - 1.0f; }
	// degree-8 polynomial in Estrin form with explicit fused multiply-adds (4 dependent steps)
	const float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
	const float a0 = fmaf(-2.4999993993e-1f, x, 3.3333331174e-1f);
	const float a1 = fmaf(-1.6668057665e-1f, x, 2.0000714765e-1f);
	const float a2 = fmaf(-1.2420140846e-1f, x, 1.4249322787e-1f);
	const float a3 = fmaf(-1.1514610310e-1f, x, 1.1676998740e-1f);
	const float b0 = fmaf(a1, x2, a0), b1 = fmaf(a3, x2, a2);
	const float p = fmaf(7.0376836292e-2f, x8, fmaf(b1, x4, b0));
	const float fe = (float)e;
	float y = (x * x2) * p;
	y = fmaf(-2.12194440e-4f, fe, y);
	y = fmaf(-0.5f, x2, y);
	return fmaf(0.693359375f, fe, x + y);
}
float c_expf(float x) {
	if (x > 88.0f) x = 88.0f;
	if (x < -87.0f) x = -87.0f;
	const float z = floorf(fmaf(1.44269504088896341f, x, 0.5f));
	x = fmaf(z, -0.693359375f, x);
	x = fmaf(z, 2.12194440e-4f, x);
	const int n = (int)z;
	const float x2 = x * x;
	const float a0 = fmaf(1.6666665459e-1f, x, 5.0000001201e-1f);
	const float a1 = fmaf(8.3334519073e-3f, x, 4.1665795894e-2f);
	const float a2 = fmaf(1.9875691500e-4f, x, 1.3981999507e-3f);
	const float p = fmaf(fmaf(a2, x2, a1), x2, a0);
	return scalbnf(fmaf(p, x2, x) + 1.0f, n);
}

// nerf_device.cuh:372-424, evaluated per call exactly as written there.
// Divisions by MIN/MAX_CONE_STEPSIZE and log1p_c are multiplications by their once-rounded
// reciprocals (what --use_fast_math does to the reference's own build, CMakeLists.txt:82).
float to_stepping_space(float t, float cone_angle) {
	const float inv_min = 1.0f / MIN_CONE_STEPSIZE, inv_max = 1.0f / MAX_CONE_STEPSIZE;
	if (cone_angle <= 1e-5f) return t * inv_min;
	const float log1p_c = c_logf(1.0f + cone_angle);
	const float inv_log1p_c = 1.0f / log1p_c;
	const float a = (c_logf(MIN_CONE_STEPSIZE) - c_logf(log1p_c)) / log1p_c;
	const float b = (c_logf(MAX_CONE_STEPSIZE) - c_logf(log1p_c)) / log1p_c;
	const float at = c_expf(a * log1p_c), bt = c_expf(b * log1p_c);
	if (t <= at) return (t - at) * inv_min + a;
	if (t <= bt) return c_logf(t) * inv_log1p_c;
	return (t - bt) * inv_max + b;
}
float from_stepping_space(float n, float cone_angle) {
	if (cone_angle <= 1e-5f) return n * MIN_CONE_STEPSIZE;
	const float log1p_c = c_logf(1.0f + cone_angle);
	const float a = (c_logf(MIN_CONE_STEPSIZE) - c_logf(log1p_c)) / log1p_c;
	const float b = (c_logf(MAX_CONE_STEPSIZE) - c_logf(log1p_c)) / log1p_c;
	const float at = c_expf(a * log1p_c), bt = c_expf(b * log1p_c);
	if (n <= a) return (n - a) * MIN_CONE_STEPSIZE + at;
	if (n <= b) return c_expf(n * log1p_c);
	return (n - b) * MAX_CONE_STEPSIZE + bt;
}
float advance_n_steps(float t, float cone_angle, float n) { return from_stepping_space(to_stepping_space(t, cone_angle) + n, cone_angle); }
float calc_dt(float t, float cone_angle) { return advance_n_steps(t, cone_angle, 1.0f) - t; }
inline float sgn(float x) { return copysignf(1.0f, x); }
float distance_to_next_voxel(V pos, V dir, V idir, float res) {
	const V p = res * (pos - 0.5f);
	const float tx = (floorf(p.x + 0.5f + 0.5f * sgn(dir.x)) - p.x) * idir.x;
	const float ty = (floorf(p.y + 0.5f + 0.5f * sgn(dir.y)) - p.y) * idir.y;
	const float tz = (floorf(p.z + 0.5f + 0.5f * sgn(dir.z)) - p.z) * idir.z;
	const float t = fminf(fminf(tx, ty), tz);
	return fmaxf(t / res, 0.0f);
}
float advance_to_next_voxel(float t, float cone_angle, V pos, V dir, V idir, uint32_t mip) {
	const float res = scalbnf((float)GRIDSIZE, -(int)mip);
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	t = to_stepping_space(t, cone_angle);
	t_target = to_stepping_space(t_target, cone_angle);
	return from_stepping_space(t + ceilf(fmaxf(t_target - t, 0.5f)), cone_angle);
}
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
uint32_t mip_from_pos(V pos, uint32_t max_cascade) {
	int exponent;
	const float maxval = fmaxf(fmaxf(fabsf(pos.x - 0.5f), fabsf(pos.y - 0.5f)), fabsf(pos.z - 0.5f));
	frexpf(maxval, &exponent);
	return (uint32_t)clampi(exponent + 1, 0, (int)max_cascade);
}
uint32_t mip_from_dt(float dt, V pos, uint32_t max_cascade) {
	const uint32_t mip = mip_from_pos(pos, max_cascade);
	dt *= 2 * GRIDSIZE;
	if (dt < 1.0f) return mip;
	int exponent;
	frexpf(dt, &exponent);
	return (uint32_t)clampi((int)mip, exponent, (int)max_cascade);
}
uint32_t cascaded_grid_idx_at(V pos, uint32_t mip) {
	const float mip_scale = scalbnf(1.0f, -(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	const int ix = (int)(pos.x * (float)GRIDSIZE), iy = (int)(pos.y * (float)GRIDSIZE), iz = (int)(pos.z * (float)GRIDSIZE);
	if (ix < 0 || ix >= (int)GRIDSIZE || iy < 0 || iy >= (int)GRIDSIZE || iz < 0 || iz >= (int)GRIDSIZE) return 0xFFFFFFFFu;
	return morton3D(ix, iy, iz);
}
bool occupied_at(V pos, const uint8_t* bitfield, uint32_t mip) {
	const uint32_t idx = cascaded_grid_idx_at(pos, mip);
	if (idx == 0xFFFFFFFFu) return false;
	return (bitfield[idx / 8 + (size_t)GRID_CELLS * mip / 8] & (1u << (idx % 8))) != 0;
}
inline float warp_dt(float dt) {
	const float max_stepsize = MIN_CONE_STEPSIZE * (1 << (CASCADES - 1));
	return (dt - MIN_CONE_STEPSIZE) / (max_stepsize - MIN_CONE_STEPSIZE);
}
inline float unwarp_dt(float dt) {
	const float max_stepsize = MIN_CONE_STEPSIZE * (1 << (CASCADES - 1));
	return dt * (max_stepsize - MIN_CONE_STEPSIZE) + MIN_CONE_STEPSIZE;
}

// ---- images, lens, rays ------------------------------------------------------------------------------
inline float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
inline float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

struct Rgba { float r, g, b, a; };
Rgba read_rgba(float u, float v, const oracle_nerf_image& m) {
	const int px = clampi((int)(u * (float)m.res_x), 0, m.res_x - 1), py = clampi((int)(v * (float)m.res_y), 0, m.res_y - 1);
	const size_t idx = (size_t)px + (size_t)py * m.res_x;
	switch (m.image_data_type) {
		case 1: {
			const uint32_t val = ((const uint32_t*)m.pixels)[idx];
			if (val == 0x00FF00FFu) return {-1.0f, -1.0f, -1.0f, -1.0f};
			const float alpha = (float)(val >> 24) * (1.0f / 255.0f);
			return {srgb_to_linear((float)(val & 0xFF) * (1.0f / 255.0f)) * alpha, srgb_to_linear((float)((val >> 8) & 0xFF) * (1.0f / 255.0f)) * alpha,
				srgb_to_linear((float)((val >> 16) & 0xFF) * (1.0f / 255.0f)) * alpha, alpha};
		}
		case 2: {
			const uint16_t* p = (const uint16_t*)m.pixels + idx * 4;
			return {(float)nbits_to_half(p[0]), (float)nbits_to_half(p[1]), (float)nbits_to_half(p[2]), (float)nbits_to_half(p[3])};
		}
		case 3: {
			const float* p = (const float*)m.pixels + idx * 4;
			return {p[0], p[1], p[2], p[3]};
		}
		default: return {5.0f, 0.0f, 0.0f, 1.0f};
	}
}

void opencv_lens_distortion_delta(const float* p, float u, float v, float* du, float* dv) {
	const float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
	const float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
	const float radial = k1 * r2 + k2 * r2 * r2;
	*du = u * radial + 2.0f * p1 * uv + p2 * (r2 + 2.0f * u2);
	*dv = v * radial + 2.0f * p2 * uv + p1 * (r2 + 2.0f * v2);
}
void iterative_opencv_lens_undistortion(const float* params, float* u, float* v) {
	const float kMaxStepNorm = 1e-10f, kRelStepSize = 1e-6f;
	const float x0 = *u, y0 = *v;
	float x = x0, y = y0;
	for (uint32_t i = 0; i < 100; ++i) {
		const float step0 = fmaxf(FLT_EPSILON, fabsf(kRelStepSize * x));
		const float step1 = fmaxf(FLT_EPSILON, fabsf(kRelStepSize * y));
		float dx0, dx1, b00, b01, f00, f01, b10, b11, f10, f11;
		opencv_lens_distortion_delta(params, x, y, &dx0, &dx1);
		opencv_lens_distortion_delta(params, x - step0, y, &b00, &b01);
		opencv_lens_distortion_delta(params, x + step0, y, &f00, &f01);
		opencv_lens_distortion_delta(params, x, y - step1, &b10, &b11);
		opencv_lens_distortion_delta(params, x, y + step1, &f10, &f11);
		const float j00 = 1 + (f00 - b00) / (2 * step0);
		const float j10 = (f10 - b10) / (2 * step1);
		const float j01 = (f01 - b01) / (2 * step0);
		const float j11 = 1 + (f11 - b11) / (2 * step1);
		const float rx = x + dx0 - x0, ry = y + dx1 - y0;
		const float det = j00 * j11 - j10 * j01;
		// inverse(J) * r as tcnn computes it (vec.h:802-807, 861-863, 587-597): the adjugate divided
		// by the determinant element by element, then column 0 * r.x + column 1 * r.y
		const float sx = (j11 / det) * rx + ((-j10) / det) * ry;
		const float sy = ((-j01) / det) * rx + (j00 / det) * ry;
		x -= sx;
		y -= sy;
		if (sx * sx + sy * sy < kMaxStepNorm) break;
	}
	*u = x;
	*v = y;
}
// The lattice march of csrc/nerf_train.cu, serially: the reference's loop (src/testbed_nerf.cu:760-790) with every
// position written as from(n0 + k), k an integer: an occupied candidate is a sample and k += 1, an empty one skips
// ceil(max(to(t + distance_to_next_voxel) - n_k, 0.5)) candidates (= advance_to_next_voxel, nerf_device.cuh:426-437).
void lattice_march(const Box& aabb, uint32_t max_mip, const uint8_t* bitfield, float cone_angle, V o, V dn, V idir, float n0, std::vector<uint32_t>& ks) {
	uint32_t k = 0;
	for (;;) {
		const float nk = n0 + (float)k;
		const float t = from_stepping_space(nk, cone_angle);
		const V pos = o + t * dn;
		if (!contains(aabb, pos) || ks.size() >= NERF_STEPS || k > 65535u - 64u) break;
		const float dt = from_stepping_space(nk + 1.0f, cone_angle) - t;
		const uint32_t mip = mip_from_dt(dt, pos, max_mip);
		if (occupied_at(pos, bitfield, mip)) {
			ks.push_back(k);
			k += 1;
		} else {
			const float res = scalbnf((float)GRIDSIZE, -(int)mip);
			const float n_target = to_stepping_space(t + distance_to_next_voxel(pos, dn, idir, res), cone_angle);
			k += (uint32_t)fminf(ceilf(fmaxf(n_target - nk, 0.5f)), 1048576.0f);
		}
	}
}
struct Cam { V c[4]; };
Cam cam_of(const oracle_nerf_image& m) {
	Cam c;
	for (int k = 0; k < 4; ++k) c.c[k] = {m.xform[3 * k + 0], m.xform[3 * k + 1], m.xform[3 * k + 2]};
	return c;
}
void uv_to_ray(float u, float v, const oracle_nerf_image& m, V& origin, V& dir) {
	const Cam cam = cam_of(m);
	V d = {(u - m.principal_x) * (float)m.res_x / m.focal_x, (v - m.principal_y) * (float)m.res_y / m.focal_y, 1.0f};
	if (m.lens_mode == 1) iterative_opencv_lens_undistortion(m.lens_params, &d.x, &d.y);
	dir = cam.c[0] * d.x + cam.c[1] * d.y + cam.c[2] * d.z;
	origin = cam.c[3];
}
uint32_t image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images) {
	return (uint32_t)((uint32_t)(base_idx * n_images) / n_rays) % n_images;
}
void random_image_pos(OraclePcg32& rng, int res_x, int res_y, bool snap, float& u, float& v) {
	u = rng.next_float();
	v = rng.next_float();
	if (snap) {
		u = ((float)clampi((int)(u * (float)res_x), 0, res_x - 1) + 0.5f) / (float)res_x;
		v = ((float)clampi((int)(v * (float)res_y), 0, res_y - 1) + 0.5f) / (float)res_y;
	}
}
OraclePcg32 rng_of(const uint64_t* s) {
	OraclePcg32 r;
	r.state = s[0];
	r.inc = s[1];
	return r;
}
Box box_of(const oracle_nerf_config& c) { return {{c.aabb_min[0], c.aabb_min[1], c.aabb_min[2]}, {c.aabb_max[0], c.aabb_max[1], c.aabb_max[2]}}; }

} // namespace

// generate_training_samples_nerf for rays [0, n_rays): numsteps[i] (0 = no samples), rays [i][6] and
// the samples of ray i appended in ray order to coords [..][7] (up to max_samples rows). Returns the
// total number of samples over all rays (may exceed max_samples; rows beyond it are not written).
extern "C" uint64_t oracle_nerf_generate_samples(const oracle_nerf_config* cfg, const oracle_nerf_image* images, const uint8_t* bitfield,
	const uint64_t* rng_state, uint32_t* numsteps, float* rays, float* coords, uint64_t max_samples)
{
	const Box aabb = box_of(*cfg);
	uint64_t total = 0;
	for (uint32_t i = 0; i < cfg->n_rays; ++i) {
		numsteps[i] = 0;
		const uint32_t img = image_idx(i, cfg->n_rays, cfg->n_training_images);
		const oracle_nerf_image& m = images[img];
		OraclePcg32 rng = rng_of(rng_state);
		rng.advance((int64_t)i * MAX_RANDOM_SAMPLES_PER_RAY);
		float u, v;
		random_image_pos(rng, m.res_x, m.res_y, cfg->snap_to_pixel_centers != 0, u, v);
		if (read_rgba(u, v, m).r < 0.0f) continue;
		rng.next_float(); // motionblur_time
		V o, d;
		uv_to_ray(u, v, m, o, d);
		if (d.x == 0.0f && d.y == 0.0f && d.z == 0.0f) { const Cam c = cam_of(m); o = c.c[3]; d = c.c[2]; }
		const V dn = normalize(d);
		float tmin, tmax;
		ray_intersect(aabb, o, dn, tmin, tmax);
		tmin = fmaxf(tmin, 0.0f);
		const float cone_angle = cfg->cone_angle_constant;
		const float startt = advance_n_steps(tmin, cone_angle, rng.next_float());
		const V idir = V{1.0f, 1.0f, 1.0f} / dn;
		// the lattice march (csrc/nerf_train.cu): candidate k sits at n0 + k in stepping space
		const float n0 = to_stepping_space(startt, cone_angle);
		std::vector<uint32_t> ks;
		lattice_march(aabb, cfg->max_mip, bitfield, cone_angle, o, dn, idir, n0, ks);
		const uint32_t j = (uint32_t)ks.size();
		if (j == 0) continue;
		numsteps[i] = j;
		float* r = rays + (size_t)i * 6;
		r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
		const V wdir = (dn + 1.0f) * 0.5f;
		for (uint32_t k = 0; k < j; ++k) {
			if (total + k >= max_samples) break;
			const float nk = n0 + (float)ks[k];
			const float t = from_stepping_space(nk, cone_angle);
			const float dt = from_stepping_space(nk + 1.0f, cone_angle) - t;
			const V wp = (o + t * dn - aabb.min) / diag(aabb);
			float* c = coords + (size_t)(total + k) * 7;
			c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt); c[4] = wdir.x; c[5] = wdir.y; c[6] = wdir.z;
		}
		total += j;
	}
	return total;
}

namespace {
inline float logistic(float x) { return 1.0f / (1.0f + c_expf(-x)); }
inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
float net_to_rgb(float v, int act) {
	switch (act) { case 1: return v > 0 ? v : 0; case 2: return logistic(v); case 3: return c_expf(clampf(v, -10.0f, 10.0f)); default: return v; }
}
float net_to_rgb_d(float v, int act) {
	switch (act) { case 1: return v > 0 ? 1.0f : 0.0f; case 2: { float d = logistic(v); return d * (1 - d); } case 3: return c_expf(clampf(v, -10.0f, 10.0f)); default: return 1.0f; }
}
float net_to_density(float v, int act) {
	switch (act) { case 1: return v > 0 ? v : 0; case 2: return logistic(v); case 3: return c_expf(v); default: return v; }
}
float net_to_density_d(float v, int act) {
	switch (act) { case 1: return v > 0 ? 1.0f : 0.0f; case 2: { float d = logistic(v); return d * (1 - d); } case 3: return c_expf(clampf(v, -15.0f, 15.0f)); default: return 1.0f; }
}
void loss1(float target, float pred, int type, float& loss, float& grad) {
	const float d = pred - target;
	switch (type) {
		case 6: { const float den = pred * pred + 1e-2f; loss = d * d / den; grad = 2.0f * d / den; } break;
		case 1: loss = fabsf(d); grad = copysignf(1.0f, d); break;
		case 2: { const float den = fabsf(pred) + 1e-2f; loss = fabsf(d) / den; grad = copysignf(1.0f / den, d); } break;
		case 3: { const float den = 0.5f * (fabsf(pred) + fabsf(target)) + 1e-2f; loss = fabsf(d) / den; grad = copysignf(1.0f / den, d); } break;
		case 4: {
			const float alpha = 0.1f, ad = fabsf(d), sq = 0.5f / alpha * d * d;
			loss = (ad > alpha ? (ad - 0.5f * alpha) : sq) / 5.0f;
			grad = (ad > alpha ? (d > 0 ? 1.0f : -1.0f) : (d / alpha)) / 5.0f;
		} break;
		case 5: { const float div = fabsf(d) + 1.0f; loss = c_logf(div); grad = copysignf(1.0f / div, d); } break;
		default: loss = d * d; grad = 2.0f * d; break;
	}
}
} // namespace

extern "C" void oracle_nerf_compute_loss(const oracle_nerf_config* cfg, const oracle_nerf_image* images, float mean_density, uint32_t n_active,
	const uint32_t* ray_indices, const uint32_t* numsteps_in, const float* rays, const float* coords, const uint16_t* rgbsigma,
	const uint64_t* rng_state, uint32_t* compacted_numsteps_out, float* loss_out, float* dloss)
{
	const Box aabb = box_of(*cfg);
	for (uint32_t i = 0; i < n_active; ++i) {
		const uint32_t numsteps = numsteps_in[i * 2 + 0], base = numsteps_in[i * 2 + 1];
		const float* c_in = coords + (size_t)base * 7;
		const uint16_t* out = rgbsigma + (size_t)base * 4;
		float T = 1.f;
		const float EPSILON = 1e-4f;
		V rgb_ray = {0, 0, 0};
		const V ray_o = {rays[(size_t)i * 6 + 0], rays[(size_t)i * 6 + 1], rays[(size_t)i * 6 + 2]};
		uint32_t compacted = 0;
		for (; compacted < numsteps; ++compacted) {
			if (T < EPSILON) break;
			const float o[4] = {(float)nbits_to_half(out[compacted * 4 + 0]), (float)nbits_to_half(out[compacted * 4 + 1]), (float)nbits_to_half(out[compacted * 4 + 2]), (float)nbits_to_half(out[compacted * 4 + 3])};
			const V rgb = {net_to_rgb(o[0], cfg->rgb_activation), net_to_rgb(o[1], cfg->rgb_activation), net_to_rgb(o[2], cfg->rgb_activation)};
			const float dt = unwarp_dt(c_in[(size_t)compacted * 7 + 3]);
			const float density = net_to_density(o[3], cfg->density_activation);
			const float alpha = 1.f - c_expf(-density * dt);
			const float weight = alpha * T;
			rgb_ray = rgb_ray + weight * rgb;
			T *= (1.f - alpha);
		}
		const uint32_t ray_idx = ray_indices[i];
		OraclePcg32 rng = rng_of(rng_state);
		rng.advance((int64_t)ray_idx * MAX_RANDOM_SAMPLES_PER_RAY);
		const uint32_t img = image_idx(ray_idx, cfg->n_rays, cfg->n_training_images);
		const oracle_nerf_image& m = images[img];
		float u, v;
		random_image_pos(rng, m.res_x, m.res_y, cfg->snap_to_pixel_centers != 0, u, v);
		rng.advance(1);
		V bg = {cfg->background_color[0], cfg->background_color[1], cfg->background_color[2]};
		if (cfg->random_bg_color) { bg.x = rng.next_float(); bg.y = rng.next_float(); bg.z = rng.next_float(); }
		bg = {srgb_to_linear(bg.x), srgb_to_linear(bg.y), srgb_to_linear(bg.z)};
		const Rgba tex = read_rgba(u, v, m);
		const V tex_rgb = {tex.r, tex.g, tex.b};
		V target;
		if (cfg->linear_colors || cfg->color_space == 0) {
			target = tex_rgb + (1.0f - tex.a) * bg;
			if (!cfg->linear_colors) {
				target = {linear_to_srgb(target.x), linear_to_srgb(target.y), linear_to_srgb(target.z)};
				bg = {linear_to_srgb(bg.x), linear_to_srgb(bg.y), linear_to_srgb(bg.z)};
			}
		} else {
			bg = {linear_to_srgb(bg.x), linear_to_srgb(bg.y), linear_to_srgb(bg.z)};
			if (tex.a > 0) {
				const V t = tex_rgb / tex.a;
				target = V{linear_to_srgb(t.x), linear_to_srgb(t.y), linear_to_srgb(t.z)} * tex.a + (1.0f - tex.a) * bg;
			} else target = bg;
		}
		if (compacted == numsteps) rgb_ray = rgb_ray + T * bg;
		compacted_numsteps_out[i] = compacted;
		loss_out[i] = 0.0f;
		if (compacted == 0) continue;
		V l, g;
		loss1(target.x, rgb_ray.x, cfg->loss_type, l.x, g.x);
		loss1(target.y, rgb_ray.y, cfg->loss_type, l.y, g.y);
		loss1(target.z, rgb_ray.z, cfg->loss_type, l.z, g.z);
		loss_out[i] = ((l.x + l.y + l.z) / 3.0f) / (float)cfg->n_rays;
		const float loss_scale = cfg->loss_scale / (float)cfg->n_rays;
		const float output_l2_reg = cfg->rgb_activation == 3 ? 1e-4f : 0.0f;
		const float output_l1_reg_density = mean_density < MIN_OPTICAL_THICKNESS ? 1e-4f : 0.0f;
		V rgb_ray2 = {0, 0, 0};
		T = 1.f;
		for (uint32_t j = 0; j < compacted; ++j) {
			const float* c = c_in + (size_t)j * 7;
			const V pos = aabb.min + V{c[0], c[1], c[2]} * diag(aabb);
			const float depth = length(pos - ray_o);
			const float dt = unwarp_dt(c[3]);
			const float o[4] = {(float)nbits_to_half(out[j * 4 + 0]), (float)nbits_to_half(out[j * 4 + 1]), (float)nbits_to_half(out[j * 4 + 2]), (float)nbits_to_half(out[j * 4 + 3])};
			const V rgb = {net_to_rgb(o[0], cfg->rgb_activation), net_to_rgb(o[1], cfg->rgb_activation), net_to_rgb(o[2], cfg->rgb_activation)};
			const float density = net_to_density(o[3], cfg->density_activation);
			const float alpha = 1.f - c_expf(-density * dt);
			const float weight = alpha * T;
			rgb_ray2 = rgb_ray2 + weight * rgb;
			T *= (1.f - alpha);
			const V suffix = rgb_ray - rgb_ray2;
			const V dl = weight * g;
			float* d = dloss + (size_t)(base + j) * 4;
			d[0] = loss_scale * (dl.x * net_to_rgb_d(o[0], cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o[0]));
			d[1] = loss_scale * (dl.y * net_to_rgb_d(o[1], cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o[1]));
			d[2] = loss_scale * (dl.z * net_to_rgb_d(o[2], cfg->rgb_activation) + fmaxf(0.0f, output_l2_reg * o[2]));
			const float dd = net_to_density_d(o[3], cfg->density_activation);
			const float dmlp = dd * (dt * dot(g, T * rgb - suffix));
			d[3] = loss_scale * dmlp + (o[3] < 0.0f ? -output_l1_reg_density : 0.0f) + (o[3] > -10.0f && depth < cfg->near_distance ? 1e-4f : 0.0f);
		}
	}
}

extern "C" void oracle_nerf_bitfield(uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density) {
	double sum = 0.0;
	for (uint32_t i = 0; i < GRID_CELLS; ++i) sum += (double)(fmaxf(grid[i], 0.f) / (float)GRID_CELLS);
	const float mean = (float)sum;
	*mean_density = mean;
	const float thresh = fminf(MIN_OPTICAL_THICKNESS, mean);
	const uint32_t n = GRID_CELLS / 8 * CASCADES, n_nonzero = GRID_CELLS / 8 * (max_cascade + 1);
	for (uint32_t i = 0; i < n; ++i) {
		if (i >= n_nonzero) { bitfield[i] = 0; continue; }
		uint8_t bits = 0;
		for (uint8_t j = 0; j < 8; ++j) bits |= grid[(size_t)i * 8 + j] > thresh ? ((uint8_t)1 << j) : 0;
		bitfield[i] = bits;
	}
	for (uint32_t level = 1; level < CASCADES; ++level) {
		const uint8_t* prev = bitfield + (size_t)GRID_CELLS * (level - 1) / 8;
		uint8_t* next = bitfield + (size_t)GRID_CELLS * level / 8;
		for (uint32_t i = 0; i < GRID_CELLS / 64; ++i) {
			uint8_t bits = 0;
			for (uint8_t j = 0; j < 8; ++j) bits |= prev[(size_t)i * 8 + j] > 0 ? ((uint8_t)1 << j) : 0;
			const uint32_t x = morton3D_invert(i >> 0) + GRIDSIZE / 8, y = morton3D_invert(i >> 1) + GRIDSIZE / 8, z = morton3D_invert(i >> 2) + GRIDSIZE / 8;
			next[morton3D(x, y, z)] |= bits;
		}
	}
}

// ---- occupancy-grid maintenance (src/testbed_nerf.cu:74-282) ---------------------------------------
namespace {
// pos_to_uv (common_device.cuh:497-531) without parallax / foveation: the inverse of the camera's
// 3x3 part by cofactors, in the operation order of csrc/nerf_grid.cu so that both sides round alike.
void pos_to_uv(V pos, const oracle_nerf_image& m, float& u, float& v) {
	const Cam cam = cam_of(m);
	const V d = pos - cam.c[3];
	const V c0 = cam.c[0], c1 = cam.c[1], c2 = cam.c[2];
	const float det = c0.x * (c1.y * c2.z - c2.y * c1.z) - c1.x * (c0.y * c2.z - c2.y * c0.z) + c2.x * (c0.y * c1.z - c1.y * c0.z);
	const float inv = 1.0f / det;
	V dir;
	dir.x = inv * (d.x * (c1.y * c2.z - c2.y * c1.z) - c1.x * (d.y * c2.z - c2.y * d.z) + c2.x * (d.y * c1.z - c1.y * d.z));
	dir.y = inv * (c0.x * (d.y * c2.z - c2.y * d.z) - d.x * (c0.y * c2.z - c2.y * c0.z) + c2.x * (c0.y * d.z - d.y * c0.z));
	dir.z = inv * (c0.x * (c1.y * d.z - d.y * c1.z) - c1.x * (c0.y * d.z - d.y * c0.z) + d.x * (c0.y * c1.z - c1.y * c0.z));
	dir = dir / dir.z;
	float du = 0.0f, dv = 0.0f;
	if (m.lens_mode == 1) opencv_lens_distortion_delta(m.lens_params, dir.x, dir.y, &du, &dv);
	dir.x += du;
	dir.y += dv;
	u = dir.x * m.focal_x / (float)m.res_x + m.principal_x;
	v = dir.y * m.focal_y / (float)m.res_y + m.principal_y;
}
} // namespace

// mark_untrained_density_grid (:74-147): a cell no training camera sees is marked -1 (never sampled).
extern "C" void oracle_nerf_mark_untrained(uint32_t n_elements, float* grid, uint32_t n_images, const oracle_nerf_image* images, int clear_visible_voxels) {
#pragma omp parallel for schedule(static)
	for (int64_t ii = 0; ii < (int64_t)n_elements; ++ii) {
		const uint32_t i = (uint32_t)ii;
		const uint32_t level = i / GRID_CELLS, pos_idx = i % GRID_CELLS;
		const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
		const float voxel_size = scalbnf(1.0f / GRIDSIZE, (int)level);
		const V pos = (V{(float)x, (float)y, (float)z} / (float)GRIDSIZE - 0.5f) * scalbnf(1.0f, (int)level) + 0.5f;
		const uint32_t min_count = 1;
		uint32_t count = 0;
		for (uint32_t j = 0; j < n_images && count < min_count; ++j) {
			const oracle_nerf_image& m = images[j];
			const Cam cam = cam_of(m);
			for (uint32_t k = 0; k < 8; ++k) {
				const V corner = pos + V{(k & 1) ? voxel_size : 0.0f, (k & 2) ? voxel_size : 0.0f, (k & 4) ? voxel_size : 0.0f};
				const V dir = normalize(corner - cam.c[3]);
				if (dot(dir, cam.c[2]) < 1e-4f) continue;
				float u, v;
				pos_to_uv(corner, m, u, v);
				V ro, rd;
				uv_to_ray(u, v, m, ro, rd);
				if (length(normalize(rd) - dir) < 1e-3f && u > 0.0f && v > 0.0f && u < 1.0f && v < 1.0f) { ++count; break; }
			}
		}
		if (clear_visible_voxels || (grid[i] < 0) != (count < min_count)) grid[i] = (count >= min_count) ? 0.f : -1.f;
	}
}

// generate_grid_samples_nerf_nonuniform (:149-184): positions [n][4] (warped xyz, warped dt) and cell indices.
extern "C" void oracle_nerf_generate_grid_samples(uint32_t n, const uint64_t* rng_state, uint32_t step, const float* aabb6, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh)
{
	const Box aabb = {{aabb6[0], aabb6[1], aabb6[2]}, {aabb6[3], aabb6[4], aabb6[5]}};
	for (uint32_t i = 0; i < n; ++i) {
		OraclePcg32 rng = rng_of(rng_state);
		rng.advance((int64_t)i * 4);
		const uint32_t level = (uint32_t)(rng.next_float() * n_cascades) % n_cascades;
		uint32_t idx = 0;
		for (uint32_t j = 0; j < 10; ++j) {
			idx = ((i + step * n) * 56924617u + j * 19349663u + 96925573u) % GRID_CELLS;
			idx += level * GRID_CELLS;
			if (grid_in[idx] > thresh) break;
		}
		const uint32_t pos_idx = idx % GRID_CELLS;
		const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
		const float rx = rng.next_float(), ry = rng.next_float(), rz = rng.next_float();
		const V pos = ((V{(float)x, (float)y, (float)z} + V{rx, ry, rz}) / (float)GRIDSIZE - 0.5f) * scalbnf(1.0f, (int)level) + 0.5f;
		const V wp = (pos - aabb.min) / diag(aabb);
		positions[(size_t)i * 4 + 0] = wp.x; positions[(size_t)i * 4 + 1] = wp.y; positions[(size_t)i * 4 + 2] = wp.z;
		positions[(size_t)i * 4 + 3] = warp_dt(MIN_CONE_STEPSIZE);
		indices[i] = idx;
	}
}

// splat_grid_samples_nerf_max_nearest_neighbor (:186-204): running maximum of the optical thickness per cell.
extern "C" void oracle_nerf_splat_grid_samples(uint32_t n, const uint32_t* indices, const uint16_t* density_raw, float* grid_tmp, int density_activation) {
	for (uint32_t i = 0; i < n; ++i) {
		const float mlp = net_to_density(oracle_half_to_float(density_raw[i]), density_activation);
		const float optical_thickness = mlp * MIN_CONE_STEPSIZE;
		uint32_t bits, cur;
		memcpy(&bits, &optical_thickness, 4);
		memcpy(&cur, &grid_tmp[indices[i]], 4);
		if (bits > cur) memcpy(&grid_tmp[indices[i]], &bits, 4); // atomicMax on the bit patterns, as the reference does
	}
}

// ema_grid_samples_nerf (:264-282)
extern "C" void oracle_nerf_ema_grid_samples(uint32_t n_elements, float decay, float* grid, const float* grid_tmp) {
	for (uint32_t i = 0; i < n_elements; ++i) {
		const float importance = grid_tmp[i], prev_val = grid[i];
		grid[i] = (prev_val < 0.f) ? prev_val : fmaxf(prev_val * decay, importance);
	}
}

// ---- render-buffer epilogue (src/render_buffer.cu:232-346 accumulate_kernel + tonemap, :533-565 tonemap_kernel)
// frames: spp x n x 4 linear premultiplied RGBA; out: n x 4. color_space / output_color_space: 0 Linear, 1 SRGB;
// curve: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard; background in sRGB.
extern "C" void oracle_render_buffer_epilogue(uint32_t n, uint32_t spp, const float* frames, float exposure, const float* background, int color_space,
	int output_color_space, int curve, float* out)
{
	for (uint32_t i = 0; i < n; ++i) {
		float acc[4] = {0, 0, 0, 0};
		for (uint32_t s = 0; s < spp; ++s) {
			const float* f = frames + ((size_t)s * n + i) * 4;
			float c[3] = {f[0], f[1], f[2]};
			if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
			const float cnt = (float)s;
			for (int k = 0; k < 3; ++k) acc[k] = (acc[k] * cnt + c[k]) / (cnt + 1);
			acc[3] = (acc[3] * cnt + f[3]) / (cnt + 1);
		}
		float bg[3] = {background[0], background[1], background[2]};
		if (color_space != 1) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
		const float weight = (1 - acc[3]) * background[3];
		float col[3];
		for (int k = 0; k < 3; ++k) col[k] = acc[k] + bg[k] * weight;
		const float alpha = acc[3] + weight;
		if (color_space == 1) for (int k = 0; k < 3; ++k) col[k] = srgb_to_linear(col[k]);
		const float gain = powf(2.0f, exposure);
		for (int k = 0; k < 3; ++k) col[k] *= gain;
		if (curve != 0) {
			for (int k = 0; k < 3; ++k) col[k] = fmaxf(col[k], 0.0f);
			if (curve == 3) {
				const float Y = 0.2126f * col[0] + 0.7152f * col[1] + 0.0722f * col[2];
				for (int k = 0; k < 3; ++k) col[k] = col[k] * (1.f / (Y + 1.0f));
			} else {
				float k0, k1, k2, k3, k4, k5;
				if (curve == 1) {
					k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
				} else {
					const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
					k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
					const float W = 11.2f;
					const float white_scale = (k3 * (W * W) + k4 * W + k5) / (k0 * (W * W) + k1 * W + k2);
					k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
				}
				for (int k = 0; k < 3; ++k) {
					const float sq = col[k] * col[k];
					col[k] = (sq * k0 + k1 * col[k] + k2) / (k3 * sq + k4 * col[k] + k5);
				}
			}
		}
		if (output_color_space == 1) for (int k = 0; k < 3; ++k) col[k] = linear_to_srgb(col[k]);
		out[(size_t)i * 4 + 0] = col[0]; out[(size_t)i * 4 + 1] = col[1]; out[(size_t)i * 4 + 2] = col[2]; out[(size_t)i * 4 + 3] = alpha;
	}
}

extern "C" float oracle_det_logf(float x) { return c_logf(x); }
extern "C" float oracle_det_expf(float x) { return c_expf(x); }
extern "C" float oracle_advance_n_steps(float t, float cone_angle, float n) { return advance_n_steps(t, cone_angle, n); }
extern "C" uint32_t oracle_morton3d(uint32_t x, uint32_t y, uint32_t z) { return morton3D(x, y, z); }
extern "C" uint32_t oracle_morton3d_invert(uint32_t x) { return morton3D_invert(x); }

// ================================== rendering ====================================================
namespace {
// random_val.cuh:161-291 (Burley 2020): Sobol dims 0/1 with nested uniform (Owen) scrambling.
inline uint32_t reverse_bits(uint32_t x) {
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
}
uint32_t sobol(uint32_t index, uint32_t dim) {
	// dimension 0: identity generator matrix (radical inverse); dimension 1: Pascal matrix mod 2
	uint32_t x = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1u) x ^= dim == 0 ? (0x80000000u >> bit) : v;
		v ^= v >> 1;
	}
	return x;
}
inline uint32_t laine_karras(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
inline uint32_t owen_scramble(uint32_t x, uint32_t seed) { return reverse_bits(laine_karras(reverse_bits(x), seed)); }
inline uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) {
	const float S = 1.0f / 4294967296.0f;
	index = owen_scramble(index, seed);
	return (float)owen_scramble(sobol(index, dim), hash_combine(seed, dim)) * S;
}
// nerf_device.cuh:461-494 with MIP_FROM_DT = false, min_mip = 0
float advance_to_next_occupied_voxel(float t, float cone_angle, V o, V d, V idir, const uint8_t* bitfield, uint32_t max_mip, const Box& box) {
	while (true) {
		const V pos = o + d * t;
		if (t >= MAX_DEPTH || !contains(box, pos)) return MAX_DEPTH;
		uint32_t mip = mip_from_pos(pos, CASCADES - 1);
		if (mip > max_mip) mip = max_mip;
		if (occupied_at(pos, bitfield, mip)) return t;
		while (mip < CASCADES - 1 && !occupied_at(pos, bitfield, mip + 1)) ++mip;   // past the scene's last cascade: csrc/nerf_device.cuh kSkipMaxMip
		t = advance_to_next_voxel(t, cone_angle, pos, d, idir, mip);
	}
}
Box box_of(const oracle_render_config& c) { return {{c.aabb_min[0], c.aabb_min[1], c.aabb_min[2]}, {c.aabb_max[0], c.aabb_max[1], c.aabb_max[2]}}; }
// One ray from its first possible sample to the far side of the box: init_rays_with_payload_kernel_nerf
// (:1475-1488), advance_pos_nerf (:333-362), then the loop of generate_next_nerf_network_inputs (:452-465)
// run to exhaustion. emit(j, t, dt) sees every sample. d must be normalised.
template <typename Emit>
uint32_t march_one_ray(const Box& box, V o, V d, uint32_t pix, uint32_t spp_index, float cone_angle, uint32_t max_mip, const uint8_t* bitfield, uint32_t max_steps, Emit emit) {
	float tmin, tmax;
	ray_intersect(box, o, d, tmin, tmax);
	float t = fmaxf(tmin, 0.0f) + 1e-6f;
	if (!contains(box, o + d * t)) return 0;
	const V idir = V{1.0f, 1.0f, 1.0f} / d;
	t = advance_n_steps(t, cone_angle, ld_random_val(spp_index, pix * 786433u, 0));
	uint32_t count = 0;
	while (count < max_steps) {
		t = advance_to_next_occupied_voxel(t, cone_angle, o, d, idir, bitfield, max_mip, box);
		if (t >= MAX_DEPTH) break;
		const float dt = calc_dt(t, cone_angle);
		emit(count, t, dt);
		++count;
		t += dt;
	}
	return count;
}
} // namespace

extern "C" float oracle_ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) { return ld_random_val(index, seed, dim); }

extern "C" uint64_t oracle_nerf_render_march(const oracle_render_config* cfg, const uint8_t* bitfield, uint32_t* numsteps, float* coords, uint64_t max_samples) {
	const Box box = box_of(*cfg);
	const uint32_t n = cfg->pixel_end - cfg->pixel_begin;
	V cam[4];
	for (int k = 0; k < 4; ++k) cam[k] = {cfg->camera[3 * k + 0], cfg->camera[3 * k + 1], cfg->camera[3 * k + 2]};
	// ld_random_pixel_offset (random_val.cuh:320-326)
	const uint32_t spp = cfg->snap_to_pixel_centers ? 0u : cfg->spp_index;
	float ox = 0.5f - ld_random_val(0, 0xdeadbeefu, 0) + ld_random_val(spp, 0xdeadbeefu, 0);
	float oy = 0.5f - ld_random_val(0, 0xdeadbeefu, 1) + ld_random_val(spp, 0xdeadbeefu, 1);
	ox -= floorf(ox);
	oy -= floorf(oy);
	uint64_t total = 0;
	for (int pass = 0; pass < 2; ++pass) {
		// pass 0 counts (serial prefix afterwards), pass 1 emits
		if (pass == 1 && !coords) break;
#pragma omp parallel for schedule(dynamic, 64)
		for (uint32_t r = 0; r < n; ++r) {
			const uint32_t pix = cfg->pixel_begin + r;
			const uint32_t x = pix % cfg->width, y = pix / cfg->width;
			const float u = ((float)x + ox) / (float)cfg->width, v = ((float)y + oy) / (float)cfg->height;
			V dd = {(u - cfg->screen_cx) * (float)cfg->width / cfg->focal_x, (v - cfg->screen_cy) * (float)cfg->height / cfg->focal_y, 1.0f};
			if (cfg->lens_mode == 1) iterative_opencv_lens_undistortion(cfg->lens_params, &dd.x, &dd.y);
			const V o = cam[3];
			const V d = normalize(cam[0] * dd.x + cam[1] * dd.y + cam[2] * dd.z);
			const V wd = (d + 1.0f) * 0.5f;
			const uint64_t base = pass == 1 ? numsteps[2 * r + 1] : 0;
			const uint32_t count = march_one_ray(box, o, d, pix, cfg->spp_index, cfg->cone_angle_constant, cfg->max_mip, bitfield, cfg->max_steps_per_ray,
				[&](uint32_t j, float t, float dt) {
					if (pass == 1 && base + j < max_samples) {
						const V wp = (o + d * t - box.min) / diag(box);
						float* c = coords + (base + j) * 7;
						c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
					}
				});
			if (pass == 0) numsteps[2 * r] = count;
		}
		if (pass == 0) {
			total = 0;
			for (uint32_t r = 0; r < n; ++r) { numsteps[2 * r + 1] = (uint32_t)total; total += numsteps[2 * r]; }
		}
	}
	return total;
}

extern "C" void oracle_nerf_render_composite(const oracle_render_config* cfg, const uint32_t* numsteps, const float* coords, const uint16_t* rgbsigma,
	float* rgba_out, float* depth_out, uint32_t* steps_out)
{
	const Box box = box_of(*cfg);
	const uint32_t n = cfg->pixel_end - cfg->pixel_begin;
	const V cam_fwd = {cfg->camera[6], cfg->camera[7], cfg->camera[8]}, cam_o = {cfg->camera[9], cfg->camera[10], cfg->camera[11]};
#pragma omp parallel for schedule(dynamic, 64)
	for (uint32_t r = 0; r < n; ++r) {
		const uint32_t count = numsteps[2 * r];
		const uint64_t base = numsteps[2 * r + 1];
		float rgba[4] = {0, 0, 0, 0};
		float depth = 0.0f, max_weight = 0.0f;
		uint32_t j = 0, taken = 0;
		for (; j < count; ++j) {
			const uint16_t* o16 = rgbsigma + (base + j) * 4;
			const float* c = coords + (base + j) * 7;
			const V pos = box.min + V{c[0], c[1], c[2]} * diag(box);
			const float T = 1.0f - rgba[3];
			const float dt = unwarp_dt(c[3]);
			const float alpha = 1.0f - c_expf(-net_to_density(oracle_half_to_float(o16[3]), cfg->density_activation) * dt);
			const float weight = alpha * T;
			for (int k = 0; k < 3; ++k) rgba[k] += net_to_rgb(oracle_half_to_float(o16[k]), cfg->rgb_activation) * weight;
			rgba[3] += weight;
			if (weight > max_weight) { max_weight = weight; depth = dot(cam_fwd, pos - cam_o); }
			taken = j + 1;
			if (rgba[3] > 1.0f - cfg->min_transmittance) {
				for (int k = 0; k < 3; ++k) rgba[k] /= rgba[3];
				rgba[3] = 1.0f;
				break;
			}
		}
		steps_out[r] = taken;
		// compact_kernel_nerf keeps rays with alpha > 0.001 (:1365); shade_kernel_nerf (:1341-1348)
		float out[4] = {0, 0, 0, 0};
		float dout = MAX_DEPTH;
		if (rgba[3] > 0.001f) {
			for (int k = 0; k < 3; ++k) out[k] = cfg->linear_colors ? rgba[k] : srgb_to_linear(rgba[k]);
			out[3] = rgba[3];
			if (out[3] > 0.2f) dout = depth;
		}
		for (int k = 0; k < 4; ++k) rgba_out[4 * r + k] = out[k];
		depth_out[r] = dout;
	}
}

// The march of explicit rays (o, d normalised; rays [n][6]), for the comparison with the reference's own functions
// composed the same way (oracle/ref_host.cu ref_march_rays): per ray the sample count, the first
// and the last sample distance.
extern "C" void oracle_march_rays(uint32_t n, const float* rays, const uint32_t* pix, uint32_t spp_index, const float* aabb, float cone_angle,
	uint32_t max_mip, const uint8_t* bitfield, uint32_t max_steps, uint32_t* counts, float* t_first, float* t_last)
{
	const Box box{{aabb[0], aabb[1], aabb[2]}, {aabb[3], aabb[4], aabb[5]}};
#pragma omp parallel for schedule(dynamic, 64)
	for (uint32_t r = 0; r < n; ++r) {
		const V o{rays[6 * r + 0], rays[6 * r + 1], rays[6 * r + 2]}, d{rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5]};
		float first = 0.0f, last = 0.0f;
		counts[r] = march_one_ray(box, o, d, pix[r], spp_index, cone_angle, max_mip, bitfield, max_steps, [&](uint32_t j, float t, float) { if (j == 0) first = t; last = t; });
		t_first[r] = first; t_last[r] = last;
	}
}

// ================================== function-grain probes ========================================
// The oracle's answer to oracle/ref_host.cu's ref_probe: same names, same rows of 32-bit words in
// and out (floats by bit pattern). tests/test_ref_pin_cpu.py compares the two word for word - against
// oracle/_ref/libref_host.so (the reference's own header functions compiled for the host) where
// it is present, and against the committed vectors it produced (tests/golden/ref_host_vectors.npz).
namespace {
inline float wf(uint32_t w) { float v; memcpy(&v, &w, 4); return v; }
inline uint32_t fw(float v) { uint32_t w; memcpy(&w, &v, 4); return w; }
} // namespace

extern "C" int oracle_probe(const char* name_c, uint32_t n, const uint32_t* in_all, uint32_t k_in, uint32_t* out_all, uint32_t k_out) {
	const std::string name = name_c;
	for (uint32_t i = 0; i < n; ++i) {
		const uint32_t* in = in_all + (size_t)i * k_in;
		uint32_t* out = out_all + (size_t)i * k_out;
		if (name == "cascaded_grid_idx_at") {
			out[0] = cascaded_grid_idx_at(V{wf(in[0]), wf(in[1]), wf(in[2])}, in[3]);
		} else if (name == "mip_from_pos") {
			out[0] = mip_from_pos(V{wf(in[0]), wf(in[1]), wf(in[2])}, in[3]);
		} else if (name == "mip_from_dt") {
			out[0] = mip_from_dt(wf(in[0]), V{wf(in[1]), wf(in[2]), wf(in[3])}, in[4]);
		} else if (name == "distance_to_next_voxel") {
			const V dir{wf(in[3]), wf(in[4]), wf(in[5])};
			out[0] = fw(distance_to_next_voxel(V{wf(in[0]), wf(in[1]), wf(in[2])}, dir, V{1.0f, 1.0f, 1.0f} / dir, wf(in[6])));
		} else if (name == "ray_intersect") {
			const Box box{V{wf(in[0]), wf(in[1]), wf(in[2])}, V{wf(in[3]), wf(in[4]), wf(in[5])}};
			float tmin, tmax;
			ray_intersect(box, V{wf(in[6]), wf(in[7]), wf(in[8])}, V{wf(in[9]), wf(in[10]), wf(in[11])}, tmin, tmax);
			out[0] = fw(tmin); out[1] = fw(tmax);
		} else if (name == "opencv_undistort") {
			const float params[7] = {wf(in[0]), wf(in[1]), wf(in[2]), wf(in[3]), 0, 0, 0};
			float x = wf(in[4]), y = wf(in[5]);
			iterative_opencv_lens_undistortion(params, &x, &y);
			out[0] = fw(x); out[1] = fw(y);
		} else if (name == "uv_to_ray") {
			oracle_nerf_image m{};
			m.res_x = (int)in[2]; m.res_y = (int)in[3];
			m.focal_x = wf(in[4]); m.focal_y = wf(in[5]); m.principal_x = wf(in[6]); m.principal_y = wf(in[7]);
			for (int k = 0; k < 12; ++k) m.xform[k] = wf(in[8 + k]);
			m.lens_mode = (int)in[20];
			for (int k = 0; k < 4; ++k) m.lens_params[k] = wf(in[21 + k]);
			V o, d;
			uv_to_ray(wf(in[0]), wf(in[1]), m, o, d);
			out[0] = fw(o.x); out[1] = fw(o.y); out[2] = fw(o.z); out[3] = fw(d.x); out[4] = fw(d.y); out[5] = fw(d.z);
		} else if (name == "srgb_to_linear") {
			out[0] = fw(srgb_to_linear(wf(in[0])));
		} else if (name == "linear_to_srgb") {
			out[0] = fw(linear_to_srgb(wf(in[0])));
		} else if (name == "loss") {
			for (int k = 0; k < 3; ++k) {
				float l, g;
				loss1(wf(in[k]), wf(in[3 + k]), (int)in[6], l, g);
				out[k] = fw(l); out[3 + k] = fw(g);
			}
		} else if (name == "network_to_rgb") {
			out[0] = fw(net_to_rgb(wf(in[0]), (int)in[1])); out[1] = fw(net_to_rgb_d(wf(in[0]), (int)in[1]));
		} else if (name == "network_to_density") {
			out[0] = fw(net_to_density(wf(in[0]), (int)in[1])); out[1] = fw(net_to_density_d(wf(in[0]), (int)in[1]));
		} else if (name == "image_idx") {
			out[0] = image_idx(in[0], in[1], in[2]);
		} else if (name == "ld_random_val") {
			out[0] = fw(ld_random_val(in[0], in[1], in[2]));
		} else if (name == "warp_dt") {
			out[0] = fw(warp_dt(wf(in[0]))); out[1] = fw(unwarp_dt(wf(in[0])));
		} else if (name == "stepping") {
			const float t = wf(in[0]), cone = wf(in[1]), steps = wf(in[2]);
			out[0] = fw(to_stepping_space(t, cone)); out[1] = fw(from_stepping_space(steps, cone));
			out[2] = fw(calc_dt(t, cone)); out[3] = fw(advance_n_steps(t, cone, steps));
		} else if (name == "morton3D") {
			out[0] = morton3D(in[0], in[1], in[2]);
		} else if (name == "morton3D_invert") {
			out[0] = morton3D_invert(in[0]);
		} else {
			return -1;
		}
	}
	return 0;
}
