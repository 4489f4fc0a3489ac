// Device-side NeRF math shared by the training, occupancy-grid and render kernels.
//
// Restates (as arithmetic, not as code) the helpers the reference keeps in
//   include/neural-graphics-primitives/nerf_device.cuh   (stepping, cascades, activations, losses)
//   include/neural-graphics-primitives/common_device.cuh (lens, uv_to_ray, read_rgba, sRGB)
//   include/neural-graphics-primitives/bounding_box.cuh  (ray/box intersection)
//   tcnn common_device.h:759-785                         (Morton codes)
//
// Determinism contract: every function that decides WHERE a sample lies (ray set-up, box
// clipping, stepping, mip selection, bitfield lookup) is written with IEEE add/mul/div/sqrt
// only, the files that include this header are compiled with -fmad=false, and the two
// transcendentals the exponential stepping needs (log, exp) are the fixed polynomial
// implementations below instead of libm / fast-math intrinsics. The CPU oracle restates the
// same arithmetic, so sample positions and per-ray sample counts agree bit for bit between
// the B200 and the host (the reference itself is built with --use_fast_math, CMakeLists.txt:82,
// so its own last bits are not reproducible across GPUs either).
#pragma once

#include "common.cuh"

#include <cfloat>
#include <cstring>

namespace ngpb {

// ---- constants (nerf_device.cuh:20-40, common_device.cuh:32) --------------------------------
constexpr uint32_t kNerfGridSize = 128;
constexpr uint32_t kNerfGridCells = kNerfGridSize * kNerfGridSize * kNerfGridSize;
constexpr uint32_t kNerfSteps = 1024;
constexpr uint32_t kNerfCascades = 8;
constexpr float kSqrt3 = 1.73205080757f;
constexpr float kMinConeStepsize = kSqrt3 / kNerfSteps;
constexpr float kMaxConeStepsize = kMinConeStepsize * (1 << (kNerfCascades - 1)) * kNerfSteps / kNerfGridSize;
constexpr uint32_t kMaxRandomSamplesPerRay = 16;
constexpr float kNerfMinOpticalThickness = 0.01f;
constexpr float kMaxDepth = 16384.0f;

// ---- small vector type ----------------------------------------------------------------------
struct V3 {
	float x, y, z;
};
NGPB_HD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
NGPB_HD V3 v3(float s) { return V3{s, s, s}; }
NGPB_HD V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
NGPB_HD V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
NGPB_HD V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
NGPB_HD V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
NGPB_HD V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
NGPB_HD V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
NGPB_HD V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
NGPB_HD V3 operator+(V3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
NGPB_HD V3 operator-(V3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
NGPB_HD V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
NGPB_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
NGPB_HD float length(V3 a) { return sqrtf(dot(a, a)); }
NGPB_HD V3 normalize(V3 a) { return a / length(a); }
NGPB_HD float distance(V3 a, V3 b) { return length(a - b); }
NGPB_HD float mean(V3 a) { return (a.x + a.y + a.z) / 3.0f; }

// 4x3 camera matrix, column-major like tcnn's mat4x3: c[0..2] rotation columns, c[3] origin.
struct Mat4x3 {
	V3 c[4];
};
NGPB_HD V3 mul3(const Mat4x3& m, V3 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z; }

struct Ray3 {
	V3 o, d;
};

struct Aabb {
	V3 min, max;
};
NGPB_HD bool aabb_contains(const Aabb& b, V3 p) {
	return p.x >= b.min.x && p.x <= b.max.x && p.y >= b.min.y && p.y <= b.max.y && p.z >= b.min.z && p.z <= b.max.z;
}
NGPB_HD V3 aabb_diag(const Aabb& b) { return b.max - b.min; }
// bounding_box.cuh:172-219
NGPB_HD void aabb_ray_intersect(const Aabb& b, V3 pos, V3 dir, float& out_tmin, float& out_tmax) {
	float tmin = (b.min.x - pos.x) / dir.x, tmax = (b.max.x - pos.x) / dir.x;
	if (tmin > tmax) { float t = tmin; tmin = tmax; tmax = t; }
	float tymin = (b.min.y - pos.y) / dir.y, tymax = (b.max.y - pos.y) / dir.y;
	if (tymin > tymax) { float t = tymin; tymin = tymax; tymax = t; }
	if (tmin > tymax || tymin > tmax) { out_tmin = out_tmax = FLT_MAX; return; }
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = (b.min.z - pos.z) / dir.z, tzmax = (b.max.z - pos.z) / dir.z;
	if (tzmin > tzmax) { float t = tzmin; tzmin = tzmax; tzmax = t; }
	if (tmin > tzmax || tzmin > tmax) { out_tmin = out_tmax = FLT_MAX; return; }
	if (tzmin > tmin) tmin = tzmin;
	if (tzmax < tmax) tmax = tzmax;
	out_tmin = tmin;
	out_tmax = tmax;
}

// ---- Morton codes (tcnn common_device.h:759-785) ----------------------------------------------
NGPB_HD uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
NGPB_HD uint32_t morton3d(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
NGPB_HD uint32_t morton3d_invert(uint32_t x) {
	x = x & 0x49249249;
	x = (x | (x >> 2)) & 0xc30c30c3;
	x = (x | (x >> 4)) & 0x0f00f00f;
	x = (x | (x >> 8)) & 0xff0000ff;
	x = (x | (x >> 16)) & 0x0000ffff;
	return x;
}

// ---- deterministic log / exp -------------------------------------------------------------------
// Single-precision Cephes algorithms (S. Moshier, cephes/single/logf.c, expf.c: range reduction by
// frexp / round-to-nearest multiple of ln 2 split in two constants, degree-8 / degree-5
// polynomials), evaluated in Estrin form with explicit fmaf(): a fused multiply-add is correctly
// rounded on the device (FFMA) and on the host (FMA3 or glibc), so both give the same bits, and
// the dependent chain is 9 operations instead of 25 - the serial per-ray march is latency bound.
// Bit-level frexp / ldexp for positive normal numbers (every caller's range): the same values as
// frexpf / scalbnf without their special-case handling, which dominates the instruction count of
// the serial march loop.
NGPB_HD uint32_t f32_bits(float x) {
#ifdef __CUDA_ARCH__
	return __float_as_uint(x);
#else
	uint32_t u; memcpy(&u, &x, 4); return u;
#endif
}
NGPB_HD float bits_f32(uint32_t u) {
#ifdef __CUDA_ARCH__
	return __uint_as_float(u);
#else
	float x; memcpy(&x, &u, 4); return x;
#endif
}
NGPB_HD float frexp_pos(float x, int& e) {
	const uint32_t u = f32_bits(x);
	e = (int)(u >> 23) - 126;
	return bits_f32((u & 0x007FFFFFu) | 0x3F000000u);
}
NGPB_HD float pow2i(int n) { return bits_f32((uint32_t)(n + 127) << 23); } // 2^n, -126 <= n <= 127
// exponent frexpf would report, for x >= 0 (0 -> 0; subnormals report -126, callers clamp at 0)
NGPB_HD int frexp_exponent(float x) { return x == 0.0f ? 0 : (int)(f32_bits(x) >> 23) - 126; }

NGPB_HD float det_logf(float xx) {
	int e;
	float x = frexp_pos(xx, e);
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
NGPB_HD float det_expf(float x) {
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
	return (fmaf(p, x2, x) + 1.0f) * pow2i(n);
}

// ---- exponential stepping (nerf_device.cuh:372-437) ------------------------------------------
// The reference recomputes log(1+c), a, b, exp(a log(1+c)), exp(b log(1+c)) on every call; they
// only depend on the cone angle, so they are tabulated once per launch.
// Divisions by the three per-launch constants are multiplications by their (once rounded)
// reciprocals, here and in the oracle; the reference is built with --use_fast_math, which makes
// the same substitution.
constexpr float kInvMinConeStepsize = 1.0f / kMinConeStepsize;
constexpr float kInvMaxConeStepsize = 1.0f / kMaxConeStepsize;
struct Stepping {
	float cone_angle;
	float log1p_c, a, b, at, bt, inv_log1p_c;
};
NGPB_HD Stepping make_stepping(float cone_angle) {
	Stepping s{};
	s.cone_angle = cone_angle;
	if (cone_angle <= 1e-5f) return s;
	s.log1p_c = det_logf(1.0f + cone_angle);
	s.a = (det_logf(kMinConeStepsize) - det_logf(s.log1p_c)) / s.log1p_c;
	s.b = (det_logf(kMaxConeStepsize) - det_logf(s.log1p_c)) / s.log1p_c;
	s.inv_log1p_c = 1.0f / s.log1p_c;
	s.at = det_expf(s.a * s.log1p_c);
	s.bt = det_expf(s.b * s.log1p_c);
	return s;
}
NGPB_HD float to_stepping_space(float t, const Stepping& s) {
	if (s.cone_angle <= 1e-5f) return t * kInvMinConeStepsize;
	if (t <= s.at) return (t - s.at) * kInvMinConeStepsize + s.a;
	if (t <= s.bt) return det_logf(t) * s.inv_log1p_c;
	return (t - s.bt) * kInvMaxConeStepsize + s.b;
}
NGPB_HD float from_stepping_space(float n, const Stepping& s) {
	if (s.cone_angle <= 1e-5f) return n * kMinConeStepsize;
	if (n <= s.a) return (n - s.a) * kMinConeStepsize + s.at;
	if (n <= s.b) return det_expf(n * s.log1p_c);
	return (n - s.b) * kMaxConeStepsize + s.bt;
}
// Select forms for the march loops: the three regimes are evaluated side by side and selected, so
// the compiler can interleave independent calls instead of emitting one branchy block per call
// (the serial march is bound by dependent-issue latency). Each regime is the expression above:
// same bits. kCone = false is the cone_angle <= 1e-5 case.
template <bool kCone> NGPB_HD float to_stepping_space_sel(float t, const Stepping& s) {
	if (!kCone) return t * kInvMinConeStepsize;
	const float lo = (t - s.at) * kInvMinConeStepsize + s.a;
	const float mid = det_logf(t) * s.inv_log1p_c;
	const float hi = (t - s.bt) * kInvMaxConeStepsize + s.b;
	return t <= s.at ? lo : (t <= s.bt ? mid : hi);
}
template <bool kCone> NGPB_HD float from_stepping_space_sel(float n, const Stepping& s) {
	if (!kCone) return n * kMinConeStepsize;
	const float lo = (n - s.a) * kMinConeStepsize + s.at;
	const float mid = det_expf(n * s.log1p_c);
	const float hi = (n - s.b) * kMaxConeStepsize + s.bt;
	return n <= s.a ? lo : (n <= s.b ? mid : hi);
}
NGPB_HD float advance_n_steps(float t, const Stepping& s, float n) { return from_stepping_space(to_stepping_space(t, s) + n, s); }
NGPB_HD float calc_dt(float t, const Stepping& s) { return advance_n_steps(t, s, 1.0f) - t; }

NGPB_HD float sign_f(float x) { return copysignf(1.0f, x); }
// nerf_device.cuh:362-370
NGPB_HD float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, float res, float inv_res) {
	const V3 p = res * (pos - 0.5f);
	const float tx = (floorf(p.x + 0.5f + 0.5f * sign_f(dir.x)) - p.x) * idir.x;
	const float ty = (floorf(p.y + 0.5f + 0.5f * sign_f(dir.y)) - p.y) * idir.y;
	const float tz = (floorf(p.z + 0.5f + 0.5f * sign_f(dir.z)) - p.z) * idir.z;
	const float t = fminf(fminf(tx, ty), tz);
	return fmaxf(t * inv_res, 0.0f); // res is a power of two: same bits as t / res
}
// nerf_device.cuh:426-437
// The same with to_stepping_space(t) supplied by the caller (the march loop already has it from
// calc_dt): identical operations, one logarithm less on the serial path.
NGPB_HD float advance_to_next_voxel_n(float t, float n, const Stepping& s, V3 pos, V3 dir, V3 idir, uint32_t mip) {
	const float res = pow2i(7 - (int)mip);
	const float t_target = to_stepping_space(t + distance_to_next_voxel(pos, dir, idir, res, pow2i((int)mip - 7)), s);
	return from_stepping_space(n + ceilf(fmaxf(t_target - n, 0.5f)), s);
}
NGPB_HD float advance_to_next_voxel(float t, const Stepping& s, V3 pos, V3 dir, V3 idir, uint32_t mip) {
	const float res = pow2i(7 - (int)mip); // kNerfGridSize / 2^mip
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res, pow2i((int)mip - 7));
	t = to_stepping_space(t, s);
	t_target = to_stepping_space(t_target, s);
	return from_stepping_space(t + ceilf(fmaxf(t_target - t, 0.5f)), s);
}

// ---- cascaded occupancy grid (nerf_device.cuh:316-340, 439-458) -------------------------------
NGPB_HD int clamp_i(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
NGPB_HD uint32_t mip_from_pos(V3 pos, uint32_t max_cascade = kNerfCascades - 1) {
	const float maxval = fmaxf(fmaxf(fabsf(pos.x - 0.5f), fabsf(pos.y - 0.5f)), fabsf(pos.z - 0.5f));
	return (uint32_t)clamp_i(frexp_exponent(maxval) + 1, 0, (int)max_cascade);
}
NGPB_HD uint32_t mip_from_dt(float dt, V3 pos, uint32_t max_cascade = kNerfCascades - 1) {
	const uint32_t mip = mip_from_pos(pos, max_cascade);
	dt *= 2 * kNerfGridSize;
	if (dt < 1.0f) return mip;
	return (uint32_t)clamp_i((int)mip, frexp_exponent(dt), (int)max_cascade);
}
NGPB_HD uint32_t cascaded_grid_idx_at(V3 pos, uint32_t mip) {
	const float mip_scale = pow2i(-(int)mip);
	pos = pos - 0.5f;
	pos = pos * mip_scale;
	pos = pos + 0.5f;
	const int ix = (int)(pos.x * (float)kNerfGridSize), iy = (int)(pos.y * (float)kNerfGridSize), iz = (int)(pos.z * (float)kNerfGridSize);
	if (ix < 0 || ix >= (int)kNerfGridSize || iy < 0 || iy >= (int)kNerfGridSize || iz < 0 || iz >= (int)kNerfGridSize) return 0xFFFFFFFFu;
	return morton3d((uint32_t)ix, (uint32_t)iy, (uint32_t)iz);
}
NGPB_HD uint32_t grid_mip_offset(uint32_t mip) { return kNerfGridCells * mip; }
NGPB_HD bool density_grid_occupied_at(V3 pos, const uint8_t* __restrict__ bitfield, uint32_t mip) {
	const uint32_t idx = cascaded_grid_idx_at(pos, mip);
	if (idx == 0xFFFFFFFFu) return false;
	return (bitfield[idx / 8 + grid_mip_offset(mip) / 8] & (1u << (idx % 8))) != 0;
}

// ---- warps (nerf_device.cuh:265-314) -----------------------------------------------------------
NGPB_HD V3 warp_position(V3 pos, const Aabb& aabb) { return (pos - aabb.min) / aabb_diag(aabb); }
NGPB_HD V3 unwarp_position(V3 pos, const Aabb& aabb) { return aabb.min + pos * aabb_diag(aabb); }
NGPB_HD V3 warp_direction(V3 dir) { return (dir + 1.0f) * 0.5f; }
NGPB_HD float warp_dt(float dt) {
	const float max_stepsize = kMinConeStepsize * (1 << (kNerfCascades - 1));
	return (dt - kMinConeStepsize) / (max_stepsize - kMinConeStepsize);
}
NGPB_HD float unwarp_dt(float dt) {
	const float max_stepsize = kMinConeStepsize * (1 << (kNerfCascades - 1));
	return dt * (max_stepsize - kMinConeStepsize) + kMinConeStepsize;
}

// ---- activations (nerf_device.cuh:203-263) -----------------------------------------------------
// Every exponential on the compositing path is the fixed polynomial det_expf, here and in the oracle:
// the transmittance a ray has accumulated decides where it stops (T < 1e-4 in training,
// alpha > 1 - min_transmittance in rendering), so per-ray sample counts can only be bit-exact between
// B200 and the host if the exponentials are. The reference uses __expf (its build has
// --use_fast_math, CMakeLists.txt:82), whose bits are not reproducible off the GPU either.
enum NerfActivation : int { kActNone = 0, kActReLU = 1, kActLogistic = 2, kActExponential = 3 };
NGPB_HD float logistic_f(float x) { return 1.0f / (1.0f + det_expf(-x)); }
NGPB_HD float clamp_f(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
NGPB_HD float network_to_rgb(float val, int act) {
	switch (act) {
		case kActReLU: return val > 0.0f ? val : 0.0f;
		case kActLogistic: return logistic_f(val);
		case kActExponential: return det_expf(clamp_f(val, -10.0f, 10.0f));
		default: return val;
	}
}
NGPB_HD float network_to_rgb_derivative(float val, int act) {
	switch (act) {
		case kActReLU: return val > 0.0f ? 1.0f : 0.0f;
		case kActLogistic: { const float d = logistic_f(val); return d * (1 - d); }
		case kActExponential: return det_expf(clamp_f(val, -10.0f, 10.0f));
		default: return 1.0f;
	}
}
NGPB_HD float network_to_density(float val, int act) {
	switch (act) {
		case kActReLU: return val > 0.0f ? val : 0.0f;
		case kActLogistic: return logistic_f(val);
		case kActExponential: return det_expf(val);
		default: return val;
	}
}
NGPB_HD float network_to_density_derivative(float val, int act) {
	switch (act) {
		case kActReLU: return val > 0.0f ? 1.0f : 0.0f;
		case kActLogistic: { const float d = logistic_f(val); return d * (1 - d); }
		case kActExponential: return det_expf(clamp_f(val, -15.0f, 15.0f));
		default: return 1.0f;
	}
}

// alpha of one sample (src/testbed_nerf.cu:926, :524): 1 - exp(-density dt)
NGPB_HD float composite_alpha(float density, float dt) { return 1.0f - det_expf(-density * dt); }

// ---- losses (nerf_device.cuh:74-143, 600-615) ---------------------------------------------------
enum NerfLoss : int { kLossL2 = 0, kLossL1 = 1, kLossMape = 2, kLossSmape = 3, kLossHuber = 4, kLossLogL1 = 5, kLossRelativeL2 = 6 };
struct LossAndGradient {
	V3 loss, gradient;
};
NGPB_HD void loss1(float target, float pred, int type, float& loss, float& grad) {
	const float d = pred - target;
	switch (type) {
		case kLossRelativeL2: { const float den = pred * pred + 1e-2f; loss = d * d / den; grad = 2.0f * d / den; } break;
		case kLossL1: loss = fabsf(d); grad = copysignf(1.0f, d); break;
		case kLossMape: { const float den = fabsf(pred) + 1e-2f; loss = fabsf(d) / den; grad = copysignf(1.0f / den, d); } break;
		case kLossSmape: { const float den = 0.5f * (fabsf(pred) + fabsf(target)) + 1e-2f; loss = fabsf(d) / den; grad = copysignf(1.0f / den, d); } break;
		case kLossHuber: {
			// huber_loss(alpha = 0.1) / 5 (nerf_device.cuh:607-610)
			const float alpha = 0.1f, ad = fabsf(d);
			const float sq = 0.5f / alpha * d * d;
			loss = (ad > alpha ? (ad - 0.5f * alpha) : sq) / 5.0f;
			grad = (ad > alpha ? (d > 0 ? 1.0f : -1.0f) : (d / alpha)) / 5.0f;
		} break;
		case kLossLogL1: { const float div = fabsf(d) + 1.0f; loss = det_logf(div); grad = copysignf(1.0f / div, d); } break;
		default: loss = d * d; grad = 2.0f * d; break;
	}
}
NGPB_HD LossAndGradient loss_and_gradient(V3 target, V3 pred, int type) {
	LossAndGradient lg;
	loss1(target.x, pred.x, type, lg.loss.x, lg.gradient.x);
	loss1(target.y, pred.y, type, lg.loss.y, lg.gradient.y);
	loss1(target.z, pred.z, type, lg.loss.z, lg.gradient.z);
	return lg;
}

// ---- colour (common_device.cuh:34-80) -----------------------------------------------------------
NGPB_HD float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
NGPB_HD float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }
NGPB_HD V3 srgb_to_linear(V3 c) { return {srgb_to_linear(c.x), srgb_to_linear(c.y), srgb_to_linear(c.z)}; }
NGPB_HD V3 linear_to_srgb(V3 c) { return {linear_to_srgb(c.x), linear_to_srgb(c.y), linear_to_srgb(c.z)}; }

// ---- training images (nerf_device.cuh:45-61, common_device.cuh:730-806) --------------------------
enum ImageDataType : int { kImageNone = 0, kImageByte = 1, kImageHalf = 2, kImageFloat = 3 };
enum LensMode : int { kLensPerspective = 0, kLensOpenCV = 1, kLensFTheta = 2, kLensLatLong = 3, kLensOpenCVFisheye = 4, kLensEquirectangular = 5 };
enum ColorSpace : int { kColorLinear = 0, kColorSRGB = 1 };

// POD mirrored by the C ABI (ngpb_nerf_image_t) and by TrainingImageMetadata in the reference.
struct ImageMeta {
	const void* pixels;
	int image_data_type;
	int lens_mode;
	float lens_params[7];
	int res_x, res_y;
	float principal_x, principal_y;
	float focal_x, focal_y;
	Mat4x3 xform; // TrainingXForm::start; rolling shutter / motion blur are not supported
};

struct Rgba {
	float r, g, b, a;
};
NGPB_D Rgba read_rgba(int px, int py, int res_x, const void* pixels, int type) {
	const size_t idx = (size_t)px + (size_t)py * res_x;
	switch (type) {
		case kImageByte: {
			const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(pixels) + idx);
			if (v == 0x00FF00FFu) return {-1.0f, -1.0f, -1.0f, -1.0f};
			const float alpha = (float)(v >> 24) * (1.0f / 255.0f);
			return {srgb_to_linear((float)(v & 0xFF) * (1.0f / 255.0f)) * alpha, srgb_to_linear((float)((v >> 8) & 0xFF) * (1.0f / 255.0f)) * alpha,
				srgb_to_linear((float)((v >> 16) & 0xFF) * (1.0f / 255.0f)) * alpha, alpha};
		}
		case kImageHalf: {
			const uint2 v = __ldg(reinterpret_cast<const uint2*>(pixels) + idx);
			const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
			return {a.x, a.y, b.x, b.y};
		}
		case kImageFloat: {
			const float4 v = __ldg(reinterpret_cast<const float4*>(pixels) + idx);
			return {v.x, v.y, v.z, v.w};
		}
		default: return {5.0f, 0.0f, 0.0f, 1.0f};
	}
}
NGPB_D Rgba read_rgba_uv(float u, float v, int res_x, int res_y, const void* pixels, int type) {
	const int px = clamp_i((int)(u * (float)res_x), 0, res_x - 1), py = clamp_i((int)(v * (float)res_y), 0, res_y - 1);
	return read_rgba(px, py, res_x, pixels, type);
}

// ---- lens (common_device.cuh:248-338, 393-460) ----------------------------------------------------
NGPB_HD void opencv_lens_distortion_delta(const float* p, float u, float v, float* du, float* dv) {
	const float k1 = p[0], k2 = p[1], p1 = p[2], p2 = p[3];
	const float u2 = u * u, uv = u * v, v2 = v * v, r2 = u2 + v2;
	const float radial = k1 * r2 + k2 * r2 * r2;
	*du = u * radial + 2.0f * p1 * uv + p2 * (r2 + 2.0f * u2);
	*dv = v * radial + 2.0f * p2 * uv + p1 * (r2 + 2.0f * v2);
}
// Newton iteration with central differences, 100 iterations max (common_device.cuh:288-329).
NGPB_HD void iterative_opencv_lens_undistortion(const float* params, float* u, float* v) {
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
		// J[col][row], as tcnn mat2
		const float j00 = 1 + (f00 - b00) / (2 * step0);
		const float j10 = (f10 - b10) / (2 * step1);
		const float j01 = (f01 - b01) / (2 * step0);
		const float j11 = 1 + (f11 - b11) / (2 * step1);
		const float rx = x + dx0 - x0, ry = y + dx1 - y0;
		// inverse(J) * r with the adjugate / determinant (tcnn vec.h inverse(mat2))
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

// uv_to_ray for the training / render cameras: perspective and OpenCV lenses, no foveation, no
// aperture, no parallax (all zero / identity in the headless paths).
NGPB_HD Ray3 uv_to_ray(float u, float v, int res_x, int res_y, float focal_x, float focal_y, const Mat4x3& cam, float screen_cx, float screen_cy,
	int lens_mode, const float* lens_params, float near_distance)
{
	V3 dir = {(u - screen_cx) * (float)res_x / focal_x, (v - screen_cy) * (float)res_y / focal_y, 1.0f};
	if (lens_mode == kLensOpenCV) iterative_opencv_lens_undistortion(lens_params, &dir.x, &dir.y);
	dir = mul3(cam, dir);
	V3 origin = cam.c[3];
	origin = origin + dir * near_distance;
	return {origin, dir};
}

// ---- per-ray random choices (nerf_device.cuh:552-598, src/testbed_nerf.cu:712-727) ------------------
NGPB_HD uint32_t image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_training_images) {
	return (uint32_t)((((uint64_t)base_idx * n_training_images) & 0xFFFFFFFFu) / n_rays) % n_training_images;
}
NGPB_HD void random_image_pos_training(Pcg32& rng, int res_x, int res_y, bool snap_to_pixel_centers, float& u, float& v) {
	u = rng.next_float();
	v = rng.next_float();
	if (snap_to_pixel_centers) {
		u = ((float)clamp_i((int)(u * (float)res_x), 0, res_x - 1) + 0.5f) / (float)res_x;
		v = ((float)clamp_i((int)(v * (float)res_y), 0, res_y - 1) + 0.5f) / (float)res_y;
	}
}

// Empty-space skipping climbs past the scene's last cascade. The bitfield holds all 8 cascades whatever aabb_scale is,
// and update_density_grid_mean_and_bitfield max-pools every one from the next finer (src/testbed_nerf.cu:2362-2379), so
// cascade m > max_mip is a valid, coarser occupancy of the same scene: a 2^m / 128 voxel that is empty there contains no
// occupied finer voxel. The reference stops the climb at max_mip (nerf_device.cuh:485-490) and crosses a large empty
// region in steps of its coarsest voxel - up to ~220 dependent iterations on a ray through an aabb_scale 4 box, which
// is what a render window waits for. The samples are the same: a skipped stretch holds no occupied candidate either
// way, and both routes land on the first lattice point behind it.
constexpr uint32_t kSkipMaxMip = kNerfCascades - 1;

// ---- render-time marching (nerf_device.cuh:461-494) --------------------------------------------
// if_unoccupied_advance_to_next_occupied_voxel<MIP_FROM_DT=false>: skip to the next occupied cell,
// climbing to the coarsest empty cascade around the sample first. Returns kMaxDepth once the ray
// leaves the box (render_aabb == train aabb, identity render_aabb_to_local in the headless path).
NGPB_HD float advance_to_next_occupied_voxel(float t, const Stepping& s, V3 o, V3 d, V3 idir, const uint8_t* __restrict__ bitfield,
	uint32_t min_mip, uint32_t max_mip, const Aabb& aabb)
{
	while (true) {
		const V3 pos = o + d * t;
		if (t >= kMaxDepth || !aabb_contains(aabb, pos)) return kMaxDepth;
		uint32_t mip = mip_from_pos(pos);
		mip = mip < min_mip ? min_mip : (mip > max_mip ? max_mip : mip);
		if (density_grid_occupied_at(pos, bitfield, mip)) return t;
		while (mip < kSkipMaxMip && !density_grid_occupied_at(pos, bitfield, mip + 1)) ++mip;
		t = advance_to_next_voxel(t, s, pos, d, idir, mip);
	}
}

// The same loop on the select forms of the stepping-space maps (the two logarithms of a skip are
// independent and interleave): same operations, same operands, same t.
template <bool kCone>
NGPB_HD float advance_to_next_occupied_voxel_sel(float t, const Stepping& s, V3 o, V3 d, V3 idir, const uint8_t* __restrict__ bitfield,
	uint32_t min_mip, uint32_t max_mip, const Aabb& aabb, uint32_t skip_mip = kSkipMaxMip)
{
	while (true) {
		const V3 pos = o + d * t;
		if (t >= kMaxDepth || !aabb_contains(aabb, pos)) return kMaxDepth;
		uint32_t mip = mip_from_pos(pos);
		mip = mip < min_mip ? min_mip : (mip > max_mip ? max_mip : mip);
		if (density_grid_occupied_at(pos, bitfield, mip)) return t;
		while (mip < skip_mip && !density_grid_occupied_at(pos, bitfield, mip + 1)) ++mip;
		const float n_target = to_stepping_space_sel<kCone>(t + distance_to_next_voxel(pos, d, idir, pow2i(7 - (int)mip), pow2i((int)mip - 7)), s);
		const float n = to_stepping_space_sel<kCone>(t, s);
		t = from_stepping_space_sel<kCone>(n + ceilf(fmaxf(n_target - n, 0.5f)), s);
	}
}

// ---- low-discrepancy values (random_val.cuh:161-327; Burley 2020, "Practical Hash-based Owen
// Scrambling", JCGT 9(4)). Only Sobol dimensions 0 and 1 are needed: dimension 0 is the bit
// reversal of the index, the direction numbers of dimension 1 follow v_k = v_{k-1} ^ (v_{k-1} >> 1).
NGPB_HD uint32_t reverse_bits32(uint32_t x) {
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
}
NGPB_HD uint32_t sobol01(uint32_t index, uint32_t dim) {
	if (dim == 0) return reverse_bits32(index);
	uint32_t x = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		if ((index >> bit) & 1u) x ^= v;
		v ^= v >> 1;
	}
	return x;
}
NGPB_HD uint32_t laine_karras_permutation(uint32_t x, uint32_t seed) {
	x += seed;
	x ^= x * 0x6c50b47cu;
	x ^= x * 0xb82f1e52u;
	x ^= x * 0xc7afe638u;
	x ^= x * 0x8d22f6e6u;
	return x;
}
NGPB_HD uint32_t nested_uniform_scramble_base2(uint32_t x, uint32_t seed) {
	return reverse_bits32(laine_karras_permutation(reverse_bits32(x), seed));
}
NGPB_HD uint32_t hash_combine_u32(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
NGPB_HD float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim = 0) {
	const float S = 1.0f / 4294967296.0f;
	index = nested_uniform_scramble_base2(index, seed);
	return (float)nested_uniform_scramble_base2(sobol01(index, dim), hash_combine_u32(seed, dim)) * S;
}
// ld_random_pixel_offset (random_val.cuh:321-326)
NGPB_HD void ld_random_pixel_offset(uint32_t spp, float& ox, float& oy) {
	const uint32_t seed = 0xdeadbeefu;
	ox = 0.5f - ld_random_val(0, seed, 0) + ld_random_val(spp, seed, 0);
	oy = 0.5f - ld_random_val(0, seed, 1) + ld_random_val(spp, seed, 1);
	ox -= floorf(ox);
	oy -= floorf(oy);
}

} // namespace ngpb
