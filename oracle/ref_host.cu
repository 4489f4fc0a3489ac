// oracle/_ref/libref_host.so: the REFERENCE'S OWN device functions, compiled for the host and run on
// the CPU - test infrastructure that pins the oracle restatement (oracle/*.cpp), nothing else.
//
// The reference has no CPU path and its kernels need a GPU, but the arithmetic its kernels are made
// of lives in header functions: the NGP_HOST_DEVICE helpers of
// include/neural-graphics-primitives/{nerf_device.cuh, common_device.cuh, bounding_box.cuh, random_val.cuh}
// and the __device__ helpers of dependencies/tiny-cuda-nn/include/tiny-cuda-nn/common_device.h
// (grid_index, the hashes, pos_fract, sh_enc), which become callable on the host when __device__ is
// widened to __host__ __device__ around the include. This file only INCLUDES those headers from
// where they lie under /root/reference (nothing is copied) and exposes them through one probe call:
// rows of 32-bit words in, rows of 32-bit words out (floats by bit pattern), so that a comparison
// of the words is a bit-exact comparison. Built by oracle/build_ref.sh when /root/reference exists;
// tests/golden/make_ref_vectors.py turns its outputs into the committed golden vectors.
#include <cstdio>
#include <cmath>
#include <cstring>
#include <cstdint>
#include <string>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <tiny-cuda-nn/common.h>
#undef __device__
#define __device__ __attribute__((host)) __attribute__((device))
#include <tiny-cuda-nn/common_device.h>
#undef __device__
#define __device__ __attribute__((device))
#include <pcg32/pcg32.h>
#include <neural-graphics-primitives/bounding_box.cuh>
#include <neural-graphics-primitives/nerf_device.cuh>

using namespace ngp;

namespace {
inline float f(uint32_t w) { float v; memcpy(&v, &w, 4); return v; }
inline uint32_t u(float v) { uint32_t w; memcpy(&w, &v, 4); return w; }

// corner loop of kernel_grid (tcnn grid.h:130-150): pos_fract per dimension, then grid_index per corner
template <uint32_t D>
void grid_corner_indices(const uint32_t* in, uint32_t* out) {
	const float scale = f(in[D]);
	const uint32_t resolution = in[D + 1], hashmap_size = in[D + 2];
	float pos[D];
	tcnn::uvec<D> pos_grid;
	for (uint32_t d = 0; d < D; ++d) tcnn::pos_fract(f(in[d]), &pos[d], &pos_grid[d], scale, tcnn::identity_fun);
	for (uint32_t idx = 0; idx < (1u << D); ++idx) {
		tcnn::uvec<D> local;
		for (uint32_t d = 0; d < D; ++d) local[d] = (idx & (1u << d)) == 0 ? pos_grid[d] : pos_grid[d] + 1;
		out[idx] = tcnn::grid_index<D, tcnn::HashType::CoherentPrime>(tcnn::GridType::Hash, hashmap_size, resolution, local);
	}
	for (uint32_t d = 0; d < D; ++d) out[(1u << D) + d] = u(pos[d]);
}
} // namespace

extern "C" int ref_probe(const char* name_c, uint32_t n, const uint32_t* in_all, uint32_t k_in, uint32_t* out_all, uint32_t k_out) {
	const std::string name = name_c;
	for (uint32_t i = 0; i < n; ++i) {
		const uint32_t* in = in_all + (size_t)i * k_in;
		uint32_t* out = out_all + (size_t)i * k_out;
		if (name == "grid_level") { // tcnn common_device.h:709-718
			// log2 of the per-level scale as GridEncodingTemplated hands it to the kernels (grid.h: std::log2(m_per_level_scale))
			const float s = tcnn::grid_scale(in[0], std::log2(f(in[1])), in[2]);
			out[0] = u(s); out[1] = tcnn::grid_resolution(s);
		} else if (name == "grid_index3") {
			grid_corner_indices<3>(in, out);
		} else if (name == "grid_index2") {
			grid_corner_indices<2>(in, out);
		} else if (name == "morton3D") {
			out[0] = tcnn::morton3D(in[0], in[1], in[2]);
		} else if (name == "morton3D_invert") {
			out[0] = tcnn::morton3D_invert(in[0]);
		} else if (name == "sh4") { // kernel_sh (spherical_harmonics.h:46-72): inputs in [0,1] are mapped to [-1,1]
			struct { float v[16]; float& operator()(uint32_t k) { return v[k]; } } o;
			tcnn::sh_enc<float>(4, f(in[0]) * 2.f - 1.f, f(in[1]) * 2.f - 1.f, f(in[2]) * 2.f - 1.f, o);
			for (int k = 0; k < 16; ++k) out[k] = u(o.v[k]);
		} else if (name == "pcg32") { // pcg32 rng{seed}; advance; next_uint, next_float
			pcg32 rng{(uint64_t)in[0] | ((uint64_t)in[1] << 32)};
			rng.advance((int64_t)in[2]);
			out[0] = rng.next_uint(); out[1] = u(rng.next_float());
		} else if (name == "cascaded_grid_idx_at") {
			out[0] = cascaded_grid_idx_at(vec3{f(in[0]), f(in[1]), f(in[2])}, in[3]);
		} else if (name == "mip_from_pos") {
			out[0] = mip_from_pos(vec3{f(in[0]), f(in[1]), f(in[2])}, in[3]);
		} else if (name == "mip_from_dt") {
			out[0] = mip_from_dt(f(in[0]), vec3{f(in[1]), f(in[2]), f(in[3])}, in[4]);
		} else if (name == "distance_to_next_voxel") {
			const vec3 dir{f(in[3]), f(in[4]), f(in[5])};
			out[0] = u(distance_to_next_voxel(vec3{f(in[0]), f(in[1]), f(in[2])}, dir, vec3(1.0f) / dir, f(in[6])));
		} else if (name == "ray_intersect") {
			BoundingBox box{vec3{f(in[0]), f(in[1]), f(in[2])}, vec3{f(in[3]), f(in[4]), f(in[5])}};
			const vec2 t = box.ray_intersect(vec3{f(in[6]), f(in[7]), f(in[8])}, vec3{f(in[9]), f(in[10]), f(in[11])});
			out[0] = u(t.x); out[1] = u(t.y);
		} else if (name == "opencv_undistort") {
			const float params[4] = {f(in[0]), f(in[1]), f(in[2]), f(in[3])};
			float x = f(in[4]), y = f(in[5]);
			iterative_opencv_lens_undistortion(params, &x, &y);
			out[0] = u(x); out[1] = u(y);
		} else if (name == "uv_to_ray") { // the training / render camera: uv_to_ray (common_device.cuh:393-461) as generate_training_samples_nerf calls it
			mat4x3 cam;
			for (int c = 0; c < 4; ++c) cam[c] = vec3{f(in[8 + 3 * c]), f(in[9 + 3 * c]), f(in[10 + 3 * c])};
			Lens lens;
			lens.mode = (ELensMode)in[20];
			for (int k = 0; k < 4; ++k) lens.params[k] = f(in[21 + k]);
			const Ray ray = uv_to_ray(0, vec2{f(in[0]), f(in[1])}, ivec2{(int)in[2], (int)in[3]}, vec2{f(in[4]), f(in[5])}, cam, vec2{f(in[6]), f(in[7])},
				vec3(0.0f), 0.0f, 1.0f, 0.0f, {}, {}, lens);
			for (int k = 0; k < 3; ++k) { out[k] = u(ray.o[k]); out[3 + k] = u(ray.d[k]); }
		} else if (name == "srgb_to_linear") {
			out[0] = u(srgb_to_linear(f(in[0])));
		} else if (name == "linear_to_srgb") {
			out[0] = u(linear_to_srgb(f(in[0])));
		} else if (name == "loss") {
			const LossAndGradient lg = loss_and_gradient(vec3{f(in[0]), f(in[1]), f(in[2])}, vec3{f(in[3]), f(in[4]), f(in[5])}, (ELossType)in[6]);
			for (int k = 0; k < 3; ++k) { out[k] = u(lg.loss[k]); out[3 + k] = u(lg.gradient[k]); }
		} else if (name == "network_to_rgb") {
			out[0] = u(network_to_rgb(f(in[0]), (ENerfActivation)in[1])); out[1] = u(network_to_rgb_derivative(f(in[0]), (ENerfActivation)in[1]));
		} else if (name == "network_to_density") {
			out[0] = u(network_to_density(f(in[0]), (ENerfActivation)in[1])); out[1] = u(network_to_density_derivative(f(in[0]), (ENerfActivation)in[1]));
		} else if (name == "image_idx") {
			out[0] = image_idx(in[0], in[1], 0, in[2]);
		} else if (name == "ld_random_val") {
			out[0] = u(ld_random_val(in[0], in[1], in[2]));
		} else if (name == "warp_dt") {
			out[0] = u(warp_dt(f(in[0]))); out[1] = u(unwarp_dt(f(in[0])));
		} else if (name == "stepping") { // libm log / exp here; the oracle uses fixed polynomials: compared with a tolerance
			const float t = f(in[0]), cone = f(in[1]), steps = f(in[2]);
			out[0] = u(to_stepping_space(t, cone)); out[1] = u(from_stepping_space(steps, cone));
			out[2] = u(calc_dt(t, cone)); out[3] = u(advance_n_steps(t, cone, steps));
		} else {
			return -1;
		}
	}
	return 0;
}

// The reference's render march of one ray, composed from its own functions exactly as its kernels
// compose them: init_rays_with_payload_kernel_nerf (src/testbed_nerf.cu:1475-1488: t = max(tmin, 0) + 1e-6,
// dead if outside), advance_pos_nerf (:333-362: jittered start, skip to the first occupied voxel),
// generate_next_nerf_network_inputs (:452-465: skip, emit at t with dt = calc_dt, t += dt), run
// until the ray leaves the box. Same outputs as oracle_march_rays.
extern "C" void ref_march_rays(uint32_t n, const float* rays, const uint32_t* pix, uint32_t spp_index, const float* aabb, float cone_angle_constant,
	uint32_t max_mip, const uint8_t* bitfield, uint32_t max_steps, uint32_t* counts, float* t_first, float* t_last)
{
	const BoundingBox render_aabb{vec3{aabb[0], aabb[1], aabb[2]}, vec3{aabb[3], aabb[4], aabb[5]}};
	for (uint32_t r = 0; r < n; ++r) {
		Ray ray{vec3{rays[6 * r + 0], rays[6 * r + 1], rays[6 * r + 2]}, vec3{rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5]}};
		counts[r] = 0; t_first[r] = 0.0f; t_last[r] = 0.0f;
		float t = fmaxf(render_aabb.ray_intersect(ray.o, ray.d).x, 0.0f) + 1e-6f;
		if (!render_aabb.contains(ray(t))) continue;
		const vec3 idir = vec3(1.0f) / ray.d;
		const float cone_angle = calc_cone_angle(1.0f, vec2(1.0f), cone_angle_constant);
		t = advance_n_steps(t, cone_angle, ld_random_val(spp_index, pix[r] * 786433));
		uint32_t j = 0;
		for (; j < max_steps; ++j) {
			t = if_unoccupied_advance_to_next_occupied_voxel(t, cone_angle, ray, idir, bitfield, 0, max_mip, render_aabb);
			if (t >= MAX_DEPTH()) break;
			const float dt = calc_dt(t, cone_angle);
			if (j == 0) t_first[r] = t;
			t_last[r] = t;
			t += dt;
		}
		counts[r] = j;
	}
}
