// NeRF training kernels: ray sampling through the occupancy bitfield (generate_training_samples_nerf,
// src/testbed_nerf.cu:679-838). Compiled with -fmad=false (see nerf_device.cuh).
//
// The reference gives a ray to a thread, marches it twice (count, then emit) and lets every ray claim its sample
// range with a global atomicAdd (:800,807). Here a WARP marches a ray once (see "the lattice march" below), keeps the
// lattice indices of its samples in shared memory (2 bytes each), claims the range with one atomic and writes the
// rows with a lane per sample. A step has a few thousand rays: a warp each fills the machine.
#include "nerf.h"
#include "block_scan.cuh"
#include "module.h"

#include <algorithm>
#include <cstdlib>

namespace ngpb {

constexpr uint32_t kMarchWarps = kRayBlock / 32;   // rays per block
constexpr float kMaxLatticeSkip = 1048576.0f;

// ---- the lattice march ---------------------------------------------------------------------------
// The reference's loop (src/testbed_nerf.cu:760-790) is `t += calc_dt(t)` on occupied samples and
// advance_to_next_voxel(t) on empty ones. Both are moves on a lattice in stepping space: calc_dt(t) =
// from(to(t) + 1) - t and advance_to_next_voxel(t) = from(to(t) + ceil(max(to(t_next_voxel) - to(t), 0.5)))
// (nerf_device.cuh:398-437), i.e. every position a ray ever visits is from(n0 + k) with n0 = to(t_start) and an
// integer k - up to the rounding of to(from(.)). Written that way,
//     candidate k:  n_k = n0 + k,  t_k = from(n_k),  dt_k = from(n_k + 1) - t_k,  pos_k = o + t_k d
//                   occupied(pos_k, mip(dt_k, pos_k)) ? sample, k += 1
//                                                     : k += ceil(max(to(t_k + distance_to_next_voxel) - n_k, 0.5))
// a candidate no longer depends on its predecessor's floating-point value, only on its index. So a WARP marches one
// ray: 32 lanes evaluate candidates k0 .. k0 + 31 side by side (two exponentials, a logarithm and one bitfield byte
// each), the set the serial walk visits is the orbit of lane 0 under c -> c + step_c, found by pointer doubling
// with five shuffle rounds, and the occupied visited lanes are the samples (ballot + popc = compaction). The serial
// chain per ray drops from ~270 dependent iterations to ~10-40 windows. The CPU checker restates exactly this rule
// (its lattice_march); against the reference's own functions composed the serial way the sample
// counts agree on the same ~99 % of the rays as before (the difference is the rounding of to(from(.)), of the
// size of the polynomial log/exp already in use; tests/test_ref_pin_cpu.py reports the fraction).
struct LatticeCandidate {
	float t, dt;
	V3 pos;
	bool inside, occupied;
	uint32_t step;
};
template <bool kCone>
NGPB_D LatticeCandidate lattice_candidate(const Aabb& aabb, uint32_t max_mip, const uint8_t* __restrict__ bitfield, const Stepping& stepping,
	const V3 o, const V3 dir_n, const V3 idir, const float n0, const uint32_t k)
{
	LatticeCandidate c;
	const float nk = n0 + (float)k;
	c.t = from_stepping_space_sel<kCone>(nk, stepping);
	c.pos = o + c.t * dir_n;
	c.inside = aabb_contains(aabb, c.pos);
	c.dt = from_stepping_space_sel<kCone>(nk + 1.0f, stepping) - c.t;
	const uint32_t mip = mip_from_dt(c.dt, c.pos, max_mip);
	c.occupied = c.inside && density_grid_occupied_at(c.pos, bitfield, mip);
	c.step = 1;
	if (!c.occupied) {
		const float n_target = to_stepping_space_sel<kCone>(c.t + distance_to_next_voxel(c.pos, dir_n, idir, pow2i(7 - (int)mip), pow2i((int)mip - 7)), stepping);
		c.step = (uint32_t)fminf(ceilf(fmaxf(n_target - nk, 0.5f)), kMaxLatticeSkip);
	}
	return c;
}

// One warp, one ray: lattice indices of the ray's samples into ks[0 .. count), count returned (warp-uniform).
template <bool kCone>
NGPB_D uint32_t lattice_march(const NerfTrainConfig& cfg, const uint8_t* __restrict__ bitfield, const Stepping& stepping,
	const V3 o, const V3 dir_n, const V3 idir, const float n0, uint16_t* __restrict__ ks)
{
	const uint32_t lane = threadIdx.x % 32;
	const uint32_t lt_mask = (1u << lane) - 1u;
	uint32_t k_base = 0, j = 0;
	for (;;) {
		const uint32_t k = k_base + lane;
		const LatticeCandidate c = lattice_candidate<kCone>(cfg.aabb, cfg.max_mip, bitfield, stepping, o, dir_n, idir, n0, k);
		// lanes the serial walk visits = orbit of lane 0 under lane -> lane + step
		uint32_t jump = min(lane + c.step, 32u), reach = 1u << lane;
#pragma unroll
		for (uint32_t r = 0; r < 5; ++r) {
			const uint32_t reach2 = __shfl_sync(0xFFFFFFFFu, reach, jump & 31u), jump2 = __shfl_sync(0xFFFFFFFFu, jump, jump & 31u);
			if (jump < 32u) { reach |= reach2; jump = jump2; }
		}
		const uint32_t visited = __shfl_sync(0xFFFFFFFFu, reach, 0);
		const uint32_t outside = __ballot_sync(0xFFFFFFFFu, !c.inside);
		const uint32_t before_exit = outside ? ((1u << (__ffs(outside) - 1)) - 1u) : 0xFFFFFFFFu;   // candidates in front of the first one outside the box
		uint32_t emit = visited & before_exit & __ballot_sync(0xFFFFFFFFu, c.occupied);
		bool done = (visited & ~before_exit) != 0 || k_base > 65535u - 64u;   // the walk left the box (or the 16-bit index space: never on a real ray)
		const uint32_t room = kNerfSteps - j;
		if ((uint32_t)__popc(emit) >= room) {
			while ((uint32_t)__popc(emit) > room) emit &= ~(0x80000000u >> __clz(emit));
			done = true; // j < NERF_STEPS fails next
		}
		if ((emit >> lane) & 1u) ks[j + __popc(emit & lt_mask)] = (uint16_t)k;
		j += __popc(emit);
		if (done) break;
		const uint32_t last = 31u - (uint32_t)__clz(visited);
		k_base += last + __shfl_sync(0xFFFFFFFFu, c.step, last);
	}
	return j;
}

// Warp per ray. All lanes set the ray up redundantly (same values, no divergence), march it together, claim the
// ray's sample range with one atomic, then write the coordinate rows with a lane per sample.
__global__ void __launch_bounds__(kRayBlock) generate_training_samples_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf,
	const __grid_constant__ Stepping stepping, Pcg32 rng)
{
	__shared__ uint16_t ks_all[kMarchWarps][kNerfSteps];

	const uint32_t lane = threadIdx.x % 32, warp = threadIdx.x / 32;
	const uint32_t local = blockIdx.x * kMarchWarps + warp;
	if (local >= local_ray_count(cfg)) return;
	const uint32_t i = cfg.ray_begin + local; // global ray index
	uint16_t* ks = ks_all[warp];

	const uint32_t img = image_idx(i, cfg.n_rays, cfg.n_training_images);
	const ImageMeta& m = buf.metadata[img];
	rng.advance((int64_t)i * kMaxRandomSamplesPerRay);
	float u, v;
	random_image_pos_training(rng, m.res_x, m.res_y, cfg.snap_to_pixel_centers != 0, u, v);
	// negative values mark masked-away pixels
	if (read_rgba_uv(u, v, m.res_x, m.res_y, m.pixels, m.image_data_type).r < 0.0f) return;
	rng.next_float(); // motionblur_time
	Ray3 ray = uv_to_ray(u, v, m.res_x, m.res_y, m.focal_x, m.focal_y, m.xform, m.principal_x, m.principal_y, m.lens_mode, m.lens_params, 0.0f);
	if (ray.d.x == 0.0f && ray.d.y == 0.0f && ray.d.z == 0.0f) ray = {m.xform.c[3], m.xform.c[2]};
	const V3 dir_n = normalize(ray.d);
	float tmin, tmax;
	aabb_ray_intersect(cfg.aabb, ray.o, dir_n, tmin, tmax);
	tmin = fmaxf(tmin, 0.0f);
	const float startt = advance_n_steps(tmin, stepping, rng.next_float());
	const V3 idir = v3(1.0f) / dir_n;
	const float n0 = to_stepping_space(startt, stepping);
	const bool cone = stepping.cone_angle > 1e-5f;

	const uint32_t numsteps = cone
		? lattice_march<true>(cfg, buf.density_bitfield, stepping, ray.o, dir_n, idir, n0, ks)
		: lattice_march<false>(cfg, buf.density_bitfield, stepping, ray.o, dir_n, idir, n0, ks);
	if (numsteps == 0) return;

	uint32_t base = 0, ray_idx = 0xFFFFFFFFu;
	if (lane == 0) {
		base = atomicAdd(&buf.counters->numsteps, numsteps);
		// a ray that does not fit the sample buffer is dropped whole (:804-806)
		if (base + numsteps <= cfg.max_samples) ray_idx = atomicAdd(&buf.counters->rays, 1u);
	}
	base = __shfl_sync(0xFFFFFFFFu, base, 0);
	ray_idx = __shfl_sync(0xFFFFFFFFu, ray_idx, 0);
	if (ray_idx == 0xFFFFFFFFu) return;

	if (lane == 0) {
		buf.ray_indices[ray_idx] = i;
		buf.numsteps[ray_idx * 2 + 0] = numsteps;
		buf.numsteps[ray_idx * 2 + 1] = base;
	}
	if (lane < 6) buf.rays_unnormalized[(size_t)ray_idx * 6 + lane] = lane == 0 ? ray.o.x : lane == 1 ? ray.o.y : lane == 2 ? ray.o.z : lane == 3 ? ray.d.x : lane == 4 ? ray.d.y : ray.d.z;
	__syncwarp();

	const V3 warped_dir = warp_direction(dir_n);
	float* out = buf.coords + (size_t)base * kCoordFloats;
	for (uint32_t jj = lane; jj < numsteps; jj += 32) {
		const float nk = n0 + (float)ks[jj];
		const float t = cone ? from_stepping_space_sel<true>(nk, stepping) : from_stepping_space_sel<false>(nk, stepping);
		const float dt = (cone ? from_stepping_space_sel<true>(nk + 1.0f, stepping) : from_stepping_space_sel<false>(nk + 1.0f, stepping)) - t;
		const V3 wp = warp_position(ray.o + t * dir_n, cfg.aabb);
		float* c = out + (size_t)jj * kCoordFloats;
		c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt);
		c[4] = warped_dir.x; c[5] = warped_dir.y; c[6] = warped_dir.z;
	}
}

void nerf_generate_training_samples(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng) {
	if (local_ray_count(cfg) == 0) return;
	const Stepping stepping = make_stepping(cfg.cone_angle_constant);
	generate_training_samples_kernel<<<div_round_up(local_ray_count(cfg), kMarchWarps), kRayBlock, 0, stream>>>(cfg, buf, stepping, rng);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
