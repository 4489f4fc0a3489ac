// NeRF training kernels: ray sampling through the occupancy bitfield, and compositing + loss +
// gradient + compaction. Compiled with -fmad=false (see nerf_device.cuh).
//
// Slot allocation: the reference lets every ray claim its sample range with a global atomicAdd
// (src/testbed_nerf.cu:800,807,997). Here a block scans its rays' counts (warp shuffles + one
// shared-memory step) and claims one contiguous range per block, so a 2^18-ray batch costs 2^11
// atomics instead of 2^19 and the rays of a block stay contiguous in the sample buffers.
//
// Emission: the reference marches every ray twice in the same thread (count, then emit). Here the
// counting pass also records t at every 32nd sample, and a second kernel re-marches each 32-sample
// segment from its checkpoint in its own thread, so a 600-sample ray is emitted by 19 threads
// instead of one. The arithmetic from a checkpoint on is the serial loop's, bit for bit.
#include "nerf.h"
#include "block_scan.cuh"
#include "module.h"

namespace ngpb {

constexpr uint32_t kEmitSegment = 32; // samples per emission thread; kNerfSteps / 32 = 32 checkpoints per ray

__global__ void __launch_bounds__(kRayBlock) generate_training_samples_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf,
	const __grid_constant__ Stepping stepping, Pcg32 rng, float* __restrict__ checkpoints)
{
	__shared__ uint32_t warp_sums[kRayBlock / 32 + 1];
	__shared__ uint32_t block_base, block_ray_base;

	const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
	bool valid = local < local_ray_count(cfg);
	const uint32_t i = cfg.ray_begin + local; // global ray index

	Ray3 ray{};
	V3 dir_n{}, idir{};
	float startt = 0.0f;
	uint32_t numsteps = 0;

	if (valid) {
		const uint32_t img = image_idx(i, cfg.n_rays, cfg.n_training_images);
		const ImageMeta& m = buf.metadata[img];
		rng.advance((int64_t)i * kMaxRandomSamplesPerRay);
		float u, v;
		random_image_pos_training(rng, m.res_x, m.res_y, cfg.snap_to_pixel_centers != 0, u, v);
		// negative values mark masked-away pixels
		if (read_rgba_uv(u, v, m.res_x, m.res_y, m.pixels, m.image_data_type).r < 0.0f) valid = false;
		rng.next_float(); // motionblur_time
		if (valid) {
			ray = uv_to_ray(u, v, m.res_x, m.res_y, m.focal_x, m.focal_y, m.xform, m.principal_x, m.principal_y, m.lens_mode, m.lens_params, 0.0f);
			if (ray.d.x == 0.0f && ray.d.y == 0.0f && ray.d.z == 0.0f) ray = {m.xform.c[3], m.xform.c[2]};
			dir_n = normalize(ray.d);
			float tmin, tmax;
			aabb_ray_intersect(cfg.aabb, ray.o, dir_n, tmin, tmax);
			tmin = fmaxf(tmin, 0.0f);
			startt = advance_n_steps(tmin, stepping, rng.next_float());
			idir = v3(1.0f) / dir_n;

			// first pass: count the samples
			uint32_t j = 0;
			float t = startt;
			V3 pos;
			while (aabb_contains(cfg.aabb, pos = ray.o + t * dir_n) && j < kNerfSteps) {
				const float dt = calc_dt(t, stepping);
				const uint32_t mip = mip_from_dt(dt, pos, cfg.max_mip);
				if (density_grid_occupied_at(pos, buf.density_bitfield, mip)) {
					if (j % kEmitSegment == 0) checkpoints[(size_t)local * kEmitSegment + j / kEmitSegment] = t;
					++j;
					t += dt;
				} else {
					t = advance_to_next_voxel(t, stepping, pos, dir_n, idir, mip);
				}
			}
			numsteps = j;
			if (j == 0) valid = false;
		}
	}

	uint32_t block_total;
	const uint32_t prefix = block_exclusive_scan(valid ? numsteps : 0u, block_total, warp_sums);
	if (threadIdx.x == 0) block_base = block_total ? atomicAdd(&buf.counters->numsteps, block_total) : 0u;
	__syncthreads();
	const uint32_t base = block_base + prefix;
	if (valid && base + numsteps > cfg.max_samples) valid = false;

	uint32_t kept_total;
	const uint32_t ray_prefix = block_exclusive_scan(valid ? 1u : 0u, kept_total, warp_sums);
	if (threadIdx.x == 0) block_ray_base = kept_total ? atomicAdd(&buf.counters->rays, kept_total) : 0u;
	__syncthreads();
	if (!valid) return;

	const uint32_t ray_idx = block_ray_base + ray_prefix;
	buf.ray_indices[ray_idx] = i;
	float* ro = buf.rays_unnormalized + (size_t)ray_idx * 6;
	ro[0] = ray.o.x; ro[1] = ray.o.y; ro[2] = ray.o.z; ro[3] = ray.d.x; ro[4] = ray.d.y; ro[5] = ray.d.z;
	buf.numsteps[ray_idx * 2 + 0] = numsteps;
	buf.numsteps[ray_idx * 2 + 1] = base;
}

// Thread per (kept ray, 32-sample segment): samples [32 s, min(32 s + 32, numsteps)) from checkpoint s.
__global__ void __launch_bounds__(kRayBlock) emit_training_samples_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf,
	const __grid_constant__ Stepping stepping, const float* __restrict__ checkpoints)
{
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t ray_idx = tid / kEmitSegment, seg = tid % kEmitSegment;
	if (ray_idx >= buf.counters->rays) return;
	const uint32_t numsteps = buf.numsteps[ray_idx * 2 + 0], base = buf.numsteps[ray_idx * 2 + 1];
	uint32_t j = seg * kEmitSegment;
	if (j >= numsteps) return;
	const uint32_t end = min(numsteps, j + kEmitSegment);
	const uint32_t local = buf.ray_indices[ray_idx] - cfg.ray_begin;
	const float* ro = buf.rays_unnormalized + (size_t)ray_idx * 6;
	const V3 o = {ro[0], ro[1], ro[2]};
	const V3 dir_n = normalize(v3(ro[3], ro[4], ro[5]));
	const V3 idir = v3(1.0f) / dir_n;
	const V3 warped_dir = warp_direction(dir_n);
	float* out = buf.coords + (size_t)base * kCoordFloats;
	float t = checkpoints[(size_t)local * kEmitSegment + seg];
	V3 pos;
	while (aabb_contains(cfg.aabb, pos = o + t * dir_n) && j < end) {
		const float dt = calc_dt(t, stepping);
		const uint32_t mip = mip_from_dt(dt, pos, cfg.max_mip);
		if (density_grid_occupied_at(pos, buf.density_bitfield, mip)) {
			const V3 wp = warp_position(pos, cfg.aabb);
			float* c = out + (size_t)j * kCoordFloats;
			c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt);
			c[4] = warped_dir.x; c[5] = warped_dir.y; c[6] = warped_dir.z;
			++j;
			t += dt;
		} else {
			t = advance_to_next_voxel(t, stepping, pos, dir_n, idir, mip);
		}
	}
}

void nerf_generate_training_samples(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng) {
	if (local_ray_count(cfg) == 0) return;
	const Stepping stepping = make_stepping(cfg.cone_angle_constant);
	const uint32_t n_local = local_ray_count(cfg);
	DeviceBuffer checkpoints((size_t)n_local * kEmitSegment * sizeof(float), stream); // stream-ordered, returned to the pool at scope exit
	generate_training_samples_kernel<<<div_round_up(n_local, kRayBlock), kRayBlock, 0, stream>>>(cfg, buf, stepping, rng, checkpoints.as<float>());
	NGPB_CUDA_CHECK(cudaGetLastError());
	emit_training_samples_kernel<<<(uint32_t)div_round_up64((uint64_t)n_local * kEmitSegment, kRayBlock), kRayBlock, 0, stream>>>(cfg, buf, stepping, checkpoints.as<float>());
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
