// NeRF training kernels: ray sampling through the occupancy bitfield, and compositing + loss +
// gradient + compaction. Compiled with -fmad=false (see nerf_device.cuh).
//
// Slot allocation: the reference lets every ray claim its sample range with a global atomicAdd
// (src/testbed_nerf.cu:800,807,997). Here a block scans its rays' counts (warp shuffles + one
// shared-memory step) and claims one contiguous range per block, so a 2^18-ray batch costs 2^11
// atomics instead of 2^19 and the rays of a block stay contiguous in the sample buffers.
//
// Emission: the reference marches every ray twice in the same thread (count, then emit). Here the
// counting pass stores t of every sample it finds (32-sample chunks claimed from a pool, one atomic
// per chunk), and a second kernel writes the coordinates with one LANE per sample - positions, dt
// and warps are pure functions of (ray, t), evaluated with the serial loop's arithmetic, so the
// rows are bit-identical to a second march.
//
// Lane packing: a step has a few thousand rays and the march is a serial, divergent loop, so a warp
// full of rays runs for as long as the union of its 32 loops while most of the GPU idles. The
// counting pass therefore spreads rays over warps (down to one ray per warp) until the grid fills
// the machine.
#include "nerf.h"
#include "block_scan.cuh"
#include "module.h"

#include <algorithm>
#include <cstdlib>

namespace ngpb {

constexpr uint32_t kChunk = 32;                       // samples per pool chunk
constexpr uint32_t kChunksPerRay = kNerfSteps / kChunk; // 32
constexpr uint32_t kNoChunk = 0xFFFFFFFFu;

// Scratch of one generate call: chunk pool of sample distances + per-ray chunk table.
struct MarchScratch {
	uint32_t* pool_counter;
	uint32_t n_chunks;
	float* pool;        // [n_chunks][kChunk]
	uint32_t* chunks;   // [n_rays_local][kChunksPerRay]
};

#ifdef NGPB_MARCH_STATS
// Diagnostic build only (scripts/march_stats.py): histogram of loop iterations per ray, 16 per bin.
__device__ uint32_t g_march_hist[256];
#endif

// The counting march of one ray (src/testbed_nerf.cu:760-790): distances of its samples into the
// chunk pool, number of samples returned.
//
// The reference's iteration is one serial chain: position -> log -> exp (dt) -> mip -> occupancy bit
// -> log -> exp (skip target). Here the mip level is predicted from dt ~ t * cone_angle (exact up to
// rounding), so the occupancy load and the skip target's logarithm start beside log(t) and both
// exponentials beside each other: two chains per iteration instead of four. The prediction is
// checked against the real dt and the speculated values are redone when it was wrong, so every
// value is produced by the reference's operations on the reference's operands.
template <bool kCone>
NGPB_D uint32_t march_and_record(const NerfTrainConfig& cfg, const uint8_t* __restrict__ bitfield, const Stepping& stepping, const MarchScratch& scratch,
	const uint32_t local, const V3 o, const V3 dir_n, const V3 idir, const float startt, bool& pool_ok)
{
	uint32_t j = 0, chunk = kNoChunk;
#ifdef NGPB_MARCH_STATS
	uint32_t iters = 0;
#endif
	float t = startt;
	V3 pos;
	while (aabb_contains(cfg.aabb, pos = o + t * dir_n) && j < kNerfSteps) {
#ifdef NGPB_MARCH_STATS
		++iters;
#endif
		const float dt_guess = kCone ? fminf(fmaxf(t * stepping.cone_angle, kMinConeStepsize), kMaxConeStepsize) : kMinConeStepsize;
		const uint32_t mip_guess = mip_from_dt(dt_guess, pos, cfg.max_mip);
		bool occupied = density_grid_occupied_at(pos, bitfield, mip_guess);
		const float n_target = to_stepping_space_sel<kCone>(t + distance_to_next_voxel(pos, dir_n, idir, pow2i(7 - (int)mip_guess), pow2i((int)mip_guess - 7)), stepping);
		const float n = to_stepping_space_sel<kCone>(t, stepping);
		const float dt = from_stepping_space_sel<kCone>(n + 1.0f, stepping) - t; // calc_dt
		float t_skip = from_stepping_space_sel<kCone>(n + ceilf(fmaxf(n_target - n, 0.5f)), stepping); // advance_to_next_voxel
		const uint32_t mip = mip_from_dt(dt, pos, cfg.max_mip);
		if (mip != mip_guess) {
			occupied = density_grid_occupied_at(pos, bitfield, mip);
			t_skip = advance_to_next_voxel_n(t, n, stepping, pos, dir_n, idir, mip);
		}
		if (occupied) {
			if (j % kChunk == 0) {
				chunk = atomicAdd(scratch.pool_counter, 1u);
				if (chunk >= scratch.n_chunks) { chunk = kNoChunk; pool_ok = false; }
				scratch.chunks[(size_t)local * kChunksPerRay + j / kChunk] = chunk;
			}
			if (chunk != kNoChunk) scratch.pool[(size_t)chunk * kChunk + j % kChunk] = t;
			++j;
			t += dt;
		} else {
			t = t_skip;
		}
	}
#ifdef NGPB_MARCH_STATS
	atomicAdd(&g_march_hist[min(iters / 16u, 255u)], 1u);
#endif
	return j;
}

__global__ void __launch_bounds__(kRayBlock) generate_training_samples_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf,
	const __grid_constant__ Stepping stepping, Pcg32 rng, const __grid_constant__ MarchScratch scratch, const uint32_t rays_per_warp)
{
	__shared__ uint32_t warp_sums[kRayBlock / 32 + 1];
	__shared__ uint32_t block_base, block_ray_base;

	const uint32_t lane = threadIdx.x % 32;
	const uint32_t local = ((blockIdx.x * blockDim.x + threadIdx.x) / 32) * rays_per_warp + lane;
	bool valid = lane < rays_per_warp && local < local_ray_count(cfg);
	const uint32_t i = cfg.ray_begin + local; // global ray index

	Ray3 ray{};
	V3 dir_n{}, idir{};
	float startt = 0.0f;
	uint32_t numsteps = 0;
	bool pool_ok = true;

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
			numsteps = stepping.cone_angle <= 1e-5f
				? march_and_record<false>(cfg, buf.density_bitfield, stepping, scratch, local, ray.o, dir_n, idir, startt, pool_ok)
				: march_and_record<true>(cfg, buf.density_bitfield, stepping, scratch, local, ray.o, dir_n, idir, startt, pool_ok);
			if (numsteps == 0) valid = false;
		}
	}

	uint32_t block_total;
	const uint32_t prefix = block_exclusive_scan(valid ? numsteps : 0u, block_total, warp_sums);
	if (threadIdx.x == 0) block_base = block_total ? atomicAdd(&buf.counters->numsteps, block_total) : 0u;
	__syncthreads();
	const uint32_t base = block_base + prefix;
	// a ray that does not fit the sample buffer (or found the chunk pool exhausted, which only happens
	// when the batch overflows the buffer anyway) is dropped whole (:804-806)
	if (valid && (base + numsteps > cfg.max_samples || !pool_ok)) valid = false;

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

// Warp per kept ray, lane per sample: coordinates from the stored distances.
__global__ void __launch_bounds__(kRayBlock) emit_training_samples_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf,
	const __grid_constant__ Stepping stepping, const __grid_constant__ MarchScratch scratch)
{
	const uint32_t ray_idx = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
	if (ray_idx >= buf.counters->rays) return;
	const uint32_t numsteps = buf.numsteps[ray_idx * 2 + 0], base = buf.numsteps[ray_idx * 2 + 1];
	const uint32_t local = buf.ray_indices[ray_idx] - cfg.ray_begin;
	const float* ro = buf.rays_unnormalized + (size_t)ray_idx * 6;
	const V3 o = {ro[0], ro[1], ro[2]};
	const V3 dir_n = normalize(v3(ro[3], ro[4], ro[5]));
	const V3 warped_dir = warp_direction(dir_n);
	float* out = buf.coords + (size_t)base * kCoordFloats;
	for (uint32_t j = lane; j < numsteps; j += 32) {
		const uint32_t chunk = scratch.chunks[(size_t)local * kChunksPerRay + j / kChunk];
		const float t = scratch.pool[(size_t)chunk * kChunk + j % kChunk];
		const V3 pos = o + t * dir_n;
		const float dt = calc_dt(t, stepping);
		const V3 wp = warp_position(pos, cfg.aabb);
		float* c = out + (size_t)j * kCoordFloats;
		c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt);
		c[4] = warped_dir.x; c[5] = warped_dir.y; c[6] = warped_dir.z;
	}
}

void nerf_generate_training_samples(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng) {
	if (local_ray_count(cfg) == 0) return;
	// The testbed runs these two kernels on a second stream under the fused MLP kernels of the
	// previous step. An SM keeps one shared-memory / L1 split while any block is resident, and the MLP
	// kernels need nearly all of it as shared memory: ask for the same split here, or the MLP CTAs
	// wait until the last march block has left the SM.
	static const bool carveout_set = [] {
		if (getenv("NGPB_MARCH_SPLIT_DEFAULT") && atoi(getenv("NGPB_MARCH_SPLIT_DEFAULT"))) return false;
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(generate_training_samples_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(emit_training_samples_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		return true;
	}();
	(void)carveout_set;
	const Stepping stepping = make_stepping(cfg.cone_angle_constant);
	const uint32_t n_local = local_ray_count(cfg);
	// rays per warp: few enough that the serial loops of a warp's rays do not pile up, many enough
	// that the warp instructions (one per iteration whatever the lane count) do not saturate issue
	static const uint32_t max_warps = getenv("NGPB_MARCH_WARPS") ? (uint32_t)atoi(getenv("NGPB_MARCH_WARPS")) : 512u;
	uint32_t rays_per_warp = 1;
	while (rays_per_warp < 32 && n_local / rays_per_warp > max_warps) rays_per_warp *= 2;
	MarchScratch scratch{};
	// every ray can hold its kNerfSteps samples: the pool never decides which rays are kept (at most
	// 2^18 rays x 4 KiB = 1 GiB of address space, touched only where samples exist)
	scratch.n_chunks = n_local * kChunksPerRay;
	const size_t pool_bytes = (size_t)scratch.n_chunks * kChunk * sizeof(float), table_bytes = (size_t)n_local * kChunksPerRay * sizeof(uint32_t);
	DeviceBuffer mem(256 + pool_bytes + table_bytes, stream); // stream-ordered, returned to the pool at scope exit
	scratch.pool_counter = mem.as<uint32_t>();
	scratch.pool = reinterpret_cast<float*>(mem.as<char>() + 256);
	scratch.chunks = reinterpret_cast<uint32_t*>(mem.as<char>() + 256 + pool_bytes);
	NGPB_CUDA_CHECK(cudaMemsetAsync(scratch.pool_counter, 0, sizeof(uint32_t), stream));
	const uint32_t n_warps = div_round_up(n_local, rays_per_warp);
	generate_training_samples_kernel<<<div_round_up(n_warps, kRayBlock / 32), kRayBlock, 0, stream>>>(cfg, buf, stepping, rng, scratch, rays_per_warp);
	NGPB_CUDA_CHECK(cudaGetLastError());
	emit_training_samples_kernel<<<div_round_up(n_local, kRayBlock / 32), kRayBlock, 0, stream>>>(cfg, buf, stepping, scratch);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

#ifdef NGPB_MARCH_STATS
extern "C" int ngpb_debug_march_hist(uint32_t* out256, int reset) {
	if (cudaMemcpyFromSymbol(out256, g_march_hist, sizeof(g_march_hist)) != cudaSuccess) return 1;
	if (reset) { static const uint32_t zero[256] = {}; cudaMemcpyToSymbol(g_march_hist, zero, sizeof(zero)); }
	return 0;
}
#endif

} // namespace ngpb
