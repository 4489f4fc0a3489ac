// NeRF rendering: Testbed::render_nerf / NerfTracer (src/testbed_nerf.cu:333-677, 1309-1489,
// 1556-1987) as a wavefront that never returns to the host inside a window.
//
// The reference marches all rays in lock-step windows of 1..8 samples: compact_kernel_nerf
// (one global atomic per ray) -> host reads the live count (cudaStreamSynchronize every window,
// :1685-1686) -> generate_next_nerf_network_inputs -> network -> composite_kernel_nerf ->
// finally shade_kernel_nerf over the "hit" list. Here
//   * the live count, the window length and the sample count stay in device memory
//     (RenderCounters); the encoder and the fused MLP read the sample count there
//     (NerfNetworkModule::inference_counted), so a chunk of windows is queued back to back and the
//     host looks at the live count once per chunk;
//   * compositing, compaction and shading are ONE kernel: a ray that is still alive is appended to
//     the next list with a block scan (one atomic per 128 rays), a ray that finished is shaded
//     straight into the frame buffer - there is no "hit" list and no separate compaction pass;
//   * ray state is four float4 planes (origin|t, dir|pixel, rgba, depth|max_weight|steps), so every
//     access is a coalesced 16-byte load / store instead of the 40-byte NerfPayload AoS.
// Per-ray arithmetic (jittered start, occupied-voxel skipping, dt, compositing order, early
// termination at 1 - min_transmittance, depth at the heaviest sample) follows the reference, so a
// pixel does not depend on how rays are grouped into windows.
#include "testbed.h"
#include "block_scan.cuh"

#include <algorithm>

namespace ngpb {

struct RenderCounters {
	uint32_t n_alive[2];
	uint32_t n_samples;       // samples of the current window: n_alive * window_steps
	uint32_t window_steps;
	unsigned long long total_samples;
};

struct RenderParams {
	Aabb aabb;
	Mat4x3 camera;
	uint32_t width, height;
	float focal_x, focal_y, screen_cx, screen_cy;
	uint32_t spp_index;
	uint32_t pixel_begin, pixel_end, n_local;
	uint32_t max_mip;
	uint32_t skip_mip;        // coarsest cascade the empty-space skip may climb to (kSkipMaxMip; max_mip = the reference's rule)
	uint32_t shard_rank, shard_world; // tile interleave over ranks (kRenderTile pixels per tile)
	uint32_t capacity;        // rays per plane
	uint32_t target_queries, min_steps, max_steps;
	float min_transmittance;
	int rgb_activation, density_activation, linear_colors, snap_to_pixel_centers;
	int lens_mode;            // Testbed::m_nerf.render_lens (testbed.h), used when render_with_lens_distortion
	float lens_params[7];
};

constexpr uint32_t kRenderTile = 256;

NGPB_D uint32_t window_length(const RenderParams& p, uint32_t n_alive) {
	// clamp(target_n_queries / n_alive, MIN, MAX_STEPS_INBETWEEN_COMPACTION) (:1693-1694)
	const uint32_t s = p.target_queries / max(n_alive, 1u);
	return min(max(s, p.min_steps), p.max_steps);
}

// init_rays_with_payload_kernel_nerf (:1375-1489) + advance_pos_nerf_kernel (:333-379); live rays
// are appended to list 0.
__global__ void __launch_bounds__(kRayBlock) render_init_rays_kernel(const __grid_constant__ RenderParams p, const __grid_constant__ Stepping stepping,
	const uint8_t* __restrict__ bitfield, float4* __restrict__ rays, RenderCounters* __restrict__ counters, float4* __restrict__ frame, float* __restrict__ depth_out)
{
	__shared__ uint32_t warp_sums[kRayBlock / 32 + 1];
	__shared__ uint32_t block_base;
	const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t pix = p.pixel_begin + ((local / kRenderTile) * p.shard_world + p.shard_rank) * kRenderTile + local % kRenderTile;
	bool alive = local < p.n_local && pix < p.pixel_end;
	V3 o{}, d{};
	float t = 0.0f;
	if (alive) {
		frame[pix] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
		if (depth_out) depth_out[pix] = kMaxDepth;
		const uint32_t x = pix % p.width, y = pix / p.width;
		float ox, oy;
		ld_random_pixel_offset(p.snap_to_pixel_centers ? 0u : p.spp_index, ox, oy);
		const float u = ((float)x + ox) / (float)p.width, v = ((float)y + oy) / (float)p.height;
		const Ray3 ray = uv_to_ray(u, v, (int)p.width, (int)p.height, p.focal_x, p.focal_y, p.camera, p.screen_cx, p.screen_cy, p.lens_mode, p.lens_params, 0.0f);
		o = ray.o;
		d = normalize(ray.d);
		float tmin, tmax;
		aabb_ray_intersect(p.aabb, o, d, tmin, tmax);
		t = fmaxf(tmin, 0.0f) + 1e-6f;
		if (!aabb_contains(p.aabb, o + d * t)) alive = false;
		if (alive) {
			const V3 idir = v3(1.0f) / d;
			t = advance_n_steps(t, stepping, ld_random_val(p.spp_index, pix * 786433u));
			t = stepping.cone_angle <= 1e-5f ? advance_to_next_occupied_voxel_sel<false>(t, stepping, o, d, idir, bitfield, 0, p.max_mip, p.aabb, p.skip_mip)
			                                 : advance_to_next_occupied_voxel_sel<true>(t, stepping, o, d, idir, bitfield, 0, p.max_mip, p.aabb, p.skip_mip);
			if (t >= kMaxDepth) alive = false;
		}
	}
	uint32_t total;
	const uint32_t prefix = block_exclusive_scan(alive ? 1u : 0u, total, warp_sums);
	if (threadIdx.x == 0) block_base = total ? atomicAdd(&counters->n_alive[0], total) : 0u;
	__syncthreads();
	if (!alive) return;
	const uint32_t slot = block_base + prefix;
	rays[slot] = make_float4(o.x, o.y, o.z, t);
	rays[slot + p.capacity] = make_float4(d.x, d.y, d.z, __uint_as_float(pix));
	rays[slot + 2 * p.capacity] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
	rays[slot + 3 * p.capacity] = make_float4(0.0f, 0.0f, __uint_as_float(0u), 0.0f); // depth, max_weight, steps taken
}

// generate_next_nerf_network_inputs (:417-470). Sample j of ray i lands at i + j * n_alive.
template <bool kCone>
__global__ void __launch_bounds__(kRayBlock) render_march_kernel(const __grid_constant__ RenderParams p, const __grid_constant__ Stepping stepping,
	const uint8_t* __restrict__ bitfield, const uint32_t cur, float4* __restrict__ rays, uint32_t* __restrict__ ray_steps,
	RenderCounters* __restrict__ counters, float* __restrict__ coords)
{
	const uint32_t n_alive = counters->n_alive[cur];
	const uint32_t n_steps = window_length(p, n_alive);
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i == 0) {
		counters->n_samples = n_alive * n_steps;
		counters->window_steps = n_steps;
		counters->n_alive[cur ^ 1u] = 0;
		counters->total_samples += (unsigned long long)n_alive * n_steps;
	}
	if (i >= n_alive) return;
	const float4 r0 = rays[i], r1 = rays[i + p.capacity];
	const V3 o = {r0.x, r0.y, r0.z}, d = {r1.x, r1.y, r1.z};
	const V3 idir = v3(1.0f) / d;
	const V3 wd = warp_direction(d);
	float t = r0.w;
	uint32_t j = 0;
	// One flat loop: an iteration either skips one empty voxel or emits one sample. The reference
	// nests the skip loop inside the step loop (:436-447); in a warp of unrelated rays that shape
	// serialises every ray's skips at every step (ncu, round 1: 5.6 of 32 lanes active). Flat, a
	// lane that skips and a lane that emits share the iteration - including its exponential: both
	// continuations are from_stepping_space of (n + something), evaluated once on a per-lane
	// argument. Same sequence of t per ray, same operations on the same operands.
	while (j < n_steps) {
		const V3 pos = o + d * t;
		if (t >= kMaxDepth || !aabb_contains(p.aabb, pos)) break;
		uint32_t mip = min(mip_from_pos(pos), p.max_mip);
		const float n = to_stepping_space_sel<kCone>(t, stepping);
		const bool occupied = density_grid_occupied_at(pos, bitfield, mip);
		float n_next = n + 1.0f; // calc_dt
		if (!occupied) {
			// largest empty voxel around the sample (nerf_device.cuh:485-490), then advance_to_next_voxel
			// (measured and not kept: the up-to-seven coarser-mip loads issued together instead of this early-exit chain - the
			// chain usually ends at its first load, and the extra loads cost more than the latency they hide: march 1.88 -> 2.46 ms)
			while (mip < p.skip_mip && !density_grid_occupied_at(pos, bitfield, mip + 1)) ++mip;
			const float n_target = to_stepping_space_sel<kCone>(t + distance_to_next_voxel(pos, d, idir, pow2i(7 - (int)mip), pow2i((int)mip - 7)), stepping);
			n_next = n + ceilf(fmaxf(n_target - n, 0.5f));
		}
		const float t_next = from_stepping_space_sel<kCone>(n_next, stepping);
		if (occupied) {
			const float dt = t_next - t;
			const V3 wp = warp_position(pos, p.aabb);
			float* c = coords + (size_t)(i + j * n_alive) * kCoordFloats;
			c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt);
			c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
			t += dt;
			++j;
		} else {
			t = t_next;
		}
	}
	ray_steps[i] = j;
	if (j == n_steps) rays[i].w = t;
}

// composite_kernel_nerf (:471-677, Shade mode) + compact_kernel_nerf (:1351-1373) +
// shade_kernel_nerf (:1309-1349).
__global__ void __launch_bounds__(kRayBlock) render_composite_kernel(const __grid_constant__ RenderParams p, const uint32_t cur,
	const float4* __restrict__ rays_cur, float4* __restrict__ rays_next, const uint32_t* __restrict__ ray_steps, RenderCounters* __restrict__ counters,
	const float* __restrict__ coords, const __half* __restrict__ rgbsigma, float4* __restrict__ frame, float* __restrict__ depth_out, uint32_t* __restrict__ steps_out)
{
	__shared__ uint32_t warp_sums[kRayBlock / 32 + 1];
	__shared__ uint32_t block_base;
	const uint32_t n_alive = counters->n_alive[cur];
	const uint32_t n_steps = counters->window_steps;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (blockIdx.x * blockDim.x >= n_alive) return;
	bool keep = false;
	float4 r0{}, r1{}, rgba{}, r3{};
	if (i < n_alive) {
		r0 = rays_cur[i];
		r1 = rays_cur[i + p.capacity];
		rgba = rays_cur[i + 2 * p.capacity];
		r3 = rays_cur[i + 3 * p.capacity];
		const uint32_t my_steps = ray_steps[i];
		const V3 cam_fwd = p.camera.c[2], cam_o = p.camera.c[3];
		float depth = r3.x, max_weight = r3.y;
		uint32_t j = 0;
		for (; j < my_steps; ++j) {
			const size_t s = (size_t)i + (size_t)j * n_alive;
			const uint2 raw = __ldg(reinterpret_cast<const uint2*>(rgbsigma) + s);
			const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), bs = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
			const float* c = coords + s * kCoordFloats;
			const V3 pos = unwarp_position(v3(c[0], c[1], c[2]), p.aabb);
			const float T = 1.0f - rgba.w;
			const float dt = unwarp_dt(c[3]);
			const float alpha = composite_alpha(network_to_density(bs.y, p.density_activation), dt);
			const float weight = alpha * T;
			rgba.x += network_to_rgb(rg.x, p.rgb_activation) * weight;
			rgba.y += network_to_rgb(rg.y, p.rgb_activation) * weight;
			rgba.z += network_to_rgb(bs.x, p.rgb_activation) * weight;
			rgba.w += weight;
			if (weight > max_weight) {
				max_weight = weight;
				depth = dot(cam_fwd, pos - cam_o);
			}
			if (rgba.w > 1.0f - p.min_transmittance) {
				const float inv = 1.0f / rgba.w;
				rgba.x *= inv; rgba.y *= inv; rgba.z *= inv; rgba.w = 1.0f;
				break;
			}
		}
		const uint32_t taken = __float_as_uint(r3.z) + (j < my_steps ? j + 1 : j);
		r3 = make_float4(depth, max_weight, __uint_as_float(taken), 0.0f);
		keep = j >= n_steps;
		if (!keep) {
			const uint32_t pix = __float_as_uint(r1.w);
			if (steps_out) steps_out[pix] = taken;
			if (rgba.w > 0.001f) {
				float4 tmp = rgba;
				if (!p.linear_colors) { tmp.x = srgb_to_linear(tmp.x); tmp.y = srgb_to_linear(tmp.y); tmp.z = srgb_to_linear(tmp.z); }
				frame[pix] = tmp;
				if (depth_out && tmp.w > 0.2f) depth_out[pix] = depth;
			}
		}
	}
	uint32_t total;
	const uint32_t prefix = block_exclusive_scan(keep ? 1u : 0u, total, warp_sums);
	if (threadIdx.x == 0) block_base = total ? atomicAdd(&counters->n_alive[cur ^ 1u], total) : 0u;
	__syncthreads();
	if (!keep) return;
	const uint32_t slot = block_base + prefix;
	rays_next[slot] = r0;
	rays_next[slot + p.capacity] = r1;
	rays_next[slot + 2 * p.capacity] = rgba;
	rays_next[slot + 3 * p.capacity] = r3;
}

void Testbed::render_nerf(uint32_t width, uint32_t height, const float* camera_3x4, float focal_x, float focal_y, float principal_x, float principal_y,
	uint32_t spp_index, uint32_t pixel_begin, uint32_t pixel_end, float* rgba_out, float* depth_out)
{
	if (!m_network) throw std::runtime_error{"ngp-b200: render_nerf needs a network (reload_network_from_json first)"};
	if (width == 0 || height == 0) throw std::runtime_error{"ngp-b200: render resolution must be positive"};
	const uint32_t n_pixels = width * height;
	pixel_end = std::min(pixel_end, n_pixels);
	last_render_samples = 0;
	if (pixel_begin >= pixel_end) return;
	if (render_shard_world == 0 || render_shard_rank >= render_shard_world) throw std::runtime_error{"ngp-b200: render_shard_rank must be < render_shard_world"};
	const uint32_t n_tiles = div_round_up(pixel_end - pixel_begin, kRenderTile);
	const uint32_t my_tiles = n_tiles > render_shard_rank ? div_round_up(n_tiles - render_shard_rank, render_shard_world) : 0;
	const uint32_t n_rays = render_shard_world == 1 ? pixel_end - pixel_begin : my_tiles * kRenderTile;
	if (n_rays == 0) return;

	RenderParams p{};
	p.aabb = m_aabb;
	for (int c = 0; c < 4; ++c) p.camera.c[c] = {camera_3x4[0 * 4 + c], camera_3x4[1 * 4 + c], camera_3x4[2 * 4 + c]};
	p.width = width; p.height = height;
	p.focal_x = focal_x; p.focal_y = focal_y; p.screen_cx = principal_x; p.screen_cy = principal_y;
	p.spp_index = spp_index;
	p.lens_mode = render_with_lens_distortion ? render_lens_mode : kLensPerspective;
	for (int i = 0; i < 7; ++i) p.lens_params[i] = render_lens_params[i];
	p.pixel_begin = pixel_begin; p.pixel_end = pixel_end;
	p.max_mip = m_max_cascade;
	p.skip_mip = render_skip_past_last_cascade ? kSkipMaxMip : m_max_cascade;
	p.n_local = n_rays;
	p.shard_rank = render_shard_rank; p.shard_world = render_shard_world;
	p.capacity = next_multiple(n_rays, kBatchSizeGranularity);
	// A shard holds 1 / world of the frame's rays. Its windows are latency-bound (a march kernel over 1/8 of the rays
	// takes a third of the whole frame's, not an eighth), so what a shard wants is FEW windows, not the whole frame's
	// window lengths: measured on one of 8 shards of the 1080p fox frame (scripts/profile_render.py), budget / 8 runs 20
	// windows = 3.33 ms, the frame's budget 8 windows = 2.44 ms, twice the budget 4 windows = 2.24 ms at 33.6 instead of
	// 30.5 queries per ray (max_steps caps a window at 32 steps). The whole frame on one GPU is best at 1x (12.2 ms
	// against 12.6 at 2x) and so is half a frame (2 shards: 7.76 ms at 1x, 7.94 at 2x), so the doubled budget goes to ray sets
	// of at most 2^19 rays - a quarter of a 1080p frame or less (4 shards: 4.83 ms, 8 shards: 2.92 ms).
	p.target_queries = std::max(render_target_queries * (n_rays <= (1u << 19) ? 2u : 1u), 1u);
	p.min_steps = std::max(render_min_steps, 1u);
	p.max_steps = std::max(render_max_steps, p.min_steps);
	p.min_transmittance = render_min_transmittance;
	p.rgb_activation = rgb_activation; p.density_activation = density_activation;
	p.linear_colors = linear_colors; p.snap_to_pixel_centers = render_snap_to_pixel_centers;
	const Stepping stepping = make_stepping(m_cone_angle_constant);

	// one allocation: 2 x 4 ray planes | per-ray window steps | coords | rgb sigma | counters
	// (a window holds n_alive * clamp(target / n_alive, min_steps, max_steps) samples: never more than capacity * max_steps)
	const size_t max_samples = next_multiple((uint32_t)std::min<size_t>(std::max<size_t>(p.target_queries, (size_t)p.capacity * p.min_steps), (size_t)p.capacity * p.max_steps) + kTile, kTile);
	const size_t plane_bytes = (size_t)p.capacity * sizeof(float4);
	size_t off = 0;
	auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) / 256 * 256; return o; };
	const size_t o_rays0 = take(4 * plane_bytes), o_rays1 = take(4 * plane_bytes), o_steps = take((size_t)p.capacity * sizeof(uint32_t));
	const size_t o_coords = take(max_samples * kCoordFloats * sizeof(float)), o_out = take(max_samples * 4 * sizeof(__half)), o_counters = take(sizeof(RenderCounters));
	if (m_render_scratch.bytes < off) {
		m_render_scratch.reserve(off, m_stream);
		// padding samples of a window's last tile are read by the encoder: keep them finite
		NGPB_CUDA_CHECK(cudaMemsetAsync(m_render_scratch.ptr, 0, off, m_stream));
	}
	char* base = m_render_scratch.as<char>();
	float4* rays[2] = {reinterpret_cast<float4*>(base + o_rays0), reinterpret_cast<float4*>(base + o_rays1)};
	uint32_t* ray_steps = reinterpret_cast<uint32_t*>(base + o_steps);
	float* coords = reinterpret_cast<float*>(base + o_coords);
	__half* rgbsigma = reinterpret_cast<__half*>(base + o_out);
	RenderCounters* counters = reinterpret_cast<RenderCounters*>(base + o_counters);
	float4* frame = reinterpret_cast<float4*>(rgba_out);
	const uint8_t* bitfield = m_density_bitfield.as<uint8_t>();
	const __half* net_params = params_inference();

	NGPB_CUDA_CHECK(cudaMemsetAsync(counters, 0, sizeof(RenderCounters), m_stream));
	prof_mark(m_stream, "render.init_rays");
	render_init_rays_kernel<<<div_round_up(n_rays, kRayBlock), kRayBlock, 0, m_stream>>>(p, stepping, bitfield, rays[0], counters, frame, depth_out);
	NGPB_CUDA_CHECK(cudaGetLastError());

	RenderCounters* host = reinterpret_cast<RenderCounters*>(m_counters_host_render);
	uint32_t bound = n_rays; // upper bound of the live count known to the host
	uint32_t cur = 0, steps_done = 1, chunk = std::max(render_windows_per_sync, 1u);
	const uint32_t kMarchIter = 10000; // MARCH_ITER (:37)
	while (steps_done < kMarchIter && bound > 0) {
		for (uint32_t k = 0; k < chunk && steps_done < kMarchIter; ++k) {
			const uint32_t blocks = div_round_up(bound, kRayBlock);
			prof_mark(m_stream, "render.march");
			if (stepping.cone_angle <= 1e-5f) render_march_kernel<false><<<blocks, kRayBlock, 0, m_stream>>>(p, stepping, bitfield, cur, rays[cur], ray_steps, counters, coords);
			else render_march_kernel<true><<<blocks, kRayBlock, 0, m_stream>>>(p, stepping, bitfield, cur, rays[cur], ray_steps, counters, coords);
			const uint64_t hi = (uint64_t)bound * p.max_steps, lo = std::max<uint64_t>(p.target_queries, (uint64_t)bound * p.min_steps);
			const uint32_t max_n = next_multiple((uint32_t)std::min<uint64_t>(std::min(hi, lo), max_samples - kTile), kTile);
			m_network->inference_counted(m_stream, max_n, &counters->n_samples, coords, kCoordFloats, rgbsigma, 4, net_params);
			prof_mark(m_stream, "render.composite");
			render_composite_kernel<<<blocks, kRayBlock, 0, m_stream>>>(p, cur, rays[cur], rays[cur ^ 1u], ray_steps, counters, coords, rgbsigma, frame, depth_out, m_render_steps_out);
			NGPB_CUDA_CHECK(cudaGetLastError());
			cur ^= 1u;
			// the window length is only known on the device; the reference advances its iteration
			// budget by it (:1789), here by the minimum so the budget is never exceeded early
			steps_done += p.min_steps;
		}
		prof_mark(m_stream, "render.sync");
		NGPB_CUDA_CHECK(cudaMemcpyAsync(host, counters, sizeof(RenderCounters), cudaMemcpyDeviceToHost, m_stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
		bound = host->n_alive[cur];
	}
	prof_mark(m_stream, "end");
	last_render_samples = host->total_samples;
}

} // namespace ngpb
