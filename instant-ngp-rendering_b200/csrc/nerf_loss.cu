// compute_loss_kernel_train_nerf (src/testbed_nerf.cu:841-1160) and the roll-over padding of the
// compacted batch (tcnn common_device.h:940-960, launched at src/testbed_nerf.cu:2862-2870).
// One warp per ray: composite with early stop at T < 1e-4, target lookup with the ray's
// random background, loss, then the per-sample dL/d(rgb, sigma) and the compacted coordinates.
// Network output and gradient are packed 4 halves per sample (the reference strides through
// 16-wide padded rows and reads the same 4 values, :915, :1075).
#include "nerf.h"
#include "block_scan.cuh"

namespace ngpb {

struct Sample4 {
	float r, g, b, s;
};
NGPB_D Sample4 load_rgbsigma(const __half* p) {
	const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
	const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
	return {a.x, a.y, b.x, b.y};
}

// One WARP per ray. The reference walks a ray's samples three times in one thread
// (:913-935 composite, :1065-1119 gradients); a training step only has a few thousand rays, so that
// shape leaves a B200 empty. Here the lanes take consecutive samples: loads, activations and
// exponentials run 32 wide, transmittance is carried through the chunk by a 32-step shuffle chain
// in the reference's order (T *= 1 - alpha, one rounding per sample, so the early-stop index is
// the one the serial loop finds), colour sums are warp reductions / scans.
constexpr uint32_t kRaysPerBlock = kRayBlock / 32;

struct SampleEval {
	V3 rgb;
	float alpha, dt, density;
	Sample4 o;
};
NGPB_D SampleEval eval_sample(const NerfTrainConfig& cfg, const __half* out, const float* coords, uint32_t j) {
	SampleEval e;
	e.o = load_rgbsigma(out + (size_t)j * 4);
	e.rgb = {network_to_rgb(e.o.r, cfg.rgb_activation), network_to_rgb(e.o.g, cfg.rgb_activation), network_to_rgb(e.o.b, cfg.rgb_activation)};
	e.dt = unwarp_dt(__ldg(coords + (size_t)j * kCoordFloats + 3));
	e.density = network_to_density(e.o.s, cfg.density_activation);
	e.alpha = composite_alpha(e.density, e.dt);
	return e;
}
// Transmittance in front of every lane's sample given T in front of the chunk; returns T behind it.
NGPB_D float chunk_transmittance(float one_minus_alpha, float T_in, float& T_before) {
	const uint32_t lane = threadIdx.x % 32;
	float T = T_in;
#pragma unroll
	for (uint32_t k = 0; k < 32; ++k) {
		if (lane == k) T_before = T;
		T *= __shfl_sync(0xFFFFFFFFu, one_minus_alpha, k);
	}
	return T;
}
NGPB_D float warp_sum(float v) {
#pragma unroll
	for (uint32_t o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
	return v;
}
NGPB_D float warp_inclusive_sum(float v) {
	const uint32_t lane = threadIdx.x % 32;
#pragma unroll
	for (uint32_t o = 1; o < 32; o <<= 1) {
		const float t = __shfl_up_sync(0xFFFFFFFFu, v, o);
		if (lane >= o) v += t;
	}
	return v;
}

__global__ void __launch_bounds__(kRayBlock) compute_loss_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf, Pcg32 rng)
{
	__shared__ uint32_t ray_counts[kRaysPerBlock];
	__shared__ uint32_t block_base;

	const uint32_t lane = threadIdx.x % 32, warp = threadIdx.x / 32;
	const uint32_t i = blockIdx.x * kRaysPerBlock + warp;
	const bool active = i < buf.counters->rays;

	uint32_t numsteps = 0, base = 0, compacted_numsteps = 0;
	float T = 1.0f;
	const float EPSILON = 1e-4f;
	V3 rgb_ray = v3(0.0f);
	V3 ray_o{};
	const float* coords_in = nullptr;
	const __half* out = nullptr;
	if (active) {
		numsteps = buf.numsteps[i * 2 + 0];
		base = buf.numsteps[i * 2 + 1];
		const float* ro = buf.rays_unnormalized + (size_t)i * 6;
		ray_o = {ro[0], ro[1], ro[2]};
		coords_in = buf.coords + (size_t)base * kCoordFloats;
		out = buf.rgbsigma + (size_t)base * 4;
		for (uint32_t c0 = 0; c0 < numsteps; c0 += 32) {
			const uint32_t j = c0 + lane;
			SampleEval e{};
			e.alpha = 0.0f;
			if (j < numsteps) e = eval_sample(cfg, out, coords_in, j);
			float T_before = 1.0f;
			const float T_out = chunk_transmittance(1.f - e.alpha, T, T_before);
			// the serial loop stops in front of the first sample that sees T < EPSILON
			const uint32_t stop = __ballot_sync(0xFFFFFFFFu, j >= numsteps || T_before < EPSILON);
			const uint32_t n_take = stop ? (uint32_t)__ffs(stop) - 1 : 32u;
			const float weight = lane < n_take ? e.alpha * T_before : 0.0f;
			rgb_ray.x += warp_sum(weight * e.rgb.x);
			rgb_ray.y += warp_sum(weight * e.rgb.y);
			rgb_ray.z += warp_sum(weight * e.rgb.z);
			compacted_numsteps += n_take;
			if (n_take < 32) {
				// T behind the last processed sample = T in front of the first rejected one
				T = __shfl_sync(0xFFFFFFFFu, T_before, n_take);
				break;
			}
			T = T_out;
		}
	}

	// Compacted slot: every active ray takes part, in ray order within the block.
	if (lane == 0) ray_counts[warp] = active ? compacted_numsteps : 0u;
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t total = 0;
#pragma unroll
		for (uint32_t w = 0; w < kRaysPerBlock; ++w) total += ray_counts[w];
		block_base = total ? atomicAdd(&buf.counters->numsteps_compacted, total) : 0u;
	}
	__syncthreads();
	if (!active) return;
	uint32_t prefix = 0;
	for (uint32_t w = 0; w < warp; ++w) prefix += ray_counts[w];

	// Same generator position as the sampler to obtain the same pixel and background colour.
	const uint32_t ray_idx = buf.ray_indices[i];
	rng.advance((int64_t)ray_idx * kMaxRandomSamplesPerRay);
	const uint32_t img = image_idx(ray_idx, cfg.n_rays, cfg.n_training_images);
	const ImageMeta& m = buf.metadata[img];
	float u, v;
	random_image_pos_training(rng, m.res_x, m.res_y, cfg.snap_to_pixel_centers != 0, u, v);
	rng.advance(1); // motionblur_time

	V3 background_color = {cfg.background_color[0], cfg.background_color[1], cfg.background_color[2]};
	if (cfg.random_bg_color) {
		background_color.x = rng.next_float();
		background_color.y = rng.next_float();
		background_color.z = rng.next_float();
	}
	background_color = srgb_to_linear(background_color);

	const Rgba texsamp = read_rgba_uv(u, v, m.res_x, m.res_y, m.pixels, m.image_data_type);
	const V3 tex_rgb = {texsamp.r, texsamp.g, texsamp.b};
	V3 rgbtarget;
	if (cfg.linear_colors || cfg.color_space == kColorLinear) {
		rgbtarget = tex_rgb + (1.0f - texsamp.a) * background_color;
		if (!cfg.linear_colors) {
			rgbtarget = linear_to_srgb(rgbtarget);
			background_color = linear_to_srgb(background_color);
		}
	} else {
		background_color = linear_to_srgb(background_color);
		if (texsamp.a > 0) rgbtarget = linear_to_srgb(tex_rgb / texsamp.a) * texsamp.a + (1.0f - texsamp.a) * background_color;
		else rgbtarget = background_color;
	}

	if (compacted_numsteps == numsteps) rgb_ray += T * background_color;

	const uint32_t compacted_base = block_base + prefix;
	compacted_numsteps = min(cfg.max_samples_compacted - min(cfg.max_samples_compacted, compacted_base), compacted_numsteps);
	if (lane == 0) {
		buf.numsteps[i * 2 + 0] = compacted_numsteps;
		buf.numsteps[i * 2 + 1] = compacted_base;
	}
	if (compacted_numsteps == 0) return;

	LossAndGradient lg = loss_and_gradient(rgbtarget, rgb_ray, cfg.loss_type);
	const float mean_loss = mean(lg.loss);
	if (buf.loss && lane == 0) buf.loss[i] = mean_loss / (float)cfg.n_rays;

	const float loss_scale = cfg.loss_scale / (float)cfg.n_rays;
	const float output_l2_reg = cfg.rgb_activation == kActExponential ? 1e-4f : 0.0f;
	const float output_l1_reg_density = *buf.mean_density < kNerfMinOpticalThickness ? 1e-4f : 0.0f;

	float* coords_out = buf.coords_compacted + (size_t)compacted_base * kCoordFloats;
	__half* dout = buf.dloss_doutput + (size_t)compacted_base * 4;

	// the compacted copy of the coordinates: the ray's rows are contiguous on both sides
	for (uint32_t k = lane; k < compacted_numsteps * kCoordFloats; k += 32) coords_out[k] = coords_in[k];
	if (buf.compact_src) for (uint32_t k = lane; k < compacted_numsteps; k += 32) buf.compact_src[compacted_base + k] = base + k;

	V3 rgb_ray2 = v3(0.0f); // colour accumulated in front of the chunk
	T = 1.f;
	for (uint32_t c0 = 0; c0 < compacted_numsteps; c0 += 32) {
		const uint32_t j = c0 + lane;
		const bool mine = j < compacted_numsteps;
		SampleEval e{};
		e.alpha = 0.0f;
		if (mine) e = eval_sample(cfg, out, coords_in, j);
		float T_before = 1.0f;
		const float T_out = chunk_transmittance(1.f - e.alpha, T, T_before);
		const float weight = mine ? e.alpha * T_before : 0.0f;
		const V3 acc = {rgb_ray2.x + warp_inclusive_sum(weight * e.rgb.x), rgb_ray2.y + warp_inclusive_sum(weight * e.rgb.y), rgb_ray2.z + warp_inclusive_sum(weight * e.rgb.z)};
		if (mine) {
			const float T_after = T_before * (1.f - e.alpha);
			const float* c = coords_in + (size_t)j * kCoordFloats;
			const V3 pos = unwarp_position({c[0], c[1], c[2]}, cfg.aabb);
			const float depth = distance(pos, ray_o);
			// suffix = colour contributed by everything behind this sample
			const V3 suffix = rgb_ray - acc;
			const V3 dloss_by_drgb = weight * lg.gradient;

			const float d0 = loss_scale * (dloss_by_drgb.x * network_to_rgb_derivative(e.o.r, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * e.o.r));
			const float d1 = loss_scale * (dloss_by_drgb.y * network_to_rgb_derivative(e.o.g, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * e.o.g));
			const float d2 = loss_scale * (dloss_by_drgb.z * network_to_rgb_derivative(e.o.b, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * e.o.b));

			const float density_derivative = network_to_density_derivative(e.o.s, cfg.density_activation);
			const float dloss_by_dmlp = density_derivative * (e.dt * dot(lg.gradient, T_after * e.rgb - suffix));
			const float d3 = loss_scale * dloss_by_dmlp + (e.o.s < 0.0f ? -output_l1_reg_density : 0.0f) + (e.o.s > -10.0f && depth < cfg.near_distance ? 1e-4f : 0.0f);

			const __half2 h01 = __floats2half2_rn(d0, d1), h23 = __floats2half2_rn(d2, d3);
			*reinterpret_cast<uint2*>(dout + (size_t)j * 4) = make_uint2(*reinterpret_cast<const uint32_t*>(&h01), *reinterpret_cast<const uint32_t*>(&h23));
		}
		rgb_ray2 = {__shfl_sync(0xFFFFFFFFu, acc.x, 31), __shfl_sync(0xFFFFFFFFu, acc.y, 31), __shfl_sync(0xFFFFFFFFu, acc.z, 31)};
		T = T_out;
	}
}

// fill_rollover_and_rescale on the gradients, fill_rollover on the coordinates: the compacted
// batch is padded to the full batch size by wrapping around; gradients of the padding (and only
// of the padding, as in the reference) are scaled by n_input / n_elements.
__global__ void fill_rollover_kernel(uint32_t n_elements, const NerfStepCounters* __restrict__ counters, float* __restrict__ coords, __half* __restrict__ dout, uint32_t* __restrict__ compact_src) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t n_in = min(counters->numsteps_compacted, n_elements);
	if (i < n_in || i >= n_elements || n_in == 0) return;
	const uint32_t src = i % n_in;
#pragma unroll
	for (uint32_t k = 0; k < kCoordFloats; ++k) coords[(size_t)i * kCoordFloats + k] = coords[(size_t)src * kCoordFloats + k];
#pragma unroll
	for (uint32_t k = 0; k < 4; ++k) dout[(size_t)i * 4 + k] = __float2half(__half2float(dout[(size_t)src * 4 + k]) * n_in / n_elements);
	if (compact_src) compact_src[i] = compact_src[src];
}

void nerf_compute_loss_and_compact(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng) {
	if (local_ray_count(cfg) == 0) return;
	compute_loss_kernel<<<div_round_up(local_ray_count(cfg), kRaysPerBlock), kRayBlock, 0, stream>>>(cfg, buf, rng);
	NGPB_CUDA_CHECK(cudaGetLastError());
	fill_rollover_kernel<<<div_round_up(cfg.max_samples_compacted, 256), 256, 0, stream>>>(cfg.max_samples_compacted, buf.counters, buf.coords_compacted, buf.dloss_doutput, buf.compact_src);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
