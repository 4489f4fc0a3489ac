// compute_loss_kernel_train_nerf (src/testbed_nerf.cu:841-1160) and the roll-over padding of the
// compacted batch (tcnn common_device.h:940-960, launched at src/testbed_nerf.cu:2862-2870).
// One thread per ray: composite with early stop at T < 1e-4, target lookup with the ray's
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

__global__ void __launch_bounds__(kRayBlock) compute_loss_kernel(const __grid_constant__ NerfTrainConfig cfg, const __grid_constant__ NerfTrainBuffers buf, Pcg32 rng)
{
	__shared__ uint32_t warp_sums[kRayBlock / 32 + 1];
	__shared__ uint32_t block_base;

	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const bool active = i < buf.counters->rays;

	uint32_t numsteps = 0, base = 0, compacted_numsteps = 0;
	float T = 1.0f;
	const float EPSILON = 1e-4f;
	V3 rgb_ray = v3(0.0f);
	V3 ray_o{};
	if (active) {
		numsteps = buf.numsteps[i * 2 + 0];
		base = buf.numsteps[i * 2 + 1];
		const float* ro = buf.rays_unnormalized + (size_t)i * 6;
		ray_o = {ro[0], ro[1], ro[2]};
		const float* coords = buf.coords + (size_t)base * kCoordFloats;
		const __half* out = buf.rgbsigma + (size_t)base * 4;
		for (; compacted_numsteps < numsteps; ++compacted_numsteps) {
			if (T < EPSILON) break;
			const Sample4 o = load_rgbsigma(out + (size_t)compacted_numsteps * 4);
			const V3 rgb = {network_to_rgb(o.r, cfg.rgb_activation), network_to_rgb(o.g, cfg.rgb_activation), network_to_rgb(o.b, cfg.rgb_activation)};
			const float dt = unwarp_dt(coords[(size_t)compacted_numsteps * kCoordFloats + 3]);
			const float density = network_to_density(o.s, cfg.density_activation);
			const float alpha = 1.f - __expf(-density * dt);
			const float weight = alpha * T;
			rgb_ray += weight * rgb;
			T *= (1.f - alpha);
		}
	}

	// Compacted slot: every active ray takes part, in ray order within the block.
	uint32_t block_total;
	const uint32_t prefix = block_exclusive_scan(active ? compacted_numsteps : 0u, block_total, warp_sums);
	if (threadIdx.x == 0) block_base = block_total ? atomicAdd(&buf.counters->numsteps_compacted, block_total) : 0u;
	__syncthreads();
	if (!active) return;

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
	buf.numsteps[i * 2 + 0] = compacted_numsteps;
	buf.numsteps[i * 2 + 1] = compacted_base;
	if (compacted_numsteps == 0) return;

	LossAndGradient lg = loss_and_gradient(rgbtarget, rgb_ray, cfg.loss_type);
	const float mean_loss = mean(lg.loss);
	if (buf.loss) buf.loss[i] = mean_loss / (float)cfg.n_rays;

	const float loss_scale = cfg.loss_scale / (float)cfg.n_rays;
	const float output_l2_reg = cfg.rgb_activation == kActExponential ? 1e-4f : 0.0f;
	const float output_l1_reg_density = *buf.mean_density < kNerfMinOpticalThickness ? 1e-4f : 0.0f;

	const float* coords_in = buf.coords + (size_t)base * kCoordFloats;
	const __half* out = buf.rgbsigma + (size_t)base * 4;
	float* coords_out = buf.coords_compacted + (size_t)compacted_base * kCoordFloats;
	__half* dout = buf.dloss_doutput + (size_t)compacted_base * 4;

	V3 rgb_ray2 = v3(0.0f);
	T = 1.f;
	for (uint32_t j = 0; j < compacted_numsteps; ++j) {
		float c[kCoordFloats];
#pragma unroll
		for (uint32_t k = 0; k < kCoordFloats; ++k) {
			c[k] = coords_in[(size_t)j * kCoordFloats + k];
			coords_out[(size_t)j * kCoordFloats + k] = c[k];
		}
		const V3 pos = unwarp_position({c[0], c[1], c[2]}, cfg.aabb);
		const float depth = distance(pos, ray_o);
		const float dt = unwarp_dt(c[3]);
		const Sample4 o = load_rgbsigma(out + (size_t)j * 4);
		const V3 rgb = {network_to_rgb(o.r, cfg.rgb_activation), network_to_rgb(o.g, cfg.rgb_activation), network_to_rgb(o.b, cfg.rgb_activation)};
		const float density = network_to_density(o.s, cfg.density_activation);
		const float alpha = 1.f - __expf(-density * dt);
		const float weight = alpha * T;
		rgb_ray2 += weight * rgb;
		T *= (1.f - alpha);

		// suffix = colour contributed by everything behind this sample
		const V3 suffix = rgb_ray - rgb_ray2;
		const V3 dloss_by_drgb = weight * lg.gradient;

		const float d0 = loss_scale * (dloss_by_drgb.x * network_to_rgb_derivative(o.r, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o.r));
		const float d1 = loss_scale * (dloss_by_drgb.y * network_to_rgb_derivative(o.g, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o.g));
		const float d2 = loss_scale * (dloss_by_drgb.z * network_to_rgb_derivative(o.b, cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o.b));

		const float density_derivative = network_to_density_derivative(o.s, cfg.density_activation);
		const float dloss_by_dmlp = density_derivative * (dt * dot(lg.gradient, T * rgb - suffix));
		const float d3 = loss_scale * dloss_by_dmlp + (o.s < 0.0f ? -output_l1_reg_density : 0.0f) + (o.s > -10.0f && depth < cfg.near_distance ? 1e-4f : 0.0f);

		const __half2 h01 = __floats2half2_rn(d0, d1), h23 = __floats2half2_rn(d2, d3);
		*reinterpret_cast<uint2*>(dout + (size_t)j * 4) = make_uint2(*reinterpret_cast<const uint32_t*>(&h01), *reinterpret_cast<const uint32_t*>(&h23));
	}
}

// fill_rollover_and_rescale on the gradients, fill_rollover on the coordinates: the compacted
// batch is padded to the full batch size by wrapping around; gradients of the padding (and only
// of the padding, as in the reference) are scaled by n_input / n_elements.
__global__ void fill_rollover_kernel(uint32_t n_elements, const NerfStepCounters* __restrict__ counters, float* __restrict__ coords, __half* __restrict__ dout) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t n_in = min(counters->numsteps_compacted, n_elements);
	if (i < n_in || i >= n_elements || n_in == 0) return;
	const uint32_t src = i % n_in;
#pragma unroll
	for (uint32_t k = 0; k < kCoordFloats; ++k) coords[(size_t)i * kCoordFloats + k] = coords[(size_t)src * kCoordFloats + k];
#pragma unroll
	for (uint32_t k = 0; k < 4; ++k) dout[(size_t)i * 4 + k] = __float2half(__half2float(dout[(size_t)src * 4 + k]) * n_in / n_elements);
}

void nerf_compute_loss_and_compact(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng) {
	if (local_ray_count(cfg) == 0) return;
	compute_loss_kernel<<<div_round_up(local_ray_count(cfg), kRayBlock), kRayBlock, 0, stream>>>(cfg, buf, rng);
	NGPB_CUDA_CHECK(cudaGetLastError());
	fill_rollover_kernel<<<div_round_up(cfg.max_samples_compacted, 256), 256, 0, stream>>>(cfg.max_samples_compacted, buf.counters, buf.coords_compacted, buf.dloss_doutput);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
