// Neural-image mode of the Testbed: Testbed::train_image / render_image
// (src/testbed_image.cu:60-74 stratify2_kernel, :164-210 eval_image_kernel_and_snap, :212-275
// train_image, :277-340 render_image) and the tcnn losses the image configs use
// (tcnn losses/l2.h:38-75, relative_l2.h:38-76, l1.h) behind tcnn::Trainer::training_step
// (trainer.h:163-190): forward, loss, backward, optimizer step.
#include "testbed.h"

#include <algorithm>

namespace ngpb {

// stratify2_kernel: sample i of a 2^k batch is moved into cell (i mod 2^(k/2), i div 2^(k/2)).
__global__ void stratify2_kernel(uint32_t n, uint32_t log2_batch_size, float* __restrict__ inout) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t log2_size = log2_batch_size / 2;
	const uint32_t size = 1u << log2_size;
	const uint32_t in_batch = i & ((1u << log2_batch_size) - 1);
	const uint32_t x = in_batch & (size - 1), y = in_batch >> log2_size;
	const float vx = inout[2 * i], vy = inout[2 * i + 1];
	inout[2 * i] = vx / (float)size + ((float)x / (float)size);
	inout[2 * i + 1] = vy / (float)size + ((float)y / (float)size);
}

template <typename T> NGPB_D float4 load_texel(const void* texture, size_t idx);
template <> NGPB_D float4 load_texel<float>(const void* texture, size_t idx) { return __ldg(reinterpret_cast<const float4*>(texture) + idx); }
template <> NGPB_D float4 load_texel<__half>(const void* texture, size_t idx) {
	const uint2 v = __ldg(reinterpret_cast<const uint2*>(texture) + idx);
	const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
	return make_float4(a.x, a.y, b.x, b.y);
}

// eval_image_kernel_and_snap<T, 3>: target colour at each position (nearest texel when snapping -
// the position is moved to the texel centre - else bilinear), sRGB-encoded unless linear_colors.
template <typename T>
__global__ void eval_image_and_snap_kernel(uint32_t n, const void* __restrict__ texture, float* __restrict__ positions, int res_x, int res_y,
	float* __restrict__ result, int snap_to_pixel_centers, int linear_colors)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float px = positions[2 * i], py = positions[2 * i + 1];
	auto read_val = [&](int x, int y) {
		float4 v = load_texel<T>(texture, (size_t)y * res_x + x);
		if (!linear_colors) { v.x = linear_to_srgb(v.x); v.y = linear_to_srgb(v.y); v.z = linear_to_srgb(v.z); }
		return v;
	};
	float4 val;
	if (snap_to_pixel_centers) {
		int ix = (int)floorf(px * (float)res_x), iy = (int)floorf(py * (float)res_y);
		positions[2 * i] = ((float)ix + 0.5f) / (float)res_x;
		positions[2 * i + 1] = ((float)iy + 0.5f) / (float)res_y;
		ix = clamp_i(ix, 0, res_x - 1);
		iy = clamp_i(iy, 0, res_y - 1);
		val = read_val(ix, iy);
	} else {
		px = clamp_f(px * (float)res_x - 0.5f, 0.0f, (float)res_x - (1.0f + 1e-4f));
		py = clamp_f(py * (float)res_y - 0.5f, 0.0f, (float)res_y - (1.0f + 1e-4f));
		const int ix0 = (int)px, iy0 = (int)py;
		const float wx = px - (float)ix0, wy = py - (float)iy0;
		const int ix = clamp_i(ix0, 0, res_x - 2), iy = clamp_i(iy0, 0, res_y - 2);
		const float4 a = read_val(ix, iy), b = read_val(ix + 1, iy), c = read_val(ix, iy + 1), d = read_val(ix + 1, iy + 1);
		const float w00 = (1 - wx) * (1 - wy), w10 = wx * (1 - wy), w01 = (1 - wx) * wy, w11 = wx * wy;
		val = make_float4(w00 * a.x + w10 * b.x + w01 * c.x + w11 * d.x, w00 * a.y + w10 * b.y + w01 * c.y + w11 * d.y,
			w00 * a.z + w10 * b.z + w01 * c.z + w11 * d.z, 0.0f);
	}
	result[3 * i + 0] = val.x;
	result[3 * i + 1] = val.y;
	result[3 * i + 2] = val.z;
}

// l2_loss / relative_l2_loss / l1_loss over fp16 predictions [n][stride] and fp32 targets [n][dims]:
// per-element loss values and dL/dprediction = loss_scale * gradient / (n * dims); padding
// columns get zero.
__global__ void image_loss_kernel(uint32_t n_elements, uint32_t stride, uint32_t dims, float loss_scale, int loss_type,
	const __half* __restrict__ predictions, const float* __restrict__ targets, float* __restrict__ values, __half* __restrict__ gradients)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const uint32_t intra = i % stride, inter = i / stride;
	if (intra >= dims) {
		values[i] = 0.0f;
		gradients[i] = __float2half(0.0f);
		return;
	}
	const float n_total = (float)(n_elements / stride * dims);
	const float prediction = __half2float(predictions[i]);
	const float difference = prediction - targets[inter * dims + intra];
	float value, gradient;
	if (loss_type == kLossRelativeL2) {
		const float denom = prediction * prediction + 0.01f;
		value = difference * difference / denom;
		gradient = 2 * difference / denom;
	} else if (loss_type == kLossL1) {
		value = fabsf(difference);
		gradient = copysignf(1.0f, difference);
	} else {
		value = difference * difference;
		gradient = 2 * difference;
	}
	values[i] = value / n_total;
	gradients[i] = __float2half(loss_scale * gradient / n_total);
}

__global__ void sum_kernel(uint32_t n, const float* __restrict__ v, float* __restrict__ out) {
	float s = 0.0f;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += v[i];
#pragma unroll
	for (uint32_t o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
	if (threadIdx.x % 32 == 0) atomicAdd(out, s);
}

// render_image without the camera / foveation indirection of the GUI path: network colour at every
// pixel centre, linearised unless the model was trained in linear colours (:131-162).
__global__ void image_positions_kernel(uint32_t n, uint32_t n_padded, int res_x, int res_y, float* __restrict__ positions) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_padded) return;
	const uint32_t j = i < n ? i : 0;
	positions[2 * i] = ((float)(j % res_x) + 0.5f) / (float)res_x;
	positions[2 * i + 1] = ((float)(j / res_x) + 0.5f) / (float)res_y;
}
__global__ void image_shade_kernel(uint32_t n, const __half* __restrict__ out, uint32_t stride, int linear_colors, float4* __restrict__ frame) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float r = __half2float(out[(size_t)i * stride]), g = __half2float(out[(size_t)i * stride + 1]), b = __half2float(out[(size_t)i * stride + 2]);
	if (!linear_colors) { r = srgb_to_linear(r); g = srgb_to_linear(g); b = srgb_to_linear(b); }
	frame[i] = make_float4(r, g, b, 1.0f);
}

// ---------------------------------------------------------------------------------------------
void Testbed::set_image(const void* pixels, int image_data_type, int width, int height) {
	if (image_data_type != kImageHalf && image_data_type != kImageFloat) throw std::runtime_error{"ngp-b200: image data must be Half or Float RGBA"};
	if (width < 2 || height < 2) throw std::runtime_error{"Image resolution must be at least 2x2."};
	const size_t bytes = (size_t)width * height * (image_data_type == kImageHalf ? 8 : 16);
	m_image_data.alloc(bytes, m_stream);
	NGPB_CUDA_CHECK(cudaMemcpyAsync(m_image_data.ptr, pixels, bytes, cudaMemcpyDefault, m_stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	m_image_type = image_data_type;
	m_image_res_x = width;
	m_image_res_y = height;
	m_mode = kModeImage;
	m_aabb = {{0, 0, 0}, {1, 1, 1}};
}

void Testbed::reset_network_image() {
	Json config = m_network_config;
	if (!config.contains("encoding") || !config.contains("network")) throw std::runtime_error{"Network config must contain 'encoding' and 'network'."};
	Json encoding_config = config["encoding"];
	const Json optimizer_config = config.value("optimizer", Json::object());
	// hash-grid hyper-parameters (src/testbed.cu:3679-3713) with n_pos = 2, desired resolution max(res) / 2
	const std::string otype = encoding_config.value("otype", "OneBlob");
	if (equals_case_insensitive(otype, "HashGrid") || equals_case_insensitive(otype, "Grid") || equals_case_insensitive(otype, "DenseGrid") || equals_case_insensitive(otype, "TiledGrid")) {
		const uint32_t n_pos = 2;
		const uint32_t F = encoding_config.value("n_features_per_level", 2u);
		uint32_t n_levels;
		if (encoding_config.contains("n_features") && encoding_config["n_features"].as_uint() > 0) n_levels = encoding_config["n_features"].as_uint() / F;
		else n_levels = encoding_config.value("n_levels", 16u);
		const uint32_t log2_hashmap_size = encoding_config.value("log2_hashmap_size", 15u);
		uint32_t base_resolution = encoding_config.value("base_resolution", 0u);
		if (!base_resolution) {
			base_resolution = 1u << (log2_hashmap_size / n_pos);
			encoding_config["base_resolution"] = base_resolution;
		}
		const float desired_resolution = (float)std::max(m_image_res_x, m_image_res_y) / 2.0f;
		float per_level_scale = encoding_config.value("per_level_scale", 0.0f);
		if (per_level_scale <= 0.0f && n_levels > 1) {
			per_level_scale = std::exp(std::log(desired_resolution * (float)aabb_scale / (float)base_resolution) / (n_levels - 1));
			encoding_config["per_level_scale"] = per_level_scale;
		}
	}
	m_network_config["encoding"] = encoding_config;
	m_image_network = std::make_unique<NetworkWithInputEncodingModule>(2u, 3u, encoding_config, config["network"]);
	m_image_network->set_ordered_batches(true);   // stratified positions walk the image in order (stratify2)
	m_optimizer = std::make_unique<Optimizer>(optimizer_config);

	m_n_params = m_image_network->n_params();
	m_optimizer->allocate((uint32_t)m_n_params, m_image_network->layer_sizes());
	const size_t n_pad = next_multiple((uint32_t)m_n_params, 8u);
	m_params_fp32.alloc(n_pad * sizeof(float), m_stream);
	m_params.alloc(n_pad * sizeof(__half), m_stream);
	m_gradients.alloc(n_pad * sizeof(__half), m_stream);
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_params_fp32.ptr, 0, n_pad * sizeof(float), m_stream));
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_gradients.ptr, 0, n_pad * sizeof(__half), m_stream));
	Pcg32 trainer_rng{seed};
	m_image_network->initialize_params(m_stream, trainer_rng, m_params_fp32.as<float>(), 1.0f);
	cast_fp32_to_fp16(m_stream, m_n_params, m_params_fp32.as<float>(), m_params.as<__half>());
	m_counters.reserve(sizeof(NerfStepCounters), m_stream);
	m_training_step = 0;
	m_loss_scalar = 0.0f;
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

void Testbed::train_image(uint32_t batch_size, bool get_loss_scalar) {
	const uint32_t n = batch_size;
	const uint32_t stride = kMlpPaddedOut;
	m_image_positions.reserve((size_t)n * 2 * sizeof(float), m_stream);
	m_image_targets.reserve((size_t)n * 3 * sizeof(float), m_stream);
	m_image_output.reserve((size_t)n * stride * sizeof(__half), m_stream);
	m_image_dloss.reserve((size_t)n * stride * sizeof(__half), m_stream);
	m_image_loss_values.reserve((size_t)n * stride * sizeof(float), m_stream);
	float* positions = m_image_positions.as<float>();

	// generate_training_data (:217-260), ERandomMode::Stratified
	prof_mark(m_stream, "image.generate");
	random_uniform(m_stream, m_rng, (size_t)n * 2, positions, 0.0f, 1.0f);
	uint32_t log2_batch = 0;
	while ((1u << log2_batch) < n) ++log2_batch;
	if ((1u << log2_batch) == n && log2_batch % 2 == 0) stratify2_kernel<<<div_round_up(n, 256), 256, 0, m_stream>>>(n, log2_batch, positions);
	if (m_image_type == kImageFloat) {
		eval_image_and_snap_kernel<float><<<div_round_up(n, 256), 256, 0, m_stream>>>(n, m_image_data.ptr, positions, m_image_res_x, m_image_res_y,
			m_image_targets.as<float>(), image_snap_to_pixel_centers, linear_colors);
	} else {
		eval_image_and_snap_kernel<__half><<<div_round_up(n, 256), 256, 0, m_stream>>>(n, m_image_data.ptr, positions, m_image_res_x, m_image_res_y,
			m_image_targets.as<float>(), image_snap_to_pixel_centers, linear_colors);
	}
	NGPB_CUDA_CHECK(cudaGetLastError());

	// Trainer::training_step (trainer.h:163-190)
	prof_mark(m_stream, "image.forward");
	auto ctx = m_image_network->forward(m_stream, n, positions, m_image_output.as<__half>(), params(), false);
	prof_mark(m_stream, "image.loss");
	image_loss_kernel<<<div_round_up(n * stride, 256), 256, 0, m_stream>>>(n * stride, stride, 3, kLossScaleFp16, loss_type, m_image_output.as<__half>(),
		m_image_targets.as<float>(), m_image_loss_values.as<float>(), m_image_dloss.as<__half>());
	NGPB_CUDA_CHECK(cudaGetLastError());
	prof_mark(m_stream, "image.backward");
	m_image_network->backward(m_stream, *ctx, n, nullptr, m_image_dloss.as<__half>(), gradients(), positions, m_image_output.as<__half>(), params());
	prof_mark(m_stream, "optimizer");
	m_optimizer->step(m_stream, kLossScaleFp16, params_fp32(), params(), gradients());
	prof_mark(m_stream, "end");
	++m_training_step;
	measured_batch_size = n;
	measured_batch_size_before_compaction = n;
	if (get_loss_scalar) {
		float* sum = &m_counters.as<NerfStepCounters>()->loss_sum;
		NGPB_CUDA_CHECK(cudaMemsetAsync(sum, 0, sizeof(float), m_stream));
		sum_kernel<<<sm_count(), 256, 0, m_stream>>>(n * stride, m_image_loss_values.as<float>(), sum);
		NGPB_CUDA_CHECK(cudaMemcpyAsync(m_loss_host, sum, sizeof(float), cudaMemcpyDeviceToHost, m_stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
		m_loss_scalar = *m_loss_host;
	}
}

void Testbed::render_image(uint32_t width, uint32_t height, float* rgba_out) {
	if (!m_image_network) throw std::runtime_error{"ngp-b200: render_image needs a network (reload_network_from_json first)"};
	const uint32_t n = width * height, n_padded = next_multiple(n, kBatchSizeGranularity);
	m_image_positions.reserve((size_t)n_padded * 2 * sizeof(float), m_stream);
	m_image_output.reserve((size_t)n_padded * kMlpPaddedOut * sizeof(__half), m_stream);
	image_positions_kernel<<<div_round_up(n_padded, 256), 256, 0, m_stream>>>(n, n_padded, (int)width, (int)height, m_image_positions.as<float>());
	m_image_network->inference(m_stream, n_padded, m_image_positions.as<float>(), m_image_output.as<__half>(), params_inference());
	image_shade_kernel<<<div_round_up(n, 256), 256, 0, m_stream>>>(n, m_image_output.as<__half>(), kMlpPaddedOut, linear_colors, reinterpret_cast<float4*>(rgba_out));
	NGPB_CUDA_CHECK(cudaGetLastError());
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

} // namespace ngpb
