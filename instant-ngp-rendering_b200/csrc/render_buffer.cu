// Headless render buffer: what CudaRenderBuffer does between the tracer and the pixels pyngp's
// render() returns (src/render_buffer.cu:232-346 accumulate / tonemap, :533-565 tonemap_kernel,
// :635-697 host side; src/python_api.cu:124-221 render_to_cpu). The GL surface becomes a plain
// device buffer; DLSS, depth splat and the VR overlays are not part of the path.
#include "testbed.h"

namespace ngpb {

enum : int { kTonemapIdentity = 0, kTonemapACES = 1, kTonemapHable = 2, kTonemapReinhard = 3 }; // ETonemapCurve

// accumulate_kernel: running mean over the samples per pixel, colours optionally kept in sRGB.
__global__ void accumulate_kernel(uint32_t n, const float4* __restrict__ frame, float4* __restrict__ accum, float sample_count, int color_space) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 color = frame[i];
	float4 tmp = accum[i];
	if (color_space == kColorSRGB) { color.x = linear_to_srgb(color.x); color.y = linear_to_srgb(color.y); color.z = linear_to_srgb(color.z); }
	tmp.x = (tmp.x * sample_count + color.x) / (sample_count + 1);
	tmp.y = (tmp.y * sample_count + color.y) / (sample_count + 1);
	tmp.z = (tmp.z * sample_count + color.z) / (sample_count + 1);
	tmp.w = (tmp.w * sample_count + color.w) / (sample_count + 1);
	accum[i] = tmp;
}

NGPB_D V3 tonemap_curve(V3 x, int curve) {
	if (curve == kTonemapIdentity) return x;
	x = {fmaxf(x.x, 0.0f), fmaxf(x.y, 0.0f), fmaxf(x.z, 0.0f)};
	float k0, k1, k2, k3, k4, k5;
	if (curve == kTonemapACES) {
		// Narkowicz's fit with the pre-exposure folded into the constants
		k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f;
		k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
	} else if (curve == kTonemapHable) {
		const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
		k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f;
		k3 = A * F; k4 = B * F; k5 = D * F * F;
		const float W = 11.2f;
		const float nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5;
		const float white_scale = denom / nom;
		k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale;
		k3 = 4.0f * k3; k4 = 2.0f * k4;
	} else { // Reinhard on the luminance
		const float Y = 0.2126f * x.x + 0.7152f * x.y + 0.0722f * x.z;
		return x * (1.f / (Y + 1.0f));
	}
	const V3 sq = x * x;
	const V3 nom = sq * k0 + k1 * x + k2, denom = k3 * sq + k4 * x + k5;
	return nom / denom;
}

// tonemap_kernel: background behind the premultiplied colour, exposure, curve, output colour space.
__global__ void tonemap_kernel(uint32_t n, float exposure, float4 background, const float4* __restrict__ accum, int color_space, int output_color_space,
	int curve, bool clamp_output_color, bool unmultiply_alpha, float4* __restrict__ out)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	// the background colour is given in sRGB
	V3 bg = {background.x, background.y, background.z};
	if (color_space != kColorSRGB) bg = srgb_to_linear(bg);
	const float4 a = accum[i];
	const float weight = (1 - a.w) * background.w;
	V3 col = v3(a.x, a.y, a.z) + bg * weight;
	float alpha = a.w + weight;
	if (color_space == kColorSRGB) col = srgb_to_linear(col);
	col = col * powf(2.0f, exposure);
	col = tonemap_curve(col, curve);
	if (output_color_space == kColorSRGB) col = linear_to_srgb(col);
	if (unmultiply_alpha && alpha > 0.0f) col = col / alpha;
	if (clamp_output_color) {
		col = {fminf(fmaxf(col.x, 0.0f), 1.0f), fminf(fmaxf(col.y, 0.0f), 1.0f), fminf(fmaxf(col.z, 0.0f), 1.0f)};
		alpha = fminf(fmaxf(alpha, 0.0f), 1.0f);
	}
	out[i] = make_float4(col.x, col.y, col.z, alpha);
}

// overlay_image_kernel (src/render_buffer.cu:348-411): the training image of a view resampled
// (nearest texel) onto the output grid along fov_axis, background, exposure, curve and transfer as in
// tonemap_kernel, blended over what the buffer holds. With alpha = 1 this IS the frame pyngp returns
// when render_ground_truth is set (Testbed::render_frame_epilogue, src/testbed.cu:4592-4608; the trace
// itself is skipped, :4383).
__global__ void overlay_image_kernel(uint32_t width, uint32_t height, float alpha, float exposure, float4 background, const void* __restrict__ image,
	int image_data_type, int image_res_x, int image_res_y, int curve, int color_space, int output_color_space, int fov_axis, float zoom,
	float screen_center_x, float screen_center_y, float4* __restrict__ out)
{
	const uint32_t x = threadIdx.x + blockDim.x * blockIdx.x, y = threadIdx.y + blockDim.y * blockIdx.y;
	if (x >= width || y >= height) return;
	const float scale = (float)(fov_axis == 0 ? image_res_x : image_res_y) / (float)(fov_axis == 0 ? width : height);
	float fx = x + 0.5f, fy = y + 0.5f;
	fx -= width * 0.5f; fx /= zoom; fx += screen_center_x * width;
	fy -= height * 0.5f; fy /= zoom; fy += screen_center_y * height;
	const float u = (fx - width * 0.5f) * scale + image_res_x * 0.5f;
	const float v = (fy - height * 0.5f) * scale + image_res_y * 0.5f;
	const int srcx = (int)floorf(u), srcy = (int)floorf(v);
	Rgba val{0.0f, 0.0f, 0.0f, 0.0f};
	if (srcx < image_res_x && srcy < image_res_y && srcx >= 0 && srcy >= 0) val = read_rgba(srcx, srcy, image_res_x, image, image_data_type);
	V3 col = {val.r, val.g, val.b};
	float a = val.a;
	V3 bg = {background.x, background.y, background.z};
	if (color_space != kColorSRGB) {
		bg = srgb_to_linear(bg);
	} else {
		col = a > 0 ? linear_to_srgb(col / a) * a : v3(0.0f);
	}
	const float weight = (1 - a) * background.w;
	col = col + bg * weight;
	a += weight;
	if (color_space == kColorSRGB) col = srgb_to_linear(col);
	col = col * powf(2.0f, exposure);
	col = tonemap_curve(col, curve);
	if (output_color_space == kColorSRGB) col = linear_to_srgb(col);
	const float4 prev = alpha < 1.f ? out[(size_t)y * width + x] : make_float4(0.f, 0.f, 0.f, 0.f); // uninitialised memory may hold NaNs
	out[(size_t)y * width + x] = make_float4(col.x * alpha + prev.x * (1.f - alpha), col.y * alpha + prev.y * (1.f - alpha),
		col.z * alpha + prev.z * (1.f - alpha), a * alpha + prev.w * (1.f - alpha));
}

void Testbed::set_render_lens(int lens_mode, const float* params) {
	if (lens_mode != kLensPerspective && lens_mode != kLensOpenCV) throw std::runtime_error{"ngp-b200: lens mode must be Perspective or OpenCV"};
	render_lens_mode = lens_mode;
	for (int i = 0; i < 7; ++i) render_lens_params[i] = params ? params[i] : 0.0f;
}

void Testbed::overlay_ground_truth(uint32_t width, uint32_t height, bool linear, uint32_t view, int fov_axis, float zoom, float alpha, float* rgba_inout) {
	if (m_mode != kModeNerf) throw std::runtime_error{"ngp-b200: the ground-truth overlay is implemented for the NeRF mode"};
	if (view >= m_images.size()) throw std::runtime_error{"ngp-b200: ground-truth view index out of range"};
	if (width == 0 || height == 0 || !rgba_inout) throw std::runtime_error{"ngp-b200: overlay_ground_truth needs a positive resolution and a buffer"};
	if (fov_axis != 0 && fov_axis != 1) throw std::runtime_error{"ngp-b200: fov_axis must be 0 or 1"};
	if (tonemap_curve < kTonemapIdentity || tonemap_curve > kTonemapReinhard) throw std::runtime_error{"Invalid tonemap curve."};
	const ImageMeta& m = m_images[view].meta;
	if (!m.pixels) throw std::runtime_error{"ngp-b200: the view has no image"};
	const float4 bg = make_float4(background_color[0], background_color[1], background_color[2], background_color[3]);
	const dim3 threads = {16, 8, 1}, blocks = {div_round_up(width, threads.x), div_round_up(height, threads.y), 1};
	overlay_image_kernel<<<blocks, threads, 0, m_stream>>>(width, height, alpha, exposure, bg, m.pixels, m.image_data_type, m.res_x, m.res_y, tonemap_curve,
		render_color_space, linear ? kColorLinear : kColorSRGB, fov_axis, zoom, 0.5f, 0.5f, reinterpret_cast<float4*>(rgba_inout));
	NGPB_CUDA_CHECK(cudaGetLastError());
}

// render_to_cpu without the camera-path / motion-blur options: `spp` traced frames with sample
// indices 0..spp-1 averaged, then the epilogue. rgba_out: DEVICE float4 [h][w].
void Testbed::render(uint32_t width, uint32_t height, uint32_t spp, bool linear, const float* camera_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, float* rgba_out)
{
	if (m_mode != kModeNerf) throw std::runtime_error{"ngp-b200: render() is implemented for the NeRF mode; image mode renders with render_image"};
	if (width == 0 || height == 0 || spp == 0) throw std::runtime_error{"ngp-b200: render needs a positive resolution and sample count"};
	if (tonemap_curve < kTonemapIdentity || tonemap_curve > kTonemapReinhard) throw std::runtime_error{"Invalid tonemap curve."};
	const uint32_t n = width * height;
	m_frame_buffer.reserve((size_t)n * sizeof(float4), m_stream);
	m_accumulate_buffer.reserve((size_t)n * sizeof(float4), m_stream);
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_accumulate_buffer.ptr, 0, (size_t)n * sizeof(float4), m_stream)); // reset_accumulation
	for (uint32_t s = 0; s < spp; ++s) {
		render_nerf(width, height, camera_3x4, focal_x, focal_y, principal_x, principal_y, s, 0, n, m_frame_buffer.as<float>(), nullptr);
		accumulate_kernel<<<div_round_up(n, 256), 256, 0, m_stream>>>(n, m_frame_buffer.as<float4>(), m_accumulate_buffer.as<float4>(), (float)s, render_color_space);
	}
	const float4 bg = make_float4(background_color[0], background_color[1], background_color[2], background_color[3]);
	tonemap_kernel<<<div_round_up(n, 256), 256, 0, m_stream>>>(n, exposure, bg, m_accumulate_buffer.as<float4>(), render_color_space,
		linear ? kColorLinear : kColorSRGB, tonemap_curve, false, false, reinterpret_cast<float4*>(rgba_out));
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
