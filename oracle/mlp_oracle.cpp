// CPU oracle for the fused MLP and the NerfNetwork glue. TEST INFRASTRUCTURE ONLY (see oracle.h).
// Restates:
//   kernel_mlp_fused (forward / inference)   tcnn src/fully_fused_mlp.cu:499-557 (+ :82-148, 315-476)
//   kernel_mlp_fused_backward + dW GEMMs      tcnn src/fully_fused_mlp.cu:150-259, 735-836
//   weight layout                             tcnn src/fully_fused_mlp.cu:656-678
//   xavier init                               tcnn include/tiny-cuda-nn/gpu_matrix.h:284-299
//   sh_enc (degree <= 4)                      tcnn include/tiny-cuda-nn/common_device.h:339-365
//   kernel_sh input mapping / padding         tcnn encodings/spherical_harmonics.h:46-72
// Activations are stored as fp16 between layers exactly as the reference stores them. The
// accumulator is fp32 (what the TMEM kernels use) or, with accumulate_fp16 = 1, rounded to fp16
// after every 16-wide K block, which models the reference's wmma fp16 accumulator
// (fully_fused_mlp.cu:66-68) closely enough to bound its error in the tolerance tests.
#include "oracle.h"
#include "pcg32_oracle.h"

#include <cmath>
#include <cstring>
#include <vector>

typedef _Float16 half_t;
static inline half_t bits_to_half(uint16_t b) { half_t h; memcpy(&h, &b, 2); return h; }
static inline uint16_t half_to_bits(half_t h) { uint16_t b; memcpy(&b, &h, 2); return b; }

namespace {

// y[rows] = W[rows x cols] x[cols]; W, x fp16.
inline void matvec(const uint16_t* W, uint32_t rows, uint32_t cols, const half_t* x, float* y, int accumulate_fp16) {
	for (uint32_t r = 0; r < rows; ++r) {
		const uint16_t* w = W + (size_t)r * cols;
		if (accumulate_fp16) {
			half_t acc = (half_t)0.0f;
			for (uint32_t k0 = 0; k0 < cols; k0 += 16) {
				float s = 0.0f;
				for (uint32_t k = k0; k < k0 + 16 && k < cols; ++k) s += (float)bits_to_half(w[k]) * (float)x[k];
				acc = (half_t)((float)acc + s);
			}
			y[r] = (float)acc;
		} else {
			float s = 0.0f;
			for (uint32_t k = 0; k < cols; ++k) s += (float)bits_to_half(w[k]) * (float)x[k];
			y[r] = s;
		}
	}
}

} // namespace

extern "C" size_t oracle_mlp_n_params(uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out) {
	return (size_t)width * in_width + (size_t)(n_hidden - 1) * width * width + (size_t)padded_out * width;
}

extern "C" void oracle_mlp_forward(uint32_t n, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, int relu,
	const uint16_t* weights, const uint16_t* x, uint32_t x_stride, uint16_t* out, uint32_t out_stride, uint16_t* hidden, int accumulate_fp16)
{
#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		std::vector<half_t> cur(in_width), nxt(width);
		std::vector<float> acc(width > padded_out ? width : padded_out);
		for (uint32_t k = 0; k < in_width; ++k) cur[k] = bits_to_half(x[(size_t)i * x_stride + k]);
		const uint16_t* W = weights;
		uint32_t cols = in_width;
		for (uint32_t l = 0; l < n_hidden; ++l) {
			matvec(W, width, cols, cur.data(), acc.data(), accumulate_fp16);
			for (uint32_t r = 0; r < width; ++r) {
				float v = acc[r];
				if (relu) v = v > 0.0f ? v : 0.0f;
				nxt[r] = (half_t)v;
				if (hidden) hidden[((size_t)l * n + i) * width + r] = half_to_bits(nxt[r]);
			}
			W += (size_t)width * cols;
			cols = width;
			cur.assign(nxt.begin(), nxt.end());
		}
		matvec(W, padded_out, width, cur.data(), acc.data(), accumulate_fp16);
		for (uint32_t r = 0; r < padded_out; ++r) out[(size_t)i * out_stride + r] = half_to_bits((half_t)acc[r]);
	}
}

// Backward of the chain above. `hidden` are the forward activations [n_hidden][n][width] (fp16).
// dL_dx (optional): fp16 [n][in_width]. dW: double [n_params], summed over the batch from fp16
// operands (the implementations sum in fp32 tensor-core accumulators or fp16 split-K). Gradients
// are rounded to fp16 at every layer boundary, where both implementations store them as fp16.
extern "C" void oracle_mlp_backward(uint32_t n, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, int relu,
	const uint16_t* weights, const uint16_t* x, uint32_t x_stride, const uint16_t* hidden,
	const uint16_t* dL_dout, uint32_t dy_stride, uint16_t* dL_dx, uint32_t dx_stride, double* dW)
{
	const size_t n_params = oracle_mlp_n_params(in_width, width, n_hidden, padded_out);
	for (size_t k = 0; k < n_params; ++k) dW[k] = 0.0;
	std::vector<size_t> w_off(n_hidden + 1);
	{
		size_t o = 0;
		w_off[0] = 0; o += (size_t)width * in_width;
		for (uint32_t l = 1; l < n_hidden; ++l) { w_off[l] = o; o += (size_t)width * width; }
		w_off[n_hidden] = o;
	}
	for (uint32_t i = 0; i < n; ++i) {
		std::vector<half_t> g(padded_out), d(width), dprev(width);
		for (uint32_t r = 0; r < padded_out; ++r) g[r] = bits_to_half(dL_dout[(size_t)i * dy_stride + r]);
		// output layer
		const uint16_t* hl = hidden + ((size_t)(n_hidden - 1) * n + i) * width;
		for (uint32_t r = 0; r < padded_out; ++r)
			for (uint32_t k = 0; k < width; ++k) dW[w_off[n_hidden] + (size_t)r * width + k] += (double)g[r] * (double)bits_to_half(hl[k]);
		for (uint32_t k = 0; k < width; ++k) {
			float s = 0.0f;
			for (uint32_t r = 0; r < padded_out; ++r) s += (float)g[r] * (float)bits_to_half(weights[w_off[n_hidden] + (size_t)r * width + k]);
			if (relu && !((float)bits_to_half(hl[k]) > 0.0f)) s = 0.0f;
			d[k] = (half_t)s;
		}
		for (uint32_t l = n_hidden - 1; l >= 1; --l) {
			const uint16_t* hprev = hidden + ((size_t)(l - 1) * n + i) * width;
			for (uint32_t r = 0; r < width; ++r)
				for (uint32_t k = 0; k < width; ++k) dW[w_off[l] + (size_t)r * width + k] += (double)d[r] * (double)bits_to_half(hprev[k]);
			for (uint32_t k = 0; k < width; ++k) {
				float s = 0.0f;
				for (uint32_t r = 0; r < width; ++r) s += (float)d[r] * (float)bits_to_half(weights[w_off[l] + (size_t)r * width + k]);
				if (relu && !((float)bits_to_half(hprev[k]) > 0.0f)) s = 0.0f;
				dprev[k] = (half_t)s;
			}
			d.assign(dprev.begin(), dprev.end());
		}
		for (uint32_t r = 0; r < width; ++r)
			for (uint32_t k = 0; k < in_width; ++k) dW[w_off[0] + (size_t)r * in_width + k] += (double)d[r] * (double)bits_to_half(x[(size_t)i * x_stride + k]);
		if (dL_dx) {
			for (uint32_t k = 0; k < in_width; ++k) {
				float s = 0.0f;
				for (uint32_t r = 0; r < width; ++r) s += (float)d[r] * (float)bits_to_half(weights[w_off[0] + (size_t)r * in_width + k]);
				dL_dx[(size_t)i * dx_stride + k] = half_to_bits((half_t)s);
			}
		}
	}
}

// Xavier-uniform init of one MLP with a pcg32 stream, as the reference draws it on the host.
// `rng_state` = {state, inc}, advanced in place.
extern "C" void oracle_mlp_init(uint64_t* rng_state, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, float scale, float* params)
{
	OraclePcg32 rng;
	rng.state = rng_state[0];
	rng.inc = rng_state[1];
	size_t pos = 0;
	auto fill = [&](uint32_t rows, uint32_t cols) {
		const float sc = scale * std::sqrt(6.0f / (float)(rows + cols));
		for (uint32_t k = 0; k < rows * cols; ++k) params[pos++] = rng.next_float() * 2.0f * sc - sc;
	};
	fill(width, in_width);
	for (uint32_t l = 1; l < n_hidden; ++l) fill(width, width);
	fill(padded_out, width);
	rng_state[0] = rng.state;
	rng_state[1] = rng.inc;
}

// kernel_sh for degree <= 4: dirs are [0,1]^3-encoded, mapped to [-1,1]; output fp16
// [n][out_stride] with (16 - degree^2) leading ones.
extern "C" void oracle_sh_encode(uint32_t n, uint32_t degree, const float* dirs, uint32_t dir_stride, uint16_t* out, uint32_t out_stride)
{
	const uint32_t n_sh = degree * degree;
	const uint32_t pad = 16 - n_sh;
	for (uint32_t i = 0; i < n; ++i) {
		const float x = dirs[(size_t)i * dir_stride + 0] * 2.f - 1.f;
		const float y = dirs[(size_t)i * dir_stride + 1] * 2.f - 1.f;
		const float z = dirs[(size_t)i * dir_stride + 2] * 2.f - 1.f;
		const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
		float s[16];
		s[0] = 0.28209479177387814f;
		s[1] = -0.48860251190291987f * y;
		s[2] = 0.48860251190291987f * z;
		s[3] = -0.48860251190291987f * x;
		s[4] = 1.0925484305920792f * xy;
		s[5] = -1.0925484305920792f * yz;
		s[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
		s[7] = -1.0925484305920792f * xz;
		s[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
		s[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
		s[10] = 2.8906114426405538f * xy * z;
		s[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
		s[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
		s[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
		s[14] = 1.4453057213202769f * z * (x2 - y2);
		s[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
		uint16_t* o = out + (size_t)i * out_stride;
		for (uint32_t k = 0; k < pad; ++k) o[k] = half_to_bits((half_t)1.0f);
		for (uint32_t k = 0; k < n_sh; ++k) o[pad + k] = half_to_bits((half_t)s[k]);
	}
}

// pcg32(seed) -> {state, inc}; advance by `delta` draws.
extern "C" void oracle_pcg32_seed(uint64_t seed, uint64_t* rng_state) {
	OraclePcg32 rng(seed);
	rng_state[0] = rng.state;
	rng_state[1] = rng.inc;
}
extern "C" void oracle_pcg32_advance(uint64_t* rng_state, int64_t delta) {
	OraclePcg32 rng;
	rng.state = rng_state[0];
	rng.inc = rng_state[1];
	rng.advance(delta);
	rng_state[0] = rng.state;
}
