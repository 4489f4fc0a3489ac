// Encoding modules behind ngpb_create_encoding (tcnn/src/encoding.cu:54-141).
#include "encoding.h"

namespace ngpb {

static bool device_available() {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
	return n > 0;
}

GridEncodingModule::GridEncodingModule(uint32_t n_dims, const Json& encoding)
: m_cfg{make_grid_config(n_dims, encoding, device_available())} {}

void GridEncodingModule::inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) {
	grid_forward(stream, m_cfg.p, n, params, input, n_input_dims(), output, padded_output_width(), m_max_level);
}

std::unique_ptr<Context> GridEncodingModule::forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
	const __half* params, bool /*prepare_input_gradients*/)
{
	// Nothing is kept: the backward pass recomputes cell coordinates from `input`
	// (the reference keeps positions + dy_dx, grid.h:1115-1118).
	if (output) grid_forward(stream, m_cfg.p, n, params, input, n_input_dims(), output, padded_output_width(), m_max_level);
	return std::make_unique<Context>();
}

void GridEncodingModule::backward(cudaStream_t stream, const Context&, uint32_t n, float* dL_dinput, const __half* dL_doutput,
	__half* dL_dparams, const float* input, const __half* /*output*/, const __half* params)
{
	if (dL_dparams) {
		// GradientMode::Overwrite (cpp_api.cu:103): zero, then scatter (grid.h:857-880).
		NGPB_CUDA_CHECK(cudaMemsetAsync(dL_dparams, 0, n_params() * sizeof(__half), stream));
		grid_backward(stream, m_cfg.p, n, dL_dparams, input, n_input_dims(), dL_doutput, padded_output_width(), m_max_level);
	}
	if (dL_dinput) {
		grid_backward_input(stream, m_cfg.p, n, params, input, n_input_dims(), dL_doutput, padded_output_width(),
			dL_dinput, n_input_dims(), m_max_level);
	}
}

void GridEncodingModule::initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) {
	random_uniform(stream, rng, n_params(), params_fp32, -1e-4f * scale, 1e-4f * scale);
}

Module* create_encoding_module(uint32_t n_input_dims, const Json& encoding) {
	const std::string otype = encoding.value("otype", "OneBlob");
	if (equals_case_insensitive(otype, "Grid") || equals_case_insensitive(otype, "HashGrid") ||
		equals_case_insensitive(otype, "TiledGrid") || equals_case_insensitive(otype, "DenseGrid")) {
		return new GridEncodingModule{n_input_dims, encoding};
	}
	throw std::runtime_error{"Invalid encoding type: " + otype};
}

} // namespace ngpb
