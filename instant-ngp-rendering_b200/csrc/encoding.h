// Encoding modules (tcnn Encoding<T> subclasses that the shipped NeRF / image configs use).
#pragma once

#include "grid.h"
#include "module.h"

namespace ngpb {

class GridEncodingModule : public Module {
public:
	GridEncodingModule(uint32_t n_dims, const Json& encoding);

	void inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) override;
	std::unique_ptr<Context> forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
		const __half* params, bool prepare_input_gradients) override;
	void backward(cudaStream_t stream, const Context& ctx, uint32_t n, float* dL_dinput, const __half* dL_doutput,
		__half* dL_dparams, const float* input, const __half* output, const __half* params) override;

	uint32_t n_input_dims() const override { return m_cfg.p.n_dims; }
	uint32_t n_output_dims() const override { return m_cfg.n_output_dims(); }
	uint32_t padded_output_width() const override { return m_cfg.n_output_dims(); }
	size_t n_params() const override { return m_cfg.n_params; }
	void initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) override;
	Json hyperparams() const override { return m_cfg.hyperparams(); }
	std::string name() const override { return "Grid"; }   // GridEncodingTemplated::name (grid.h)

	const GridConfig& config() const { return m_cfg; }
	void set_max_level(float v) { m_max_level = v; }
	float max_level() const { return m_max_level; }

private:
	GridConfig m_cfg;
	float m_max_level = 1.0f;
};

} // namespace ngpb
