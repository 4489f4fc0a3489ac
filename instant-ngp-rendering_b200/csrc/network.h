// Network modules behind the C ABI: tcnn::NetworkWithInputEncoding
// (tcnn/include/tiny-cuda-nn/network_with_input_encoding.h:45-165) and ngp::NerfNetwork
// (include/neural-graphics-primitives/nerf_network.h:75-505), both on the fused tcgen05 kernels.
#pragma once

#include "grid.h"
#include "mlp.h"
#include "module.h"

namespace ngpb {

// tcnn network JSON (tcnn/src/network.cu:98-136) -> shape. Only FullyFusedMLP-class networks with
// 64 neurons exist here; the otype strings MLP / CutlassMLP are accepted and run on the same
// kernels.
MlpShape parse_mlp(const Json& network, uint32_t in_width, uint32_t out_width);

class NetworkWithInputEncodingModule : public Module {
public:
	NetworkWithInputEncodingModule(uint32_t n_input_dims, uint32_t n_output_dims, const Json& encoding, const Json& network);

	void inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) override;
	std::unique_ptr<Context> forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
		const __half* params, bool prepare_input_gradients) override;
	void backward(cudaStream_t stream, const Context& ctx, uint32_t n, float* dL_dinput, const __half* dL_doutput,
		__half* dL_dparams, const float* input, const __half* output, const __half* params) override;

	uint32_t n_input_dims() const override { return m_n_input_dims; }
	uint32_t n_output_dims() const override { return m_mlp.out_width; }
	uint32_t padded_output_width() const override { return kMlpPaddedOut; }
	size_t n_params() const override { return (size_t)m_mlp.n_params() + (m_has_grid ? m_grid.n_params : 0); }
	std::vector<std::pair<uint32_t, uint32_t>> layer_sizes() const override;
	void initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) override;
	Json hyperparams() const override;
	std::string name() const override { return "NetworkWithInputEncoding"; }

	const GridConfig* grid_config() const { return m_has_grid ? &m_grid : nullptr; }
	size_t encoding_param_offset() const { return m_mlp.n_params(); }
	const FusedNet& net() const { return m_net; }
	// Consecutive samples of a batch are neighbours in space (stratified image positions): grid_backward's `ordered`.
	void set_ordered_batches(bool ordered) { m_ordered_batches = ordered; }

	// features in the tile layout of mlp.h; n is rounded up to whole tiles inside.
	void encode_tiled(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, const __half* params, __half* feat);

	// inference() with the feature-tile scratch supplied by the caller instead of a stream-ordered allocation
	// (capi_static_scratch.cu). scratch must hold inference_scratch_bytes(n) bytes, 256-byte aligned.
	size_t inference_scratch_bytes(uint32_t n) const { return tiled_bytes(n, m_enc_width); }
	void inference_with_scratch(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params, void* scratch, size_t scratch_bytes);

private:
	uint32_t m_n_input_dims;
	bool m_has_grid = false;
	float m_identity_scale = 1.0f, m_identity_offset = 0.0f;   // IdentityEncoding (tcnn encodings/identity.h:60-63): y = x * scale + offset
	bool m_ordered_batches = false;
	GridConfig m_grid;
	uint32_t m_enc_width = 0; // padded to 16
	MlpShape m_mlp;
	FusedNet m_net{};
	Json m_encoding_json, m_network_json;
};

class NerfNetworkModule : public Module {
public:
	NerfNetworkModule(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
		const Json& pos_encoding, const Json& dir_encoding, const Json& density_network, const Json& rgb_network);

	void inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) override;
	std::unique_ptr<Context> forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
		const __half* params, bool prepare_input_gradients) override;
	void backward(cudaStream_t stream, const Context& ctx, uint32_t n, float* dL_dinput, const __half* dL_doutput,
		__half* dL_dparams, const float* input, const __half* output, const __half* params) override;

	uint32_t n_input_dims() const override { return m_dir_offset + m_n_dir_dims + m_n_extra_dims; } // nerf_network.h:381-383
	uint32_t n_output_dims() const override { return 4; }                                            // :385-387
	uint32_t padded_output_width() const override { return kMlpPaddedOut; }                           // :377-379
	size_t n_params() const override { return (size_t)m_density.n_params() + m_rgb.n_params() + m_grid.n_params; }
	std::vector<std::pair<uint32_t, uint32_t>> layer_sizes() const override;
	void initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) override;
	Json hyperparams() const override;
	std::string name() const override { return "NerfNetwork"; }

	// Strided variants used by the NeRF driver: rows of `in_stride` floats, outputs of `out_stride`
	// halves per sample (16 = tcnn layout, 4 = packed rgb sigma).
	void inference_strided(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params);
	// As inference_strided over min(*n_dev, max_n) samples, *n_dev read on the device.
	void inference_counted(cudaStream_t stream, uint32_t max_n, const uint32_t* n_dev, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params);
	// before_mlp / after_mlp (optional): recorded right before the fused MLP kernel is launched (the
	// encoder pass is done) and once it is done, before the grid scatter.
	void backward_strided(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, const __half* dL_doutput, uint32_t dy_stride,
		__half* dL_dparams, const __half* params, cudaEvent_t after_mlp = nullptr, cudaEvent_t before_mlp = nullptr, const uint32_t* reuse_src = nullptr);
	// reuse_src (optional, DEVICE [n]): sample i of this batch is a copy of row reuse_src[i] of the batch the LAST inference_counted
	// call ran on, with the same parameters. Its encoded features are then copied out of that call's feature tiles instead of
	// being gathered from the hash table again (the training step: the compacted batch is a subset of the marched samples).
	// NerfNetwork::density (nerf_network.h:270-279): out[i*out_stride] = density-MLP output 0 when
	// out_stride == 1, else the padded 16 outputs.
	void density(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params);

	const GridConfig& grid_config() const { return m_grid; }
	size_t pos_encoding_param_offset() const { return (size_t)m_density.n_params() + m_rgb.n_params(); } // :356-371
	size_t n_mlp_params() const { return (size_t)m_density.n_params() + m_rgb.n_params(); }
	const FusedNet& net() const { return m_net; }

private:
	uint32_t m_n_pos_dims, m_n_dir_dims, m_n_extra_dims, m_dir_offset;
	GridConfig m_grid;
	uint32_t m_pos_width = 0;
	uint32_t m_sh_degree = 4;
	MlpShape m_density, m_rgb;
	FusedNet m_net{};
	Json m_pos_json, m_dir_json, m_density_json, m_rgb_json;
	DeviceBuffer m_scratch_feat, m_scratch_dfeat, m_scratch_dw, m_scratch_feat_train;
	uint32_t m_infer_tiles_rows = 0;   // rows of the feature tiles inference_counted left in m_scratch_feat (0: none)
};

} // namespace ngpb
