// Fused MLP chains on tcgen05 / TMEM.
//
// Replaces tcnn::FullyFusedMLP (tcnn/src/fully_fused_mlp.cu:499-836, wmma HMMA) and, for the
// NeRF case, the whole of ngp::NerfNetwork's wiring (nerf_network.h:105-268) including its
// three glue kernels. Weight layout in the flat parameter buffer is the reference's
// (fully_fused_mlp.cu:656-678): W0 [64 x in] | hidden [64 x 64]... | last [padded_out x 64],
// row-major fp16, so snapshots and initialisers stay interchangeable.
//
// Feature tiles ("tiled" layout): the encoders hand their output to the MLP kernels as
//   feat[tile][chunk][row 0..127][8 halves]      tile = sample / 128, chunk = column / 8
// which is byte for byte the shared-memory image the UMMA descriptors in tc_device.cuh read, so
// a tile arrives with one 1-D bulk copy and the producing kernel's stores are fully coalesced
// (consecutive samples write consecutive 16-byte chunks).
#pragma once

#include "common.cuh"

namespace ngpb {

constexpr uint32_t kMlpWidth = 64;      // n_neurons of every shipped NeRF / image config
constexpr uint32_t kMlpPaddedOut = 16;  // FullyFusedMLP::REQUIRED_ALIGNMENT (fully_fused_mlp.h:108-110)
constexpr uint32_t kTile = 128;         // samples per MMA tile (UMMA M)

struct MlpShape {
	uint32_t in_width = 0;    // multiple of 16, <= 64
	uint32_t n_hidden = 1;    // n_hidden_layers >= 1
	uint32_t out_width = 0;   // unpadded
	uint32_t relu = 1;        // hidden activation: 1 ReLU, 0 None
	uint32_t n_params() const { return kMlpWidth * in_width + (n_hidden - 1) * kMlpWidth * kMlpWidth + kMlpPaddedOut * kMlpWidth; }
};

// POD describing one or two chained MLPs to the kernels.
struct FusedNet {
	uint32_t nerf;            // 0: chain A only. 1: A (density) -> [A_out(16) | SH(dir)(16)] -> B (rgb)
	uint32_t a_in, a_hidden, a_relu;
	uint32_t b_in, b_hidden, b_relu;
	uint32_t a_param_off;     // offsets in halves into the params buffer
	uint32_t b_param_off;
	uint32_t sh_degree;       // 1..4
	uint32_t dir_offset;      // column of the first direction component in the coords rows
};

NGPB_HD uint32_t chain_params(uint32_t in_width, uint32_t n_hidden) {
	return kMlpWidth * in_width + (n_hidden - 1) * kMlpWidth * kMlpWidth + kMlpPaddedOut * kMlpWidth;
}
NGPB_HD uint32_t fused_net_mlp_params(const FusedNet& n) {
	return chain_params(n.a_in, n.a_hidden) + (n.nerf ? chain_params(n.b_in, n.b_hidden) : 0);
}

// Bytes of one feature tile of `width` columns.
inline size_t tiled_bytes(uint32_t n, uint32_t width) { return (size_t)div_round_up(n, kTile) * kTile * width * sizeof(__half); }

struct FusedInferenceArgs {
	FusedNet net;
	uint32_t n;                 // samples, multiple of 128
	const uint32_t* n_dev;      // optional device-side sample count: min(*n_dev, n) samples are processed
	const __half* params;       // flat parameter buffer (chain offsets inside `net`)
	const __half* feat;         // tiled chain-A input, a_in columns
	const float* coords;        // nerf: rows with the view direction at [dir_offset, dir_offset+3)
	uint32_t coord_stride;      // floats per coords row
	__half* out;                // [n][out_stride]
	uint32_t out_stride;        // 16 (tcnn padded layout) or 4 (nerf: rgb sigma packed) or 1 (chain A column 0 only)
};

struct FusedTrainArgs {
	FusedNet net;
	uint32_t n;
	const __half* params;
	const __half* feat;
	const float* coords;
	uint32_t coord_stride;
	const __half* dL_dout;      // [n][dy_stride]; nerf: columns 0-2 rgb, 3 sigma (nerf_network.h:45-74)
	uint32_t dy_stride;
	__half* dfeat;              // tiled dL/d(chain-A input), a_in columns; may be null
	float* dw_partials;         // [grid][mlp params] fp32 scratch
	__half* dL_dparams;         // MLP part of the gradient buffer, overwritten (sum of the partials)
};

// Launchers. Throw std::runtime_error on unsupported shapes.
void fused_inference(cudaStream_t stream, const FusedInferenceArgs& args);
// Returns nothing; scratch for the partials is taken from `dw_partials` which must hold
// fused_train_grid() * fused_net_mlp_params(net) floats.
void fused_train(cudaStream_t stream, const FusedTrainArgs& args);
uint32_t fused_train_grid(const FusedNet& net, uint32_t n);
void fused_check_supported(const FusedNet& net);

// fp32 rows [n][in_stride] -> tiled fp16 features of `width` columns: first n_dims columns
// are x*scale+offset, the rest 1 (tcnn identity.h:46-67).
void identity_encode_tiled(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, uint32_t n_dims,
	float scale, float offset, __half* feat, uint32_t width);
// tiled fp16 [n][width] <-> AoS fp16 [n][stride]
void tiled_to_aos(cudaStream_t stream, uint32_t n, const __half* tiled, uint32_t width, __half* aos, uint32_t stride);
void aos_to_tiled(cudaStream_t stream, uint32_t n, const __half* aos, uint32_t stride, __half* tiled, uint32_t width);

} // namespace ngpb
