// Hash-grid encoder + fused MLP chain(s) + (training) gradient scatter in ONE persistent, warp-specialised
// kernel per pass: "tcgen05/TMEM tiles fed directly from the encoding output".
//
// Replaces, for the shipped shapes, the reference's seam between GridEncoding and FullyFusedMLP
// (tcnn network_with_input_encoding.h:60-110; ngp nerf_network.h:105-268), where the encoded features,
// every hidden activation and dL/d(features) travel through global memory between five to eight kernels.
// Here a 128-sample tile is gathered straight into the shared-memory operand image of the first UMMA,
// runs forward (and backward) on chip, and in training its dL/d(features) rows leave TMEM only as
// warp-aggregated red.global.add.f16x2 into the gradient table. Nothing per-sample is written except
// the 8-byte network output (inference).
#pragma once

#include "grid.h"
#include "mlp.h"

namespace ngpb {

struct FusedGridNetArgs {
	FusedNet net;
	GridParams grid;            // encoding in front of chain A (Linear interpolation, all levels)
	uint32_t n;                 // capacity in samples, multiple of 128
	const uint32_t* n_dev;      // optional device-side count: min(*n_dev, n) samples are processed
	const __half* params;       // MLP parameters (chain offsets inside `net`)
	const __half* grid_params;  // table of the encoding
	const float* input;         // rows of in_stride floats: position first; nerf: direction at net.dir_offset
	uint32_t in_stride;
	// inference
	__half* out;                // [n][out_stride], out_stride 16 | 4 (nerf: rgb sigma) | 1 (column 0)
	uint32_t out_stride;
	// training
	const __half* dL_dout;      // [n][dy_stride]
	uint32_t dy_stride;
	__half* grid_grad;          // gradient table, zeroed by the caller; null: no scatter
	float* dw_partials;         // [fused_grid_train_ctas()][mlp params] fp32 scratch
	__half* dL_dparams;         // MLP part of the gradients, overwritten
	// Tile-fed mode: chain-A input tiles made by grid_forward(..., tiled = true) instead of the in-kernel gather,
	// and (training) dL/d(chain-A input) written as tiles for grid_backward(..., tiled = true) instead of the
	// in-kernel scatter. The gather and the scatter are L2-latency kernels that want 2048 threads per SM; this
	// kernel keeps 544.
	const __half* feat_tiles;
	__half* dfeat_tiles;
};

// Whether (net, grid) is one of the shapes the fused kernels are built for; otherwise the callers use the
// separate grid_forward / fused_inference / fused_train / grid_backward kernels.
bool fused_grid_supported(const FusedNet& net, const GridParams& grid);
void fused_grid_inference(cudaStream_t stream, const FusedGridNetArgs& args);
void fused_grid_train(cudaStream_t stream, const FusedGridNetArgs& args);
uint32_t fused_grid_train_ctas(const FusedNet& net, uint32_t n);
// Default: tile-fed (gather / scatter in their own full-occupancy kernels). NGPB_FUSED_GATHER=1: in-kernel.
bool fused_grid_gather_in_kernel();
// Whether the training pass uses this kernel too (default: no, see fused.cu; NGPB_FUSED_TRAIN=1).
bool fused_grid_train_enabled();

} // namespace ngpb
