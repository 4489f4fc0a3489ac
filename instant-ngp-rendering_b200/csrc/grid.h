// Multiresolution grid encoding: configuration, level table and launchers.
//
// Replaces tcnn::GridEncodingTemplated (tcnn/include/tiny-cuda-nn/encodings/grid.h:652-1141)
// and its JSON factory (grid.h:1143-1208). The JSON keys, defaults and error texts follow
// the reference; the data layout and kernels do not (see grid.cu).
#pragma once

#include "common.cuh"
#include "json.h"

namespace ngpb {

enum class GridType : int { Hash = 0, Dense = 1, Tiled = 2 };
enum class HashType : int { Prime = 0, CoherentPrime = 1, ReversedPrime = 2, Rng = 3 };
enum class InterpolationType : int { Nearest = 0, Linear = 1, Smoothstep = 2 };

constexpr uint32_t kMaxGridLevels = 32; // tcnn allows 128 (grid_interface.h:84); no shipped config exceeds 16
constexpr uint32_t kMaxGridDims = 4;

// POD handed to the kernels by value (__grid_constant__). Everything that
// tcnn recomputes per thread but that only depends on the level is tabulated here.
struct GridParams {
	uint32_t n_dims;
	uint32_t n_features_per_level;
	uint32_t n_levels;
	uint32_t base_resolution;
	float log2_per_level_scale;
	int grid_type;
	int hash_type;
	int interpolation_type;
	uint32_t stochastic_interpolation;
	uint32_t primes[kMaxGridDims];
	// per level
	uint32_t offset[kMaxGridLevels + 1];     // entry offsets (grid.h:692-723)
	float scale[kMaxGridLevels];             // grid_scale(level) (common_device.h:709-714)
	uint32_t resolution[kMaxGridLevels];     // grid_resolution(scale) (common_device.h:716-718)
	uint32_t size[kMaxGridLevels];           // entries in the level ("hashmap_size" in kernel_grid)
	uint32_t stride[kMaxGridLevels][kMaxGridDims]; // dense strides; 0 for dims the reference's index loop never reaches
	uint32_t use_hash[kMaxGridLevels];       // grid_index() falls through to the hash (common_device.h:702-704)
	uint32_t size_is_pow2[kMaxGridLevels];
};

struct GridConfig {
	GridParams p{};
	float per_level_scale = 2.0f;
	uint32_t log2_hashmap_size = 19;
	uint32_t n_params = 0; // entries * F

	uint32_t n_output_dims() const { return p.n_levels * p.n_features_per_level; }
	uint32_t n_entries() const { return p.offset[p.n_levels]; }
	Json hyperparams() const;
};

// Builds the configuration from a tcnn encoding JSON object ("otype": Grid | HashGrid |
// DenseGrid | TiledGrid). Throws std::runtime_error with the reference's messages.
// `device_scales`: when a CUDA device is present the per-level scale is evaluated on the
// device (device exp2f, as kernel_grid does at grid.h:97) so that cell boundaries fall where
// the reference's do; without a device the host exp2f value is used.
GridConfig make_grid_config(uint32_t n_dims_to_encode, const Json& encoding, bool device_scales);

std::string to_string(GridType t);
std::string to_string(HashType t);
std::string to_string(InterpolationType t);

// ---- launchers (grid.cu) -----------------------------------------------------------------
// positions: fp32, sample i at positions[i*in_stride + d].
// out:       fp16 AoS, sample i at out[i*out_stride + level*F + f]; columns
//            [L*F, out_stride) are set to `pad_value` (the reference pads with 1,
//            grid.h:806-811 via Encoding::set_padded_output_width).
// max_level: as tcnn's m_max_level (1.0 == all levels).
// tiled:     out / dL_dy use the tile layout of mlp.h ([tile][chunk][128][8 halves], out_stride
//            = padded feature width) instead of AoS rows; rows up to the next multiple of 128
//            are written (zeros) so that a consumer may read whole tiles.
// n_dev:     optional device-side count; min(*n_dev, n) samples are encoded.
void grid_forward(cudaStream_t stream, const GridParams& p, uint32_t n, const __half* grid,
	const float* positions, uint32_t in_stride, __half* out, uint32_t out_stride, float max_level, bool tiled = false, const uint32_t* n_dev = nullptr);

// dL_dy: fp16 AoS with the same layout as `out`. grad: fp16 [entries][F], must be zeroed by
// the caller when overwrite semantics are wanted (the reference zeroes at grid.h:857-859;
// ngpb_module_backward does the memset).
// ordered: consecutive samples are neighbours in space (ray-ordered NeRF batches): the warp-aggregated scatter alone is
// used. Otherwise (uniform / stratified points) the coarsest dense levels are accumulated in shared memory first.
void grid_backward(cudaStream_t stream, const GridParams& p, uint32_t n, __half* grad,
	const float* positions, uint32_t in_stride, const __half* dL_dy, uint32_t dy_stride, float max_level, bool tiled = false, bool ordered = false);

// dL_dx[i][d] = sum_k dL_dy[i][k] * dy_k/dx_d, recomputed from the table (the reference
// stores dy_dx in the forward pass and contracts in kernel_grid_backward_input, grid.h:322-349).
void grid_backward_input(cudaStream_t stream, const GridParams& p, uint32_t n, const __half* grid,
	const float* positions, uint32_t in_stride, const __half* dL_dy, uint32_t dy_stride,
	float* dL_dx, uint32_t dx_stride, float max_level);

// Integer view for parity tests: idx[(i*L + level)*2^D + corner] = entry index within the level.
void grid_indices(cudaStream_t stream, const GridParams& p, uint32_t n, const float* positions,
	uint32_t in_stride, uint32_t* idx);

} // namespace ngpb
