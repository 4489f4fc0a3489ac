// Standalone hash-grid encoding kernels (forward, parameter-gradient scatter, input gradient)
// and the host-side level table. Replaces tcnn kernel_grid / kernel_grid_backward /
// kernel_grid_backward_input (tcnn/include/tiny-cuda-nn/encodings/grid.h:48-349).
//
// Layout choice (differs from the reference on purpose): one thread owns one sample and walks
// all levels, producing the sample's L*F features as contiguous fp16 (AoS). On B200 the whole
// table (24-26 MB for the shipped NeRF configs) sits in the 126 MB L2, so the reference's
// one-level-per-blockIdx.y launch, which exists to keep a single level cache-resident on small
// L2s (grid.h:761-766), buys nothing; walking levels per thread instead makes the output row
// directly usable as an MMA A-operand row by the fused MLP kernel and removes the SoA<->AoS
// transposes (grid_interface.h:41-83).
#include "grid_device.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <limits>

namespace ngpb {

std::string to_string(GridType t) {
	switch (t) {
		case GridType::Hash: return "Hash";
		case GridType::Dense: return "Dense";
		case GridType::Tiled: return "Tiled";
	}
	throw std::runtime_error{"Invalid grid type."};
}
std::string to_string(HashType t) {
	switch (t) {
		case HashType::Prime: return "Prime";
		case HashType::CoherentPrime: return "CoherentPrime";
		case HashType::ReversedPrime: return "ReversedPrime";
		case HashType::Rng: return "Rng";
	}
	throw std::runtime_error{"Invalid hash type."};
}
std::string to_string(InterpolationType t) {
	switch (t) {
		case InterpolationType::Nearest: return "Nearest";
		case InterpolationType::Linear: return "Linear";
		case InterpolationType::Smoothstep: return "Smoothstep";
	}
	throw std::runtime_error{"Invalid interpolation type."};
}

static GridType string_to_grid_type(const std::string& s) {
	if (equals_case_insensitive(s, "Hash")) return GridType::Hash;
	if (equals_case_insensitive(s, "Dense")) return GridType::Dense;
	if (equals_case_insensitive(s, "Tiled") || equals_case_insensitive(s, "Tile")) return GridType::Tiled;
	throw std::runtime_error{"Invalid grid type: " + s};
}
static HashType string_to_hash_type(const std::string& s) {
	if (equals_case_insensitive(s, "Prime")) return HashType::Prime;
	if (equals_case_insensitive(s, "CoherentPrime")) return HashType::CoherentPrime;
	if (equals_case_insensitive(s, "ReversedPrime")) return HashType::ReversedPrime;
	if (equals_case_insensitive(s, "Rng")) return HashType::Rng;
	throw std::runtime_error{"Invalid hash type: " + s};
}
static InterpolationType string_to_interpolation_type(const std::string& s) {
	if (equals_case_insensitive(s, "Nearest")) return InterpolationType::Nearest;
	if (equals_case_insensitive(s, "Linear")) return InterpolationType::Linear;
	if (equals_case_insensitive(s, "Smoothstep")) return InterpolationType::Smoothstep;
	throw std::runtime_error{"Invalid interpolation type: " + s};
}

static uint32_t powi_u32(uint32_t base, uint32_t exponent) {
	uint32_t r = 1;
	for (uint32_t i = 0; i < exponent; ++i) r *= base;
	return r;
}

__global__ void level_scale_kernel(uint32_t n_levels, float log2_per_level_scale, uint32_t base_resolution, float* out) {
	const uint32_t l = threadIdx.x;
	// Same expression, on the same unit, as grid_scale() inside the reference's kernels.
	if (l < n_levels) out[l] = exp2f(l * log2_per_level_scale) * base_resolution - 1.0f;
}

GridConfig make_grid_config(uint32_t n_dims, const Json& enc, bool device_scales) {
	GridConfig cfg;
	GridParams& p = cfg.p;

	const HashType hash_type = string_to_hash_type(enc.value("hash", "CoherentPrime"));
	const uint32_t F = enc.value("n_features_per_level", 2u);
	if (F != 1 && F != 2 && F != 4 && F != 8) throw std::runtime_error{"GridEncoding: n_features_per_level must be 1, 2, 4, or 8."};

	const uint32_t log2_hashmap_size = enc.value("log2_hashmap_size", 19u);
	const std::string otype = enc.value("otype", "Grid");
	const std::string default_type = equals_case_insensitive(otype, "TiledGrid") ? "Tiled" : (equals_case_insensitive(otype, "DenseGrid") ? "Dense" : "Hash");

	uint32_t n_features;
	if (enc.contains("n_features") || enc.contains("n_grid_features")) {
		n_features = enc.contains("n_features") ? enc["n_features"].as_uint() : enc["n_grid_features"].as_uint();
		if (enc.contains("n_levels")) {
			throw std::runtime_error{"GridEncoding: may not specify n_features and n_levels simultaneously (one determines the other)"};
		}
	} else {
		n_features = F * enc.value("n_levels", 16u);
	}
	const uint32_t n_levels = (n_features + F - 1) / F;
	const GridType grid_type = string_to_grid_type(enc.value("type", default_type));
	const uint32_t base_resolution = enc.value("base_resolution", 16u);
	const float per_level_scale = enc.value("per_level_scale",
		grid_type == GridType::Dense ? std::exp(std::log(256.0f / (float)base_resolution) / (n_levels - 1)) : 2.0f);

	if (n_dims < 2 || n_dims > 4) throw std::runtime_error{"GridEncoding: number of input dims must be 2 or 3."};
	if (n_levels > kMaxGridLevels) {
		throw std::runtime_error{"GridEncoding: m_n_levels=" + std::to_string(n_levels) + " must be at most MAX_N_LEVELS=" + std::to_string(kMaxGridLevels)};
	}
	if (n_features % F != 0) {
		throw std::runtime_error{"GridEncoding: n_features=" + std::to_string(n_features) + " must be a multiple of N_FEATURES_PER_LEVEL=" + std::to_string(F)};
	}
	if (hash_type == HashType::Rng) throw std::runtime_error{"GridEncoding: hash type Rng is not available in ngp-b200."};

	p.n_dims = n_dims;
	p.n_features_per_level = F;
	p.n_levels = n_levels;
	p.base_resolution = base_resolution;
	p.log2_per_level_scale = std::log2(per_level_scale);
	p.grid_type = (int)grid_type;
	p.hash_type = (int)hash_type;
	p.interpolation_type = (int)string_to_interpolation_type(enc.value("interpolation", "Linear"));
	p.stochastic_interpolation = enc.value("stochastic_interpolation", false) ? 1u : 0u;
	if (p.stochastic_interpolation) throw std::runtime_error{"GridEncoding: stochastic_interpolation is not available in ngp-b200."};
	cfg.per_level_scale = per_level_scale;
	cfg.log2_hashmap_size = log2_hashmap_size;

	static const uint32_t prime[7] = {1958374283u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
	for (uint32_t d = 0; d < kMaxGridDims; ++d) {
		switch (hash_type) {
			case HashType::Prime: p.primes[d] = prime[d]; break;
			case HashType::CoherentPrime: p.primes[d] = d == 0 ? 1u : prime[d]; break;
			case HashType::ReversedPrime: p.primes[d] = prime[6 - d]; break;
			default: p.primes[d] = 0;
		}
	}

	// Per-level scale. The offsets below use the host value, exactly as the reference's
	// constructor does (grid.h:694); the kernels use the device value, as kernel_grid does.
	float host_scale[kMaxGridLevels];
	for (uint32_t l = 0; l < n_levels; ++l) host_scale[l] = exp2f(l * p.log2_per_level_scale) * base_resolution - 1.0f;
	for (uint32_t l = 0; l < n_levels; ++l) p.scale[l] = host_scale[l];
	if (device_scales) {
		float* d_scale = nullptr;
		NGPB_CUDA_CHECK(cudaMalloc(&d_scale, sizeof(float) * kMaxGridLevels));
		level_scale_kernel<<<1, kMaxGridLevels>>>(n_levels, p.log2_per_level_scale, base_resolution, d_scale);
		NGPB_CUDA_CHECK(cudaMemcpy(p.scale, d_scale, sizeof(float) * n_levels, cudaMemcpyDeviceToHost));
		NGPB_CUDA_CHECK(cudaFree(d_scale));
	}

	uint32_t offset = 0;
	for (uint32_t l = 0; l < n_levels; ++l) {
		const uint32_t res_host = (uint32_t)ceilf(host_scale[l]) + 1;
		const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
		uint32_t params_in_level = std::pow((float)res_host, (float)n_dims) > (float)max_params ? max_params : powi_u32(res_host, n_dims);
		params_in_level = next_multiple(params_in_level, 8u);
		if (grid_type == GridType::Tiled) params_in_level = std::min(params_in_level, powi_u32(base_resolution, n_dims));
		else if (grid_type == GridType::Hash) params_in_level = std::min(params_in_level, 1u << log2_hashmap_size);

		p.offset[l] = offset;
		p.size[l] = params_in_level;
		p.size_is_pow2[l] = (params_in_level & (params_in_level - 1)) == 0 ? 1u : 0u;
		offset += params_in_level;

		// What the kernels see: resolution from the (device) scale, and grid_index()'s loop
		// `for (dim < D && stride <= hashmap_size)`, unrolled into a stride table.
		const uint32_t res = (uint32_t)ceilf(p.scale[l]) + 1;
		p.resolution[l] = res;
		uint32_t stride = 1;
		for (uint32_t d = 0; d < kMaxGridDims; ++d) p.stride[l][d] = 0;
		for (uint32_t d = 0; d < n_dims && stride <= params_in_level; ++d) {
			p.stride[l][d] = stride;
			stride *= res;
		}
		p.use_hash[l] = (grid_type == GridType::Hash && params_in_level < stride) ? 1u : 0u;
	}
	p.offset[n_levels] = offset;
	cfg.n_params = offset * F;
	return cfg;
}

Json GridConfig::hyperparams() const {
	Json j = Json::object();
	j["otype"] = "Grid";
	j["type"] = to_string((GridType)p.grid_type);
	j["n_levels"] = p.n_levels;
	j["n_features_per_level"] = p.n_features_per_level;
	j["base_resolution"] = p.base_resolution;
	j["per_level_scale"] = per_level_scale;
	j["interpolation"] = to_string((InterpolationType)p.interpolation_type);
	j["hash"] = to_string((HashType)p.hash_type);
	if ((GridType)p.grid_type == GridType::Hash) j["log2_hashmap_size"] = log2_hashmap_size;
	return j;
}

// =================================== kernels =============================================

NGPB_D uint32_t next_multiple_d(uint32_t a, uint32_t b) { return (a + b - 1) / b * b; }

template <uint32_t D>
NGPB_D void load_position(const float* __restrict__ positions, uint32_t in_stride, uint32_t i, float (&x)[D]) {
	const float* src = positions + (size_t)i * in_stride;
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) x[d] = __ldg(src + d);
}

// Forward: thread per sample, 8/F levels per 16-byte output store.
template <uint32_t D, uint32_t F, bool VEC_OUT>
__global__ void __launch_bounds__(256) grid_forward_kernel(
	const __grid_constant__ GridParams p, const uint32_t n_max,
	const __half* __restrict__ grid, const float* __restrict__ positions, const uint32_t in_stride,
	__half* __restrict__ out, const uint32_t out_stride, const float max_level, const uint32_t tiled, const uint32_t* __restrict__ n_dev)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t n = n_dev ? min(__ldg(n_dev), n_max) : n_max;
	constexpr uint32_t C = 8 / F; // levels per 16 B
	const uint32_t L = p.n_levels;

	if (tiled) {
		// Tiled output (mlp.h): chunk ch of sample i lives at ((i/128)*(W/8) + ch)*128 + i%128 (in
		// 16-byte units), so a warp's stores are contiguous. Rows past n (tile padding) are zeroed.
		if (i >= next_multiple_d(n, 128u)) return;
		uint4* base = reinterpret_cast<uint4*>(out) + (size_t)(i / 128) * (out_stride / 8) * 128 + (i % 128);
		if (i >= n) {
			for (uint32_t ch = 0; ch < out_stride / 8; ++ch) base[ch * 128] = make_uint4(0u, 0u, 0u, 0u);
			return;
		}
		float x[D];
		load_position<D>(positions, in_stride, i, x);
		for (uint32_t l0 = 0; l0 < L; l0 += C) {
			alignas(16) __half vals[8];
#pragma unroll
			for (uint32_t c = 0; c < C; ++c) {
				const uint32_t level = l0 + c;
				if (level < L && !grid_level_masked_fwd(p, level, max_level)) {
					grid_encode_level<D, F>(p, level, x, grid, &vals[c * F]);
				} else {
#pragma unroll
					for (uint32_t f = 0; f < F; ++f) vals[c * F + f] = __float2half(level < L ? 0.0f : 1.0f);
				}
			}
			base[(l0 * F / 8) * 128] = *reinterpret_cast<const uint4*>(vals);
		}
		for (uint32_t ch = (L * F + 7) / 8; ch < out_stride / 8; ++ch) base[ch * 128] = make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u);
		return;
	}

	if (i >= n) return;

	float x[D];
	load_position<D>(positions, in_stride, i, x);

	__half* dst = out + (size_t)i * out_stride;

	for (uint32_t l0 = 0; l0 < L; l0 += C) {
		alignas(16) __half vals[8];
#pragma unroll
		for (uint32_t c = 0; c < C; ++c) {
			const uint32_t level = l0 + c;
			if (level < L && !grid_level_masked_fwd(p, level, max_level)) {
				grid_encode_level<D, F>(p, level, x, grid, &vals[c * F]);
			} else {
#pragma unroll
				for (uint32_t f = 0; f < F; ++f) vals[c * F + f] = __float2half(0.0f);
			}
		}
		if (VEC_OUT && l0 + C <= L) {
			*reinterpret_cast<uint4*>(dst + l0 * F) = *reinterpret_cast<const uint4*>(vals);
		} else {
#pragma unroll
			for (uint32_t k = 0; k < 8; ++k) {
				if (l0 * F + k < L * F) dst[l0 * F + k] = vals[k];
			}
		}
	}
	// Padding columns carry 1 (tcnn Encoding padding, grid.h:806-811).
	for (uint32_t k = L * F; k < out_stride; ++k) dst[k] = __float2half(1.0f);
}

// Parameter gradients: thread per sample; per level 2^D packed reductions into L2.
//
// Warp-aggregated scatter: NeRF batches are ray-ordered, so on the coarse levels runs of
// consecutive samples fall into the same cell and would hit the same 2^D entries with 2^D * F/2
// atomics each - the L2 serialises same-address reductions, and the ncu capture of round 1 shows
// this kernel at 9 % L2 / 5 % SM throughput for exactly that reason. Per level the warp finds the
// runs of equal cells among neighbouring lanes (one ballot), sums each run's contributions with
// a segmented shuffle reduction in fp32 and lets the run's first lane issue one
// red.global.add.f16x2 per entry. Levels where no two neighbouring lanes share a cell (the fine,
// hashed ones) take the direct path. Each contribution is still half(w) * half(dy) rounded to
// fp16 as in the reference (`(GRAD_T)weight * grad`, grid.h:302-311); the run sum is rounded once.
template <uint32_t D, uint32_t F, bool VEC_IN>
__global__ void __launch_bounds__(256) grid_backward_kernel(
	const __grid_constant__ GridParams p, const uint32_t n,
	__half* __restrict__ grad, const float* __restrict__ positions, const uint32_t in_stride,
	const __half* __restrict__ dL_dy, const uint32_t dy_stride, const float max_level, const uint32_t tiled, const uint32_t first_level)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x % 32;
	const bool active = i < n;
	if (__ballot_sync(0xFFFFFFFFu, active) == 0) return;

	float x[D];
	if (active) load_position<D>(positions, in_stride, i, x);
	else {
#pragma unroll
		for (uint32_t d = 0; d < D; ++d) x[d] = 0.0f;
	}

	constexpr uint32_t C = 8 / F;
	const uint32_t L = p.n_levels;
	const __half* src = dL_dy + (size_t)i * dy_stride;
	const uint4* tbase = reinterpret_cast<const uint4*>(dL_dy) + (size_t)(i / 128) * (dy_stride / 8) * 128 + (i % 128);

	for (uint32_t l0 = (first_level / C) * C; l0 < L; l0 += C) {
		alignas(16) __half g[8];
		if (!active) {
			*reinterpret_cast<uint4*>(g) = make_uint4(0u, 0u, 0u, 0u);
		} else if (tiled) {
			*reinterpret_cast<uint4*>(g) = __ldg(tbase + (l0 * F / 8) * 128);
		} else if (VEC_IN && l0 + C <= L) {
			*reinterpret_cast<uint4*>(g) = __ldg(reinterpret_cast<const uint4*>(src + l0 * F));
		} else {
#pragma unroll
			for (uint32_t k = 0; k < 8; ++k) g[k] = (l0 * F + k < L * F) ? src[l0 * F + k] : __float2half(0.0f);
		}
		grid_scatter_chunk<D, F>(p, l0, x, g, grad, max_level, active, lane, first_level);
	}
}

// Coarse dense levels, privatised in shared memory. On unordered batches (config 1: 2^20 uniform points) every
// sample sends 2^D reductions into the few thousand entries of each of the coarsest levels, and the L2 serialises
// reductions to one address: levels 0 and 1 of the L16 grid cost 0.21 ms of a 0.78 ms backward pass, five times a
// fine level (scripts/grid_probe.py). Here a block keeps an fp32 copy of the leading `n_coarse` dense levels in its
// shared memory (level 0: 17^3 x F floats; together at most kCoarseSmemBytes), its threads add their samples'
// contributions there with shared-memory atomics, and at the end the block sends each touched group of 8 values to
// the gradient table with one 16-byte vector reduction. grid_backward_kernel then starts at level n_coarse.
// Each contribution is half(w) * half(dy) rounded to fp16 as everywhere else; the block's sum is fp32, rounded once.
// Shared memory has native atomics for 32-bit integers only; a float add is a compare-and-swap loop either way
// (ATOMS.CAST.SPIN). One 64-bit loop carries a feature pair, which halves the number of loops.
NGPB_D void smem_add_float2(float2* addr, float2 v) {
	unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
	unsigned long long old = *a, assumed;
	do {
		assumed = old;
		const float2 cur = make_float2(__uint_as_float((uint32_t)assumed), __uint_as_float((uint32_t)(assumed >> 32)));
		const unsigned long long next = (unsigned long long)__float_as_uint(cur.x + v.x) | ((unsigned long long)__float_as_uint(cur.y + v.y) << 32);
		old = atomicCAS(a, assumed, next);
	} while (old != assumed);
}

constexpr uint32_t kCoarseSmemBytes = 200u * 1024u;
constexpr uint32_t kCoarseThreads = 1024;

template <uint32_t D, uint32_t F>
__global__ void __launch_bounds__(kCoarseThreads) grid_backward_coarse_kernel(
	const __grid_constant__ GridParams p, const uint32_t n, const uint32_t n_coarse,
	__half* __restrict__ grad, const float* __restrict__ positions, const uint32_t in_stride,
	const __half* __restrict__ dL_dy, const uint32_t dy_stride, const float max_level, const uint32_t tiled)
{
	extern __shared__ __align__(16) float coarse_acc[];
	constexpr uint32_t NC = 1u << D;
	const uint32_t total = (p.offset[n_coarse - 1] + p.size[n_coarse - 1]) * F;   // floats; a multiple of 8
	for (uint32_t e = threadIdx.x; e < total / 4; e += blockDim.x) reinterpret_cast<float4*>(coarse_acc)[e] = make_float4(0.f, 0.f, 0.f, 0.f);
	__syncthreads();

	const uint32_t lane = threadIdx.x % 32;
	for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < n; i0 += gridDim.x * blockDim.x) {   // block-uniform trip count
		const uint32_t i = i0 + threadIdx.x;
		const bool active = i < n;
		float x[D];
#pragma unroll
		for (uint32_t d = 0; d < D; ++d) x[d] = 0.0f;
		if (active) load_position<D>(positions, in_stride, i, x);
		for (uint32_t level = 0; level < n_coarse; ++level) {
			if (grid_level_masked_bwd(p, level, max_level)) continue;
			__half2 g2[F / 2];
#pragma unroll
			for (uint32_t f = 0; f < F / 2; ++f) g2[f] = __float2half2_rn(0.0f);
			if (active) {
				const __half* src = tiled ? dL_dy + (((size_t)(i / 128) * (dy_stride / 8) + (level * F) / 8) * 128 + (i % 128)) * 8 + (level * F) % 8
				                          : dL_dy + (size_t)i * dy_stride + level * F;
#pragma unroll
				for (uint32_t f = 0; f < F / 2; ++f) g2[f] = __halves2half2(src[2 * f], src[2 * f + 1]);
			}
			GridCellT<D> cell;
			grid_cell<D>(p, level, x, cell);
			float2 acc[NC][F / 2];
#pragma unroll
			for (uint32_t k = 0; k < NC; ++k) {
				const __half2 w = __float2half2_rn(active ? grid_corner_weight<D>(cell, k) : 0.0f);
#pragma unroll
				for (uint32_t f = 0; f < F / 2; ++f) acc[k][f] = __half22float2(__hmul2(w, g2[f]));
			}
			// ordered batches (stratified image samples, rays): neighbouring lanes share a cell; one lane adds the run's sum
			const uint32_t heads = grid_run_heads<D>(cell, active, lane);
			if (heads != 0xFFFFFFFFu) grid_run_reduce<NC, F / 2>(acc, heads, lane);
			if (active && ((heads >> lane) & 1u)) {
				float* level_acc = coarse_acc + (size_t)p.offset[level] * F;
#pragma unroll
				for (uint32_t k = 0; k < NC; ++k) {
					float* dst = level_acc + (size_t)grid_corner_index<D>(p, level, cell, k) * F;
#pragma unroll
					for (uint32_t f = 0; f < F / 2; ++f) smem_add_float2(reinterpret_cast<float2*>(dst) + f, acc[k][f]);
				}
			}
		}
	}
	__syncthreads();

	for (uint32_t e = threadIdx.x; e < total / 8; e += blockDim.x) {
		const float4 a = reinterpret_cast<const float4*>(coarse_acc)[2 * e], b = reinterpret_cast<const float4*>(coarse_acc)[2 * e + 1];
		if (a.x == 0.f && a.y == 0.f && a.z == 0.f && a.w == 0.f && b.x == 0.f && b.y == 0.f && b.z == 0.f && b.w == 0.f) continue;
		red_add_half2x4(grad + (size_t)e * 8, __floats2half2_rn(a.x, a.y), __floats2half2_rn(a.z, a.w), __floats2half2_rn(b.x, b.y), __floats2half2_rn(b.z, b.w));
	}
}

// Leading dense levels that fit the block's shared memory together (0: the coarse kernel is not used).
static uint32_t coarse_levels(const GridParams& p, uint32_t n) {
	static const bool off = getenv("NGPB_GRID_COARSE") && atoi(getenv("NGPB_GRID_COARSE")) == 0;
	if (off || n < (1u << 16) || p.interpolation_type == (int)InterpolationType::Nearest || p.n_features_per_level == 1) return 0;
	uint32_t k = 0;
	while (k < p.n_levels && !p.use_hash[k] && (size_t)(p.offset[k] + p.size[k]) * p.n_features_per_level * sizeof(float) <= kCoarseSmemBytes) ++k;
	return k;
}

// Input gradient, recomputing dy/dx from the table (fp32 math as grid.h:172-211, 322-349).
template <uint32_t D, uint32_t F>
__global__ void __launch_bounds__(256) grid_backward_input_kernel(
	const __grid_constant__ GridParams p, const uint32_t n,
	const __half* __restrict__ grid, const float* __restrict__ positions, const uint32_t in_stride,
	const __half* __restrict__ dL_dy, const uint32_t dy_stride,
	float* __restrict__ dL_dx, const uint32_t dx_stride, const float max_level)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float x[D];
	load_position<D>(positions, in_stride, i, x);
	float result[D];
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) result[d] = 0.0f;

	const __half* dy = dL_dy + (size_t)i * dy_stride;
	for (uint32_t level = 0; level < p.n_levels; ++level) {
		if (grid_level_masked_fwd(p, level, max_level)) continue;
		if (p.interpolation_type == (int)InterpolationType::Nearest) continue;
		GridCellT<D> c;
		grid_cell<D, true>(p, level, x, c);
		const __half* base = grid + (size_t)p.offset[level] * F;
		const float scale = p.scale[level];
		float gy[F];
#pragma unroll
		for (uint32_t f = 0; f < F; ++f) gy[f] = __half2float(dy[level * F + f]);

#pragma unroll
		for (uint32_t gd = 0; gd < D; ++gd) {
			float grads[F];
#pragma unroll
			for (uint32_t f = 0; f < F; ++f) grads[f] = 0.0f;
#pragma unroll
			for (uint32_t idx = 0; idx < (1u << (D - 1)); ++idx) {
				float weight = scale;
				uint32_t corner = 0;
#pragma unroll
				for (uint32_t ngd = 0; ngd < D - 1; ++ngd) {
					const uint32_t dim = ngd >= gd ? ngd + 1 : ngd;
					if ((idx & (1u << ngd)) == 0) weight *= 1 - c.w[dim];
					else { weight *= c.w[dim]; corner |= 1u << dim; }
				}
				GridEntry<F> left = grid_load<F>(base, grid_corner_index<D>(p, level, c, corner));
				GridEntry<F> right = grid_load<F>(base, grid_corner_index<D>(p, level, c, corner | (1u << gd)));
#pragma unroll
				for (uint32_t f = 0; f < F; ++f) {
					float vl, vr;
					if constexpr (F == 1) { vl = __half2float(left.v); vr = __half2float(right.v); }
					else {
						vl = (f & 1) ? __high2float(left.v[f / 2]) : __low2float(left.v[f / 2]);
						vr = (f & 1) ? __high2float(right.v[f / 2]) : __low2float(right.v[f / 2]);
					}
					grads[f] += weight * (vr - vl) * c.dw[gd];
				}
			}
#pragma unroll
			for (uint32_t f = 0; f < F; ++f) result[gd] += gy[f] * grads[f];
		}
	}
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) dL_dx[(size_t)i * dx_stride + d] = result[d];
}

template <uint32_t D>
__global__ void __launch_bounds__(256) grid_indices_kernel(
	const __grid_constant__ GridParams p, const uint32_t n,
	const float* __restrict__ positions, const uint32_t in_stride, uint32_t* __restrict__ idx)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float x[D];
	load_position<D>(positions, in_stride, i, x);
	for (uint32_t level = 0; level < p.n_levels; ++level) {
		GridCellT<D> c;
		grid_cell<D>(p, level, x, c);
#pragma unroll
		for (uint32_t k = 0; k < (1u << D); ++k) {
			idx[((size_t)i * p.n_levels + level) * (1u << D) + k] = grid_corner_index<D>(p, level, c, k);
		}
	}
}

// =================================== launchers ===========================================

#define NGPB_GRID_DISPATCH(D_, F_, ...) \
	do { \
		bool _ok = true; \
		switch ((D_) * 16 + (F_)) { \
			case 2 * 16 + 1: { constexpr uint32_t D = 2, F = 1; __VA_ARGS__; } break; \
			case 2 * 16 + 2: { constexpr uint32_t D = 2, F = 2; __VA_ARGS__; } break; \
			case 2 * 16 + 4: { constexpr uint32_t D = 2, F = 4; __VA_ARGS__; } break; \
			case 2 * 16 + 8: { constexpr uint32_t D = 2, F = 8; __VA_ARGS__; } break; \
			case 3 * 16 + 1: { constexpr uint32_t D = 3, F = 1; __VA_ARGS__; } break; \
			case 3 * 16 + 2: { constexpr uint32_t D = 3, F = 2; __VA_ARGS__; } break; \
			case 3 * 16 + 4: { constexpr uint32_t D = 3, F = 4; __VA_ARGS__; } break; \
			case 3 * 16 + 8: { constexpr uint32_t D = 3, F = 8; __VA_ARGS__; } break; \
			case 4 * 16 + 1: { constexpr uint32_t D = 4, F = 1; __VA_ARGS__; } break; \
			case 4 * 16 + 2: { constexpr uint32_t D = 4, F = 2; __VA_ARGS__; } break; \
			case 4 * 16 + 4: { constexpr uint32_t D = 4, F = 4; __VA_ARGS__; } break; \
			case 4 * 16 + 8: { constexpr uint32_t D = 4, F = 8; __VA_ARGS__; } break; \
			default: _ok = false; \
		} \
		if (!_ok) throw std::runtime_error{"GridEncoding: unsupported (n_dims, n_features_per_level)"}; \
	} while (0)

static bool aligned16(const void* ptr, uint32_t stride_halves) {
	return ((uintptr_t)ptr % 16 == 0) && ((stride_halves * 2) % 16 == 0);
}

void grid_forward(cudaStream_t stream, const GridParams& p, uint32_t n, const __half* grid,
	const float* positions, uint32_t in_stride, __half* out, uint32_t out_stride, float max_level, bool tiled, const uint32_t* n_dev)
{
	if (n == 0) return;
	const uint32_t blocks = div_round_up(tiled ? next_multiple(n, 128u) : n, 256);
	const bool vec = aligned16(out, out_stride);
	if (tiled && (!vec || out_stride % 8 != 0 || (p.n_levels * p.n_features_per_level) % 8 != 0)) {
		throw std::runtime_error{"GridEncoding: tiled output needs a 16-byte aligned buffer and a feature count that is a multiple of 8"};
	}
	NGPB_GRID_DISPATCH(p.n_dims, p.n_features_per_level, {
		if (vec) grid_forward_kernel<D, F, true><<<blocks, 256, 0, stream>>>(p, n, grid, positions, in_stride, out, out_stride, max_level, tiled ? 1u : 0u, n_dev);
		else grid_forward_kernel<D, F, false><<<blocks, 256, 0, stream>>>(p, n, grid, positions, in_stride, out, out_stride, max_level, 0u, n_dev);
	});
	NGPB_CUDA_CHECK(cudaGetLastError());
}

void grid_backward(cudaStream_t stream, const GridParams& p, uint32_t n, __half* grad,
	const float* positions, uint32_t in_stride, const __half* dL_dy, uint32_t dy_stride, float max_level, bool tiled, bool ordered)
{
	if (n == 0) return;
	const uint32_t blocks = div_round_up(n, 256);
	const bool vec = aligned16(dL_dy, dy_stride);
	if (tiled && (!vec || dy_stride % 8 != 0 || (p.n_levels * p.n_features_per_level) % 8 != 0)) {
		throw std::runtime_error{"GridEncoding: tiled gradients need a 16-byte aligned buffer and a feature count that is a multiple of 8"};
	}
	static const bool coarse_for_ordered = getenv("NGPB_GRID_COARSE_ORDERED") && atoi(getenv("NGPB_GRID_COARSE_ORDERED")) != 0;
	const uint32_t n_coarse = (ordered && !coarse_for_ordered) ? 0u : coarse_levels(p, n);
	NGPB_GRID_DISPATCH(p.n_dims, p.n_features_per_level, {
		if constexpr (F >= 2) {
			if (n_coarse > 0) {
				const uint32_t smem = (p.offset[n_coarse - 1] + p.size[n_coarse - 1]) * F * (uint32_t)sizeof(float);
				static bool attr_set = false;   // per instantiation
				if (!attr_set) {
					NGPB_CUDA_CHECK(cudaFuncSetAttribute(grid_backward_coarse_kernel<D, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCoarseSmemBytes));
					attr_set = true;
				}
				const uint32_t coarse_blocks = std::min((uint32_t)sm_count(), div_round_up(n, kCoarseThreads));
				grid_backward_coarse_kernel<D, F><<<coarse_blocks, kCoarseThreads, smem, stream>>>(p, n, n_coarse, grad, positions, in_stride, dL_dy, dy_stride, max_level, tiled ? 1u : 0u);
			}
		}
		if (vec) grid_backward_kernel<D, F, true><<<blocks, 256, 0, stream>>>(p, n, grad, positions, in_stride, dL_dy, dy_stride, max_level, tiled ? 1u : 0u, n_coarse);
		else grid_backward_kernel<D, F, false><<<blocks, 256, 0, stream>>>(p, n, grad, positions, in_stride, dL_dy, dy_stride, max_level, 0u, n_coarse);
	});
	NGPB_CUDA_CHECK(cudaGetLastError());
}

void grid_backward_input(cudaStream_t stream, const GridParams& p, uint32_t n, const __half* grid,
	const float* positions, uint32_t in_stride, const __half* dL_dy, uint32_t dy_stride,
	float* dL_dx, uint32_t dx_stride, float max_level)
{
	if (n == 0) return;
	const uint32_t blocks = div_round_up(n, 256);
	NGPB_GRID_DISPATCH(p.n_dims, p.n_features_per_level, {
		grid_backward_input_kernel<D, F><<<blocks, 256, 0, stream>>>(p, n, grid, positions, in_stride, dL_dy, dy_stride, dL_dx, dx_stride, max_level);
	});
	NGPB_CUDA_CHECK(cudaGetLastError());
}

void grid_indices(cudaStream_t stream, const GridParams& p, uint32_t n, const float* positions,
	uint32_t in_stride, uint32_t* idx)
{
	if (n == 0) return;
	const uint32_t blocks = div_round_up(n, 256);
	switch (p.n_dims) {
		case 2: grid_indices_kernel<2><<<blocks, 256, 0, stream>>>(p, n, positions, in_stride, idx); break;
		case 3: grid_indices_kernel<3><<<blocks, 256, 0, stream>>>(p, n, positions, in_stride, idx); break;
		case 4: grid_indices_kernel<4><<<blocks, 256, 0, stream>>>(p, n, positions, in_stride, idx); break;
		default: throw std::runtime_error{"GridEncoding: number of input dims must be 2 or 3."};
	}
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
