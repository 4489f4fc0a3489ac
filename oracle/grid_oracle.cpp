// CPU oracle for the multiresolution grid encoding. TEST INFRASTRUCTURE ONLY (see oracle.h).
// Restates, routine by routine:
//   grid_scale / grid_resolution        tcnn common_device.h:709-718
//   GridEncodingTemplated ctor offsets  tcnn encodings/grid.h:692-723
//   pos_fract                           tcnn common_device.h:856-868
//   lcg_hash / *_prime_hash / grid_index tcnn common_device.h:631-707
//   kernel_grid                         tcnn encodings/grid.h:48-170
//   kernel_grid (dy_dx) + backward_input grid.h:172-211, 322-349
//   kernel_grid_backward                grid.h:214-320
// fp16 arithmetic is emulated exactly: a half FMA is round_to_half(a*b + c) with a*b + c
// evaluated in double (22-bit product + 11-bit addend fit without intermediate rounding for
// every value the tests use).
#include "oracle.h"
#include "pcg32_oracle.h"

#include <cmath>
#include <cstring>
#include <limits>
#include <omp.h>
#include <vector>

typedef _Float16 half_t;

static inline half_t bits_to_half(uint16_t b) { half_t h; memcpy(&h, &b, 2); return h; }
static inline uint16_t half_to_bits(half_t h) { uint16_t b; memcpy(&b, &h, 2); return b; }
static inline half_t half_fma(half_t a, half_t b, half_t c) { return (half_t)((double)a * (double)b + (double)c); }
static inline half_t half_mul(half_t a, half_t b) { return (half_t)((double)a * (double)b); }

extern "C" uint16_t oracle_float_to_half(float f) { return half_to_bits((half_t)f); }
extern "C" float oracle_half_to_float(uint16_t h) { return (float)bits_to_half(h); }
extern "C" int oracle_num_threads(void) { return omp_get_max_threads(); }
// The thread count of the oracle's OpenMP loops (torchrun pins OMP_NUM_THREADS=1 for its workers; the timed CPU legs of
// bench.py want the host's cores without changing the variable for the rest of the process).
extern "C" void oracle_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

namespace {

constexpr uint32_t MAX_DIMS = 4;
constexpr uint32_t MAX_LEVELS = 128;

const uint32_t PRIME_FACTORS[7] = {1958374283u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};

float host_grid_scale(uint32_t level, float log2_per_level_scale, uint32_t base_resolution) {
	return exp2f(level * log2_per_level_scale) * base_resolution - 1.0f;
}
uint32_t grid_resolution_of(float scale) { return (uint32_t)ceilf(scale) + 1; }

uint32_t next_multiple_u32(uint32_t v, uint32_t m) { return ((v + m - 1) / m) * m; }
uint32_t powi(uint32_t b, uint32_t e) { uint32_t r = 1; for (uint32_t i = 0; i < e; ++i) r *= b; return r; }

struct Levels {
	uint32_t offsets[MAX_LEVELS + 1];
	float scale[MAX_LEVELS];
	uint32_t resolution[MAX_LEVELS];
};

void build_levels(const oracle_grid_config& c, Levels& lv) {
	const float log2_pls = std::log2(c.per_level_scale);
	uint32_t offset = 0;
	for (uint32_t i = 0; i < c.n_levels; ++i) {
		// Host-side table (constructor): always the host's exp2f.
		const uint32_t resolution = grid_resolution_of(host_grid_scale(i, log2_pls, c.base_resolution));
		const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
		uint32_t params_in_level = std::pow((float)resolution, (float)c.n_dims) > (float)max_params ? max_params : powi(resolution, c.n_dims);
		params_in_level = next_multiple_u32(params_in_level, 8u);
		if (c.grid_type == 2) params_in_level = std::min(params_in_level, powi(c.base_resolution, c.n_dims));
		else if (c.grid_type == 0) params_in_level = std::min(params_in_level, 1u << c.log2_hashmap_size);
		lv.offsets[i] = offset;
		offset += params_in_level;
		// Kernel-side values.
		lv.scale[i] = c.scale_override ? c.scale_override[i] : host_grid_scale(i, log2_pls, c.base_resolution);
		lv.resolution[i] = grid_resolution_of(lv.scale[i]);
	}
	lv.offsets[c.n_levels] = offset;
}

uint32_t hash_coords(const oracle_grid_config& c, const uint32_t* pos_grid) {
	uint32_t result = 0;
	for (uint32_t i = 0; i < c.n_dims; ++i) {
		uint32_t f;
		switch (c.hash_type) {
			case 0: f = PRIME_FACTORS[i]; break;
			case 1: f = i == 0 ? 1u : PRIME_FACTORS[i]; break;
			default: f = PRIME_FACTORS[6 - i]; break;
		}
		result ^= pos_grid[i] * f;
	}
	return result;
}

uint32_t grid_index(const oracle_grid_config& c, uint32_t hashmap_size, uint32_t resolution, const uint32_t* pos_grid) {
	uint32_t stride = 1;
	uint32_t index = 0;
	for (uint32_t dim = 0; dim < c.n_dims && stride <= hashmap_size; ++dim) {
		index += pos_grid[dim] * stride;
		stride *= resolution;
	}
	if (c.grid_type == 0 && hashmap_size < stride) index = hash_coords(c, pos_grid);
	return index % hashmap_size;
}

float smoothstep(float v) { return v * v * (3.0f - 2.0f * v); }
float smoothstep_derivative(float v) { return 6 * v * (1.0f - v); }

void pos_fract(const oracle_grid_config& c, float input, float scale, float* pos, float* pos_derivative, uint32_t* pos_grid) {
	*pos = fmaf(scale, input, 0.5f);
	float tmp = floorf(*pos);
	*pos_grid = (uint32_t)(int)tmp;
	*pos -= tmp;
	if (c.interpolation == 2) {
		*pos_derivative = smoothstep_derivative(*pos);
		*pos = smoothstep(*pos);
	} else {
		*pos_derivative = 1.0f;
	}
}

bool masked_fwd(const oracle_grid_config& c, uint32_t level, float max_level) {
	const float ml = (max_level * (float)(c.n_levels * c.n_features_per_level)) / (float)c.n_features_per_level;
	return (float)level >= ml + 1e-3f;
}
bool masked_bwd(const oracle_grid_config& c, uint32_t level, float max_level) {
	const float ml = (max_level * (float)(c.n_levels * c.n_features_per_level)) / (float)c.n_features_per_level;
	return (float)level > ml + 1e-3f;
}

} // namespace

extern "C" uint32_t oracle_grid_offsets(const oracle_grid_config* cfg, uint32_t* offsets) {
	Levels lv;
	build_levels(*cfg, lv);
	for (uint32_t i = 0; i <= cfg->n_levels; ++i) offsets[i] = lv.offsets[i];
	return lv.offsets[cfg->n_levels];
}

extern "C" void oracle_grid_levels(const oracle_grid_config* cfg, float* scales, uint32_t* resolutions) {
	Levels lv;
	build_levels(*cfg, lv);
	for (uint32_t i = 0; i < cfg->n_levels; ++i) { scales[i] = lv.scale[i]; resolutions[i] = lv.resolution[i]; }
}

extern "C" void oracle_grid_indices(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride, uint32_t* idx) {
	const oracle_grid_config& c = *cfg;
	Levels lv;
	build_levels(c, lv);
	const uint32_t D = c.n_dims, NC = 1u << D;
#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		for (uint32_t level = 0; level < c.n_levels; ++level) {
			const uint32_t hashmap_size = lv.offsets[level + 1] - lv.offsets[level];
			float pos[MAX_DIMS], der[MAX_DIMS];
			uint32_t pos_grid[MAX_DIMS];
			for (uint32_t d = 0; d < D; ++d) pos_fract(c, positions[(size_t)i * in_stride + d], lv.scale[level], &pos[d], &der[d], &pos_grid[d]);
			for (uint32_t k = 0; k < NC; ++k) {
				uint32_t local[MAX_DIMS];
				for (uint32_t d = 0; d < D; ++d) local[d] = (k & (1u << d)) ? pos_grid[d] + 1 : pos_grid[d];
				idx[((size_t)i * c.n_levels + level) * NC + k] = grid_index(c, hashmap_size, lv.resolution[level], local);
			}
		}
	}
}

extern "C" void oracle_grid_forward(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* params, uint16_t* out, uint32_t out_stride, float max_level)
{
	const oracle_grid_config& c = *cfg;
	Levels lv;
	build_levels(c, lv);
	const uint32_t D = c.n_dims, F = c.n_features_per_level, NC = 1u << D;
	const uint16_t one = half_to_bits((half_t)1.0f);
#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		uint16_t* dst = out + (size_t)i * out_stride;
		for (uint32_t level = 0; level < c.n_levels; ++level) {
			if (masked_fwd(c, level, max_level)) {
				for (uint32_t f = 0; f < F; ++f) dst[level * F + f] = 0;
				continue;
			}
			const uint16_t* grid = params + (size_t)lv.offsets[level] * F;
			const uint32_t hashmap_size = lv.offsets[level + 1] - lv.offsets[level];
			float pos[MAX_DIMS], der[MAX_DIMS];
			uint32_t pos_grid[MAX_DIMS];
			for (uint32_t d = 0; d < D; ++d) pos_fract(c, positions[(size_t)i * in_stride + d], lv.scale[level], &pos[d], &der[d], &pos_grid[d]);

			if (c.interpolation == 0) {
				const uint32_t index = grid_index(c, hashmap_size, lv.resolution[level], pos_grid) * F;
				for (uint32_t f = 0; f < F; ++f) dst[level * F + f] = grid[index + f];
				continue;
			}
			half_t result[8] = {};
			for (uint32_t k = 0; k < NC; ++k) {
				float weight = 1;
				uint32_t local[MAX_DIMS];
				for (uint32_t d = 0; d < D; ++d) {
					if ((k & (1u << d)) == 0) { weight *= 1 - pos[d]; local[d] = pos_grid[d]; }
					else { weight *= pos[d]; local[d] = pos_grid[d] + 1; }
				}
				const uint32_t index = grid_index(c, hashmap_size, lv.resolution[level], local) * F;
				const half_t w = (half_t)weight;
				for (uint32_t f = 0; f < F; ++f) result[f] = half_fma(w, bits_to_half(grid[index + f]), result[f]);
			}
			for (uint32_t f = 0; f < F; ++f) dst[level * F + f] = half_to_bits(result[f]);
		}
		for (uint32_t k = c.n_levels * F; k < out_stride; ++k) dst[k] = one;
	}
}

extern "C" void oracle_grid_backward(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* dL_dy, uint32_t dy_stride, double* grad, float max_level)
{
	const oracle_grid_config& c = *cfg;
	Levels lv;
	build_levels(c, lv);
	const uint32_t D = c.n_dims, F = c.n_features_per_level, NC = 1u << D;
	memset(grad, 0, sizeof(double) * (size_t)lv.offsets[c.n_levels] * F);
	// Levels write disjoint ranges: parallel over levels, serial (deterministic) over samples.
#pragma omp parallel for schedule(dynamic, 1)
	for (int level = 0; level < (int)c.n_levels; ++level) {
		if (masked_bwd(c, level, max_level)) continue;
		double* g = grad + (size_t)lv.offsets[level] * F;
		const uint32_t hashmap_size = lv.offsets[level + 1] - lv.offsets[level];
		for (uint32_t i = 0; i < n; ++i) {
			float pos[MAX_DIMS], der[MAX_DIMS];
			uint32_t pos_grid[MAX_DIMS];
			for (uint32_t d = 0; d < D; ++d) pos_fract(c, positions[(size_t)i * in_stride + d], lv.scale[level], &pos[d], &der[d], &pos_grid[d]);
			const uint16_t* dy = dL_dy + (size_t)i * dy_stride + level * F;
			if (c.interpolation == 0) {
				const uint32_t index = grid_index(c, hashmap_size, lv.resolution[level], pos_grid) * F;
				for (uint32_t f = 0; f < F; ++f) g[index + f] += (double)half_mul((half_t)1.0f, bits_to_half(dy[f]));
				continue;
			}
			for (uint32_t k = 0; k < NC; ++k) {
				float weight = 1;
				uint32_t local[MAX_DIMS];
				for (uint32_t d = 0; d < D; ++d) {
					if ((k & (1u << d)) == 0) { weight *= 1 - pos[d]; local[d] = pos_grid[d]; }
					else { weight *= pos[d]; local[d] = pos_grid[d] + 1; }
				}
				const uint32_t index = grid_index(c, hashmap_size, lv.resolution[level], local) * F;
				const half_t w = (half_t)weight;
				for (uint32_t f = 0; f < F; ++f) g[index + f] += (double)half_mul(w, bits_to_half(dy[f]));
			}
		}
	}
}

extern "C" void oracle_grid_backward_input(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* params, const uint16_t* dL_dy, uint32_t dy_stride, float* dL_dx, float max_level)
{
	const oracle_grid_config& c = *cfg;
	Levels lv;
	build_levels(c, lv);
	const uint32_t D = c.n_dims, F = c.n_features_per_level;
#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		float result[MAX_DIMS] = {};
		// The reference sums over k = level*F + f in ascending order (grid.h:335-343).
		for (uint32_t level = 0; level < c.n_levels; ++level) {
			float grads[8][MAX_DIMS] = {};
			if (!masked_fwd(c, level, max_level) && c.interpolation != 0) {
				const uint16_t* grid = params + (size_t)lv.offsets[level] * F;
				const uint32_t hashmap_size = lv.offsets[level + 1] - lv.offsets[level];
				const float scale = lv.scale[level];
				float pos[MAX_DIMS], der[MAX_DIMS];
				uint32_t pos_grid[MAX_DIMS];
				for (uint32_t d = 0; d < D; ++d) pos_fract(c, positions[(size_t)i * in_stride + d], scale, &pos[d], &der[d], &pos_grid[d]);
				for (uint32_t grad_dim = 0; grad_dim < D; ++grad_dim) {
					for (uint32_t idx = 0; idx < (1u << (D - 1)); ++idx) {
						float weight = scale;
						uint32_t local[MAX_DIMS];
						for (uint32_t ngd = 0; ngd < D - 1; ++ngd) {
							const uint32_t dim = ngd >= grad_dim ? ngd + 1 : ngd;
							if ((idx & (1u << ngd)) == 0) { weight *= 1 - pos[dim]; local[dim] = pos_grid[dim]; }
							else { weight *= pos[dim]; local[dim] = pos_grid[dim] + 1; }
						}
						local[grad_dim] = pos_grid[grad_dim];
						const uint32_t il = grid_index(c, hashmap_size, lv.resolution[level], local) * F;
						local[grad_dim] = pos_grid[grad_dim] + 1;
						const uint32_t ir = grid_index(c, hashmap_size, lv.resolution[level], local) * F;
						for (uint32_t f = 0; f < F; ++f) {
							grads[f][grad_dim] += weight * ((float)bits_to_half(grid[ir + f]) - (float)bits_to_half(grid[il + f])) * der[grad_dim];
						}
					}
				}
			}
			for (uint32_t f = 0; f < F; ++f) {
				const float dy = (float)bits_to_half(dL_dy[(size_t)i * dy_stride + level * F + f]);
				for (uint32_t d = 0; d < D; ++d) result[d] += dy * grads[f][d];
			}
		}
		for (uint32_t d = 0; d < D; ++d) dL_dx[(size_t)i * D + d] = result[d];
	}
}

extern "C" void oracle_pcg32_floats(uint64_t seed, uint64_t advance, uint32_t n, float* out) {
	OraclePcg32 rng(seed);
	rng.advance((int64_t)advance);
	for (uint32_t i = 0; i < n; ++i) out[i] = rng.next_float();
}

// Same mapping continuing from an arbitrary generator state; advances the state by n
// (generate_random_uniform, tcnn random.h:66-73).
extern "C" void oracle_random_uniform_state(uint64_t* rng_state, size_t n, float* out, float lower, float upper) {
	const size_t n_threads_needed = (n + 3) / 4;
	const size_t n_threads = 128 * ((n_threads_needed + 127) / 128);
	for (size_t t = 0; t < n_threads; ++t) {
		OraclePcg32 rng;
		rng.state = rng_state[0];
		rng.inc = rng_state[1];
		rng.advance((int64_t)(t * 4));
		for (size_t j = 0; j < 4; ++j) {
			const size_t idx = t + n_threads * j;
			if (idx >= n) break;
			out[idx] = fmaf(rng.next_float(), upper - lower, lower);
		}
	}
	OraclePcg32 rng;
	rng.state = rng_state[0];
	rng.inc = rng_state[1];
	rng.advance((int64_t)n);
	rng_state[0] = rng.state;
}

extern "C" void oracle_random_uniform(uint64_t seed, size_t n, float* out, float lower, float upper) {
	// generate_random_kernel<T, RNG, 4> with 128-thread blocks (tcnn random.h:39-64).
	const size_t n_threads_needed = (n + 3) / 4;
	const size_t n_threads = 128 * ((n_threads_needed + 127) / 128);
	for (size_t t = 0; t < n_threads; ++t) {
		OraclePcg32 rng(seed);
		rng.advance((int64_t)(t * 4));
		for (size_t j = 0; j < 4; ++j) {
			const size_t idx = t + n_threads * j;
			if (idx >= n) break;
			out[idx] = fmaf(rng.next_float(), upper - lower, lower);
		}
	}
}
