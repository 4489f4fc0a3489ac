// Device-side pieces of the grid encoding, shared by the standalone kernels (grid.cu) and
// the fused encode+MLP kernels (mlp.cu).
//
// Arithmetic contract (what must agree with the reference bit for bit):
//   * cell coordinate and fraction: pos_fract, tcnn common_device.h:856-868
//   * corner index: grid_index / lcg_hash, tcnn common_device.h:631-707
//   * interpolation: fp16 FMA chain over corners in ascending corner order,
//     `result = fma((T)weight, val, result)`, tcnn grid.h:143-163
// The level-only quantities (scale, resolution, strides, hash-or-dense) come from the
// GridParams table instead of being recomputed per thread.
#pragma once

#include "grid.h"

namespace ngpb {

NGPB_D float smoothstep_f(float v) { return v * v * (3.0f - 2.0f * v); }
NGPB_D float smoothstep_derivative_f(float v) { return 6 * v * (1.0f - v); }

template <uint32_t D>
struct GridCellT {
	float w[D];        // interpolation weight toward the +1 corner per dim
	float dw[D];       // derivative of w w.r.t. the fractional coordinate
	uint32_t lo[D];    // per-dim index term of the lower corner (dense: pos*stride, hash: pos*prime)
	uint32_t hi[D];    // same for the +1 corner
};

// Per-(sample, level) set-up. `x` are the D input coordinates.
template <uint32_t D, bool WITH_DERIV = false>
NGPB_D void grid_cell(const GridParams& p, uint32_t level, const float (&x)[D], GridCellT<D>& c) {
	const float scale = p.scale[level];
	const bool hashed = p.use_hash[level] != 0;
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) {
		float pos = fmaf(scale, x[d], 0.5f);
		float fl = floorf(pos);
		uint32_t pg = (uint32_t)(int)fl;
		pos -= fl;
		if (p.interpolation_type == (int)InterpolationType::Smoothstep) {
			if (WITH_DERIV) c.dw[d] = smoothstep_derivative_f(pos);
			pos = smoothstep_f(pos);
		} else if (WITH_DERIV) {
			c.dw[d] = 1.0f;
		}
		c.w[d] = pos;
		const uint32_t m = hashed ? p.primes[d] : p.stride[level][d];
		c.lo[d] = pg * m;
		c.hi[d] = (pg + 1) * m;
	}
}

// Entry index (within the level) of corner `corner` (bit d set = +1 along dim d).
template <uint32_t D>
NGPB_D uint32_t grid_corner_index(const GridParams& p, uint32_t level, const GridCellT<D>& c, uint32_t corner) {
	const bool hashed = p.use_hash[level] != 0;
	uint32_t idx = 0;
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) {
		const uint32_t t = (corner & (1u << d)) ? c.hi[d] : c.lo[d];
		idx = hashed ? (idx ^ t) : (idx + t);
	}
	const uint32_t size = p.size[level];
	if (p.size_is_pow2[level]) return idx & (size - 1);
	return idx < size ? idx : idx % size;
}

template <uint32_t D>
NGPB_D float grid_corner_weight(const GridCellT<D>& c, uint32_t corner) {
	float weight = 1;
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) {
		weight *= (corner & (1u << d)) ? c.w[d] : 1 - c.w[d];
	}
	return weight;
}

#ifndef NGPB_GRID_PAIR_LOADS
#define NGPB_GRID_PAIR_LOADS 1
#endif
constexpr bool kGridPairLoads = NGPB_GRID_PAIR_LOADS != 0;

// One table entry = F halves, loaded with one instruction of F*2 bytes.
template <uint32_t F> struct GridEntry;
template <> struct GridEntry<1> { __half v; };
template <> struct alignas(4) GridEntry<2> { __half2 v[1]; };
template <> struct alignas(8) GridEntry<4> { __half2 v[2]; };
template <> struct alignas(16) GridEntry<8> { __half2 v[4]; };

template <uint32_t F>
NGPB_D GridEntry<F> grid_load(const __half* __restrict__ level_base, uint32_t idx) {
	GridEntry<F> e;
	if constexpr (F == 1) {
		e.v = __ldg(level_base + idx);
	} else if constexpr (F == 2) {
		e.v[0] = __ldg(reinterpret_cast<const __half2*>(level_base) + idx);
	} else if constexpr (F == 4) {
		uint2 r = __ldg(reinterpret_cast<const uint2*>(level_base) + idx);
		e.v[0] = *reinterpret_cast<__half2*>(&r.x);
		e.v[1] = *reinterpret_cast<__half2*>(&r.y);
	} else {
		uint4 r = __ldg(reinterpret_cast<const uint4*>(level_base) + idx);
		e.v[0] = *reinterpret_cast<__half2*>(&r.x);
		e.v[1] = *reinterpret_cast<__half2*>(&r.y);
		e.v[2] = *reinterpret_cast<__half2*>(&r.z);
		e.v[3] = *reinterpret_cast<__half2*>(&r.w);
	}
	return e;
}

// Interpolates level `level` at `x`; writes F halves to res. All 2^D gathers are issued
// before the FMA chain so that they are in flight together.
template <uint32_t D, uint32_t F>
NGPB_D void grid_encode_level(const GridParams& p, uint32_t level, const float (&x)[D],
	const __half* __restrict__ grid, __half* res)
{
	GridCellT<D> c;
	grid_cell<D>(p, level, x, c);
	const __half* base = grid + (size_t)p.offset[level] * F;

	if (p.interpolation_type == (int)InterpolationType::Nearest) {
		GridEntry<F> e = grid_load<F>(base, grid_corner_index<D>(p, level, c, 0));
		if constexpr (F == 1) res[0] = e.v;
		else {
#pragma unroll
			for (uint32_t f = 0; f < F / 2; ++f) reinterpret_cast<__half2*>(res)[f] = e.v[f];
		}
		return;
	}

	constexpr uint32_t NC = 1u << D;
	GridEntry<F> e[NC];
	if constexpr (kGridPairLoads && (F == 2 || F == 4)) {
		// x-neighbours that form one aligned entry pair (see grid_red_pair) come with one load of 2 F halves
#pragma unroll
		for (uint32_t k = 0; k < NC; k += 2) {
			const uint32_t i0 = grid_corner_index<D>(p, level, c, k), i1 = grid_corner_index<D>(p, level, c, k + 1);
			if ((i0 ^ i1) == 1u) {
				const bool swap = i0 > i1;
				if constexpr (F == 2) {
					const uint2 r = __ldg(reinterpret_cast<const uint2*>(base) + (i0 >> 1));
					e[k].v[0] = *reinterpret_cast<const __half2*>(swap ? &r.y : &r.x);
					e[k + 1].v[0] = *reinterpret_cast<const __half2*>(swap ? &r.x : &r.y);
				} else {
					const uint4 r = __ldg(reinterpret_cast<const uint4*>(base) + (i0 >> 1));
					const uint2 lo = make_uint2(r.x, r.y), hi = make_uint2(r.z, r.w);
					const uint2 a = swap ? hi : lo, b = swap ? lo : hi;
					e[k].v[0] = *reinterpret_cast<const __half2*>(&a.x); e[k].v[1] = *reinterpret_cast<const __half2*>(&a.y);
					e[k + 1].v[0] = *reinterpret_cast<const __half2*>(&b.x); e[k + 1].v[1] = *reinterpret_cast<const __half2*>(&b.y);
				}
			} else {
				e[k] = grid_load<F>(base, i0);
				e[k + 1] = grid_load<F>(base, i1);
			}
		}
	} else {
#pragma unroll
		for (uint32_t k = 0; k < NC; ++k) e[k] = grid_load<F>(base, grid_corner_index<D>(p, level, c, k));
	}

	if constexpr (F == 1) {
		__half acc = __float2half(0.0f);
#pragma unroll
		for (uint32_t k = 0; k < NC; ++k) acc = __hfma(__float2half(grid_corner_weight<D>(c, k)), e[k].v, acc);
		res[0] = acc;
	} else {
		__half2 acc[F / 2];
#pragma unroll
		for (uint32_t f = 0; f < F / 2; ++f) acc[f] = __float2half2_rn(0.0f);
#pragma unroll
		for (uint32_t k = 0; k < NC; ++k) {
			const __half2 w = __float2half2_rn(grid_corner_weight<D>(c, k));
#pragma unroll
			for (uint32_t f = 0; f < F / 2; ++f) acc[f] = __hfma2(w, e[k].v[f], acc[f]);
		}
#pragma unroll
		for (uint32_t f = 0; f < F / 2; ++f) reinterpret_cast<__half2*>(res)[f] = acc[f];
	}
}

// The reference zeroes levels at or above max_level (grid.h:69-92).
NGPB_D bool grid_level_masked_fwd(const GridParams& p, uint32_t level, float max_level) {
	const float ml = (max_level * (float)(p.n_levels * p.n_features_per_level)) / (float)p.n_features_per_level;
	return (float)level >= ml + 1e-3f;
}
// ... and skips the scatter for levels strictly above it (grid.h:241-243).
NGPB_D bool grid_level_masked_bwd(const GridParams& p, uint32_t level, float max_level) {
	const float ml = (max_level * (float)(p.n_levels * p.n_features_per_level)) / (float)p.n_features_per_level;
	return (float)level > ml + 1e-3f;
}

// Gradient scatter of one packed pair: the reference's atomic_add_gmem(__half2*) compiles
// to the same instruction (tcnn vec.h:326-349).
NGPB_D void red_add_half2(__half* addr, __half2 v) {
	asm volatile("red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;" ::"l"(addr), "r"(*reinterpret_cast<uint32_t*>(&v)) : "memory");
}
NGPB_D void red_add_half2x2(__half* addr, __half2 a, __half2 b) {
	asm volatile("red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(addr),
		"r"(*reinterpret_cast<uint32_t*>(&a)), "r"(*reinterpret_cast<uint32_t*>(&b)) : "memory");
}
NGPB_D void red_add_half2x4(__half* addr, __half2 a, __half2 b, __half2 c, __half2 d) {
	asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(addr),
		"r"(*reinterpret_cast<uint32_t*>(&a)), "r"(*reinterpret_cast<uint32_t*>(&b)),
		"r"(*reinterpret_cast<uint32_t*>(&c)), "r"(*reinterpret_cast<uint32_t*>(&d)) : "memory");
}

// Scatter w * g (F halves) into one table entry.
template <uint32_t F>
NGPB_D void grid_scatter_entry(__half* level_grad, uint32_t idx, float weight, const __half* g) {
	if constexpr (F == 1) {
		atomicAdd(level_grad + idx, __hmul(__float2half(weight), g[0]));
	} else {
		const __half2 w = __float2half2_rn(weight);
		const __half2* g2 = reinterpret_cast<const __half2*>(g);
		__half* addr = level_grad + (size_t)idx * F;
		if constexpr (F == 2) {
			red_add_half2(addr, __hmul2(w, g2[0]));
		} else if constexpr (F == 4) {
			red_add_half2x2(addr, __hmul2(w, g2[0]), __hmul2(w, g2[1]));
		} else {
			red_add_half2x4(addr, __hmul2(w, g2[0]), __hmul2(w, g2[1]), __hmul2(w, g2[2]), __hmul2(w, g2[3]));
		}
	}
}

// The two x-neighbours of a corner pair. Under the coherent-prime hash (x is multiplied by 1 and XORed in) and in the
// dense layout (x has stride 1) they are the two halves of one ALIGNED entry pair whenever the lower x is even - half
// of all cells: idx0 ^ idx1 == 1. Then one vector reduction carries both entries. The L2 retires vector reductions
// at the rate of scalar ones (scripts/grid_probe.py: 208-219 G instr/s for 4-, 8- and 16-byte reds alike), so this
// removes a quarter of the scatter's L2 operations. Values and rounding are those of two separate reductions.
template <uint32_t F>
NGPB_D void grid_red_pair(__half* level_grad, uint32_t idx0, uint32_t idx1, const __half2* v0, const __half2* v1) {
	if constexpr (F == 2 || F == 4) {
		if ((idx0 ^ idx1) == 1u) {
			const bool swap = idx0 > idx1;
			const __half2* lo = swap ? v1 : v0;
			const __half2* hi = swap ? v0 : v1;
			__half* addr = level_grad + (size_t)(idx0 & ~1u) * F;
			if constexpr (F == 2) red_add_half2x2(addr, lo[0], hi[0]);
			else red_add_half2x4(addr, lo[0], lo[1], hi[0], hi[1]);
			return;
		}
	}
	if constexpr (F == 2) {
		red_add_half2(level_grad + (size_t)idx0 * F, v0[0]);
		red_add_half2(level_grad + (size_t)idx1 * F, v1[0]);
	} else if constexpr (F == 4) {
		red_add_half2x2(level_grad + (size_t)idx0 * F, v0[0], v0[1]);
		red_add_half2x2(level_grad + (size_t)idx1 * F, v1[0], v1[1]);
	} else {
		red_add_half2x4(level_grad + (size_t)idx0 * F, v0[0], v0[1], v0[2], v0[3]);
		red_add_half2x4(level_grad + (size_t)idx1 * F, v1[0], v1[1], v1[2], v1[3]);
	}
}

// Runs of consecutive lanes whose samples fall into the same cell of a level: bit l of the result is set when lane l
// starts a run (inactive lanes are runs of their own). Must be called by all 32 lanes.
template <uint32_t D>
NGPB_D uint32_t grid_run_heads(const GridCellT<D>& cell, bool active, uint32_t lane) {
	bool same = lane != 0;
#pragma unroll
	for (uint32_t d = 0; d < D; ++d) same = (__shfl_up_sync(0xFFFFFFFFu, cell.lo[d], 1) == cell.lo[d]) && same;
	const bool prev_active = __shfl_up_sync(0xFFFFFFFFu, (int)active, 1) != 0;
	return __ballot_sync(0xFFFFFFFFu, !(same && active && prev_active));
}

// Segmented shuffle reduction over those runs: on exit the first lane of every run holds the run's sums.
template <uint32_t NC, uint32_t H>
NGPB_D void grid_run_reduce(float2 (&acc)[NC][H], uint32_t heads, uint32_t lane) {
	// lane takes the partial sum of lane + o while no run starts in (lane, lane + o]
	const uint32_t later_heads = lane == 31 ? 0xFFFFFFFFu : (heads >> (lane + 1));
#pragma unroll
	for (uint32_t o = 1; o < 32; o <<= 1) {
		const bool take = lane + o < 32 && (later_heads & ((1u << o) - 1u)) == 0;
#pragma unroll
		for (uint32_t k = 0; k < NC; ++k) {
#pragma unroll
			for (uint32_t f = 0; f < H; ++f) {
				const float tx = __shfl_down_sync(0xFFFFFFFFu, acc[k][f].x, o), ty = __shfl_down_sync(0xFFFFFFFFu, acc[k][f].y, o);
				if (take) { acc[k][f].x += tx; acc[k][f].y += ty; }
			}
		}
	}
}

// Gradient scatter of one 16-byte chunk of dL/dy (levels l0 .. l0 + 8/F - 1) of one sample, warp-aggregated:
// lanes of a warp whose samples fall into the same cell of a level (runs of consecutive lanes: NeRF batches are
// ray-ordered) are found with one ballot, their contributions summed in fp32 by a segmented shuffle reduction,
// and the run's first lane issues one red.global.add.f16x2 per entry. Must be called by all 32 lanes.
// Each contribution is half(w) * half(dy) rounded to fp16 as in the reference (grid.h:302-311).
// Levels below first_level are skipped (grid_backward_coarse_kernel has taken them).
template <uint32_t D, uint32_t F>
NGPB_D void grid_scatter_chunk(const GridParams& p, uint32_t l0, const float (&x)[D], const __half* g, __half* __restrict__ grad,
	float max_level, bool active, uint32_t lane, uint32_t first_level = 0)
{
	constexpr uint32_t C = 8 / F;
	constexpr uint32_t NC = 1u << D;
	constexpr bool kAggregate = NC * F <= 32 && F >= 2;
	const uint32_t L = p.n_levels;
#pragma unroll
	for (uint32_t c = 0; c < C; ++c) {
		const uint32_t level = l0 + c;
		if (level >= L || level < first_level || grid_level_masked_bwd(p, level, max_level)) continue; // warp-uniform
		GridCellT<D> cell;
		grid_cell<D>(p, level, x, cell);
		__half* level_grad = grad + (size_t)p.offset[level] * F;
		if (p.interpolation_type == (int)InterpolationType::Nearest) {
			if (active) grid_scatter_entry<F>(level_grad, grid_corner_index<D>(p, level, cell, 0), 1.0f, &g[c * F]);
			continue;
		}
		if constexpr (kAggregate) {
			// runs of equal cells among consecutive lanes
			const uint32_t heads = grid_run_heads<D>(cell, active, lane);
			if (heads != 0xFFFFFFFFu) {
				const __half2* g2 = reinterpret_cast<const __half2*>(&g[c * F]);
				float2 acc[NC][F / 2];
#pragma unroll
				for (uint32_t k = 0; k < NC; ++k) {
					const __half2 w = __float2half2_rn(active ? grid_corner_weight<D>(cell, k) : 0.0f);
#pragma unroll
					for (uint32_t f = 0; f < F / 2; ++f) acc[k][f] = __half22float2(__hmul2(w, g2[f]));
				}
				grid_run_reduce<NC, F / 2>(acc, heads, lane);
				if (active && ((heads >> lane) & 1u)) {
#pragma unroll
					for (uint32_t k = 0; k < NC; k += 2) {
						__half2 v0[F / 2], v1[F / 2];
#pragma unroll
						for (uint32_t f = 0; f < F / 2; ++f) { v0[f] = __float22half2_rn(acc[k][f]); v1[f] = __float22half2_rn(acc[k + 1][f]); }
						grid_red_pair<F>(level_grad, grid_corner_index<D>(p, level, cell, k), grid_corner_index<D>(p, level, cell, k + 1), v0, v1);
					}
				}
				continue;
			}
		}
		if (!active) continue;
		if constexpr (F >= 2) {
			const __half2* g2 = reinterpret_cast<const __half2*>(&g[c * F]);
#pragma unroll
			for (uint32_t k = 0; k < NC; k += 2) {
				const __half2 w0 = __float2half2_rn(grid_corner_weight<D>(cell, k)), w1 = __float2half2_rn(grid_corner_weight<D>(cell, k + 1));
				__half2 v0[F / 2], v1[F / 2];
#pragma unroll
				for (uint32_t f = 0; f < F / 2; ++f) { v0[f] = __hmul2(w0, g2[f]); v1[f] = __hmul2(w1, g2[f]); }
				grid_red_pair<F>(level_grad, grid_corner_index<D>(p, level, cell, k), grid_corner_index<D>(p, level, cell, k + 1), v0, v1);
			}
		} else {
#pragma unroll
			for (uint32_t k = 0; k < NC; ++k) {
				grid_scatter_entry<F>(level_grad, grid_corner_index<D>(p, level, cell, k), grid_corner_weight<D>(cell, k), &g[c * F]);
			}
		}
	}
}

} // namespace ngpb
