// Persistent warp-specialised kernel: hash-grid gather -> tcgen05 MLP chain(s) -> (training) backward chain
// -> warp-aggregated gradient scatter, one launch per pass (see fused.h).
//
// One CTA per SM. Warps 0 .. 4G-1 form G worker groups of 128 threads; the last warp is the MMA warp.
//
//   worker group g (thread = sample row r of the group's current tile = TMEM lane r):
//       gathers the tile's features from the hash table into the group's X0 buffer (the chunk-major image a
//       K-major UMMA descriptor reads), then for every layer: arrive on ready[g] -> wait on done[g] -> epilogue
//       (tcgen05.ld of the group's own 64-column accumulator, activation, fp16, store of the next operand);
//       in training the backward chain follows and the tile's dL/d(features) rows go from TMEM straight into
//       red.global.add.f16x2 on the gradient table.
//   MMA warp (one elected thread): polls ready[0..G), and for whichever group is ready issues the tcgen05.mma
//       batch of that group's next layer and commits it to done[g].
//
// G tiles are in flight per SM at different stages, so one group's gathers (L2 latency) and epilogues
// (tcgen05.ld, ALU, st.shared) overlap the other groups' MMAs: the tensor pipe no longer idles behind a single
// serial chain (round 1: 13 % tensor-pipe active, 12 % warps active). Weight-gradient accumulators live in TMEM
// columns shared by all groups - one thread issues every MMA, and MMAs of one thread execute in order, so
// accumulating into the same columns from different groups' tiles is race-free - and leave the SM once.
#include "fused.h"
#include "grid_device.cuh"
#include "mlp_device.cuh"

#include <algorithm>
#include <cstdlib>

namespace ngpb {

constexpr uint32_t kMaxGroups = 7;          // inference: 7 x 128 workers + the MMA warp = 928 threads
constexpr uint32_t kMaxTrainGroups = 4;     // training: shared memory and TMEM hold at most 3-4 tiles
constexpr uint32_t kFusedMaxThreads = kMaxGroups * 128 + 32;
constexpr uint32_t kFusedMaxTrainThreads = kMaxTrainGroups * 128 + 32;
constexpr uint32_t kHiddenBytes = kTile * kMlpWidth * 2;

struct FusedLayout {
	uint32_t groups;
	uint32_t w_a, w_b;            // smem offsets of the chains' weights
	uint32_t prog, n_ops_max;     // the tile's MMA program (MmaOp table)
	uint32_t group0, group_stride;
	uint32_t x0, ha, x1, hb, gy;  // offsets inside a group's region
	uint32_t total;
};

// MMAs of one tile (what the kernel's emit_round() produces, summed over the rounds)
static uint32_t count_ops(const FusedNet& n, bool train) {
	auto fwd = [](uint32_t in, uint32_t hidden) { return in / 16 + hidden * 4; };              // layer 0, then hidden-1 layers + output layer
	auto bwd = [](uint32_t hidden) { return (hidden + 1) * 8 + 1 + (hidden - 1) * 4 + 4; };   // dW per layer + input gradients
	uint32_t c = fwd(n.a_in, n.a_hidden) + (n.nerf ? fwd(n.b_in, n.b_hidden) : 0);
	if (train) c += bwd(n.a_hidden) + (n.nerf ? bwd(n.b_hidden) : 0);
	return c;
}

static FusedLayout make_layout(const FusedNet& n, bool train, uint32_t groups) {
	FusedLayout l{};
	uint32_t o = 0;
	l.w_a = o; o += 2 * chain_params(n.a_in, n.a_hidden);
	l.w_b = o; if (n.nerf) o += 2 * chain_params(n.b_in, n.b_hidden);
	l.prog = o;
	l.n_ops_max = count_ops(n, train);
	o += l.n_ops_max * 32u;
	o = (o + 127u) & ~127u;
	l.group0 = o;
	uint32_t g = 0;
	l.x0 = g; g += kTile * n.a_in * 2;
	l.ha = g; g += train ? n.a_hidden * kHiddenBytes : kHiddenBytes;
	if (train) {
		l.x1 = g; if (n.nerf) g += kTile * n.b_in * 2;
		l.hb = g; if (n.nerf) g += n.b_hidden * kHiddenBytes;
		l.gy = g; g += kTile * 16 * 2;
	} else {
		l.x1 = l.ha; l.hb = l.ha; l.gy = l.ha;
	}
	l.group_stride = (g + 127u) & ~127u;
	l.groups = groups;
	l.total = l.group0 + groups * l.group_stride;
	return l;
}

static uint32_t dw_columns(const FusedNet& n) {
	uint32_t c = n.a_in + (n.a_hidden - 1) * 64 + 16;
	if (n.nerf) c += n.b_in + (n.b_hidden - 1) * 64 + 16;
	return c;
}

static uint32_t pick_groups(const FusedNet& n, bool train) {
	static const int forced = getenv("NGPB_FUSED_GROUPS") ? atoi(getenv("NGPB_FUSED_GROUPS")) : 0;
	uint32_t best = 0;
	for (uint32_t g = 1; g <= (train ? kMaxTrainGroups : kMaxGroups); ++g) {
		if (make_layout(n, train, g).total > 227u * 1024u - 512u) break;   // 227 KB per CTA less the kernel's static shared memory
		if (64 * g + (train ? dw_columns(n) : 0) > 512) break;
		best = g;
	}
	if (forced > 0) best = std::min(best, (uint32_t)forced);
	return best;
}

bool fused_grid_supported(const FusedNet& n, const GridParams& g) {
	static const bool off = getenv("NGPB_FUSED") && atoi(getenv("NGPB_FUSED")) == 0;
	if (off) return false;
	const uint32_t F = g.n_features_per_level;
	if (!((g.n_dims == 3 && (F == 2 || F == 4)) || (g.n_dims == 2 && F == 2 && !n.nerf))) return false;
	if (g.interpolation_type != (int)InterpolationType::Linear) return false;
	if ((g.n_levels * F) % 8 != 0 || g.n_levels * F > n.a_in) return false;
	return pick_groups(n, true) >= 1 && pick_groups(n, false) >= 1;
}

// ---------------------------------------------------------------------------------------------------------
NGPB_D void mbar_arrive(uint64_t* bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
NGPB_D bool mbar_test(uint64_t* bar, uint32_t parity) {
	uint32_t ok;
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
		"selp.u32 %0, 1, 0, p;\n"
		"}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
	return ok != 0;
}

struct Chain {
	uint32_t in_width, n_hidden, relu;
	uint32_t w;          // smem address of the chain's weights
	uint32_t h;          // smem address of hidden buffer 0
	uint32_t h_stride;   // bytes between hidden buffers (0: every layer shares one)
	uint8_t* h_ptr;      // generic pointer of hidden buffer 0
	uint32_t dw_col;     // first TMEM column of the chain's dW accumulators
};
NGPB_D uint32_t chain_w(const Chain& ch, uint32_t layer) {
	return ch.w + (layer == 0 ? 0u : 2u * kMlpWidth * ch.in_width + (layer - 1) * 2u * kMlpWidth * kMlpWidth);
}
NGPB_D uint32_t chain_dw_col(const Chain& ch, uint32_t layer) {
	return ch.dw_col + (layer == 0 ? 0u : ch.in_width + (layer - 1) * kMlpWidth);
}

// ---- MMA warp: the program of one tile -------------------------------------------------------------------
// Building descriptors and walking the layer structure costs a few hundred dependent instructions per round; with
// one thread issuing for every group that latency would sit on the critical path of every layer of every tile
// (measured: the first version of this kernel, which did exactly that, ran 10 k cycles per round). So the CTA
// compiles the tile's MMAs ONCE into a table in shared memory - operand descriptors for group 0, with flags that
// say which fields move with the group - and the MMA thread's inner loop is load, relocate, tcgen05.mma.
struct alignas(16) MmaOp {
	uint64_t desc_a, desc_b;
	uint32_t tmem_d, idesc;
	uint32_t flags;      // bit 0/1/2: desc_a / desc_b / tmem_d move with the group; bit 3: accumulate; bit 4: accumulate iff
	                     // the dW accumulator `layer_bit` (bits 8..15) has been written before; bit 5: this op marks it written
	uint32_t pad;
};
constexpr uint32_t kOpARel = 1u, kOpBRel = 2u, kOpDRel = 4u, kOpAcc = 8u, kOpAccIfStarted = 16u, kOpMarksStarted = 32u;
constexpr uint32_t kMaxRounds = 24;

struct Emitter {
	MmaOp* ops;
	uint32_t n;
	NGPB_D void emit(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t flags) {
		MmaOp o;
		o.desc_a = da; o.desc_b = db; o.tmem_d = d; o.idesc = idesc; o.flags = flags; o.pad = 0;
		if (ops) ops[n] = o;
		++n;
	}
	// D[128 x N] = X[128 x K] W[N x K]^T  (mma_xwT)
	NGPB_D void xwT(uint32_t acc, uint32_t x, uint32_t w, uint32_t K, uint32_t N) {
		const uint32_t idesc = instr_desc(128, N, false, false);
		for (uint32_t k = 0; k < K / 16; ++k)
			emit(acc, smem_desc(x + 2 * k * (kTileRows * 16), kTileRows * 16, 128), smem_desc(w + 2 * k * (N * 16), N * 16, 128), idesc, kOpARel | kOpDRel | (k ? kOpAcc : 0u));
	}
	// D[128 x Kin] = dY[128 x Nout] W[Nout x Kin]  (mma_dyW)
	NGPB_D void dyW(uint32_t acc, uint32_t dy, uint32_t w, uint32_t Nout, uint32_t Kin) {
		const uint32_t idesc = instr_desc(128, Kin, false, true);
		for (uint32_t k = 0; k < Nout / 16; ++k)
			emit(acc, smem_desc(dy + 2 * k * (kTileRows * 16), kTileRows * 16, 128), smem_desc(w + 16 * k * 16, 128, Nout * 16), idesc, kOpARel | kOpDRel | (k ? kOpAcc : 0u));
	}
	// D[64 x Nf] (+)= P[128 x 64]^T Q[128 x Nf]  (mma_pTq), accumulated over every tile of the CTA
	NGPB_D void pTq(uint32_t dw, uint32_t p, uint32_t q, uint32_t Nf, uint32_t layer_bit) {
		const uint32_t idesc = instr_desc(64, Nf, true, true);
		for (uint32_t k = 0; k < kTileRows / 16; ++k)
			emit(dw, smem_desc(p + 16 * k * 16, 128, kTileRows * 16), smem_desc(q + 16 * k * 16, 128, kTileRows * 16), idesc,
				kOpARel | kOpBRel | (layer_bit << 8) | (k ? kOpAcc : (kOpAccIfStarted | kOpMarksStarted)));
	}
};

// ---- worker side ----------------------------------------------------------------------------------------
struct Worker {
	uint32_t tmem_row;   // TMEM address of this thread's row in the group's accumulator
	uint32_t row;
	uint64_t* ready;
	uint64_t* done;
	uint32_t phase;
	uint32_t group;
	uint32_t round;      // round of the current tile (self-issuing groups)
	bool self_issue;     // the group's own thread 0 issues its MMAs (inference: no accumulator is shared between groups)
};
NGPB_D void group_barrier(uint32_t id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
// publish this thread's smem writes / finished TMEM reads, then run the round's MMAs: handed to the MMA warp
// (training), or issued by the group's own thread 0 (inference)
template <typename I>
NGPB_D void grp_round(Worker& w, I& issue) {
	fence_proxy_async();
	tc_fence_before();
	if (w.self_issue) {
		group_barrier(1 + w.group);
		if (w.row == 0) {
			tc_fence_after();
			issue(w.round);
			umma_commit(w.done);
		}
	} else {
		__syncwarp();
		if ((w.row & 31u) == 0) mbar_arrive(w.ready);   // one arrival per warp (128 serialised arrivals cost ~400 cycles a round)
	}
	++w.round;
	mbar_wait(w.done, w.phase);
	w.phase ^= 1;
	tc_fence_after();
}
NGPB_D void epi_store(const Worker& w, uint8_t* dst_tile, uint32_t ncols, bool relu) {
	for (uint32_t c0 = 0; c0 < ncols; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(w.tmem_row + c0, v);
		tmem_wait_ld();
		uint32_t h[8];
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) {
			float a = __uint_as_float(v[2 * k]), b = __uint_as_float(v[2 * k + 1]);
			if (relu) { a = fmaxf(a, 0.0f); b = fmaxf(b, 0.0f); }
			h[k] = pack2(a, b);
		}
		*reinterpret_cast<uint4*>(dst_tile + chunk_offset(kTile, w.row, c0 / 8)) = make_uint4(h[0], h[1], h[2], h[3]);
		*reinterpret_cast<uint4*>(dst_tile + chunk_offset(kTile, w.row, c0 / 8 + 1)) = make_uint4(h[4], h[5], h[6], h[7]);
	}
}
// in place: tile row <- acc * (tile row > 0) (ReLU) or acc
NGPB_D void epi_mask(const Worker& w, uint8_t* act_tile, bool relu) {
	for (uint32_t c0 = 0; c0 < kMlpWidth; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(w.tmem_row + c0, v);
		uint4* p0 = reinterpret_cast<uint4*>(act_tile + chunk_offset(kTile, w.row, c0 / 8));
		uint4* p1 = reinterpret_cast<uint4*>(act_tile + chunk_offset(kTile, w.row, c0 / 8 + 1));
		const uint4 a0 = *p0, a1 = *p1;
		const uint32_t act[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
		tmem_wait_ld();
		uint32_t h[8];
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) {
			float a = __uint_as_float(v[2 * k]), b = __uint_as_float(v[2 * k + 1]);
			if (relu) {
				const float2 f = unpack2(act[k]);
				a = f.x > 0.0f ? a : 0.0f;
				b = f.y > 0.0f ? b : 0.0f;
			}
			h[k] = pack2(a, b);
		}
		*p0 = make_uint4(h[0], h[1], h[2], h[3]);
		*p1 = make_uint4(h[4], h[5], h[6], h[7]);
	}
}
// chain input tile written by the caller -> the chain's 16 output columns in acc[0, 16)
template <typename I>
NGPB_D void worker_forward(Worker& w, const Chain& ch, I& issue) {
	for (uint32_t j = 0; j < ch.n_hidden; ++j) {
		grp_round(w, issue);
		epi_store(w, ch.h_ptr + j * ch.h_stride, kMlpWidth, ch.relu);
	}
	grp_round(w, issue);
}
// gradient seed tile written by the caller -> dL/d(chain input) in acc; hidden buffers become their gradients
template <typename I>
NGPB_D void worker_backward(Worker& w, const Chain& ch, I& issue) {
	for (uint32_t s = 0; s < ch.n_hidden; ++s) {
		grp_round(w, issue);
		epi_mask(w, ch.h_ptr + (ch.n_hidden - 1 - s) * ch.h_stride, ch.relu);
	}
	grp_round(w, issue);
}

// dW accumulators -> this CTA's slice of the partials. M = 64 accumulators keep row r in TMEM lane (r % 16) + 32 (r / 16).
NGPB_D void store_dw(uint32_t tmem_lane_base, uint32_t t, uint32_t col, uint32_t ncols, float* __restrict__ dst, bool transposed) {
	const uint32_t warp = t / 32, lane = t % 32;
	const uint32_t r = warp * 16 + lane;
	for (uint32_t c0 = 0; c0 < ncols; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(tmem_lane_base + col + c0, v);
		tmem_wait_ld();
		if (lane < 16) {
#pragma unroll
			for (uint32_t k = 0; k < 16; ++k) {
				if (transposed) dst[(c0 + k) * kMlpWidth + r] = __uint_as_float(v[k]);
				else dst[r * ncols + c0 + k] = __uint_as_float(v[k]);
			}
		}
	}
}
NGPB_D void store_chain_dw(uint32_t tmem_lane_base, uint32_t t, const Chain& ch, float* __restrict__ dst) {
	store_dw(tmem_lane_base, t, chain_dw_col(ch, 0), ch.in_width, dst, false);
	dst += kMlpWidth * ch.in_width;
	for (uint32_t j = 1; j < ch.n_hidden; ++j) {
		store_dw(tmem_lane_base, t, chain_dw_col(ch, j), kMlpWidth, dst, false);
		dst += kMlpWidth * kMlpWidth;
	}
	store_dw(tmem_lane_base, t, chain_dw_col(ch, ch.n_hidden), kMlpPaddedOut, dst, true);
}

// ---------------------------------------------------------------------------------------------------------
// SHAPE: 0 = any supported network (the MMA thread runs the precompiled table); 1 = the NeRF network of
// configs/nerf/base.json (32 -> 64 -> 16 | 32 -> 64 -> 64 -> 16); 2 = the image network of configs/image/base.json
// (32 -> 64 -> 64 -> 16). For 1 and 2 the widths are compile-time constants, so the MMA thread's code for a round is
// straight-line: descriptors are a uniform base plus immediates, and an MMA costs its issuer ~10 instructions.
template <uint32_t SHAPE>
NGPB_D FusedNet specialise_net(const FusedNet& in) {
	FusedNet n = in;
	if (SHAPE == 1) { n.nerf = 1; n.a_in = 32; n.a_hidden = 1; n.b_in = 32; n.b_hidden = 2; }
	if (SHAPE == 2) { n.nerf = 0; n.a_in = 32; n.a_hidden = 2; }
	return n;
}

// Issues a round's MMAs right away for one group (same interface as Emitter, which records them).
struct DirectIssuer {
	uint32_t rel;        // byte offset of the group's buffers from group 0's
	uint32_t acc;        // TMEM address of the group's activation accumulator
	uint32_t tmem;       // TMEM allocation base (dW accumulators are shared)
	uint32_t& dw_mask;
	NGPB_D void xwT(uint32_t acc_col, uint32_t x, uint32_t w, uint32_t K, uint32_t N) { mma_xwT(acc + acc_col, x + rel, w, K, N, false); }
	NGPB_D void dyW(uint32_t acc_col, uint32_t dy, uint32_t w, uint32_t Nout, uint32_t Kin) { mma_dyW(acc + acc_col, dy + rel, w, Nout, Kin); }
	NGPB_D void pTq(uint32_t dw_col, uint32_t p, uint32_t q, uint32_t Nf, uint32_t layer_bit) {
		const bool accumulate = (dw_mask >> layer_bit) & 1u;
		dw_mask |= 1u << layer_bit;
		mma_pTq(tmem + dw_col, p + rel, q + rel, Nf, accumulate);
	}
};

template <uint32_t D, uint32_t F, bool NERF, bool TRAIN, uint32_t SHAPE>
__global__ void __launch_bounds__(TRAIN ? kFusedMaxTrainThreads : kFusedMaxThreads, 1) fused_grid_net_kernel(
	const __grid_constant__ FusedNet net_in, const __grid_constant__ GridParams gp, const __grid_constant__ FusedLayout lay,
	const uint32_t n_max, const uint32_t* __restrict__ n_dev,
	const __half* __restrict__ params, const __half* __restrict__ table,
	const float* __restrict__ input, const uint32_t in_stride,
	__half* __restrict__ out, const uint32_t out_stride,
	const __half* __restrict__ dL_dout, const uint32_t dy_stride,
	__half* __restrict__ grid_grad, float* __restrict__ dw_partials,
	const __half* __restrict__ feat_tiles, __half* __restrict__ dfeat_tiles)
{
	extern __shared__ __align__(128) uint8_t smem[];   // SWIZZLE_NONE operands need 16-byte alignment only
	__shared__ uint64_t bar_ready[kMaxGroups];
	__shared__ uint64_t bar_done[kMaxGroups];
	__shared__ uint32_t tmem_slot;
	__shared__ uint32_t round_begin[kMaxRounds + 1];

	const FusedNet net = specialise_net<SHAPE>(net_in);
	const uint32_t G = lay.groups;
	const uint32_t t = threadIdx.x, warp = t / 32, lane = t % 32;
	const bool is_mma_warp = warp == 4 * G;
	const uint32_t n = n_dev ? min(__ldg(n_dev), n_max) : n_max;
	const uint32_t n_tiles = (n + kTile - 1) / kTile;
	const uint32_t tile_stride = gridDim.x * G;

	if (is_mma_warp) tmem_alloc<512>(&tmem_slot);
	if (t == 0) {
		for (uint32_t g = 0; g < kMaxGroups; ++g) {
			mbar_init(&bar_ready[g], 4);
			mbar_init(&bar_done[g], 1);
		}
		fence_barrier_init();
	}

	const uint32_t sbase = smem_u32(smem);
	const uint32_t dw_base_col = 64 * G;
	const uint32_t RA = net.a_hidden + 1, RB = NERF ? net.b_hidden + 1 : 0;
	const uint32_t rounds_per_tile = TRAIN ? 2 * (RA + RB) : RA + RB;
	const bool want_dx = TRAIN && (grid_grad != nullptr || dfeat_tiles != nullptr);

	// Chains with group g's buffers. TMEM addresses are relative to the allocation (column 0).
	auto make_chains = [&](uint32_t g, Chain& A, Chain& B) {
		const uint32_t off = lay.group0 + g * lay.group_stride;
		A.in_width = net.a_in; A.n_hidden = net.a_hidden; A.relu = net.a_relu;
		A.w = sbase + lay.w_a; A.h = sbase + off + lay.ha; A.h_ptr = smem + off + lay.ha;
		A.h_stride = TRAIN ? kHiddenBytes : 0;
		A.dw_col = dw_base_col;
		B = A;
		if (NERF) {
			B.in_width = net.b_in; B.n_hidden = net.b_hidden; B.relu = net.b_relu;
			B.w = sbase + lay.w_b; B.h = sbase + off + lay.hb; B.h_ptr = smem + off + lay.hb;
			B.dw_col = dw_base_col + net.a_in + (net.a_hidden - 1) * 64 + 16;
		}
	};

	// The MMAs of round r of a tile, for group 0 with TMEM column 0 as the allocation base (relocated at issue time).
	auto emit_round = [&](auto& em, uint32_t r) {
		Chain A, B;
		make_chains(0, A, B);
		const uint32_t off = sbase + lay.group0;
		const uint32_t x0 = off + lay.x0, x1 = off + lay.x1, gy = off + lay.gy;
		auto forward = [&](const Chain& ch, uint32_t x, uint32_t j) {
			if (j == 0) em.xwT(0, x, chain_w(ch, 0), ch.in_width, kMlpWidth);
			else if (j < ch.n_hidden) em.xwT(0, ch.h + (j - 1) * ch.h_stride, chain_w(ch, j), kMlpWidth, kMlpWidth);
			else em.xwT(0, ch.h + (ch.n_hidden - 1) * ch.h_stride, chain_w(ch, ch.n_hidden), kMlpWidth, kMlpPaddedOut);
		};
		// backward step s (0 = output layer ... n_hidden = first layer): dW of the layer, then the gradient of its input
		auto backward = [&](const Chain& ch, uint32_t x, uint32_t g_seed, uint32_t dx_cols, uint32_t s, uint32_t bit_base) {
			const uint32_t layer = ch.n_hidden - s;
			const uint32_t dw = chain_dw_col(ch, layer);
			if (s == 0) {
				em.pTq(dw, ch.h + (ch.n_hidden - 1) * ch.h_stride, g_seed, kMlpPaddedOut, bit_base + layer);   // dW_out^T = H^T dY
				em.dyW(0, g_seed, chain_w(ch, ch.n_hidden), kMlpPaddedOut, kMlpWidth);
			} else if (s < ch.n_hidden) {
				const uint32_t d = ch.h + layer * ch.h_stride, hprev = ch.h + (layer - 1) * ch.h_stride;
				em.pTq(dw, d, hprev, kMlpWidth, bit_base + layer);                                           // dW_j = D^T H_{j-1}
				em.dyW(0, d, chain_w(ch, layer), kMlpWidth, kMlpWidth);
			} else {
				em.pTq(dw, ch.h, x, ch.in_width, bit_base + layer);                                          // dW_0 = D^T X
				if (dx_cols) em.dyW(0, ch.h, chain_w(ch, 0), kMlpWidth, dx_cols);
			}
		};
		if (r < RA) forward(A, x0, r);
		else if (r < RA + RB) forward(B, x1, r - RA);
		else {
			r -= RA + RB;
			if (NERF) {
				if (r < RB) backward(B, x1, gy, 16, r, RA);
				else backward(A, x0, x1, want_dx ? net.a_in : 0, r - RB, 0);
			} else {
				backward(A, x0, gy, want_dx ? net.a_in : 0, r, 0);
			}
		}
	};

	MmaOp* ops = reinterpret_cast<MmaOp*>(smem + lay.prog);
	if (TRAIN && SHAPE == 0 && t < rounds_per_tile) {
		// thread r compiles round r: its first op index is the op count of the rounds before it (dry runs)
		uint32_t begin = 0;
		for (uint32_t r = 0; r < t; ++r) {
			Emitter dry{nullptr, 0};
			emit_round(dry, r);
			begin += dry.n;
		}
		Emitter em{ops + begin, 0};
		emit_round(em, t);
		round_begin[t] = begin;
		if (t + 1 == rounds_per_tile) round_begin[t + 1] = begin + em.n;
	}
	// weights: every thread helps
	for (uint32_t pass = 0; pass < (NERF ? 2u : 1u); ++pass) {
		uint8_t* dst = smem + (pass ? lay.w_b : lay.w_a);
		const __half* src = params + (pass ? net.b_param_off : net.a_param_off);
		const uint32_t in_w = pass ? net.b_in : net.a_in, n_hidden = pass ? net.b_hidden : net.a_hidden;
		load_chain_weights(dst, src, in_w, n_hidden);
	}
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();

	const uint32_t tmem = tmem_slot;

	if (is_mma_warp) {
		if (TRAIN && SHAPE != 0 && lane == 0) {
			// per-group state packed into registers so that the group index can stay a run-time value (one copy of
			// the straight-line round code instead of one per group)
			uint32_t round_pk = 0, parity = 0, live = 0, dw_mask = 0;
			uint64_t iter_pk = 0;
			for (uint32_t g = 0; g < G; ++g) if (blockIdx.x * G + g < n_tiles) live |= 1u << g;
			uint32_t g = 0;
			while (live) {
				if ((live >> g) & 1u) {
					// Training serves the groups in a fixed cyclic order (bit-reproducible dW); inference serves whoever is ready.
					bool ready = true;
					if (TRAIN) mbar_wait(&bar_ready[g], (parity >> g) & 1u);
					else ready = mbar_test(&bar_ready[g], (parity >> g) & 1u);
					if (ready) {
						parity ^= 1u << g;
						tc_fence_after();
						DirectIssuer di{g * lay.group_stride, tmem + 64 * g, tmem, dw_mask};
						const uint32_t r = (round_pk >> (8 * g)) & 0xFFu;
						switch (r) {
							case 0: emit_round(di, 0); break;
							case 1: emit_round(di, 1); break;
							case 2: emit_round(di, 2); break;
							case 3: emit_round(di, 3); break;
							case 4: emit_round(di, 4); break;
							case 5: emit_round(di, 5); break;
							case 6: emit_round(di, 6); break;
							case 7: emit_round(di, 7); break;
							case 8: emit_round(di, 8); break;
							case 9: emit_round(di, 9); break;
							default: break;
						}
						umma_commit(&bar_done[g]);
						if (r + 1 == rounds_per_tile) {
							round_pk &= ~(0xFFu << (8 * g));
							iter_pk += 1ull << (16 * g);
							const uint32_t it = (uint32_t)(iter_pk >> (16 * g)) & 0xFFFFu;
							if (blockIdx.x * G + g + it * tile_stride >= n_tiles) live &= ~(1u << g);
						} else {
							round_pk += 1u << (8 * g);
						}
					}
				}
				g = g + 1 == G ? 0 : g + 1;
			}
		}
		if (TRAIN && SHAPE == 0 && lane == 0) {
			uint32_t tile[kMaxGroups], round[kMaxGroups], parity[kMaxGroups];
			uint32_t live = 0, dw_mask = 0;
#pragma unroll
			for (uint32_t g = 0; g < kMaxGroups; ++g) {
				tile[g] = blockIdx.x * G + g;
				round[g] = 0;
				parity[g] = 0;
				if (g < G && tile[g] < n_tiles) live |= 1u << g;
			}
			while (live) {
				bool any = false;
#pragma unroll
				for (uint32_t g = 0; g < kMaxGroups; ++g) {
					if (!((live >> g) & 1u)) continue;
					// Training serves the groups in a fixed cyclic order: the weight-gradient accumulators are shared, and a
					// fixed order of accumulation makes dW bit-reproducible run to run. Inference serves whoever is ready.
					if (TRAIN) mbar_wait(&bar_ready[g], parity[g]);
					else if (!mbar_test(&bar_ready[g], parity[g])) continue;
					any = true;
					parity[g] ^= 1;
					tc_fence_after();
					const uint64_t smem_rel = (uint64_t)((g * lay.group_stride) >> 4);   // start-address field of a descriptor
					const uint32_t acc = tmem + 64 * g;
					const uint32_t e1 = round_begin[round[g] + 1];
					for (uint32_t e = round_begin[round[g]]; e < e1; e += 4) {
						// four table entries per trip: the loads are issued together, so their latency is paid once
						uint4 lo[4], hi[4];
#pragma unroll
						for (uint32_t u = 0; u < 4; ++u) {
							const uint32_t eu = min(e + u, e1 - 1);
							lo[u] = *reinterpret_cast<const uint4*>(&ops[eu]);
							hi[u] = *(reinterpret_cast<const uint4*>(&ops[eu]) + 1);
						}
#pragma unroll
						for (uint32_t u = 0; u < 4; ++u) {
							if (e + u >= e1) break;
							const uint32_t flags = hi[u].z;
							uint64_t da = (uint64_t)lo[u].x | ((uint64_t)lo[u].y << 32), db = (uint64_t)lo[u].z | ((uint64_t)lo[u].w << 32);
							if (flags & kOpARel) da += smem_rel;
							if (flags & kOpBRel) db += smem_rel;
							const uint32_t d = (flags & kOpDRel) ? acc + hi[u].x : tmem + hi[u].x;
							const uint32_t bit = (flags >> 8) & 0xFFu;
							const bool accumulate = (flags & kOpAcc) || ((flags & kOpAccIfStarted) && ((dw_mask >> bit) & 1u));
							if (flags & kOpMarksStarted) dw_mask |= 1u << bit;
							umma_f16(d, da, db, hi[u].y, accumulate);
						}
					}
					umma_commit(&bar_done[g]);
					if (++round[g] == rounds_per_tile) {
						round[g] = 0;
						tile[g] += tile_stride;
						if (tile[g] >= n_tiles) live &= ~(1u << g);
					}
				}
				(void)any;   // spin: __nanosleep rounds up to about a microsecond, which is a whole round
			}
		}
		__syncwarp();
	} else {
		const uint32_t g = warp / 4;
		const uint32_t row = t - g * 128;
		Worker w;
		w.tmem_row = tmem + ((row & ~31u) << 16) + 64 * g;
		w.row = row;
		w.ready = &bar_ready[g];
		w.done = &bar_done[g];
		w.phase = 0;
		w.group = g;
		w.round = 0;
		w.self_issue = !TRAIN;
		uint32_t unused_mask = 0;
		DirectIssuer own{g * lay.group_stride, tmem + 64 * g, tmem, unused_mask};
		auto issue = [&](uint32_t r) {
			if (SHAPE != 0) {
				switch (r) {
					case 0: emit_round(own, 0); break;
					case 1: emit_round(own, 1); break;
					case 2: emit_round(own, 2); break;
					case 3: emit_round(own, 3); break;
					case 4: emit_round(own, 4); break;
					default: break;
				}
			} else {
				emit_round(own, r);
			}
		};
		Chain A, B;
		make_chains(g, A, B);
		uint8_t* gbase = smem + lay.group0 + g * lay.group_stride;
		uint8_t* x0 = gbase + lay.x0;
		uint8_t* x1 = gbase + lay.x1;
		uint8_t* gy = gbase + lay.gy;
		constexpr uint32_t C = 8 / F;           // levels per 16-byte chunk
		const uint32_t L = gp.n_levels;
		const uint32_t n_feat_chunks = L * F / 8;

		for (uint32_t tile = blockIdx.x * G + g; tile < n_tiles; tile += tile_stride) {
			const uint32_t i = tile * kTile + row;
			const bool valid = i < n;
			w.round = 0;
			const float* in_row = input + (size_t)i * in_stride;
			float x[D];
#pragma unroll
			for (uint32_t d = 0; d < D; ++d) x[d] = valid ? __ldg(in_row + d) : 0.0f;

			if (feat_tiles) {
				// ---- tile-fed: the features were gathered by grid_forward_kernel at full occupancy into the tiled
				// layout, which is this buffer's image: one coalesced 16-byte load per chunk ----
				const uint4* src = reinterpret_cast<const uint4*>(feat_tiles) + (size_t)tile * (net.a_in / 8) * kTile + row;
				uint4 q[8];
#pragma unroll
				for (uint32_t ch = 0; ch < 8; ++ch) if (ch < net.a_in / 8) q[ch] = __ldg(src + ch * kTile);
#pragma unroll
				for (uint32_t ch = 0; ch < 8; ++ch) if (ch < net.a_in / 8) *reinterpret_cast<uint4*>(x0 + chunk_offset(kTile, row, ch)) = q[ch];
			} else {
				// ---- gather: the tile's encoded features, straight into the first UMMA operand ----
#pragma unroll 2
				for (uint32_t ch = 0; ch < n_feat_chunks; ++ch) {
					alignas(16) __half vals[8];
					if (valid) {
#pragma unroll
						for (uint32_t c = 0; c < C; ++c) grid_encode_level<D, F>(gp, ch * C + c, x, table, &vals[c * F]);
					} else {
						*reinterpret_cast<uint4*>(vals) = make_uint4(0u, 0u, 0u, 0u);
					}
					*reinterpret_cast<uint4*>(x0 + chunk_offset(kTile, row, ch)) = *reinterpret_cast<const uint4*>(vals);
				}
				for (uint32_t ch = n_feat_chunks; ch < net.a_in / 8; ++ch) {
					// padding columns carry 1 (tcnn Encoding padding, grid.h:806-811)
					const uint32_t one2 = valid ? 0x3C003C00u : 0u;
					*reinterpret_cast<uint4*>(x0 + chunk_offset(kTile, row, ch)) = make_uint4(one2, one2, one2, one2);
				}

			}

			// ---- forward ----
			worker_forward(w, A, issue);
			uint32_t v[16];
			float sigma_raw = 0.0f;
			if (NERF) {
				// chain-B input row: [density-MLP output (16) | SH(dir) (16)]
				tmem_ld16(w.tmem_row, v);
				const float* crow = in_row + net.dir_offset;
				const float dx = valid ? __ldg(crow) * 2.0f - 1.0f : 0.0f, dy = valid ? __ldg(crow + 1) * 2.0f - 1.0f : 0.0f, dz = valid ? __ldg(crow + 2) * 2.0f - 1.0f : 0.0f;
				float sh[16];
				sh_encode16(net.sh_degree, dx, dy, dz, sh);
				tmem_wait_ld();
				uint32_t h[8];
#pragma unroll
				for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
				sigma_raw = unpack2(h[0]).x; // the fp16 value the reference copies out (nerf_network.h:31-43)
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 0)) = make_uint4(h[0], h[1], h[2], h[3]);
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 1)) = make_uint4(h[4], h[5], h[6], h[7]);
#pragma unroll
				for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(sh[2 * k], sh[2 * k + 1]);
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 2)) = make_uint4(h[0], h[1], h[2], h[3]);
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 3)) = make_uint4(h[4], h[5], h[6], h[7]);
				worker_forward(w, B, issue);
			}

			if (!TRAIN) {
				tmem_ld16(w.tmem_row, v);
				tmem_wait_ld();
				uint32_t h[8];
#pragma unroll
				for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
				if (NERF) {
					// channel 3 carries the density (nerf_network.h:132-138)
					const float2 b = unpack2(h[1]);
					h[1] = pack2(b.x, sigma_raw);
				}
				if (valid) {
					__half* o = out + (size_t)i * out_stride;
					if (out_stride >= 16) {
						reinterpret_cast<uint4*>(o)[0] = make_uint4(h[0], h[1], h[2], h[3]);
						reinterpret_cast<uint4*>(o)[1] = make_uint4(h[4], h[5], h[6], h[7]);
					} else if (out_stride == 4) {
						*reinterpret_cast<uint2*>(o) = make_uint2(h[0], h[1]);
					} else {
						o[0] = __float2half(unpack2(h[0]).x);
					}
				}
				continue;
			}

			// ---- backward ----
			if (NERF) {
				// dL/d(rgb network output): channels 0-2, rest zero (nerf_network.h:45-60)
				const uint2 gq = __ldg(reinterpret_cast<const uint2*>(dL_dout + (size_t)i * dy_stride));
				const float dsigma = unpack2(gq.y).y;
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, row, 0)) = make_uint4(gq.x, gq.y & 0xFFFFu, 0u, 0u);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, row, 1)) = make_uint4(0u, 0u, 0u, 0u);
				worker_backward(w, B, issue);
				// dL/d(density-MLP output) = chain-B input gradient, columns 0-15, + dsigma on column 0
				// (nerf_network.h:62-74); overwrites the first two chunks of x1 in place.
				tmem_ld16(w.tmem_row, v);
				tmem_wait_ld();
				uint32_t h[8];
				h[0] = pack2(__uint_as_float(v[0]) + dsigma, __uint_as_float(v[1]));
#pragma unroll
				for (uint32_t k = 1; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 0)) = make_uint4(h[0], h[1], h[2], h[3]);
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, row, 1)) = make_uint4(h[4], h[5], h[6], h[7]);
				worker_backward(w, A, issue);
			} else {
				const uint4* gsrc = reinterpret_cast<const uint4*>(dL_dout + (size_t)i * dy_stride);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, row, 0)) = __ldg(gsrc);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, row, 1)) = __ldg(gsrc + 1);
				worker_backward(w, A, issue);
			}

			// ---- dL/d(features): tiled to global memory for grid_backward_kernel, or scattered from here ----
			if (dfeat_tiles) {
				uint4* dst = reinterpret_cast<uint4*>(dfeat_tiles) + (size_t)tile * (net.a_in / 8) * kTile + row;
				for (uint32_t c0 = 0; c0 < net.a_in; c0 += 16) {
					tmem_ld16(w.tmem_row + c0, v);
					tmem_wait_ld();
					uint32_t h[8];
#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
					dst[(c0 / 8) * kTile] = make_uint4(h[0], h[1], h[2], h[3]);
					dst[(c0 / 8 + 1) * kTile] = make_uint4(h[4], h[5], h[6], h[7]);
				}
			} else if (grid_grad) {
				for (uint32_t c0 = 0; c0 < L * F; c0 += 16) {
					tmem_ld16(w.tmem_row + c0, v);
					tmem_wait_ld();
					alignas(16) __half gh[16];
#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) reinterpret_cast<uint32_t*>(gh)[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
					grid_scatter_chunk<D, F>(gp, (c0 / 8) * C, x, gh, grid_grad, 1.0f, true, lane);
					if (c0 + 8 < L * F) grid_scatter_chunk<D, F>(gp, (c0 / 8 + 1) * C, x, gh + 8, grid_grad, 1.0f, true, lane);
				}
			}
		}
	}

	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	if (TRAIN && t < 128) {
		// every MMA of this CTA was waited for by the group that owned it
		Chain A, B;
		make_chains(0, A, B);
		const uint32_t lane_base = tmem + ((t & ~31u) << 16);
		float* dst = dw_partials + (size_t)blockIdx.x * (chain_params(net.a_in, net.a_hidden) + (NERF ? chain_params(net.b_in, net.b_hidden) : 0u));
		store_chain_dw(lane_base, t, A, dst);
		if (NERF) store_chain_dw(lane_base, t, B, dst + chain_params(net.a_in, net.a_hidden));
	}
	tc_fence_before();
	__syncthreads();
	if (is_mma_warp) tmem_dealloc<512>(tmem);
}

// Deterministic sum of the per-CTA partials -> fp16 gradients.
__global__ void fused_reduce_dw_kernel(uint32_t n_params, uint32_t n_parts, const float* __restrict__ partials, __half* __restrict__ grad) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_params) return;
	float s = 0.0f;
	for (uint32_t p = 0; p < n_parts; ++p) s += partials[(size_t)p * n_params + i];
	grad[i] = __float2half(s);
}

// Training: the warp-specialised kernel needs one thread to issue every group's MMAs (the weight-gradient accumulators
// are shared), and that thread's instruction latency (~9 cycles x ~140 instructions a round) caps it at the rate of
// the one-tile-per-CTA kernel of mlp.cu with two CTAs per SM (measured: 0.206 vs 0.187 ms for the whole backward pass
// at 2^18 samples). So mlp.cu stays the default there; NGPB_FUSED_TRAIN=1 selects this kernel.
bool fused_grid_train_enabled() {
	static const bool on = getenv("NGPB_FUSED_TRAIN") && atoi(getenv("NGPB_FUSED_TRAIN")) != 0;
	return on;
}

bool fused_grid_gather_in_kernel() {
	static const bool in_kernel = getenv("NGPB_FUSED_GATHER") && atoi(getenv("NGPB_FUSED_GATHER")) != 0;
	return in_kernel;
}

// ---------------------------------------------------------------------------------------------------------
uint32_t fused_grid_train_ctas(const FusedNet& net, uint32_t n) {
	const uint32_t G = std::max(1u, pick_groups(net, true));
	return std::max(1u, std::min((uint32_t)sm_count(), div_round_up(div_round_up(n, kTile), G)));
}

template <bool TRAIN>
static void launch(cudaStream_t stream, const FusedGridNetArgs& a) {
	if (a.n == 0) return;
	if (a.n % kTile != 0) throw std::runtime_error{"ngp-b200 fused MLP: batch size must be a multiple of 128"};
	const uint32_t G = pick_groups(a.net, TRAIN);
	if (G == 0 || !fused_grid_supported(a.net, a.grid)) throw std::runtime_error{"ngp-b200 fused encoder + MLP: unsupported shape"};
	const FusedLayout lay = make_layout(a.net, TRAIN, G);
	const uint32_t n_tiles = a.n / kTile;
	const uint32_t grid = TRAIN ? fused_grid_train_ctas(a.net, a.n) : std::max(1u, std::min((uint32_t)sm_count(), div_round_up(n_tiles, G)));
	const uint32_t threads = 128 * G + 32;
#define NGPB_FUSED_LAUNCH(D_, F_, NERF_, SHAPE_) \
	do { \
		auto k = fused_grid_net_kernel<D_, F_, NERF_, TRAIN, SHAPE_>; \
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lay.total)); \
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)); \
		k<<<grid, threads, lay.total, stream>>>(a.net, a.grid, lay, a.n, a.n_dev, a.params, a.grid_params, a.input, a.in_stride, \
			a.out, a.out_stride, a.dL_dout, a.dy_stride, a.grid_grad, a.dw_partials, a.feat_tiles, a.dfeat_tiles); \
	} while (0)
	const uint32_t D = a.grid.n_dims, F = a.grid.n_features_per_level;
	const bool nerf_base = a.net.nerf && a.net.a_in == 32 && a.net.a_hidden == 1 && a.net.b_in == 32 && a.net.b_hidden == 2;
	const bool image_base = !a.net.nerf && a.net.a_in == 32 && a.net.a_hidden == 2;
	static const bool generic_only = getenv("NGPB_FUSED_GENERIC") && atoi(getenv("NGPB_FUSED_GENERIC"));
	if (D == 3 && F == 4) {
		if (a.net.nerf) { if (nerf_base && !generic_only) NGPB_FUSED_LAUNCH(3, 4, true, 1); else NGPB_FUSED_LAUNCH(3, 4, true, 0); }
		else NGPB_FUSED_LAUNCH(3, 4, false, 0);
	} else if (D == 3 && F == 2) {
		if (a.net.nerf) { if (nerf_base && !generic_only) NGPB_FUSED_LAUNCH(3, 2, true, 1); else NGPB_FUSED_LAUNCH(3, 2, true, 0); }
		else NGPB_FUSED_LAUNCH(3, 2, false, 0);
	} else if (D == 2 && F == 2 && !a.net.nerf) {
		if (image_base && !generic_only) NGPB_FUSED_LAUNCH(2, 2, false, 2); else NGPB_FUSED_LAUNCH(2, 2, false, 0);
	}
	else throw std::runtime_error{"ngp-b200 fused encoder + MLP: unsupported shape"};
#undef NGPB_FUSED_LAUNCH
	NGPB_CUDA_CHECK(cudaGetLastError());
	if (TRAIN) {
		const uint32_t n_w = fused_net_mlp_params(a.net);
		fused_reduce_dw_kernel<<<div_round_up(n_w, 256), 256, 0, stream>>>(n_w, grid, a.dw_partials, a.dL_dparams);
		NGPB_CUDA_CHECK(cudaGetLastError());
	}
}

void fused_grid_inference(cudaStream_t stream, const FusedGridNetArgs& a) { launch<false>(stream, a); }
void fused_grid_train(cudaStream_t stream, const FusedGridNetArgs& a) { launch<true>(stream, a); }

} // namespace ngpb
