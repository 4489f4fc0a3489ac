// Fused MLP kernels (tcgen05.mma, accumulators in TMEM, one 128-sample tile per CTA at a time).
//
// fused_net_kernel<NERF, false>  inference: features -> chain A [-> SH concat -> chain B] -> out
// fused_net_kernel<NERF, true>   training:  the same forward, kept in shared memory, followed by
//                                 the whole backward pass of the tile: dL/dW of every layer is
//                                 accumulated in TMEM across all tiles a CTA processes and leaves
//                                 the SM once, dL/d(features) goes out tiled for the grid scatter.
//
// The reference runs forward and backward as separate kernels with every hidden activation
// written to and re-read from global memory, plus one CUTLASS split-K GEMM per layer for dW
// (fully_fused_mlp.cu:735-836). Here no activation leaves the SM.
//
// Thread t of the 128-thread CTA owns sample row t of the tile == TMEM lane t.
#include "mlp_device.cuh"

#include <algorithm>

namespace ngpb {

// ---------------------------------------------------------------------------------------------
// shared-memory carve-up (bytes from the 1024-aligned dynamic base)
struct SmemPlan {
	uint32_t w_a, w_b;      // weights of chain A / B (chunk-major per matrix)
	uint32_t x0[2];         // double-buffered chain-A input tiles
	uint32_t ha;            // chain-A hidden activations (training: a_hidden buffers, inference: 1)
	uint32_t x1;            // chain-B input tile (training only; inference reuses `ha`)
	uint32_t hb;            // chain-B hidden activations (training only)
	uint32_t gy;            // 128 x 16 gradient seed tile (training only)
	uint32_t total;
};

NGPB_HD SmemPlan make_plan(const FusedNet& n, bool train) {
	SmemPlan p{};
	uint32_t o = 0;
	p.w_a = o; o += 2 * (kMlpWidth * n.a_in + (n.a_hidden - 1) * kMlpWidth * kMlpWidth + kMlpPaddedOut * kMlpWidth);
	p.w_b = o; if (n.nerf) o += 2 * (kMlpWidth * n.b_in + (n.b_hidden - 1) * kMlpWidth * kMlpWidth + kMlpPaddedOut * kMlpWidth);
	p.x0[0] = o; o += kTile * n.a_in * 2;
	p.x0[1] = o; o += kTile * n.a_in * 2;
	const uint32_t hbuf = kTile * kMlpWidth * 2;
	p.ha = o; o += train ? n.a_hidden * hbuf : hbuf;
	if (train) {
		p.x1 = o; if (n.nerf) o += kTile * n.b_in * 2;
		p.hb = o; if (n.nerf) o += n.b_hidden * hbuf;
		p.gy = o; o += kTile * 16 * 2;
	} else {
		p.x1 = p.ha; p.hb = p.ha; p.gy = p.ha;
	}
	p.total = o;
	return p;
}

// TMEM columns: [0,64) activation accumulator, then one dW accumulator per layer (training).
NGPB_HD uint32_t tmem_cols_needed(const FusedNet& n, bool train) {
	if (!train) return 64;
	uint32_t c = 64 + n.a_in + (n.a_hidden - 1) * 64 + 16;
	if (n.nerf) c += n.b_in + (n.b_hidden - 1) * 64 + 16;
	return c;
}

// ---------------------------------------------------------------------------------------------
struct TileCtx {
	uint32_t tmem;       // TMEM base (column 0 of the allocation, lane 0)
	uint32_t tmem_row;   // this thread's lane
	uint64_t* bar_mma;
	uint32_t phase;
	uint32_t row;        // threadIdx.x
};

// One "round": thread 0 issues a batch of MMAs and commits; everyone waits for completion.
template <typename F>
NGPB_D void mma_round(TileCtx& c, F&& issue) {
	if (c.row == 0) {
		tc_fence_after();
		issue();
		umma_commit(c.bar_mma);
	}
	mbar_wait(c.bar_mma, c.phase);
	c.phase ^= 1;
	tc_fence_after();
}
// Makes this thread's smem writes visible to the tensor core and orders TMEM reads before the
// next round's MMAs.
NGPB_D void publish() {
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
}

// Accumulator columns [0, ncols) of this thread's row -> (optional ReLU) -> fp16 -> smem tile row.
NGPB_D void epilogue_store(const TileCtx& c, uint8_t* dst_tile, uint32_t ncols, bool relu) {
	for (uint32_t c0 = 0; c0 < ncols; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(c.tmem_row + c0, v);
		tmem_wait_ld();
		uint32_t h[8];
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) {
			float a = __uint_as_float(v[2 * k]), b = __uint_as_float(v[2 * k + 1]);
			if (relu) { a = fmaxf(a, 0.0f); b = fmaxf(b, 0.0f); }
			h[k] = pack2(a, b);
		}
		*reinterpret_cast<uint4*>(dst_tile + chunk_offset(kTile, c.row, c0 / 8)) = make_uint4(h[0], h[1], h[2], h[3]);
		*reinterpret_cast<uint4*>(dst_tile + chunk_offset(kTile, c.row, c0 / 8 + 1)) = make_uint4(h[4], h[5], h[6], h[7]);
	}
}

// Backward activation transfer, in place: tile row <- acc * (tile row > 0) (ReLU) or acc (None).
NGPB_D void epilogue_mask(const TileCtx& c, uint8_t* act_tile, bool relu) {
	for (uint32_t c0 = 0; c0 < kMlpWidth; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(c.tmem_row + c0, v);
		uint4* p0 = reinterpret_cast<uint4*>(act_tile + chunk_offset(kTile, c.row, c0 / 8));
		uint4* p1 = reinterpret_cast<uint4*>(act_tile + chunk_offset(kTile, c.row, c0 / 8 + 1));
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

struct Chain {
	uint32_t in_width, n_hidden, relu;
	uint32_t w;          // smem address (u32) of the chain's weights
	uint32_t h;          // smem address of hidden buffer 0
	uint32_t h_stride;   // bytes between hidden buffers (0 for inference: all layers share one)
	uint8_t* h_ptr;      // generic pointer of hidden buffer 0
	uint32_t dw_col;     // first TMEM column of the chain's dW accumulators (training)
};
NGPB_D uint32_t chain_w(const Chain& ch, uint32_t layer) { // layer 0..n_hidden
	return ch.w + (layer == 0 ? 0u : 2u * kMlpWidth * ch.in_width + (layer - 1) * 2u * kMlpWidth * kMlpWidth);
}
NGPB_D uint32_t chain_dw_col(const Chain& ch, uint32_t layer) {
	return ch.dw_col + (layer == 0 ? 0u : ch.in_width + (layer - 1) * kMlpWidth);
}

// X (in_width cols at smem address x) -> ... -> 16 output columns left in TMEM columns [0,16).
NGPB_D void chain_forward(TileCtx& c, const Chain& ch, uint32_t x) {
	mma_round(c, [&] { mma_xwT(c.tmem, x, chain_w(ch, 0), ch.in_width, kMlpWidth, false); });
	epilogue_store(c, ch.h_ptr, kMlpWidth, ch.relu);
	publish();
	for (uint32_t j = 1; j < ch.n_hidden; ++j) {
		mma_round(c, [&] { mma_xwT(c.tmem, ch.h + (j - 1) * ch.h_stride, chain_w(ch, j), kMlpWidth, kMlpWidth, false); });
		epilogue_store(c, ch.h_ptr + j * ch.h_stride, kMlpWidth, ch.relu);
		publish();
	}
	mma_round(c, [&] { mma_xwT(c.tmem, ch.h + (ch.n_hidden - 1) * ch.h_stride, chain_w(ch, ch.n_hidden), kMlpWidth, kMlpPaddedOut, false); });
}

// gy: 128 x 16 tile holding dL/d(chain output). On return TMEM columns [0, dx_cols) hold
// dL/d(chain input); every hidden buffer has been overwritten by its gradient.
NGPB_D void chain_backward(TileCtx& c, const Chain& ch, uint32_t x, uint32_t gy, uint32_t dx_cols, bool accumulate) {
	const uint32_t h_last = ch.h + (ch.n_hidden - 1) * ch.h_stride;
	mma_round(c, [&] {
		mma_pTq(c.tmem + chain_dw_col(ch, ch.n_hidden), h_last, gy, kMlpPaddedOut, accumulate); // dW_last^T = H^T dY
		mma_dyW(c.tmem, gy, chain_w(ch, ch.n_hidden), kMlpPaddedOut, kMlpWidth);
	});
	epilogue_mask(c, ch.h_ptr + (ch.n_hidden - 1) * ch.h_stride, ch.relu);
	publish();
	for (uint32_t j = ch.n_hidden - 1; j >= 1; --j) {
		const uint32_t d = ch.h + j * ch.h_stride, hprev = ch.h + (j - 1) * ch.h_stride;
		mma_round(c, [&] {
			mma_pTq(c.tmem + chain_dw_col(ch, j), d, hprev, kMlpWidth, accumulate); // dW_j = D^T H_{j-1}
			mma_dyW(c.tmem, d, chain_w(ch, j), kMlpWidth, kMlpWidth);
		});
		epilogue_mask(c, ch.h_ptr + (j - 1) * ch.h_stride, ch.relu);
		publish();
	}
	mma_round(c, [&] {
		mma_pTq(c.tmem + chain_dw_col(ch, 0), ch.h, x, ch.in_width, accumulate); // dW_0 = D^T X
		if (dx_cols) mma_dyW(c.tmem, ch.h, chain_w(ch, 0), kMlpWidth, dx_cols);
	});
}

// dW accumulators -> this CTA's slice of the partials buffer. M = 64 accumulators keep row r in
// TMEM lane (r % 16) + 32 * (r / 16).
NGPB_D void store_dw(const TileCtx& c, uint32_t col, uint32_t ncols, float* __restrict__ dst, bool transposed) {
	const uint32_t warp = c.row / 32, lane = c.row % 32;
	const uint32_t r = warp * 16 + lane;
	for (uint32_t c0 = 0; c0 < ncols; c0 += 16) {
		uint32_t v[16];
		tmem_ld16(c.tmem_row + col + c0, v);
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
NGPB_D void store_chain_dw(const TileCtx& c, const Chain& ch, float* __restrict__ dst) {
	store_dw(c, chain_dw_col(ch, 0), ch.in_width, dst, false);
	dst += kMlpWidth * ch.in_width;
	for (uint32_t j = 1; j < ch.n_hidden; ++j) {
		store_dw(c, chain_dw_col(ch, j), kMlpWidth, dst, false);
		dst += kMlpWidth * kMlpWidth;
	}
	store_dw(c, chain_dw_col(ch, ch.n_hidden), kMlpPaddedOut, dst, true);
}

// ---------------------------------------------------------------------------------------------
template <bool NERF, bool TRAIN, uint32_t TMEM_COLS>
__global__ void __launch_bounds__(128, TRAIN ? 2 : 4) fused_net_kernel(
	const __grid_constant__ FusedNet net, const uint32_t n_tiles_max, const uint32_t* __restrict__ n_dev,
	const __half* __restrict__ params, const __half* __restrict__ feat,
	const float* __restrict__ coords, const uint32_t coord_stride,
	__half* __restrict__ out, const uint32_t out_stride,
	const __half* __restrict__ dL_dout, const uint32_t dy_stride,
	__half* __restrict__ dfeat, float* __restrict__ dw_partials)
{
	extern __shared__ __align__(1024) uint8_t smem[];
	__shared__ uint64_t bar_mma;
	__shared__ uint64_t bar_x[2];
	__shared__ uint32_t tmem_slot;

	const SmemPlan plan = make_plan(net, TRAIN);
	const uint32_t t = threadIdx.x;
	uint32_t n_tiles = n_tiles_max;
	if (n_dev) n_tiles = min(n_tiles_max, (__ldg(n_dev) + kTile - 1) / kTile);

	load_chain_weights(smem + plan.w_a, params + net.a_param_off, net.a_in, net.a_hidden);
	if (NERF) load_chain_weights(smem + plan.w_b, params + net.b_param_off, net.b_in, net.b_hidden);

	if (t / 32 == 0) tmem_alloc<TMEM_COLS>(&tmem_slot);
	if (t == 0) {
		mbar_init(&bar_mma, 1);
		mbar_init(&bar_x[0], 1);
		mbar_init(&bar_x[1], 1);
		fence_barrier_init();
	}
	fence_proxy_async();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();

	TileCtx c;
	c.tmem = tmem_slot;
	c.tmem_row = c.tmem + ((t & ~31u) << 16); // lane field = first lane of this warp's quarter
	c.bar_mma = &bar_mma;
	c.phase = 0;
	c.row = t;

	const uint32_t sbase = smem_u32(smem);
	Chain A;
	A.in_width = net.a_in; A.n_hidden = net.a_hidden; A.relu = net.a_relu;
	A.w = sbase + plan.w_a; A.h = sbase + plan.ha; A.h_ptr = smem + plan.ha;
	A.h_stride = TRAIN ? kTile * kMlpWidth * 2 : 0;
	A.dw_col = 64;
	Chain B = A;
	if (NERF) {
		B.in_width = net.b_in; B.n_hidden = net.b_hidden; B.relu = net.b_relu;
		B.w = sbase + plan.w_b; B.h = sbase + plan.hb; B.h_ptr = smem + plan.hb;
		B.dw_col = 64 + net.a_in + (net.a_hidden - 1) * 64 + 16;
	}
	uint8_t* x1 = smem + plan.x1;
	uint8_t* gy = smem + plan.gy;

	const uint32_t x_bytes = kTile * net.a_in * 2;
	uint32_t x_phase[2] = {0, 0};
	if (t == 0 && blockIdx.x < n_tiles) {
		mbar_expect_tx(&bar_x[0], x_bytes);
		bulk_load(smem + plan.x0[0], reinterpret_cast<const uint8_t*>(feat) + (size_t)blockIdx.x * x_bytes, x_bytes, &bar_x[0]);
	}

	uint32_t it = 0;
	for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
		const uint32_t buf = it & 1;
		const uint32_t i = tile * kTile + t; // this thread's sample
		if (t == 0) {
			const uint32_t next = tile + gridDim.x;
			if (next < n_tiles) {
				mbar_expect_tx(&bar_x[buf ^ 1], x_bytes);
				bulk_load(smem + plan.x0[buf ^ 1], reinterpret_cast<const uint8_t*>(feat) + (size_t)next * x_bytes, x_bytes, &bar_x[buf ^ 1]);
			}
			mbar_wait(&bar_x[buf], x_phase[buf]);
			x_phase[buf] ^= 1;
		}
		const uint32_t x0 = sbase + plan.x0[buf];

		// ------------------------------ forward ------------------------------
		chain_forward(c, A, x0);

		uint32_t v[16];
		float sigma_raw = 0.0f;
		if (NERF) {
			// chain-B input row: [density-MLP output (16) | SH(dir) (16)]
			tmem_ld16(c.tmem_row, v);
			const float* crow = coords + (size_t)i * coord_stride + net.dir_offset;
			const float dx = __ldg(crow) * 2.0f - 1.0f, dy = __ldg(crow + 1) * 2.0f - 1.0f, dz = __ldg(crow + 2) * 2.0f - 1.0f;
			float sh[16];
			sh_encode16(net.sh_degree, dx, dy, dz, sh);
			tmem_wait_ld();
			uint32_t h[8];
#pragma unroll
			for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
			sigma_raw = unpack2(h[0]).x; // the fp16 value the reference copies out (nerf_network.h:31-43)
			*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 0)) = make_uint4(h[0], h[1], h[2], h[3]);
			*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 1)) = make_uint4(h[4], h[5], h[6], h[7]);
#pragma unroll
			for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(sh[2 * k], sh[2 * k + 1]);
			*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 2)) = make_uint4(h[0], h[1], h[2], h[3]);
			*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 3)) = make_uint4(h[4], h[5], h[6], h[7]);
			publish();
			chain_forward(c, B, sbase + plan.x1);
		}

		if (!TRAIN) {
			tmem_ld16(c.tmem_row, v);
			tmem_wait_ld();
			uint32_t h[8];
#pragma unroll
			for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
			if (NERF) {
				// channel 3 carries the density (nerf_network.h:132-138)
				const float2 b = unpack2(h[1]);
				h[1] = pack2(b.x, sigma_raw);
			}
			__half* o = out + (size_t)i * out_stride;
			if (out_stride >= 16) {
				reinterpret_cast<uint4*>(o)[0] = make_uint4(h[0], h[1], h[2], h[3]);
				reinterpret_cast<uint4*>(o)[1] = make_uint4(h[4], h[5], h[6], h[7]);
			} else if (out_stride == 4) {
				*reinterpret_cast<uint2*>(o) = make_uint2(h[0], h[1]);
			} else {
				const float2 a = unpack2(h[0]);
				o[0] = __float2half(a.x);
			}
			publish(); // TMEM reads done before the next tile's first MMA overwrites the accumulator
			continue;
		}

		// ------------------------------ backward ------------------------------
		if (TRAIN) {
			const bool accumulate = it > 0;
			float dsigma = 0.0f;
			if (NERF) {
				// dL/d(rgb network output): channels 0-2, rest zero (nerf_network.h:45-60)
				const uint2 g = __ldg(reinterpret_cast<const uint2*>(dL_dout + (size_t)i * dy_stride));
				dsigma = unpack2(g.y).y;
				const uint32_t gb = g.y & 0xFFFFu;
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, t, 0)) = make_uint4(g.x, gb, 0u, 0u);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, t, 1)) = make_uint4(0u, 0u, 0u, 0u);
				publish();
				chain_backward(c, B, sbase + plan.x1, sbase + plan.gy, 16, accumulate);
				// dL/d(density-MLP output) = chain-B input gradient, columns 0-15, + dsigma on column 0
				// (nerf_network.h:62-74); overwrites the first two chunks of x1 in place.
				tmem_ld16(c.tmem_row, v);
				tmem_wait_ld();
				uint32_t h[8];
				h[0] = pack2(__uint_as_float(v[0]) + dsigma, __uint_as_float(v[1]));
#pragma unroll
				for (uint32_t k = 1; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 0)) = make_uint4(h[0], h[1], h[2], h[3]);
				*reinterpret_cast<uint4*>(x1 + chunk_offset(kTile, t, 1)) = make_uint4(h[4], h[5], h[6], h[7]);
				publish();
				chain_backward(c, A, x0, sbase + plan.x1, dfeat ? net.a_in : 0, accumulate);
			} else {
				const uint4* g = reinterpret_cast<const uint4*>(dL_dout + (size_t)i * dy_stride);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, t, 0)) = __ldg(g);
				*reinterpret_cast<uint4*>(gy + chunk_offset(kTile, t, 1)) = __ldg(g + 1);
				publish();
				chain_backward(c, A, x0, sbase + plan.gy, dfeat ? net.a_in : 0, accumulate);
			}
			if (dfeat) {
				uint8_t* dst = reinterpret_cast<uint8_t*>(dfeat) + (size_t)tile * x_bytes;
				for (uint32_t c0 = 0; c0 < net.a_in; c0 += 16) {
					tmem_ld16(c.tmem_row + c0, v);
					tmem_wait_ld();
					uint32_t h[8];
#pragma unroll
					for (uint32_t k = 0; k < 8; ++k) h[k] = pack2(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
					*reinterpret_cast<uint4*>(dst + chunk_offset(kTile, t, c0 / 8)) = make_uint4(h[0], h[1], h[2], h[3]);
					*reinterpret_cast<uint4*>(dst + chunk_offset(kTile, t, c0 / 8 + 1)) = make_uint4(h[4], h[5], h[6], h[7]);
				}
			}
			publish();
		}
	}

	if (TRAIN) {
		// All MMAs of this CTA were waited for at the end of the last round.
		float* dst = dw_partials + (size_t)blockIdx.x * (chain_params(net.a_in, net.a_hidden) + (NERF ? chain_params(net.b_in, net.b_hidden) : 0u));
		if (it > 0) {
			store_chain_dw(c, A, dst);
			if (NERF) store_chain_dw(c, B, dst + chain_params(net.a_in, net.a_hidden));
		}
	}

	tc_fence_before();
	__syncthreads();
	if (t / 32 == 0) tmem_dealloc<TMEM_COLS>(c.tmem);
}

// Deterministic sum of the per-CTA partials -> fp16 gradients.
__global__ void reduce_dw_kernel(uint32_t n_params, uint32_t n_parts, const float* __restrict__ partials, __half* __restrict__ grad) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_params) return;
	float s = 0.0f;
	for (uint32_t p = 0; p < n_parts; ++p) s += partials[(size_t)p * n_params + i];
	grad[i] = __float2half(s);
}

// ---------------------------------------------------------------------------------------------
void fused_check_supported(const FusedNet& n) {
	auto bad_in = [](uint32_t w) { return w == 0 || w % 16 != 0 || w > 64; };
	if (bad_in(n.a_in) || (n.nerf && bad_in(n.b_in))) throw std::runtime_error{"ngp-b200 fused MLP: input width must be a multiple of 16 and at most 64"};
	if (n.a_hidden < 1 || (n.nerf && n.b_hidden < 1)) throw std::runtime_error{"FullyFusedMLP requires at least 1 hidden layer (3 layers in total)."};
	if (n.nerf && (n.sh_degree < 1 || n.sh_degree > 4)) throw std::runtime_error{"ngp-b200 NerfNetwork: SphericalHarmonics degree must be 1..4"};
	if (tmem_cols_needed(n, true) > 512) throw std::runtime_error{"ngp-b200 fused MLP: too many layers for the TMEM gradient accumulators"};
	if (make_plan(n, true).total > 220 * 1024) throw std::runtime_error{"ngp-b200 fused MLP: too many layers for shared memory"};
}

template <typename K>
static void set_smem(K kernel, uint32_t bytes) {
	NGPB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
	// Always the largest shared-memory split: two training CTAs need 196 KB to the byte - exactly one
	// of the driver's carve-out steps - which would leave no room for a block of the ray-march kernel
	// that runs beside them on the prefetch stream (an SM keeps its split while any block is resident).
	NGPB_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
}

void fused_inference(cudaStream_t stream, const FusedInferenceArgs& a) {
	if (a.n == 0) return;
	if (a.n % kTile != 0) throw std::runtime_error{"ngp-b200 fused MLP: batch size must be a multiple of 128"};
	const uint32_t n_tiles = a.n / kTile;
	const SmemPlan plan = make_plan(a.net, false);
	const uint32_t ctas_per_sm = std::max(1u, std::min(8u, (220u * 1024u) / (plan.total + 1024u)));
	const uint32_t grid = std::min(n_tiles, (uint32_t)sm_count() * ctas_per_sm);
	if (a.net.nerf) {
		auto k = fused_net_kernel<true, false, 64>;
		set_smem(k, plan.total);
		k<<<grid, 128, plan.total, stream>>>(a.net, n_tiles, a.n_dev, a.params, a.feat, a.coords, a.coord_stride, a.out, a.out_stride, nullptr, 0, nullptr, nullptr);
	} else {
		auto k = fused_net_kernel<false, false, 64>;
		set_smem(k, plan.total);
		k<<<grid, 128, plan.total, stream>>>(a.net, n_tiles, a.n_dev, a.params, a.feat, a.coords, a.coord_stride, a.out, a.out_stride, nullptr, 0, nullptr, nullptr);
	}
	NGPB_CUDA_CHECK(cudaGetLastError());
}

static uint32_t train_ctas_per_sm(const FusedNet& net) {
	const SmemPlan plan = make_plan(net, true);
	const uint32_t by_smem = std::max(1u, (220u * 1024u) / (plan.total + 1024u));
	const uint32_t by_tmem = tmem_cols_needed(net, true) <= 256 ? 2u : 1u;
	return std::min(by_smem, by_tmem);
}

uint32_t fused_train_grid(const FusedNet& net, uint32_t n) {
	return std::max(1u, std::min(div_round_up(n, kTile), (uint32_t)sm_count() * train_ctas_per_sm(net)));
}

void fused_train(cudaStream_t stream, const FusedTrainArgs& a) {
	if (a.n == 0) return;
	if (a.n % kTile != 0) throw std::runtime_error{"ngp-b200 fused MLP: batch size must be a multiple of 128"};
	const uint32_t n_tiles = a.n / kTile;
	const SmemPlan plan = make_plan(a.net, true);
	const uint32_t grid = fused_train_grid(a.net, a.n);
	const bool big = tmem_cols_needed(a.net, true) > 256;
#define NGPB_LAUNCH_TRAIN(NERF_, COLS_) \
	do { \
		auto k = fused_net_kernel<NERF_, true, COLS_>; \
		set_smem(k, plan.total); \
		k<<<grid, 128, plan.total, stream>>>(a.net, n_tiles, nullptr, a.params, a.feat, a.coords, a.coord_stride, nullptr, 0, a.dL_dout, a.dy_stride, a.dfeat, a.dw_partials); \
	} while (0)
	if (a.net.nerf) { if (big) NGPB_LAUNCH_TRAIN(true, 512); else NGPB_LAUNCH_TRAIN(true, 256); }
	else { if (big) NGPB_LAUNCH_TRAIN(false, 512); else NGPB_LAUNCH_TRAIN(false, 256); }
#undef NGPB_LAUNCH_TRAIN
	NGPB_CUDA_CHECK(cudaGetLastError());
	const uint32_t n_w = fused_net_mlp_params(a.net);
	reduce_dw_kernel<<<div_round_up(n_w, 256), 256, 0, stream>>>(n_w, grid, a.dw_partials, a.dL_dparams);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
NGPB_D size_t tiled_index(uint32_t i, uint32_t col, uint32_t width) {
	return ((size_t)(i / kTile) * (width / 8) + col / 8) * (kTile * 8) + (i % kTile) * 8 + col % 8;
}

__global__ void identity_encode_tiled_kernel(uint32_t n, uint32_t n_pad_rows, const float* __restrict__ input, uint32_t in_stride, uint32_t n_dims,
	float scale, float offset, __half* __restrict__ feat, uint32_t width)
{
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; // one 16-byte chunk each
	const uint32_t cpr = width / 8;
	if (e >= n_pad_rows * cpr) return;
	// consecutive threads -> consecutive rows of the same chunk column (coalesced stores)
	const uint32_t tile = e / (kTile * cpr), rem = e % (kTile * cpr);
	const uint32_t ch = rem / kTile, r = rem % kTile;
	const uint32_t i = tile * kTile + r;
	alignas(16) __half v[8];
#pragma unroll
	for (uint32_t k = 0; k < 8; ++k) {
		const uint32_t col = ch * 8 + k;
		float f = 1.0f;
		if (i >= n) f = 0.0f;
		else if (col < n_dims) f = input[(size_t)i * in_stride + col] * scale + offset;
		v[k] = __float2half(f);
	}
	*reinterpret_cast<uint4*>(feat + tiled_index(i, ch * 8, width)) = *reinterpret_cast<const uint4*>(v);
}

void identity_encode_tiled(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, uint32_t n_dims,
	float scale, float offset, __half* feat, uint32_t width)
{
	if (n == 0) return;
	const uint32_t rows = next_multiple(n, kTile);
	const uint32_t total = rows * (width / 8);
	identity_encode_tiled_kernel<<<div_round_up(total, 256), 256, 0, stream>>>(n, rows, input, in_stride, n_dims, scale, offset, feat, width);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

__global__ void tiled_aos_kernel(uint32_t n, uint32_t n_pad_rows, __half* __restrict__ tiled, uint32_t width, __half* __restrict__ aos, uint32_t stride, bool to_aos) {
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t cpr = width / 8;
	if (e >= n_pad_rows * cpr) return;
	const uint32_t tile = e / (kTile * cpr), rem = e % (kTile * cpr);
	const uint32_t ch = rem / kTile, r = rem % kTile;
	const uint32_t i = tile * kTile + r;
	__half* tp = tiled + tiled_index(i, ch * 8, width);
	if (to_aos) {
		if (i >= n) return;
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) aos[(size_t)i * stride + ch * 8 + k] = tp[k];
	} else {
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) tp[k] = i < n ? aos[(size_t)i * stride + ch * 8 + k] : __float2half(0.0f);
	}
}

void tiled_to_aos(cudaStream_t stream, uint32_t n, const __half* tiled, uint32_t width, __half* aos, uint32_t stride) {
	if (n == 0) return;
	const uint32_t rows = next_multiple(n, kTile);
	tiled_aos_kernel<<<div_round_up(rows * (width / 8), 256), 256, 0, stream>>>(n, rows, const_cast<__half*>(tiled), width, aos, stride, true);
	NGPB_CUDA_CHECK(cudaGetLastError());
}
void aos_to_tiled(cudaStream_t stream, uint32_t n, const __half* aos, uint32_t stride, __half* tiled, uint32_t width) {
	if (n == 0) return;
	const uint32_t rows = next_multiple(n, kTile);
	tiled_aos_kernel<<<div_round_up(rows * (width / 8), 256), 256, 0, stream>>>(n, rows, tiled, width, const_cast<__half*>(aos), stride, false);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
