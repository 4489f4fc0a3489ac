// tcgen05 / TMEM / mbarrier building blocks for the fused MLP kernels (sm_100a inline PTX).
//
// Shared-memory operand layout used everywhere ("chunk-major"):
//   a tile of R rows x W halves (W % 8 == 0) stores the 16-byte chunk (row r, chunk c = col/8)
//   at byte  c * (R*16) + r * 16.
// Read as a K-major UMMA operand (rows = M or N, cols = K) this is the SWIZZLE_NONE canonical
// layout with 8x16B core matrices: LBO (next K chunk) = R*16, SBO (next 8 rows) = 128.
// Read as an MN-major operand (cols = M or N, rows = K) the very same bytes are canonical too:
// LBO (next 8 K-rows) = 128, SBO (next 8 MN-cols) = R*16. So one activation tile feeds
//   Y = X W^T      (X K-major A, W K-major B)
//   dX = dY W      (dY K-major A, W MN-major B)
//   dW = dY^T X    (dY MN-major A, X MN-major B, K = the tile's 128 samples)
// without any transposed copies. A thread that owns row r writes whole 16-byte chunks; the 8
// threads of a quarter-warp cover 128 contiguous bytes, so the stores are conflict-free.
#pragma once

#include "common.cuh"

namespace ngpb {
namespace tc {

constexpr uint32_t kTileRows = 128;

NGPB_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ---------------------------------------------------------------------------
NGPB_D void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
NGPB_D void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
NGPB_D void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Generic-proxy smem writes -> visible to the async proxy (UMMA operand reads).
NGPB_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM -------------------------------------------------------------------------------
template <uint32_t COLS>
NGPB_D void tmem_alloc(uint32_t* dst_smem) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "n"(COLS) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t COLS>
NGPB_D void tmem_dealloc(uint32_t taddr) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
NGPB_D void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
NGPB_D void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
NGPB_D void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 bit, 16 consecutive columns: thread (lane l of warp w) reads TMEM lane
// 32*(w%4)+l, columns [col, col+16).
NGPB_D void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
		: "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
		  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
		: "r"(taddr) : "memory");
}

// ---- UMMA descriptors ---------------------------------------------------------------------
// Shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (Blackwell).
NGPB_D uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
	uint64_t d = 0;
	d |= (uint64_t)((saddr >> 4) & 0x3FFF);
	d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
	d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
	d |= (uint64_t)1 << 46;
	return d;
}
// Instruction descriptor for kind::f16: fp16 A/B, fp32 accumulate.
NGPB_HD constexpr uint32_t instr_desc(uint32_t M, uint32_t N, bool a_mn_major, bool b_mn_major) {
	return (1u << 4) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

NGPB_D void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, bool accumulate) {
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"setp.ne.b32 p, %4, 0;\n"
		"tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
		"}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"((uint32_t)accumulate) : "memory");
}
// Arrives on `bar` once every MMA issued so far by this thread has completed
// (implies tcgen05.fence::before_thread_sync).
NGPB_D void umma_commit(uint64_t* bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- tile addressing ----------------------------------------------------------------------
NGPB_HD constexpr uint32_t tile_bytes(uint32_t rows, uint32_t width) { return rows * width * 2; }
NGPB_D uint32_t chunk_offset(uint32_t rows, uint32_t row, uint32_t chunk) { return chunk * (rows * 16) + row * 16; }

// One tcgen05.mma costs its issuing thread as many cycles as the instructions around it: a descriptor built from
// scratch per K step (shifts, masks, 64-bit ors) made an MMA ~105 cycles of issue time against an 8-32 cycle
// execution floor (scripts/umma_rate.py), and the issuing thread is the serial resource of these kernels. So each
// view builds its two descriptors once and steps their 14-bit start-address field with one add per MMA; the K loop
// is unrolled with predicated tails.
NGPB_D uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }

// D[128 x N] (+)= X[128 x K] * W[N x K]^T ; X tile has kTileRows rows, W tile has N rows.
NGPB_D void mma_xwT(uint32_t tmem_d, uint32_t x_saddr, uint32_t w_saddr, uint32_t K, uint32_t N, bool accumulate_first) {
	const uint32_t idesc = instr_desc(128, N, false, false);
	const uint64_t da = smem_desc(x_saddr, kTileRows * 16, 128);
	const uint64_t db = smem_desc(w_saddr, N * 16, 128);
	const uint32_t nk = K / 16, b_step = 2 * N * 16;
#pragma unroll
	for (uint32_t k = 0; k < 4; ++k) {
		if (k < nk) umma_f16(tmem_d, desc_advance(da, k * 2 * kTileRows * 16), desc_advance(db, k * b_step), idesc, accumulate_first || k > 0);
	}
}
// D[128 x Kin] = dY[128 x Nout] * W[Nout x Kin] ; W tile (Nout rows, Kin cols) read MN-major.
NGPB_D void mma_dyW(uint32_t tmem_d, uint32_t dy_saddr, uint32_t w_saddr, uint32_t Nout, uint32_t Kin) {
	const uint32_t idesc = instr_desc(128, Kin, false, true);
	const uint64_t da = smem_desc(dy_saddr, kTileRows * 16, 128);
	// B: N = Kin (MN, contiguous), K = Nout rows [16k, 16k+16): start at row 16k.
	const uint64_t db = smem_desc(w_saddr, 128, Nout * 16);
	const uint32_t nk = Nout / 16;
#pragma unroll
	for (uint32_t k = 0; k < 4; ++k) {
		if (k < nk) umma_f16(tmem_d, desc_advance(da, k * 2 * kTileRows * 16), desc_advance(db, k * 16 * 16), idesc, k > 0);
	}
}
// D[Mf x Nf] (+)= P[128 x Mf]^T * Q[128 x Nf]  (sum over the tile's 128 samples). Mf == 64.
NGPB_D void mma_pTq(uint32_t tmem_d, uint32_t p_saddr, uint32_t q_saddr, uint32_t Nf, bool accumulate_first) {
	const uint32_t idesc = instr_desc(64, Nf, true, true);
	const uint64_t da = smem_desc(p_saddr, 128, kTileRows * 16);
	const uint64_t db = smem_desc(q_saddr, 128, kTileRows * 16);
#pragma unroll
	for (uint32_t k = 0; k < kTileRows / 16; ++k) {
		umma_f16(tmem_d, desc_advance(da, k * 16 * 16), desc_advance(db, k * 16 * 16), idesc, accumulate_first || k > 0);
	}
}

} // namespace tc
} // namespace ngpb
