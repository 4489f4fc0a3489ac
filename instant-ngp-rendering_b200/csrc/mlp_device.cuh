// Device helpers shared by the fused MLP kernels (mlp.cu: one tile per CTA, features from global memory;
// fused.cu: warp-specialised, hash-grid gather and gradient scatter inside the kernel).
#pragma once

#include "mlp.h"
#include "tc_device.cuh"

namespace ngpb {

using namespace tc;

// ---------------------------------------------------------------------------------------------
NGPB_D void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
NGPB_D void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

NGPB_D uint32_t pack2(float a, float b) {
	__half2 h = __floats2half2_rn(a, b);
	return *reinterpret_cast<uint32_t*>(&h);
}
NGPB_D float2 unpack2(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }

// Global row-major [rows][cols] fp16 -> chunk-major smem tile.
NGPB_D void load_weights_chunk_major(uint8_t* dst, const __half* __restrict__ src, uint32_t rows, uint32_t cols) {
	const uint32_t cpr = cols / 8;
	for (uint32_t i = threadIdx.x; i < rows * cpr; i += blockDim.x) {
		const uint32_t r = i / cpr, c = i % cpr;
		*reinterpret_cast<uint4*>(dst + chunk_offset(rows, r, c)) = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r * cols + c * 8));
	}
}
NGPB_D void load_chain_weights(uint8_t* dst, const __half* __restrict__ src, uint32_t in_width, uint32_t n_hidden) {
	load_weights_chunk_major(dst, src, kMlpWidth, in_width);
	dst += 2 * kMlpWidth * in_width; src += kMlpWidth * in_width;
	for (uint32_t j = 1; j < n_hidden; ++j) {
		load_weights_chunk_major(dst, src, kMlpWidth, kMlpWidth);
		dst += 2 * kMlpWidth * kMlpWidth; src += kMlpWidth * kMlpWidth;
	}
	load_weights_chunk_major(dst, src, kMlpPaddedOut, kMlpWidth);
}

// Spherical harmonics of degree <= 4 in the reference's order and constants
// (tcnn common_device.h:339-365); padding columns come first (spherical_harmonics.h:59-63).
NGPB_D void sh_encode16(uint32_t degree, float x, float y, float z, float (&o)[16]) {
	const uint32_t n = degree * degree;
	const uint32_t pad = 16 - n;
	float s[16];
	const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	s[0] = 0.28209479177387814f;
	s[1] = -0.48860251190291987f * y;
	s[2] = 0.48860251190291987f * z;
	s[3] = -0.48860251190291987f * x;
	s[4] = 1.0925484305920792f * xy;
	s[5] = -1.0925484305920792f * yz;
	s[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
	s[7] = -1.0925484305920792f * xz;
	s[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
	s[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
	s[10] = 2.8906114426405538f * xy * z;
	s[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
	s[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
	s[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
	s[14] = 1.4453057213202769f * z * (x2 - y2);
	s[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
#pragma unroll
	for (uint32_t k = 0; k < 16; ++k) o[k] = k < pad ? 1.0f : s[(k - pad) & 15];
}


} // namespace ngpb
