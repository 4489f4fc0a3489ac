// Block-wide exclusive scan used by the NeRF kernels for slot allocation.
#pragma once

#include "common.cuh"

namespace ngpb {

constexpr uint32_t kRayBlock = 128;

// Exclusive scan of `v` over the block; returns this thread's prefix, `total` = block sum.
NGPB_D uint32_t block_exclusive_scan(uint32_t v, uint32_t& total, uint32_t* warp_sums /* [kRayBlock/32 + 1] */) {
	const uint32_t lane = threadIdx.x % 32, warp = threadIdx.x / 32;
	uint32_t inc = v;
#pragma unroll
	for (uint32_t o = 1; o < 32; o <<= 1) {
		const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) warp_sums[warp] = inc;
	__syncthreads();
	uint32_t warp_prefix = 0, sum = 0;
#pragma unroll
	for (uint32_t w = 0; w < kRayBlock / 32; ++w) {
		if (w < warp) warp_prefix += warp_sums[w];
		sum += warp_sums[w];
	}
	total = sum;
	__syncthreads(); // warp_sums may be reused
	return warp_prefix + inc - v;
}

} // namespace ngpb
