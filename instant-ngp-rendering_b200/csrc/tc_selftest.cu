// Self-test kernels for the tcgen05 building blocks in tc_device.cuh: each of the three
// operand views (X W^T, dY W, P^T Q) is run once on caller-provided data and the fp32
// accumulator is copied out of TMEM, so that tests can compare with a plain matmul.
#include "../../include/ngp_b200.h"
#include "tc_device.cuh"

namespace ngpb {

// Copies a row-major [rows][width] fp16 matrix into a chunk-major smem tile.
__device__ void load_tile_chunk_major(uint8_t* smem_tile, const __half* __restrict__ src, uint32_t rows, uint32_t width) {
	const uint32_t chunks_per_row = width / 8;
	for (uint32_t i = threadIdx.x; i < rows * chunks_per_row; i += blockDim.x) {
		const uint32_t r = i / chunks_per_row, c = i % chunks_per_row;
		const uint4 v = *reinterpret_cast<const uint4*>(src + (size_t)r * width + c * 8);
		*reinterpret_cast<uint4*>(smem_tile + tc::chunk_offset(rows, r, c)) = v;
	}
}

// mode 0: D[128 x N] = X[128 x K] W[N x K]^T
// mode 1: D[128 x N] = X[128 x K] W[K x N]        (W read MN-major; "dY W")
// mode 2: D[64 x N]  = X[128 x 64]^T W[128 x N]   ("P^T Q"; accumulator in the M=64 TMEM layout)
__global__ void __launch_bounds__(128) umma_selftest_kernel(int mode, uint32_t K, uint32_t N,
	const __half* __restrict__ X, const __half* __restrict__ W, float* __restrict__ D)
{
	extern __shared__ __align__(1024) uint8_t smem[];
	__shared__ uint64_t bar;
	__shared__ uint32_t tmem_base_slot;

	uint8_t* xs = smem;
	uint8_t* ws = smem + 32768;
	const uint32_t warp = threadIdx.x / 32, lane = threadIdx.x % 32;

	if (mode == 0) { load_tile_chunk_major(xs, X, 128, K); load_tile_chunk_major(ws, W, N, K); }
	else if (mode == 1) { load_tile_chunk_major(xs, X, 128, K); load_tile_chunk_major(ws, W, K, N); }
	else { load_tile_chunk_major(xs, X, 128, 64); load_tile_chunk_major(ws, W, 128, N); }

	if (warp == 0) tc::tmem_alloc<64>(&tmem_base_slot);
	if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
	tc::fence_proxy_async();
	tc::tc_fence_before();
	__syncthreads();
	tc::tc_fence_after();
	const uint32_t tmem = tmem_base_slot;

	if (threadIdx.x == 0) {
		if (mode == 0) tc::mma_xwT(tmem, tc::smem_u32(xs), tc::smem_u32(ws), K, N, false);
		else if (mode == 1) tc::mma_dyW(tmem, tc::smem_u32(xs), tc::smem_u32(ws), K, N);
		else tc::mma_pTq(tmem, tc::smem_u32(xs), tc::smem_u32(ws), N, false);
		tc::umma_commit(&bar);
	}
	tc::mbar_wait(&bar, 0);
	tc::tc_fence_after();

	for (uint32_t c0 = 0; c0 < N; c0 += 16) {
		uint32_t v[16];
		tc::tmem_ld16(tmem + ((warp * 32) << 16) + c0, v);
		tc::tmem_wait_ld();
		if (mode == 2) {
			// M = 64: row r lives in TMEM lane (r % 16) + 32 * (r / 16).
			if (lane < 16) {
				const uint32_t r = warp * 16 + lane;
				for (uint32_t j = 0; j < 16; ++j) D[r * N + c0 + j] = __uint_as_float(v[j]);
			}
		} else {
			const uint32_t r = warp * 32 + lane;
			for (uint32_t j = 0; j < 16; ++j) D[r * N + c0 + j] = __uint_as_float(v[j]);
		}
	}
	tc::tc_fence_before();
	__syncthreads();
	if (warp == 0) tc::tmem_dealloc<64>(tmem);
}

// Rate probe: thread 0 issues `reps` back-to-back batches of one operand view (same operands, same accumulator),
// commits once and waits. out[0] = cycles from first issue to completion, out[1] = cycles spent issuing.
__global__ void __launch_bounds__(128) umma_rate_kernel(int mode, uint32_t K, uint32_t N, uint32_t reps, unsigned long long* __restrict__ out) {
	extern __shared__ __align__(1024) uint8_t smem[];
	__shared__ uint64_t bar;
	__shared__ uint32_t tmem_base_slot;
	uint8_t* xs = smem;
	uint8_t* ws = smem + 32768;
	for (uint32_t i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u;
	if (threadIdx.x / 32 == 0) tc::tmem_alloc<64>(&tmem_base_slot);
	if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); }
	tc::fence_proxy_async();
	tc::tc_fence_before();
	__syncthreads();
	tc::tc_fence_after();
	const uint32_t tmem = tmem_base_slot;
	if (threadIdx.x == 0) {
		const unsigned long long t0 = clock64();
		for (uint32_t r = 0; r < reps; ++r) {
			if (mode == 0) tc::mma_xwT(tmem, tc::smem_u32(xs), tc::smem_u32(ws), K, N, false);
			else if (mode == 1) tc::mma_dyW(tmem, tc::smem_u32(xs), tc::smem_u32(ws), K, N);
			else tc::mma_pTq(tmem, tc::smem_u32(xs), tc::smem_u32(ws), N, false);
		}
		const unsigned long long t1 = clock64();
		tc::umma_commit(&bar);
		tc::mbar_wait(&bar, 0);
		const unsigned long long t2 = clock64();
		out[0] = t2 - t0;
		out[1] = t1 - t0;
	}
	tc::tc_fence_before();
	__syncthreads();
	if (threadIdx.x / 32 == 0) tc::tmem_dealloc<64>(tmem);
}

} // namespace ngpb

extern "C" int ngpb_probe_umma_rate(void* stream, int mode, uint32_t K, uint32_t N, uint32_t reps, unsigned long long* out) {
	using namespace ngpb;
	try {
		if (K % 16 != 0 || N % 16 != 0 || K > 64 || N > 64 || mode < 0 || mode > 2) throw std::runtime_error{"ngpb_probe_umma_rate: bad shape"};
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
		umma_rate_kernel<<<1, 128, 65536, (cudaStream_t)stream>>>(mode, K, N, reps, out);
		NGPB_CUDA_CHECK(cudaGetLastError());
	} catch (const std::exception& e) {
		set_last_error(e.what());
		return NGPB_ERROR;
	}
	return NGPB_OK;
}

extern "C" int ngpb_selftest_umma(void* stream, int mode, uint32_t K, uint32_t N, const void* X, const void* W, float* D) {
	using namespace ngpb;
	try {
		if (K % 16 != 0 || N % 16 != 0 || K > 64 || N > 64 || mode < 0 || mode > 2) throw std::runtime_error{"ngpb_selftest_umma: bad shape"};
		NGPB_CUDA_CHECK(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
		umma_selftest_kernel<<<1, 128, 65536, (cudaStream_t)stream>>>(mode, K, N, (const __half*)X, (const __half*)W, D);
		NGPB_CUDA_CHECK(cudaGetLastError());
	} catch (const std::exception& e) {
		set_last_error(e.what());
		return NGPB_ERROR;
	}
	return NGPB_OK;
}
