// Micro-probes behind the gather / scatter rooflines of the hash-grid encoding ("peak =
// measured L2-resident random-gather bandwidth on the box"). No reference counterpart. Each thread
// issues `per_thread` independent random accesses (LCG indices, not data dependent, 8 in flight)
// into a table of 2^k entries; the caller times the launch with CUDA events.
#include "common.cuh"
#include "../../include/ngp_b200.h"

namespace ngpb {

template <typename T> NGPB_D uint32_t fold(T v);
template <> NGPB_D uint32_t fold<uint32_t>(uint32_t v) { return v; }
template <> NGPB_D uint32_t fold<uint2>(uint2 v) { return v.x ^ v.y; }
template <> NGPB_D uint32_t fold<uint4>(uint4 v) { return v.x ^ v.y ^ v.z ^ v.w; }

template <typename T>
__global__ void __launch_bounds__(256) gather_probe_kernel(const T* __restrict__ table, uint32_t mask, uint32_t per_thread, uint32_t* __restrict__ sink) {
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t idx = tid * 2654435761u + 12345u, acc = 0;
	for (uint32_t k = 0; k < per_thread; k += 8) {
		uint32_t j[8];
#pragma unroll
		for (uint32_t u = 0; u < 8; ++u) { idx = idx * 1664525u + 1013904223u; j[u] = (idx >> 7) & mask; }
#pragma unroll
		for (uint32_t u = 0; u < 8; ++u) acc ^= fold<T>(__ldg(table + j[u]));
	}
	if (acc == 0x9E3779B9u) sink[0] = acc; // keeps the loads alive
}

__global__ void __launch_bounds__(256) scatter_probe_kernel(__half* __restrict__ table, uint32_t mask, uint32_t per_thread) {
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t idx = tid * 2654435761u + 12345u;
	const uint32_t one = 0x00010001u; // two fp16 subnormals: the sums stay finite
	for (uint32_t k = 0; k < per_thread; ++k) {
		idx = idx * 1664525u + 1013904223u;
		const uint32_t j = (idx >> 7) & mask;
		asm volatile("red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;" ::"l"(table + 2 * (size_t)j), "r"(one) : "memory");
	}
}

// Blocks that stay resident for `ns` nanoseconds without issuing work (they sleep): what a kernel
// loses by SHARING its SMs with other blocks, as opposed to competing with them.
__global__ void resident_probe_kernel(unsigned long long ns) {
	unsigned long long t0, t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
	do {
		__nanosleep(2000);
		asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	} while (t - t0 < ns);
}

} // namespace ngpb

extern "C" int ngpb_probe_resident(void* stream, uint32_t blocks, uint32_t threads, uint64_t nanoseconds) {
	using namespace ngpb;
	try {
		if (blocks == 0 || threads == 0 || threads > 1024 || nanoseconds > 2000000000ull) throw std::runtime_error{"ngpb_probe_resident: 1..1024 threads, at most 2 s"};
		resident_probe_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(nanoseconds);
		NGPB_CUDA_CHECK(cudaGetLastError());
		return NGPB_OK;
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_probe_gather(void* stream, const void* table, uint32_t log2_entries, uint32_t entry_bytes, uint32_t n_threads, uint32_t per_thread, uint32_t* sink) {
	using namespace ngpb;
	try {
		if (log2_entries == 0 || log2_entries > 30 || per_thread % 8 != 0) throw std::runtime_error{"ngpb_probe_gather: log2_entries in 1..30, per_thread a multiple of 8"};
		const uint32_t mask = (1u << log2_entries) - 1, blocks = div_round_up(n_threads, 256);
		cudaStream_t s = (cudaStream_t)stream;
		if (entry_bytes == 4) gather_probe_kernel<uint32_t><<<blocks, 256, 0, s>>>((const uint32_t*)table, mask, per_thread, sink);
		else if (entry_bytes == 8) gather_probe_kernel<uint2><<<blocks, 256, 0, s>>>((const uint2*)table, mask, per_thread, sink);
		else if (entry_bytes == 16) gather_probe_kernel<uint4><<<blocks, 256, 0, s>>>((const uint4*)table, mask, per_thread, sink);
		else throw std::runtime_error{"ngpb_probe_gather: entry_bytes must be 4, 8 or 16"};
		NGPB_CUDA_CHECK(cudaGetLastError());
		return NGPB_OK;
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_probe_scatter(void* stream, void* table, uint32_t log2_entries, uint32_t n_threads, uint32_t per_thread) {
	using namespace ngpb;
	try {
		if (log2_entries == 0 || log2_entries > 30) throw std::runtime_error{"ngpb_probe_scatter: log2_entries in 1..30"};
		scatter_probe_kernel<<<div_round_up(n_threads, 256), 256, 0, (cudaStream_t)stream>>>((__half*)table, (1u << log2_entries) - 1, per_thread);
		NGPB_CUDA_CHECK(cudaGetLastError());
		return NGPB_OK;
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
}
