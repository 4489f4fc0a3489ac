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

// Vector reductions: V packed f16x2 words (4 V bytes, aligned) per instruction - what merging the two x-neighbours of
// a corner pair (adjacent entries under the coherent-prime hash when the lower x is even) buys at the L2.
template <uint32_t V>
__global__ void __launch_bounds__(256) scatter_vec_probe_kernel(__half* __restrict__ table, uint32_t mask, uint32_t per_thread) {
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t idx = tid * 2654435761u + 12345u;
	const uint32_t one = 0x00010001u;
	for (uint32_t k = 0; k < per_thread; ++k) {
		idx = idx * 1664525u + 1013904223u;
		const uint32_t j = ((idx >> 7) & mask) & ~(V - 1u);
		__half* addr = table + 2 * (size_t)j;
		if constexpr (V == 1) asm volatile("red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;" ::"l"(addr), "r"(one) : "memory");
		else if constexpr (V == 2) asm volatile("red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1, %1};" ::"l"(addr), "r"(one) : "memory");
		else asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %1, %1, %1};" ::"l"(addr), "r"(one) : "memory");
	}
}

// Streaming read of an L2-resident buffer (ld.global.cg: no L1), 16 bytes per lane, fully coalesced: the L2 -> SM
// bandwidth peak that a sector count is measured against (grid.roofline in bench.py).
__global__ void __launch_bounds__(256) l2_stream_probe_kernel(const uint4* __restrict__ buf, uint32_t n_vec, uint32_t passes, uint32_t* __restrict__ sink) {
	uint32_t acc = 0;
	for (uint32_t pass = 0; pass < passes; ++pass) {
		for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += 4 * gridDim.x * blockDim.x) {
			uint4 v[4];
#pragma unroll
			for (uint32_t u = 0; u < 4; ++u) {
				const uint32_t j = i + u * gridDim.x * blockDim.x;
				v[u] = j < n_vec ? __ldcg(buf + j) : make_uint4(0u, 0u, 0u, 0u);
			}
#pragma unroll
			for (uint32_t u = 0; u < 4; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
		}
	}
	if (acc == 0x9E3779B9u) sink[0] = acc;
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

extern "C" int ngpb_probe_scatter_vec(void* stream, void* table, uint32_t log2_entries, uint32_t vec, uint32_t n_threads, uint32_t per_thread) {
	using namespace ngpb;
	try {
		if (log2_entries < 2 || log2_entries > 30) throw std::runtime_error{"ngpb_probe_scatter_vec: log2_entries in 2..30"};
		const uint32_t mask = (1u << log2_entries) - 1, blocks = div_round_up(n_threads, 256);
		cudaStream_t s = (cudaStream_t)stream;
		if (vec == 1) scatter_vec_probe_kernel<1><<<blocks, 256, 0, s>>>((__half*)table, mask, per_thread);
		else if (vec == 2) scatter_vec_probe_kernel<2><<<blocks, 256, 0, s>>>((__half*)table, mask, per_thread);
		else if (vec == 4) scatter_vec_probe_kernel<4><<<blocks, 256, 0, s>>>((__half*)table, mask, per_thread);
		else throw std::runtime_error{"ngpb_probe_scatter_vec: vec must be 1, 2 or 4"};
		NGPB_CUDA_CHECK(cudaGetLastError());
		return NGPB_OK;
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_probe_l2_stream(void* stream, const void* buf, size_t bytes, uint32_t passes, uint32_t blocks, uint32_t* sink) {
	using namespace ngpb;
	try {
		if (bytes < 16 || bytes > (1ull << 32) || passes == 0 || blocks == 0) throw std::runtime_error{"ngpb_probe_l2_stream: 16 B .. 4 GiB, passes and blocks >= 1"};
		l2_stream_probe_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)buf, (uint32_t)(bytes / 16), passes, sink);
		NGPB_CUDA_CHECK(cudaGetLastError());
		return NGPB_OK;
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
}
