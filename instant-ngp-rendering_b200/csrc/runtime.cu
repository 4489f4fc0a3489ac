// Small runtime pieces: error string, SM count, pcg32 fills.
#include "module.h"

#include <map>
#include <sstream>
#include <vector>

namespace ngpb {

static thread_local std::string g_last_error;

void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* last_error() { return g_last_error.c_str(); }

// ---- phase timer ---------------------------------------------------------------------------------
namespace {
struct ProfState {
	bool enabled = false;
	std::vector<std::pair<std::string, cudaEvent_t>> marks;
	std::vector<cudaEvent_t> pool;
	std::map<std::string, std::pair<double, uint64_t>> totals; // ms, count
};
ProfState& prof() { static thread_local ProfState s; return s; }
}
void prof_enable(bool on) { prof().enabled = on; }
bool prof_enabled() { return prof().enabled; }
void prof_mark(cudaStream_t stream, const char* name) {
	ProfState& p = prof();
	if (!p.enabled) return;
	cudaEvent_t e;
	if (!p.pool.empty()) { e = p.pool.back(); p.pool.pop_back(); }
	else NGPB_CUDA_CHECK(cudaEventCreate(&e));
	NGPB_CUDA_CHECK(cudaEventRecord(e, stream));
	p.marks.emplace_back(name, e);
}
std::string prof_collect(bool reset) {
	ProfState& p = prof();
	for (size_t i = 0; i + 1 < p.marks.size(); ++i) {
		if (p.marks[i].first == "end") continue;
		float ms = 0.0f;
		if (cudaEventElapsedTime(&ms, p.marks[i].second, p.marks[i + 1].second) == cudaSuccess) {
			auto& t = p.totals[p.marks[i].first];
			t.first += ms;
			t.second += 1;
		} else cudaGetLastError();
	}
	for (auto& m : p.marks) p.pool.push_back(m.second);
	p.marks.clear();
	std::ostringstream o;
	for (auto& kv : p.totals) o << kv.first << "=" << kv.second.first << ":" << kv.second.second << ";";
	if (reset) p.totals.clear();
	return o.str();
}

// Stream-ordered scratch (DeviceBuffer) is freed and re-taken every step; without a release
// threshold the driver returns the pool's pages to the OS at every synchronisation and the next
// cudaMallocAsync maps them again (hundreds of microseconds per step).
void keep_pool_memory() {
	static bool done[64] = {};
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
	cudaMemPool_t pool;
	if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
		uint64_t threshold = UINT64_MAX;
		cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
	}
	cudaGetLastError();
	done[dev] = true;
}

int sm_count() {
	static int count = -1;
	if (count < 0) {
		int dev = 0;
		cudaDeviceProp prop;
		if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
			cudaGetLastError();
			return 148;
		}
		count = prop.multiProcessorCount;
	}
	return count;
}

// The reference fills with generate_random_kernel<T, RNG, 4> (tcnn random.h:39-56): thread t of
// T = 128 * ceil(ceil(n/4) / 128) threads advances the generator by 4*t and writes elements
// t, t+T, t+2T, t+3T. So element e takes stream position 4*(e % T) + e / T. One thread per
// element here (coalesced writes), jumping straight to that position.
__global__ void random_uniform_kernel(size_t n, size_t n_threads_ref, Pcg32 rng, float* __restrict__ out, float lower, float upper) {
	const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= n) return;
	const size_t t = e % n_threads_ref;
	const size_t j = e / n_threads_ref;
	rng.advance((int64_t)(t * 4 + j));
	out[e] = rng.next_float() * (upper - lower) + lower;
}

void random_uniform(cudaStream_t stream, Pcg32& rng, size_t n, float* out, float lower, float upper) {
	if (n > 0) {
		const size_t ref_threads_needed = (n + 3) / 4;
		const size_t n_threads_ref = 128 * ((ref_threads_needed + 127) / 128);
		const uint32_t blocks = (uint32_t)((n + 255) / 256);
		random_uniform_kernel<<<blocks, 256, 0, stream>>>(n, n_threads_ref, rng, out, lower, upper);
		NGPB_CUDA_CHECK(cudaGetLastError());
	}
	rng.advance((int64_t)n);
}

__global__ void cast_fp32_to_fp16_kernel(size_t n, const float* __restrict__ src, __half* __restrict__ dst) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = __float2half(src[i]);
}
void cast_fp32_to_fp16(cudaStream_t stream, size_t n, const float* src, __half* dst) {
	if (!n) return;
	cast_fp32_to_fp16_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, stream>>>(n, src, dst);
	NGPB_CUDA_CHECK(cudaGetLastError());
}


__global__ void cast_fp16_to_fp32_kernel(size_t n, const __half* __restrict__ src, float* __restrict__ dst) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = __half2float(src[i]);
}
void cast_fp16_to_fp32(cudaStream_t stream, size_t n, const __half* src, float* dst) {
	if (!n) return;
	cast_fp16_to_fp32_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, stream>>>(n, src, dst);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
