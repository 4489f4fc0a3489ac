// ngpb_module_inference_static_scratch: ngpb_module_inference for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference.
//
// ngpb_module_inference keeps the encoder's feature tiles (the operand image the first UMMA reads, mlp.h) in a
// stream-ordered allocation made per call. Here the tiles live in a file-scope device array instead: 4 MiB that
// every GPU which loads the library holds until exit. The entry point makes no cudaMalloc / cudaFree / pool call.
//
// Every call on a device shares the one array, and it keeps the last call's contents. The region a call uses is
// cleared first (the tile padding must read as zero). Two calls that overlap on the GPU - different streams, or
// graph branches - race on it; callers serialise them.
#include "../../include/ngp_b200.h"

#include "network.h"

#include <mutex>

using namespace ngpb;

struct ngpb_module { std::unique_ptr<Module> impl; }; // same layout as in capi.cu

constexpr size_t kStaticScratchBytes = NGPB_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address the kernels may use
// comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kMaxDevices = 64;
static void* static_scratch_for_current_device() {
	static std::mutex mutex;
	static void* ptrs[kMaxDevices] = {};
	int dev = -1;
	NGPB_CUDA_CHECK(cudaGetDevice(&dev));
	if (dev < 0 || dev >= kMaxDevices) throw std::runtime_error{"ngpb_module_inference_static_scratch: device ordinal out of range"};
	std::lock_guard<std::mutex> lock{mutex};
	if (!ptrs[dev]) {
		void* p = nullptr;
		const cudaError_t err = cudaGetSymbolAddress(&p, g_static_scratch);
		if (err != cudaSuccess || !p) {
			cudaGetLastError();
			throw std::runtime_error{std::string{"ngpb_module_inference_static_scratch: cudaGetSymbolAddress failed: "} + cudaGetErrorString(err)};
		}
		ptrs[dev] = p;
	}
	return ptrs[dev];
}

extern "C" {

size_t ngpb_static_scratch_bytes(void) { return kStaticScratchBytes; }

int ngpb_module_inference_static_scratch(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output, const void* params) {
	try {
		int n_dev = 0;
		if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
			cudaGetLastError();
			throw std::runtime_error{"ngp-b200: no CUDA device; this library has no CPU fallback"};
		}
		if (!m || !m->impl) throw std::runtime_error{"ngpb_module_inference_static_scratch: null module"};
		auto* net = dynamic_cast<NetworkWithInputEncodingModule*>(m->impl.get());
		if (!net) throw std::runtime_error{"ngpb_module_inference_static_scratch: the module must be a NetworkWithInputEncoding (ngpb_create_network_with_input_encoding, ngpb_create_network)"};
		const size_t need = net->inference_scratch_bytes(n);
		if (need > kStaticScratchBytes) {
			throw std::runtime_error{"ngpb_module_inference_static_scratch: a batch of " + std::to_string(n) + " samples needs " + std::to_string(need) +
				" bytes of feature tiles, the static scratch holds " + std::to_string(kStaticScratchBytes) + "; split the batch or use ngpb_module_inference"};
		}
		if (n == 0) return NGPB_OK;
		void* scratch = static_scratch_for_current_device();
		NGPB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, need, (cudaStream_t)stream));
		net->inference_with_scratch((cudaStream_t)stream, n, input, (__half*)output, (const __half*)params, scratch, need);
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; }
	catch (...) { set_last_error("unknown error"); return NGPB_ERROR; }
	return NGPB_OK;
}

} // extern "C"
