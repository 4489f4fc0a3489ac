// Host-side object model behind the C ABI: the counterpart of
// tcnn::DifferentiableObject<float, __half, __half> + ParametricObject (tcnn object.h:64-374),
// reduced to what tcnn::cpp::Module exposes (cpp_api.h:86-115). Raw device pointers in,
// raw device pointers out; a module owns no parameter memory.
#pragma once

#include "common.cuh"
#include "json.h"

#include <memory>
#include <utility>
#include <vector>

namespace ngpb {

constexpr uint32_t kBatchSizeGranularity = 256; // tcnn common.h:235
constexpr float kLossScaleFp16 = 128.0f;        // tcnn common.h:232

// Forward activations kept for the backward pass (tcnn::Context).
struct Context {
	virtual ~Context() {}
};

void keep_pool_memory(); // runtime.cu: sets the default mempool's release threshold once per device

// Device scratch owned by a context or a module call; freed on the stream it was made on.
struct DeviceBuffer {
	void* ptr = nullptr;
	size_t bytes = 0;
	cudaStream_t stream = nullptr;
	bool plain = false;   // cudaMalloc instead of the stream-ordered pool: memory that is exported to peer processes (cudaIpc)
	DeviceBuffer() {}
	DeviceBuffer(size_t n, cudaStream_t s) { alloc(n, s); }
	DeviceBuffer(const DeviceBuffer&) = delete;
	DeviceBuffer& operator=(const DeviceBuffer&) = delete;
	DeviceBuffer(DeviceBuffer&& o) noexcept { *this = std::move(o); }
	DeviceBuffer& operator=(DeviceBuffer&& o) noexcept {
		release();
		ptr = o.ptr; bytes = o.bytes; stream = o.stream; plain = o.plain;
		o.ptr = nullptr; o.bytes = 0;
		return *this;
	}
	~DeviceBuffer() { release(); }
	void alloc(size_t n, cudaStream_t s) {
		release();
		if (n == 0) return;
		stream = s;
		bytes = n;
		keep_pool_memory();
		NGPB_CUDA_CHECK(cudaMallocAsync(&ptr, n, s));
	}
	// Grows only; contents are not preserved.
	void reserve(size_t n, cudaStream_t s) { if (n > bytes) alloc(n, s); }
	void alloc_plain(size_t n, cudaStream_t s) {
		release();
		if (n == 0) return;
		stream = s;
		bytes = n;
		plain = true;
		NGPB_CUDA_CHECK(cudaMalloc(&ptr, n));
	}
	void release() {
		if (ptr) { if (plain) cudaFree(ptr); else cudaFreeAsync(ptr, stream); }
		plain = false;
		ptr = nullptr;
		bytes = 0;
	}
	template <typename T> T* as() const { return reinterpret_cast<T*>(ptr); }
};

class Module {
public:
	virtual ~Module() {}

	virtual void inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) = 0;
	virtual std::unique_ptr<Context> forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
		const __half* params, bool prepare_input_gradients) = 0;
	virtual void backward(cudaStream_t stream, const Context& ctx, uint32_t n, float* dL_dinput, const __half* dL_doutput,
		__half* dL_dparams, const float* input, const __half* output, const __half* params) = 0;

	virtual uint32_t n_input_dims() const = 0;
	virtual uint32_t n_output_dims() const = 0;          // unpadded
	virtual uint32_t padded_output_width() const = 0;
	virtual size_t n_params() const = 0;
	virtual std::vector<std::pair<uint32_t, uint32_t>> layer_sizes() const { return {}; }

	// params_fp32: device pointer. The generator is advanced as the reference advances its pcg32.
	virtual void initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) = 0;

	virtual Json hyperparams() const = 0;
	virtual std::string name() const = 0;
};

// tcnn random.h:39-73, same element -> stream-position mapping.
void random_uniform(cudaStream_t stream, Pcg32& rng, size_t n, float* out, float lower, float upper);

// ---- phase timer -----------------------------------------------------------------------------
// When enabled (ngpb_profile_enable), prof_mark(stream, name) records a CUDA event; the time
// between a mark and the next one on the same thread is charged to the FIRST mark's name when
// prof_collect() runs (after a stream synchronisation). Disabled: a branch and nothing else.
void prof_enable(bool on);
bool prof_enabled();
void prof_mark(cudaStream_t stream, const char* name);
// Sums pending intervals into the per-name totals; returns "name=ms;name=ms;..." of the totals.
std::string prof_collect(bool reset);

// params_fp16[i] = (half)params_fp32[i] (tcnn trainer.h:82-85)
void cast_fp32_to_fp16(cudaStream_t stream, size_t n, const float* src, __half* dst);
void cast_fp16_to_fp32(cudaStream_t stream, size_t n, const __half* src, float* dst);

// Factories (tcnn/src/encoding.cu:132, tcnn/src/network.cu:98, tcnn/src/cpp_api.cu:140-165).
Module* create_encoding_module(uint32_t n_input_dims, const Json& encoding);
Module* create_network_with_input_encoding_module(uint32_t n_input_dims, uint32_t n_output_dims, const Json& encoding, const Json& network);

Module* create_nerf_network_module(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
	const Json& pos_encoding, const Json& dir_encoding, const Json& density_network, const Json& rgb_network);

} // namespace ngpb
