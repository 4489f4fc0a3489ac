// The reference-side binding of INTEGRATION.md section 1 as a file that compiles: a second implementation of tiny-cuda-nn's
// raw-pointer facade (tcnn/include/tiny-cuda-nn/cpp_api.h:58-119) over libngp_b200.so. It takes the place of
// tcnn/src/cpp_api.cu in a build of the reference: every caller of tcnn::cpp::Module - tcnn's own PyTorch binding
// (tcnn/bindings/torch/tinycudann/bindings.cpp) among them - then runs on the sm_100a kernels without a source change.
//
// Nothing here touches CUDA: device pointers and the stream are forwarded as they come. Needs the reference's headers
// on the include path (cpp_api.h and its json.hpp), so it is built only where /root/reference exists
// (integration/build.sh); it is not part of libngp_b200.so.
#include <tiny-cuda-nn/cpp_api.h>

#include <ngp_b200.h>

#include <functional>
#include <stdexcept>
#include <string>

namespace tcnn { namespace cpp {

namespace {
// C ABI status -> the exception type the reference throws (CHECK_THROW / CUDA_CHECK_THROW raise std::runtime_error)
void check(int rc) {
	if (rc != NGPB_OK) throw std::runtime_error{ngpb_last_error()};
}

struct B200Context : public tcnn::Context {
	ngpb_context_t* handle = nullptr;
	~B200Context() override { ngpb_context_free(handle); }
};

template <typename F> std::string text_of(F&& writer) {
	std::string buf(writer(nullptr, 0), '\0');
	if (buf.empty()) throw std::runtime_error{ngpb_last_error()};
	writer(&buf[0], buf.size());
	buf.resize(buf.size() - 1);   // the C side counts the terminating NUL
	return buf;
}

class B200Module : public Module {
public:
	explicit B200Module(ngpb_module_t* handle)
	: Module{(Precision)ngpb_module_param_precision(handle), (Precision)ngpb_module_output_precision(handle)}, m_handle{handle} {}
	~B200Module() override { ngpb_module_free(m_handle); }

	void inference(cudaStream_t stream, uint32_t n_elements, const float* input, void* output, void* params) override {
		check(ngpb_module_inference(m_handle, stream, n_elements, input, output, params));
	}

	Context forward(cudaStream_t stream, uint32_t n_elements, const float* input, void* output, void* params, bool prepare_input_gradients) override {
		auto ctx = std::make_unique<B200Context>();
		check(ngpb_module_forward(m_handle, stream, n_elements, input, output, params, prepare_input_gradients ? 1 : 0, &ctx->handle));
		Context result;
		result.ctx = std::move(ctx);
		return result;
	}

	void backward(cudaStream_t stream, const Context& ctx, uint32_t n_elements, float* dL_dinput, const void* dL_doutput, void* dL_dparams,
			const float* input, const void* output, const void* params) override {
		const auto* mine = dynamic_cast<const B200Context*>(ctx.ctx.get());
		if (!mine) throw std::runtime_error{"backward: the context was not produced by this module's forward"};
		check(ngpb_module_backward(m_handle, stream, mine->handle, n_elements, dL_dinput, dL_doutput, dL_dparams, input, output, params));
	}

	// Second derivatives (tcnn object.h:300-374) are outside the NeRF / image path: fail loudly.
	void backward_backward_input(cudaStream_t, const Context&, uint32_t, const float*, const float*, const void*, void*, void*, float*, const void*) override {
		throw std::runtime_error{"DifferentiableObject::backward_backward_input_impl: not implemented error"};
	}

	uint32_t n_input_dims() const override { return ngpb_module_n_input_dims(m_handle); }
	uint32_t n_output_dims() const override { return ngpb_module_n_output_dims(m_handle); }   // padded, as cpp_api.cu:132
	size_t n_params() const override { return ngpb_module_n_params(m_handle); }

	void initialize_params(size_t seed, float* params_full_precision, float scale) override {
		check(ngpb_module_initialize_params(m_handle, nullptr, (uint64_t)seed, params_full_precision, scale));
	}

	json hyperparams() const override {
		return json::parse(text_of([this](char* buf, size_t cap) { return ngpb_module_hyperparams(m_handle, buf, cap); }));
	}
	std::string name() const override {
		return text_of([this](char* buf, size_t cap) { return ngpb_module_name(m_handle, buf, cap); });
	}

private:
	ngpb_module_t* m_handle;
};
} // namespace

uint32_t batch_size_granularity() { return ngpb_batch_size_granularity(); }

int cuda_device() {
	int device = 0;
	if (cudaGetDevice(&device) != cudaSuccess) throw std::runtime_error{"cudaGetDevice failed"};
	return device;
}
void set_cuda_device(int device) {
	if (cudaSetDevice(device) != cudaSuccess) throw std::runtime_error{"cudaSetDevice failed"};
}

// libngp_b200.so allocates scratch stream-ordered per call and returns it before the call ends: there is no arena to drain
void free_temporary_memory() {}

bool has_networks() { return true; }

float default_loss_scale(Precision p) { return ngpb_default_loss_scale((int)p); }
Precision preferred_precision() { return Precision::Fp16; }

// the library does not log; errors arrive as exceptions
void set_log_callback(const std::function<void(LogSeverity, const std::string&)>&) {}

Module* create_network_with_input_encoding(uint32_t n_input_dims, uint32_t n_output_dims, const json& encoding, const json& network) {
	ngpb_module_t* handle = nullptr;
	check(ngpb_create_network_with_input_encoding(n_input_dims, n_output_dims, encoding.dump().c_str(), network.dump().c_str(), &handle));
	return new B200Module{handle};
}

Module* create_network(uint32_t n_input_dims, uint32_t n_output_dims, const json& network) {
	ngpb_module_t* handle = nullptr;
	check(ngpb_create_network(n_input_dims, n_output_dims, network.dump().c_str(), &handle));
	return new B200Module{handle};
}

Module* create_encoding(uint32_t n_input_dims, const json& encoding, Precision requested_precision) {
	ngpb_module_t* handle = nullptr;
	check(ngpb_create_encoding(n_input_dims, encoding.dump().c_str(), (int)requested_precision, &handle));
	return new B200Module{handle};
}

}} // namespace tcnn::cpp
