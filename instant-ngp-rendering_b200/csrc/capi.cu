// extern "C" surface declared in include/ngp_b200.h. Every function catches C++ exceptions,
// stores the message for ngpb_last_error() and returns NGPB_ERROR.
#include "../../include/ngp_b200.h"

#include "encoding.h"
#include "module.h"

#include <cstring>

using namespace ngpb;

struct ngpb_module { std::unique_ptr<Module> impl; };
struct ngpb_context { std::unique_ptr<Context> impl; };

#define NGPB_API_BEGIN try {
#define NGPB_API_END \
	} catch (const std::exception& e) { set_last_error(e.what()); return NGPB_ERROR; } \
	catch (...) { set_last_error("unknown error"); return NGPB_ERROR; } \
	return NGPB_OK;

static size_t copy_out(const std::string& s, char* buf, size_t cap) {
	if (buf && cap > 0) {
		const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
		memcpy(buf, s.data(), n);
		buf[n] = 0;
	}
	return s.size() + 1;
}

static void require_device() {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
		cudaGetLastError();
		throw std::runtime_error{"ngp-b200: no CUDA device; this library has no CPU fallback"};
	}
}

extern "C" {

const char* ngpb_last_error(void) { return last_error(); }
int ngpb_version(void) { return 100; }
int ngpb_has_device(void) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n > 0 ? 1 : 0;
}
uint32_t ngpb_batch_size_granularity(void) { return kBatchSizeGranularity; }
float ngpb_default_loss_scale(int precision) { return precision == NGPB_PRECISION_FP16 ? kLossScaleFp16 : 1.0f; }

int ngpb_create_encoding(uint32_t n_input_dims, const char* encoding_json, int precision, ngpb_module_t** out) {
	NGPB_API_BEGIN
	if (!out || !encoding_json) throw std::runtime_error{"ngpb_create_encoding: null argument"};
	if (precision != NGPB_PRECISION_FP16) throw std::runtime_error{"ngp-b200 encodings produce fp16 (network_precision_t == __half, tcnn common.h:114-124)"};
	*out = new ngpb_module{std::unique_ptr<Module>{create_encoding_module(n_input_dims, Json::parse(encoding_json))}};
	NGPB_API_END
}


int ngpb_create_network_with_input_encoding(uint32_t n_input_dims, uint32_t n_output_dims,
	const char* encoding_json, const char* network_json, ngpb_module_t** out)
{
	NGPB_API_BEGIN
	if (!out || !encoding_json || !network_json) throw std::runtime_error{"ngpb_create_network_with_input_encoding: null argument"};
	*out = new ngpb_module{std::unique_ptr<Module>{create_network_with_input_encoding_module(
		n_input_dims, n_output_dims, Json::parse(encoding_json), Json::parse(network_json))}};
	NGPB_API_END
}

int ngpb_create_network(uint32_t n_input_dims, uint32_t n_output_dims, const char* network_json, ngpb_module_t** out) {
	NGPB_API_BEGIN
	if (!out || !network_json) throw std::runtime_error{"ngpb_create_network: null argument"};
	Json identity = Json::object();
	identity["otype"] = "Identity";
	*out = new ngpb_module{std::unique_ptr<Module>{create_network_with_input_encoding_module(
		n_input_dims, n_output_dims, identity, Json::parse(network_json))}};
	NGPB_API_END
}

int ngpb_create_nerf_network(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
	const char* pos_encoding_json, const char* dir_encoding_json,
	const char* density_network_json, const char* rgb_network_json, ngpb_module_t** out)
{
	NGPB_API_BEGIN
	if (!out) throw std::runtime_error{"ngpb_create_nerf_network: null argument"};
	*out = new ngpb_module{std::unique_ptr<Module>{create_nerf_network_module(n_pos_dims, n_dir_dims, n_extra_dims, dir_offset,
		Json::parse(pos_encoding_json), Json::parse(dir_encoding_json), Json::parse(density_network_json), Json::parse(rgb_network_json))}};
	NGPB_API_END
}

void ngpb_module_free(ngpb_module_t* m) { delete m; }

uint32_t ngpb_module_n_input_dims(const ngpb_module_t* m) { return m->impl->n_input_dims(); }
uint32_t ngpb_module_n_output_dims(const ngpb_module_t* m) { return m->impl->padded_output_width(); }
uint32_t ngpb_module_n_output_dims_unpadded(const ngpb_module_t* m) { return m->impl->n_output_dims(); }
size_t ngpb_module_n_params(const ngpb_module_t* m) { return m->impl->n_params(); }
int ngpb_module_param_precision(const ngpb_module_t*) { return NGPB_PRECISION_FP16; }
int ngpb_module_output_precision(const ngpb_module_t*) { return NGPB_PRECISION_FP16; }

size_t ngpb_module_hyperparams(const ngpb_module_t* m, char* buf, size_t cap) {
	try { return copy_out(m->impl->hyperparams().dump(), buf, cap); } catch (...) { return 0; }
}
size_t ngpb_module_name(const ngpb_module_t* m, char* buf, size_t cap) {
	try { return copy_out(m->impl->name(), buf, cap); } catch (...) { return 0; }
}
size_t ngpb_module_layer_sizes(const ngpb_module_t* m, uint32_t* rows_cols, size_t cap) {
	auto ls = m->impl->layer_sizes();
	for (size_t i = 0; i < ls.size() && i < cap; ++i) {
		rows_cols[2 * i] = ls[i].first;
		rows_cols[2 * i + 1] = ls[i].second;
	}
	return ls.size();
}

int ngpb_module_initialize_params(ngpb_module_t* m, void* stream, uint64_t seed, float* params_fp32, float scale) {
	NGPB_API_BEGIN
	require_device();
	Pcg32 rng{seed};
	m->impl->initialize_params((cudaStream_t)stream, rng, params_fp32, scale);
	NGPB_API_END
}

int ngpb_module_inference(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output, const void* params) {
	NGPB_API_BEGIN
	require_device();
	m->impl->inference((cudaStream_t)stream, n, input, (__half*)output, (const __half*)params);
	NGPB_API_END
}

int ngpb_module_forward(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output,
	const void* params, int prepare_input_gradients, ngpb_context_t** ctx)
{
	NGPB_API_BEGIN
	require_device();
	if (!ctx) throw std::runtime_error{"ngpb_module_forward: ctx must not be null"};
	auto c = m->impl->forward((cudaStream_t)stream, n, input, (__half*)output, (const __half*)params, prepare_input_gradients != 0);
	*ctx = new ngpb_context{std::move(c)};
	NGPB_API_END
}

int ngpb_module_backward(ngpb_module_t* m, void* stream, const ngpb_context_t* ctx, uint32_t n,
	float* dL_dinput, const void* dL_doutput, void* dL_dparams,
	const float* input, const void* output, const void* params)
{
	NGPB_API_BEGIN
	require_device();
	if (!ctx || !ctx->impl) throw std::runtime_error{"ngpb_module_backward: missing forward context"};
	m->impl->backward((cudaStream_t)stream, *ctx->impl, n, dL_dinput, (const __half*)dL_doutput, (__half*)dL_dparams,
		input, (const __half*)output, (const __half*)params);
	NGPB_API_END
}

void ngpb_context_free(ngpb_context_t* ctx) { delete ctx; }

uint32_t ngpb_grid_level_table(const ngpb_module_t* m, uint32_t* offsets, float* scales, uint32_t* resolutions) {
	auto* g = dynamic_cast<const GridEncodingModule*>(m->impl.get());
	if (!g) return 0;
	const GridParams& p = g->config().p;
	for (uint32_t l = 0; l < p.n_levels; ++l) {
		if (offsets) offsets[l] = p.offset[l];
		if (scales) scales[l] = p.scale[l];
		if (resolutions) resolutions[l] = p.resolution[l];
	}
	if (offsets) offsets[p.n_levels] = p.offset[p.n_levels];
	return p.n_levels;
}

int ngpb_grid_indices(const ngpb_module_t* m, void* stream, uint32_t n, const float* input, uint32_t* idx) {
	NGPB_API_BEGIN
	require_device();
	auto* g = dynamic_cast<const GridEncodingModule*>(m->impl.get());
	if (!g) throw std::runtime_error{"ngpb_grid_indices: module is not a grid encoding"};
	grid_indices((cudaStream_t)stream, g->config().p, n, input, g->n_input_dims(), idx);
	NGPB_API_END
}

int ngpb_grid_set_max_level(ngpb_module_t* m, float max_level) {
	NGPB_API_BEGIN
	auto* g = dynamic_cast<GridEncodingModule*>(m->impl.get());
	if (!g) throw std::runtime_error{"ngpb_grid_set_max_level: module is not a grid encoding"};
	g->set_max_level(max_level);
	NGPB_API_END
}

int ngpb_random_uniform(void* stream, uint64_t* rng_state, size_t n, float* out, float lower, float upper) {
	NGPB_API_BEGIN
	require_device();
	Pcg32 rng;
	rng.state = rng_state[0];
	rng.inc = rng_state[1];
	random_uniform((cudaStream_t)stream, rng, n, out, lower, upper);
	rng_state[0] = rng.state;
	rng_state[1] = rng.inc;
	NGPB_API_END
}

} // extern "C"
