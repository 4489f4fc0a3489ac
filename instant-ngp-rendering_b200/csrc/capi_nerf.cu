// extern "C" surface for the optimizer, the NeRF kernels and the Testbed (include/ngp_b200.h).
#include "../../include/ngp_b200.h"

#include "testbed.h"

#include <cstring>

using namespace ngpb;

struct ngpb_module { std::unique_ptr<Module> impl; }; // same layout as in capi.cu
struct ngpb_optimizer { std::unique_ptr<Optimizer> impl; };

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

// The C structs are the kernels' PODs under another name.
static_assert(sizeof(ngpb_nerf_image_t) == sizeof(ImageMeta), "ngpb_nerf_image_t layout");
static_assert(sizeof(ngpb_nerf_counters_t) == sizeof(NerfStepCounters), "ngpb_nerf_counters_t layout");
static_assert(sizeof(ngpb_nerf_train_config_t) == sizeof(NerfTrainConfig), "ngpb_nerf_train_config_t layout");
static_assert(sizeof(ngpb_nerf_train_buffers_t) == sizeof(NerfTrainBuffers), "ngpb_nerf_train_buffers_t layout");

static Pcg32 rng_from(const uint64_t* s) {
	Pcg32 r;
	r.state = s[0];
	r.inc = s[1];
	return r;
}

extern "C" {

// ---- optimizer ---------------------------------------------------------------------------------
int ngpb_optimizer_create(const char* optimizer_json, ngpb_optimizer_t** out) {
	NGPB_API_BEGIN
	if (!out || !optimizer_json) throw std::runtime_error{"ngpb_optimizer_create: null argument"};
	*out = new ngpb_optimizer{std::make_unique<Optimizer>(Json::parse(optimizer_json))};
	NGPB_API_END
}
void ngpb_optimizer_free(ngpb_optimizer_t* o) { delete o; }
int ngpb_optimizer_allocate(ngpb_optimizer_t* o, uint32_t n_weights, const uint32_t* layer_rows_cols, size_t n_layers) {
	NGPB_API_BEGIN
	require_device();
	std::vector<std::pair<uint32_t, uint32_t>> layers;
	for (size_t i = 0; i < n_layers; ++i) layers.emplace_back(layer_rows_cols[2 * i], layer_rows_cols[2 * i + 1]);
	o->impl->allocate(n_weights, layers);
	NGPB_API_END
}
int ngpb_optimizer_step(ngpb_optimizer_t* o, void* stream, float loss_scale, float* weights_fp32, void* weights, const void* gradients) {
	NGPB_API_BEGIN
	require_device();
	o->impl->step((cudaStream_t)stream, loss_scale, weights_fp32, (__half*)weights, (const __half*)gradients);
	NGPB_API_END
}
float ngpb_optimizer_learning_rate(const ngpb_optimizer_t* o) { return o->impl->learning_rate(); }
int ngpb_optimizer_set_learning_rate(ngpb_optimizer_t* o, float lr) {
	NGPB_API_BEGIN
	o->impl->set_learning_rate(lr);
	NGPB_API_END
}
uint32_t ngpb_optimizer_step_count(const ngpb_optimizer_t* o) { return o->impl->step_count(); }
void* ngpb_optimizer_custom_weights(const ngpb_optimizer_t* o) { return o->impl->custom_weights(); }
int ngpb_optimizer_update_hyperparams(ngpb_optimizer_t* o, const char* optimizer_json) {
	NGPB_API_BEGIN
	o->impl->update_hyperparams(Json::parse(optimizer_json));
	NGPB_API_END
}
size_t ngpb_optimizer_hyperparams(const ngpb_optimizer_t* o, char* buf, size_t cap) {
	try { return copy_out(o->impl->hyperparams().dump(), buf, cap); } catch (...) { return 0; }
}
int ngpb_optimizer_state(const ngpb_optimizer_t* o, float** first_moments, float** second_moments, uint32_t** param_steps) {
	NGPB_API_BEGIN
	if (first_moments) *first_moments = o->impl->first_moments();
	if (second_moments) *second_moments = o->impl->second_moments();
	if (param_steps) *param_steps = o->impl->param_steps();
	NGPB_API_END
}

// ---- NerfNetwork extras -------------------------------------------------------------------------
int ngpb_nerf_network_density(ngpb_module_t* m, void* stream, uint32_t n, const float* input, uint32_t in_stride, void* output, uint32_t out_stride, const void* params) {
	NGPB_API_BEGIN
	require_device();
	auto* net = dynamic_cast<NerfNetworkModule*>(m->impl.get());
	if (!net) throw std::runtime_error{"ngpb_nerf_network_density: module is not a NerfNetwork"};
	if (out_stride != 1 && out_stride != 16) throw std::runtime_error{"ngpb_nerf_network_density: out_stride must be 1 or 16"};
	net->density((cudaStream_t)stream, n, input, in_stride, (__half*)output, out_stride, (const __half*)params);
	NGPB_API_END
}
int ngpb_nerf_network_param_layout(const ngpb_module_t* m, size_t* offsets) {
	NGPB_API_BEGIN
	auto* net = dynamic_cast<const NerfNetworkModule*>(m->impl.get());
	if (!net) throw std::runtime_error{"ngpb_nerf_network_param_layout: module is not a NerfNetwork"};
	offsets[0] = 0;
	offsets[1] = net->net().b_param_off;
	offsets[2] = net->pos_encoding_param_offset();
	offsets[3] = net->n_params();
	NGPB_API_END
}

// ---- NeRF kernels ----------------------------------------------------------------------------------
int ngpb_nerf_generate_training_samples(void* stream, const ngpb_nerf_train_config_t* cfg, const ngpb_nerf_train_buffers_t* buf, const uint64_t* rng_state) {
	NGPB_API_BEGIN
	require_device();
	NerfTrainConfig c; NerfTrainBuffers b;
	memcpy(&c, cfg, sizeof(c)); memcpy(&b, buf, sizeof(b));
	nerf_generate_training_samples((cudaStream_t)stream, c, b, rng_from(rng_state));
	NGPB_API_END
}
int ngpb_nerf_compute_loss_and_compact(void* stream, const ngpb_nerf_train_config_t* cfg, const ngpb_nerf_train_buffers_t* buf, const uint64_t* rng_state) {
	NGPB_API_BEGIN
	require_device();
	NerfTrainConfig c; NerfTrainBuffers b;
	memcpy(&c, cfg, sizeof(c)); memcpy(&b, buf, sizeof(b));
	nerf_compute_loss_and_compact((cudaStream_t)stream, c, b, rng_from(rng_state));
	NGPB_API_END
}
int ngpb_nerf_update_bitfield(void* stream, uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density) {
	NGPB_API_BEGIN
	require_device();
	if (max_cascade >= kNerfCascades) throw std::runtime_error{"ngpb_nerf_update_bitfield: max_cascade must be < 8"};
	nerf_update_bitfield((cudaStream_t)stream, max_cascade, grid, bitfield, mean_density);
	NGPB_API_END
}
size_t ngpb_nerf_bitfield_bytes(void) { return nerf_bitfield_bytes(); }

int ngpb_nerf_mark_untrained_density_grid(void* stream, uint32_t n_elements, float* grid, uint32_t n_images, const ngpb_nerf_image_t* metadata, int clear_visible_voxels) {
	NGPB_API_BEGIN
	require_device();
	static_assert(sizeof(ngpb_nerf_image_t) == sizeof(ImageMeta), "ngpb_nerf_image_t must mirror ImageMeta");
	nerf_mark_untrained_density_grid((cudaStream_t)stream, n_elements, grid, n_images, reinterpret_cast<const ImageMeta*>(metadata), clear_visible_voxels != 0);
	NGPB_API_END
}
int ngpb_nerf_generate_grid_samples(void* stream, uint32_t n, const uint64_t* rng_state, uint32_t step, const float* aabb6, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh)
{
	NGPB_API_BEGIN
	require_device();
	if (n_cascades == 0 || n_cascades > kNerfCascades) throw std::runtime_error{"ngpb_nerf_generate_grid_samples: n_cascades must be 1..8"};
	const Aabb aabb = {{aabb6[0], aabb6[1], aabb6[2]}, {aabb6[3], aabb6[4], aabb6[5]}};
	nerf_generate_grid_samples((cudaStream_t)stream, n, rng_from(rng_state), step, aabb, grid_in, positions, indices, n_cascades, thresh);
	NGPB_API_END
}
int ngpb_nerf_splat_grid_samples(void* stream, uint32_t n, const uint32_t* indices, const void* density_raw, float* grid_tmp, int density_activation) {
	NGPB_API_BEGIN
	require_device();
	nerf_splat_grid_samples((cudaStream_t)stream, n, indices, (const __half*)density_raw, grid_tmp, density_activation);
	NGPB_API_END
}
int ngpb_nerf_ema_grid_samples(void* stream, uint32_t n_elements, float decay, float* grid, const float* grid_tmp) {
	NGPB_API_BEGIN
	require_device();
	nerf_ema_grid_samples((cudaStream_t)stream, n_elements, decay, grid, grid_tmp);
	NGPB_API_END
}

} // extern "C"
