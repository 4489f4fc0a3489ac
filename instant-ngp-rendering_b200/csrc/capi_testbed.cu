// extern "C" surface of the Testbed (include/ngp_b200.h).
#include "../../include/ngp_b200.h"

#include "testbed.h"
#include "dist.h"

#include <cstring>

using namespace ngpb;

struct ngpb_testbed {
	std::unique_ptr<Testbed> impl;
};

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

int ngpb_testbed_create(ngpb_testbed_t** out) {
	NGPB_API_BEGIN
	require_device();
	if (!out) throw std::runtime_error{"ngpb_testbed_create: null argument"};
	*out = new ngpb_testbed{std::make_unique<Testbed>()};
	NGPB_API_END
}
void ngpb_testbed_free(ngpb_testbed_t* t) { delete t; }

void ngpb_profile_enable(int on) { prof_enable(on != 0); }
size_t ngpb_profile_collect(char* buf, size_t cap, int reset) {
	try { return copy_out(prof_collect(reset != 0), buf, cap); } catch (...) { return 0; }
}
int ngpb_testbed_set_stream(ngpb_testbed_t* t, void* stream) {
	NGPB_API_BEGIN
	t->impl->set_stream((cudaStream_t)stream);
	NGPB_API_END
}

int ngpb_testbed_create_empty_nerf_dataset(ngpb_testbed_t* t, uint32_t n_images, int aabb_scale, int is_hdr) {
	NGPB_API_BEGIN
	t->impl->create_empty_nerf_dataset(n_images, aabb_scale, is_hdr != 0);
	NGPB_API_END
}
int ngpb_testbed_set_nerf_image(ngpb_testbed_t* t, uint32_t idx, const void* pixels, int image_data_type, int width, int height) {
	NGPB_API_BEGIN
	t->impl->set_nerf_image(idx, pixels, image_data_type, width, height);
	NGPB_API_END
}
int ngpb_testbed_set_nerf_camera(ngpb_testbed_t* t, uint32_t idx, const float* xform_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, int lens_mode, const float* lens_params)
{
	NGPB_API_BEGIN
	t->impl->set_nerf_camera(idx, xform_3x4, focal_x, focal_y, principal_x, principal_y, lens_mode, lens_params);
	NGPB_API_END
}
int ngpb_testbed_reload_network_from_json(ngpb_testbed_t* t, const char* config_json) {
	NGPB_API_BEGIN
	if (!config_json) throw std::runtime_error{"ngpb_testbed_reload_network_from_json: null argument"};
	t->impl->reload_network_from_json(Json::parse(config_json));
	NGPB_API_END
}
int ngpb_testbed_reload_network_from_file(ngpb_testbed_t* t, const char* path) {
	NGPB_API_BEGIN
	if (!path) throw std::runtime_error{"ngpb_testbed_reload_network_from_file: null argument"};
	t->impl->reload_network_from_file(path);
	NGPB_API_END
}
int ngpb_testbed_load_file(ngpb_testbed_t* t, const char* path) {
	NGPB_API_BEGIN
	if (!path) throw std::runtime_error{"ngpb_testbed_load_file: null argument"};
	t->impl->load_file(path);
	NGPB_API_END
}
int ngpb_testbed_load_training_data(ngpb_testbed_t* t, const char* path, uint32_t max_images) {
	NGPB_API_BEGIN
	if (!path) throw std::runtime_error{"ngpb_testbed_load_training_data: null argument"};
	t->impl->load_training_data(path, max_images);
	NGPB_API_END
}
int ngpb_testbed_save_snapshot(ngpb_testbed_t* t, const char* path, int include_optimizer_state, int compress) {
	NGPB_API_BEGIN
	if (!path) throw std::runtime_error{"ngpb_testbed_save_snapshot: null argument"};
	t->impl->save_snapshot(path, include_optimizer_state != 0, compress != 0);
	NGPB_API_END
}
int ngpb_testbed_load_snapshot(ngpb_testbed_t* t, const char* path) {
	NGPB_API_BEGIN
	if (!path) throw std::runtime_error{"ngpb_testbed_load_snapshot: null argument"};
	t->impl->load_snapshot(std::string{path});
	NGPB_API_END
}
int ngpb_testbed_image_metadata(const ngpb_testbed_t* t, uint32_t idx, ngpb_nerf_image_t* out) {
	NGPB_API_BEGIN
	if (!out) throw std::runtime_error{"ngpb_testbed_image_metadata: null argument"};
	const ImageMeta m = t->impl->image_meta(idx);
	static_assert(sizeof(ngpb_nerf_image_t) == sizeof(ImageMeta), "ngpb_nerf_image_t must mirror ImageMeta");
	memcpy(out, &m, sizeof(m));
	NGPB_API_END
}
int ngpb_testbed_view_state(ngpb_testbed_t* t, float* values24, int set) {
	NGPB_API_BEGIN
	if (!values24) throw std::runtime_error{"ngpb_testbed_view_state: null argument"};
	ViewState& v = t->impl->view;
	float* fields[] = {v.camera, v.relative_focal_length, v.screen_center, &v.zoom, v.up_dir, v.sun_dir};
	const int counts[] = {12, 2, 2, 1, 3, 3};
	int k = 0;
	for (int f = 0; f < 6; ++f) for (int i = 0; i < counts[f]; ++i, ++k) { if (set) fields[f][i] = values24[k]; else values24[k] = fields[f][i]; }
	if (set) v.fov_axis = (int)values24[23]; else values24[23] = (float)v.fov_axis;
	NGPB_API_END
}
size_t ngpb_testbed_dataset_info(const ngpb_testbed_t* t, char* buf, size_t cap) {
	try {
		const DatasetInfo& d = t->impl->dataset_info;
		Json j = Json::object();
		j["scale"] = (double)d.scale;
		Json off = Json::array(), up = Json::array(), paths = Json::array();
		for (int k = 0; k < 3; ++k) { off.push_back(Json{(double)d.offset[k]}); up.push_back(Json{(double)d.up[k]}); }
		for (const std::string& p : d.paths) paths.push_back(Json{p});
		j["offset"] = off; j["up"] = up; j["paths"] = paths;
		j["from_mitsuba"] = d.from_mitsuba; j["is_hdr"] = d.is_hdr;
		if (d.has_render_aabb) {
			Json a = Json::array();
			for (int k = 0; k < 6; ++k) a.push_back(Json{(double)d.render_aabb[k]});
			j["render_aabb"] = a;
		}
		const std::string s = j.dump();
		if (buf && cap > 0) { const size_t n = s.size() < cap - 1 ? s.size() : cap - 1; memcpy(buf, s.data(), n); buf[n] = 0; }
		return s.size() + 1;
	} catch (...) { return 0; }
}
int ngpb_testbed_set_dataset_info(ngpb_testbed_t* t, const char* info_json) {
	NGPB_API_BEGIN
	if (!info_json) throw std::runtime_error{"ngpb_testbed_set_dataset_info: null argument"};
	const Json j = Json::parse(info_json);
	DatasetInfo d = t->impl->dataset_info;
	if (j.contains("scale")) d.scale = j["scale"].as_float();
	for (int k = 0; k < 3; ++k) {
		if (j.contains("offset")) d.offset[k] = j["offset"][k].as_float();
		if (j.contains("up")) d.up[k] = j["up"][k].as_float();
	}
	if (j.contains("paths")) { d.paths.clear(); for (const Json& p : j["paths"].elements()) d.paths.push_back(p.as_string()); }
	if (j.contains("from_mitsuba")) d.from_mitsuba = j["from_mitsuba"].as_bool();
	if (j.contains("is_hdr")) d.is_hdr = j["is_hdr"].as_bool();
	if (j.contains("render_aabb") && j["render_aabb"].is_array()) { d.has_render_aabb = true; for (int k = 0; k < 6; ++k) d.render_aabb[k] = j["render_aabb"][k].as_float(); }
	else if (j.contains("render_aabb")) d.has_render_aabb = false;
	t->impl->dataset_info = d;
	NGPB_API_END
}
int ngpb_testbed_reset_network(ngpb_testbed_t* t, int clear_density_grid) {
	NGPB_API_BEGIN
	t->impl->reset_network(clear_density_grid != 0);
	NGPB_API_END
}
int ngpb_testbed_train(ngpb_testbed_t* t, uint32_t batch_size) {
	NGPB_API_BEGIN
	t->impl->train(batch_size);
	NGPB_API_END
}

int ngpb_testbed_train_begin(ngpb_testbed_t* t, uint32_t batch_size) {
	NGPB_API_BEGIN
	t->impl->train_begin(batch_size);
	NGPB_API_END
}
int ngpb_testbed_train_forward(ngpb_testbed_t* t, uint32_t batch_size) {
	NGPB_API_BEGIN
	t->impl->train_forward(batch_size);
	NGPB_API_END
}
int ngpb_testbed_train_backward(ngpb_testbed_t* t, uint32_t batch_size) {
	NGPB_API_BEGIN
	t->impl->train_backward(batch_size, true);
	NGPB_API_END
}
int ngpb_nccl_get_unique_id(void* out128) {
	NGPB_API_BEGIN
	if (!out128) throw std::runtime_error{"ngpb_nccl_get_unique_id: null argument"};
	static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
	ncclUniqueId id;
	NGPB_NCCL_CHECK(NcclApi::get().GetUniqueId(&id));
	memcpy(out128, &id, sizeof(id));
	NGPB_API_END
}
int ngpb_testbed_init_data_parallel(ngpb_testbed_t* t, const void* id_main, const void* id_aux, int rank, int world) {
	NGPB_API_BEGIN
	if (world > 1 && (!id_main || !id_aux)) throw std::runtime_error{"ngpb_testbed_init_data_parallel: null unique id"};
	t->impl->init_data_parallel(id_main, id_aux, rank, world);
	NGPB_API_END
}
int ngpb_testbed_train_end(ngpb_testbed_t* t, uint32_t batch_size) {
	NGPB_API_BEGIN
	t->impl->train_end(batch_size);
	NGPB_API_END
}

#define NGPB_PROPS(X) \
	X(rays_per_batch) X(measured_batch_size) X(measured_batch_size_before_compaction) X(seed) X(random_bg_color) X(linear_colors) \
	X(snap_to_pixel_centers) X(near_distance) X(density_grid_decay) X(render_min_transmittance) X(loss_type) X(rgb_activation) \
	X(density_activation) X(color_space) X(train_network) X(train_encoding) X(aabb_scale) \
	X(render_target_queries) X(render_min_steps) X(render_max_steps) X(render_windows_per_sync) X(render_skip_past_last_cascade) X(render_snap_to_pixel_centers) \
	X(shard_rank) X(shard_world) X(render_shard_rank) X(render_shard_world) X(image_snap_to_pixel_centers) X(train_prefetch) \
	X(exposure) X(tonemap_curve) X(render_color_space) X(render_with_lens_distortion)

int ngpb_testbed_get(const ngpb_testbed_t* t, const char* name, double* value) {
	NGPB_API_BEGIN
	const Testbed& tb = *t->impl;
	const std::string n = name;
	if (n == "training_step") *value = tb.training_step();
	else if (n == "loss") *value = tb.loss();
	else if (n == "n_params") *value = (double)tb.n_params();
	else if (n == "max_cascade") *value = tb.max_cascade();
	else if (n == "cone_angle_constant") *value = tb.cone_angle_constant();
	else if (n == "n_images") *value = tb.n_images();
	else if (n == "last_render_samples") *value = (double)tb.last_render_samples;
	else if (n == "prefetches_adopted") *value = (double)tb.prefetches_adopted;
	else if (n == "l2_window_bytes") *value = tb.l2_window_bytes;
	else if (n == "l2_window_hit_ratio") *value = tb.l2_window_hit_ratio;
	else if (n == "optimizer_step") *value = const_cast<Testbed&>(tb).optimizer() ? const_cast<Testbed&>(tb).optimizer()->step_count() : 0;
	else if (n == "optimizer_learning_rate_factor") *value = const_cast<Testbed&>(tb).optimizer() ? const_cast<Testbed&>(tb).optimizer()->learning_rate_factor() : 1.0;
	else if (n == "optimizer_base_learning_rate") *value = const_cast<Testbed&>(tb).optimizer() ? const_cast<Testbed&>(tb).optimizer()->base_learning_rate() : 0.0;
	else if (n == "background_color_r") *value = tb.background_color[0];
	else if (n == "background_color_g") *value = tb.background_color[1];
	else if (n == "background_color_b") *value = tb.background_color[2];
	else if (n == "background_color_a") *value = tb.background_color[3];
#define X(f) else if (n == #f) *value = (double)tb.f;
	NGPB_PROPS(X)
#undef X
	else throw std::runtime_error{"ngpb_testbed_get: unknown property " + n};
	NGPB_API_END
}
int ngpb_testbed_set(ngpb_testbed_t* t, const char* name, double value) {
	NGPB_API_BEGIN
	Testbed& tb = *t->impl;
	const std::string n = name;
	if (n == "optimizer_step" || n == "optimizer_learning_rate_factor" || n == "optimizer_base_learning_rate") {
		// Adam's current_step; ExponentialDecay's learning_rate_factor and learning_rate (tcnn adam.h:288-299, exponential_decay.h:144-148)
		Optimizer* o = tb.optimizer();
		if (!o) throw std::runtime_error{"ngpb_testbed_set: no optimizer (reload_network_from_json first)"};
		if (n == "optimizer_step") o->set_step_count((uint32_t)value, o->learning_rate_factor());
		else if (n == "optimizer_learning_rate_factor") o->set_step_count(o->step_count(), (float)value);
		else o->set_base_learning_rate((float)value);
	}
	else if (n == "cone_angle_constant") {
		if (!(value >= 0.0)) throw std::runtime_error{"ngpb_testbed_set: cone_angle_constant must be >= 0"};
		tb.set_cone_angle_constant((float)value);
	}
	else if (n == "background_color_r") tb.background_color[0] = (float)value;
	else if (n == "background_color_g") tb.background_color[1] = (float)value;
	else if (n == "background_color_b") tb.background_color[2] = (float)value;
	else if (n == "background_color_a") tb.background_color[3] = (float)value;
#define X(f) else if (n == #f) tb.f = (decltype(tb.f))value;
	NGPB_PROPS(X)
#undef X
	else throw std::runtime_error{"ngpb_testbed_set: unknown or read-only property " + n};
	NGPB_API_END
}

int ngpb_testbed_device_ptr(ngpb_testbed_t* t, const char* name, void** ptr, size_t* bytes) {
	NGPB_API_BEGIN
	Testbed& tb = *t->impl;
	const std::string n = name;
	void* p = nullptr;
	size_t b = 0;
	const size_t n_grid = (size_t)kNerfGridCells * (tb.max_cascade() + 1);
	// train() returns with the backward pass and the optimizer still in flight on the testbed's stream
	NGPB_CUDA_CHECK(cudaStreamSynchronize(tb.stream()));
	if (n == "params") { p = tb.params(); b = tb.n_params() * 2; }
	else if (n == "params_fp32") { p = tb.params_fp32(); b = tb.n_params() * 4; }
	else if (n == "params_inference") { p = tb.params_inference(); b = tb.n_params() * 2; }
	else if (n == "gradients") { p = tb.gradients(); b = tb.n_params() * 2; }
	else if (n == "optimizer_first_moments" && tb.optimizer()) { p = tb.optimizer()->first_moments(); b = tb.n_params() * 4; }
	else if (n == "optimizer_second_moments" && tb.optimizer()) { p = tb.optimizer()->second_moments(); b = tb.n_params() * 4; }
	else if (n == "optimizer_param_steps" && tb.optimizer()) { p = tb.optimizer()->param_steps(); b = tb.n_params() * 4; }
	else if (n == "density_grid") { p = tb.density_grid(); b = n_grid * 4; }
	else if (n == "density_grid_bitfield") { p = tb.density_grid_bitfield(); b = nerf_bitfield_bytes(); }
	else if (n == "density_grid_mean") { p = tb.density_grid_mean(); b = 4; }
	else if (n == "image_positions") { p = tb.image_buffer(0, &b); }
	else if (n == "image_targets") { p = tb.image_buffer(1, &b); }
	else if (n == "image_output") { p = tb.image_buffer(2, &b); }
	else if (n == "image_dloss") { p = tb.image_buffer(3, &b); }
	else if (n == "counters") { p = tb.counters(); b = sizeof(NerfStepCounters); }
	else if (n == "metadata") { tb.upload_metadata(); p = (void*)tb.metadata_gpu(); b = tb.n_images() * sizeof(ImageMeta); }
	else throw std::runtime_error{"ngpb_testbed_device_ptr: unknown buffer " + n};
	if (ptr) *ptr = p;
	if (bytes) *bytes = b;
	NGPB_API_END
}

size_t ngpb_testbed_network_config(const ngpb_testbed_t* t, char* buf, size_t cap) {
	try { return copy_out(t->impl->network_config().dump(), buf, cap); } catch (...) { return 0; }
}

int ngpb_testbed_update_density_grid(ngpb_testbed_t* t, float decay, uint32_t n_uniform, uint32_t n_nonuniform) {
	NGPB_API_BEGIN
	t->impl->update_density_grid_nerf(decay, n_uniform, n_nonuniform);
	NGPB_CUDA_CHECK(cudaStreamSynchronize(t->impl->stream()));
	NGPB_API_END
}

int ngpb_testbed_set_params(ngpb_testbed_t* t, const void* host_params, int full_precision, size_t n_params) {
	NGPB_API_BEGIN
	t->impl->set_params(host_params, full_precision != 0, n_params);
	NGPB_API_END
}
int ngpb_testbed_set_density_grid_fp16(ngpb_testbed_t* t, const void* host_grid, size_t n_cells) {
	NGPB_API_BEGIN
	t->impl->set_density_grid_fp16(host_grid, n_cells);
	NGPB_API_END
}
int ngpb_testbed_restore_training_state(ngpb_testbed_t* t, uint32_t training_step, float loss) {
	NGPB_API_BEGIN
	t->impl->restore_training_state(training_step, loss);
	NGPB_API_END
}

int ngpb_testbed_render(ngpb_testbed_t* t, uint32_t width, uint32_t height, uint32_t spp, int linear, const float* camera_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, float* rgba)
{
	NGPB_API_BEGIN
	if (!camera_3x4 || !rgba) throw std::runtime_error{"ngpb_testbed_render: null argument"};
	t->impl->render(width, height, spp, linear != 0, camera_3x4, focal_x, focal_y, principal_x, principal_y, rgba);
	NGPB_API_END
}

int ngpb_testbed_set_render_lens(ngpb_testbed_t* t, int lens_mode, const float* lens_params) {
	NGPB_API_BEGIN
	t->impl->set_render_lens(lens_mode, lens_params);
	NGPB_API_END
}

int ngpb_testbed_overlay_ground_truth(ngpb_testbed_t* t, uint32_t width, uint32_t height, int linear, uint32_t view, int fov_axis, float zoom, float alpha, float* rgba) {
	NGPB_API_BEGIN
	t->impl->overlay_ground_truth(width, height, linear != 0, view, fov_axis, zoom, alpha, rgba);
	NGPB_API_END
}

int ngpb_testbed_set_image(ngpb_testbed_t* t, const void* pixels, int image_data_type, int width, int height) {
	NGPB_API_BEGIN
	t->impl->set_image(pixels, image_data_type, width, height);
	NGPB_API_END
}
int ngpb_testbed_render_image(ngpb_testbed_t* t, uint32_t width, uint32_t height, float* rgba) {
	NGPB_API_BEGIN
	t->impl->render_image(width, height, rgba);
	NGPB_API_END
}

int ngpb_testbed_set_render_steps_out(ngpb_testbed_t* t, uint32_t* steps) {
	NGPB_API_BEGIN
	t->impl->set_render_steps_out(steps);
	NGPB_API_END
}

int ngpb_testbed_render_nerf(ngpb_testbed_t* t, uint32_t width, uint32_t height, const float* camera_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, uint32_t spp_index, uint32_t pixel_begin, uint32_t pixel_end, float* rgba, float* depth)
{
	NGPB_API_BEGIN
	t->impl->render_nerf(width, height, camera_3x4, focal_x, focal_y, principal_x, principal_y, spp_index, pixel_begin, pixel_end, rgba, depth);
	NGPB_API_END
}

} // extern "C"
