// Testbed::save_snapshot / load_snapshot in host C++ (src/testbed.cu:4775-4972). See host_io.cu for the container.
#include "testbed.h"
#include "host_io.h"

#include <algorithm>

namespace ngpb {

constexpr int kSnapshotFormatVersion = 1;   // src/testbed.cu:4773

static Json jvec(const float* v, int n) {
	Json a = Json::array();
	for (int i = 0; i < n; ++i) a.push_back(Json{(double)v[i]});
	return a;
}
// a tmat<float, C, R> is a list of ROWS (tcnn vec_json.h)
static Json jmat_rows(const float* row_major, int rows, int cols) {
	Json m = Json::array();
	for (int r = 0; r < rows; ++r) m.push_back(jvec(row_major + r * cols, cols));
	return m;
}
static Json jaabb(int aabb_scale) {
	// m_aabb of a NeRF scene: the unit cube's centre inflated by half the scale (load_nerf_post, src/testbed_nerf.cu:2700-2712)
	const float h = 0.5f * (float)aabb_scale;
	const float lo[3] = {0.5f - h, 0.5f - h, 0.5f - h}, hi[3] = {0.5f + h, 0.5f + h, 0.5f + h};
	Json j = Json::object();
	j["min"] = jvec(lo, 3);
	j["max"] = jvec(hi, 3);
	return j;
}
static const float kIdentity3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};

// to_json(Lens), json_binding.h:37-63: Perspective leaves the value null
static Json lens_to_json(const ImageMeta& m) {
	if (m.lens_mode == kLensOpenCV) {
		Json j = Json::object();
		j["is_fisheye"] = false;
		j["k1"] = (double)m.lens_params[0]; j["k2"] = (double)m.lens_params[1];
		j["p1"] = (double)m.lens_params[2]; j["p2"] = (double)m.lens_params[3];
		return j;
	}
	return Json{};
}
// from_json(Lens), json_binding.h:65-96
static void lens_from_json(const Json& j, int& mode, float (&params)[7]) {
	mode = kLensPerspective;
	for (float& p : params) p = 0.0f;
	if (!j.is_object()) return;
	if (j.contains("k1")) {
		if (j.value("is_fisheye", false)) throw std::runtime_error{"OpenCV fisheye lenses are not part of the path"};
		mode = kLensOpenCV;
		params[0] = j["k1"].as_float(); params[1] = j["k2"].as_float(); params[2] = j["p1"].as_float(); params[3] = j["p2"].as_float();
	} else if (j.contains("ftheta_p0") || j.contains("latlong") || j.contains("equirectangular")) {
		throw std::runtime_error{"F-theta / lat-long / equirectangular lenses are not part of the path"};
	}
}

template <typename T>
static Json device_to_binary(const T* dev, size_t n, cudaStream_t stream) {
	std::vector<uint8_t> host(n * sizeof(T));
	if (n) {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(host.data(), dev, host.size(), cudaMemcpyDeviceToHost, stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
	}
	return Json::binary(std::move(host));
}

// from_json(GPUMemory<T>) (tcnn gpu_memory_json.h:52-71): msgpack bin, or the {"bytes": [...]} object of text JSON
static std::vector<uint8_t> binary_of(const Json& j) {
	if (j.is_binary()) return j.as_binary();
	if (j.is_object() && j.contains("bytes")) {
		std::vector<uint8_t> out;
		for (const Json& b : j["bytes"].elements()) out.push_back((uint8_t)b.as_uint());
		return out;
	}
	throw std::runtime_error{"Invalid json type: must be either binary or object"};
}

static std::string lower(std::string s) {
	for (char& c : s) c = (char)tolower((unsigned char)c);
	return s;
}

// Ema{ExponentialDecay{Adam}}::serialize, following the chain the network config names
static Json optimizer_to_json(Testbed& tb, const Json& cfg) {
	Optimizer& o = *tb.optimizer();
	const std::string otype = lower(cfg.value("otype", "Adam"));
	const size_t n = tb.n_params();
	Json j = Json::object();
	if (otype == "ema") {
		j["nested"] = optimizer_to_json(tb, cfg.value("nested", Json::object()));
		j["weights_ema_binary"] = device_to_binary(tb.params_inference(), n, tb.stream());
	} else if (otype == "exponentialdecay") {
		j["nested"] = optimizer_to_json(tb, cfg.value("nested", Json::object()));
		j["learning_rate"] = (double)o.base_learning_rate();
		j["learning_rate_factor"] = (double)o.learning_rate_factor();
	} else if (otype == "adam") {
		j["current_step"] = o.step_count();
		j["base_learning_rate"] = (double)(o.base_learning_rate() * o.learning_rate_factor());
		j["first_moments_binary"] = device_to_binary(o.first_moments(), n, tb.stream());
		j["second_moments_binary"] = device_to_binary(o.second_moments(), n, tb.stream());
		j["param_steps_binary"] = device_to_binary(o.param_steps(), n, tb.stream());
	} else {
		throw std::runtime_error{"optimizer '" + cfg.value("otype", "Adam") + "' is not part of the path"};
	}
	return j;
}

static void put_device(void* dst, const Json& blob, size_t bytes, const char* what, cudaStream_t stream) {
	const std::vector<uint8_t> src = binary_of(blob);
	if (src.size() != bytes) throw std::runtime_error{std::string{"optimizer state '"} + what + "' has " + std::to_string(src.size()) + " bytes, the network needs " + std::to_string(bytes)};
	NGPB_CUDA_CHECK(cudaMemcpyAsync(dst, src.data(), bytes, cudaMemcpyHostToDevice, stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
}

// the deserialize() counterparts (ema.h:197-206, exponential_decay.h:144-148, adam.h:288-299). The EMA weights need no
// copy of their own: Trainer::set_params has already written the same bytes (params_binary IS the EMA copy, trainer.h:281).
static void optimizer_from_json(Testbed& tb, const Json& cfg, const Json& d, bool has_decay) {
	Optimizer& o = *tb.optimizer();
	const std::string otype = lower(cfg.value("otype", "Adam"));
	const size_t n = tb.n_params();
	if (otype == "ema") {
		optimizer_from_json(tb, cfg.value("nested", Json::object()), d["nested"], has_decay);
	} else if (otype == "exponentialdecay") {
		o.set_base_learning_rate(d["learning_rate"].as_float());
		o.set_step_count(o.step_count(), d.value("learning_rate_factor", 1.0f));
		optimizer_from_json(tb, cfg.value("nested", Json::object()), d["nested"], true);
	} else if (otype == "adam") {
		put_device(o.first_moments(), d["first_moments_binary"], n * sizeof(float), "first_moments_binary", tb.stream());
		put_device(o.second_moments(), d["second_moments_binary"], n * sizeof(float), "second_moments_binary", tb.stream());
		if (d.contains("param_steps_binary")) put_device(o.param_steps(), d["param_steps_binary"], n * sizeof(uint32_t), "param_steps_binary", tb.stream());
		else NGPB_CUDA_CHECK(cudaMemsetAsync(o.param_steps(), 0, n * sizeof(uint32_t), tb.stream()));
		o.set_step_count(d["current_step"].as_uint(), o.learning_rate_factor());
		if (!has_decay) o.set_base_learning_rate(d["base_learning_rate"].as_float());
	} else {
		throw std::runtime_error{"optimizer '" + cfg.value("otype", "Adam") + "' is not part of the path"};
	}
}

// to_json(NerfDataset), json_binding.h:105-132
static Json dataset_to_json(Testbed& tb, const std::vector<ImageMeta>& metas) {
	const DatasetInfo& ds = tb.dataset_info;
	const uint32_t n = (uint32_t)metas.size();
	Json paths = Json::array(), meta = Json::array(), xforms = Json::array();
	for (uint32_t i = 0; i < n; ++i) {
		const ImageMeta& m = metas[i];
		paths.push_back(Json{i < ds.paths.size() ? ds.paths[i] : std::string{}});
		Json e = Json::object();
		const float focal[2] = {m.focal_x, m.focal_y > 0.0f ? m.focal_y : m.focal_x};
		const float principal[2] = {m.principal_x, m.principal_y};
		const float shutter[4] = {0, 0, 0, 0};
		e["focal_length"] = jvec(focal, 2);
		e["lens"] = lens_to_json(m);
		e["principal_point"] = jvec(principal, 2);
		e["rolling_shutter"] = jvec(shutter, 4);
		Json res = Json::array();
		res.push_back(Json{m.res_x}); res.push_back(Json{m.res_y});
		e["resolution"] = res;
		meta.push_back(e);
		float rows[12];
		for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) rows[r * 4 + c] = (&m.xform.c[c].x)[r];
		Json x = Json::object();
		x["start"] = jmat_rows(rows, 3, 4);
		x["end"] = jmat_rows(rows, 3, 4);
		xforms.push_back(x);
	}
	Json j = Json::object();
	j["n_images"] = n;
	j["paths"] = paths;
	j["metadata"] = meta;
	j["xforms"] = xforms;
	if (ds.has_render_aabb) {
		Json a = Json::object();
		a["min"] = jvec(ds.render_aabb, 3);
		a["max"] = jvec(ds.render_aabb + 3, 3);
		j["render_aabb"] = a;
	} else {
		j["render_aabb"] = jaabb(tb.aabb_scale);
	}
	j["render_aabb_to_local"] = jmat_rows(kIdentity3, 3, 3);
	j["up"] = jvec(ds.up, 3);
	j["offset"] = jvec(ds.offset, 3);
	Json env = Json::array();
	env.push_back(Json{0}); env.push_back(Json{0});
	j["envmap_resolution"] = env;
	j["scale"] = (double)ds.scale;
	j["aabb_scale"] = tb.aabb_scale;
	j["from_mitsuba"] = ds.from_mitsuba;
	j["is_hdr"] = ds.is_hdr;
	j["wants_importance_sampling"] = true;
	j["n_extra_learnable_dims"] = 0;
	return j;
}

// per-image pose / latent optimizers of the reference (AdamOptimizer<vec3>, adam_optimizer.h): never stepped on this path,
// written in their initial state as the reference writes them for an untouched camera
static Json idle_adam(int dims) {
	Json j = Json::object();
	j["iter"] = 0;
	j["beta1"] = (double)0.9f; j["beta2"] = (double)0.99f; j["epsilon"] = (double)1e-8f; j["learning_rate"] = (double)1e-4f;
	Json zeros = Json::array();
	for (int i = 0; i < dims; ++i) zeros.push_back(Json{0.0});
	j["first_moment"] = zeros; j["second_moment"] = zeros; j["variable"] = zeros;
	return j;
}

// The config object save_snapshot writes: m_network_config with ["snapshot"] filled in (src/testbed.cu:4776-4824;
// Trainer::serialize, tcnn trainer.h:275-288).
Json Testbed::snapshot_config(bool include_optimizer_state) {
	if (m_mode != kModeNerf) throw std::runtime_error{"ngp-b200: snapshots are written for the NeRF mode"};
	if (!m_network || m_n_params == 0) throw std::runtime_error{"save_snapshot: there is no network to save (reload_network_from_json first)"};
	join_prefetch();
	Json config = m_network_config;
	Json snap = Json::object();
	snap["n_params"] = (uint64_t)m_n_params;
	snap["params_type"] = "__half";
	snap["params_binary"] = device_to_binary(params_inference(), m_n_params, m_stream);
	if (include_optimizer_state) snap["optimizer"] = optimizer_to_json(*this, m_network_config.value("optimizer", Json::object()));
	snap["version"] = kSnapshotFormatVersion;
	snap["mode"] = "nerf";
	snap["density_grid_size"] = (uint32_t)kNerfGridSize;
	{
		const size_t n_grid = (size_t)kNerfGridCells * (m_max_cascade + 1);
		m_density_grid_tmp.reserve(n_grid * sizeof(float), m_stream);
		cast_fp32_to_fp16(m_stream, n_grid, m_density_grid.as<float>(), m_density_grid_tmp.as<__half>());
		snap["density_grid_binary"] = device_to_binary(m_density_grid_tmp.as<__half>(), n_grid, m_stream);
	}
	std::vector<ImageMeta> metas(m_images.size());
	for (size_t i = 0; i < m_images.size(); ++i) metas[i] = m_images[i].meta;
	Json nerf = Json::object();
	nerf["aabb_scale"] = aabb_scale;
	Json pos = Json::array(), rot = Json::array(), extra = Json::array();
	for (size_t i = 0; i < metas.size(); ++i) { pos.push_back(idle_adam(3)); rot.push_back(idle_adam(3)); extra.push_back(idle_adam(0)); }
	nerf["cam_pos_offset"] = pos; nerf["cam_rot_offset"] = rot; nerf["extra_dims_opt"] = extra;
	Json rgb = Json::object();
	rgb["rays_per_batch"] = rays_per_batch;
	rgb["measured_batch_size"] = measured_batch_size;
	rgb["measured_batch_size_before_compaction"] = measured_batch_size_before_compaction;
	nerf["rgb"] = rgb;
	nerf["dataset"] = dataset_to_json(*this, metas);
	snap["nerf"] = nerf;
	snap["training_step"] = m_training_step;
	snap["loss"] = (double)m_loss_scalar;
	snap["aabb"] = jaabb(aabb_scale);
	snap["bounding_radius"] = 1.0;
	snap["render_aabb_to_local"] = jmat_rows(kIdentity3, 3, 3);
	snap["render_aabb"] = nerf["dataset"]["render_aabb"];
	snap["up_dir"] = jvec(view.up_dir, 3);
	snap["sun_dir"] = jvec(view.sun_dir, 3);
	snap["exposure"] = (double)exposure;
	snap["background_color"] = jvec(background_color, 4);
	Json cam = Json::object();
	cam["matrix"] = jmat_rows(view.camera, 3, 4);
	cam["fov_axis"] = view.fov_axis;
	cam["relative_focal_length"] = jvec(view.relative_focal_length, 2);
	cam["screen_center"] = jvec(view.screen_center, 2);
	cam["zoom"] = (double)view.zoom;
	cam["scale"] = 1.5;
	cam["aperture_size"] = 0.0;
	cam["autofocus"] = false;
	const float target[3] = {0.5f, 0.5f, 0.5f};
	cam["autofocus_target"] = jvec(target, 3);
	cam["autofocus_depth"] = 0.5;
	snap["camera"] = cam;
	config["snapshot"] = snap;
	return config;
}

void Testbed::save_snapshot(const std::string& path, bool include_optimizer_state, bool compress) {
	write_snapshot_container(path, snapshot_config(include_optimizer_state), compress);
}

void Testbed::load_snapshot(const std::string& path) {
	const Json config = load_network_config(path);
	if (!config.contains("snapshot")) throw std::runtime_error{"File '" + path + "' does not contain a snapshot."};
	load_snapshot(config);
}

// Testbed::load_snapshot(nlohmann::json config), src/testbed.cu:4841-4955
void Testbed::load_snapshot(const Json& config) {
	if (!config.contains("snapshot")) throw std::runtime_error{"Given stream does not contain a snapshot."};
	const Json& snap = config["snapshot"];
	if (snap.value("version", 0) < kSnapshotFormatVersion) throw std::runtime_error{"Snapshot uses an old format and can not be loaded."};
	if (snap.contains("mode")) {
		const std::string mode = snap["mode"].as_string();
		if (mode != "nerf" && mode != "sdf" && mode != "image" && mode != "volume") throw std::runtime_error{"Unknown snapshot mode. Snapshot must be regenerated with a new version of instant-ngp."};
		if (mode != "nerf") throw std::runtime_error{"snapshot mode '" + mode + "' is not part of the path (NeRF snapshots only)"};
	}
	if (m_mode == kModeImage) throw std::runtime_error{"ngp-b200: this testbed is in image mode; create a new Testbed for a NeRF snapshot"};
	if (snap["density_grid_size"].as_uint() != kNerfGridSize) throw std::runtime_error{"Incompatible grid size."};
	const Json& nerf = snap["nerf"];
	if (n_images() == 0 && nerf.contains("dataset")) {
		// no scene loaded: the dataset's metadata comes from the snapshot, enough to render (:4867-4870)
		const Json& ds = nerf["dataset"];
		const uint32_t n = ds["n_images"].as_uint();
		create_empty_nerf_dataset(n, ds["aabb_scale"].as_int(), ds.value("is_hdr", false));
		DatasetInfo info;
		for (uint32_t i = 0; i < n; ++i) {
			const Json& m = ds["metadata"][i];
			ImageMeta& meta = m_images[i].meta;
			meta.res_x = m["resolution"][0].as_int();
			meta.res_y = m["resolution"][1].as_int();
			int lens_mode; float lens[7];
			lens_from_json(m.contains("lens") ? m["lens"] : m.value("camera_distortion", Json{}), lens_mode, lens);
			float x[12];
			const Json& start = ds["xforms"][i]["start"];
			for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) x[r * 4 + c] = start[r][c].as_float();
			set_nerf_camera(i, x, m["focal_length"][0].as_float(), m["focal_length"][1].as_float(),
				m["principal_point"][0].as_float(), m["principal_point"][1].as_float(), lens_mode, lens);
			info.paths.push_back(ds.contains("paths") && ds["paths"].is_array() && i < ds["paths"].size() ? ds["paths"][i].as_string() : std::string{});
		}
		info.scale = ds["scale"].as_float();
		for (int k = 0; k < 3; ++k) { info.offset[k] = ds["offset"][k].as_float(); info.up[k] = ds["up"][k].as_float(); }
		if (ds.contains("render_aabb") && ds["render_aabb"].is_object()) {
			info.has_render_aabb = true;
			for (int k = 0; k < 3; ++k) { info.render_aabb[k] = ds["render_aabb"]["min"][k].as_float(); info.render_aabb[3 + k] = ds["render_aabb"]["max"][k].as_float(); }
		}
		info.is_hdr = ds.value("is_hdr", false);
		info.from_mitsuba = ds.value("from_mitsuba", false);
		dataset_info = info;
	} else if (nerf.contains("aabb_scale") && nerf["aabb_scale"].as_int() != aabb_scale) {
		// the reference re-derives the scene box from the snapshot's scale here; the images on the device were laid out for the other one
		throw std::runtime_error{"ngp-b200: the snapshot's aabb_scale (" + std::to_string(nerf["aabb_scale"].as_int()) + ") differs from the loaded scene's (" + std::to_string(aabb_scale) + ")"};
	}
	exposure = snap.value("exposure", exposure);
	if (snap.contains("background_color")) for (int k = 0; k < 4; ++k) background_color[k] = snap["background_color"][k].as_float();
	if (snap.contains("camera")) {
		const Json& cam = snap["camera"];
		if (cam.contains("matrix")) for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) view.camera[r * 4 + c] = cam["matrix"][r][c].as_float();
		view.fov_axis = cam.value("fov_axis", view.fov_axis);
		if (cam.contains("relative_focal_length")) for (int k = 0; k < 2; ++k) view.relative_focal_length[k] = cam["relative_focal_length"][k].as_float();
		if (cam.contains("screen_center")) for (int k = 0; k < 2; ++k) view.screen_center[k] = cam["screen_center"][k].as_float();
		view.zoom = cam.value("zoom", view.zoom);
	}
	if (snap.contains("up_dir")) for (int k = 0; k < 3; ++k) view.up_dir[k] = snap["up_dir"][k].as_float();
	if (snap.contains("sun_dir")) for (int k = 0; k < 3; ++k) view.sun_dir[k] = snap["sun_dir"][k].as_float();

	// m_network_config = config; reset_network(false) rebuilds the trainer; the occupancy grid is the snapshot's
	Json network_config = config;
	network_config.erase("snapshot");
	reload_network_from_json(network_config);
	{
		const std::vector<uint8_t> grid = binary_of(snap["density_grid_binary"]);
		set_density_grid_fp16(grid.data(), grid.size() / sizeof(__half));
	}
	restore_training_state(snap["training_step"].as_uint(), snap["loss"].as_float());
	// Trainer::deserialize (tcnn trainer.h:290-314)
	const std::string kind = snap.value("params_type", "__half");
	if (kind != "float" && kind != "__half") throw std::runtime_error{"Trainer: snapshot parameters must be of type float of __half"};
	{
		const std::vector<uint8_t> params = binary_of(snap["params_binary"]);
		set_params(params.data(), kind == "float", params.size() / (kind == "float" ? sizeof(float) : sizeof(__half)));
	}
	if (snap.contains("optimizer")) {
		optimizer_from_json(*this, m_network_config.value("optimizer", Json::object()), snap["optimizer"], false);
		NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	}
	const Json& rgb = nerf["rgb"];
	rays_per_batch = rgb["rays_per_batch"].as_uint();
	measured_batch_size = rgb["measured_batch_size"].as_uint();
	measured_batch_size_before_compaction = rgb["measured_batch_size_before_compaction"].as_uint();
}

} // namespace ngpb
