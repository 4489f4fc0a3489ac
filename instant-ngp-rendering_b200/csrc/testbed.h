// Testbed: the host object behind the `ngpb_testbed_*` C ABI. Mirrors the train / render entry
// points of ngp::Testbed for the NeRF mode (include/neural-graphics-primitives/testbed.h:70-655,
// src/testbed.cu:3624-3868 reset_network, :4020-4109 train, src/testbed_nerf.cu:2271-2946) with
// the state the hot path needs and nothing of the GUI / camera-path / VR shell.
#pragma once

#include "nerf.h"
#include "network.h"
#include "optimizer.h"

#include <memory>
#include <string>
#include <vector>

namespace ngpb {

struct DistContext;
// Parameter buffers are padded to a multiple of this, so that any rank count dividing 256 gets 16-byte aligned shards.
constexpr uint32_t kParamPadding = 2048;

struct NerfImage {
	DeviceBuffer pixels;
	ImageMeta meta{};
};

// Host-side facts about the loaded scene that only the file-level entry points need (NerfDataset's scalar members,
// include/neural-graphics-primitives/nerf_loader.h:60-99): written by load_training_data / load_snapshot, saved by save_snapshot.
struct DatasetInfo {
	std::vector<std::string> paths;
	float scale = 0.33f;                 // NERF_SCALE
	float offset[3] = {0.5f, 0.5f, 0.5f};
	float up[3] = {0.0f, 1.0f, 0.0f};
	bool has_render_aabb = false;
	float render_aabb[6] = {};           // min xyz, max xyz
	bool from_mitsuba = false;
	bool is_hdr = false;
};

// The interactive camera of the reference's Testbed (m_camera, m_fov_axis, m_relative_focal_length, m_screen_center, m_zoom,
// m_up_dir, m_sun_dir; testbed.h:700-740), kept because snapshots carry it (src/testbed.cu:4806-4822, :4935-4950).
struct ViewState {
	float camera[12] = {1.0f, 0.0f, 0.0f, 0.5f, 0.0f, -1.0f, 0.0f, 0.5f, 0.0f, 0.0f, -1.0f, 0.5f};   // 3x4, rows
	int fov_axis = 1;
	float relative_focal_length[2] = {1.0f, 1.0f};
	float screen_center[2] = {0.5f, 0.5f};
	float zoom = 1.0f;
	float up_dir[3] = {0.0f, 1.0f, 0.0f};
	float sun_dir[3] = {0.57735027f, 0.57735027f, 0.57735027f};
};

class Testbed {
public:
	Testbed();
	~Testbed();

	// ---- dataset (Testbed::create_empty_nerf_dataset / set_image / set_camera_*, python_api.cu:383-567)
	void create_empty_nerf_dataset(uint32_t n_images, int aabb_scale, bool is_hdr);
	// pixels: HOST or DEVICE pointer to w*h RGBA texels of `image_data_type` (Byte = sRGB8 straight
	// alpha, Half / Float = linear premultiplied), copied to the device.
	void set_nerf_image(uint32_t idx, const void* pixels, int image_data_type, int width, int height);
	// camera-to-world 3x4 in the NGP convention (already passed through nerf_matrix_to_ngp),
	// row-major [3][4]; focal length in pixels; principal point in [0,1]; lens as ELensMode + params.
	void set_nerf_camera(uint32_t idx, const float* xform_3x4, float focal_x, float focal_y, float principal_x, float principal_y, int lens_mode, const float* lens_params);

	// ---- neural image mode (Testbed::load_image / train_image / render_image, src/testbed_image.cu)
	// pixels: HOST or DEVICE RGBA texels (Half or Float, linear colours); switches the testbed to
	// ETestbedMode::Image: reload_network_from_json then builds the 2-D NetworkWithInputEncoding and
	// train(batch) runs train_image.
	void set_image(const void* pixels, int image_data_type, int width, int height);
	// Network colour at every pixel centre of a width x height grid: DEVICE float4 [h][w], linear.
	void render_image(uint32_t width, uint32_t height, float* rgba_out);
	NetworkWithInputEncodingModule* image_network() { return m_image_network.get(); }
	bool image_snap_to_pixel_centers = true; // Image::Training::snap_to_pixel_centers (testbed.h:878)
	int mode() const { return m_mode; }
	// Buffers of the last train_image step, for parity tests: 0 positions fp32 [n][2], 1 targets fp32
	// [n][3], 2 network output fp16 [n][16], 3 dL/doutput fp16 [n][16].
	void* image_buffer(int which, size_t* bytes) {
		DeviceBuffer* b[4] = {&m_image_positions, &m_image_targets, &m_image_output, &m_image_dloss};
		const size_t per[4] = {8, 12, 32, 32};
		if (bytes) *bytes = (size_t)measured_batch_size * per[which];
		return b[which]->ptr;
	}

	// ---- network (reload_network_from_json, reset_network)
	void reload_network_from_json(const Json& config);
	void reset_network(bool clear_density_grid = true);

	// ---- training (Testbed::train)
	void train(uint32_t batch_size);
	// Data-parallel form of train(): train_begin leaves this rank's gradients (rays
	// [rank R/P, (rank+1) R/P) of the step's R rays, batch_size / P compacted samples) in
	// gradients() and its counters in counters(); the caller sums both over the ranks (NCCL
	// all-reduce on the same stream); train_end runs the optimizer and the batch-size control loop
	// on the summed values. batch_size is the GLOBAL batch in both calls.
	void train_begin(uint32_t batch_size);
	void train_end(uint32_t batch_size);
	// The same step in three calls, for sharded runs that want the sample prefetch: train_forward
	// (rays, inference, loss: counters() final for this rank), the caller sums counters() over the
	// ranks, train_backward (gradients() final for this rank), the caller sums gradients(),
	// train_end. train_begin == train_forward + train_backward.
	void train_forward(uint32_t batch_size);
	void train_backward(uint32_t batch_size, bool counters_final = true);
	uint32_t shard_rank = 0, shard_world = 1;
	// The whole sharded step inside the library (dist.cu): NCCL communicators from two unique ids, then train() runs
	// forward / counter all-reduce / backward / reduce-scatter / sharded Adam / all-gather on every rank.
	void init_data_parallel(const void* id_main, const void* id_aux, int rank, int world);
	void train_dp(uint32_t batch_size);
	bool data_parallel() const { return (bool)m_dist; }
	NerfStepCounters* counters() { return m_counters.as<NerfStepCounters>() + (shard_world > 1 ? 2u : m_counter_slot); }
	// Sample prefetch (no reference counterpart): once the loss kernel of step k has reported its
	// counts, train() marches the rays of step k+1 on a second stream, under the backward pass and
	// the optimizer of step k. The samples depend on the dataset, the occupancy bitfield, the step's
	// RNG and rays_per_batch only - never on the parameters - so the result is the serial one; a
	// prefetch whose inputs changed before it is used is dropped and redone (tests/test_nerf_gpu.py).
	bool train_prefetch = true;
	uint64_t prefetches_adopted = 0; // steps whose rays came from a prefetch
	float loss() const { return m_loss_scalar; }
	uint32_t training_step() const { return m_training_step; }

	// ---- introspection for tests / tools
	NerfNetworkModule* nerf_network() { return m_network.get(); }
	Optimizer* optimizer() { return m_optimizer.get(); }
	size_t n_params() const { return m_n_params; }
	void apply_l2_access_policy();           // testbed.cu; readable as properties l2_window_bytes / l2_window_hit_ratio
	double l2_window_bytes = 0, l2_window_hit_ratio = 0;
	float* params_fp32() { return m_params_fp32.as<float>(); }
	__half* params() { return m_params.as<__half>(); }
	__half* params_inference() { return m_optimizer && m_optimizer->custom_weights() ? m_optimizer->custom_weights() : m_params.as<__half>(); }
	__half* gradients() { return m_gradients.as<__half>(); }
	float* density_grid() { return m_density_grid.as<float>(); }
	uint8_t* density_grid_bitfield() { return m_density_bitfield.as<uint8_t>(); }
	float* density_grid_mean() { return m_density_mean.as<float>(); }
	const Aabb& aabb() const { return m_aabb; }
	uint32_t max_cascade() const { return m_max_cascade; }
	float cone_angle_constant() const { return m_cone_angle_constant; }
	// Testbed::m_nerf.cone_angle_constant (python_api.cu:640): 0 = uniform steps, as scripts/run.py --nerf_compatibility sets
	void set_cone_angle_constant(float v) { m_cone_angle_constant = v; ++m_epoch; }
	const ImageMeta* metadata_gpu() const { return m_metadata_gpu.as<ImageMeta>(); }
	uint32_t n_images() const { return (uint32_t)m_images.size(); }
	ImageMeta image_meta(uint32_t idx) const {
		if (idx >= m_images.size()) throw std::runtime_error{"Invalid frame index"};
		return m_images[idx].meta;
	}
	cudaStream_t stream() const { return m_stream; }
	void set_stream(cudaStream_t s);
	const Json& network_config() const { return m_network_config; }

	// counters of the last step (NerfCounters)
	uint32_t rays_per_batch = 1u << 12;
	uint32_t measured_batch_size = 0;
	uint32_t measured_batch_size_before_compaction = 0;
	uint32_t n_rays_total = 0;

	// knobs (testbed.h / nerf.h defaults)
	uint64_t seed = 1337;
	int rgb_activation = kActLogistic;
	int density_activation = kActExponential;
	int loss_type = kLossL2;
	int color_space = kColorLinear;
	bool random_bg_color = true;
	bool linear_colors = false;
	bool snap_to_pixel_centers = true;
	float near_distance = 0.1f;
	float density_grid_decay = 0.95f;
	float background_color[4] = {0.0f, 0.0f, 0.0f, 1.0f};
	bool train_network = true;
	bool train_encoding = true;
	float render_min_transmittance = 0.01f;
	int aabb_scale = 1;

	// Steps of update_density_grid_nerf, exposed for tests.
	void update_density_grid_nerf(float decay, uint32_t n_uniform, uint32_t n_nonuniform);
	void update_density_grid_mean_and_bitfield();
	void upload_metadata();

	// ---- snapshots (Testbed::load_snapshot, src/testbed.cu:4841-4955; Trainer::deserialize, tcnn trainer.h:290-314)
	// Trainer::set_params / set_params_full_precision from HOST memory: the inference (EMA) weights,
	// the training weights and the fp32 master copy all take the given values.
	void set_params(const void* host_params, bool full_precision, size_t n_params);
	// the density grid as the snapshot holds it (fp16, HOST) -> float grid, mean and bitfield
	void set_density_grid_fp16(const void* host_grid, size_t n_cells);
	void restore_training_state(uint32_t training_step, float loss);

	// ---- file-level entry points, host C++ as in the reference (host_io.cu, host_dataset.cu)
	// Testbed::reload_network_from_file (src/testbed.cu:274-308) with load_network_config / "parent" merge (:86-97, :246-272)
	void reload_network_from_file(const std::string& path);
	// Testbed::save_snapshot / load_snapshot (src/testbed.cu:4775-4972): msgpack, zlib stream for *.ingp
	void save_snapshot(const std::string& path, bool include_optimizer_state, bool compress);
	void load_snapshot(const std::string& path);
	void load_snapshot(const Json& config);
	Json snapshot_config(bool include_optimizer_state);
	// Testbed::load_training_data for NeRF scenes (src/testbed.cu:125-152, src/nerf_loader.cu:273-740): transforms.json
	// (or a directory of them); JPEG decoded on the GPU by nvjpeg, PNG on the host. max_images 0 = all.
	void load_training_data(const std::string& path, uint32_t max_images = 0);
	void load_image_file(const std::string& path);
	void load_file(const std::string& path);         // Testbed::load_file (src/testbed.cu:318-391)   // Testbed::load_image (src/testbed_image.cu:362-428): *.exr, *.png, *.jpg
	DatasetInfo dataset_info;
	ViewState view;

	// ---- rendering (Testbed::render_nerf via NerfTracer; see nerf_render.cu)
	// camera 3x4 row-major NGP convention, focal length in pixels of the render resolution.
	// rgba_out: DEVICE float4 [h][w] (linear, premultiplied); depth_out: DEVICE float [h][w] or null.
	// Pixels [pixel_begin, pixel_end) in row-major order are rendered (tile sharding).
	void render_nerf(uint32_t width, uint32_t height, const float* camera_3x4, float focal_x, float focal_y, float principal_x, float principal_y,
		uint32_t spp_index, uint32_t pixel_begin, uint32_t pixel_end, float* rgba_out, float* depth_out);
	uint64_t last_render_samples = 0;
	// Window policy of the wavefront: a window marches clamp(target_queries / n_alive, min, max)
	// samples per ray (reference: 2 Mi queries, 1..8 steps, src/testbed_nerf.cu:36-38,1693-1694)
	// and the host reads the live count back every `render_windows_per_sync` windows.
	// A pixel does not depend on the policy (tests/test_render_gpu.py), so the defaults are the
	// fastest measured on B200 (scripts/profile_render.py), not the reference's.
	uint32_t render_target_queries = 12u << 20; // sweep on B200 (scripts/render_shard_sweep.py): 14.7 ms per 1080p frame against 15.5 at 8M; shards of 8 unchanged
	uint32_t render_min_steps = 4;
	uint32_t render_max_steps = 32;
	uint32_t render_windows_per_sync = 4;
	bool render_skip_past_last_cascade = true;   // nerf_device.cuh kSkipMaxMip; false = the reference's climb, up to max_cascade
	// Tile sharding: with render_shard_world > 1 only the 256-pixel tiles t of the pixel range with
	// t % render_shard_world == render_shard_rank are traced (the others stay untouched).
	uint32_t render_shard_rank = 0, render_shard_world = 1;
	// Testbed::m_nerf.render_with_lens_distortion / render_lens (testbed.h; set_camera_to_training_view, src/testbed.cu:477-478):
	// the lens the render camera's rays are undistorted with (Perspective or OpenCV).
	bool render_with_lens_distortion = false;
	int render_lens_mode = kLensPerspective;
	float render_lens_params[7] = {};
	void set_render_lens(int lens_mode, const float* params);
	bool render_snap_to_pixel_centers = false; // Testbed::m_snap_to_pixel_centers (testbed.h:995)
	// ---- headless render buffer (CudaRenderBuffer accumulate + tonemap, render_buffer.cu): `spp`
	// traced frames averaged, background blended in, exposure, tonemap curve, sRGB unless `linear`.
	// What pyngp's Testbed.render(width, height, spp, linear) returns, in a DEVICE float4 [h][w].
	void render(uint32_t width, uint32_t height, uint32_t spp, bool linear, const float* camera_3x4, float focal_x, float focal_y,
		float principal_x, float principal_y, float* rgba_out);
	// Ground-truth overlay (CudaRenderBuffer::overlay_image, render_buffer.cu:348-411, 700-733): the
	// training image of `view` blended with weight `alpha` over the DEVICE float4 [h][w] buffer.
	void overlay_ground_truth(uint32_t width, uint32_t height, bool linear, uint32_t view, int fov_axis, float zoom, float alpha, float* rgba_inout);
	float exposure = 0.0f;              // Testbed::m_exposure (testbed.h:735)
	int tonemap_curve = 0;              // ETonemapCurve: Identity, ACES, Hable, Reinhard (common.h:112-117)
	int render_color_space = kColorLinear; // CudaRenderBuffer::m_color_space: the space frames are averaged in
	// Optional DEVICE buffer [height * width] receiving the number of samples composited per pixel.
	void set_render_steps_out(uint32_t* p) { m_render_steps_out = p; }

private:
	enum : int { kModeNerf = 0, kModeImage = 1 };
	int m_mode = kModeNerf;
	void reset_network_image();
	void train_image(uint32_t batch_size, bool get_loss_scalar);
	std::unique_ptr<NetworkWithInputEncodingModule> m_image_network;
	DeviceBuffer m_image_data, m_image_positions, m_image_targets, m_image_output, m_image_dloss, m_image_loss_values;
	int m_image_type = kImageFloat, m_image_res_x = 0, m_image_res_y = 0;

	void training_prep_nerf();
	bool m_get_loss_scalar = false;
	void train_nerf_forward(uint32_t target_batch_size);
	NerfTrainConfig make_train_config(uint32_t target_batch_size) const;
	NerfTrainBuffers make_train_buffers(uint32_t counter_slot);
	void reserve_step_buffers(uint32_t target_batch_size);
	void control_loop(uint32_t target_batch_size);
	void launch_prefetch(uint32_t target_batch_size);
	void join_prefetch();   // orders m_stream after an in-flight prefetch
	void touch_generator_inputs() { ++m_epoch; }
	struct Prefetch {
		bool pending = false;
		NerfTrainConfig cfg{};
		Pcg32 rng{};
		uint64_t epoch = 0;
	} m_prefetch;
	uint64_t m_epoch = 0;          // bumped whenever something the ray generator reads changes
	uint32_t m_counter_slot = 0;   // m_counters holds three structs: this step's, the prefetched step's, the sharded sum
	bool m_counters_read = false;  // the step's counters are already on their way to the host
	cudaStream_t m_side_stream = nullptr;
	cudaEvent_t m_counters_ready = nullptr, m_prefetch_done = nullptr, m_mlp_done = nullptr, m_encoder_done = nullptr;

	std::unique_ptr<DistContext> m_dist;
	cudaStream_t m_stream = nullptr;
	cudaStream_t m_own_stream = nullptr;
	Json m_network_config;
	std::unique_ptr<NerfNetworkModule> m_network;
	std::unique_ptr<Optimizer> m_optimizer;
	size_t m_n_params = 0;
	DeviceBuffer m_params_fp32, m_params, m_gradients;
	DeviceBuffer m_compact_src;   // [batch] marched row each compacted sample was copied from (feature-tile reuse, network.h)

	std::vector<NerfImage> m_images;
	DeviceBuffer m_metadata_gpu;
	bool m_metadata_dirty = true;
	bool m_is_hdr = false;

	Aabb m_aabb{{0, 0, 0}, {1, 1, 1}};
	uint32_t m_max_cascade = 0;
	float m_cone_angle_constant = 0.0f;

	Pcg32 m_rng;
	Pcg32 m_density_grid_rng;
	uint32_t m_density_grid_ema_step = 0;
	uint32_t m_training_step = 0;
	float m_loss_scalar = 0.0f;

	DeviceBuffer m_density_grid, m_density_grid_tmp, m_density_bitfield, m_density_mean;
	DeviceBuffer m_grid_positions, m_grid_indices, m_grid_mlp_out;

	// per-step scratch
	DeviceBuffer m_counters, m_ray_indices, m_rays, m_numsteps, m_coords, m_rgbsigma, m_coords_compacted, m_dloss, m_loss;
	DeviceBuffer m_render_scratch, m_frame_buffer, m_accumulate_buffer;
	NerfStepCounters* m_counters_host = nullptr; // pinned
	float* m_loss_host = nullptr;                // pinned
	void* m_counters_host_render = nullptr;      // pinned
	uint32_t* m_render_steps_out = nullptr;
};

} // namespace ngpb
