/* ngp_b200.h - C ABI of libngp_b200.so (B200 / sm_100a implementation of instant-ngp's
 * hash-grid + fused-MLP + NeRF ray-march hot path).
 *
 * Every entry point takes plain pointers and sizes: device pointers are raw CUDA device
 * addresses owned by the caller, `stream` is a cudaStream_t passed as void*. Functions
 * return NGPB_OK (0) or NGPB_ERROR (1); the message of the last failure on the calling
 * thread is available from ngpb_last_error(). The texts match the std::runtime_error
 * messages of the reference where a corresponding check exists.
 *
 * Each block below names the reference interface it replaces
 * (paths relative to the reference root; tcnn/ = dependencies/tiny-cuda-nn/).
 */
#ifndef NGP_B200_H
#define NGP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NGPB_OK 0
#define NGPB_ERROR 1

#define NGPB_PRECISION_FP32 0
#define NGPB_PRECISION_FP16 1

typedef struct ngpb_module ngpb_module_t;   /* tcnn::cpp::Module, tcnn/include/tiny-cuda-nn/cpp_api.h:86-115 */
typedef struct ngpb_context ngpb_context_t; /* tcnn::cpp::Context, cpp_api.h:82-84 */

/* ---- library ------------------------------------------------------------------------- */
/* Text of the last failure on the calling thread: what the reference throws as std::runtime_error from CHECK_THROW /
 * CUDA_CHECK_THROW (tcnn common_host.h) arrives here, behind the int status. No reference counterpart as a function. */
const char* ngpb_last_error(void);
/* Library version (no reference counterpart; NGP_VERSION of src/main.cu:186 is a build string). */
int ngpb_version(void);
/* 1 when a CUDA device is usable from this process, else 0 (never an error). */
int ngpb_has_device(void);
/* tcnn::cpp::batch_size_granularity(), cpp_api.h:62; tcnn common.h:235 (256). */
uint32_t ngpb_batch_size_granularity(void);
/* tcnn::cpp::default_loss_scale(), cpp_api.h:75; tcnn common.h:232 (128 for fp16). */
float ngpb_default_loss_scale(int precision);

/* ---- phase timer (no reference counterpart; the reference keeps wall-clock EMAs, testbed.h:936-940) ----
 * When enabled, the library brackets its kernel groups with CUDA events on the calling thread.
 * ngpb_profile_collect writes "name=total_ms:count;..." (call after synchronising the stream). */
void ngpb_profile_enable(int on);
size_t ngpb_profile_collect(char* buf, size_t cap, int reset);

/* ---- module factories -----------------------------------------------------------------
 * tcnn::cpp::create_encoding / create_network_with_input_encoding / create_network,
 * cpp_api.h:117-119, implemented at tcnn/src/cpp_api.cu:140-165. JSON strings use the tcnn
 * schema (tcnn/src/encoding.cu:54-119, tcnn/src/network.cu:48-136, grid.h:1143-1208). */
int ngpb_create_encoding(uint32_t n_input_dims, const char* encoding_json, int precision, ngpb_module_t** out);
int ngpb_create_network_with_input_encoding(uint32_t n_input_dims, uint32_t n_output_dims,
	const char* encoding_json, const char* network_json, ngpb_module_t** out);
int ngpb_create_network(uint32_t n_input_dims, uint32_t n_output_dims, const char* network_json, ngpb_module_t** out);
/* ngp::NerfNetwork<T> ctor, include/neural-graphics-primitives/nerf_network.h:81-101. */
int ngpb_create_nerf_network(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
	const char* pos_encoding_json, const char* dir_encoding_json,
	const char* density_network_json, const char* rgb_network_json, ngpb_module_t** out);
void ngpb_module_free(ngpb_module_t* m);

/* ---- module introspection (cpp_api.h:98-111) ------------------------------------------ */
uint32_t ngpb_module_n_input_dims(const ngpb_module_t* m);
uint32_t ngpb_module_n_output_dims(const ngpb_module_t* m);        /* padded width, as cpp_api.cu:132 */
uint32_t ngpb_module_n_output_dims_unpadded(const ngpb_module_t* m);
size_t ngpb_module_n_params(const ngpb_module_t* m);
int ngpb_module_param_precision(const ngpb_module_t* m);
int ngpb_module_output_precision(const ngpb_module_t* m);
/* Writes the hyperparams JSON (Module::hyperparams) into buf; returns needed size incl. NUL. */
size_t ngpb_module_hyperparams(const ngpb_module_t* m, char* buf, size_t cap);
/* Module::name. */
size_t ngpb_module_name(const ngpb_module_t* m, char* buf, size_t cap);
/* ParametricObject::layer_sizes, tcnn object.h:102: writes up to cap (rows, cols) pairs;
 * returns the number of layers. Grid parameters are not a layer (as in the reference). */
size_t ngpb_module_layer_sizes(const ngpb_module_t* m, uint32_t* rows_cols, size_t cap);

/* Module::initialize_params(seed, float* params_full_precision, scale), cpp_api.cu:134-137.
 * params_fp32 is a DEVICE pointer to n_params floats. Same seed gives the same values as the
 * reference's pcg32-driven initialisers (grid: grid.h:1059-1062; MLP: fully_fused_mlp.cu:866-891). */
int ngpb_module_initialize_params(ngpb_module_t* m, void* stream, uint64_t seed, float* params_fp32, float scale);

/* ---- module execution (cpp_api.h:91-94) -----------------------------------------------
 * input : fp32 [n][n_input_dims]   (sample-major; tcnn ColumnMajor dims x n)
 * output: fp16 [n][n_output_dims]  (padded width)
 * params: fp16 [n_params] in the module's param precision
 * n must be a multiple of ngpb_batch_size_granularity() for networks (tcnn object.h:130). */
int ngpb_module_inference(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output, const void* params);
int ngpb_module_forward(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output,
	const void* params, int prepare_input_gradients, ngpb_context_t** ctx);
/* dL_dinput may be NULL; dL_dparams may be NULL (GradientMode::Ignore) else it is overwritten. */
int ngpb_module_backward(ngpb_module_t* m, void* stream, const ngpb_context_t* ctx, uint32_t n,
	float* dL_dinput, const void* dL_doutput, void* dL_dparams,
	const float* input, const void* output, const void* params);
void ngpb_context_free(ngpb_context_t* ctx);

/* ngpb_module_inference without any allocation, for callers that cannot call cudaMalloc (CUDA graph capture,
 * tight serving loops). NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels, same result as ngpb_module_inference for a module made by
 * ngpb_create_network_with_input_encoding or ngpb_create_network; the encoder's feature tiles live in a
 * file-scope device array of NGPB_STATIC_SCRATCH_BYTES (held by every GPU that loads the library, until
 * exit) instead of a per-call stream-ordered allocation. Limit: ceil(n / 128) * 128 * padded_encoding_width * 2 bytes must fit
 * (n <= 65536 samples at 32 encoded features); larger batches are rejected with NGPB_ERROR and a message.
 * All calls on one device share the array: calls that overlap on the GPU (two streams) race. */
#define NGPB_STATIC_SCRATCH_BYTES (4u << 20)
size_t ngpb_static_scratch_bytes(void);
int ngpb_module_inference_static_scratch(ngpb_module_t* m, void* stream, uint32_t n, const float* input, void* output, const void* params);

/* ---- grid introspection for parity tests ----------------------------------------------
 * GridOffsetTable (grid_interface.h:84-88) and per-level scale/resolution
 * (common_device.h:709-718). Arrays must hold n_levels+1 / n_levels / n_levels items.
 * Returns the number of levels, or 0 if the module is not a grid encoding. */
uint32_t ngpb_grid_level_table(const ngpb_module_t* m, uint32_t* offsets, float* scales, uint32_t* resolutions);
/* idx[(i*L + level)*2^D + corner] = grid_index() of each corner (common_device.h:690-707). */
int ngpb_grid_indices(const ngpb_module_t* m, void* stream, uint32_t n, const float* input, uint32_t* idx);
/* Encoding::set_max_level analogue (grid_interface.h:60-66). */
int ngpb_grid_set_max_level(ngpb_module_t* m, float max_level);

/* ---- random fills (tcnn random.h:39-87) -------------------------------------------------
 * Fills a device array with pcg32 uniforms exactly as generate_random_uniform does, and
 * advances the caller's generator state (state[0] = state, state[1] = inc). */
int ngpb_random_uniform(void* stream, uint64_t* rng_state, size_t n, float* out, float lower, float upper);


/* ---- optimizer ------------------------------------------------------------------------------
 * tcnn::create_optimizer<T>(json) (tcnn/src/optimizer.cu:48-82) for the chain the shipped configs
 * use, [Ema o] [ExponentialDecay o] Adam, and tcnn::Optimizer<T> (tcnn optimizer.h):
 * allocate / step / learning_rate / step / custom_weights / update_hyperparams / hyperparams. */
typedef struct ngpb_optimizer ngpb_optimizer_t;
int ngpb_optimizer_create(const char* optimizer_json, ngpb_optimizer_t** out);
void ngpb_optimizer_free(ngpb_optimizer_t* o);
/* layer_rows_cols: n_layers (rows, cols) pairs = ParametricObject::layer_sizes of the model. */
int ngpb_optimizer_allocate(ngpb_optimizer_t* o, uint32_t n_weights, const uint32_t* layer_rows_cols, size_t n_layers);
/* weights_fp32 / weights (fp16) / gradients (fp16): device pointers to n_weights elements. */
int ngpb_optimizer_step(ngpb_optimizer_t* o, void* stream, float loss_scale, float* weights_fp32, void* weights, const void* gradients);
float ngpb_optimizer_learning_rate(const ngpb_optimizer_t* o);
int ngpb_optimizer_set_learning_rate(ngpb_optimizer_t* o, float lr);
uint32_t ngpb_optimizer_step_count(const ngpb_optimizer_t* o);
/* EMA weights (device, fp16) or NULL when the chain has no Ema stage. */
void* ngpb_optimizer_custom_weights(const ngpb_optimizer_t* o);
int ngpb_optimizer_update_hyperparams(ngpb_optimizer_t* o, const char* optimizer_json);
size_t ngpb_optimizer_hyperparams(const ngpb_optimizer_t* o, char* buf, size_t cap);
/* Device pointers to the Adam state (first moments, second moments: float; steps: uint32). */
int ngpb_optimizer_state(const ngpb_optimizer_t* o, float** first_moments, float** second_moments, uint32_t** param_steps);

/* ---- NerfNetwork extras (nerf_network.h:270-279, 356-371) --------------------------------------- */
/* density-MLP output 0 per sample: input fp32 rows of in_stride floats (position first), output
 * fp16 [n][out_stride] with out_stride 1 or 16; n multiple of 128. */
int ngpb_nerf_network_density(ngpb_module_t* m, void* stream, uint32_t n, const float* input, uint32_t in_stride, void* output, uint32_t out_stride, const void* params);
/* offsets[4] (in parameters) of density MLP, rgb MLP, pos encoding, dir encoding in the flat buffer. */
int ngpb_nerf_network_param_layout(const ngpb_module_t* m, size_t* offsets);

/* ---- NeRF kernels (src/testbed_nerf.cu) ---------------------------------------------------------
 * POD views of the buffers the reference passes to generate_training_samples_nerf (:679-838) and
 * compute_loss_kernel_train_nerf (:841-1160). All pointers are device pointers. */
typedef struct {
	const void* pixels;          /* TrainingImageMetadata::pixels */
	int image_data_type;         /* EImageDataType: 1 Byte (sRGB8, straight alpha), 2 Half, 3 Float (linear, premultiplied) */
	int lens_mode;               /* ELensMode: 0 Perspective, 1 OpenCV */
	float lens_params[7];
	int res_x, res_y;
	float principal_x, principal_y;
	float focal_x, focal_y;
	float xform[12];             /* camera-to-world, column-major 4 columns x 3 (tcnn mat4x3) */
} ngpb_nerf_image_t;

typedef struct {
	uint32_t numsteps, numsteps_compacted, rays;
	float loss_sum;
} ngpb_nerf_counters_t;

typedef struct {
	float aabb_min[3], aabb_max[3];
	uint32_t n_rays, n_rays_total, max_samples, max_samples_compacted, n_training_images, max_mip;
	float cone_angle_constant, near_distance, loss_scale;
	int snap_to_pixel_centers, random_bg_color, linear_colors, color_space, loss_type, rgb_activation, density_activation;
	float background_color[3];
	/* Ray sharding for data-parallel training (no reference counterpart: the reference trains on one
	 * GPU, README.md:250-252): the call handles n_rays_local rays from global index ray_begin of the
	 * n_rays-ray batch; n_rays_local == 0 means the whole batch. */
	uint32_t ray_begin, n_rays_local;
} ngpb_nerf_train_config_t;

typedef struct {
	const ngpb_nerf_image_t* metadata;
	const uint8_t* density_bitfield;
	const float* mean_density;
	ngpb_nerf_counters_t* counters;
	uint32_t* ray_indices;
	float* rays_unnormalized;    /* [n_rays][6] */
	uint32_t* numsteps;          /* [n_rays][2] */
	float* coords;               /* [max_samples][7] NerfCoordinate */
	const void* rgbsigma;        /* fp16 [max_samples][4] */
	float* coords_compacted;     /* [max_samples_compacted][7] */
	void* dloss_doutput;         /* fp16 [max_samples_compacted][4] */
	float* loss;                 /* [n_rays] */
	uint32_t* compact_src;       /* [max_samples_compacted] or NULL: row of `coords` each compacted sample was copied from (no reference counterpart) */
} ngpb_nerf_train_buffers_t;

/* generate_training_samples_nerf (src/testbed_nerf.cu:679-838) and compute_loss_kernel_train_nerf with its compaction
 * (:841-1160), with the buffers Testbed::train_nerf_step hands them (:2730-2870).
 * rng_state = {state, inc} of the step's pcg32 (m_rng). Counters must be zeroed by the caller. */
int ngpb_nerf_generate_training_samples(void* stream, const ngpb_nerf_train_config_t* cfg, const ngpb_nerf_train_buffers_t* buf, const uint64_t* rng_state);
int ngpb_nerf_compute_loss_and_compact(void* stream, const ngpb_nerf_train_config_t* cfg, const ngpb_nerf_train_buffers_t* buf, const uint64_t* rng_state);
/* update_density_grid_mean_and_bitfield (:2362-2379): grid float [(max_cascade+1) * 128^3] in Morton
 * order, bitfield uint8 [8 * 128^3 / 8], mean_density 1 float. */
int ngpb_nerf_update_bitfield(void* stream, uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density);
size_t ngpb_nerf_bitfield_bytes(void);
/* The kernels of update_density_grid_nerf (src/testbed_nerf.cu:2271-2360), one call each:
 * mark_untrained_density_grid (:74-147; cells no camera sees become -1), in place on grid [n_elements];
 * generate_grid_samples_nerf_nonuniform (:149-184): n cells chosen by the (index, step) hash among those
 *   above `thresh`, a random position in each -> positions float [n][4] (warped xyz, warped dt), indices [n];
 *   aabb6 = min xyz, max xyz; rng_state = {state, inc} of m_density_grid_rng;
 * splat_grid_samples_nerf_max_nearest_neighbor (:186-204): grid_tmp[indices[i]] = max(., density * min step),
 *   density_raw = fp16 density-network output per sample;
 * ema_grid_samples_nerf (:264-282): grid = grid < 0 ? grid : max(grid * decay, grid_tmp). */
int ngpb_nerf_mark_untrained_density_grid(void* stream, uint32_t n_elements, float* grid, uint32_t n_images, const ngpb_nerf_image_t* metadata, int clear_visible_voxels);
int ngpb_nerf_generate_grid_samples(void* stream, uint32_t n, const uint64_t* rng_state, uint32_t step, const float* aabb6, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh);
int ngpb_nerf_splat_grid_samples(void* stream, uint32_t n, const uint32_t* indices, const void* density_raw, float* grid_tmp, int density_activation);
int ngpb_nerf_ema_grid_samples(void* stream, uint32_t n_elements, float decay, float* grid, const float* grid_tmp);

/* ---- Testbed (include/neural-graphics-primitives/testbed.h, src/python_api.cu:383-567) -------------
 * NeRF mode of ngp::Testbed: dataset set-up as pyngp's create_empty_nerf_dataset / set_image /
 * set_camera_extrinsics+intrinsics, reload_network_from_json, train(batch), render. */
typedef struct ngpb_testbed ngpb_testbed_t;
int ngpb_testbed_create(ngpb_testbed_t** out);
void ngpb_testbed_free(ngpb_testbed_t* t);
int ngpb_testbed_create_empty_nerf_dataset(ngpb_testbed_t* t, uint32_t n_images, int aabb_scale, int is_hdr);
/* pixels: host or device pointer, w*h RGBA texels of image_data_type. */
int ngpb_testbed_set_nerf_image(ngpb_testbed_t* t, uint32_t idx, const void* pixels, int image_data_type, int width, int height);
/* xform_3x4: camera-to-world in NGP convention, row-major [3][4]. */
int ngpb_testbed_set_nerf_camera(ngpb_testbed_t* t, uint32_t idx, const float* xform_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, int lens_mode, const float* lens_params);
/* The stream every call of this testbed runs on (default: an internal non-blocking stream). */
int ngpb_testbed_set_stream(ngpb_testbed_t* t, void* stream);
int ngpb_testbed_reload_network_from_json(ngpb_testbed_t* t, const char* config_json);
int ngpb_testbed_reset_network(ngpb_testbed_t* t, int clear_density_grid);

/* ---- file-level entry points, host C++ as in the reference (csrc/host_io.cu, host_snapshot.cu, host_dataset.cu) ----
 * Testbed::reload_network_from_file (src/testbed.cu:274-308): *.json with comments and "parent" inheritance
 * (load_network_config + merge_parent_network_config, :86-97, :246-272), or a snapshot file whose network config is used. */
int ngpb_testbed_reload_network_from_file(ngpb_testbed_t* t, const char* path);
/* Testbed::load_training_data for NeRF scenes (src/testbed.cu:125-152 -> load_nerf, src/nerf_loader.cu:273-740): a
 * transforms.json or a directory of them. JPEG frames are decoded on the GPU (nvjpeg), PNG and EXR frames on the host;
 * Perspective and OpenCV lenses; alpha files, dynamic masks, white / black_transparent. max_images 0 = all.
 * An image file (*.exr, *.png, *.jpg) instead selects the neural-image mode (Testbed::load_image, src/testbed_image.cu:362-428). */
int ngpb_testbed_load_training_data(ngpb_testbed_t* t, const char* path, uint32_t max_images);
/* Testbed::load_file (src/testbed.cu:318-391): *.ingp / *.msgpack -> snapshot, a network config *.json -> network, anything else -> training data. */
int ngpb_testbed_load_file(ngpb_testbed_t* t, const char* path);
/* Testbed::save_snapshot(path, include_optimizer_state, compress) / load_snapshot(path), src/testbed.cu:4775-4972:
 * json::to_msgpack of the network config with a "snapshot" object (json_binding.h key names), zlib stream for *.ingp.
 * Files interchange with the reference's for the shipped NeRF configs. */
int ngpb_testbed_save_snapshot(ngpb_testbed_t* t, const char* path, int include_optimizer_state, int compress);
int ngpb_testbed_load_snapshot(ngpb_testbed_t* t, const char* path);
/* TrainingImageMetadata of one view as the testbed holds it (python_api.cu:578-590); `pixels` is the device pointer. */
int ngpb_testbed_image_metadata(const ngpb_testbed_t* t, uint32_t idx, ngpb_nerf_image_t* out);
/* The interactive camera a snapshot carries (m_camera 3x4 rows [0..12), relative_focal_length [12..14), screen_center
 * [14..16), zoom [16], up_dir [17..20), sun_dir [20..23), fov_axis [23]); set != 0 writes, else reads. */
int ngpb_testbed_view_state(ngpb_testbed_t* t, float* values24, int set);
/* NerfDataset's scalar members as JSON (scale, offset, up, paths, from_mitsuba, is_hdr, render_aabb); returns the size needed. */
size_t ngpb_testbed_dataset_info(const ngpb_testbed_t* t, char* buf, size_t cap);
/* The same record written (keys that are present): for hosts that fill the dataset through set_nerf_image / set_nerf_camera. */
int ngpb_testbed_set_dataset_info(ngpb_testbed_t* t, const char* info_json);
/* Host-only helpers of the same layer (no device needed). load_network_config as JSON text (binaries print as nlohmann's
 * {"bytes": [...]}); returns the size needed incl. NUL, 0 on failure (ngpb_last_error). */
size_t ngpb_host_load_network_config(const char* path, char* buf, size_t cap);
/* msgpack -> value tree -> msgpack with nlohmann's encoding choices; returns the size needed, 0 on failure. */
size_t ngpb_host_msgpack_reencode(const void* in, size_t n, void* out, size_t cap);
/* The loader's PNG decoder: 8-bit RGBA rows top to bottom into rgba (width * height * 4 bytes when cap allows). */
/* What ngpb_testbed_load_training_data decides before it touches a pixel, as JSON: frame order after natural sort and
 * sharpness culling, resolved image / alpha / mask paths, NGP camera matrices, lens, scale / offset / up / aabb_scale, flags.
 * Returns the size needed incl. NUL, 0 on failure. */
size_t ngpb_host_describe_dataset(const char* path, char* buf, size_t cap);
/* The frame order of the loader: SI::natural::compare of dependencies/NaturalSort as src/nerf_loader.cu:347-349 uses it. 1 when a sorts before b. */
int ngpb_host_natural_less(const char* a, const char* b);
int ngpb_host_decode_png(const char* path, void* rgba, size_t cap, int* width, int* height);
/* The loader's host decoders for 8-bit files, chosen from the file's content as stb_image does (load_stbi,
 * src/common_host.cu:214-217): JPEG (baseline / progressive, with stb_image's IDCT / upsampling / colour arithmetic), PNG, BMP,
 * GIF (first frame), PSD (composite), PIC, TGA, binary PNM -> RGBA8 rows top to bottom, the texels stbi_load(..., 4) returns.
 * (ngpb_testbed_load_training_data decodes JPEG on the device with nvjpeg unless NGPB_JPEG_DECODER=host.) */
int ngpb_host_decode_image(const char* path, void* rgba, size_t cap, int* width, int* height);
/* The loader's OpenEXR reader (single-part scanline or tiled, HALF / FLOAT, NONE / RLE / ZIPS / ZIP / PIZ): float RGBA rows top to bottom. */
int ngpb_host_decode_exr(const char* path, float* rgba, size_t cap_floats, int* width, int* height);
/* The same reader with the channel rule of NeRF frames (load_exr_to_gpu, src/tinyexr_wrapper.cu:153-258): R / G / B / A by name,
 * absent colours 0, absent alpha 1; ngpb_host_decode_exr follows the image mode's LoadEXR (a one-channel file fills all four). */
int ngpb_host_decode_exr_frame(const char* path, float* rgba, size_t cap_floats, int* width, int* height);
/* Radiance HDR (RGBE) as float RGBA, alpha 1: stbi_loadf(..., 4), the image mode's reader for such files (load_stbi_gpu,
 * src/common_host.cu:268-299). ngpb_host_decode_image gives the same file as bytes the way stbi_load does (gamma 1 / 2.2). */
int ngpb_host_decode_hdr(const char* path, float* rgba, size_t cap_floats, int* width, int* height);
/* Testbed::train(batch_size), src/testbed.cu:4020-4109. Returns once the step's counters (loss,
 * measured batch sizes) are on the host; the backward pass and the optimizer may still be running on
 * the testbed's stream - later calls on this testbed are ordered behind them. */
int ngpb_testbed_train(ngpb_testbed_t* t, uint32_t batch_size);
/* Data-parallel form of ngpb_testbed_train (no reference counterpart; the reference trains on one GPU):
 * with properties shard_rank / shard_world set, train_begin computes this rank's share of the step
 * (rays [rank R/P, (rank+1) R/P) of the step's R rays) and leaves gradients in "gradients" and the
 * step counters in "counters" (ngpb_nerf_counters_t); the caller all-reduces (sum) both over the
 * ranks on the testbed's stream; train_end applies the optimizer and the ray-count control loop.
 * batch_size is the global batch in both calls. ngpb_testbed_train == begin + end for one rank. */
int ngpb_testbed_train_begin(ngpb_testbed_t* t, uint32_t batch_size);
int ngpb_testbed_train_end(ngpb_testbed_t* t, uint32_t batch_size);
/* The same step in three calls, which lets a sharded run keep the sample prefetch (train_prefetch):
 * train_forward (rays, inference, loss) leaves this rank's counters in "counters"; the caller sums
 * them over the ranks; train_backward starts their copy to the host and computes this rank's
 * "gradients"; the caller sums those; train_end steps the optimizer and, with the global counts
 * already on the host, marches the next step's rays under it. train_begin = forward + backward. */
/* Data-parallel training inside the library (csrc/dist.cu): one process per GPU, NCCL bound at run time from the
 * process's libnccl.so.2. Rank 0 obtains two unique ids (128 bytes each) with ngpb_nccl_get_unique_id and hands them to
 * every rank over the caller's bootstrap (torch.distributed store, MPI, a file); every rank then calls
 * ngpb_testbed_init_data_parallel with the same ids, and ngpb_testbed_train(global batch) runs the sharded step: rays
 * [rank R/P, (rank+1) R/P), the step counters all-reduced beside the backward pass, gradients reduce-scattered, Adam +
 * EMA on this rank's 1/P of the parameters, fp16 weights all-gathered. world == 1 returns to single-GPU training. */
int ngpb_nccl_get_unique_id(void* out128);
int ngpb_testbed_init_data_parallel(ngpb_testbed_t* t, const void* id_main, const void* id_aux, int rank, int world);
int ngpb_testbed_train_forward(ngpb_testbed_t* t, uint32_t batch_size);
int ngpb_testbed_train_backward(ngpb_testbed_t* t, uint32_t batch_size);
/* Named scalar properties (the public members pyngp binds, src/python_api.cu:481-566, 645-697):
 * training_step, loss, rays_per_batch, measured_batch_size, measured_batch_size_before_compaction,
 * n_params, max_cascade, cone_angle_constant, aabb_scale, seed, random_bg_color, linear_colors,
 * snap_to_pixel_centers, near_distance, density_grid_decay, render_min_transmittance, loss_type,
 * rgb_activation, density_activation, color_space, train_network, train_encoding, last_render_samples,
 * render_target_queries, render_min_steps, render_max_steps (the window policy of NerfTracer::trace,
 * src/testbed_nerf.cu:1693-1694), render_windows_per_sync, render_snap_to_pixel_centers,
 * image_snap_to_pixel_centers, shard_rank, shard_world (training ray shards), render_shard_rank, render_shard_world (render_nerf
 * then traces only the 256-pixel tiles t of [pixel_begin, pixel_end) with t % world == rank),
 * train_prefetch (default 1: ngpb_testbed_train marches the rays of the next step on a second stream
 * while the backward pass and the optimizer of this step run; same samples as the serial order),
 * prefetches_adopted (read-only: steps whose rays came from such a prefetch), exposure, tonemap_curve,
 * render_color_space (0 Linear, 1 SRGB: the space samples are averaged in), background_color_r/g/b/a. */
int ngpb_testbed_get(const ngpb_testbed_t* t, const char* name, double* value);
int ngpb_testbed_set(ngpb_testbed_t* t, const char* name, double value);
/* Device buffers by name: params (fp16), params_fp32, params_inference (fp16), gradients (fp16),
 * density_grid (float), density_grid_bitfield (uint8), density_grid_mean (float), metadata,
 * counters (ngpb_nerf_counters_t of the last step); image mode: image_positions (float [n][2]),
 * image_targets (float [n][3]), image_output (fp16 [n][16]), image_dloss (fp16 [n][16]) of the last step. */
int ngpb_testbed_device_ptr(ngpb_testbed_t* t, const char* name, void** ptr, size_t* bytes);
size_t ngpb_testbed_network_config(const ngpb_testbed_t* t, char* buf, size_t cap);
/* update_density_grid_nerf(decay, n_uniform, n_nonuniform), src/testbed_nerf.cu:2271-2360. */
int ngpb_testbed_update_density_grid(ngpb_testbed_t* t, float decay, uint32_t n_uniform, uint32_t n_nonuniform);
/* Snapshot restore, the device side of Testbed::load_snapshot (src/testbed.cu:4841-4955) after the host
 * has read the msgpack container: Trainer::set_params / set_params_full_precision (tcnn trainer.h:242-269)
 * from HOST memory (fp16 "__half" or fp32 "float" params_binary, n_params must match the network);
 * density_grid_binary (fp16, HOST; 128^3 x cascades cells, or 0 cells for a model that never trained)
 * -> float grid + update_density_grid_mean_and_bitfield; training_step and loss (:4931-4932).
 * Optimizer state travels through ngpb_testbed_device_ptr ("optimizer_first_moments",
 * "optimizer_second_moments", "optimizer_param_steps") and the properties "optimizer_step",
 * "optimizer_base_learning_rate", "optimizer_learning_rate_factor". */
int ngpb_testbed_set_params(ngpb_testbed_t* t, const void* host_params, int full_precision, size_t n_params);
int ngpb_testbed_set_density_grid_fp16(ngpb_testbed_t* t, const void* host_grid, size_t n_cells);
int ngpb_testbed_restore_training_state(ngpb_testbed_t* t, uint32_t training_step, float loss);
/* Testbed::render_nerf (src/testbed_nerf.cu:1827-1987) into caller-owned DEVICE buffers: rgba
 * float4 [h][w] (linear, premultiplied alpha), depth float [h][w] or NULL. Only pixels
 * [pixel_begin, pixel_end) (row-major) are traced: the unit of tile sharding across GPUs. */
int ngpb_testbed_render_nerf(ngpb_testbed_t* t, uint32_t width, uint32_t height, const float* camera_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, uint32_t spp_index, uint32_t pixel_begin, uint32_t pixel_end, float* rgba, float* depth);

/* Testbed::render_to_cpu (src/python_api.cu:124-221) minus camera paths and motion blur, with the
 * headless CudaRenderBuffer epilogue (src/render_buffer.cu:232-346, 533-565, 635-697): `spp` frames
 * traced with sample indices 0..spp-1 are averaged, the background colour (properties
 * background_color_r/g/b/a, given in sRGB) is blended behind the premultiplied result, then
 * 2^exposure, the tonemap curve (property tonemap_curve: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard) and
 * the sRGB transfer unless `linear`. rgba: DEVICE float4 [h][w]. */
int ngpb_testbed_render(ngpb_testbed_t* t, uint32_t width, uint32_t height, uint32_t spp, int linear, const float* camera_3x4, float focal_x, float focal_y,
	float principal_x, float principal_y, float* rgba);

/* The ground-truth frame pyngp's render() returns while render_ground_truth is set
 * (Testbed::render_frame_epilogue, src/testbed.cu:4592-4608 -> CudaRenderBuffer::overlay_image,
 * src/render_buffer.cu:348-411, 700-733): the training image of `view` resampled (nearest texel) to
 * width x height along fov_axis (0 = x, 1 = y) at `zoom`, background / exposure / tonemap curve /
 * transfer as in ngpb_testbed_render, blended with weight `alpha` over rgba (DEVICE float4 [h][w],
 * read and written; with alpha = 1 its old content does not matter, as in the reference, which
 * skips the trace, src/testbed.cu:4383). */
/* Testbed::m_nerf.render_lens (set by set_camera_to_training_view, src/testbed.cu:477-478): lens the
 * render camera's rays are undistorted with while property render_with_lens_distortion is non-zero.
 * lens_mode 0 Perspective, 1 OpenCV (k1, k2, p1, p2 in lens_params[0..3]; 7 floats or NULL). */
int ngpb_testbed_set_render_lens(ngpb_testbed_t* t, int lens_mode, const float* lens_params);

/* Testbed::render_ground_truth as scripts/run.py uses it for PSNR (src/testbed.cu:4592-4608): the training image of `view`
 * drawn over the frame by overlay_image_kernel (src/render_buffer.cu:348-411). rgba: device float [height][width][4]. */
int ngpb_testbed_overlay_ground_truth(ngpb_testbed_t* t, uint32_t width, uint32_t height, int linear, uint32_t view, int fov_axis, float zoom, float alpha, float* rgba);

/* Neural-image mode (ETestbedMode::Image; Testbed::load_image, train_image, render_image,
 * src/testbed_image.cu:212-340, 360-410): the image is RGBA Half (2) or Float (3), linear colours,
 * host or device pointer. After this call reload_network_from_json builds the 2-D
 * NetworkWithInputEncoding (per_level_scale from max(w, h) / 2, src/testbed.cu:3701-3713),
 * ngpb_testbed_train runs train_image (stratified positions, snapped / bilinear targets, the
 * config's loss: L2, RelativeL2 or L1) and render_image evaluates the network at the pixel centres
 * of a width x height grid into a DEVICE float4 buffer (linear colours, alpha 1). */
int ngpb_testbed_set_image(ngpb_testbed_t* t, const void* pixels, int image_data_type, int width, int height);
int ngpb_testbed_render_image(ngpb_testbed_t* t, uint32_t width, uint32_t height, float* rgba);

/* Optional DEVICE buffer [height*width] that render_nerf fills with the number of samples composited
 * per pixel (NerfPayload::n_steps, the value ERenderMode::Cost shows, src/testbed_nerf.cu:1329-1331);
 * NULL switches it off. Pixels whose ray never enters occupied space are not written. */
int ngpb_testbed_set_render_steps_out(ngpb_testbed_t* t, uint32_t* steps);

/* ---- roofline probes (no reference counterpart) -------------------------------------------------
 * Random-access micro-kernels that measure what the hash-grid kernels are bounded by on this GPU:
 * independent random loads of entry_bytes (4, 8, 16) from a table of 2^log2_entries entries, and
 * random red.global.add.f16x2 into a table of 2^log2_entries 4-byte entries. n_threads x per_thread
 * accesses per launch (per_thread a multiple of 8 for the gather); time the launch with CUDA events. */
/* `blocks` x `threads` that stay resident for `nanoseconds` and only sleep: measures what a kernel
 * loses by sharing its SMs with resident blocks of another stream (scripts/corun_probe.py). */
int ngpb_probe_resident(void* stream, uint32_t blocks, uint32_t threads, uint64_t nanoseconds);
int ngpb_probe_gather(void* stream, const void* table, uint32_t log2_entries, uint32_t entry_bytes, uint32_t n_threads, uint32_t per_thread, uint32_t* sink);
int ngpb_probe_scatter(void* stream, void* table, uint32_t log2_entries, uint32_t n_threads, uint32_t per_thread);
/* vec packed f16x2 words (1, 2 or 4: 4 / 8 / 16 aligned bytes) per reduction instruction */
int ngpb_probe_scatter_vec(void* stream, void* table, uint32_t log2_entries, uint32_t vec, uint32_t n_threads, uint32_t per_thread);
/* `passes` coalesced 16-byte-per-lane reads of buf through L2 only (ld.global.cg): the L2 -> SM bandwidth peak */
int ngpb_probe_l2_stream(void* stream, const void* buf, size_t bytes, uint32_t passes, uint32_t blocks, uint32_t* sink);

/* ---- self test of the tcgen05 operand views (tc_device.cuh); diagnostics, no reference counterpart ----
 * mode 0: D[128xN] = X[128xK] W[NxK]^T; mode 1: D[128xN] = X[128xK] W[KxN];
 * mode 2: D[64xN] = X[128x64]^T W[128xN]. fp16 row-major inputs, fp32 row-major D (device). */
int ngpb_selftest_umma(void* stream, int mode, uint32_t K, uint32_t N, const void* X, const void* W, float* D);
/* Issue-to-completion cycles of `reps` back-to-back batches of one operand view on one SM (same modes):
 * out[0] = cycles from the first tcgen05.mma to the commit's arrival, out[1] = cycles spent issuing. DEVICE out[2]. */
int ngpb_probe_umma_rate(void* stream, int mode, uint32_t K, uint32_t N, uint32_t reps, unsigned long long* out);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
