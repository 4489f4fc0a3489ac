/* oracle.h - CPU restatement of the reference's hot-path arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is linked into, imported by or called from
 * the product (libngp_b200.so / the Python package). Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load liboracle.so, and only as the
 * checker or the timed CPU baseline.
 *
 * PARITY: pinned at function grain against the reference's OWN code - oracle/ref_host.cu compiles
 * the reference's header functions (hash / index / Morton / cascade / mip / voxel stepping / box /
 * lens / sRGB / losses / activations / image choice / Sobol / pcg32 / SH) for the host from
 * /root/reference into oracle/_ref/, tests/golden/ref_host_vectors.npz holds their outputs, and
 * tests/test_ref_pin_cpu.py compares word for word (bit-exact; the stepping-space maps and the
 * ray-grain march within the tolerance stated there).
 * Since round 2 also pinned at kernel grain against the reference's OWN CUDA KERNELS run on a B200: the
 * reference is built headless from its sources (baseline/build_ref.sh), baseline/harness.cu runs tiny-cuda-nn's
 * kernel_grid / kernel_grid_backward, kernel_mlp_fused (+ backward), ngp::NerfNetwork and adam_step / ema_step on
 * the seeded inputs of tests/ref_gpu_cases.py, tests/golden/ref_gpu_vectors.npz holds their outputs, and
 * tests/test_ref_gpu_fixture_cpu.py checks this oracle against them (grid forward bit-identical; MLP, gradients and
 * Adam within the tolerances stated there - the reference accumulates in fp16 and sums atomics in arbitrary order).
 * The container formats (msgpack, merge_patch) are pinned to the reference's own JSON library
 * (oracle/ref_json.cc, tests/golden/ref_json_vectors.json.gz), snapshots and loader metadata to a snapshot the
 * reference wrote (tests/golden/ref_fox_step5000.snapshot.json), the loader's frame order to the reference's vendored
 * NaturalSort header (tests/golden/ref_natural_sort.json.gz), and the host image decoders to the reference's vendored
 * stb_image and tinyexr compiled as they are (oracle/ref_images.cc, tests/test_host_io_cpu.py).
 * One stated departure: the training march places samples on the stepping-space lattice with fixed-polynomial
 * log / exp; against the reference's functions its per-ray counts agree on 99.6 - 99.8 % of rays.
 * What still rests on the restatement alone: the loss kernel's glue and the occupancy update as whole kernels
 * (their arithmetic is pinned function by function above).
 */
#ifndef NGPB_ORACLE_H
#define NGPB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Grid configuration (tcnn grid.h:668-730). grid_type: 0 Hash, 1 Dense, 2 Tiled.
 * hash_type: 0 Prime, 1 CoherentPrime, 2 ReversedPrime. interpolation: 0 Nearest, 1 Linear,
 * 2 Smoothstep. `scale_override`: optional per-level scales (the library reports the values it
 * evaluated with the device's exp2f); NULL = host exp2f. */
typedef struct {
	uint32_t n_dims;
	uint32_t n_features_per_level;
	uint32_t n_levels;
	uint32_t log2_hashmap_size;
	uint32_t base_resolution;
	float per_level_scale;
	int grid_type;
	int hash_type;
	int interpolation;
	const float* scale_override;
} oracle_grid_config;

/* offsets[n_levels+1] in entries; returns total entries (grid.h:692-723). */
uint32_t oracle_grid_offsets(const oracle_grid_config* cfg, uint32_t* offsets);
/* scales[n_levels], resolutions[n_levels] as the kernels see them (common_device.h:709-718). */
void oracle_grid_levels(const oracle_grid_config* cfg, float* scales, uint32_t* resolutions);
/* idx[(i*L + level)*2^D + corner] (common_device.h:690-707, 856-868). */
void oracle_grid_indices(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride, uint32_t* idx);
/* kernel_grid (grid.h:48-170): fp16 params, fp16 FMA chain; out AoS [n][out_stride] halves
 * (bit patterns); padding columns set to 1.0. max_level as tcnn (1.0 = all). */
void oracle_grid_forward(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* params, uint16_t* out, uint32_t out_stride, float max_level);
/* kernel_grid_backward (grid.h:214-320): each contribution is half(weight)*half(dy) rounded to
 * fp16 as the reference's `(GRAD_T)weight * grad`; contributions are summed in double (the
 * reference sums them with fp16 atomics in arbitrary order). grad: [entries*F] doubles. */
void oracle_grid_backward(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* dL_dy, uint32_t dy_stride, double* grad, float max_level);
/* grid.h:172-211 + 322-349: dL_dx[n][D] in fp32. */
void oracle_grid_backward_input(const oracle_grid_config* cfg, uint32_t n, const float* positions, uint32_t in_stride,
	const uint16_t* params, const uint16_t* dL_dy, uint32_t dy_stride, float* dL_dx, float max_level);

/* pcg32 (tcnn/dependencies/pcg32/pcg32.h) helpers for tests. */
void oracle_pcg32_floats(uint64_t seed, uint64_t advance, uint32_t n, float* out);
/* generate_random_uniform (tcnn random.h:39-73): element -> stream position mapping. */
void oracle_random_uniform(uint64_t seed, size_t n, float* out, float lower, float upper);
void oracle_random_uniform_state(uint64_t* rng_state, size_t n, float* out, float lower, float upper);

/* ---- fused MLP (mlp_oracle.cpp) ------------------------------------------------------------
 * Weight layout of tcnn FullyFusedMLP (fully_fused_mlp.cu:656-678): W0 [width x in] | hidden
 * [width x width] x (n_hidden-1) | last [padded_out x width], row-major fp16. */
size_t oracle_mlp_n_params(uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out);
/* kernel_mlp_fused (fully_fused_mlp.cu:499-557): x fp16 [n][x_stride], out fp16 [n][out_stride]
 * (padded_out columns), hidden (optional) fp16 [n_hidden][n][width]. accumulate_fp16: 0 = fp32
 * accumulator (TMEM), 1 = fp16 accumulator per 16-wide K block (models wmma, :66-68). */
void oracle_mlp_forward(uint32_t n, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, int relu,
	const uint16_t* weights, const uint16_t* x, uint32_t x_stride, uint16_t* out, uint32_t out_stride, uint16_t* hidden, int accumulate_fp16);
/* kernel_mlp_fused_backward + weight-gradient GEMMs (fully_fused_mlp.cu:150-259, 735-836). */
void oracle_mlp_backward(uint32_t n, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, int relu,
	const uint16_t* weights, const uint16_t* x, uint32_t x_stride, const uint16_t* hidden,
	const uint16_t* dL_dout, uint32_t dy_stride, uint16_t* dL_dx, uint32_t dx_stride, double* dW);
/* initialize_xavier_uniform per matrix (gpu_matrix.h:284-299); rng_state = {state, inc}, advanced. */
void oracle_mlp_init(uint64_t* rng_state, uint32_t in_width, uint32_t width, uint32_t n_hidden, uint32_t padded_out, float scale, float* params);
/* kernel_sh, degree <= 4 (spherical_harmonics.h:46-72, common_device.h:339-365). */
void oracle_sh_encode(uint32_t n, uint32_t degree, const float* dirs, uint32_t dir_stride, uint16_t* out, uint32_t out_stride);
void oracle_pcg32_seed(uint64_t seed, uint64_t* rng_state);
void oracle_pcg32_advance(uint64_t* rng_state, int64_t delta);

/* ---- NeRF driver kernels (nerf_oracle.cpp) ------------------------------------------------------- */
typedef struct {
	const void* pixels;          /* HOST pixels */
	int image_data_type;         /* 1 Byte, 2 Half, 3 Float */
	int lens_mode;               /* 0 Perspective, 1 OpenCV */
	float lens_params[7];
	int res_x, res_y;
	float principal_x, principal_y;
	float focal_x, focal_y;
	float xform[12];             /* column-major 4 x 3 */
} oracle_nerf_image;
typedef struct {
	float aabb_min[3], aabb_max[3];
	uint32_t n_rays, n_rays_total, max_samples, max_samples_compacted, n_training_images, max_mip;
	float cone_angle_constant, near_distance, loss_scale;
	int snap_to_pixel_centers, random_bg_color, linear_colors, color_space, loss_type, rgb_activation, density_activation;
	float background_color[3];
	uint32_t ray_begin, n_rays_local; /* layout parity with ngpb_nerf_train_config_t; the oracle always walks the whole batch */
} oracle_nerf_config;
/* generate_training_samples_nerf (src/testbed_nerf.cu:679-838), rays in index order. */
uint64_t oracle_nerf_generate_samples(const oracle_nerf_config* cfg, const oracle_nerf_image* images, const uint8_t* bitfield,
	const uint64_t* rng_state, uint32_t* numsteps, float* rays, float* coords, uint64_t max_samples);
/* compute_loss_kernel_train_nerf (src/testbed_nerf.cu:841-1160) for n_active rays given as
 * (ray_indices, numsteps_in {count, base}, rays, coords, rgbsigma fp16 [..][4]). Outputs per ray:
 * compacted sample count and loss; per INPUT sample (base + j): dL/d(rgb sigma) as float [..][4]. */
void oracle_nerf_compute_loss(const oracle_nerf_config* cfg, const oracle_nerf_image* images, float mean_density, uint32_t n_active,
	const uint32_t* ray_indices, const uint32_t* numsteps_in, const float* rays, const float* coords, const uint16_t* rgbsigma,
	const uint64_t* rng_state, uint32_t* compacted_numsteps, float* loss, float* dloss);
/* grid_to_bitfield + bitfield_max_pool + mean (src/testbed_nerf.cu:284-331, 2362-2379). */
void oracle_nerf_bitfield(uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density);
/* Occupancy-grid maintenance (src/testbed_nerf.cu:74-282): mark_untrained_density_grid (in place on `grid`),
 * generate_grid_samples_nerf_nonuniform (positions [n][4], indices [n]; aabb6 = min xyz, max xyz),
 * splat_grid_samples_nerf_max_nearest_neighbor (density_raw fp16 bits), ema_grid_samples_nerf. */
void oracle_nerf_mark_untrained(uint32_t n_elements, float* grid, uint32_t n_images, const oracle_nerf_image* images, int clear_visible_voxels);
void oracle_nerf_generate_grid_samples(uint32_t n, const uint64_t* rng_state, uint32_t step, const float* aabb6, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh);
void oracle_nerf_splat_grid_samples(uint32_t n, const uint32_t* indices, const uint16_t* density_raw, float* grid_tmp, int density_activation);
void oracle_nerf_ema_grid_samples(uint32_t n_elements, float decay, float* grid, const float* grid_tmp);
/* CudaRenderBuffer accumulate + tonemap (src/render_buffer.cu:232-346, 533-565): frames [spp][n][4] -> out [n][4]. */
void oracle_render_buffer_epilogue(uint32_t n, uint32_t spp, const float* frames, float exposure, const float* background, int color_space,
	int output_color_space, int curve, float* out);
/* deterministic log / exp used by the stepping (Cephes single precision) */
float oracle_det_logf(float x);
float oracle_det_expf(float x);
/* advance_n_steps / calc_dt (nerf_device.cuh:414-424) */
float oracle_advance_n_steps(float t, float cone_angle, float n);
uint32_t oracle_morton3d(uint32_t x, uint32_t y, uint32_t z);
uint32_t oracle_morton3d_invert(uint32_t x);

/* ---- NeRF rendering (nerf_oracle.cpp) -----------------------------------------------------------
 * Per-pixel restatement of NerfTracer (src/testbed_nerf.cu:333-677, 1375-1489): a pixel's result
 * does not depend on how the reference groups rays into windows, so the oracle walks ray by ray. */
typedef struct {
	float aabb_min[3], aabb_max[3];
	float camera[12];            /* column-major 4 x 3, NGP convention */
	uint32_t width, height;
	float focal_x, focal_y, screen_cx, screen_cy;
	uint32_t spp_index, pixel_begin, pixel_end, max_mip;
	float cone_angle_constant, min_transmittance;
	int rgb_activation, density_activation, linear_colors, snap_to_pixel_centers;
	uint32_t max_steps_per_ray;
	int lens_mode;               /* render lens: 0 Perspective, 1 OpenCV (src/testbed.cu:477-478) */
	float lens_params[7];
} oracle_render_config;
/* init_rays_with_payload_kernel_nerf + advance_pos_nerf + generate_next_nerf_network_inputs: ALL
 * samples of every pixel in [pixel_begin, pixel_end) until the ray leaves the box (or
 * max_steps_per_ray). numsteps[2*(pix - begin)] = {count, base}; coords [..][7] may be NULL to count
 * only. Returns the total number of samples. */
uint64_t oracle_nerf_render_march(const oracle_render_config* cfg, const uint8_t* bitfield, uint32_t* numsteps, float* coords, uint64_t max_samples);
/* composite_kernel_nerf (Shade) + shade_kernel_nerf: rgbsigma fp16 [..][4] per marched sample ->
 * rgba [n][4] (linear, premultiplied), depth [n], steps [n] (samples composited before termination). */
void oracle_nerf_render_composite(const oracle_render_config* cfg, const uint32_t* numsteps, const float* coords, const uint16_t* rgbsigma,
	float* rgba, float* depth, uint32_t* steps);
/* ld_random_val (random_val.cuh:287-291), dims 0 and 1. */
float oracle_ld_random_val(uint32_t index, uint32_t seed, uint32_t dim);

/* fp16 <-> fp32 (round to nearest even), exposed so tests can build fp16 inputs. */
uint16_t oracle_float_to_half(float f);
float oracle_half_to_float(uint16_t h);

/* The render march of explicit rays [n][6] (origin, normalised direction): count, first and last
 * sample distance per ray; the body of oracle_nerf_render_march. Compared with the reference's own
 * functions composed the same way (oracle/ref_host.cu ref_march_rays). aabb = min3, max3. */
void oracle_march_rays(uint32_t n, const float* rays, const uint32_t* pix, uint32_t spp_index, const float* aabb, float cone_angle,
	uint32_t max_mip, const uint8_t* bitfield, uint32_t max_steps, uint32_t* counts, float* t_first, float* t_last);

/* Function-grain probes with the interface of oracle/ref_host.cu's ref_probe (rows of 32-bit words,
 * floats by bit pattern): cascaded_grid_idx_at, mip_from_pos, mip_from_dt, distance_to_next_voxel,
 * ray_intersect, opencv_undistort, uv_to_ray, srgb_to_linear, linear_to_srgb, loss, network_to_rgb,
 * network_to_density, image_idx, ld_random_val, warp_dt, stepping, morton3D, morton3D_invert.
 * Returns 0, or -1 for an unknown name. */
int oracle_probe(const char* name, uint32_t n, const uint32_t* in, uint32_t k_in, uint32_t* out, uint32_t k_out);

int oracle_num_threads(void);
void oracle_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
