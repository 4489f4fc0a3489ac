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
const char* ngpb_last_error(void);
int ngpb_version(void);
/* 1 when a CUDA device is usable from this process, else 0 (never an error). */
int ngpb_has_device(void);
/* tcnn::cpp::batch_size_granularity(), cpp_api.h:62; tcnn common.h:235 (256). */
uint32_t ngpb_batch_size_granularity(void);
/* tcnn::cpp::default_loss_scale(), cpp_api.h:75; tcnn common.h:232 (128 for fp16). */
float ngpb_default_loss_scale(int precision);

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

/* ---- self test of the tcgen05 operand views (tc_device.cuh) --------------------------------
 * mode 0: D[128xN] = X[128xK] W[NxK]^T; mode 1: D[128xN] = X[128xK] W[KxN];
 * mode 2: D[64xN] = X[128x64]^T W[128xN]. fp16 row-major inputs, fp32 row-major D (device). */
int ngpb_selftest_umma(void* stream, int mode, uint32_t K, uint32_t N, const void* X, const void* W, float* D);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
