// NeRF driver kernels: training-sample generation, loss + gradient + compaction, occupancy-grid
// maintenance and rendering. Replaces the kernels of src/testbed_nerf.cu listed in SURVEY.md
// section 2.2-F / 8a (a10-a13). All buffers are caller-owned device memory.
#pragma once

#include "nerf_device.cuh"

namespace ngpb {

constexpr uint32_t kCoordFloats = 7; // NerfCoordinate: pos3, dt, dir3 (nerf_device.cuh:175-201)

// Counters of one training step, in device memory (NerfCounters, testbed.h; the reference keeps
// them in three separate GPUMemory objects).
struct NerfStepCounters {
	uint32_t numsteps;            // samples written by the generator (before compaction)
	uint32_t numsteps_compacted;  // samples kept by the loss kernel (may exceed the batch size)
	uint32_t rays;                // rays that produced samples
	float loss_sum;               // sum of the per-ray losses (filled on request)
};

struct NerfTrainConfig {
	Aabb aabb;
	uint32_t n_rays;              // rays_per_batch
	uint32_t n_rays_total;
	uint32_t max_samples;         // capacity of the coords / rgbsigma buffers
	uint32_t max_samples_compacted; // target batch size
	uint32_t n_training_images;
	uint32_t max_mip;
	float cone_angle_constant;
	float near_distance;
	float loss_scale;
	int snap_to_pixel_centers;
	int random_bg_color;
	int linear_colors;
	int color_space;
	int loss_type;
	int rgb_activation;
	int density_activation;
	float background_color[3];
	// Ray sharding (data-parallel training): this call handles the n_rays_local rays starting at
	// global index ray_begin of the n_rays-ray batch; image choice, per-ray RNG position and the
	// 1/n_rays loss normalisation use the GLOBAL index and count, so the union of the shards is the
	// single-GPU batch. n_rays_local == 0 means the whole batch.
	uint32_t ray_begin;
	uint32_t n_rays_local;
};
NGPB_HD uint32_t local_ray_count(const NerfTrainConfig& c) { return c.n_rays_local ? c.n_rays_local : c.n_rays; }

struct NerfTrainBuffers {
	const ImageMeta* metadata;        // [n_training_images]
	const uint8_t* density_bitfield;  // cascaded occupancy bitfield
	const float* mean_density;        // 1 float
	NerfStepCounters* counters;
	uint32_t* ray_indices;            // [n_rays]
	float* rays_unnormalized;         // [n_rays][6] origin, direction
	uint32_t* numsteps;               // [n_rays][2] count, base (rewritten by the loss kernel)
	float* coords;                    // [max_samples][7]
	const __half* rgbsigma;           // [max_samples][4] network output (packed)
	float* coords_compacted;          // [max_samples_compacted][7]
	__half* dloss_doutput;            // [max_samples_compacted][4]
	float* loss;                      // [n_rays]
	uint32_t* compact_src;            // [max_samples_compacted] or null: index in `coords` each compacted sample was copied from
};

// src/testbed_nerf.cu:679-838. `rng` is the step's generator (m_rng); ray i uses rng.advance(16 i).
void nerf_generate_training_samples(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng);
// src/testbed_nerf.cu:841-1160 followed by the three fill_rollover launches (:2862-2870).
void nerf_compute_loss_and_compact(cudaStream_t stream, const NerfTrainConfig& cfg, const NerfTrainBuffers& buf, Pcg32 rng);

// ---- occupancy grid (src/testbed_nerf.cu:74-331, 2271-2379) ----------------------------------------
// mark_untrained_density_grid
void nerf_mark_untrained_density_grid(cudaStream_t stream, uint32_t n_elements, float* grid, uint32_t n_images, const ImageMeta* metadata, bool clear_visible_voxels);
// generate_grid_samples_nerf_nonuniform: positions [n][4] (pos3, dt) and cell indices.
void nerf_generate_grid_samples(cudaStream_t stream, uint32_t n, Pcg32 rng, uint32_t step, const Aabb& aabb, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh);
// splat_grid_samples_nerf_max_nearest_neighbor; density_raw: fp16 density-network output per sample.
void nerf_splat_grid_samples(cudaStream_t stream, uint32_t n, const uint32_t* indices, const __half* density_raw, float* grid_tmp, int density_activation);
// ema_grid_samples_nerf
void nerf_ema_grid_samples(cudaStream_t stream, uint32_t n_elements, float decay, float* grid, const float* grid_tmp);
// update_density_grid_mean_and_bitfield: mean over cascade 0, grid_to_bitfield, bitfield_max_pool x7.
void nerf_update_bitfield(cudaStream_t stream, uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density);

inline size_t nerf_bitfield_bytes() { return (size_t)kNerfGridCells * kNerfCascades / 8; }

} // namespace ngpb
