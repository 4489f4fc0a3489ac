// Testbed::train and the NeRF training step (src/testbed.cu:4020-4109, src/testbed_nerf.cu:2271-2946).
#include "testbed.h"
#include "dist.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace ngpb {

__global__ void reduce_loss_kernel(uint32_t n, const float* __restrict__ loss, float* __restrict__ out) {
	__shared__ float warp_part[8];
	float s = 0.0f;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += loss[i];
#pragma unroll
	for (uint32_t o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
	if (threadIdx.x % 32 == 0) warp_part[threadIdx.x / 32] = s;
	__syncthreads();
	if (threadIdx.x == 0) {
		float t = 0.0f;
		for (uint32_t w = 0; w < blockDim.x / 32; ++w) t += warp_part[w];
		atomicAdd(out, t);
	}
}

void Testbed::update_density_grid_mean_and_bitfield() {
	join_prefetch();
	touch_generator_inputs();
	nerf_update_bitfield(m_stream, m_max_cascade, m_density_grid.as<float>(), m_density_bitfield.as<uint8_t>(), m_density_mean.as<float>());
}

void Testbed::update_density_grid_nerf(float decay, uint32_t n_uniform, uint32_t n_nonuniform) {
	join_prefetch();
	upload_metadata();
	const uint32_t n_elements = kNerfGridCells * (m_max_cascade + 1);
	const uint32_t n_samples = n_uniform + n_nonuniform;
	const uint32_t n_padded = next_multiple(n_samples, kTile);
	m_grid_positions.reserve((size_t)n_padded * 4 * sizeof(float), m_stream);
	m_grid_indices.reserve((size_t)n_padded * sizeof(uint32_t), m_stream);
	m_grid_mlp_out.reserve((size_t)n_padded * sizeof(__half), m_stream);

	if (m_training_step == 0) {
		m_density_grid_ema_step = 0;
		nerf_mark_untrained_density_grid(m_stream, n_elements, m_density_grid.as<float>(), n_images(), metadata_gpu(), true);
	}

	NGPB_CUDA_CHECK(cudaMemsetAsync(m_density_grid_tmp.ptr, 0, sizeof(float) * n_elements, m_stream));
	float* positions = m_grid_positions.as<float>();
	uint32_t* indices = m_grid_indices.as<uint32_t>();
	nerf_generate_grid_samples(m_stream, n_uniform, m_density_grid_rng, m_density_grid_ema_step, m_aabb, m_density_grid.as<float>(), positions, indices, m_max_cascade + 1, -0.01f);
	m_density_grid_rng.advance(1ll << 32);
	nerf_generate_grid_samples(m_stream, n_nonuniform, m_density_grid_rng, m_density_grid_ema_step, m_aabb, m_density_grid.as<float>(),
		positions + (size_t)n_uniform * 4, indices + n_uniform, m_max_cascade + 1, kNerfMinOpticalThickness);
	m_density_grid_rng.advance(1ll << 32);
	if (n_padded > n_samples) NGPB_CUDA_CHECK(cudaMemsetAsync(positions + (size_t)n_samples * 4, 0, (size_t)(n_padded - n_samples) * 4 * sizeof(float), m_stream));

	// density network on the training parameters (use_inference_params = false, :2350)
	m_network->density(m_stream, n_padded, positions, 4, m_grid_mlp_out.as<__half>(), 1, params());
	nerf_splat_grid_samples(m_stream, n_samples, indices, m_grid_mlp_out.as<__half>(), m_density_grid_tmp.as<float>(), density_activation);
	nerf_ema_grid_samples(m_stream, n_elements, decay, m_density_grid.as<float>(), m_density_grid_tmp.as<float>());
	++m_density_grid_ema_step;
	update_density_grid_mean_and_bitfield();
}

void Testbed::training_prep_nerf() {
	if (n_images() == 0) return;
	const uint32_t n_cascades = m_max_cascade + 1;
	if (m_training_step < 256) update_density_grid_nerf(density_grid_decay, kNerfGridCells * n_cascades, 0);
	else update_density_grid_nerf(density_grid_decay, kNerfGridCells / 4 * n_cascades, kNerfGridCells / 4 * n_cascades);
}

static uint32_t prep_interval(uint32_t training_step) { return std::min(std::max(training_step / 16u, 1u), 16u); }

void Testbed::train(uint32_t batch_size) {
	if (m_mode == kModeImage) {
		if (!m_image_network) throw std::runtime_error{"Unable to create a neural network trainer."};
		if (batch_size == 0 || batch_size % kBatchSizeGranularity != 0) throw std::runtime_error{"ngp-b200: batch size must be a positive multiple of 256"};
		Json leaf = Json::object();
		leaf["otype"] = "Adam";
		leaf["optimize_matrix_params"] = train_network;
		leaf["optimize_non_matrix_params"] = train_encoding;
		m_optimizer->update_hyperparams(leaf);
		train_image(batch_size, m_training_step % 16 == 0);
		return;
	}
	if (m_dist) { train_dp(batch_size); return; }
	if (shard_world > 1) throw std::runtime_error{"ngp-b200: a sharded testbed trains with train_begin / all-reduce / train_end (or init_data_parallel, then train)"};
	train_begin(batch_size);
	train_end(batch_size);
}

void Testbed::train_begin(uint32_t batch_size) {
	train_forward(batch_size);
	train_backward(batch_size, shard_world == 1);
}

void Testbed::train_forward(uint32_t batch_size) {
	if (n_images() == 0) throw std::runtime_error{"Cannot train without training data."};
	if (!m_network) throw std::runtime_error{"Unable to create a neural network trainer."};
	if (batch_size == 0 || batch_size % kBatchSizeGranularity != 0) throw std::runtime_error{"ngp-b200: batch size must be a positive multiple of 256"};
	if (shard_world == 0 || shard_rank >= shard_world || (batch_size / shard_world) % kBatchSizeGranularity != 0 || batch_size % shard_world != 0 || kBatchSizeGranularity % shard_world != 0)
		throw std::runtime_error{"ngp-b200: shard_world must divide 256 and leave a per-rank batch that is a multiple of 256"};
	upload_metadata();

	if (m_training_step % prep_interval(m_training_step) == 0) {
		prof_mark(m_stream, "density_grid_update");
		training_prep_nerf();
	}

	Json leaf = Json::object();
	leaf["otype"] = "Adam";
	leaf["optimize_matrix_params"] = train_network;
	leaf["optimize_non_matrix_params"] = train_encoding;
	m_optimizer->update_hyperparams(leaf);

	m_get_loss_scalar = m_training_step % 16 == 0;
	train_nerf_forward(batch_size / shard_world);
}

void Testbed::reserve_step_buffers(uint32_t target_batch_size) {
	const uint32_t max_samples = target_batch_size * 16; // "somewhat of a worst case" (:2685)
	m_counters.reserve(3 * sizeof(NerfStepCounters), m_stream);
	m_ray_indices.reserve((size_t)(1u << 18) * sizeof(uint32_t), m_stream);
	m_rays.reserve((size_t)(1u << 18) * 6 * sizeof(float), m_stream);
	m_numsteps.reserve((size_t)(1u << 18) * 2 * sizeof(uint32_t), m_stream);
	m_loss.reserve((size_t)(1u << 18) * sizeof(float), m_stream);
	m_coords.reserve((size_t)max_samples * kCoordFloats * sizeof(float), m_stream);
	m_rgbsigma.reserve((size_t)max_samples * 4 * sizeof(__half), m_stream);
	m_coords_compacted.reserve((size_t)target_batch_size * kCoordFloats * sizeof(float), m_stream);
	m_dloss.reserve((size_t)target_batch_size * 4 * sizeof(__half), m_stream);
	m_compact_src.reserve((size_t)target_batch_size * sizeof(uint32_t), m_stream);
}

// Configuration of the step that starts from the current host state (rays_per_batch, n_rays_total).
NerfTrainConfig Testbed::make_train_config(uint32_t target_batch_size) const {
	NerfTrainConfig cfg;
	memset(&cfg, 0, sizeof(cfg));
	const uint32_t n_rays_local = rays_per_batch / shard_world; // rays_per_batch is a multiple of 256, shard_world divides 256
	cfg.aabb = m_aabb;
	cfg.n_rays = rays_per_batch;
	cfg.ray_begin = shard_rank * n_rays_local;
	cfg.n_rays_local = n_rays_local;
	cfg.n_rays_total = m_training_step == 0 ? 0 : n_rays_total;
	// The reference bounds the generator by last step's measured sample count (:2730-2735) because
	// its inference batch is sized on the host; here the sample count stays on the device and the
	// consumers read it there, so the full buffer is available.
	cfg.max_samples = target_batch_size * 16;
	cfg.max_samples_compacted = target_batch_size;
	cfg.n_training_images = n_images();
	cfg.max_mip = m_max_cascade;
	cfg.cone_angle_constant = m_cone_angle_constant;
	cfg.near_distance = near_distance;
	cfg.loss_scale = kLossScaleFp16;
	cfg.snap_to_pixel_centers = snap_to_pixel_centers;
	cfg.random_bg_color = random_bg_color;
	cfg.linear_colors = linear_colors;
	cfg.color_space = color_space;
	cfg.loss_type = loss_type;
	cfg.rgb_activation = rgb_activation;
	cfg.density_activation = density_activation;
	for (int k = 0; k < 3; ++k) cfg.background_color[k] = background_color[k];
	return cfg;
}

NerfTrainBuffers Testbed::make_train_buffers(uint32_t counter_slot) {
	NerfTrainBuffers buf{};
	buf.metadata = metadata_gpu();
	buf.density_bitfield = m_density_bitfield.as<uint8_t>();
	buf.mean_density = m_density_mean.as<float>();
	buf.counters = m_counters.as<NerfStepCounters>() + counter_slot;
	buf.ray_indices = m_ray_indices.as<uint32_t>();
	buf.rays_unnormalized = m_rays.as<float>();
	buf.numsteps = m_numsteps.as<uint32_t>();
	buf.coords = m_coords.as<float>();
	buf.rgbsigma = m_rgbsigma.as<__half>();
	buf.coords_compacted = m_coords_compacted.as<float>();
	buf.dloss_doutput = m_dloss.as<__half>();
	buf.loss = m_loss.as<float>();
	buf.compact_src = m_compact_src.as<uint32_t>();
	return buf;
}

void Testbed::join_prefetch() {
	if (!m_prefetch.pending) return;
	NGPB_CUDA_CHECK(cudaStreamWaitEvent(m_stream, m_prefetch_done, 0));
	m_prefetch.pending = false;
}

// Rays of the NEXT step on the side stream. Called after the host has seen this step's counters,
// i.e. after the loss kernel - the last reader of the sample buffers - has finished.
void Testbed::launch_prefetch(uint32_t target_batch_size) {
	const NerfTrainConfig cfg = make_train_config(target_batch_size);
	const uint32_t slot = m_counter_slot ^ 1u;
	const NerfTrainBuffers buf = make_train_buffers(slot);
	// The march (a warp per ray, ~70 us) is an issue-bound kernel; beside the fused MLP kernel, which is a latency
	// chain of small MMAs and epilogues, both lose (measured: the MLP kernel 88 -> 132 us with the march on its SMs).
	// The gradient scatter and the optimizer that follow are bound by L2 atomics and HBM and leave the issue slots
	// idle, so the march starts when the MLP kernel has finished and hides under those two (~150 us).
	// NGPB_PREFETCH_EARLY=1 starts it right after the loss kernel instead (the round-1 schedule, when the march
	// was a 250-340 us serial loop and needed the whole backward pass to hide under).
	static const bool early = getenv("NGPB_PREFETCH_EARLY") && atoi(getenv("NGPB_PREFETCH_EARLY"));
	if (!early) NGPB_CUDA_CHECK(cudaStreamWaitEvent(m_side_stream, m_mlp_done, 0));
	NGPB_CUDA_CHECK(cudaMemsetAsync(buf.counters, 0, sizeof(NerfStepCounters), m_side_stream));
	nerf_generate_training_samples(m_side_stream, cfg, buf, m_rng);
	NGPB_CUDA_CHECK(cudaEventRecord(m_prefetch_done, m_side_stream));
	m_prefetch.pending = true;
	m_prefetch.cfg = cfg;
	m_prefetch.rng = m_rng;
	m_prefetch.epoch = m_epoch;
}

void Testbed::train_nerf_forward(uint32_t target_batch_size) {
	const uint32_t max_samples = target_batch_size * 16;
	const NerfTrainConfig cfg = make_train_config(target_batch_size);
	n_rays_total = cfg.n_rays_total + rays_per_batch;

	// Adopt the prefetched samples when they were made from exactly this step's inputs.
	const bool adopt = m_prefetch.pending && m_prefetch.epoch == m_epoch && memcmp(&m_prefetch.cfg, &cfg, sizeof(cfg)) == 0
		&& m_prefetch.rng.state == m_rng.state && m_prefetch.rng.inc == m_rng.inc;
	join_prefetch(); // either way the side stream is done with the buffers before this stream touches them
	reserve_step_buffers(target_batch_size);
	if (adopt) { m_counter_slot ^= 1u; ++prefetches_adopted; }
	const NerfTrainBuffers buf = make_train_buffers(m_counter_slot);

	NGPB_CUDA_CHECK(cudaMemsetAsync(m_loss.ptr, 0, sizeof(float) * cfg.n_rays_local, m_stream));
	// dloss / coords of a batch that the loss kernel does not fill completely and whose roll-over
	// finds no input must not carry stale gradients
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_dloss.ptr, 0, (size_t)target_batch_size * 4 * sizeof(__half), m_stream));
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_compact_src.ptr, 0, (size_t)target_batch_size * sizeof(uint32_t), m_stream));   // slots nothing is compacted into carry zero gradients; their source must still be a valid row

	prof_mark(m_stream, "generate_training_samples");
	if (!adopt) {
		NGPB_CUDA_CHECK(cudaMemsetAsync(buf.counters, 0, sizeof(NerfStepCounters), m_stream));
		nerf_generate_training_samples(m_stream, cfg, buf, m_rng);
	}
	// network on the training parameters over exactly the generated samples (device-side count)
	m_network->inference_counted(m_stream, max_samples, &buf.counters->numsteps, m_coords.as<float>(), kCoordFloats, m_rgbsigma.as<__half>(), 4, params());
	prof_mark(m_stream, "loss_and_compact");
	nerf_compute_loss_and_compact(m_stream, cfg, buf, m_rng);
	if (m_get_loss_scalar) reduce_loss_kernel<<<sm_count(), 256, 0, m_stream>>>(cfg.n_rays_local, m_loss.as<float>(), &buf.counters->loss_sum);
	// Sharded: the caller sums the counters over the ranks in a buffer whose address never changes.
	if (shard_world > 1) NGPB_CUDA_CHECK(cudaMemcpyAsync(counters(), buf.counters, sizeof(NerfStepCounters), cudaMemcpyDeviceToDevice, m_stream));
}

void Testbed::train_backward(uint32_t batch_size, bool counters_final) {
	if (batch_size == 0 || shard_world == 0 || batch_size % shard_world != 0) throw std::runtime_error{"ngp-b200: train_backward after train_forward with the same batch size"};
	const uint32_t target_batch_size = batch_size / shard_world;
	// With the step's global counts known (single GPU, or already summed over the ranks) their copy
	// to the host overtakes the backward pass; otherwise train_end reads them.
	m_counters_read = counters_final;
	if (m_counters_read) {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(m_counters_host, counters(), sizeof(NerfStepCounters), cudaMemcpyDeviceToHost, m_stream));
		NGPB_CUDA_CHECK(cudaEventRecord(m_counters_ready, m_stream));
	}
	m_network->backward_strided(m_stream, target_batch_size, m_coords_compacted.as<float>(), kCoordFloats, m_dloss.as<__half>(), 4, gradients(), params(), m_mlp_done, m_encoder_done, m_compact_src.as<uint32_t>());
	m_rng.advance(1ll << 32);
}

void Testbed::train_end(uint32_t target_batch_size) {
	// In a sharded run the caller has all-reduced the gradients and the counters between
	// train_begin and train_end, so every rank steps the same parameters with the same counts.
	prof_mark(m_stream, "optimizer");
	m_optimizer->step(m_stream, kLossScaleFp16, params_fp32(), params(), gradients());
	prof_mark(m_stream, "end");
	++m_training_step;

	if (m_counters_read) {
		// the backward pass and the optimizer stay in flight behind this wait
		NGPB_CUDA_CHECK(cudaEventSynchronize(m_counters_ready));
	} else {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(m_counters_host, counters(), sizeof(NerfStepCounters), cudaMemcpyDeviceToHost, m_stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	}
	control_loop(target_batch_size);
	// No prefetch into a step that refreshes the occupancy grid first (its rays must see the new
	// bitfield), under the phase timer (phases are timed back to back), or when the counters were
	// only read at the end of the step (two-phase sharded protocol).
	const bool next_step_preps = m_training_step % prep_interval(m_training_step) == 0;
	if (train_prefetch && m_counters_read && !next_step_preps && !prof_enabled()) launch_prefetch(target_batch_size / shard_world);
}

// NerfCounters::update_after_training (:2422-2446)
void Testbed::control_loop(uint32_t target_batch_size) {
	measured_batch_size = 0;
	measured_batch_size_before_compaction = 0;
	if (m_counters_host->numsteps == 0 || m_counters_host->numsteps_compacted == 0) {
		m_loss_scalar = 0.0f;
		throw std::runtime_error{"Nerf training generated 0 samples. Aborting training."};
	}
	measured_batch_size_before_compaction = m_counters_host->numsteps;
	measured_batch_size = m_counters_host->numsteps_compacted;
	if (m_get_loss_scalar) m_loss_scalar = m_counters_host->loss_sum * (float)measured_batch_size / (float)target_batch_size;
	rays_per_batch = (uint32_t)((float)rays_per_batch * (float)target_batch_size / (float)measured_batch_size);
	rays_per_batch = std::min(next_multiple(rays_per_batch, kBatchSizeGranularity), 1u << 18);
}

} // namespace ngpb
