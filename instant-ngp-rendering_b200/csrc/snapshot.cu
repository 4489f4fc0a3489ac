// Device side of snapshot restore: what Testbed::load_snapshot (src/testbed.cu:4841-4955) and
// Trainer::deserialize (tcnn trainer.h:290-314) do to the trainer and the occupancy grid once the
// msgpack container has been read. The container itself (msgpack + zlib, key names) is host work
// and lives in the Python mirror (snapshot.py).
#include "testbed.h"

#include <vector>

namespace ngpb {

// Trainer::set_params (trainer.h:256-269) / set_params_full_precision (:242-254): the inference
// weights take the values, the training weights copy them, the fp32 master follows.
void Testbed::set_params(const void* host_params, bool full_precision, size_t n_params) {
	if (m_mode == kModeImage ? !m_image_network : !m_network) throw std::runtime_error{"ngp-b200: set_params needs a network (reload_network_from_json first)"};
	if (n_params != m_n_params) throw std::runtime_error{full_precision ? "Can't set fp params because buffer has the wrong size." : "Can't set params because buffer has the wrong size."};
	if (!host_params) throw std::runtime_error{"ngp-b200: set_params: null pointer"};
	join_prefetch();
	if (full_precision) {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(m_params_fp32.ptr, host_params, n_params * sizeof(float), cudaMemcpyHostToDevice, m_stream));
		cast_fp32_to_fp16(m_stream, n_params, m_params_fp32.as<float>(), m_params.as<__half>());
	} else {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(m_params.ptr, host_params, n_params * sizeof(__half), cudaMemcpyHostToDevice, m_stream));
		cast_fp16_to_fp32(m_stream, n_params, m_params.as<__half>(), m_params_fp32.as<float>());
	}
	if (params_inference() != m_params.as<__half>()) {
		NGPB_CUDA_CHECK(cudaMemcpyAsync(params_inference(), m_params.ptr, n_params * sizeof(__half), cudaMemcpyDeviceToDevice, m_stream));
		m_optimizer->reseed_ema_accumulator(m_stream);
	}
	NGPB_CUDA_CHECK(cudaMemsetAsync(m_gradients.ptr, 0, n_params * sizeof(__half), m_stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream)); // the host buffer may go away after the call
}

// load_snapshot :4885-4897: fp16 grid -> float grid, then update_density_grid_mean_and_bitfield.
// A size of 0 is the valid state of a model that never trained.
void Testbed::set_density_grid_fp16(const void* host_grid, size_t n_cells) {
	if (m_mode != kModeNerf) throw std::runtime_error{"ngp-b200: the density grid belongs to the NeRF mode"};
	const size_t n_grid = (size_t)kNerfGridCells * (m_max_cascade + 1);
	if (n_cells == 0) return;
	if (n_cells != n_grid) throw std::runtime_error{"Incompatible number of grid cascades."};
	if (!host_grid) throw std::runtime_error{"ngp-b200: set_density_grid_fp16: null pointer"};
	join_prefetch();
	m_density_grid.reserve(n_grid * sizeof(float), m_stream);
	m_density_grid_tmp.reserve(n_grid * sizeof(float), m_stream);
	m_density_bitfield.reserve(nerf_bitfield_bytes(), m_stream);
	m_density_mean.reserve(sizeof(float), m_stream);
	// staged through the scratch grid: it is as large as the float grid, the halves need half of it
	NGPB_CUDA_CHECK(cudaMemcpyAsync(m_density_grid_tmp.ptr, host_grid, n_grid * sizeof(__half), cudaMemcpyHostToDevice, m_stream));
	cast_fp16_to_fp32(m_stream, n_grid, m_density_grid_tmp.as<__half>(), m_density_grid.as<float>());
	update_density_grid_mean_and_bitfield();
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

// load_snapshot :4931-4932 (m_training_step, m_loss_scalar) after reset_network(false).
void Testbed::restore_training_state(uint32_t training_step, float loss) {
	join_prefetch();
	touch_generator_inputs();
	m_training_step = training_step;
	m_loss_scalar = loss;
}

} // namespace ngpb
