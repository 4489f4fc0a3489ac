// Optimizer chain of the shipped configs: [Ema ∘] [ExponentialDecay ∘] Adam
// (configs/nerf/base.json:5-22, configs/image/base.json:5-21). Replaces tcnn's AdamOptimizer
// (optimizers/adam.h:48-329), EmaOptimizer (optimizers/ema.h:45-217) and
// ExponentialDecayOptimizer (optimizers/exponential_decay.h:44-162) with one object and ONE
// streaming kernel per step: the reference makes two full passes over the parameters
// (adam_step, then ema_step); here Adam and the EMA update read and write each parameter once.
#pragma once

#include "module.h"

namespace ngpb {

struct AdamArgs;

// Peer views of a sharded step over NVLink (Optimizer::step_p2p): base pointers of every rank's gradient buffer, fp16
// weight buffer and rendezvous flags (uint32 [2][8], zero-initialised), mapped into this process with cudaIpc.
struct P2PPeers {
	int rank, world;
	const __half* grads[8];
	__half* params[8];
	uint32_t* flags[8];
	uint32_t epoch;      // 1, 2, 3 ... one per step, the same on every rank
	uint32_t* error;     // device word of this rank, set when a rendezvous timed out
};

class Optimizer {
public:
	explicit Optimizer(const Json& config);
	~Optimizer();
	Optimizer(const Optimizer&) = delete;
	Optimizer& operator=(const Optimizer&) = delete;

	// Optimizer::allocate(n_weights, layer_sizes): the first sum(rows*cols) weights are "matrix"
	// weights (L2 regularised, never skipped), the rest are not (adam.h:136-157).
	void allocate(uint32_t n_weights, const std::vector<std::pair<uint32_t, uint32_t>>& layer_sizes);
	// Optimizer::step (adam.h:159-197 inside ema.h:103-137 inside exponential_decay.h:60-71).
	void step(cudaStream_t stream, float loss_scale, float* weights_fp32, __half* weights, const __half* gradients);
	void step_range(cudaStream_t stream, float loss_scale, float* weights_fp32, __half* weights, const __half* gradients,
		uint32_t begin, uint32_t end, bool adam, bool advance);
	// EMA filter of [begin, end) alone, for weights another rank has stepped (after the step has been opened on this rank).
	void ema_range(cudaStream_t stream, const __half* weights, uint32_t begin, uint32_t end);
	// The step of shard [begin, end) fused with the gradient reduce-scatter and the weight all-gather over peer memory.
	void step_p2p(cudaStream_t stream, float loss_scale, float* weights_fp32, const P2PPeers& peers, uint32_t begin, uint32_t end);

	float learning_rate() const { return m_base_learning_rate * m_learning_rate_factor; }
	void set_learning_rate(float v);
	uint32_t step_count() const { return m_current_step; }
	uint32_t n_weights() const { return m_n_weights; }
	// EMA weights (used for inference, trainer.h:329-333) or nullptr without an Ema stage.
	__half* custom_weights() const { return m_has_ema ? m_weights_ema : nullptr; }
	void update_hyperparams(const Json& config);
	Json hyperparams() const;

	// state access for snapshots / tests (device pointers, n_weights elements each)
	float* first_moments() const { return m_first_moments; }
	float* second_moments() const { return m_second_moments; }
	uint32_t* param_steps() const { return m_param_steps; }
	void set_step_count(uint32_t s, float lr_factor) { m_current_step = s; m_learning_rate_factor = lr_factor; }
	float learning_rate_factor() const { return m_learning_rate_factor; }
	float base_learning_rate() const { return m_base_learning_rate; }
	void set_base_learning_rate(float v) { m_base_learning_rate = v; }
	// EmaOptimizer::deserialize (ema.h:197-204): with full_precision the fp32 accumulator restarts from the
	// restored fp16 EMA weights (otherwise the next step would filter from zero).
	void reseed_ema_accumulator(cudaStream_t stream) {
		if (m_has_ema && m_ema_full_precision && m_ema_tmp) cast_fp16_to_fp32(stream, m_n_weights, m_weights_ema, m_ema_tmp);
	}

private:
	AdamArgs make_args(float loss_scale, uint32_t begin, uint32_t end, bool adam) const;
	void parse(const Json& config, bool construct);
	void release();

	// chain structure
	bool m_has_ema = false;
	bool m_has_decay = false;
	// Adam (defaults adam.h:311-326)
	float m_non_matrix_learning_rate_factor = 1.0f;
	float m_base_learning_rate = 1e-3f;
	float m_beta1 = 0.9f, m_beta2 = 0.999f, m_epsilon = 1e-8f, m_l2_reg = 1e-8f;
	float m_relative_weight_decay = 0.0f, m_absolute_weight_decay = 0.0f, m_weight_clipping_magnitude = 0.0f;
	bool m_adabound = false, m_optimize_matrix_params = true, m_optimize_non_matrix_params = true;
	// ExponentialDecay (defaults exponential_decay.h:152-157)
	float m_learning_rate_factor = 1.0f;
	float m_decay_base = 0.1f;
	uint32_t m_decay_interval = 10000, m_decay_start = 10000, m_decay_end = 10000000;
	// Ema (defaults ema.h:209-210)
	float m_ema_decay = 0.99f;
	bool m_ema_full_precision = false;

	uint32_t m_n_weights = 0, m_n_matrix_weights = 0, m_current_step = 0;
	uint32_t m_capacity = 0;
	float* m_first_moments = nullptr;
	float* m_second_moments = nullptr;
	uint32_t* m_param_steps = nullptr;
	__half* m_weights_ema = nullptr;
	float* m_ema_tmp = nullptr;
};

} // namespace ngpb
