// Fused Adam + EMA parameter update (see optimizer.h).
#include "optimizer.h"

#include <cooperative_groups.h>

#include <cmath>
#include <limits>

namespace ngpb {

struct AdamArgs {
	uint32_t n_elements, n_matrix_weights;
	float relative_weight_decay, absolute_weight_decay, weight_clipping_magnitude;
	float loss_scale, learning_rate, non_matrix_learning_rate_factor;
	uint32_t optimize_matrix_params, optimize_non_matrix_params;
	float beta1, beta2, epsilon, lower_lr_bound, upper_lr_bound, l2_reg;
	// EMA
	uint32_t begin;          // first element of the launch (multiple of 4); n_elements is one past the last
	uint32_t adam;           // 0: EMA only (elements whose Adam step another rank ran, after the all-gather of the weights)
	uint32_t has_ema, ema_full_precision;
	float ema_decay, ema_debias_old, ema_debias_new;
	float inv_loss_scale, log2_beta1, log2_beta2;
};

// beta^step for the bias correction. Every parameter has its own step count (hash-grid entries
// only step when they receive a gradient, adam.h:76-78,93), so the power cannot be hoisted to the
// host; powf costs ~60 instructions twice per parameter and made this streaming kernel issue
// bound (ncu, round 1: 73 % SM / 24 % DRAM throughput). ex2.approx of step * log2(beta) is 2
// instructions and exact to ~2 ulp over the range that matters (the term vanishes as step grows).
NGPB_D float pow_beta(float log2_beta, uint32_t step) {
	float r;
	asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(log2_beta * (float)step));
	return r;
}

// ema_step_half_precision / ema_step_full_precision (ema.h:46-78) with every rounding spelled out: the owner of a shard
// (update_one) and the other ranks (ema_only_kernel) must produce the same bits, whatever the compiler would contract.
NGPB_D float ema_filter(float prev, float w, float decay, float debias_old, float debias_new) {
	return __fmul_rn(__fadd_rn(__fmul_rn(__fmul_rn(prev, decay), debias_old), __fmul_rn(w, __fsub_rn(1.0f, decay))), debias_new);
}

// One parameter: the arithmetic of adam_step (adam.h:72-118) followed by ema_step_half_precision /
// ema_step_full_precision (ema.h:46-78), expression for expression.
NGPB_D void update_one(const AdamArgs& a, uint32_t i, float grad_h, float& w_fp, __half& w_h, float& m1, float& m2, uint32_t& steps,
	__half& ema_h, float* ema_tmp)
{
	float gradient = grad_h * a.inv_loss_scale;
	bool skip;
	if (i >= a.n_matrix_weights) skip = !a.optimize_non_matrix_params || gradient == 0;
	else skip = !a.optimize_matrix_params;
	if (!a.adam) skip = true;

	if (!skip) {
		const float weight_fp = w_fp;
		if (i < a.n_matrix_weights) gradient += a.l2_reg * weight_fp;
		const float gradient_sq = gradient * gradient;
		const float first_moment = m1 = a.beta1 * m1 + (1 - a.beta1) * gradient;
		const float second_moment = m2 = a.beta2 * m2 + (1 - a.beta2) * gradient_sq;
		float learning_rate = a.learning_rate;
		if (i >= a.n_matrix_weights) learning_rate *= a.non_matrix_learning_rate_factor;
		const uint32_t current_step = ++steps;
		learning_rate *= sqrtf(1 - pow_beta(a.log2_beta2, current_step)) / (1 - pow_beta(a.log2_beta1, current_step));
		const float effective_learning_rate = fminf(fmaxf(learning_rate / (sqrtf(second_moment) + a.epsilon), a.lower_lr_bound), a.upper_lr_bound);
		const float decayed_weight = (1 - a.relative_weight_decay * learning_rate) * weight_fp - copysignf(a.absolute_weight_decay * learning_rate, weight_fp);
		float new_weight = decayed_weight - effective_learning_rate * first_moment;
		if (a.weight_clipping_magnitude != 0.0f) new_weight = fminf(fmaxf(new_weight, -a.weight_clipping_magnitude), a.weight_clipping_magnitude);
		w_fp = new_weight;
		w_h = __float2half(new_weight);
	}
	if (a.has_ema) {
		const float prev = a.ema_full_precision ? *ema_tmp : __half2float(ema_h);
		const float filtered = ema_filter(prev, __half2float(w_h), a.ema_decay, a.ema_debias_old, a.ema_debias_new);
		if (a.ema_full_precision) *ema_tmp = filtered;
		ema_h = __float2half(filtered);
	}
}

// Four parameters per thread: 8/16-byte accesses on every stream. Parameters whose gradient is
// zero skip the Adam state entirely (only grad + fp16 weight + EMA are touched), as in the
// reference, so untouched hash-table entries cost 8 bytes instead of 38.
// Measured and not kept (round 2, fox step, 0.096 ms): streaming ld.cs / st.cs on the Adam state and the EMA copy,
// meant to keep the fp16 table in L2 for the next step's gathers: 0.130 ms and no faster gathers; the fp16 weight /
// EMA loads issued ahead of the gradient test: 0.099 ms; an L2 access-policy window over the table: 0.162 ms.
__global__ void __launch_bounds__(256) adam_ema_step_kernel(const __grid_constant__ AdamArgs a,
	float* __restrict__ weights_fp32, __half* __restrict__ weights, const __half* __restrict__ gradients,
	float* __restrict__ first_moments, float* __restrict__ second_moments, uint32_t* __restrict__ param_steps,
	__half* __restrict__ weights_ema, float* __restrict__ ema_tmp)
{
	const uint32_t i0 = a.begin + (blockIdx.x * blockDim.x + threadIdx.x) * 4;
	if (i0 >= a.n_elements) return;
	if (i0 + 4 <= a.n_elements) {
		const uint2 graw = __ldg(reinterpret_cast<const uint2*>(gradients + i0));
		__half g[4];
		*reinterpret_cast<uint2*>(g) = graw;
		bool any_adam = false;
#pragma unroll
		for (uint32_t k = 0; k < 4; ++k) {
			const uint32_t i = i0 + k;
			const bool skip = i >= a.n_matrix_weights ? (!a.optimize_non_matrix_params || __half2float(g[k]) * a.inv_loss_scale == 0) : !a.optimize_matrix_params;
			any_adam |= !skip && a.adam;
		}
		if (!any_adam && !a.has_ema) return;
		__half wh[4], eh[4];
		*reinterpret_cast<uint2*>(wh) = *reinterpret_cast<const uint2*>(weights + i0);
		if (a.has_ema) *reinterpret_cast<uint2*>(eh) = *reinterpret_cast<const uint2*>(weights_ema + i0);
		float wf[4], m1[4], m2[4], et[4];
		uint32_t st[4];
		if (any_adam) {
			*reinterpret_cast<float4*>(wf) = *reinterpret_cast<const float4*>(weights_fp32 + i0);
			*reinterpret_cast<float4*>(m1) = *reinterpret_cast<const float4*>(first_moments + i0);
			*reinterpret_cast<float4*>(m2) = *reinterpret_cast<const float4*>(second_moments + i0);
			*reinterpret_cast<uint4*>(st) = *reinterpret_cast<const uint4*>(param_steps + i0);
		}
		if (a.has_ema && a.ema_full_precision) *reinterpret_cast<float4*>(et) = *reinterpret_cast<const float4*>(ema_tmp + i0);
#pragma unroll
		for (uint32_t k = 0; k < 4; ++k) update_one(a, i0 + k, __half2float(g[k]), wf[k], wh[k], m1[k], m2[k], st[k], eh[k], &et[k]);
		if (any_adam) {
			*reinterpret_cast<float4*>(weights_fp32 + i0) = *reinterpret_cast<const float4*>(wf);
			*reinterpret_cast<float4*>(first_moments + i0) = *reinterpret_cast<const float4*>(m1);
			*reinterpret_cast<float4*>(second_moments + i0) = *reinterpret_cast<const float4*>(m2);
			*reinterpret_cast<uint4*>(param_steps + i0) = *reinterpret_cast<const uint4*>(st);
			*reinterpret_cast<uint2*>(weights + i0) = *reinterpret_cast<const uint2*>(wh);
		}
		if (a.has_ema) {
			*reinterpret_cast<uint2*>(weights_ema + i0) = *reinterpret_cast<const uint2*>(eh);
			if (a.ema_full_precision) *reinterpret_cast<float4*>(ema_tmp + i0) = *reinterpret_cast<const float4*>(et);
		}
	} else {
		for (uint32_t i = i0; i < a.n_elements; ++i) {
			__half eh = a.has_ema ? weights_ema[i] : __float2half(0.0f);
			float et = (a.has_ema && a.ema_full_precision) ? ema_tmp[i] : 0.0f;
			update_one(a, i, __half2float(gradients[i]), weights_fp32[i], weights[i], first_moments[i], second_moments[i], param_steps[i], eh, &et);
			if (a.has_ema) {
				weights_ema[i] = eh;
				if (a.ema_full_precision) ema_tmp[i] = et;
			}
		}
	}
}

// EMA filter alone over [begin, end) (begin a multiple of 8), eight parameters per thread with 16-byte accesses: the pass a
// rank of a sharded run makes over the shards it does not own (6 bytes per parameter: weight read, EMA read + write).
__global__ void __launch_bounds__(256) ema_only_kernel(uint32_t begin, uint32_t end, float decay, float debias_old, float debias_new,
	const __half* __restrict__ weights, __half* __restrict__ weights_ema)
{
	const uint32_t i0 = begin + (blockIdx.x * blockDim.x + threadIdx.x) * 8;
	if (i0 >= end) return;
	if (i0 + 8 <= end) {
		const uint4 w = *reinterpret_cast<const uint4*>(weights + i0);
		uint4 e = *reinterpret_cast<const uint4*>(weights_ema + i0);
		const __half2* w2 = reinterpret_cast<const __half2*>(&w);
		__half2* e2 = reinterpret_cast<__half2*>(&e);
#pragma unroll
		for (uint32_t k = 0; k < 4; ++k) {
			const float2 wf = __half22float2(w2[k]), ef = __half22float2(e2[k]);
			// ema_step_half_precision (ema.h:63-78), expression for expression as in update_one
			e2[k] = __floats2half2_rn(ema_filter(ef.x, wf.x, decay, debias_old, debias_new), ema_filter(ef.y, wf.y, decay, debias_old, debias_new));
		}
		*reinterpret_cast<uint4*>(weights_ema + i0) = e;
	} else {
		for (uint32_t i = i0; i < end; ++i) {
			weights_ema[i] = __float2half(ema_filter(__half2float(weights_ema[i]), __half2float(weights[i]), decay, debias_old, debias_new));
		}
	}
}

void Optimizer::ema_range(cudaStream_t stream, const __half* weights, uint32_t begin, uint32_t end) {
	if (!m_has_ema || begin >= end) return;
	if (m_ema_full_precision || begin % 8 != 0 || (uintptr_t)weights % 16 != 0) {   // rare shapes take the general kernel
		step_range(stream, 1.0f, nullptr, const_cast<__half*>(weights), weights, begin & ~3u, end, false, false);
		return;
	}
	const float debias_old = 1 - (float)std::pow(m_ema_decay, m_current_step - 1), debias_new = 1.0f / (1 - (float)std::pow(m_ema_decay, m_current_step));
	const uint32_t n_threads = div_round_up(end - begin, 8);
	ema_only_kernel<<<div_round_up(n_threads, 256), 256, 0, stream>>>(begin, end, m_ema_decay, debias_old, debias_new, weights, m_weights_ema);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

// -------------------------------------------------------------------------------------------------
Optimizer::Optimizer(const Json& config) { parse(config, true); }
Optimizer::~Optimizer() { release(); }

void Optimizer::release() {
	cudaFree(m_first_moments); cudaFree(m_second_moments); cudaFree(m_param_steps); cudaFree(m_weights_ema); cudaFree(m_ema_tmp);
	m_first_moments = m_second_moments = m_ema_tmp = nullptr;
	m_param_steps = nullptr;
	m_weights_ema = nullptr;
	m_capacity = 0;
}

void Optimizer::parse(const Json& config, bool construct) {
	const Json* cur = &config;
	bool seen_adam = false;
	for (int depth = 0; depth < 8 && cur; ++depth) {
		const std::string otype = cur->value("otype", "Adam");
		if (equals_case_insensitive(otype, "Ema")) {
			if (construct) m_has_ema = true;
			if (cur->contains("decay")) m_ema_decay = (*cur)["decay"].as_float();
			if (cur->contains("full_precision")) m_ema_full_precision = (*cur)["full_precision"].as_bool();
		} else if (equals_case_insensitive(otype, "ExponentialDecay")) {
			if (construct) m_has_decay = true;
			if (cur->contains("decay_base")) m_decay_base = (*cur)["decay_base"].as_float();
			if (cur->contains("decay_interval")) m_decay_interval = (*cur)["decay_interval"].as_uint();
			if (cur->contains("decay_start")) m_decay_start = (*cur)["decay_start"].as_uint();
			if (cur->contains("decay_end")) m_decay_end = (*cur)["decay_end"].as_uint();
		} else if (equals_case_insensitive(otype, "Adam")) {
			const Json& p = *cur;
			if (p.contains("beta1")) m_beta1 = p["beta1"].as_float();
			if (p.contains("beta2")) m_beta2 = p["beta2"].as_float();
			if (p.contains("epsilon")) m_epsilon = p["epsilon"].as_float();
			// Under ExponentialDecay the wrapper re-imposes its own base rate every step
			// (exponential_decay.h:69), so a later "learning_rate" update has no effect there.
			if (p.contains("learning_rate") && (construct || !m_has_decay)) m_base_learning_rate = p["learning_rate"].as_float();
			if (p.contains("l2_reg")) m_l2_reg = p["l2_reg"].as_float();
			if (p.contains("adabound")) m_adabound = p["adabound"].as_bool();
			if (p.contains("relative_decay")) m_relative_weight_decay = p["relative_decay"].as_float();
			if (p.contains("absolute_decay")) m_absolute_weight_decay = p["absolute_decay"].as_float();
			if (p.contains("clipping_magnitude")) m_weight_clipping_magnitude = p["clipping_magnitude"].as_float();
			if (p.contains("non_matrix_learning_rate_factor")) m_non_matrix_learning_rate_factor = p["non_matrix_learning_rate_factor"].as_float();
			if (p.contains("optimize_matrix_params")) m_optimize_matrix_params = p["optimize_matrix_params"].as_bool();
			if (p.contains("optimize_non_matrix_params")) m_optimize_non_matrix_params = p["optimize_non_matrix_params"].as_bool();
			seen_adam = true;
			break;
		} else {
			const bool reference_type = equals_case_insensitive(otype, "Average") || equals_case_insensitive(otype, "Batched") ||
				equals_case_insensitive(otype, "Composite") || equals_case_insensitive(otype, "Lookahead") ||
				equals_case_insensitive(otype, "Novograd") || equals_case_insensitive(otype, "SGD") || equals_case_insensitive(otype, "Shampoo");
			if (reference_type) throw std::runtime_error{"ngp-b200: optimizer type " + otype + " is not available (Ema, ExponentialDecay and Adam are)"};
			throw std::runtime_error{"Invalid optimizer type: " + otype};
		}
		if (!cur->contains("nested")) { cur = nullptr; break; }
		cur = &(*cur)["nested"];
	}
	if (construct && !seen_adam && cur == nullptr) {
		// create_optimizer(params.value("nested", json::object())) -> Adam with defaults
		seen_adam = true;
	}
}

void Optimizer::update_hyperparams(const Json& config) { parse(config, false); }

void Optimizer::set_learning_rate(float v) { m_base_learning_rate = v / m_learning_rate_factor; }

Json Optimizer::hyperparams() const {
	Json adam = Json::object();
	adam["otype"] = "Adam";
	adam["beta1"] = m_beta1; adam["beta2"] = m_beta2; adam["epsilon"] = m_epsilon;
	adam["learning_rate"] = learning_rate(); adam["l2_reg"] = m_l2_reg; adam["adabound"] = m_adabound;
	adam["relative_decay"] = m_relative_weight_decay; adam["absolute_decay"] = m_absolute_weight_decay;
	adam["clipping_magnitude"] = m_weight_clipping_magnitude;
	adam["non_matrix_learning_rate_factor"] = m_non_matrix_learning_rate_factor;
	adam["optimize_matrix_params"] = m_optimize_matrix_params; adam["optimize_non_matrix_params"] = m_optimize_non_matrix_params;
	Json cur = adam;
	if (m_has_decay) {
		Json d = Json::object();
		d["otype"] = "ExponentialDecay"; d["nested"] = cur;
		d["decay_base"] = m_decay_base; d["decay_interval"] = m_decay_interval; d["decay_start"] = m_decay_start; d["decay_end"] = m_decay_end;
		cur = d;
	}
	if (m_has_ema) {
		Json e = Json::object();
		e["otype"] = "EMA"; e["nested"] = cur; e["decay"] = m_ema_decay; e["full_precision"] = m_ema_full_precision;
		cur = e;
	}
	return cur;
}

void Optimizer::allocate(uint32_t n_weights, const std::vector<std::pair<uint32_t, uint32_t>>& layer_sizes) {
	m_n_weights = n_weights;
	if (n_weights <= m_capacity) return;
	release();
	const size_t n = next_multiple(n_weights, 4u);
	NGPB_CUDA_CHECK(cudaMalloc(&m_first_moments, n * sizeof(float)));
	NGPB_CUDA_CHECK(cudaMalloc(&m_second_moments, n * sizeof(float)));
	NGPB_CUDA_CHECK(cudaMalloc(&m_param_steps, n * sizeof(uint32_t)));
	NGPB_CUDA_CHECK(cudaMemset(m_first_moments, 0, n * sizeof(float)));
	NGPB_CUDA_CHECK(cudaMemset(m_second_moments, 0, n * sizeof(float)));
	NGPB_CUDA_CHECK(cudaMemset(m_param_steps, 0, n * sizeof(uint32_t)));
	if (m_has_ema) {
		NGPB_CUDA_CHECK(cudaMalloc(&m_weights_ema, n * sizeof(__half)));
		NGPB_CUDA_CHECK(cudaMemset(m_weights_ema, 0, n * sizeof(__half)));
		if (m_ema_full_precision) {
			NGPB_CUDA_CHECK(cudaMalloc(&m_ema_tmp, n * sizeof(float)));
			NGPB_CUDA_CHECK(cudaMemset(m_ema_tmp, 0, n * sizeof(float)));
		}
	}
	m_capacity = n_weights;
	m_n_matrix_weights = 0;
	for (const auto& l : layer_sizes) m_n_matrix_weights += l.first * l.second;
}

void Optimizer::step(cudaStream_t stream, float loss_scale, float* weights_fp32, __half* weights, const __half* gradients) {
	step_range(stream, loss_scale, weights_fp32, weights, gradients, 0, m_n_weights, true, true);
}

// Elements [begin, end) of the step (begin a multiple of 4): the whole of Optimizer::step when the range is everything.
// `advance`: this call opens a new step (learning-rate schedule, step count, EMA debias factors); later calls of the
// same step pass false. `adam` false runs the EMA filter only - for weights another rank has stepped (sharded
// optimizer, Testbed::train_dp): every rank filters the same fp16 weights, so the EMA copies stay bit-identical.
void Optimizer::step_range(cudaStream_t stream, float loss_scale, float* weights_fp32, __half* weights, const __half* gradients,
	uint32_t begin, uint32_t end, bool adam, bool advance)
{
	if (m_n_weights == 0 || begin >= end) { if (!advance) return; }
	if (begin % 4 != 0 || end > m_n_weights) throw std::runtime_error{"ngp-b200 optimizer: step range must start at a multiple of 4 inside the weights"};
	if (advance) {
	// ExponentialDecayOptimizer::step (exponential_decay.h:60-71), evaluated on the step count
	// *before* Adam increments it.
	if (m_has_decay) {
		if (m_current_step == 0) m_learning_rate_factor = 1.0f;
		if (m_current_step >= m_decay_start && (m_current_step - m_decay_start) % m_decay_interval == 0 && m_current_step <= m_decay_end) {
			m_learning_rate_factor *= m_decay_base;
		}
	}
	++m_current_step;
	}
	if (begin >= end) return;
	const AdamArgs a = make_args(loss_scale, begin, end, adam);
	const bool aligned = ((uintptr_t)weights_fp32 % 16 == 0) && ((uintptr_t)weights % 8 == 0) && ((uintptr_t)gradients % 8 == 0);
	if (!aligned) throw std::runtime_error{"ngp-b200 optimizer: parameter buffers must be 16-byte (fp32) / 8-byte (fp16) aligned"};
	const uint32_t n_threads = div_round_up(end - begin, 4);
	adam_ema_step_kernel<<<div_round_up(n_threads, 256), 256, 0, stream>>>(a, weights_fp32, weights, gradients,
		m_first_moments, m_second_moments, m_param_steps, m_weights_ema, m_ema_tmp);
	NGPB_CUDA_CHECK(cudaGetLastError());
}

AdamArgs Optimizer::make_args(float loss_scale, uint32_t begin, uint32_t end, bool adam) const {
	AdamArgs a{};
	a.n_elements = end;
	a.begin = begin;
	a.adam = adam ? 1u : 0u;
	a.n_matrix_weights = m_n_matrix_weights;
	a.relative_weight_decay = m_relative_weight_decay;
	a.absolute_weight_decay = m_absolute_weight_decay;
	a.weight_clipping_magnitude = m_weight_clipping_magnitude;
	a.loss_scale = loss_scale;
	a.learning_rate = m_base_learning_rate * m_learning_rate_factor;
	a.non_matrix_learning_rate_factor = m_non_matrix_learning_rate_factor;
	a.optimize_matrix_params = m_optimize_matrix_params;
	a.optimize_non_matrix_params = m_optimize_non_matrix_params;
	a.beta1 = m_beta1; a.beta2 = m_beta2; a.epsilon = m_epsilon;
	a.lower_lr_bound = 0.0f;
	a.upper_lr_bound = std::numeric_limits<float>::max();
	if (m_adabound) { // adam.h:165-168
		a.lower_lr_bound = 0.1f - 0.1f / ((1 - m_beta2) * (float)m_current_step + 1);
		a.upper_lr_bound = 0.1f + 0.1f / ((1 - m_beta2) * (float)m_current_step);
	}
	a.l2_reg = m_l2_reg;
	a.has_ema = m_has_ema;
	a.ema_full_precision = m_ema_full_precision;
	a.ema_decay = m_ema_decay;
	a.ema_debias_old = 1 - (float)std::pow(m_ema_decay, m_current_step - 1); // ema.h:106-107
	a.ema_debias_new = 1.0f / (1 - (float)std::pow(m_ema_decay, m_current_step));

	a.inv_loss_scale = 1.0f / loss_scale;
	a.log2_beta1 = std::log2(m_beta1);
	a.log2_beta2 = std::log2(m_beta2);
	return a;
}

// -------------------------------------------------------------------------------------------------
// Sharded optimizer step fused with its two collectives over NVLink peer memory (one node, <= 8 ranks):
//
//     wait until every rank's gradients are complete            (flags in peer memory, system-scope release / acquire)
//     for my shard: g = sum over ranks of their gradient        (P2P loads, fixed rank order, rounded once to fp16)
//                   Adam + EMA on g                              (the arithmetic of adam_ema_step_kernel)
//                   new fp16 weight -> every rank's weights     (P2P stores)
//     wait until every rank has written its shard everywhere
//
// i.e. reduce-scatter + optimizer + all-gather in ONE launch, each byte crossing NVLink once in each direction while
// the Adam arithmetic runs - instead of ncclReduceScatter, a kernel, ncclAllGather back to back (three launches, each
// paying its own latency: measured 67 + 50 + 56 us at two ranks for 25.7 MB). The launch is cooperative: the two
// rendezvous are grid-wide.
NGPB_D void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
NGPB_D uint32_t ld_acquire_sys(const uint32_t* p) {
	uint32_t v;
	asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
NGPB_D uint2 ld_peer_u2(const void* p) {
	uint2 v;   // system-scope relaxed load: never served from this SM's L1 (the address held last step's gradient)
	asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
	return v;
}

// flags: every rank owns uint32 [2][8]; row 0 = "gradients of epoch e complete on rank q", row 1 = "rank q has written its
// shard of epoch e everywhere". Rank r writes column r of every peer's array and polls its own.
NGPB_D void rendezvous(const P2PPeers& p, uint32_t row) {
	if (blockIdx.x == 0 && threadIdx.x < (uint32_t)p.world) {
		const uint32_t q = threadIdx.x;
		st_release_sys(p.flags[q] + row * 8 + p.rank, p.epoch);
		const long long t0 = clock64();
		while (ld_acquire_sys(p.flags[p.rank] + row * 8 + q) < p.epoch) {
			if (clock64() - t0 > 8000000000ll) { atomicExch(p.error, 1u + row); break; }   // ~4 s: a peer died; never hang the GPU
		}
	}
}

__global__ void __launch_bounds__(256) adam_ema_step_p2p_kernel(const __grid_constant__ AdamArgs a, const __grid_constant__ P2PPeers p,
	float* __restrict__ weights_fp32, float* __restrict__ first_moments, float* __restrict__ second_moments, uint32_t* __restrict__ param_steps,
	__half* __restrict__ weights_ema, float* __restrict__ ema_tmp)
{
	cooperative_groups::grid_group grid = cooperative_groups::this_grid();
	rendezvous(p, 0);
	grid.sync();
	const uint32_t n_quads = (a.n_elements - a.begin + 3) / 4;
	for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_quads; q += gridDim.x * blockDim.x) {
		const uint32_t i0 = a.begin + q * 4;   // a.begin and the shard length are multiples of 8: whole quads, 8-byte aligned
		float gs[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll 8
		for (int r = 0; r < p.world; ++r) {
			const uint2 raw = ld_peer_u2(p.grads[r] + i0);
			const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), hi = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
			gs[0] += lo.x; gs[1] += lo.y; gs[2] += hi.x; gs[3] += hi.y;
		}
		__half wh[4], eh[4];
		float wf[4], m1[4], m2[4], et[4] = {0.0f, 0.0f, 0.0f, 0.0f}, g[4];
		uint32_t st[4];
		bool any_adam = false;
#pragma unroll
		for (uint32_t k = 0; k < 4; ++k) {
			g[k] = __half2float(__float2half(gs[k]));   // the gradient Adam sees is an fp16 value, as on one GPU
			const uint32_t i = i0 + k;
			const bool skip = i >= a.n_matrix_weights ? (!a.optimize_non_matrix_params || g[k] * a.inv_loss_scale == 0) : !a.optimize_matrix_params;
			any_adam |= !skip;
		}
		*reinterpret_cast<uint2*>(wh) = *reinterpret_cast<const uint2*>(p.params[p.rank] + i0);
		if (a.has_ema) *reinterpret_cast<uint2*>(eh) = *reinterpret_cast<const uint2*>(weights_ema + i0);
		if (any_adam) {
			*reinterpret_cast<float4*>(wf) = *reinterpret_cast<const float4*>(weights_fp32 + i0);
			*reinterpret_cast<float4*>(m1) = *reinterpret_cast<const float4*>(first_moments + i0);
			*reinterpret_cast<float4*>(m2) = *reinterpret_cast<const float4*>(second_moments + i0);
			*reinterpret_cast<uint4*>(st) = *reinterpret_cast<const uint4*>(param_steps + i0);
		}
		if (a.has_ema && a.ema_full_precision) *reinterpret_cast<float4*>(et) = *reinterpret_cast<const float4*>(ema_tmp + i0);
#pragma unroll
		for (uint32_t k = 0; k < 4; ++k) update_one(a, i0 + k, g[k], wf[k], wh[k], m1[k], m2[k], st[k], eh[k], &et[k]);
		if (any_adam) {
			*reinterpret_cast<float4*>(weights_fp32 + i0) = *reinterpret_cast<const float4*>(wf);
			*reinterpret_cast<float4*>(first_moments + i0) = *reinterpret_cast<const float4*>(m1);
			*reinterpret_cast<float4*>(second_moments + i0) = *reinterpret_cast<const float4*>(m2);
			*reinterpret_cast<uint4*>(param_steps + i0) = *reinterpret_cast<const uint4*>(st);
			const uint2 packed = *reinterpret_cast<const uint2*>(wh);
#pragma unroll 8
			for (int r = 0; r < p.world; ++r) *reinterpret_cast<uint2*>(p.params[r] + i0) = packed;   // all-gather: every rank's copy
		}
		if (a.has_ema) {
			*reinterpret_cast<uint2*>(weights_ema + i0) = *reinterpret_cast<const uint2*>(eh);
			if (a.ema_full_precision) *reinterpret_cast<float4*>(ema_tmp + i0) = *reinterpret_cast<const float4*>(et);
		}
	}
	__threadfence_system();
	grid.sync();
	rendezvous(p, 1);
	grid.sync();   // nobody leaves (and lets the stream's next kernel read the weights) before the peers' shards have landed
}

void Optimizer::step_p2p(cudaStream_t stream, float loss_scale, float* weights_fp32, const P2PPeers& peers, uint32_t begin, uint32_t end) {
	if (begin % 8 != 0) throw std::runtime_error{"ngp-b200 optimizer: a peer-to-peer shard must start at a multiple of 8"};
	// opens the step (schedule, counts) like step_range(..., advance = true); an empty shard still takes part in the rendezvous
	step_range(stream, loss_scale, weights_fp32, nullptr, nullptr, end, end, true, true);
	AdamArgs a = make_args(loss_scale, begin, std::max(begin, end), true);
	static int blocks = [] {
		int per_sm = 0;
		NGPB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, adam_ema_step_p2p_kernel, 256, 0));
		return std::max(1, std::min(per_sm, 4)) * sm_count();
	}();
	P2PPeers p = peers;
	void* args[] = {&a, &p, &weights_fp32, &m_first_moments, &m_second_moments, &m_param_steps, &m_weights_ema, &m_ema_tmp};
	NGPB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)adam_ema_step_p2p_kernel, dim3(blocks), dim3(256), args, 0, stream));
}

} // namespace ngpb
