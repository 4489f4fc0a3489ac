// Data-parallel NeRF training inside the library (see dist.h): rays of a step sharded over the ranks, and a
// SHARDED OPTIMIZER instead of "all-reduce, then every rank steps everything":
//
//   forward (rays, inference, loss) on this rank's rays ......................... main stream
//   all-reduce of the 16-byte step counters, copy to the host .................... aux stream, beside the backward pass
//   backward: this rank's fp16 gradients ......................................... main stream
//   reduce-scatter of the gradients: rank r receives the sum of shard r ........... main stream, NCCL over NVLink
//   Adam + EMA on shard r (1/P of the optimizer state is ever touched per rank) ... main stream
//   all-gather of the fp16 weights ................................................ main stream
//   EMA filter of the other ranks' shards (every rank filters the same weights) ... main stream
//
// Against the round-1 protocol (counter all-reduce + host wait between forward and backward, one 25.7 MB
// all-reduce, full Adam pass on every rank): the host never waits inside the step, the collective moves the same
// bytes as the all-reduce did (a ring all-reduce IS reduce-scatter + all-gather) and the optimizer's HBM traffic
// per rank drops by the number of ranks. Parameters stay bit-identical across ranks by construction: a weight is
// stepped on exactly one rank and everyone receives those bits.
#include "dist.h"
#include "testbed.h"

#include <dlfcn.h>

#include <cstdlib>
#include <vector>

namespace ngpb {

const NcclApi& NcclApi::get() {
	static NcclApi api = [] {
		void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // the copy the process already uses (torch's)
		if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
		if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
		if (!h) throw std::runtime_error{std::string{"ngp-b200: cannot bind NCCL (libnccl.so.2): "} + dlerror()};
		NcclApi a{};
		auto sym = [&](const char* name) {
			void* p = dlsym(h, name);
			if (!p) throw std::runtime_error{std::string{"ngp-b200: libnccl.so.2 lacks "} + name};
			return p;
		};
		a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
		a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
		a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
		a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
		a.ReduceScatter = (decltype(a.ReduceScatter))sym("ncclReduceScatter");
		a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
		a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
		a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
		a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
		return a;
	}();
	return api;
}

DistContext::~DistContext() {
	if (main || aux) {
		const NcclApi& n = NcclApi::get();
		if (main) n.CommDestroy(main);
		if (aux) n.CommDestroy(aux);
	}
	for (int r = 0; r < world && r < 8; ++r) {
		if (r == rank) continue;
		if (peer_grads[r]) cudaIpcCloseMemHandle(peer_grads[r]);
		if (peer_params[r]) cudaIpcCloseMemHandle(peer_params[r]);
		if (peer_flags[r]) cudaIpcCloseMemHandle(peer_flags[r]);
	}
	if (flags) cudaFree(flags);
	if (peer_flags_dev) cudaFree(peer_flags_dev);
	if (aux_stream) cudaStreamDestroy(aux_stream);
	if (forward_done) cudaEventDestroy(forward_done);
}

// Layout of a rank's flag allocation (uint32 words): [0, 16) the two rendezvous rows of the fused optimizer kernel
// (optimizer.cu), [32] its error word, [64, 192) the counter mailbox: [parity][source rank][8] = 4 payload words
// (the 16-byte NerfStepCounters) + a stamp word.
constexpr size_t kFlagBytes = 1024;
constexpr uint32_t kMailboxWord = 64;

// All-reduce of the 16-byte step counters over peer memory, one warp: lane q writes this rank's counters into rank q's
// mailbox and then the step's stamp (system-scope release), polls its own mailbox for rank q's stamp (acquire), and lane
// 0 adds the payloads in rank order - every rank computes the same sums, the float one included. The mailbox is double
// buffered by step parity: a rank cannot be two steps ahead of a peer, because each step needs every peer's stamp.
// Replaces a grouped ncclAllReduce pair whose kernel ran beside the fused MLP kernel and took SMs from it.
__global__ void counters_allreduce_p2p_kernel(NerfStepCounters* __restrict__ counters, uint32_t* const* __restrict__ peer_flags_dev, int rank, int world,
	uint32_t epoch, uint32_t* __restrict__ error)
{
	const uint32_t lane = threadIdx.x;
	const uint32_t parity = epoch & 1u;
	uint32_t mine[4];
	{
		const uint32_t* c = reinterpret_cast<const uint32_t*>(counters);
		for (int k = 0; k < 4; ++k) mine[k] = c[k];
	}
	if (lane < (uint32_t)world) {
		uint32_t* slot = peer_flags_dev[lane] + kMailboxWord + (parity * 8 + (uint32_t)rank) * 8;
		for (int k = 0; k < 4; ++k) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(slot + k), "r"(mine[k]) : "memory");
		asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(slot + 4), "r"(epoch) : "memory");
		const uint32_t* from = peer_flags_dev[rank] + kMailboxWord + (parity * 8 + lane) * 8;
		const long long t0 = clock64();
		for (;;) {
			uint32_t stamp;
			asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(stamp) : "l"(from + 4) : "memory");
			if (stamp == epoch) break;
			if (clock64() - t0 > 8000000000ll) { atomicExch(error, 3u); break; }   // ~4 s: a peer died; never hang the GPU
		}
	}
	__syncwarp();
	if (lane == 0) {
		uint32_t sum[3] = {0, 0, 0};
		float loss = 0.0f;
		for (int q = 0; q < world; ++q) {
			const uint32_t* from = peer_flags_dev[rank] + kMailboxWord + (parity * 8 + (uint32_t)q) * 8;
			uint32_t v[4];
			for (int k = 0; k < 4; ++k) asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v[k]) : "l"(from + k) : "memory");
			sum[0] += v[0]; sum[1] += v[1]; sum[2] += v[2];
			loss += __uint_as_float(v[3]);
		}
		counters->numsteps = sum[0];
		counters->numsteps_compacted = sum[1];
		counters->rays = sum[2];
		counters->loss_sum = loss;
	}
}

// Maps every rank's gradient buffer, fp16 weight buffer and flag array into this process (cudaIpc; the handles travel
// through an all-gather on the aux communicator). The two buffers move to plain cudaMalloc memory first: pool memory
// (cudaMallocAsync) cannot be exported. Returns false - and the step stays on NCCL - when a peer cannot be mapped.
static bool setup_p2p(Testbed& tb, DistContext& d, DeviceBuffer& params, DeviceBuffer& grads, size_t n_pad, cudaStream_t stream) {
	static const bool off = getenv("NGPB_DP_P2P") && atoi(getenv("NGPB_DP_P2P")) == 0;
	if (off || d.world > 8) return false;
	const NcclApi& n = NcclApi::get();
	for (DeviceBuffer* b : {&params, &grads}) {
		DeviceBuffer fresh;
		fresh.alloc_plain(n_pad * sizeof(__half), stream);
		NGPB_CUDA_CHECK(cudaMemcpyAsync(fresh.ptr, b->ptr, n_pad * sizeof(__half), cudaMemcpyDeviceToDevice, stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
		*b = std::move(fresh);
	}
	NGPB_CUDA_CHECK(cudaMalloc(&d.flags, kFlagBytes));
	NGPB_CUDA_CHECK(cudaMemset(d.flags, 0, kFlagBytes));
	struct Handles { cudaIpcMemHandle_t grads, params, flags; };
	Handles mine{};
	NGPB_CUDA_CHECK(cudaIpcGetMemHandle(&mine.grads, grads.ptr));
	NGPB_CUDA_CHECK(cudaIpcGetMemHandle(&mine.params, params.ptr));
	NGPB_CUDA_CHECK(cudaIpcGetMemHandle(&mine.flags, d.flags));
	DeviceBuffer all((size_t)d.world * sizeof(Handles), stream);
	NGPB_CUDA_CHECK(cudaMemcpyAsync(all.as<char>() + (size_t)d.rank * sizeof(Handles), &mine, sizeof(Handles), cudaMemcpyHostToDevice, stream));
	NGPB_NCCL_CHECK(n.AllGather(all.as<char>() + (size_t)d.rank * sizeof(Handles), all.ptr, sizeof(Handles), ncclUint8, d.aux, stream));
	std::vector<Handles> host(d.world);
	NGPB_CUDA_CHECK(cudaMemcpyAsync(host.data(), all.ptr, host.size() * sizeof(Handles), cudaMemcpyDeviceToHost, stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
	bool ok = true;
	for (int r = 0; r < d.world; ++r) {
		if (r == d.rank) { d.peer_grads[r] = grads.ptr; d.peer_params[r] = params.ptr; d.peer_flags[r] = d.flags; continue; }
		ok = ok && cudaIpcOpenMemHandle(&d.peer_grads[r], host[r].grads, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
		ok = ok && cudaIpcOpenMemHandle(&d.peer_params[r], host[r].params, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
		ok = ok && cudaIpcOpenMemHandle(&d.peer_flags[r], host[r].flags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
	}
	if (!ok) cudaGetLastError();
	if (ok) {
		NGPB_CUDA_CHECK(cudaMalloc(&d.peer_flags_dev, sizeof(d.peer_flags)));
		NGPB_CUDA_CHECK(cudaMemcpy(d.peer_flags_dev, d.peer_flags, sizeof(d.peer_flags), cudaMemcpyHostToDevice));
	}
	// every rank must take the same path: agree on the minimum
	DeviceBuffer vote(sizeof(uint32_t), stream);
	const uint32_t v = ok ? 1u : 0u;
	NGPB_CUDA_CHECK(cudaMemcpyAsync(vote.ptr, &v, sizeof(v), cudaMemcpyHostToDevice, stream));
	NGPB_NCCL_CHECK(n.AllReduce(vote.ptr, vote.ptr, 1, ncclUint32, ncclMin, d.aux, stream));
	uint32_t agreed = 0;
	NGPB_CUDA_CHECK(cudaMemcpyAsync(&agreed, vote.ptr, sizeof(agreed), cudaMemcpyDeviceToHost, stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
	(void)tb;
	return agreed == 1u;
}

void Testbed::init_data_parallel(const void* id_main, const void* id_aux, int rank, int world) {
	if (world < 1 || rank < 0 || rank >= world) throw std::runtime_error{"ngp-b200: init_data_parallel: bad rank / world"};
	if (kBatchSizeGranularity % (uint32_t)world != 0) throw std::runtime_error{"ngp-b200: the number of ranks must divide 256"};
	join_prefetch();
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	m_dist.reset();
	shard_rank = 0; shard_world = 1;
	if (world == 1) return;
	const NcclApi& n = NcclApi::get();
	auto d = std::make_unique<DistContext>();
	d->rank = rank; d->world = world;
	ncclUniqueId a, b;
	memcpy(&a, id_main, sizeof(a));
	memcpy(&b, id_aux, sizeof(b));
	NGPB_NCCL_CHECK(n.CommInitRank(&d->main, world, a, rank));
	NGPB_NCCL_CHECK(n.CommInitRank(&d->aux, world, b, rank));
	NGPB_CUDA_CHECK(cudaStreamCreateWithFlags(&d->aux_stream, cudaStreamNonBlocking));
	NGPB_CUDA_CHECK(cudaEventCreateWithFlags(&d->forward_done, cudaEventDisableTiming));
	if (m_n_params == 0) throw std::runtime_error{"ngp-b200: init_data_parallel needs a network (reload_network_from_json first)"};
	d->p2p = setup_p2p(*this, *d, m_params, m_gradients, next_multiple((uint32_t)m_n_params, kParamPadding), m_stream);
	m_dist = std::move(d);
	shard_rank = (uint32_t)rank; shard_world = (uint32_t)world;
	touch_generator_inputs();
}

// One data-parallel step; batch_size is the GLOBAL batch (batch_size / world compacted samples per rank).
void Testbed::train_dp(uint32_t batch_size) {
	DistContext& d = *m_dist;
	const NcclApi& n = NcclApi::get();
	train_forward(batch_size);

	// counters: summed beside the backward pass, on the host long before train_end asks for them
	NerfStepCounters* c = counters();
	NGPB_CUDA_CHECK(cudaEventRecord(d.forward_done, m_stream));
	NGPB_CUDA_CHECK(cudaStreamWaitEvent(d.aux_stream, d.forward_done, 0));
	static const bool counters_nccl = getenv("NGPB_DP_COUNTERS_NCCL") && atoi(getenv("NGPB_DP_COUNTERS_NCCL")) != 0;
	if (d.p2p && !counters_nccl) {
		counters_allreduce_p2p_kernel<<<1, 32, 0, d.aux_stream>>>(c, (uint32_t* const*)d.peer_flags_dev, d.rank, d.world, ++d.counter_epoch, (uint32_t*)d.flags + 32);
		NGPB_CUDA_CHECK(cudaGetLastError());
	} else {
		NGPB_NCCL_CHECK(n.GroupStart());
		NGPB_NCCL_CHECK(n.AllReduce(c, c, 3, ncclUint32, ncclSum, d.aux, d.aux_stream));
		NGPB_NCCL_CHECK(n.AllReduce(&c->loss_sum, &c->loss_sum, 1, ncclFloat32, ncclSum, d.aux, d.aux_stream));
		NGPB_NCCL_CHECK(n.GroupEnd());
	}
	NGPB_CUDA_CHECK(cudaMemcpyAsync(m_counters_host, c, sizeof(NerfStepCounters), cudaMemcpyDeviceToHost, d.aux_stream));
	NGPB_CUDA_CHECK(cudaEventRecord(m_counters_ready, d.aux_stream));
	m_counters_read = true;

	const uint32_t target = batch_size / shard_world;
	m_network->backward_strided(m_stream, target, m_coords_compacted.as<float>(), kCoordFloats, m_dloss.as<__half>(), 4, gradients(), params(), m_mlp_done, m_encoder_done, m_compact_src.as<uint32_t>());
	m_rng.advance(1ll << 32);

	// sharded optimizer
	const uint32_t n_pad = next_multiple((uint32_t)m_n_params, kParamPadding);
	const uint32_t shard = n_pad / shard_world;
	const uint32_t begin = shard_rank * shard, end = std::min(begin + shard, (uint32_t)m_n_params);
	if (d.p2p) {
		// reduce-scatter + Adam + all-gather in one kernel over NVLink peer memory (optimizer.cu adam_ema_step_p2p_kernel)
		P2PPeers peers{};
		peers.rank = d.rank; peers.world = d.world; peers.epoch = ++d.epoch;
		for (int r = 0; r < d.world; ++r) {
			peers.grads[r] = (const __half*)d.peer_grads[r];
			peers.params[r] = (__half*)d.peer_params[r];
			peers.flags[r] = (uint32_t*)d.peer_flags[r];
		}
		peers.error = (uint32_t*)d.flags + 32;
		prof_mark(m_stream, "dp.fused_reduce_scatter_adam_all_gather");
		m_optimizer->step_p2p(m_stream, kLossScaleFp16, params_fp32(), peers, std::min(begin, (uint32_t)m_n_params), end);
	} else {
		prof_mark(m_stream, "dp.reduce_scatter");
		NGPB_NCCL_CHECK(n.ReduceScatter(gradients(), gradients() + begin, shard, ncclFloat16, ncclSum, d.main, m_stream));
		prof_mark(m_stream, "optimizer");
		m_optimizer->step_range(m_stream, kLossScaleFp16, params_fp32(), params(), gradients(), std::min(begin, (uint32_t)m_n_params), end, true, true);
		prof_mark(m_stream, "dp.all_gather");
		NGPB_NCCL_CHECK(n.AllGather(params() + begin, params(), shard, ncclFloat16, d.main, m_stream));
	}
	if (m_optimizer->custom_weights()) {
		prof_mark(m_stream, "optimizer.ema_of_other_shards");
		m_optimizer->ema_range(m_stream, params(), 0, std::min(begin, (uint32_t)m_n_params));
		m_optimizer->ema_range(m_stream, params(), end, (uint32_t)m_n_params);
	}
	prof_mark(m_stream, "end");
	++m_training_step;
	// the aux stream's work must not be overtaken by the next step's reuse of the counter slot
	NGPB_CUDA_CHECK(cudaEventSynchronize(m_counters_ready));
	control_loop(batch_size);
	const bool next_step_preps = m_training_step % std::min(std::max(m_training_step / 16u, 1u), 16u) == 0;
	if (train_prefetch && !next_step_preps && !prof_enabled()) launch_prefetch(batch_size / shard_world);
}

} // namespace ngpb
