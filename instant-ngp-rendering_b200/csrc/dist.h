// NCCL for the data-parallel training step (no reference counterpart: the reference trains on one GPU,
// README.md:250-252; its only inter-GPU traffic is cudaMemcpyPeerAsync replication for VR rendering,
// src/testbed.cu:5008-5048).
//
// The library does not link NCCL: it binds the process's libnccl.so.2 at run time (the one torch has already
// loaded when the host side is Python, the system one for a C++ host), so a C++ caller linking libngp_b200.so
// gets the sharded step without any Python in the process. One process per GPU; the caller distributes the two
// unique ids (ngpb_nccl_get_unique_id on rank 0) over whatever bootstrap it has - torch.distributed's store,
// MPI, a file.
#pragma once

#include "common.cuh"

#include <nccl.h>

namespace ngpb {

struct NcclApi {
	ncclResult_t (*GetUniqueId)(ncclUniqueId*);
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
	ncclResult_t (*CommDestroy)(ncclComm_t);
	ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
	ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
	ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
	ncclResult_t (*GroupStart)();
	ncclResult_t (*GroupEnd)();
	const char* (*GetErrorString)(ncclResult_t);
	static const NcclApi& get();   // throws when no libnccl.so.2 can be bound
};

#define NGPB_NCCL_CHECK(x) \
	do { ncclResult_t _r = (x); if (_r != ncclSuccess) throw std::runtime_error{std::string{__FILE__ ":"} + std::to_string(__LINE__) + " " #x " failed: " + ::ngpb::NcclApi::get().GetErrorString(_r)}; } while (0)

// Two communicators over the same ranks: `main` carries the gradient reduce-scatter and the parameter all-gather on
// the testbed's stream; `aux` carries the 16-byte counter all-reduce on its own stream, beside the backward pass.
struct DistContext {
	ncclComm_t main = nullptr, aux = nullptr;
	int rank = 0, world = 1;
	cudaStream_t aux_stream = nullptr;
	cudaEvent_t forward_done = nullptr;
	// peer-to-peer path (one node): every rank's gradients / fp16 weights / flags mapped here (dist.cu)
	bool p2p = false;
	void* peer_grads[8] = {};
	void* peer_params[8] = {};
	void* peer_flags[8] = {};
	void* flags = nullptr;       // this rank's uint32 [2][8] + error word
	uint32_t epoch = 0;
	uint32_t counter_epoch = 0;   // step stamp of the counter mailbox (dist.cu)
	void* peer_flags_dev = nullptr; // device copy of peer_flags[] for kernels that index it by lane
	~DistContext();
};

} // namespace ngpb
