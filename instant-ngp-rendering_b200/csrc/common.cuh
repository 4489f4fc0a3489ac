// Shared host/device helpers for the ngp-b200 kernels (sm_100a only).
//
// Nothing in here is a port of the reference's runtime helpers
// (tcnn/include/tiny-cuda-nn/common_host.h, gpu_memory.h): this library takes raw
// device pointers from the caller and a stream per call, as the C ABI in
// include/ngp_b200.h describes, so the helpers are limited to error plumbing, a PCG32
// generator (published algorithm, M.E. O'Neill 2014; the reference vendors Wenzel
// Jakob's pcg32.h under tcnn/dependencies/pcg32/) and small math.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

#define NGPB_HD __host__ __device__ __forceinline__
#define NGPB_D __device__ __forceinline__

namespace ngpb {

// ---- error plumbing -------------------------------------------------------------------
// The C ABI returns an int status and keeps the message in a thread-local string; the
// C++ layer throws std::runtime_error with the same texts the reference uses
// (tcnn common_host.h CHECK_THROW / CUDA_CHECK_THROW) so that a maintainer sees familiar
// messages.
void set_last_error(const std::string& msg);
const char* last_error();

#define NGPB_CHECK_THROW(x) \
	do { if (!(x)) throw std::runtime_error{std::string{__FILE__ ":"} + std::to_string(__LINE__) + " check failed " #x}; } while (0)

#define NGPB_CUDA_CHECK(x) \
	do { cudaError_t _e = (x); if (_e != cudaSuccess) throw std::runtime_error{std::string{__FILE__ ":"} + std::to_string(__LINE__) + " " #x " failed: " + cudaGetErrorString(_e)}; } while (0)

inline uint32_t div_round_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline uint64_t div_round_up64(uint64_t a, uint64_t b) { return (a + b - 1) / b; }
inline uint32_t next_multiple(uint32_t a, uint32_t b) { return div_round_up(a, b) * b; }

// Number of SMs of the current device (148 on B200); cached per process.
int sm_count();

// ---- PCG32 ------------------------------------------------------------------------------
// XSH-RR 64/32 with the canonical default state/stream, advance() in O(log n). The
// reference seeds every sampler from this generator (src/testbed_nerf.cu:715
// `rng.advance(i * N_MAX_RANDOM_SAMPLES_PER_RAY)`), so sample positions are only
// bit-identical if this matches the published generator exactly.
struct Pcg32 {
	static constexpr uint64_t kDefaultState = 0x853c49e6748fea9bULL;
	static constexpr uint64_t kDefaultStream = 0xda3e39cb94b95bdbULL;
	static constexpr uint64_t kMult = 0x5851f42d4c957f2dULL;

	uint64_t state;
	uint64_t inc;

	NGPB_HD Pcg32() : state(kDefaultState), inc(kDefaultStream) {}
	NGPB_HD explicit Pcg32(uint64_t initstate, uint64_t initseq = 1u) { seed(initstate, initseq); }

	NGPB_HD void seed(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u;
		inc = (initseq << 1u) | 1u;
		next_uint();
		state += initstate;
		next_uint();
	}

	NGPB_HD uint32_t next_uint() {
		uint64_t old = state;
		state = old * kMult + inc;
		uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		uint32_t rot = (uint32_t)(old >> 59u);
		return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
	}

	// Uniform float in [0,1): top 23 random bits placed in the mantissa of 1.x, minus 1.
	NGPB_HD float next_float() {
		union { uint32_t u; float f; } x;
		x.u = (next_uint() >> 9) | 0x3f800000u;
		return x.f - 1.0f;
	}

	NGPB_HD void advance(int64_t delta_) {
		uint64_t cur_mult = kMult, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
		uint64_t delta = (uint64_t)delta_;
		while (delta > 0) {
			if (delta & 1) {
				acc_mult *= cur_mult;
				acc_plus = acc_plus * cur_mult + cur_plus;
			}
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta /= 2;
		}
		state = acc_mult * state + acc_plus;
	}
};

// ---- tiny vector types (device side only needs float3/float2-like PODs) ----------------
struct Vec3 { float x, y, z; };
struct Vec2 { float x, y; };

} // namespace ngpb
