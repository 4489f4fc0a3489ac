// PCG32 (XSH-RR 64/32, O'Neill 2014) as the reference uses it through the vendored
// tcnn/dependencies/pcg32/pcg32.h: default increment for `pcg32(seed)` is stream 1,
// next_float() puts the top 23 bits into [1,2) and subtracts 1, advance() is Brown's
// O(log n) jump. TEST INFRASTRUCTURE ONLY (see oracle.h).
#pragma once
#include <cstdint>

struct OraclePcg32 {
	uint64_t state, inc;
	OraclePcg32() : state(0x853c49e6748fea9bULL), inc(0xda3e39cb94b95bdbULL) {}
	explicit OraclePcg32(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u;
		inc = (initseq << 1u) | 1u;
		next_uint();
		state += initstate;
		next_uint();
	}
	uint32_t next_uint() {
		const uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		const uint32_t rot = (uint32_t)(old >> 59u);
		return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
	}
	uint32_t next_uint(uint32_t bound) {
		const uint32_t threshold = (~bound + 1u) % bound;
		for (;;) {
			const uint32_t r = next_uint();
			if (r >= threshold) return r % bound;
		}
	}
	float next_float() {
		union { uint32_t u; float f; } x;
		x.u = (next_uint() >> 9) | 0x3f800000u;
		return x.f - 1.0f;
	}
	void advance(int64_t delta_) {
		uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
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
