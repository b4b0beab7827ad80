/* default_rng_t (= pcg32, include/tiny-cuda-nn/common_device.h:331) and generate_random_uniform (random.h:58-76): what hosts of the
 * reference draw their training batches with.  The generator is the public PCG-XSH-RR 64/32 algorithm with the reference's
 * seeding (dependencies/pcg32/pcg32.h:54-60) and O(log n) jump-ahead; the device side is tcnn_b200_generate_random_uniform,
 * which reproduces the reference's element -> draw mapping, so a batch drawn here equals the batch the reference draws. */
#pragma once
#include "common.h"

namespace tcnn {

struct pcg32 {
	uint64_t state, inc;
	static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
	pcg32() : state(0x853c49e6748fea9bULL), inc(0xda3e39cb94b95bdbULL) {}
	pcg32(uint64_t initstate, uint64_t initseq = 1u) { seed(initstate, initseq); }
	void seed(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u; inc = (initseq << 1u) | 1u;
		next_uint();
		state += initstate;
		next_uint();
	}
	uint32_t next_uint() {
		const uint64_t old = state;
		state = old * MULT + inc;
		const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
		return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
	}
	float next_float() { union { uint32_t u; float f; } x; x.u = (next_uint() >> 9) | 0x3f800000u; return x.f - 1.0f; }
	void advance(int64_t delta_) {
		uint64_t cur_mult = MULT, cur_plus = inc, acc_mult = 1u, acc_plus = 0u, delta = (uint64_t)delta_;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta /= 2;
		}
		state = acc_mult * state + acc_plus;
	}
};
using default_rng_t = pcg32;

template <typename T, typename RNG>
void generate_random_uniform(cudaStream_t stream, RNG& rng, size_t n_elements, T* out, const T lower = (T)0.0, const T upper = (T)1.0) {
	static_assert(std::is_same<T, float>::value, "generate_random_uniform: float output");
	TCNN_B200_CHECK_THROW(tcnn_b200_generate_random_uniform(stream, rng.state, rng.inc, n_elements, out, lower, upper));
	rng.advance((int64_t)n_elements);
}
template <typename T, typename RNG>
void generate_random_uniform(RNG& rng, size_t n_elements, T* out, const T lower = (T)0.0, const T upper = (T)1.0) {
	generate_random_uniform(nullptr, rng, n_elements, out, lower, upper);
}

} // namespace tcnn
