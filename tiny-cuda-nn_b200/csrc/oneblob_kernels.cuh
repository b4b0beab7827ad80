// OneBlob encoding (encodings/oneblob.h:47-160, quartic kernel common_device.h:905-920) in front of the fused MLP:
// BASELINE config #2 (data/config_oneblob.json, benchmarks/image/bench_ours.cu).  No parameters.
// Output is row-major fp16 [B][width] like every encoding here (the reference's AoS layout, its preferred one for OneBlob),
// columns >= n_in * n_bins are ONE (oneblob.h:131-133).
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

#include "grid_kernels.cuh"

namespace b200 {

// |u| >= 1.25: the polynomial is at least 2 % outside [0, 1] (1.023 at u = 1.25, monotone), far more than fp32 rounding can
// move it, so the reference's clamp returns exactly 0 or 1 there: the early-out is bit-identical and lets whole warps skip
// the polynomial (two of the three wrapped terms are saturated for almost every bin).  NaN takes the polynomial path.
__device__ __forceinline__ float oneblob_quartic_cdf(float x, float inv_radius) {
	const float u = x * inv_radius;
	if (fabsf(u) >= 1.25f) return u > 0.0f ? 1.0f : 0.0f;
	const float u2 = u * u;
	const float u4 = u2 * u2;
	return fmaxf(0.0f, fminf(1.0f, (15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f));
}
// The same for a converged warp: the vote makes the skip a real (uniform) branch - left to itself the compiler if-converts the
// early-out above into a select and evaluates the polynomial for all three wrapped terms of every column (ncu r01g: 28 fp32
// instructions per column, 12 with the vote).
__device__ __forceinline__ float oneblob_quartic_cdf_warp(float x, float inv_radius) {
	const float u = x * inv_radius;
	float r = u > 0.0f ? 1.0f : 0.0f;
	if (__any_sync(0xffffffffu, !(fabsf(u) >= 1.25f))) {
		const float u2 = u * u;
		const float u4 = u2 * u2;
		const float p = fmaxf(0.0f, fminf(1.0f, (15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f));
		if (!(fabsf(u) >= 1.25f)) r = p;
	}
	return r;
}
__device__ __forceinline__ float oneblob_quartic_pdf(float x, float inv_radius) { // d/dx of the cdf
	const float u = x * inv_radius;
	const float t = fmaxf(1.0f - u * u, 0.0f);
	return (15.0f / 16.0f) * t * t * inv_radius;
}

// Block = (n_out, ceil(128 / n_out)) threads like the reference's launch (oneblob.h:118-121), so that a column sits in the
// same lane as there: thread (j, y) = encoded column j of samples (blockIdx.x * ONEBLOB_ROUNDS + r) * blockDim.y + y.  Every
// thread evaluates the wrapped CDF at the LEFT edge of its bin and takes the right edge from the neighbouring lane with a
// sub-warp shuffle of width n_bins (oneblob.h:47-70) - kept exactly, including what the shuffle does for n_bins = 64
// (> warp size), because parity is against the reference's output, not its intent.  Unlike the reference a block walks
// ONEBLOB_ROUNDS samples per thread row (262144 one-sample blocks at batch 2^18 were launch-bound: 137 us).
// image_subs > 0: the output is written as tile images for the 128-neuron MLP (grid_kernels.cuh wide_image_chunk).
constexpr uint32_t ONEBLOB_ROUNDS = 8;
__global__ void oneblob_encode_kernel(uint32_t B, uint32_t n_in, uint32_t log2_bins, uint32_t n_out, uint32_t width,
	const float* __restrict__ x, __half* __restrict__ out, uint32_t image_subs) {
	const uint32_t j = threadIdx.x;
	const uint32_t n_bins = 1u << log2_bins;
	const uint32_t bin = j & (n_bins - 1), dim = j >> log2_bins;
	const float inv_radius = (float)n_bins;
	const float left_boundary = scalbnf((float)bin, -(int)log2_bins);
	const uint32_t i0 = blockIdx.x * ONEBLOB_ROUNDS * blockDim.y + threadIdx.y;
	float xv[ONEBLOB_ROUNDS];
#pragma unroll
	for (uint32_t r = 0; r < ONEBLOB_ROUNDS; ++r) {
		const uint32_t i = i0 + r * blockDim.y;
		xv[r] = i < B ? __ldg(x + (size_t)i * n_in + dim) : 0.0f;
	}
#pragma unroll
	for (uint32_t r = 0; r < ONEBLOB_ROUNDS; ++r) {
		const uint32_t i = i0 + r * blockDim.y;
		const float d = left_boundary - xv[r];
		const float left_cdf = oneblob_quartic_cdf_warp(d, inv_radius) + oneblob_quartic_cdf_warp(d - 1.0f, inv_radius) + oneblob_quartic_cdf_warp(d + 1.0f, inv_radius);
		float right_cdf = __shfl_sync(0xffffffffu, left_cdf, bin + 1, n_bins); // whole warps stay in the loop: no lane has exited
		if (bin == n_bins - 1) right_cdf += 1.0f; // wrap-around: the last bin's right edge is the first bin's left edge + 1
		if (i >= B) continue;
		uint8_t* o = reinterpret_cast<uint8_t*>(out);
		// byte offset of column `col` of sample i: a row, or (tile images) wide_image_chunk with the tile-local part in 32 bits
		auto at = [&](uint32_t col) -> __half* {
			if (!image_subs) return reinterpret_cast<__half*>(o + ((size_t)i * width + col) * 2);
			const uint32_t r = i & 127u, local = (col >> 6) * SUB_BYTES + r * 128u + ((((col >> 3) & 7u) ^ (r & 7u)) << 4) + (col & 7u) * 2u;
			return reinterpret_cast<__half*>(o + (size_t)(i >> 7) * (image_subs * SUB_BYTES) + local);
		};
		*at(j) = __float2half(right_cdf - left_cdf);
		if (j < width - n_out) *at(n_out + j) = __float2half(1.0f);
	}
}

// dL/dx (oneblob.h:105-141): sum over the bins of dL/dy * (pdf at the left edge - pdf at the right edge), wrapped
__global__ void __launch_bounds__(256) oneblob_backward_input_kernel(uint32_t B, uint32_t n_in, uint32_t log2_bins, uint32_t width,
	const float* __restrict__ x, const __half* __restrict__ dL_dy, float* __restrict__ dL_dx) {
	const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= B * n_in) return;
	const uint32_t i = q / n_in, j = q % n_in;
	const uint32_t n_bins = 1u << log2_bins;
	const float xv = x[q], inv_radius = (float)n_bins;
	auto wrapped_pdf = [&](float boundary) {
		return oneblob_quartic_pdf(boundary - xv, inv_radius) + oneblob_quartic_pdf(boundary - xv - 1.0f, inv_radius) + oneblob_quartic_pdf(boundary - xv + 1.0f, inv_radius);
	};
	const __half* g = dL_dy + (size_t)i * width + (size_t)j * n_bins;
	float left = wrapped_pdf(0.0f), result = 0.0f;
	for (uint32_t k = 0; k < n_bins; ++k) {
		const float right = wrapped_pdf(scalbnf((float)(k + 1), -(int)log2_bins));
		result += __half2float(g[k]) * (left - right);
		left = right;
	}
	dL_dx[q] = result;
}

} // namespace b200
