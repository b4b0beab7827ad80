// OneBlob encoding (encodings/oneblob.h:47-160, quartic kernel common_device.h:905-920) in front of the fused MLP:
// BASELINE config #2 (data/config_oneblob.json, benchmarks/image/bench_ours.cu).  No parameters.
// Output is row-major fp16 [B][width] like every encoding here (the reference's AoS layout, its preferred one for OneBlob),
// columns >= n_in * n_bins are ONE (oneblob.h:131-133).
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

__device__ __forceinline__ float oneblob_quartic_cdf(float x, float inv_radius) {
	const float u = x * inv_radius;
	const float u2 = u * u;
	const float u4 = u2 * u2;
	return fmaxf(0.0f, fminf(1.0f, (15.0f / 16.0f) * u * (1.0f - (2.0f / 3.0f) * u2 + (1.0f / 5.0f) * u4) + 0.5f));
}
__device__ __forceinline__ float oneblob_quartic_pdf(float x, float inv_radius) { // d/dx of the cdf
	const float u = x * inv_radius;
	const float t = fmaxf(1.0f - u * u, 0.0f);
	return (15.0f / 16.0f) * t * t * inv_radius;
}

// Block = (n_out, ceil(128 / n_out)) threads like the reference's launch (oneblob.h:118-121): thread (j, y) = encoded column
// j of sample blockIdx.x * blockDim.y + y.  Every thread evaluates the wrapped CDF at the LEFT edge of its bin and takes the
// right edge from the neighbouring lane with a sub-warp shuffle of width n_bins (oneblob.h:47-70) - kept exactly, including
// what the shuffle does for n_bins = 64 (> warp size), because parity is against the reference's output, not its intent.
__global__ void oneblob_encode_kernel(uint32_t B, uint32_t n_in, uint32_t log2_bins, uint32_t n_out, uint32_t width,
	const float* __restrict__ x, __half* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.y + threadIdx.y;
	const uint32_t j = threadIdx.x;
	if (i >= B) return;
	const uint32_t n_bins = 1u << log2_bins;
	const uint32_t bin = j & (n_bins - 1);
	const float xv = x[(size_t)i * n_in + (j >> log2_bins)];
	const float inv_radius = (float)n_bins;
	const float left_boundary = scalbnf((float)bin, -(int)log2_bins);
	const float left_cdf = oneblob_quartic_cdf(left_boundary - xv, inv_radius) + oneblob_quartic_cdf(left_boundary - xv - 1.0f, inv_radius)
		+ oneblob_quartic_cdf(left_boundary - xv + 1.0f, inv_radius);
	float right_cdf = __shfl_sync(0xffffffffu, left_cdf, bin + 1, n_bins);
	if (bin == n_bins - 1) right_cdf += 1.0f; // wrap-around: the last bin's right edge is the first bin's left edge + 1
	out[(size_t)i * width + j] = __float2half(right_cdf - left_cdf);
	if (j < width - n_out) out[(size_t)i * width + n_out + j] = __float2half(1.0f);
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
