// Activations of the fused MLP (common_device.h:100-304 of the reference), on packed fp16 pairs.
//
// The reference applies them to fp16 fragments: the argument is the fp16 pre-activation, the function itself is evaluated in
// fp32 (expf, tanhf, ...) and rounded back to fp16 (warp_activation<__half>, common_device.h:102-160).  The derivative used
// by the backward pass is taken from the stored POST-activation (warp_activation_backward, common_device.h:240-293) and
// multiplied onto the fp16 gradient with a half multiply.  Both are restated here; ReLU keeps its own two-instruction fast
// path in the kernels (h2_relu / h2_mask_mul in ptx.cuh), everything else goes through these out-of-line functions so that
// the hot kernels carry one predicated call instead of eight inlined transcendental sequences.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

enum Activation : int {
	ACT_NONE = 0, ACT_RELU = 1, ACT_LEAKY_RELU = 2, ACT_EXPONENTIAL = 3, ACT_SIGMOID = 4, ACT_SQUAREPLUS = 5, ACT_SOFTPLUS = 6, ACT_TANH = 7,
};
constexpr float ACT_K = 10.0f; // K_ACT, common_device.h:100

__device__ __forceinline__ __half act_forward_one(__half xh, int act) {
	const float x = __half2float(xh);
	switch (act) {
		case ACT_RELU: return __hmax(xh, __float2half(0.0f));
		case ACT_LEAKY_RELU: return __hmul(xh, __float2half(__hgt(xh, __float2half(0.0f)) ? 1.0f : 0.01f));
		case ACT_EXPONENTIAL: return __float2half(expf(x));
		case ACT_SIGMOID: return __float2half(1.0f / (1.0f + expf(-x)));
		case ACT_SQUAREPLUS: { const float t = x * ACT_K; return __float2half(0.5f * (t + sqrtf(t * t + 4)) / ACT_K); }
		case ACT_SOFTPLUS: return __float2half(logf(expf(x * ACT_K) + 1.0f) / ACT_K);
		case ACT_TANH: return __float2half(tanhf(x));
		default: return xh;
	}
}
// d(activation)/d(pre-activation) expressed through the post-activation y, as an fp16 factor
__device__ __forceinline__ __half act_derivative_one(__half yh, int act) {
	const float y = __half2float(yh);
	switch (act) {
		case ACT_RELU: return __float2half(__hgt(yh, __float2half(0.0f)) ? 1.0f : 0.0f);
		case ACT_LEAKY_RELU: return __float2half(__hgt(yh, __float2half(0.0f)) ? 1.0f : 0.01f);
		case ACT_EXPONENTIAL: return yh;
		case ACT_SIGMOID: return __hmul(yh, __float2half(1.0f - y));
		case ACT_SQUAREPLUS: { const float t = y * ACT_K; return __float2half(t * t / (t * t + 1)); }
		case ACT_SOFTPLUS: return __float2half(1.0f - expf(-y * ACT_K));
		case ACT_TANH: return __float2half(1.0f - y * y);
		default: return __float2half(1.0f);
	}
}

__device__ __noinline__ uint32_t act_forward_h2(uint32_t pre, int act) {
	__half2 p = *reinterpret_cast<__half2*>(&pre);
	__half2 r = __halves2half2(act_forward_one(__low2half(p), act), act_forward_one(__high2half(p), act));
	return *reinterpret_cast<uint32_t*>(&r);
}
// gradient (fp16 pair) * derivative taken from the post-activation pair
__device__ __noinline__ uint32_t act_backward_h2(uint32_t grad, uint32_t post, int act) {
	__half2 g = *reinterpret_cast<__half2*>(&grad), y = *reinterpret_cast<__half2*>(&post);
	__half2 d = __halves2half2(act_derivative_one(__low2half(y), act), act_derivative_one(__high2half(y), act));
	__half2 r = __hmul2(g, d);
	return *reinterpret_cast<uint32_t*>(&r);
}

} // namespace b200
