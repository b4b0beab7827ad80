// Development harness (GPU box): stage-by-stage clock stamps of the fused MLP training kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DMLP_STAGE_CLOCKS -o mlp_timing tests/dev/mlp_timing.cu
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "../../tiny-cuda-nn_b200/csrc/mlp_kernels.cuh"
using namespace b200;
int main() {
	const uint32_t B = 1 << 18, E = 32, H = 2, P = 64 * E + 64 * 64 + 16 * 64;
	__half *in, *params, *dx; float *target, *dw, *lp; long long* clk;
	cudaMalloc(&in, (size_t)B * E * 2); cudaMalloc(&params, P * 2); cudaMalloc(&dx, (size_t)B * E * 2);
	cudaMalloc(&target, (size_t)B * 3 * 4); cudaMalloc(&dw, (size_t)600 * P * 4); cudaMalloc(&lp, 600 * 4); cudaMalloc(&clk, 256 * 8);
	std::vector<__half> h((size_t)B * E); for (auto& v : h) v = __float2half((rand() % 2001 - 1000) / 1000.f);
	cudaMemcpy(in, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
	std::vector<__half> hp(P); for (auto& v : hp) v = __float2half((rand() % 2001 - 1000) / 5000.f);
	cudaMemcpy(params, hp.data(), P * 2, cudaMemcpyHostToDevice);
	cudaMemset(target, 0, (size_t)B * 3 * 4); cudaMemset(clk, 0, 256 * 8);
	MlpArgs a{}; a.n_tiles = B / 128; a.in_width = E; a.n_hidden = H; a.n_out = 3; a.input = in; a.params = params; a.target = target;
	a.dL_dinput = dx; a.dW_partials = dw; a.loss_partials = lp; a.n_mlp_params = P; a.loss_type = 0; a.loss_scale = 128.f; a.n_total = B * 3.f; a.debug_clock = clk;
	GridParams gp{};
	const uint32_t smem = mlp_smem_bytes(H, E, true);
	cudaFuncSetAttribute(mlp_kernel<MODE_TRAIN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	for (int grid : {296, 148}) {
		cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
		for (int i = 0; i < 3; ++i) mlp_kernel<MODE_TRAIN, 2><<<grid, 128, smem>>>(a, gp);
		cudaEventRecord(e0);
		for (int i = 0; i < 10; ++i) mlp_kernel<MODE_TRAIN, 2><<<grid, 128, smem>>>(a, gp);
		cudaEventRecord(e1); cudaEventSynchronize(e1);
		float ms; cudaEventElapsedTime(&ms, e0, e1);
		printf("grid %d: %.1f us per launch (%s)\n", grid, ms * 100.f, cudaGetErrorString(cudaGetLastError()));
		long long c[256]; cudaMemcpy(c, clk, sizeof(c), cudaMemcpyDeviceToHost);
		printf("stage deltas (cycles), first 40 stamps:");
		for (int i = 1; i < 40 && c[i]; ++i) printf(" %lld", c[i] - c[i - 1]);
		printf("\n");
	}
	return 0;
}
