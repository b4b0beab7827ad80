// Development micro-benchmark (runs on a GPU box): the L2 ceilings the two L2-bound kernels of the hot path are measured
// against ("B200 L2 peak is not in MEASURED_PEAKS.json; measure it once and commit the number").
//   red16 / red8 / red4 : random red.global.add.noftz.{v4,v2,}f16x2 into a table (the fused scatter issues the 16-byte form)
//   ld16 / ld4          : random ld.global.nc.{v4.b32,b32} gathers from the same table (the gather kernel issues both)
// at full occupancy, for table sizes inside and outside the 126 MB L2.  One JSON object per line on stdout.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tests/dev/l2_peaks tests/dev/l2_peaks.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { // cheap integer hash: addresses are independent of neighbours' (no coalescing)
	x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
	return x;
}

template <int BYTES>
__global__ void __launch_bounds__(256) red_kernel(uint8_t* table, uint32_t slots_mask, uint32_t iters) {
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t one = 0x00010001u; // two fp16 subnormals: the adds never overflow
	for (uint32_t i = 0; i < iters; ++i) {
		uint8_t* p = table + (size_t)(mix(t * 2654435761u + i * 40503u) & slots_mask) * BYTES;
		if (BYTES == 16) asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %1, %1, %1};" ::"l"(p), "r"(one) : "memory");
		else if (BYTES == 8) asm volatile("red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1, %1};" ::"l"(p), "r"(one) : "memory");
		else asm volatile("red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;" ::"l"(p), "r"(one) : "memory");
	}
}

template <int BYTES>
__global__ void __launch_bounds__(256) ld_kernel(const uint8_t* table, uint32_t slots_mask, uint32_t iters, uint32_t* sink) {
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t acc = 0;
#pragma unroll 4
	for (uint32_t i = 0; i < iters; ++i) {
		const uint8_t* p = table + (size_t)(mix(t * 2654435761u + i * 40503u) & slots_mask) * BYTES;
		if (BYTES == 16) {
			uint4 v;
			asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
			acc += v.x ^ v.y ^ v.z ^ v.w;
		} else {
			uint32_t v;
			asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
			acc += v;
		}
	}
	if (acc == 0x12345678u) sink[0] = acc;
}

// Do gathers and reductions share one L2 budget?  Even warps gather, odd warps reduce, same op count as the pure kernels in total.
__global__ void __launch_bounds__(256) mix_kernel(uint8_t* red_table, const uint8_t* ld_table, uint32_t slots_mask, uint32_t iters, uint32_t* sink) {
	const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t one = 0x00010001u;
	uint32_t acc = 0;
	if ((threadIdx.x >> 5) & 1) {
		for (uint32_t i = 0; i < iters; ++i) {
			uint8_t* p = red_table + (size_t)(mix(t * 2654435761u + i * 40503u) & slots_mask) * 16;
			asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %1, %1, %1};" ::"l"(p), "r"(one) : "memory");
		}
	} else {
#pragma unroll 4
		for (uint32_t i = 0; i < iters; ++i) {
			const uint8_t* p = ld_table + (size_t)(mix(t * 2654435761u + i * 40503u) & slots_mask) * 16;
			uint4 v;
			asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
			acc += v.x ^ v.y ^ v.z ^ v.w;
		}
	}
	if (acc == 0x12345678u) sink[0] = acc;
}

// A streaming read-modify-write over a large buffer (the optimizer's access pattern) on half of the blocks while the other half reduce
// into an L2-resident table: does an HBM stream slow the reductions down?
__global__ void __launch_bounds__(256) stream_red_kernel(uint8_t* red_table, uint32_t slots_mask, uint32_t iters, float4* stream, size_t n_vec, int mode) {
	const uint32_t one = 0x00010001u;
	const bool streamer = (blockIdx.x & 1) != 0;
	if (streamer && (mode & 2)) {
		const size_t n_thr = (size_t)(gridDim.x / 2) * blockDim.x, me = (size_t)(blockIdx.x / 2) * blockDim.x + threadIdx.x;
		for (size_t k = me; k < n_vec; k += n_thr) { float4 v = stream[k]; v.x += 1.f; stream[k] = v; }
	} else if (!streamer && (mode & 1)) {
		const uint32_t t = (blockIdx.x / 2) * blockDim.x + threadIdx.x;
		for (uint32_t i = 0; i < iters; ++i) {
			uint8_t* p = red_table + (size_t)(mix(t * 2654435761u + i * 40503u) & slots_mask) * 16;
			asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %1, %1, %1};" ::"l"(p), "r"(one) : "memory");
		}
	}
}

template <typename L>
static float time_ms(L launch, int reps) {
	cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
	launch(); CK(cudaDeviceSynchronize());
	float best = 1e30f;
	for (int r = 0; r < reps; ++r) {
		CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
		float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = ms < best ? ms : best;
	}
	return best;
}

int main() {
	int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
	const size_t max_bytes = 512ull << 20;
	uint8_t* table; CK(cudaMalloc(&table, max_bytes)); CK(cudaMemset(table, 0, max_bytes));
	uint32_t* sink; CK(cudaMalloc(&sink, 4));
	const uint32_t blocks = sms * 8, threads = 256, iters = 256; // 8 x 256 = 2048 threads per SM: full occupancy
	const double ops = (double)blocks * threads * iters;
	const size_t sizes[] = {2ull << 20, 16ull << 20, 32ull << 20, 64ull << 20, 256ull << 20};
	for (size_t bytes : sizes) {
		auto line = [&](const char* op, int width, float ms) {
			printf("{\"op\": \"%s\", \"bytes_per_op\": %d, \"table_mib\": %zu, \"ops\": %.0f, \"ms\": %.4f, \"gops_per_s\": %.2f, \"payload_gb_per_s\": %.1f, \"sector_gb_per_s\": %.1f, \"sms\": %d}\n",
				op, width, bytes >> 20, ops, ms, ops / ms * 1e-6, ops * width / ms * 1e-6, ops * 32 / ms * 1e-6, sms);
			fflush(stdout);
		};
		CK(cudaMemset(table, 0, bytes));
		line("red", 16, time_ms([&] { red_kernel<16><<<blocks, threads>>>(table, (uint32_t)(bytes / 16 - 1), iters); }, 5));
		CK(cudaMemset(table, 0, bytes));
		line("red", 8, time_ms([&] { red_kernel<8><<<blocks, threads>>>(table, (uint32_t)(bytes / 8 - 1), iters); }, 5));
		CK(cudaMemset(table, 0, bytes));
		line("red", 4, time_ms([&] { red_kernel<4><<<blocks, threads>>>(table, (uint32_t)(bytes / 4 - 1), iters); }, 5));
		line("ld", 16, time_ms([&] { ld_kernel<16><<<blocks, threads>>>(table, (uint32_t)(bytes / 16 - 1), iters, sink); }, 5));
		line("ld", 4, time_ms([&] { ld_kernel<4><<<blocks, threads>>>(table, (uint32_t)(bytes / 4 - 1), iters, sink); }, 5));
	}
	{ // mixed kernels on a 32 MiB table (reductions) + a second 32 MiB table (gathers)
		const size_t bytes = 32ull << 20;
		const uint32_t mask = (uint32_t)(bytes / 16 - 1);
		CK(cudaMemset(table, 0, 2 * bytes));
		const float t_mix = time_ms([&] { mix_kernel<<<blocks, threads>>>(table, table + bytes, mask, iters, sink); }, 5);
		printf("{\"op\": \"mix_ld16_red16\", \"table_mib\": 32, \"ops_each\": %.0f, \"ms\": %.4f, \"note\": \"half the warps gather, half reduce; compare with the pure kernels at half their op count\"}\n", ops / 2, t_mix);
		const size_t n_vec = (384ull << 20) / 16; // 384 MiB read + written: the optimizer's traffic of one step
		float4* stream = (float4*)(table + 2 * bytes);
		for (int mode = 1; mode <= 3; ++mode) {
			const float t = time_ms([&] { stream_red_kernel<<<blocks, threads>>>(table, mask, iters, stream, n_vec, mode); }, 5);
			printf("{\"op\": \"stream_red\", \"mode\": \"%s\", \"red_ops\": %.0f, \"stream_mib_rw\": 384, \"ms\": %.4f}\n", mode == 1 ? "red only (half the blocks)" : mode == 2 ? "stream only (half the blocks)" : "both", ops / 2, t);
		}
	}
	CK(cudaGetLastError());
	return 0;
}
