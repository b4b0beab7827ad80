// Fully fused MLP, width 128, for sm_100a on tcgen05 tensor cores (BASELINE configs #2, #4, #5).
//
// Computes the reference's FullyFusedMLP<half, 128> (src/fully_fused_mlp.cu:47-259, 499-557, 735-836): ReLU hidden layers,
// linear output padded to 16, fp16 storage, plus the L2 / RelativeL2 loss (losses/l2.h:40-75, relative_l2.h:40-75), the
// weight gradients of the reference's split-K GEMMs (cutlass_matmul.h:437-479) and dL/d(input) (cutlass_matmul.h:351-408).
//
// At width 128 a tile's activations of all layers (32 KB each) plus all weight matrices (32 KB each) no longer fit one
// SM's shared memory together, and the weight-gradient accumulators (128 TMEM columns per matrix) no longer fit TMEM, so the
// training pass is three kernels with the activations / pre-activation gradients kept in HBM as the reference does:
//   mlp_wide_forward_kernel   all weights resident in shared memory; one 128-sample tile per warpgroup, two warpgroups per
//                             CTA (when shared memory allows) on independent tiles so that one's epilogue hides under the
//                             other's MMAs; the activation tile is overwritten in place layer by layer; loss +
//                             dL/d(output) in the epilogue
//   mlp_wide_backward_kernel  same structure for the dgrad chain; the ReLU mask is re-read from the stored activations
//   mlp_wide_dw_kernel        split-K over the batch: CTA (split, matrix) accumulates G^T * ACT in TMEM over its tiles,
//                             operands streamed through a 3-stage TMA pipeline
// Operand tiles are [128 rows][128 halfs] stored as TWO 128-byte-swizzled sub-tiles of [128][64 halfs] (16 KB apart): as a
// K-major operand K-step s starts at sub-tile s/4, byte 32 (s%4); as an MN-major operand (weight tiles in the dgrad MMAs,
// both operands of the weight-gradient MMAs) the sub-tiles are the two 64-wide M/N atoms, LBO = 16 KB.
//
// TILE IMAGES: every per-sample matrix these kernels exchange through HBM (the encoded input, the hidden activations, the
// pre-activation gradients) is stored as the byte image of its shared-memory tile - tile t of layer k at
// ((k * n_tiles + t) * 32 KB) - not as a row-major matrix.  A tile then moves with ONE bulk copy through the TMA unit
// (cp.async.bulk, SASS UBLKCP) in either direction, fully coalesced, asynchronous to the issuing thread, with no
// register staging and no swizzle arithmetic; the encoders write the image directly (16-byte chunk (row, c) of sub-tile s
// at s * 16 KB + row * 128 + ((c ^ (row & 7)) << 4)).
#pragma once
#include "mlp_kernels.cuh"

namespace b200 {

constexpr int WIDE_W = 128;
static_assert(SUB_BYTES == 16384, "tile image sub-tile = [128 rows][64 halfs]");
constexpr uint32_t WIDE_TILE_BYTES = 2 * SUB_BYTES;
constexpr int WIDE_MAX_HIDDEN = 8;

struct WideArgs {
	uint32_t n_tiles;              // B / 128
	uint32_t in_width;             // E_pad: multiple of 16, <= 128
	uint32_t n_hidden;             // H >= 1
	uint32_t n_out;                // real outputs (<= 16)
	uint64_t batch;                // B = row count of every per-layer buffer
	const __half* input;           // tile images of the encoded batch: n_tiles x in_sub x 16 KB (see wide_image_chunk)
	const __half* params;          // [128 x E], (H-1) x [128 x 128], [16 x 128] row-major (fully_fused_mlp.cu:654-671)
	__half* output;                // [B][16] nullable
	float* output_f32;             // [B][n_out] nullable (inference: fused trim_and_cast, object.h:174-175)
	const float* target;           // [B][n_out]
	const __half* dL_doutput;      // backward / dW: [B][16]
	__half* dL_doutput_out;        // forward with loss: [B][16] written
	__half* hidden;                // [H][n_tiles] tile images (32 KB) of the post-activation outputs of matrices 0..H-1
	__half* gtmp;                  // [H][n_tiles] tile images of the gradients w.r.t. the pre-activations of matrices 0..H-1
	__half* dL_dinput;             // [B][E] nullable
	float* dW_partials;            // [n_splits][n_mlp_params]
	float* loss_partials;          // one per forward warpgroup
	uint32_t n_mlp_params, n_splits;
	int loss_type;
	float loss_scale, n_total;
	int activation, output_activation; // Activation (activation.cuh): hidden layers / output
};

__host__ __device__ constexpr uint32_t wide_in_sub(uint32_t E) { return (E + 63u) / 64u; }
__host__ __device__ constexpr uint32_t wide_fwd_smem(uint32_t H, uint32_t E, uint32_t nwg) {
	return 1024u + nwg * WIDE_TILE_BYTES + wide_in_sub(E) * SUB_BYTES + (H - 1) * WIDE_TILE_BYTES + 4096u + 256u;
}
__host__ __device__ constexpr uint32_t wide_bwd_smem(uint32_t H, uint32_t E, uint32_t nwg, bool want_dx) {
	return 1024u + nwg * WIDE_TILE_BYTES + (want_dx ? wide_in_sub(E) * SUB_BYTES : 0u) + (H - 1) * WIDE_TILE_BYTES + 4096u + 256u;
}
constexpr uint32_t WIDE_DW_STAGES = 3;
constexpr uint32_t WIDE_DW_SMEM = 1024u + WIDE_DW_STAGES * 2 * WIDE_TILE_BYTES + 256u;

__device__ __forceinline__ uint8_t* wide_layer_image(__half* base, uint32_t layer, uint32_t n_tiles, uint32_t tile) {
	return reinterpret_cast<uint8_t*>(base) + ((size_t)layer * n_tiles + tile) * WIDE_TILE_BYTES;
}

// row-major [B][width] halfs -> tile images (for callers of the C ABI that hand over a row-major encoded batch)
__global__ void __launch_bounds__(256) wide_rowmajor_to_image_kernel(uint32_t B, uint32_t width, const __half* __restrict__ src, __half* __restrict__ dst) {
	const uint32_t cpr = width / 8, q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= B * cpr) return;
	const uint32_t i = q / cpr, c = q % cpr;
	*reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(dst) + wide_image_chunk(i, c, wide_in_sub(width))) = __ldg(reinterpret_cast<const uint4*>(src) + q);
}

// global row-major [rows][cols] halfs -> K-major swizzled sub-tiles of [rows][64 halfs] (rows * 128 bytes each); columns
// beyond `cols` up to the next multiple of 64 are zero-filled (they are read by the MN-major view of the input matrix)
__device__ __forceinline__ void wide_stage_weights(uint8_t* dst, const __half* src, uint32_t rows, uint32_t cols, uint32_t tid, uint32_t nt) {
	const uint32_t cpr = cols / 8, cprp = wide_in_sub(cols) * 8, sub_bytes = rows * 128;
	for (uint32_t q = tid; q < rows * cprp; q += nt) {
		const uint32_t r = q / cprp, c = q % cprp;
		uint4 v = make_uint4(0, 0, 0, 0);
		if (c < cpr) v = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r * cols) + c);
		*reinterpret_cast<uint4*>(dst + (c >> 3) * sub_bytes + sw128(r, c & 7)) = v;
	}
}

template <int NWG> __device__ __forceinline__ void wide_wg_sync(uint32_t wg) {
	if constexpr (NWG == 1) __syncthreads(); else named_bar_sync(1 + wg, 128);
}

// rows of `width` halfs (global, contiguous tile of 128 rows) -> swizzled tile; executed by 128 threads (index wt)
__device__ __forceinline__ void wide_load_tile(uint32_t dst_addr, const __half* src, uint32_t width, uint32_t wt) {
	const uint4* s4 = reinterpret_cast<const uint4*>(src);
	const uint32_t cpr = width / 8;
	for (uint32_t j0 = 0; j0 < cpr; j0 += 4) {
		uint4 v[4];
#pragma unroll
		for (uint32_t u = 0; u < 4; ++u) if (j0 + u < cpr) v[u] = ldg128_here(s4 + wt + 128 * (j0 + u));
#pragma unroll
		for (uint32_t u = 0; u < 4; ++u) {
			if (j0 + u < cpr) {
				const uint32_t q = wt + 128 * (j0 + u), r = q / cpr, c = q % cpr;
				sts128(dst_addr + (c >> 3) * SUB_BYTES + sw128(r, c & 7), v[u]);
			}
		}
	}
}

// ---- forward (+ loss) ----------------------------------------------------------------------------------------
// MODE_INFER: output only.  MODE_TRAIN: stores hidden activations, loss partials and dL/d(output).
// MODE_BACKWARD: stores hidden activations only (forward recomputation in front of a caller-supplied dL/d(output)).
template <int MODE, int NWG>
__global__ void __launch_bounds__(128 * NWG, 1) mlp_wide_forward_kernel(const __grid_constant__ WideArgs a) {
	constexpr bool STORE = MODE != MODE_INFER;
	extern __shared__ uint8_t smem_raw[];
	uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
	const uint32_t tid = threadIdx.x, wg = tid >> 7, wt = tid & 127, warp = wt >> 5, lane = tid & 31;
	const uint32_t E = a.in_width, H = a.n_hidden, in_sub = wide_in_sub(E);
	uint8_t* act = smem;                                    // NWG activation tiles, overwritten in place layer by layer
	uint8_t* w_in = act + NWG * WIDE_TILE_BYTES;            // in_sub x [128 out][128 B]
	uint8_t* w_hid = w_in + in_sub * SUB_BYTES;             // (H-1) x 2 x [128 out][128 B]
	uint8_t* w_out = w_hid + (H - 1) * WIDE_TILE_BYTES;     // 2 x [16 out][128 B]
	uint8_t* tail = w_out + 4096;
	uint64_t* bars = reinterpret_cast<uint64_t*>(tail);     // one per warpgroup: MMA completion
	uint64_t* bars_x = reinterpret_cast<uint64_t*>(tail + 16); // one per warpgroup: input tile landed (TMA)
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 32);
	float* red_buf = reinterpret_cast<float*>(tail + 64);   // NWG x 4

	constexpr uint32_t tmem_cols = 128 * NWG;
	if (tid < 32) tmem_alloc(tmem_slot, tmem_cols);
	if (tid == 32) {
		for (int g = 0; g < NWG; ++g) { mbar_init(bars + g, 1); mbar_init(bars_x + g, 1); }
		mbar_fence_init();
	}
	{
		const __half* p = a.params;
		wide_stage_weights(w_in, p, 128, E, tid, 128 * NWG);
		p += 128 * E;
		for (uint32_t l = 0; l + 1 < H; ++l) { wide_stage_weights(w_hid + l * WIDE_TILE_BYTES, p, 128, 128, tid, 128 * NWG); p += 128 * 128; }
		wide_stage_weights(w_out, p, 16, 128, tid, 128 * NWG);
	}
	fence_proxy_async_smem();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t t_acc = *tmem_slot + wg * 128;
	const uint32_t lane_off = (warp * 32) << 16;
	const uint32_t act_addr = smem_u32(act) + wg * WIDE_TILE_BYTES;
	const uint32_t idesc_hid = make_idesc_f16(128, 128, 0, 0, ACC_F32), idesc_out = make_idesc_f16(128, 16, 0, 0, ACC_F32);
	uint64_t* bar = bars + wg;
	uint64_t* bar_x = bars_x + wg;
	uint32_t phase = 0, phase_x = 0;
	float loss_acc = 0.f;
	const uint32_t row = wt;
	const uint32_t x_bytes = in_sub * SUB_BYTES, tile_step = gridDim.x * NWG;
	const uint8_t* x_img = reinterpret_cast<const uint8_t*>(a.input);
	// the input tile image goes straight into the activation tile with one bulk copy, requested as early as the tile is free
	if (wt == 0 && blockIdx.x * NWG + wg < a.n_tiles) {
		mbar_arrive_expect_tx(bar_x, x_bytes);
		tma_bulk_g2s_addr(act_addr, x_img + (size_t)(blockIdx.x * NWG + wg) * x_bytes, x_bytes, bar_x);
	}

	for (uint32_t tile = blockIdx.x * NWG + wg; tile < a.n_tiles; tile += tile_step) {
		const size_t sample0 = (size_t)tile * MLP_TILE;
		float tgt[4] = {0.f, 0.f, 0.f, 0.f};
		if (MODE == MODE_TRAIN) {
#pragma unroll
			for (uint32_t j = 0; j < 4; ++j) if (j < a.n_out) tgt[j] = ldg_f32_here(a.target + (sample0 + row) * a.n_out + j);
		}
		tc_fence_before();
		wide_wg_sync<NWG>(wg); // every thread has read the previous tile's output accumulator

		for (uint32_t k = 0; k < H; ++k) {
			if (wt == 0) {
				if (k == 0) mbar_wait(bar_x, phase_x); // TMA write -> MMA read: both async proxy, ordered by the mbarrier
				tc_fence_after();
				const uint32_t ksteps = (k == 0 ? E : 128u) / 16;
				const uint32_t b_addr = k == 0 ? smem_u32(w_in) : smem_u32(w_hid) + (k - 1) * WIDE_TILE_BYTES;
				for (uint32_t s = 0; s < ksteps; ++s) {
					const uint32_t off = (s >> 2) * SUB_BYTES + (s & 3) * 32;
					umma_f16(t_acc, make_desc_sw128(act_addr + off, 16, 1024), make_desc_sw128(b_addr + off, 16, 1024), idesc_hid, s > 0);
				}
				umma_commit(bar);
				if (STORE && k > 0) {
					// the tile holds hidden[k-1], complete and fenced: its image goes to HBM while the MMA reads it
					tma_bulk_s2g(wide_layer_image(a.hidden, k - 1, a.n_tiles, tile), act_addr, WIDE_TILE_BYTES);
					tma_bulk_commit();
					tma_bulk_wait_read();
				}
			}
			if (k == 0) phase_x ^= 1;
			mbar_wait(bar, phase); phase ^= 1;
			tc_fence_after();
			if (STORE && k > 0) wide_wg_sync<NWG>(wg); // thread 0 arrives once the bulk copy has read the tile: it may be overwritten
#pragma unroll
			for (int cb = 0; cb < 2; ++cb) { // 64 accumulator columns = one sub-tile row per step
				AccRow<32> acc;
				acc_issue(t_acc + lane_off + cb * 64, acc);
				acc_finish(acc);
#pragma unroll
				for (int c = 0; c < 8; ++c) {
					uint32_t pk[4];
#pragma unroll
					for (int j = 0; j < 4; ++j) pk[j] = hidden_act_forward<true>(acc.h[c * 4 + j], a.activation);
					sts128(act_addr + cb * SUB_BYTES + sw128(row, c), pk[0], pk[1], pk[2], pk[3]);
				}
			}
			fence_proxy_async_smem();
			tc_fence_before();
			wide_wg_sync<NWG>(wg);
		}
		// ---- output layer (N = 16, K = 128) ----
		if (wt == 0) {
			tc_fence_after();
			for (uint32_t s = 0; s < 8; ++s)
				umma_f16(t_acc, make_desc_sw128(act_addr + (s >> 2) * SUB_BYTES + (s & 3) * 32, 16, 1024),
					make_desc_sw128(smem_u32(w_out) + (s >> 2) * 2048 + (s & 3) * 32, 16, 1024), idesc_out, s > 0);
			umma_commit(bar);
			if (STORE) {
				tma_bulk_s2g(wide_layer_image(a.hidden, H - 1, a.n_tiles, tile), act_addr, WIDE_TILE_BYTES);
				tma_bulk_commit();
				tma_bulk_wait_read();
			}
		}
		mbar_wait(bar, phase); phase ^= 1;
		tc_fence_after();
		if (wt == 0 && tile + tile_step < a.n_tiles) {
			// the activation tile is free (output MMA complete, image copied out): fetch the next input under the output epilogue
			mbar_arrive_expect_tx(bar_x, x_bytes);
			tma_bulk_g2s_addr(act_addr, x_img + (size_t)(tile + tile_step) * x_bytes, x_bytes, bar_x);
		}
		{
			AccRow<8> acc;
			acc_issue(t_acc + lane_off, acc);
			acc_finish(acc);
			uint32_t yk[8];
#pragma unroll
			for (int j = 0; j < 8; ++j) yk[j] = acc.h[j];
			if (a.output_activation != ACT_NONE) {
#pragma unroll
				for (int j = 0; j < 8; ++j) yk[j] = act_forward_h2(yk[j], a.output_activation);
			}
			const size_t s = sample0 + row;
			if (MODE == MODE_BACKWARD && a.output_activation != ACT_NONE) {
				// forward recomputation in front of a caller-supplied dL/d(output): leave dL/d(output) * activation'(output) for
				// the backward and weight-gradient kernels (activation_backward_output_gpu, fully_fused_mlp.cu:757-762)
				const uint4* d4 = reinterpret_cast<const uint4*>(a.dL_doutput + s * MLP_OUT);
				const uint4 d0 = __ldg(d4), d1 = __ldg(d4 + 1);
				uint32_t dq[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
				for (int j = 0; j < 8; ++j) dq[j] = act_backward_h2(dq[j], yk[j], a.output_activation);
				uint4* d = reinterpret_cast<uint4*>(a.dL_doutput_out + s * MLP_OUT);
				d[0] = make_uint4(dq[0], dq[1], dq[2], dq[3]);
				d[1] = make_uint4(dq[4], dq[5], dq[6], dq[7]);
			}
			if (MODE != MODE_BACKWARD && a.output) {
				uint4* o = reinterpret_cast<uint4*>(a.output + s * MLP_OUT);
				o[0] = make_uint4(yk[0], yk[1], yk[2], yk[3]);
				o[1] = make_uint4(yk[4], yk[5], yk[6], yk[7]);
			}
			if (MODE == MODE_INFER && a.output_f32) {
				const __half* yh = reinterpret_cast<const __half*>(yk);
				for (uint32_t j = 0; j < a.n_out; ++j) a.output_f32[s * a.n_out + j] = __half2float(yh[j]);
			}
			if (MODE == MODE_TRAIN) {
				// same operation order as losses/l2.h:62-74 / relative_l2.h:62-75 (see mlp_kernels.cuh)
				const __half* yh = reinterpret_cast<const __half*>(yk);
				const bool relative = a.loss_type == LOSS_RELATIVE_L2;
				auto loss_elem = [&](float pred, float target) -> float {
					const float diff = pred - target;
					float value = diff * diff, grad = 2.f * diff;
					if (relative) { const float denom = pred * pred + 0.01f; value = value / denom; grad = grad / denom; }
					loss_acc += value / a.n_total;
					return a.loss_scale * grad / a.n_total;
				};
				float g4[4];
#pragma unroll
				for (uint32_t j = 0; j < 4; ++j) g4[j] = (j < a.n_out) ? loss_elem(__half2float(yh[j]), tgt[j]) : 0.f;
				uint32_t dyk[8] = {pack_h2(g4[0], g4[1]), pack_h2(g4[2], g4[3]), 0u, 0u, 0u, 0u, 0u, 0u};
				if (a.n_out > 4) {
#pragma unroll
					for (uint32_t j = 4; j < 16; j += 2) {
						const float ga = (j < a.n_out) ? loss_elem(__half2float(yh[j]), __ldg(a.target + s * a.n_out + j)) : 0.f;
						const float gb = (j + 1 < a.n_out) ? loss_elem(__half2float(yh[j + 1]), __ldg(a.target + s * a.n_out + j + 1)) : 0.f;
						dyk[j / 2] = pack_h2(ga, gb);
					}
				}
				if (a.output_activation != ACT_NONE) {
#pragma unroll
					for (int j = 0; j < 8; ++j) dyk[j] = act_backward_h2(dyk[j], yk[j], a.output_activation);
				}
				uint4* d = reinterpret_cast<uint4*>(a.dL_doutput_out + s * MLP_OUT);
				d[0] = make_uint4(dyk[0], dyk[1], dyk[2], dyk[3]);
				d[1] = make_uint4(dyk[4], dyk[5], dyk[6], dyk[7]);
			}
		}
		tc_fence_before(); // the next tile's barrier orders these TMEM reads before its first MMA
	}
	if (MODE == MODE_TRAIN && a.loss_partials != nullptr) {
		float s = loss_acc;
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
		if (lane == 0) red_buf[wg * 4 + warp] = s;
		wide_wg_sync<NWG>(wg);
		if (wt == 0) a.loss_partials[blockIdx.x * NWG + wg] = red_buf[wg * 4] + red_buf[wg * 4 + 1] + red_buf[wg * 4 + 2] + red_buf[wg * 4 + 3];
	}
	if (STORE && wt == 0) tma_bulk_wait_all();
	tc_fence_before();
	__syncthreads();
	if (tid < 32) tmem_dealloc(*tmem_slot, tmem_cols);
}

// ---- backward (dgrad chain) ----------------------------------------------------------------------------------
// Matrices m = 0..H (0 = input, H = output); ACT_m = input of matrix m (ACT_0 = X, ACT_m = hidden[m-1]); G_m = gradient
// w.r.t. the output of matrix m before its activation (G_H = dL/d(output)).  G_{m-1} = (G_m * W_m) (.) [ACT_m > 0]
// (fully_fused_mlp.cu:150-259, ReLU' from the stored post-activation, common_device.h:243-248), written to gtmp[m-1] for
// the weight-gradient kernel; finally dL/d(input) = G_0 * W_0.  The G tile is overwritten in place.
template <int NWG>
__global__ void __launch_bounds__(128 * NWG, 1) mlp_wide_backward_kernel(const __grid_constant__ WideArgs a) {
	extern __shared__ uint8_t smem_raw[];
	uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
	const uint32_t tid = threadIdx.x, wg = tid >> 7, wt = tid & 127, warp = wt >> 5;
	const uint32_t E = a.in_width, H = a.n_hidden, in_sub = wide_in_sub(E);
	const bool want_dx = a.dL_dinput != nullptr;
	uint8_t* gt = smem;                                                 // NWG gradient tiles
	uint8_t* w_in = gt + NWG * WIDE_TILE_BYTES;                         // only when dL/d(input) is wanted
	uint8_t* w_hid = w_in + (want_dx ? in_sub * SUB_BYTES : 0u);
	uint8_t* w_out = w_hid + (H - 1) * WIDE_TILE_BYTES;
	uint8_t* tail = w_out + 4096;
	uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 32);

	constexpr uint32_t tmem_cols = 128 * NWG;
	if (tid < 32) tmem_alloc(tmem_slot, tmem_cols);
	if (tid == 32) {
		for (int g = 0; g < NWG; ++g) mbar_init(bars + g, 1);
		mbar_fence_init();
	}
	{
		const __half* p = a.params;
		if (want_dx) wide_stage_weights(w_in, p, 128, E, tid, 128 * NWG);
		p += 128 * E;
		for (uint32_t l = 0; l + 1 < H; ++l) { wide_stage_weights(w_hid + l * WIDE_TILE_BYTES, p, 128, 128, tid, 128 * NWG); p += 128 * 128; }
		wide_stage_weights(w_out, p, 16, 128, tid, 128 * NWG);
	}
	// columns 16.. of the gradient tile are read (as zeros) by nobody before they are written, except through the K = 16
	// A operand of the output matrix, which only touches chunks 0 and 1: no clearing needed
	fence_proxy_async_smem();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t t_acc = *tmem_slot + wg * 128;
	const uint32_t lane_off = (warp * 32) << 16;
	const uint32_t g_addr = smem_u32(gt) + wg * WIDE_TILE_BYTES;
	const uint32_t idesc_dgrad = make_idesc_f16(128, 128, 0, 1, ACC_F32);         // A K-major, B = weight tile viewed MN-major
	const uint32_t idesc_dx = make_idesc_f16(128, in_sub * 64, 0, 1, ACC_F32);
	uint64_t* bar = bars + wg;
	uint32_t phase = 0;
	const uint32_t row = wt;

	for (uint32_t tile = blockIdx.x * NWG + wg; tile < a.n_tiles; tile += gridDim.x * NWG) {
		const size_t sample0 = (size_t)tile * MLP_TILE;
		{
			const uint4* d4 = reinterpret_cast<const uint4*>(a.dL_doutput + (sample0 + row) * MLP_OUT);
			const uint4 d0 = ldg128_here(d4), d1 = ldg128_here(d4 + 1);
			sts128(g_addr + sw128(row, 0), d0);
			sts128(g_addr + sw128(row, 1), d1);
		}
		fence_proxy_async_smem();
		tc_fence_before();
		wide_wg_sync<NWG>(wg);

		for (uint32_t m = H; m >= 1; --m) {
			if (wt == 0) {
				tc_fence_after();
				if (m == H) {
					// Z[128 x 128 in] = dy[128 x 16] * W_out[16 x 128]: one K step; B atoms (64 inputs each) 2 KB apart
					umma_f16(t_acc, make_desc_sw128(g_addr, 16, 1024), make_desc_sw128(smem_u32(w_out), 2048, 1024), idesc_dgrad, 0);
				} else {
					const uint32_t wb = smem_u32(w_hid) + (m - 1) * WIDE_TILE_BYTES;
					for (uint32_t s = 0; s < 8; ++s) // K = 128 outputs of W_m, 16 rows (2 KB) per step; N atoms 16 KB apart
						umma_f16(t_acc, make_desc_sw128(g_addr + (s >> 2) * SUB_BYTES + (s & 3) * 32, 16, 1024),
							make_desc_sw128(wb + s * 2048, SUB_BYTES, 1024), idesc_dgrad, s > 0);
				}
				umma_commit(bar);
				if (m < H) {
					// the tile holds G_m, complete and fenced: its image goes to HBM (for the weight-gradient kernel) under the MMA
					tma_bulk_s2g(wide_layer_image(a.gtmp, m, a.n_tiles, tile), g_addr, WIDE_TILE_BYTES);
					tma_bulk_commit();
					tma_bulk_wait_read();
				}
			}
			// ReLU mask source: this row of ACT_m = hidden[m-1], read from the tile image; requested now, in flight under the MMA
			const uint8_t* himg = wide_layer_image(a.hidden, m - 1, a.n_tiles, tile);
			uint4 av[16];
#pragma unroll
			for (int c = 0; c < 16; ++c) av[c] = ldg128_here(himg + (c >> 3) * SUB_BYTES + sw128(row, c & 7));
			mbar_wait(bar, phase); phase ^= 1;
			tc_fence_after();
			if (m < H) wide_wg_sync<NWG>(wg); // the bulk copy has read the tile (thread 0 arrives after its wait)
#pragma unroll
			for (int cb = 0; cb < 2; ++cb) {
				AccRow<32> acc;
				acc_issue(t_acc + lane_off + cb * 64, acc);
				acc_finish(acc);
#pragma unroll
				for (int c = 0; c < 8; ++c) {
					const uint4 aw4 = av[cb * 8 + c];
					const uint32_t aw[4] = {aw4.x, aw4.y, aw4.z, aw4.w};
					uint32_t pk[4];
#pragma unroll
					for (int j = 0; j < 4; ++j) pk[j] = hidden_act_backward<true>(acc.h[c * 4 + j], aw[j], a.activation);
					sts128(g_addr + cb * SUB_BYTES + sw128(row, c), pk[0], pk[1], pk[2], pk[3]);
				}
			}
			fence_proxy_async_smem();
			tc_fence_before();
			wide_wg_sync<NWG>(wg);
		}
		if (wt == 0) {
			tc_fence_after();
			if (want_dx) {
				for (uint32_t s = 0; s < 8; ++s)
					umma_f16(t_acc, make_desc_sw128(g_addr + (s >> 2) * SUB_BYTES + (s & 3) * 32, 16, 1024),
						make_desc_sw128(smem_u32(w_in) + s * 2048, SUB_BYTES, 1024), idesc_dx, s > 0);
				umma_commit(bar);
			}
			tma_bulk_s2g(wide_layer_image(a.gtmp, 0, a.n_tiles, tile), g_addr, WIDE_TILE_BYTES); // G_0
			tma_bulk_commit();
			tma_bulk_wait_read();
		}
		if (want_dx) {
			mbar_wait(bar, phase); phase ^= 1;
			tc_fence_after();
			__half* dst = a.dL_dinput + (sample0 + row) * E;
			for (uint32_t c0 = 0; c0 < E; c0 += 16) {
				AccRow<8> acc;
				acc_issue(t_acc + lane_off + c0, acc);
				acc_finish(acc);
				const uint32_t (&pk)[8] = acc.h;
				reinterpret_cast<uint4*>(dst + c0)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
				reinterpret_cast<uint4*>(dst + c0)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
			}
		}
		tc_fence_before();
		wide_wg_sync<NWG>(wg); // G_0's image has been read out (thread 0) before the next tile's dL/d(output) rows overwrite the tile
	}
	if (wt == 0) tma_bulk_wait_all();
	tc_fence_before();
	__syncthreads();
	if (tid < 32) tmem_dealloc(*tmem_slot, tmem_cols);
}

// ---- weight gradients: split-K over the batch ------------------------------------------------------------------
// CTA (split j, matrix m): D[out][in] = sum over its tiles of G_m^T[out x 128 samples] * ACT_m[128 samples x in], both
// operands MN-major views of [sample][feature] tiles; accumulated in TMEM (fp16 like the reference's split-K GEMMs, see ACC_F32) across all tiles of the CTA, one flush at the end into
// partials[j]; mlp_finalize(_adam)_kernel sums the splits (deterministic).  Replaces the reference's three split-K CUTLASS
// GEMMs (fully_fused_mlp.cu:783-829).  The kernel is a pure stream over HBM (64 KB of operands per 0.27 us of MMA): the
// tile images are fetched by bulk copies into a 3-stage ring, two tiles ahead of the MMA; only thread 0 works in the loop.
// dL/d(output) (matrix H, [B][16] row-major, 4 KB per tile) is the one operand staged through registers.
__global__ void __launch_bounds__(128, 1) mlp_wide_dw_kernel(const __grid_constant__ WideArgs a) {
	extern __shared__ uint8_t smem_raw[];
	uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
	const uint32_t tid = threadIdx.x, warp = tid >> 5;
	const uint32_t E = a.in_width, H = a.n_hidden, in_sub = wide_in_sub(E);
	const uint32_t split = blockIdx.x, m = blockIdx.y;
	constexpr uint32_t NS = WIDE_DW_STAGES, STAGE_BYTES = 2 * WIDE_TILE_BYTES; // [A 32 KB][B 32 KB]
	uint8_t* tail = smem + NS * STAGE_BYTES;
	uint64_t* full = reinterpret_cast<uint64_t*>(tail);        // NS: operands of the stage have landed
	uint64_t* done = reinterpret_cast<uint64_t*>(tail + 32);   // NS: the MMAs reading the stage have completed
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 64);
	if (warp == 0) tmem_alloc(tmem_slot, 128);
	if (tid == 32) {
		for (uint32_t s = 0; s < NS; ++s) { mbar_init(full + s, 1); mbar_init(done + s, 1); }
		mbar_fence_init();
	}
	const bool a_is_dy = m == H; // 16 columns, row-major, through registers; the other 112 columns of the A tiles stay zero
	if (a_is_dy)
		for (uint32_t q = tid; q < NS * STAGE_BYTES / 16; q += 128) reinterpret_cast<uint4*>(smem)[q] = make_uint4(0, 0, 0, 0);
	fence_proxy_async_smem();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t t_acc = *tmem_slot;
	const uint32_t lane_off = (warp * 32) << 16;
	const uint32_t bw = (m == 0) ? E : (uint32_t)WIDE_W;
	const uint32_t b_bytes = (m == 0) ? in_sub * SUB_BYTES : WIDE_TILE_BYTES;
	const uint32_t tx_bytes = b_bytes + (a_is_dy ? 0u : WIDE_TILE_BYTES);
	const uint32_t idesc = make_idesc_f16(128, (m == 0) ? in_sub * 64 : 128, 1, 1, ACC_F32);
	const uint32_t smem_addr = smem_u32(smem);
	const uint32_t n_my = split < a.n_tiles ? (a.n_tiles - split + a.n_splits - 1) / a.n_splits : 0u;
	auto tile_of = [&](uint32_t i) { return split + i * a.n_splits; };
	auto request = [&](uint32_t i) { // thread 0: bulk copies of tile i's operands into stage i % NS
		const uint32_t st = i % NS, tile = tile_of(i);
		mbar_arrive_expect_tx(full + st, tx_bytes);
		if (!a_is_dy) tma_bulk_g2s_addr(smem_addr + st * STAGE_BYTES, wide_layer_image(a.gtmp, m, a.n_tiles, tile), WIDE_TILE_BYTES, full + st);
		const void* bsrc = (m == 0) ? (const void*)(reinterpret_cast<const uint8_t*>(a.input) + (size_t)tile * b_bytes)
			: (const void*)wide_layer_image(a.hidden, m - 1, a.n_tiles, tile);
		tma_bulk_g2s_addr(smem_addr + st * STAGE_BYTES + WIDE_TILE_BYTES, bsrc, b_bytes, full + st);
	};
	uint4 dy_next[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
	auto load_dy = [&](uint32_t i) {
		const uint4* d4 = reinterpret_cast<const uint4*>(a.dL_doutput + ((size_t)tile_of(i) * MLP_TILE + tid) * MLP_OUT);
		dy_next[0] = ldg128_here(d4); dy_next[1] = ldg128_here(d4 + 1);
	};
	if (tid == 0) for (uint32_t i = 0; i < n_my && i < NS - 1; ++i) request(i);
	if (a_is_dy && n_my > 0) load_dy(0);

	for (uint32_t i = 0; i < n_my; ++i) {
		const uint32_t st = i % NS;
		if (a_is_dy) {
			// stage st was last read by the MMAs of tile i - NS; thread 0 waited for them two iterations ago (below), i.e.
			// before it arrived at the previous iteration's __syncthreads
			const uint32_t a_addr = smem_addr + st * STAGE_BYTES;
			sts128(a_addr + sw128(tid, 0), dy_next[0]);
			sts128(a_addr + sw128(tid, 1), dy_next[1]);
			if (i + 1 < n_my) load_dy(i + 1);
			fence_proxy_async_smem();
		}
		tc_fence_before();
		__syncthreads();
		if (tid == 0) {
			mbar_wait(full + st, (i / NS) & 1);
			tc_fence_after();
			const uint32_t a_addr = smem_addr + st * STAGE_BYTES, b_addr = a_addr + WIDE_TILE_BYTES;
			for (uint32_t s = 0; s < 8; ++s) // K = 128 samples, 16 rows (2 KB) per step; M / N atoms 16 KB apart
				umma_f16(t_acc, make_desc_sw128(a_addr + s * 2048, SUB_BYTES, 1024), make_desc_sw128(b_addr + s * 2048, SUB_BYTES, 1024), idesc, i > 0 || s > 0);
			umma_commit(done + st);
			if (i + NS - 1 < n_my) {
				// refill: stage of tile i + NS - 1 == stage of tile i - 1, free once that tile's MMAs have completed
				if (i >= 1) mbar_wait(done + (i - 1) % NS, ((i - 1) / NS) & 1);
				request(i + NS - 1);
			}
		}
	}
	if (n_my > 0) mbar_wait(done + (n_my - 1) % NS, ((n_my - 1) / NS) & 1); // all threads: the accumulator is final
	tc_fence_after();
	{
		const uint32_t out_rows = (m == H) ? (uint32_t)MLP_OUT : (uint32_t)WIDE_W;
		const uint32_t mat_off = (m == 0) ? 0u : 128u * E + (m - 1) * 128u * 128u;
		float* out = a.dW_partials + (size_t)split * a.n_mlp_params + mat_off + (size_t)tid * bw;
		for (uint32_t c0 = 0; c0 < bw; c0 += 16) {
			uint32_t v[16];
			tmem_ld16(t_acc + lane_off + c0, v);
			tmem_ld_wait();
			if (tid < out_rows) {
#pragma unroll
				for (int j = 0; j < 4; ++j)
					reinterpret_cast<float4*>(out + c0)[j] = n_my == 0 ? make_float4(0, 0, 0, 0)
						: make_float4(dw_value(v[4 * j]), dw_value(v[4 * j + 1]), dw_value(v[4 * j + 2]), dw_value(v[4 * j + 3]));
			}
		}
	}
	tc_fence_before();
	__syncthreads();
	if (warp == 0) tmem_dealloc(t_acc, 128);
}

} // namespace b200
