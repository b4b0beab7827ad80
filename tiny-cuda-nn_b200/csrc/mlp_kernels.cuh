// Fully fused MLP (width 64) for sm_100a on tcgen05 tensor cores.
//
// What it computes is the reference's FullyFusedMLP (src/fully_fused_mlp.cu:47-259, 499-557, 735-836): ReLU hidden
// layers, linear output padded to 16, fp16 storage; plus (MODE_TRAIN) the L2 / RelativeL2 loss (losses/l2.h:40-75,
// losses/relative_l2.h:40-75) and the weight gradients the reference obtains from CUTLASS split-K GEMMs
// (include/tiny-cuda-nn/cutlass_matmul.h:437-479) and dL/d(input) (cutlass_matmul.h:351-408).
//
// How it is built is new (not wmma/CUTLASS): one persistent CTA of 128 threads owns 128-sample tiles.
//   * every operand tile lives in shared memory exactly once, as rows of 128 bytes with the 128-byte swizzle:
//     activation / gradient tiles [128 samples][64 halfs], weight tiles [out][in].  The same bytes serve as the
//     K-major A operand of the forward / dgrad MMAs and as MN-major operands of the weight-gradient MMAs
//     (where the batch is the K dimension), so nothing is ever transposed or stored twice;
//   * accumulators live in TMEM: one 64-column block for the layer chain, and one 64-column block per weight matrix
//     whose tcgen05.mma accumulate across ALL tiles of the CTA (split-K over the batch without any HBM traffic);
//     they are flushed once per CTA as fp32 partials and reduced by mlp_finalize_kernel (deterministic);
//   * thread t owns sample row t of the tile: tcgen05.ld gives it the whole fp32 row, it applies ReLU / the ReLU mask
//     (kept as 64-bit register masks, the reference re-reads the stored activations) / the loss, and writes the fp16
//     row back to shared memory for the next MMA.  No activation ever goes to HBM (the reference moves ~0.5 GB/step).
#pragma once
#include "activation.cuh"
#include "grid_kernels.cuh"
#include "ptx.cuh"

namespace b200 {

// Development aid (tests/dev/mlp_timing.cu): per-stage clock64 stamps of CTA 0 / thread 0.
#ifdef MLP_STAGE_CLOCKS
#define STAGE_CLOCK(args) do { if (blockIdx.x == 0 && threadIdx.x == 0 && (args).debug_clock && dbg_n < 256) (args).debug_clock[dbg_n++] = clock64(); } while (0)
#else
#define STAGE_CLOCK(args) do { } while (0)
#endif

constexpr int MLP_W = 64;          // hidden width handled by this kernel
constexpr int MLP_TILE = 128;      // samples per tile = UMMA M
constexpr int MLP_OUT = 16;        // padded output width (fully_fused_mlp.cu:544: fused last layer only for <=16 outputs)
constexpr int TILE_BYTES = MLP_TILE * 128;
constexpr int MLP_MAX_HIDDEN = 5;       // 4 in the fused-scatter kernels; 5 = data/config_oneblob.json with 64 neurons (bench_ours.cu)

enum MlpMode : int { MODE_INFER = 0, MODE_TRAIN = 1, MODE_BACKWARD = 2 };
enum LossType : int { LOSS_L2 = 0, LOSS_RELATIVE_L2 = 1 };

struct MlpArgs {
	uint32_t n_tiles;              // B / 128
	uint32_t in_width;             // E_pad: 16, 32, 48 or 64
	uint32_t n_hidden;             // hidden layers H >= 1 (hidden matmuls = H-1)
	uint32_t n_out;                // real output dims (<= 16)
	uint32_t n_neurons;            // 64; 16 or 32 in the GEN instantiations (the narrow network sits zero-padded inside the 64-wide tiles)
	const __half* input;           // [B][in_width]
	const __half* params;          // [W x in], (H-1) x [W x W], [16 x W]   row-major (fully_fused_mlp.cu:654-671), W = n_neurons
	__half* output;                // [B][16]  (nullable in TRAIN)
	float* output_f32;             // [B][n_out] (INFER only, nullable): fused trim_and_cast (object.h:174-175)
	const float* target;           // [B][n_out] fp32 (TRAIN)
	const __half* dL_doutput;      // [B][16] (BACKWARD)
	__half* dL_dinput;             // [B][in_width] (nullable)
	float* dW_partials;            // [gridDim.x][n_mlp_params] fp32 (TRAIN/BACKWARD, nullable => no weight gradients)
	float* loss_partials;          // [gridDim.x] (TRAIN)
	uint32_t n_mlp_params;
	int loss_type;
	float loss_scale;
	float inv_n_total;             // 1 / (B * n_out)  (losses/l2.h:62-64: n_total = n_elements / stride * dims)
	float n_total;
	// fused gradient scatter (kernel template FD > 0): dL/d(encoded) never leaves the SM, each row's thread scatters it
	const float* positions;        // [B][D] fp32
	__half* grid_grad;             // [n_params_grid]
	__half* replicas;              // private copies of the replicated coarse levels (grid_kernels.cuh)
	int activation;                // hidden layers (Activation, activation.cuh); kernels instantiated with GEN = false assume ReLU
	int output_activation;         // applied to the (padded) output; its derivative, from the activated output, to dL/d(output)
	long long* debug_clock;        // MLP_STAGE_CLOCKS builds only
	// 128-neuron path (mlp_wide_kernels.cuh; host side only): per-layer buffers in the caller's workspace
	__half* wide_hidden; __half* wide_gtmp; __half* wide_dy;
};

// Accumulator precision.  The reference accumulates every layer in fp16 (wmma half fragments, fully_fused_mlp.cu:66-68) and its
// weight-gradient GEMMs too (cutlass_matmul.h:59-86), and its results depend on that: with fp16 accumulators in TMEM (instruction
// descriptor c_format = F16, the tensor core rounds to fp16 after every 16-deep step exactly like HMMA.16816.F16) the network
// output is BIT-IDENTICAL to the reference build's, no ReLU mask differs, and the 1000-step loss agrees to < 0.5 % where fp32
// accumulation - more accurate, but not the reference's arithmetic - ended 4 % apart.  So fp16 is the
// default; -DTCNN_B200_F32_ACC builds the fp32-accumulating kernels.
#ifdef TCNN_B200_F32_ACC
constexpr uint32_t ACC_F32 = 1;
#else
constexpr uint32_t ACC_F32 = 0;
#endif
// This thread's accumulator row, columns [0, 2 N), as N packed half2.  fp16 accumulators sit in the low half of their 32-bit TMEM
// column and are fetched two per register (tcgen05.ld .pack::16b); fp32 accumulators are fetched whole and rounded here.
template <int N> struct AccRow {
	uint32_t h[N];
	uint32_t raw[ACC_F32 ? 2 * N : 1];
};
template <int N> __device__ __forceinline__ void acc_issue(uint32_t taddr, AccRow<N>& a) {
	static_assert(N == 8 || N == 16 || N == 32, "8, 16 or 32 packed pairs");
	if constexpr (ACC_F32) {
		if constexpr (N == 8) { tmem_ld16(taddr, a.raw); }
		else if constexpr (N == 16) { tmem_ld32(taddr, a.raw); }
		else { tmem_ld32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&a.raw[0])); tmem_ld32(taddr + 32, *reinterpret_cast<uint32_t(*)[32]>(&a.raw[32])); }
	} else {
		if constexpr (N == 8) tmem_ld8_pack16(taddr, a.h);
		else if constexpr (N == 16) tmem_ld16_pack16(taddr, a.h);
		else tmem_ld32_pack16(taddr, a.h);
	}
}
template <int N> __device__ __forceinline__ void acc_finish(AccRow<N>& a) {
	tmem_ld_wait();
	if constexpr (ACC_F32) {
#pragma unroll
		for (int i = 0; i < N; ++i) a.h[i] = pack_h2(__uint_as_float(a.raw[2 * i]), __uint_as_float(a.raw[2 * i + 1]));
	}
}
// one weight-gradient accumulator element as read by the unpacked tcgen05.ld of the per-CTA flush
__device__ __forceinline__ float dw_value(uint32_t v) {
	if constexpr (ACC_F32) return __uint_as_float(v);
	else return __half2float(__ushort_as_half((unsigned short)(v & 0xFFFFu)));
}

// byte offset of 16-byte chunk `c` of row `r` inside a 128-byte-swizzled tile
__device__ __forceinline__ uint32_t sw128(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }

__host__ __device__ constexpr uint32_t mlp_smem_bytes(int n_hidden, int in_width, bool train) {
	// [ACT 0..H][G0][G1 (train)] [W_in 8K][W_inT E x 128 B][W_hid (H-1) x 8K][W_out 2K][barrier etc.]
	// The weight region (>= 8K + 2K + 2K..8K + ...) directly follows the tiles, so "tile + 16 KB" (the second, unused
	// 64-row atom of an M = 128 MN-major operand) is always mapped shared memory.
	// in_width > 64 (kernel template WIN): the input tile ACT[0] and W_in have a second 64-column sub-tile each.
	return 1024u /*align slack*/ + (uint32_t)(n_hidden + 1 + (train ? 2 : 0) + (in_width > 64 ? 1 : 0)) * TILE_BYTES + (in_width > 64 ? 16384u : 8192u)
		+ (uint32_t)in_width * 128u + (uint32_t)(n_hidden - 1) * 8192u + 2048u + 256u + (n_hidden == 1 && in_width < 48 ? 4096u : 0u);
}
// extra shared memory of the warp-specialised fused kernel: the dL/d(encoded) hand-over slot
__host__ __device__ constexpr uint32_t mlp_ws_slot_bytes(int in_width) { return 128u * (uint32_t)in_width * 2u; }
__host__ __device__ constexpr uint32_t mlp_tmem_cols(int n_hidden, bool train, bool wide_in = false) {
	// chain (128 columns if dL/d(input) can be 128 wide) + dW_in + (H-1) dW_hid + dW_out
	const uint32_t need = train ? 64u * (uint32_t)(n_hidden + 2) + (wide_in ? 64u : 0u) : 64u;
	return need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512;
}

// WS (warp-specialised scatter, fused kernels only): the CTA has a second warpgroup (threads 128..255) that does nothing
// but the hash-table scatter.  The compute warpgroup hands each tile's dL/d(encoded) over through one shared-memory slot
// (named barriers 2 = "slot full", 3 = "slot empty"), so the reductions of tile k run under the whole MMA chain of tile k+1
// instead of after it; the compute warpgroup only waits if the scatter of tile k is still running when dX of tile k+1 is ready.
#ifndef SCATTER_CHUNK_UNROLL
#define SCATTER_CHUNK_UNROLL 1
#endif
constexpr int kScatterChunkUnroll = SCATTER_CHUNK_UNROLL;
#define CTA_SYNC() do { if constexpr (NSG > 0) named_bar_sync(1, 128); else __syncthreads(); } while (0)
// GEN: hidden activation taken from a.activation at run time (any of activation.cuh); false = ReLU compiled in (hot kernels)
template <bool GEN> __device__ __forceinline__ uint32_t hidden_act_forward(uint32_t pre, int act) {
	if constexpr (GEN) return act == ACT_RELU ? h2_relu(pre) : act_forward_h2(pre, act);
	else return h2_relu(pre);
}
template <bool GEN> __device__ __forceinline__ uint32_t hidden_act_backward(uint32_t grad, uint32_t post, int act) {
	if constexpr (GEN) return act == ACT_RELU ? h2_mask_mul(grad, post) : act_backward_h2(grad, post, act);
	else return h2_mask_mul(grad, post);
}

// WIN: the input may be wider than the network (64 < in_width <= 128: OneBlob 2 x 64 in front of 64 neurons, bench_ours.cu) -
// the input tile, W_in and the dL/d(input) accumulator get a second 64-column half
template <int MODE, int H_, int FD = 0, int FF = 2, int NSG = 0, bool GEN = false, bool WIN = false>
__global__ void __launch_bounds__(128 + 128 * NSG, NSG ? 2 : 1) mlp_kernel(const __grid_constant__ MlpArgs a, const __grid_constant__ GridParams gp) {
	constexpr bool TRAIN = MODE != MODE_INFER;
	constexpr bool FUSE_SCATTER = FD > 0;
	constexpr bool WS = NSG > 0;           // NSG scatter warpgroups behind the compute warpgroup
	constexpr uint32_t NT = 128 + 128 * NSG;
	// registers per thread: the launch gives 65536 / (2 CTAs x NT) to everybody; the scatter warpgroups hand back what
	// they do not need and the compute warpgroup (128 registers of accumulator rows) takes it
	constexpr uint32_t REG_LAUNCH = NSG == 0 ? 255 : (65536 / (2 * NT)) / 8 * 8, REG_COMPUTE = 128;
	constexpr uint32_t REG_SCATTER = NSG == 0 ? 0 : ((NT * REG_LAUNCH - 128 * REG_COMPUTE) / (128 * NSG)) / 8 * 8;
	static_assert(!FUSE_SCATTER || TRAIN, "the fused scatter belongs to the backward pass");
	static_assert(!WS || FUSE_SCATTER, "warp specialisation is for the fused scatter");
	extern __shared__ uint8_t smem_raw[];
	uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
	const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	constexpr uint32_t H = H_;
	const uint32_t E = a.in_width;

	// ---- shared memory carve-up -------------------------------------------------------------------
	uint8_t* act = smem;                                        // ACT[0..H], ACT[0] = input tile
	constexpr uint32_t ACT0_EXTRA = WIN ? TILE_BYTES : 0;       // ACT[0] = input tile: columns 64.. in a second sub-tile right behind the first
	uint8_t* gbuf = act + ACT0_EXTRA + (H + 1) * TILE_BYTES;    // G[0], G[1]
	uint8_t* w_in = gbuf + (TRAIN ? 2 : 0) * TILE_BYTES;        // [64 out][128 B] (x 2 sub-tiles if WIN), in < E valid
	uint8_t* w_inT = w_in + (WIN ? 16384 : 8192);               // [E in][128 B = 64 out]   (K-major B operand of dX = G * W_in)
	uint8_t* w_hid = w_inT + E * 128;                            // (H-1) x [64 out][128 B]
	uint8_t* w_out = w_hid + (H - 1) * 8192;                    // [16 out][128 B]
	uint8_t* tail = w_out + 2048;
	uint64_t* bar = reinterpret_cast<uint64_t*>(tail);
	uint64_t* bar2 = reinterpret_cast<uint64_t*>(tail + 8); // completion of the weight-gradient MMAs of a tile's last stage
	uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 16);
	float* red_buf = reinterpret_cast<float*>(tail + 32);       // 4 floats
	uint8_t* dx_slot = tail + 256 + ((H == 1 && E < 48) ? 4096 : 0); // WS: [128 rows][E halves], 16-byte chunks swizzled

	// ---- one-time setup ---------------------------------------------------------------------------
	constexpr uint32_t tmem_cols = mlp_tmem_cols((int)H, TRAIN, WIN);
	static_assert(!WIN || (FD == 0 && NSG == 0), "inputs wider than 64 only in the unfused kernels");
	if (warp == 0) tmem_alloc(tmem_slot, tmem_cols);
	if (tid == 32) { mbar_init(bar, 1); mbar_init(bar2, 1); mbar_fence_init(); }

	// Network width: 64 compiled in, except in the GEN instantiations, which also run the reference's 16- and 32-neuron networks
	// (fully_fused_mlp.cu:862-870): their matrices sit zero-padded inside the 64-wide tiles, so every MMA keeps its shape; a padded
	// neuron has zero weights on both sides, contributes nothing forward and receives a zero gradient.
	const uint32_t W = GEN ? a.n_neurons : 64u;
	// weights: global row-major -> swizzled tiles (16-byte chunks, coalesced reads)
	{
		const __half* p = a.params;
		const uint32_t cpr_in = E / 8; // chunks per row of W_in
		const uint32_t cpr_w = W / 8;  // chunks per row of a hidden / output matrix
		if constexpr (GEN) {
			if (W < 64) for (uint8_t* z = w_in + tid * 16; z < w_out + 2048; z += NT * 16) *reinterpret_cast<uint4*>(z) = make_uint4(0, 0, 0, 0);
			__syncthreads();
		}
		for (uint32_t q = tid; q < W * cpr_in; q += NT) {
			const uint32_t r = q / cpr_in, c = q % cpr_in;
			*reinterpret_cast<uint4*>(w_in + (c >> 3) * 8192 + sw128(r, c & 7)) = __ldg(reinterpret_cast<const uint4*>(p + r * E) + c);
		}
		for (uint32_t q = tid; q < W * E; q += NT) { // transposed copy, element-wise (once per CTA)
			const uint32_t o = q / E, i = q % E;
			*reinterpret_cast<__half*>(w_inT + sw128(i, o >> 3) + (o & 7) * 2) = p[o * E + i];
		}
		p += W * E;
		for (uint32_t l = 0; l + 1 < H; ++l) {
			for (uint32_t q = tid; q < W * cpr_w; q += NT) {
				const uint32_t r = q / cpr_w, c = q % cpr_w;
				*reinterpret_cast<uint4*>(w_hid + l * 8192 + sw128(r, c)) = __ldg(reinterpret_cast<const uint4*>(p + r * W) + c);
			}
			p += W * W;
		}
		for (uint32_t q = tid; q < 16 * cpr_w; q += NT) {
			const uint32_t r = q / cpr_w, c = q % cpr_w;
			*reinterpret_cast<uint4*>(w_out + sw128(r, c)) = __ldg(reinterpret_cast<const uint4*>(p + r * W) + c);
		}
	}
	fence_proxy_async_smem();
	tc_fence_before();
	__syncthreads();
	tc_fence_after();
	const uint32_t tmem = *tmem_slot;
	const uint32_t t_acc = tmem;                 // chain accumulator, 64 columns (128 if WIN: dL/d(input) is up to 128 wide)
	const uint32_t t_dw_in = tmem + (WIN ? 128 : 64); // D[in][out]   (transposed)
	const uint32_t t_dw_hid = t_dw_in + 64;      // D[out][in], 64 columns per hidden matmul
	const uint32_t t_dw_out = t_dw_hid + 64 * (H - 1); // D[out(16 valid)][in]
	const uint32_t lane_off = (warp * 32) << 16; // this warp's TMEM lane window

	// descriptors (constant parts)
	constexpr uint32_t SBO = 1024, LBO_K = 16, LBO_MN = TILE_BYTES;
	const uint32_t act_addr = smem_u32(act), g_addr = smem_u32(gbuf);
	const uint32_t idesc_fwd64 = make_idesc_f16(128, 64, 0, 0, ACC_F32);
	const uint32_t idesc_fwd16 = make_idesc_f16(128, 16, 0, 0, ACC_F32);
	const uint32_t idesc_dgrad = make_idesc_f16(128, 64, 0, 1, ACC_F32);   // A K-major, B = weight tile viewed MN-major
	const uint32_t idesc_dx = make_idesc_f16(128, E, 0, 0, ACC_F32);       // B = W_inT K-major
	const uint32_t idesc_dw = make_idesc_f16(128, 64, 1, 1, ACC_F32);      // both operands MN-major, K = 128 samples

	// ---- scatter warpgroups (WS) ---------------------------------------------------------------------
	if constexpr (WS) {
		if (tid >= 128) {
			if constexpr (REG_SCATTER < REG_LAUNCH) setmaxnreg_dec<REG_SCATTER>();
			const uint32_t srow = tid & 127, sgroup = (tid >> 7) - 1; // row of the tile, and which chunks of the row: j = sgroup (mod NSG)
			const uint32_t cpr_s = E / 8, swz = (srow * cpr_s / 8) & (cpr_s - 1);
			const uint32_t slot_row = smem_u32(dx_slot) + srow * E * 2;
			const uint32_t rep = blockIdx.x & (gp.n_replicas - 1);
			float nxt[FD];
#pragma unroll
			for (int d = 0; d < FD; ++d) nxt[d] = ldg_f32_here(a.positions + ((size_t)blockIdx.x * MLP_TILE + srow) * FD + d);
			named_bar_arrive(3, NT); // the slot starts empty
			for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
				float pos[FD];
#pragma unroll
				for (int d = 0; d < FD; ++d) pos[d] = nxt[d];
				if (tile + gridDim.x < a.n_tiles) {
#pragma unroll
					for (int d = 0; d < FD; ++d) nxt[d] = ldg_f32_here(a.positions + ((size_t)(tile + gridDim.x) * MLP_TILE + srow) * FD + d);
				}
				named_bar_sync(2, NT); // slot full: this tile's dL/d(encoded)
#pragma unroll kScatterChunkUnroll
				for (uint32_t j = sgroup; j < cpr_s; j += NSG) // (nearly) rolled: the level constants stay in the constant bank
					scatter_chunk<FD, FF>(gp, pos, lds128(slot_row + ((j ^ swz) * 16)), j, a.grid_grad, a.replicas, rep);
				named_bar_arrive(3, NT); // slot consumed: the compute warpgroup may refill it
			}
			return;
		}
		if constexpr (REG_COMPUTE > REG_LAUNCH) setmaxnreg_inc<REG_COMPUTE>();
	}
	uint32_t phase = 0, phase2 = 0;
	bool pending2 = false; // a commit on bar2 has not been waited for yet
	uint32_t first_tile = 1;
	[[maybe_unused]] uint32_t dbg_n = 0;
	STAGE_CLOCK(a); // 0: setup done
	float loss_acc = 0.f;
	const uint32_t row = tid;

	// input tile prefetch: 128 rows x cpr 16-byte chunks, chunk q = tid + 128 j (coalesced), held in registers one tile ahead
	const uint32_t cpr = E / 8;
	constexpr uint32_t XR = WIN ? 16 : 8;
	uint4 xr[XR];
	auto act_tile = [&](uint32_t k) { return act_addr + (k ? ACT0_EXTRA : 0u) + k * TILE_BYTES; };
	auto load_x = [&](uint32_t tile) {
		const uint4* src = reinterpret_cast<const uint4*>(a.input + (size_t)tile * MLP_TILE * E);
#pragma unroll
		for (uint32_t j = 0; j < XR; ++j) if (j < cpr) xr[j] = ldg128_here(src + tid + 128 * j);
	};
	if (blockIdx.x < a.n_tiles) load_x(blockIdx.x);

	for (uint32_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
		const size_t sample0 = (size_t)tile * MLP_TILE;
		// ---- input tile -> ACT[0] ------------------------------------------------------------------
		if (pending2) { mbar_wait(bar2, phase2); phase2 ^= 1; pending2 = false; } // previous tile's dW_in MMA still reads ACT[0]
#pragma unroll
		for (uint32_t j = 0; j < XR; ++j) {
			if (j < cpr) {
				const uint32_t q = tid + 128 * j, r = q / cpr, c = q % cpr;
				sts128(act_addr + (c >> 3) * TILE_BYTES + sw128(r, c & 7), xr[j]);
			}
		}
		STAGE_CLOCK(a); // E: epilogue math + sts issued
		fence_proxy_async_smem();
		tc_fence_before();
		CTA_SYNC();
		STAGE_CLOCK(a); // S: fenced + synced
		if (tile + gridDim.x < a.n_tiles) load_x(tile + gridDim.x); // in flight during the whole chain of this tile
		// per-row side inputs, requested now and consumed after the forward chain
		float tgt[4] = {0.f, 0.f, 0.f, 0.f};
		uint4 dy_in[2];
		[[maybe_unused]] float pos[FUSE_SCATTER ? FD : 1];
		if constexpr (FUSE_SCATTER && !WS) {
#pragma unroll
			for (int d = 0; d < FD; ++d) pos[d] = ldg_f32_here(a.positions + (sample0 + row) * FD + d);
		}
		if (MODE == MODE_TRAIN) {
#pragma unroll
			for (uint32_t j = 0; j < 4; ++j) if (j < a.n_out) tgt[j] = ldg_f32_here(a.target + (sample0 + row) * a.n_out + j);
		}
		if (MODE == MODE_BACKWARD) {
			const uint4* d4 = reinterpret_cast<const uint4*>(a.dL_doutput + (sample0 + row) * MLP_OUT);
			dy_in[0] = ldg128_here(d4); dy_in[1] = ldg128_here(d4 + 1);
		}

		// ---- forward hidden layers -----------------------------------------------------------------
#pragma unroll
		for (uint32_t k = 0; k < H; ++k) {
			if (tid == 0) {
				tc_fence_after();
				const uint32_t ksteps = (k == 0 ? E : 64u) / 16;
				const uint32_t a_addr = act_tile(k);
				const uint32_t b_addr = k == 0 ? smem_u32(w_in) : smem_u32(w_hid) + (k - 1) * 8192;
				for (uint32_t s = 0; s < ksteps; ++s) // K steps 4.. (layer 0, WIN) live in the second sub-tiles of ACT[0] / W_in
					umma_f16(t_acc, make_desc_sw128(a_addr + (s >> 2) * TILE_BYTES + (s & 3) * 32, LBO_K, SBO),
						make_desc_sw128(b_addr + (s >> 2) * 8192 + (s & 3) * 32, LBO_K, SBO), idesc_fwd64, s > 0);
				umma_commit(bar);
			}
			mbar_wait(bar, phase); phase ^= 1;
			STAGE_CLOCK(a); // W: MMA complete
			tc_fence_after();
			const uint32_t dst = act_tile(k + 1);
			AccRow<32> acc;
			acc_issue(t_acc + lane_off, acc);
			acc_finish(acc);
#pragma unroll
			for (int c = 0; c < 8; ++c) {
				uint32_t pk[4];
#pragma unroll
				for (int j = 0; j < 4; ++j) pk[j] = hidden_act_forward<GEN>(acc.h[c * 4 + j], a.activation);
				sts128(dst + sw128(row, c), pk[0], pk[1], pk[2], pk[3]);
			}
			STAGE_CLOCK(a); // E: epilogue math + sts issued
			fence_proxy_async_smem();
			tc_fence_before();
			CTA_SYNC();
			STAGE_CLOCK(a); // S: fenced + synced
		}
		// ---- output layer --------------------------------------------------------------------------
		if (tid == 0) {
			tc_fence_after();
			const uint32_t a_addr = act_tile(H), b_addr = smem_u32(w_out);
			for (uint32_t s = 0; s < 4; ++s)
				umma_f16(t_acc, make_desc_sw128(a_addr + s * 32, LBO_K, SBO), make_desc_sw128(b_addr + s * 32, LBO_K, SBO), idesc_fwd16, s > 0);
			umma_commit(bar);
		}
		mbar_wait(bar, phase); phase ^= 1;
		STAGE_CLOCK(a); // W: MMA complete
		tc_fence_after();
		{
			AccRow<8> acc;
			acc_issue(t_acc + lane_off, acc);
			acc_finish(acc);
			uint32_t yk[8];
#pragma unroll
			for (int j = 0; j < 8; ++j) yk[j] = acc.h[j];
			if (a.output_activation != ACT_NONE) { // on the fp16 pre-activation, all 16 padded columns like the reference
#pragma unroll
				for (int j = 0; j < 8; ++j) yk[j] = act_forward_h2(yk[j], a.output_activation);
			}
			const size_t s = sample0 + row;
			if (MODE != MODE_BACKWARD && a.output) {
				uint4* o = reinterpret_cast<uint4*>(a.output + s * MLP_OUT);
				o[0] = make_uint4(yk[0], yk[1], yk[2], yk[3]);
				o[1] = make_uint4(yk[4], yk[5], yk[6], yk[7]);
			}
			if (MODE == MODE_INFER && a.output_f32) {
				const __half* yh = reinterpret_cast<const __half*>(yk);
#pragma unroll
				for (uint32_t j = 0; j < (uint32_t)MLP_OUT; ++j) // unrolled with a predicate: a run-time loop bound would index yk[] dynamically (local memory)
					if (j < a.n_out) a.output_f32[s * a.n_out + j] = __half2float(yh[j]);
			}
			if (MODE == MODE_TRAIN) {
				// loss on the fp16 prediction (the reference's network output is fp16), gradients as fp16 * loss_scale.
				// Same operation order as losses/l2.h:62-74 / relative_l2.h:62-75.
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
				if (a.n_out > 4) { // rare: more than 4 outputs (targets not prefetched); statically indexed to stay in registers
#pragma unroll
					for (uint32_t j = 4; j < 16; j += 2) {
						const float ga = (j < a.n_out) ? loss_elem(__half2float(yh[j]), __ldg(a.target + s * a.n_out + j)) : 0.f;
						const float gb = (j + 1 < a.n_out) ? loss_elem(__half2float(yh[j + 1]), __ldg(a.target + s * a.n_out + j + 1)) : 0.f;
						dyk[j / 2] = pack_h2(ga, gb);
					}
				}
				if (a.output_activation != ACT_NONE) { // activation_backward_output_gpu (fully_fused_mlp.cu:757-762): derivative from the activated output
#pragma unroll
					for (int j = 0; j < 8; ++j) dyk[j] = act_backward_h2(dyk[j], yk[j], a.output_activation);
				}
				sts128(g_addr + sw128(row, 0), dyk[0], dyk[1], dyk[2], dyk[3]);
				sts128(g_addr + sw128(row, 1), dyk[4], dyk[5], dyk[6], dyk[7]);
			}
			if (MODE == MODE_BACKWARD) {
				uint32_t dq[8] = {dy_in[0].x, dy_in[0].y, dy_in[0].z, dy_in[0].w, dy_in[1].x, dy_in[1].y, dy_in[1].z, dy_in[1].w};
				if (a.output_activation != ACT_NONE) {
#pragma unroll
					for (int j = 0; j < 8; ++j) dq[j] = act_backward_h2(dq[j], yk[j], a.output_activation);
				}
				sts128(g_addr + sw128(row, 0), dq[0], dq[1], dq[2], dq[3]);
				sts128(g_addr + sw128(row, 1), dq[4], dq[5], dq[6], dq[7]);
			}
		}
		if constexpr (TRAIN) {
			STAGE_CLOCK(a); // E: epilogue math + sts issued
			fence_proxy_async_smem();
			tc_fence_before();
			CTA_SYNC();
			STAGE_CLOCK(a); // S: fenced + synced
			const bool want_dw = a.dW_partials != nullptr;
			uint32_t gcur = 0; // G[gcur] holds the gradient w.r.t. the pre-activation feeding weight matrix `k+1`
			// ---- backward through output matmul and hidden matmuls: k = H .. 1 ------------------------
#pragma unroll
			for (uint32_t k = H; k >= 1; --k) {
				// weight matrix index k: k == H -> W_out (K = 16 outputs), else W_hid[k-1] (K = 64)
				if (tid == 0) {
					tc_fence_after();
					const uint32_t ga = g_addr + gcur * TILE_BYTES;
					const uint32_t ksteps = (k == H) ? 1u : 4u;
					const uint32_t wb = (k == H) ? smem_u32(w_out) : smem_u32(w_hid) + (k - 1) * 8192;
					// dA_k[128 x 64] = G[128 x K] * W[K x 64]:  B = weight tile viewed MN-major (N = in, K = out rows)
					for (uint32_t s = 0; s < ksteps; ++s)
						umma_f16(t_acc, make_desc_sw128(ga + s * 32, LBO_K, SBO), make_desc_sw128(wb + s * 2048, LBO_MN, SBO), idesc_dgrad, s > 0);
					umma_commit(bar); // the epilogue only needs dA_k ...
					if (want_dw) {
						// ... dW[out][in] += G^T[out x 128] * ACT_k[128 x in] (batch is K: both tiles viewed MN-major) runs under
						// that epilogue; the tensor pipe executes in issue order, so the next stage's commit also covers it.
						const uint32_t td = (k == H) ? t_dw_out : t_dw_hid + (k - 1) * 64;
						const uint32_t ab = act_tile(k);
						for (uint32_t s = 0; s < 8; ++s)
							umma_f16(td, make_desc_sw128(ga + s * 2048, LBO_MN, SBO), make_desc_sw128(ab + s * 2048, LBO_MN, SBO), idesc_dw, (!first_tile) || s > 0);
					}
				}
				mbar_wait(bar, phase); phase ^= 1;
				STAGE_CLOCK(a); // W: MMA complete
				tc_fence_after();
				// ReLU' from this row's stored post-activation ACT[k] (fp16 > 0), applied to the packed fp16 gradient
				const uint32_t dst = g_addr + (gcur ^ 1) * TILE_BYTES;
				const uint32_t asrc = act_tile(k);
				AccRow<32> acc;
				acc_issue(t_acc + lane_off, acc);
				uint4 av[8];
#pragma unroll
				for (int c = 0; c < 8; ++c) av[c] = lds128(asrc + sw128(row, c)); // all in flight under the TMEM load
				acc_finish(acc);
#pragma unroll
				for (int c = 0; c < 8; ++c) {
					const uint32_t aw[4] = {av[c].x, av[c].y, av[c].z, av[c].w};
					uint32_t pk[4];
#pragma unroll
					for (int j = 0; j < 4; ++j) pk[j] = hidden_act_backward<GEN>(acc.h[c * 4 + j], aw[j], a.activation);
					sts128(dst + sw128(row, c), pk[0], pk[1], pk[2], pk[3]);
				}
				gcur ^= 1;
				STAGE_CLOCK(a); // E: epilogue math + sts issued
				fence_proxy_async_smem();
				tc_fence_before();
				CTA_SYNC();
				STAGE_CLOCK(a); // S: fenced + synced
			}
			// ---- input matmul: dW_in (transposed accumulator) and dL/d(input) ---------------------------
			const bool want_dx = a.dL_dinput != nullptr || FUSE_SCATTER;
			if (want_dw || want_dx) {
				if (tid == 0) {
					tc_fence_after();
					const uint32_t ga = g_addr + gcur * TILE_BYTES;
					if (want_dx)
						for (uint32_t s = 0; s < 4; ++s)
							umma_f16(t_acc, make_desc_sw128(ga + s * 32, LBO_K, SBO), make_desc_sw128(smem_u32(w_inT) + s * 32, LBO_K, SBO), idesc_dx, s > 0);
					umma_commit(bar);
					if (want_dw)
						for (uint32_t s = 0; s < 8; ++s) // D[in][out] += X^T[in x 128] * G[128 x out], under the dX epilogue
							umma_f16(t_dw_in, make_desc_sw128(act_addr + s * 2048, LBO_MN, SBO), make_desc_sw128(ga + s * 2048, LBO_MN, SBO), idesc_dw, (!first_tile) || s > 0);
					umma_commit(bar2); // waited for before ACT[0] / the accumulators are touched again
				}
				pending2 = true;
				mbar_wait(bar, phase); phase ^= 1;
				STAGE_CLOCK(a); // W: MMA complete
				tc_fence_after();
				if (want_dx) {
					__half* dst = a.dL_dinput ? a.dL_dinput + (sample0 + row) * E : nullptr;
					for (uint32_t c0 = 0; c0 < E; c0 += 16) {
						AccRow<8> acc;
						acc_issue(t_acc + lane_off + c0, acc);
						acc_finish(acc);
						uint32_t pk[8];
#pragma unroll
						for (int j = 0; j < 8; ++j) pk[j] = acc.h[j];
						if (dst) {
							reinterpret_cast<uint4*>(dst + c0)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
							reinterpret_cast<uint4*>(dst + c0)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
						}
						if constexpr (WS) {
							// hand the row to the scatter warpgroup through the shared-memory slot
							const uint32_t cpr_s = E / 8, swz = (row * cpr_s / 8) & (cpr_s - 1), slot_row = smem_u32(dx_slot) + row * E * 2;
							if (c0 == 0) named_bar_sync(3, NT); // slot empty: the scatter warpgroup is done with the previous tile
							sts128(slot_row + (((c0 / 8) ^ swz) * 16), make_uint4(pk[0], pk[1], pk[2], pk[3]));
							sts128(slot_row + (((c0 / 8 + 1) ^ swz) * 16), make_uint4(pk[4], pk[5], pk[6], pk[7]));
						} else if constexpr (FUSE_SCATTER) {
							// this row's dL/d(encoded) goes straight into the hash-table gradient: fire-and-forget reductions that
							// overlap the weight-gradient MMA still running and the other CTA's MMA chain
							const uint32_t rep = blockIdx.x & (gp.n_replicas - 1);
							scatter_chunk<FD, FF>(gp, pos, make_uint4(pk[0], pk[1], pk[2], pk[3]), c0 / 8, a.grid_grad, a.replicas, rep);
							scatter_chunk<FD, FF>(gp, pos, make_uint4(pk[4], pk[5], pk[6], pk[7]), c0 / 8 + 1, a.grid_grad, a.replicas, rep);
						}
					}
					if constexpr (WS) named_bar_arrive(2, NT); // slot full
				}
				tc_fence_before();
				CTA_SYNC();
			}
			first_tile = 0;
		} else {
			tc_fence_before();
			CTA_SYNC();
		}
	}

	// ---- per-CTA flush: weight-gradient partials and loss partial ---------------------------------------
	if constexpr (TRAIN) {
		if (pending2) { mbar_wait(bar2, phase2); phase2 ^= 1; pending2 = false; }
		tc_fence_after();
		if (a.dW_partials != nullptr) {
			float* out = a.dW_partials + (size_t)blockIdx.x * a.n_mlp_params;
			const bool have = !first_tile; // CTA processed at least one tile
			if constexpr (WIN) {
				if (warp >= 2) { // inputs 64..127 of dW_in: accumulator rows of warps 2 and 3
					const uint32_t r = warp * 32 + lane;
#pragma unroll
					for (int half_i = 0; half_i < 2; ++half_i) {
						uint32_t v[32];
						if (have) { tmem_ld32(t_dw_in + lane_off + half_i * 32, v); tmem_ld_wait(); }
						if (r < E) {
#pragma unroll
							for (int j = 0; j < 32; ++j) if (!GEN || half_i * 32 + j < (int)W) out[(half_i * 32 + j) * E + r] = have ? dw_value(v[j]) : 0.f;
						}
					}
				}
			}
			if (warp < 2) {
				const uint32_t r = warp * 32 + lane; // accumulator row 0..63
				// dW_in: D[in = r][out = j]  ->  grads[out * E + in]
#pragma unroll
				for (int half_i = 0; half_i < 2; ++half_i) {
					uint32_t v[32];
					if (have) { tmem_ld32(t_dw_in + lane_off + half_i * 32, v); tmem_ld_wait(); }
					if (r < E) {
#pragma unroll
						for (int j = 0; j < 32; ++j) if (!GEN || half_i * 32 + j < (int)W) out[(half_i * 32 + j) * E + r] = have ? dw_value(v[j]) : 0.f;
					}
				}
				float* o = out + W * E;
				for (uint32_t l = 0; l + 1 < H; ++l) {
#pragma unroll
					for (int half_i = 0; half_i < 2; ++half_i) {
						uint32_t v[32];
						if (have) { tmem_ld32(t_dw_hid + l * 64 + lane_off + half_i * 32, v); tmem_ld_wait(); }
						if constexpr (GEN) { // run-time width: the valid W x W corner of the 64 x 64 accumulator, rows of W floats
							if (r < W) {
#pragma unroll
								for (int j = 0; j < 32; ++j) if (half_i * 32 + j < (int)W) o[r * W + half_i * 32 + j] = have ? dw_value(v[j]) : 0.f;
							}
						} else {
							float4* o4 = reinterpret_cast<float4*>(o + r * 64 + half_i * 32);
#pragma unroll
							for (int j = 0; j < 8; ++j)
								o4[j] = have ? make_float4(dw_value(v[4 * j]), dw_value(v[4 * j + 1]), dw_value(v[4 * j + 2]), dw_value(v[4 * j + 3])) : make_float4(0, 0, 0, 0);
						}
					}
					o += W * W;
				}
				if (warp == 0) {
#pragma unroll
					for (int half_i = 0; half_i < 2; ++half_i) {
						uint32_t v[32];
						if (have) { tmem_ld32(t_dw_out + lane_off + half_i * 32, v); tmem_ld_wait(); }
						if (r < 16) {
							if constexpr (GEN) {
#pragma unroll
								for (int j = 0; j < 32; ++j) if (half_i * 32 + j < (int)W) o[r * W + half_i * 32 + j] = have ? dw_value(v[j]) : 0.f;
							} else {
								float4* o4 = reinterpret_cast<float4*>(o + r * 64 + half_i * 32);
#pragma unroll
								for (int j = 0; j < 8; ++j)
									o4[j] = have ? make_float4(dw_value(v[4 * j]), dw_value(v[4 * j + 1]), dw_value(v[4 * j + 2]), dw_value(v[4 * j + 3])) : make_float4(0, 0, 0, 0);
							}
						}
					}
				}
			}
		}
		if (MODE == MODE_TRAIN && a.loss_partials != nullptr) {
			float s = loss_acc;
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
			if (lane == 0) red_buf[warp] = s;
			CTA_SYNC();
			if (tid == 0) a.loss_partials[blockIdx.x] = red_buf[0] + red_buf[1] + red_buf[2] + red_buf[3];
		}
	}
	tc_fence_before();
	CTA_SYNC();
	if (warp == 0) tmem_dealloc(tmem, tmem_cols);
}

// Sum the per-CTA fp32 partials -> fp16 weight gradients (layout of the reference's gradient matrices) and the loss.
// Block = 32 parameters x 8 slices of the partial list (coalesced 128-byte rows), tree-reduced in shared memory.
__global__ void __launch_bounds__(256) mlp_finalize_kernel(uint32_t n_params, uint32_t n_parts, const float* __restrict__ partials, __half* __restrict__ grads,
	const float* __restrict__ loss_partials, float* __restrict__ loss_out, uint32_t n_loss_parts, int accumulate) {
	__shared__ float sh[8][33];
	const uint32_t px = threadIdx.x & 31, sl = threadIdx.x >> 5;
	const uint32_t i = blockIdx.x * 32 + px;
	float s = 0.f;
	if (i < n_params && grads) {
		for (uint32_t p = sl; p < n_parts; p += 8) s += partials[(size_t)p * n_params + i];
	}
	sh[sl][px] = s;
	__syncthreads();
	if (sl == 0 && i < n_params && grads) {
		float t = 0.f;
#pragma unroll
		for (int k = 0; k < 8; ++k) t += sh[k][px];
		if (accumulate) t += __half2float(grads[i]); // GradientMode::Accumulate: D = A B + C in one rounding, like the reference's GEMM epilogue with beta = 1
		grads[i] = __float2half(t);
	}
	if (blockIdx.x == 0 && loss_partials && loss_out) {
		__syncthreads();
		float l = 0.f;
		for (uint32_t p = threadIdx.x; p < n_loss_parts; p += blockDim.x) l += loss_partials[p];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xFFFFFFFFu, l, o);
		if (px == 0) sh[sl][0] = l;
		__syncthreads();
		if (threadIdx.x == 0) {
			float t = 0.f;
			for (int w = 0; w < 8; ++w) t += sh[w][0];
			*loss_out = t;
		}
	}
}

} // namespace b200
