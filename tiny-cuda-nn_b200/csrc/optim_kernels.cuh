// Fused Adam + fp32 master weights, and the reference-compatible parameter initialisation, for sm_100a.
//
// Semantics follow the reference's adam_step (include/tiny-cuda-nn/optimizers/adam.h:47-119): unscale by loss_scale,
// L2 regularisation on matrix (MLP) weights only, skip of hash-grid entries whose gradient is exactly zero, a
// per-parameter step counter for the bias correction, AdaBound clamp, fp32 master + fp16 copy.  New here:
// 4 parameters per thread with 8/16-byte accesses, entries whose four gradients are all zero touch nothing but the
// 8 gradient bytes, the gradient is re-zeroed in the same pass (the reference spends a cudaMemsetAsync over the whole
// grid gradient every step, encodings/grid.h:857-859), and the 1/world_size of data-parallel training is folded in.
#pragma once
#include "ptx.cuh"

namespace b200 {

struct AdamArgs {
	uint32_t n;                   // end of the parameter range handled by this launch
	uint32_t begin;               // start of the range (multiple of 4)
	uint32_t n_matrix;            // leading MLP weights (L2 reg, never skipped)
	float loss_scale;             // gradient divisor: loss_scale * world_size
	float learning_rate, beta1, beta2, epsilon, l2_reg;
	float relative_decay, absolute_decay;
	float lower_lr_bound, upper_lr_bound;
	float weight_clip;
	float non_matrix_lr_factor;
	float inv_loss_scale;         // 1 / loss_scale if that is exact (power of two), else 0 => divide
	float log2_beta1, log2_beta2; // for beta^t = exp2f(t * log2(beta)) on parameters whose step lags the global step
	uint32_t global_step;         // step count of parameters updated every step (after this update)
	float global_correction;      // sqrtf(1 - beta2^t) / (1 - beta1^t) at t = global_step
	int zero_grads;               // re-zero the non-matrix gradients after consuming them
	int optimize_matrix, optimize_non_matrix; // adam.h:77-85: 0 freezes the MLP weights / the grid tables (their state does not advance)
	float* master;
	__half* params;
	__half* grads;
	float* m;
	float* v;
	uint32_t* steps;
	// single-GPU fused step: the gradient replicas of the coarse grid levels (tcnn_b200_api.cu backward_params) are summed here
	// instead of by grid_replica_reduce_kernel - one launch less on the critical path.  Parameters [rep_begin, rep_end) are the
	// replicated levels (a prefix of the grid table); replica r holds their gradients at replicas + r * (rep_end - rep_begin).
	const __half* replicas;
	uint32_t rep_begin, rep_end, n_replicas;
};

// Square root, reciprocal and 2^x of the optimizer: the SFU approximations (<= 2 ulp) unless built with -DTCNN_B200_ADAM_IEEE.
// adam_kernel was issue-bound as much as HBM-bound (ncu r02b: issue slots 62 % busy, DRAM 62 %): the IEEE division and square root
// expand to ~20 instructions per parameter and the per-parameter bias correction (every entry whose step counter lags the
// optimizer's, i.e. most of the fine levels) to ~40.  A 2-ulp error of the step size changes an update by 3e-7 of ITS size - four
// orders of magnitude below one fp32 ulp of the weight it is added to; test_adam_vs_oracle / the state-trajectory golden hold.
#ifdef TCNN_B200_ADAM_IEEE
__device__ __forceinline__ float adam_sqrt(float x) { return sqrtf(x); }
__device__ __forceinline__ float adam_div(float x, float y) { return x / y; }
__device__ __forceinline__ float adam_exp2(float x) { return exp2f(x); }
#else
__device__ __forceinline__ float adam_sqrt(float x) { float r; asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float adam_div(float x, float y) { return __fdividef(x, y); }
__device__ __forceinline__ float adam_exp2(float x) { float r; asm("ex2.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif

__device__ __forceinline__ bool adam_one(const AdamArgs& a, uint32_t i, float g_half, float& w, float& m, float& v, uint32_t& step) {
	float gradient = a.inv_loss_scale != 0.f ? g_half * a.inv_loss_scale : g_half / a.loss_scale;
	const bool is_matrix = i < a.n_matrix;
	if (is_matrix ? !a.optimize_matrix : (!a.optimize_non_matrix || gradient == 0.f)) return false;
	if (is_matrix) gradient += a.l2_reg * w;
	const float gradient_sq = gradient * gradient;
	m = a.beta1 * m + (1 - a.beta1) * gradient;
	v = a.beta2 * v + (1 - a.beta2) * gradient_sq;
	float lr = a.learning_rate;
	if (!is_matrix) lr *= a.non_matrix_lr_factor;
	step += 1;
	// adam.h:103-105: lr *= sqrtf(1 - powf(beta2, t)) / (1 - powf(beta1, t)).  Two powf per parameter made this kernel
	// SM-bound (ncu r01a: 85 % SM, 39 % DRAM); parameters in step with the optimizer use the host-evaluated factor.
	if (step == a.global_step) lr *= a.global_correction;
	else lr *= adam_div(adam_sqrt(1 - adam_exp2((float)step * a.log2_beta2)), 1 - adam_exp2((float)step * a.log2_beta1));
	const float eff = fminf(fmaxf(adam_div(lr, adam_sqrt(v) + a.epsilon), a.lower_lr_bound), a.upper_lr_bound);
	// weight_decay(relative * lr, absolute * lr, w)   (common_device.h:870-873)
	const float decayed = (1 - a.relative_decay * lr) * w - copysignf(a.absolute_decay * lr, w);
	float nw = decayed - eff * m;
	if (a.weight_clip != 0.f) nw = fminf(fmaxf(nw, -a.weight_clip), a.weight_clip);
	w = nw;
	return true;
}

__global__ void __launch_bounds__(256) adam_kernel(const __grid_constant__ AdamArgs a) {
	const uint32_t i0 = a.begin + (blockIdx.x * blockDim.x + threadIdx.x) * 4;
	if (i0 >= a.n) return;
	if (i0 + 4 <= a.n) {
		union { uint2 u; __half h[4]; } g;
		g.u = *reinterpret_cast<const uint2*>(a.grads + i0);
		if (i0 >= a.rep_begin && i0 < a.rep_end) { // same arithmetic as grid_replica_reduce_kernel: fp32 sum, one rounding
			float acc[4] = {0.f, 0.f, 0.f, 0.f};
			const size_t rep_halfs = a.rep_end - a.rep_begin;
			const __half* src = a.replicas + (i0 - a.rep_begin);
			for (uint32_t r0 = 0; r0 < a.n_replicas; r0 += 8) {
				union { uint2 u; __half2 h2[2]; } v[8];
#pragma unroll
				for (uint32_t k = 0; k < 8; ++k) v[k].u = *reinterpret_cast<const uint2*>(src + (size_t)(r0 + k) * rep_halfs);
#pragma unroll
				for (uint32_t k = 0; k < 8; ++k) {
					const float2 lo = __half22float2(v[k].h2[0]), hi = __half22float2(v[k].h2[1]);
					acc[0] += lo.x; acc[1] += lo.y; acc[2] += hi.x; acc[3] += hi.y;
				}
			}
#pragma unroll
			for (int k = 0; k < 4; ++k) g.h[k] = __float2half_rn(__half2float(g.h[k]) + acc[k]);
		}
		const bool any_nonzero = ((g.u.x | g.u.y) & 0x7FFF7FFFu) != 0;
		if (i0 >= a.n_matrix && !any_nonzero) return; // nothing to do for four untouched grid entries
		float4 w = *reinterpret_cast<float4*>(a.master + i0);
		float4 m = *reinterpret_cast<float4*>(a.m + i0);
		float4 v = *reinterpret_cast<float4*>(a.v + i0);
		uint4 s = *reinterpret_cast<uint4*>(a.steps + i0);
		adam_one(a, i0 + 0, __half2float(g.h[0]), w.x, m.x, v.x, s.x);
		adam_one(a, i0 + 1, __half2float(g.h[1]), w.y, m.y, v.y, s.y);
		adam_one(a, i0 + 2, __half2float(g.h[2]), w.z, m.z, v.z, s.z);
		adam_one(a, i0 + 3, __half2float(g.h[3]), w.w, m.w, v.w, s.w);
		*reinterpret_cast<float4*>(a.master + i0) = w;
		*reinterpret_cast<float4*>(a.m + i0) = m;
		*reinterpret_cast<float4*>(a.v + i0) = v;
		*reinterpret_cast<uint4*>(a.steps + i0) = s;
		uint2 ph;
		ph.x = pack_h2(w.x, w.y); ph.y = pack_h2(w.z, w.w);
		*reinterpret_cast<uint2*>(a.params + i0) = ph;
		if (a.zero_grads && i0 >= a.n_matrix) *reinterpret_cast<uint2*>(a.grads + i0) = make_uint2(0, 0);
	} else {
		for (uint32_t i = i0; i < a.n; ++i) {
			float w = a.master[i], m = a.m[i], v = a.v[i];
			uint32_t s = a.steps[i];
			if (adam_one(a, i, __half2float(a.grads[i]), w, m, v, s)) {
				a.master[i] = w; a.m[i] = m; a.v[i] = v; a.steps[i] = s; a.params[i] = __float2half(w);
			}
			if (a.zero_grads && i >= a.n_matrix) a.grads[i] = __float2half(0.f);
		}
	}
}

// Reduction of the per-CTA weight-gradient partials of the MLP kernel (see mlp_finalize_kernel) with the Adam update of the
// MLP weights applied in the same pass: on one GPU the 7 K weights need no separate optimizer launch on the critical path.
// The gradient is rounded to fp16 first (it is also stored), so the update equals adam_kernel on the stored gradient.
__global__ void __launch_bounds__(256) mlp_finalize_adam_kernel(uint32_t n_params, uint32_t n_parts, const float* __restrict__ partials,
	const float* __restrict__ loss_partials, float* __restrict__ loss_out, const __grid_constant__ AdamArgs a, uint32_t n_loss_parts) {
	__shared__ float sh[8][33];
	const uint32_t px = threadIdx.x & 31, sl = threadIdx.x >> 5;
	const uint32_t i = blockIdx.x * 32 + px;
	float s = 0.f;
	if (i < n_params) for (uint32_t p = sl; p < n_parts; p += 8) s += partials[(size_t)p * n_params + i];
	sh[sl][px] = s;
	__syncthreads();
	if (sl == 0 && i < n_params) {
		float t = 0.f;
#pragma unroll
		for (int k = 0; k < 8; ++k) t += sh[k][px];
		const __half gh = __float2half(t);
		a.grads[i] = gh;
		float w = a.master[i], m = a.m[i], v = a.v[i];
		uint32_t st = a.steps[i];
		adam_one(a, i, __half2float(gh), w, m, v, st); // i < n_matrix: never skipped
		a.master[i] = w; a.m[i] = m; a.v[i] = v; a.steps[i] = st; a.params[i] = __float2half(w);
	}
	if (blockIdx.x == 0 && loss_partials && loss_out) {
		__syncthreads();
		float l = 0.f;
		for (uint32_t p = threadIdx.x; p < n_loss_parts; p += blockDim.x) l += loss_partials[p];
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xFFFFFFFFu, l, o);
		if ((threadIdx.x & 31) == 0) sh[0][threadIdx.x >> 5] = l;
		__syncthreads();
		if (threadIdx.x == 0) {
			float t = 0.f;
			for (int k = 0; k < 8; ++k) t += sh[0][k];
			*loss_out = t;
		}
	}
}

// ---- data-parallel optimizer step over NVLink peer memory -----------------------------------------------
// One kernel = reduce-scatter of the fp16 gradients + Adam on this rank's shard of the parameters + all-gather of the
// updated fp16 parameters, with both inter-GPU barriers inside.  Gradient and fp16 parameter buffers are symmetric
// allocations mapped into every rank; fp32 master weights, moments and step counters exist only for the rank's own
// shard (ZeRO-1 style), so the optimizer's HBM traffic per GPU shrinks with the number of ranks.
//   entry:  every rank announces "my gradients are complete" in each peer's flag block and waits for all peers
//   body :  thread = 4 parameters: sum the 8-byte gradient slices of all ranks (peer loads, or ONE multimem.ld_reduce
//           through the NVSwitch), Adam, 8-byte store of the new fp16 parameters into all ranks (peer stores, or ONE
//           multimem.st)
//   exit :  the last CTA announces "my parameter writes are visible" and waits for the same from all peers, so that
//           when the kernel retires every rank may read its parameters and re-zero its gradients.
// The waits are on other GPUs' kernels, never on CTAs of this launch.  Ranks may be minutes apart (rank-0-only evaluation,
// checkpointing, a first-step module load, a profiler): a wait gives up only after DpComm::timeout_ns (default 10 min,
// TCNN_B200_DP_TIMEOUT_S), and then does NOT trap - a trap would kill the CUDA context of every rank - but writes the epoch
// into word DP_FLAG_ERROR of the rank's own flag block and lets the kernel return without touching the parameters; the host
// turns the flag into an error the next time it looks (Trainer.loss / Trainer.check_peers).
constexpr uint32_t DP_MAX_RANKS = 8;
constexpr uint32_t DP_FLAG_GRADS = 0, DP_FLAG_PARAMS = 8, DP_FLAG_COUNTER = 16, DP_FLAG_ERROR = 17, DP_FLAG_WORDS = 32;
struct DpComm {
	uint64_t timeout_ns;
	uint32_t rank, world, epoch, use_multicast;
	__half* peer_grads[DP_MAX_RANKS];
	__half* peer_params[DP_MAX_RANKS];
	uint32_t* peer_flags[DP_MAX_RANKS];
	__half* grads_mc;
	__half* params_mc;
};

// signal = release store, wait = acquire polling with back-off (hundreds of CTAs poll the same line), both system scope
__device__ __forceinline__ void dp_signal(uint32_t* flag, uint32_t epoch) { st_release_sys(flag, epoch); }
__device__ __forceinline__ bool dp_wait_flag(const uint32_t* flag, uint32_t epoch, uint64_t timeout_ns) {
	if ((int32_t)(ld_acquire_sys(flag) - epoch) >= 0) return true;
	const uint64_t t0 = globaltimer_ns();
	uint32_t backoff = 100;
	while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
		__nanosleep(backoff);
		if (backoff < 2000) backoff += 100; // long skews: poll the peer-written line less often
		if (globaltimer_ns() - t0 > timeout_ns) return false;
	}
	return true;
}

// Persistent: a few CTAs per SM walk the shard in rows of 256 groups (group = 4 consecutive parameters = one 8-byte
// gradient slice per rank), so the flag poll at entry and the system fence at exit are paid once per CTA, not once per
// 2048 parameters (r01d: 1.6x slower than the plain Adam kernel when they were).  MAXW = compiled bound of the world size.
template <bool MC, int MAXW>
__global__ void __launch_bounds__(256) dp_adam_kernel(const __grid_constant__ AdamArgs a, const __grid_constant__ DpComm c) {
	uint32_t* my_flags = c.peer_flags[c.rank];
#ifdef DP_STAMPS
	if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<uint64_t*>(my_flags)[12] = globaltimer_ns();
#endif
	// the scatter before this kernel is complete (stream order): thread r of the first CTA tells rank r
	if (blockIdx.x == 0 && threadIdx.x < c.world) dp_signal(c.peer_flags[threadIdx.x] + DP_FLAG_GRADS + c.rank, c.epoch);
	bool peers_ok = true;
	if (threadIdx.x < c.world) peers_ok = dp_wait_flag(my_flags + DP_FLAG_GRADS + threadIdx.x, c.epoch, c.timeout_ns);
	if (__syncthreads_or(!peers_ok)) { // a peer never arrived: leave the parameters alone, tell the host
		if (threadIdx.x == 0) st_relaxed_sys(my_flags + DP_FLAG_ERROR, c.epoch);
		return;
	}
#ifdef DP_STAMPS
	if (blockIdx.x == 0 && threadIdx.x == 0) reinterpret_cast<uint64_t*>(my_flags)[13] = globaltimer_ns();
#endif

#pragma unroll 1
	for (uint32_t i0 = a.begin + (blockIdx.x * 256 + threadIdx.x) * 4; i0 < a.n; i0 += gridDim.x * (256 * 4)) {
		float2 g01 = make_float2(0.f, 0.f), g23 = make_float2(0.f, 0.f);
		if (MC) {
			const uint2 r = multimem_ld_reduce_h2x2(c.grads_mc + i0);
			g01 = __half22float2(*reinterpret_cast<const __half2*>(&r.x));
			g23 = __half22float2(*reinterpret_cast<const __half2*>(&r.y));
		} else {
			uint2 raw[MAXW];
#pragma unroll
			for (int r = 0; r < MAXW; ++r) if (r < (int)c.world) raw[r] = ld_sys_v2(c.peer_grads[r] + i0); // all in flight at once
#pragma unroll
			for (int r = 0; r < MAXW; ++r) if (r < (int)c.world) { // fixed rank order, fp32 accumulation
				const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&raw[r].x)), hi = __half22float2(*reinterpret_cast<const __half2*>(&raw[r].y));
				g01.x += lo.x; g01.y += lo.y; g23.x += hi.x; g23.y += hi.y;
			}
		}
		const bool any_nonzero = g01.x != 0.f || g01.y != 0.f || g23.x != 0.f || g23.y != 0.f;
		if (i0 >= a.n_matrix && !any_nonzero) continue;
		float4 w = *reinterpret_cast<float4*>(a.master + i0);
		float4 m = *reinterpret_cast<float4*>(a.m + i0);
		float4 v = *reinterpret_cast<float4*>(a.v + i0);
		uint4 s = *reinterpret_cast<uint4*>(a.steps + i0);
		adam_one(a, i0 + 0, g01.x, w.x, m.x, v.x, s.x);
		adam_one(a, i0 + 1, g01.y, w.y, m.y, v.y, s.y);
		adam_one(a, i0 + 2, g23.x, w.z, m.z, v.z, s.z);
		adam_one(a, i0 + 3, g23.y, w.w, m.w, v.w, s.w);
		*reinterpret_cast<float4*>(a.master + i0) = w;
		*reinterpret_cast<float4*>(a.m + i0) = m;
		*reinterpret_cast<float4*>(a.v + i0) = v;
		*reinterpret_cast<uint4*>(a.steps + i0) = s;
		const uint2 ph = make_uint2(pack_h2(w.x, w.y), pack_h2(w.z, w.w));
		if (MC) multimem_st_v2(c.params_mc + i0, ph);
		else {
#pragma unroll
			for (int r = 0; r < MAXW; ++r) if (r < (int)c.world) st_sys_v2(c.peer_params[r] + i0, ph);
		}
	}
	// exit: count the CTA as done (its peer stores are ordered before the increment by barrier + release); the last CTA
	// tells every rank and waits for every rank, thread r taking rank r
	__shared__ uint32_t is_last;
	__syncthreads();
	if (threadIdx.x == 0) is_last = atom_inc_acq_rel_gpu(my_flags + DP_FLAG_COUNTER, gridDim.x - 1) == gridDim.x - 1; // wraps to 0 for the next launch
	__syncthreads();
	if (is_last && threadIdx.x < c.world) {
#ifdef DP_STAMPS
		if (threadIdx.x == 0) reinterpret_cast<uint64_t*>(my_flags)[14] = globaltimer_ns();
#endif
		dp_signal(c.peer_flags[threadIdx.x] + DP_FLAG_PARAMS + c.rank, c.epoch);
		if (!dp_wait_flag(my_flags + DP_FLAG_PARAMS + threadIdx.x, c.epoch, c.timeout_ns)) st_relaxed_sys(my_flags + DP_FLAG_ERROR, c.epoch);
	}
#ifdef DP_STAMPS
	__syncthreads();
	if (is_last && threadIdx.x == 0) reinterpret_cast<uint64_t*>(my_flags)[15] = globaltimer_ns();
#endif
}

// ---- parameter initialisation ---------------------------------------------------------------------
// pcg32 (PCG-XSH-RR 64/32) with O(log n) jump-ahead, as vendored by the reference (dependencies/pcg32/pcg32.h).
struct Pcg32Dev {
	uint64_t state, inc;
	__device__ void advance(uint64_t delta) {
		uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta >>= 1;
		}
		state = acc_mult * state + acc_plus;
	}
	__device__ uint32_t next_uint() {
		const uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		const uint32_t rot = (uint32_t)(old >> 59u);
		return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
	}
	__device__ float next_float() { return __uint_as_float((next_uint() >> 9) | 0x3f800000u) - 1.0f; }
};

// Xavier-uniform rows of one weight matrix; the reference does this on the host (gpu_matrix.h:284-299:
// next_float() * 2.0f * scale - scale, separate roundings), element e uses the e-th draw after `base`.
__global__ void init_xavier_kernel(uint32_t n, uint64_t state, uint64_t inc, uint64_t base, float scale, float* __restrict__ out) {
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= n) return;
	Pcg32Dev rng{state, inc};
	rng.advance(base + e);
	out[e] = __fsub_rn(__fmul_rn(__fmul_rn(rng.next_float(), 2.0f), scale), scale);
}

// Uniform(lower, upper) with the element -> draw mapping of the reference's generate_random_kernel
// (include/tiny-cuda-nn/random.h:39-63): thread i jumps ahead 4 i and writes element i + n_threads * j with draw j.
__global__ void init_uniform_kernel(uint32_t n, uint32_t n_threads, uint64_t state, uint64_t inc, uint64_t base, float lower, float upper, float* __restrict__ out) {
	const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= n) return;
	const uint32_t i = idx % n_threads, j = idx / n_threads;
	Pcg32Dev rng{state, inc};
	rng.advance(base + (uint64_t)i * 4 + j);
	out[idx] = rng.next_float() * (upper - lower) + lower;
}

__global__ void cast_f32_to_f16_kernel(size_t n, const float* __restrict__ in, __half* __restrict__ out) {
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) out[i] = __float2half(in[i]);
}

} // namespace b200
