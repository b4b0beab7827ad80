// Multiresolution hash-grid encoding for sm_100a: forward gather, backward scatter, index dump.
//
// Behaviour follows the reference's kernel_grid / kernel_grid_backward (include/tiny-cuda-nn/encodings/grid.h:48-320)
// and its index math (include/tiny-cuda-nn/common_device.h:631-718, 825-868); the kernels themselves are new:
//   * one thread owns (sample, 16-byte chunk of the encoded row) = 8/F consecutive levels, so all 2^D * 8/F gathers of
//     the chunk are in flight together and the result is one 16-byte store into the sample-major [B x E_pad] fp16 matrix
//     that the MLP kernel consumes as its K-major A operand (the reference writes SoA and re-reads it with wmma loads);
//   * per-level constants (scale, resolution, strides with the uint32 wrap of grid_index, hashed-or-not) are resolved
//     once on the host and passed by value.
#pragma once
#include "ptx.cuh"

namespace b200 {

constexpr int MAX_LEVELS = 32;
constexpr int MAX_DIMS = 4;

// Tile images (mlp_wide_kernels.cuh): the 128-neuron MLP reads its input as the byte image of its shared-memory tiles -
// 128 samples x 64 columns per 16 KB sub-tile, rows of 128 bytes, 16-byte chunks XOR-swizzled by (row & 7) - so that a tile
// moves with one TMA bulk copy.  Byte offset of 16-byte chunk c (8 halfs) of sample i, `subs` sub-tiles per tile:
constexpr uint32_t SUB_BYTES = 16384;
__host__ __device__ __forceinline__ size_t wide_image_chunk(size_t i, uint32_t c, uint32_t subs) {
	const uint32_t r = (uint32_t)(i & 127);
	return (i >> 7) * ((size_t)subs * SUB_BYTES) + (size_t)(c >> 3) * SUB_BYTES + r * 128u + (((c & 7u) ^ (r & 7u)) << 4);
}

struct GridLevel {
	float scale;            // exp2f(level * log2(per_level_scale)) * base_resolution - 1   (device-evaluated, see grid_scale_kernel)
	uint32_t resolution;    // (uint32_t)ceilf(scale) + 1
	uint32_t size;          // entries in this level (hashmap_size of grid_index)
	uint32_t offset;        // first entry of this level
	uint32_t stride[MAX_DIMS]; // dense strides incl. uint32 wrap; 0 for dims the reference loop never reaches
	uint32_t n_dense_dims;  // number of loop iterations of grid_index executed
	uint32_t hashed;        // 1 if (Hash grid && size < final stride)
	uint32_t replicated;    // backward: scatter into one of n_replicas private copies (coarse, heavily contended levels)
	uint32_t rep_offset;    // first entry of this level inside one replica
	uint64_t magic;         // ceil(2^64 / size): index % size without a division (fastmod_u32)
	uint32_t mode;          // LEVEL_HASH / LEVEL_DENSE_POW2 / LEVEL_DENSE_MOD: which of the three index computations applies (uniform per level)
	uint32_t hash_type;     // HASH_COHERENT_PRIME (the specialised kernels) / HASH_PRIME / HASH_REVERSED_PRIME / HASH_RNG (grid_index_dev only)
};
enum HashKind : uint32_t { HASH_COHERENT_PRIME = 0, HASH_PRIME = 1, HASH_REVERSED_PRIME = 2, HASH_RNG = 3 };
enum LevelMode : uint32_t { LEVEL_HASH = 0, LEVEL_DENSE_POW2 = 1, LEVEL_DENSE_MOD = 2 };

// a % d for 32-bit a, d with M = ceil(2^64 / d) (Lemire, Kaser, Kurz: "Faster remainder by direct computation", exact for all 32-bit
// operands).  The dense levels' sizes are not powers of two (grid_index takes `index % hashmap_size`, common_device.h:706); a
// hardware-less 32-bit modulo costs ~20 instructions and a branch around it, this costs four multiplies and no branch.
__host__ __device__ __forceinline__ uint64_t fastmod_magic(uint32_t d) { return 0xFFFFFFFFFFFFFFFFull / d + 1; }
__device__ __forceinline__ uint32_t fastmod_u32(uint32_t a, uint64_t M, uint32_t d) {
	// low 64 bits of M * a, then the high 32 bits of (that * d) >> 32: four 32-bit multiplies
	const uint32_t mlo = (uint32_t)M, mhi = (uint32_t)(M >> 32);
	const uint64_t t = (uint64_t)mlo * a;
	const uint32_t l_lo = (uint32_t)t, l_hi = (uint32_t)(t >> 32) + mhi * a;
	const uint64_t p = (uint64_t)l_lo * d;
	return (uint32_t)(((uint64_t)l_hi * d + (p >> 32)) >> 32);
}

struct GridParams {
	uint32_t n_levels;
	uint32_t n_features;     // n_levels * F
	uint32_t padded_width;   // encoded row width in halfs (multiple of 16)
	uint32_t n_replicas;     // power of two
	uint32_t rep_entries;    // entries per replica (sum of the replicated levels' sizes)
	uint32_t image_subs;     // 0: encoded rows [B][padded_width]; n > 0: tile images with n sub-tiles per tile (wide_image_chunk)
	uint32_t smoothstep;     // interpolation: 0 = Linear, 1 = Smoothstep (weights w -> w^2 (3 - 2 w), common_device.h:801-807)
	uint32_t stochastic;     // stochastic_interpolation (grid.h:284-303): the backward pass adds the whole gradient to ONE random corner
	uint64_t rng_state, rng_inc; // pcg32{1337} after seeding (random_val, common_device.h:333-337), filled by plan_create
	uint32_t nearest;        // interpolation Nearest (grid.h:121-139): own kernels below, never reaches the interpolating ones
	uint32_t n_levels_total; // levels of the encoding (n_levels above is the active prefix after set_max_level)
	const float* max_level_gpu; // per-sample max level (GridEncoding::set_max_level_gpu) or nullptr; read by the general kernels only
	GridLevel level[MAX_LEVELS];
};

// common_device.h:709-714, evaluated on the device with CUDA's exp2f exactly like the reference kernels do.
__device__ __forceinline__ float grid_scale_dev(uint32_t level, float log2_per_level_scale, uint32_t base_resolution) {
	return exp2f(level * log2_per_level_scale) * base_resolution - 1.0f;
}
__global__ void grid_scale_kernel(uint32_t n_levels, float log2_per_level_scale, uint32_t base_resolution, float* __restrict__ scales) {
	const uint32_t l = threadIdx.x;
	if (l < n_levels) scales[l] = grid_scale_dev(l, log2_per_level_scale, base_resolution);
}

// common_device.h:857-868 (identity interpolation function = Linear)
__device__ __forceinline__ void pos_fract_dev(float x, float scale, float& frac, uint32_t& cell) {
	const float p = fmaf(scale, x, 0.5f);
	const float f = floorf(p);
	cell = (uint32_t)(int)f;
	frac = p - f;
}

__device__ __forceinline__ float smoothstep_dev(float v) { return v * v * (3.0f - 2.0f * v); }
__device__ __forceinline__ float smoothstep_derivative_dev(float v) { return 6 * v * (1.0f - v); }

// pcg32 (pcg32.h:54-69, 145-168): the generator after `delta` steps from (state, inc), and its 32-bit output function
__device__ __forceinline__ uint64_t pcg32_state_at(uint64_t state, uint64_t inc, uint64_t delta) {
	uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
	while (delta > 0) {
		if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
		cur_plus = (cur_mult + 1) * cur_plus;
		cur_mult *= cur_mult;
		delta >>= 1;
	}
	return acc_mult * state + acc_plus;
}
__device__ __forceinline__ uint32_t pcg32_output(uint64_t old) {
	const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
	const uint32_t rot = (uint32_t)(old >> 59u);
	return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}

// grid_hash (common_device.h:631-688).  The three prime hashes differ in their factor tables; Rng draws the next output of
// pcg32{1337} advanced by xor_d(pos[d] << d * (64 / D)).
template <int D>
__device__ __forceinline__ uint32_t grid_hash_dev(uint32_t hash_type, const uint32_t (&pg)[D]) {
	constexpr uint32_t coherent[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
	uint32_t h = 0;
	if (hash_type == HASH_COHERENT_PRIME) {
#pragma unroll
		for (int d = 0; d < D; ++d) h ^= pg[d] * coherent[d];
	} else if (hash_type == HASH_PRIME) {
		constexpr uint32_t f[7] = {1958374283u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
#pragma unroll
		for (int d = 0; d < D; ++d) h ^= pg[d] * f[d];
	} else if (hash_type == HASH_REVERSED_PRIME) {
		constexpr uint32_t f[7] = {2165219737u, 1434869437u, 2097192037u, 3674653429u, 805459861u, 2654435761u, 1958374283u};
#pragma unroll
		for (int d = 0; d < D; ++d) h ^= pg[d] * f[d];
	} else {
		constexpr uint32_t bits = 64 / D;
		uint64_t step = 0;
#pragma unroll
		for (int d = 0; d < D; ++d) step ^= (uint64_t)pg[d] << (d * bits);
		// pcg32{1337}: seed(1337, 1) = state 0, inc 3, one step, += 1337, one step (pcg32.h:54-60)
		const uint64_t mult = 0x5851f42d4c957f2dULL, inc = 3u;
		uint64_t state = (0u * mult + inc + 1337u) * mult + inc;
		h = pcg32_output(pcg32_state_at(state, inc, step));
	}
	return h;
}

// common_device.h:690-707 with the level-constant part hoisted (strides, hashed flag); primes :650-654.
template <int D>
__device__ __forceinline__ uint32_t grid_index_dev(const GridLevel& lv, const uint32_t (&pg)[D]) {
	uint32_t index;
	if (lv.hashed) {
		index = grid_hash_dev<D>(lv.hash_type, pg);
	} else {
		index = 0;
#pragma unroll
		for (int d = 0; d < D; ++d) index += pg[d] * lv.stride[d]; // stride[d] == 0 for skipped dims
	}
	return fastmod_u32(index, lv.magic, lv.size);
}

// Entry indices of all 2^D corners of a cell (corner c: bit d of c selects cell[d] + 1), for a level whose index rule is known at
// compile time.  Same results as grid_index_dev per corner (uint32 wrap included), with the shared terms computed once:
// hash: (cell[d] + 1) * prime[d] = cell[d] * prime[d] + prime[d] (mod 2^32); dense: one base index plus stride sums.
template <int D, uint32_t MODE>
__device__ __forceinline__ void corner_indices(const GridLevel& lv, const uint32_t (&cell)[D], uint32_t (&idx)[1 << D]) {
	constexpr uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
	uint32_t term[D][2];
#pragma unroll
	for (int d = 0; d < D; ++d) {
		const uint32_t m = MODE == LEVEL_HASH ? primes[d] : lv.stride[d]; // stride[d] == 0 for dims the reference's loop never reaches
		term[d][0] = cell[d] * m;
		term[d][1] = term[d][0] + m;
	}
#pragma unroll
	for (int c = 0; c < (1 << D); ++c) {
		uint32_t v = term[0][c & 1];
#pragma unroll
		for (int d = 1; d < D; ++d) v = MODE == LEVEL_HASH ? (v ^ term[d][(c >> d) & 1]) : (v + term[d][(c >> d) & 1]);
		idx[c] = MODE == LEVEL_DENSE_MOD ? fastmod_u32(v, lv.magic, lv.size) : (v & (lv.size - 1));
	}
}

template <int F> struct FeatVec;
template <> struct FeatVec<1> { using type = __half; };
template <> struct FeatVec<2> { using type = __half2; };
template <> struct FeatVec<4> { using type = uint2; };
template <> struct FeatVec<8> { using type = uint4; };

template <int D>
__device__ __forceinline__ void load_position(const float* __restrict__ x, size_t i, float (&p)[D]) {
	if constexpr (D == 2) {
		const float2 v = __ldg(reinterpret_cast<const float2*>(x) + i);
		p[0] = v.x; p[1] = v.y;
	} else if constexpr (D == 4) {
		const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
		p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
	} else {
#pragma unroll
		for (int d = 0; d < D; ++d) p[d] = __ldg(x + i * D + d);
	}
}

#ifndef TCNN_B200_GATHER_QUAD
#define TCNN_B200_GATHER_QUAD 1
#endif
// One level of encode_chunk for a level whose index rule is a template parameter (uniform over the warp, see scatter_level).
// Writes the level's F interpolated halfs to out[0 .. F/2).
template <int D, int F, uint32_t MODE>
__device__ __forceinline__ void encode_level(const GridParams& gp, const GridLevel& lv, const __half* __restrict__ grid, const float (&x)[D], __half2* out) {
	constexpr int NC = 1 << D;
	float frac[D]; uint32_t cell[D];
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(x[d], lv.scale, frac[d], cell[d]);
	if (gp.smoothstep) {
#pragma unroll
		for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
	}
	const uint8_t* table = reinterpret_cast<const uint8_t*>(grid + (size_t)lv.offset * F);
	// issue every gather of the level before the fp16 fma chain (grid.h:146-163 order: corner 0..2^D-1)
	uint32_t idx[NC]; float w[NC];
	corner_indices<D, MODE>(lv, cell, idx);
#pragma unroll
	for (int c = 0; c < NC; ++c) {
		float weight = 1.0f;
#pragma unroll
		for (int d = 0; d < D; ++d) weight *= (c & (1 << d)) == 0 ? 1.0f - frac[d] : frac[d];
		w[c] = weight;
	}
	if constexpr (F == 2) {
		// Corners c and c+1 differ in dim 0 only.  For even cells their entries are the two halves of one aligned 8-byte
		// slot on every level (dense: index and index+1; coherent-prime hash: prime[0] = 1 so the indices differ in bit
		// 0 only), so one 8-byte load serves both; otherwise the second corner gets its own (predicated) load.
		__half2 val[NC];
#pragma unroll
		for (int c = 0; c < NC; c += 2) {
#if TCNN_B200_GATHER_QUAD
			// the gather is bound by L1 tag lookups (one per lane and load): a 16-byte load of the aligned slot
			// {idx & ~3 .. idx | 3} also brings the dim-0 neighbour unless idx % 4 == 3 -> 1.25 lookups per corner pair
			const uint32_t l0 = idx[c] & 3u, l1 = idx[c + 1] & 3u;
			const bool together = ((idx[c] ^ idx[c + 1]) & ~3u) == 0u;
			const uint4 a = __ldg(reinterpret_cast<const uint4*>(table + ((idx[c] >> 2) << 4)));
			const uint32_t own = l0 == 0 ? a.x : l0 == 1 ? a.y : l0 == 2 ? a.z : a.w;
			uint32_t other = l1 == 0 ? a.x : l1 == 1 ? a.y : l1 == 2 ? a.z : a.w;
			if (!together) other = __ldg(reinterpret_cast<const uint32_t*>(table + (idx[c + 1] << 2)));
#else
			const bool paired = ((idx[c] ^ idx[c + 1]) == 1u);
			const uint2 a = __ldg(reinterpret_cast<const uint2*>(table + ((idx[c] >> 1) << 3)));
			uint32_t own = (idx[c] & 1u) ? a.y : a.x, other = (idx[c] & 1u) ? a.x : a.y;
			if (!paired) other = __ldg(reinterpret_cast<const uint32_t*>(table + (idx[c + 1] << 2)));
#endif
			val[c] = *reinterpret_cast<const __half2*>(&own);
			val[c + 1] = *reinterpret_cast<const __half2*>(&other);
		}
		__half2 r = __floats2half2_rn(0.f, 0.f);
#pragma unroll
		for (int c = 0; c < NC; ++c) r = __hfma2(__half2half2(__float2half(w[c])), val[c], r);
		out[0] = r;
	} else if constexpr (F == 4) {
		uint2 val[NC];
#pragma unroll
		for (int c = 0; c < NC; c += 2) { // same pairing with 16-byte slots
			const bool paired = ((idx[c] ^ idx[c + 1]) == 1u);
			const uint4 a = __ldg(reinterpret_cast<const uint4*>(table + ((idx[c] >> 1) << 4)));
			val[c] = (idx[c] & 1u) ? make_uint2(a.z, a.w) : make_uint2(a.x, a.y);
			val[c + 1] = (idx[c] & 1u) ? make_uint2(a.x, a.y) : make_uint2(a.z, a.w);
			if (!paired) val[c + 1] = __ldg(reinterpret_cast<const uint2*>(table + (idx[c + 1] << 3)));
		}
		__half2 r0 = __floats2half2_rn(0.f, 0.f), r1 = r0;
#pragma unroll
		for (int c = 0; c < NC; ++c) {
			const __half2 wh = __half2half2(__float2half(w[c]));
			r0 = __hfma2(wh, *reinterpret_cast<__half2*>(&val[c].x), r0);
			r1 = __hfma2(wh, *reinterpret_cast<__half2*>(&val[c].y), r1);
		}
		out[0] = r0; out[1] = r1;
	} else if constexpr (F == 8) {
		uint4 val[NC];
#pragma unroll
		for (int c = 0; c < NC; ++c) val[c] = __ldg(reinterpret_cast<const uint4*>(table + (idx[c] << 4)));
		__half2 r[4];
#pragma unroll
		for (int k = 0; k < 4; ++k) r[k] = __floats2half2_rn(0.f, 0.f);
#pragma unroll
		for (int c = 0; c < NC; ++c) {
			const __half2 wh = __half2half2(__float2half(w[c]));
			const __half2* v = reinterpret_cast<const __half2*>(&val[c]);
#pragma unroll
			for (int k = 0; k < 4; ++k) r[k] = __hfma2(wh, v[k], r[k]);
		}
#pragma unroll
		for (int k = 0; k < 4; ++k) out[k] = r[k];
	}
}

// Encode the levels of one 16-byte chunk for one sample.  Returns 8 halfs packed in a uint4.
template <int D, int F>
__device__ __forceinline__ uint4 encode_chunk(const GridParams& gp, const __half* __restrict__ grid, const float (&x)[D], uint32_t chunk) {
	static_assert(F == 2 || F == 4 || F == 8, "packed half2 features");
	constexpr int LEVELS_PER_CHUNK = 8 / F;
	union { uint4 v; __half2 h2[4]; } out;
	out.v = make_uint4(0, 0, 0, 0);
#pragma unroll
	for (int s = 0; s < LEVELS_PER_CHUNK; ++s) {
		const uint32_t level = chunk * LEVELS_PER_CHUNK + s;
		if (level >= gp.n_levels) break; // padding columns stay zero (grid.h:749-759 pads with zeros too)
		const GridLevel& lv = gp.level[level];
		__half2* o = &out.h2[s * (F / 2)];
		if (lv.mode == LEVEL_HASH) encode_level<D, F, LEVEL_HASH>(gp, lv, grid, x, o);
		else if (lv.mode == LEVEL_DENSE_POW2) encode_level<D, F, LEVEL_DENSE_POW2>(gp, lv, grid, x, o);
		else encode_level<D, F, LEVEL_DENSE_MOD>(gp, lv, grid, x, o);
	}
	return out.v;
}

// grid = {ceil(B/256), E_pad/8}; chunk-major block order keeps one group of levels hot in L2 at a time.
// zero16 / n_zero16 (optional): 16-byte words this launch clears on the side (the gradient replicas of the coarse levels
// for the backward pass that follows; saves a memset launch per training step).
template <int D, int F>
__global__ void __launch_bounds__(256) grid_forward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ grid, __half* __restrict__ encoded, uint4* __restrict__ zero16, uint32_t n_zero16) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (zero16 != nullptr) {
		const uint32_t n_threads = gridDim.x * gridDim.y * blockDim.x;
		for (uint32_t k = blockIdx.y * gridDim.x * blockDim.x + i; k < n_zero16; k += n_threads) zero16[k] = make_uint4(0, 0, 0, 0);
	}
	if (i >= B) return;
	const uint32_t chunk = blockIdx.y;
	float p[D];
	load_position<D>(x, i, p);
	const uint4 v = encode_chunk<D, F>(gp, grid, p, chunk);
	const size_t off = gp.image_subs ? wide_image_chunk(i, chunk, gp.image_subs) : ((size_t)i * gp.padded_width + chunk * 8) * 2;
	*reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(encoded) + off) = v;
}

// Scatter one sample's 8-half chunk of dL/d(encoded) (fp16, loss-scaled) into the fp16 grid gradient.
// Same arithmetic as grid.h:248-318: product (half)weight * grad rounded to fp16, then red.add.noftz.f16x2.
// Shared by grid_backward_kernel and the fused MLP training kernel (mlp_kernels.cuh).
#ifndef TCNN_B200_SCATTER_QUAD
#define TCNN_B200_SCATTER_QUAD 1
#endif
// One level of scatter_chunk, for a level whose index rule (LevelMode) is a template parameter: the rule is uniform over the
// warp, so scatter_chunk branches once per level and the body carries one index computation instead of a predicated mix of all.
// g0 / g1: this level's F gradient halfs (g1 only for F >= 4, packed pairs 2, 3 in gx for F == 8).
template <int D, int F, uint32_t MODE, bool SMOOTH>
__device__ __forceinline__ void scatter_level(const GridParams& gp, const GridLevel& lv, const float (&p)[D], const __half2 (&g)[F / 2],
	__half* __restrict__ grid_grad, __half* __restrict__ replicas, uint32_t replica_id) {
	constexpr int NC = 1 << D;
	float frac[D]; uint32_t cell[D];
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
	if constexpr (SMOOTH) {
		if (gp.smoothstep) {
#pragma unroll
			for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
		}
	}
	// Same-address f16x2 reductions serialise in L2 (~80 ns each, measured: 2-D level 0 alone took 244 us at B = 2^18);
	// coarse levels therefore scatter into one of n_replicas copies chosen per block and are summed afterwards.
	uint8_t* gtable = reinterpret_cast<uint8_t*>(lv.replicated
		? replicas + ((size_t)replica_id * gp.rep_entries + lv.rep_offset) * F
		: grid_grad + (size_t)lv.offset * F);
	uint32_t idx[NC]; __half2 wh[NC];
	corner_indices<D, MODE>(lv, cell, idx);
#pragma unroll
	for (int c = 0; c < NC; ++c) {
		float weight = 1.0f;
#pragma unroll
		for (int d = 0; d < D; ++d) weight *= (c & (1 << d)) == 0 ? 1.0f - frac[d] : frac[d];
		wh[c] = __half2half2(__float2half(weight));
	}
	// byte offset of entry i inside the level: a level holds at most 2^28 entries of at most 16 bytes (checked by plan_create), so
	// 32-bit offsets + one widening add per reduction
	auto at = [&](uint32_t byte_off) { return reinterpret_cast<__half*>(gtable + byte_off); };
	// One vector reduction covers an aligned slot (see the forward kernel): for even cells it carries both dim-0 neighbours, for
	// odd cells the other lanes add +0 and the neighbour gets its own (predicated) reduction: fewer L2 reduction operations per
	// corner pair (the scatter is bound by the L2's reduction rate).
	if constexpr (F == 2) {
#pragma unroll
		for (int c = 0; c < NC; c += 2) {
			const uint32_t v0 = h2_bits(__hmul2(wh[c], g[0])), v1 = h2_bits(__hmul2(wh[c + 1], g[0]));
#if TCNN_B200_SCATTER_QUAD
			// 16-byte slot {idx & ~3 .. idx | 3}: holds the dim-0 neighbour unless idx % 4 == 3 -> 1.25 reductions per pair
			const uint32_t l0 = idx[c] & 3u, l1 = idx[c + 1] & 3u;
			const bool together = ((idx[c] ^ idx[c + 1]) & ~3u) == 0u && l0 != l1;
			const uint32_t t1 = together ? l1 : 4u;
			const uint32_t q0 = l0 == 0 ? v0 : (t1 == 0 ? v1 : 0u), q1 = l0 == 1 ? v0 : (t1 == 1 ? v1 : 0u);
			const uint32_t q2 = l0 == 2 ? v0 : (t1 == 2 ? v1 : 0u), q3 = l0 == 3 ? v0 : (t1 == 3 ? v1 : 0u);
			red_add_h2x4(at((idx[c] >> 2) << 4), q0, q1, q2, q3);
			red_add_h2_if(!together, at(idx[c + 1] << 2), v1);
#else
			const bool paired = ((idx[c] ^ idx[c + 1]) == 1u), odd = (idx[c] & 1u) != 0;
			const uint32_t nb = paired ? v1 : 0u;
			red_add_h2x2(at((idx[c] >> 1) << 3), odd ? nb : v0, odd ? v0 : nb);
			red_add_h2_if(!paired, at(idx[c + 1] << 2), v1);
#endif
		}
	} else if constexpr (F == 4) {
#pragma unroll
		for (int c = 0; c < NC; c += 2) {
			const uint32_t a0 = h2_bits(__hmul2(wh[c], g[0])), a1 = h2_bits(__hmul2(wh[c], g[1]));
			const uint32_t b0 = h2_bits(__hmul2(wh[c + 1], g[0])), b1 = h2_bits(__hmul2(wh[c + 1], g[1]));
			const bool paired = ((idx[c] ^ idx[c + 1]) == 1u), odd = (idx[c] & 1u) != 0;
			const uint32_t n0 = paired ? b0 : 0u, n1 = paired ? b1 : 0u;
			red_add_h2x4(at((idx[c] >> 1) << 4), odd ? n0 : a0, odd ? n1 : a1, odd ? a0 : n0, odd ? a1 : n1);
			red_add_h2x2_if(!paired, at(idx[c + 1] << 3), b0, b1);
		}
	} else { // F == 8: one 16-byte reduction per corner
#pragma unroll
		for (int c = 0; c < NC; ++c)
			red_add_h2x4(at(idx[c] << 4), h2_bits(__hmul2(wh[c], g[0])), h2_bits(__hmul2(wh[c], g[1])),
				h2_bits(__hmul2(wh[c], g[2])), h2_bits(__hmul2(wh[c], g[3])));
	}
}

// SMOOTH = false compiles the Smoothstep branch out: the fused MLP kernel (Linear only) keeps its register budget.
template <int D, int F, bool SMOOTH = false>
__device__ __forceinline__ void scatter_chunk(const GridParams& gp, const float (&p)[D], uint4 gv, uint32_t chunk,
	__half* __restrict__ grid_grad, __half* __restrict__ replicas, uint32_t replica_id) {
	static_assert(F == 2 || F == 4 || F == 8, "fp16 gradient scatter needs F >= 2 (grid.h:655-666)");
	constexpr int LEVELS_PER_CHUNK = 8 / F;
	union { uint4 v; __half2 h2[4]; } g;
	g.v = gv;
	// Rolled on purpose: the three index-rule variants of a level are ~120 instructions each; unrolled over the levels of a chunk the
	// scatter was 5 000 instructions of the fused kernel and ran out of the instruction cache.  The chunk's gradient halfs rotate
	// through the low words instead of being indexed by the loop counter.
#pragma unroll 1
	for (int s = 0; s < LEVELS_PER_CHUNK; ++s) {
		const uint32_t level = chunk * LEVELS_PER_CHUNK + s;
		if (level >= gp.n_levels) break;
		const GridLevel& lv = gp.level[level];
		__half2 gl[F / 2];
#pragma unroll
		for (int k = 0; k < F / 2; ++k) gl[k] = g.h2[k];
		if (lv.mode == LEVEL_HASH) scatter_level<D, F, LEVEL_HASH, SMOOTH>(gp, lv, p, gl, grid_grad, replicas, replica_id);
		else if (lv.mode == LEVEL_DENSE_POW2) scatter_level<D, F, LEVEL_DENSE_POW2, SMOOTH>(gp, lv, p, gl, grid_grad, replicas, replica_id);
		else scatter_level<D, F, LEVEL_DENSE_MOD, SMOOTH>(gp, lv, p, gl, grid_grad, replicas, replica_id);
		if constexpr (F == 2) g.v = make_uint4(g.v.y, g.v.z, g.v.w, g.v.x);
		else if constexpr (F == 4) g.v = make_uint4(g.v.z, g.v.w, g.v.x, g.v.y);
	}
}

template <int D, int F>
__global__ void __launch_bounds__(256) grid_backward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ dL_denc, __half* __restrict__ grid_grad, __half* __restrict__ replicas, uint32_t chunk_begin) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t chunk = chunk_begin + blockIdx.y;
	float p[D];
	load_position<D>(x, i, p);
	const uint4 gv = __ldg(reinterpret_cast<const uint4*>(dL_denc + (size_t)i * gp.padded_width + chunk * 8));
	scatter_chunk<D, F, true>(gp, p, gv, chunk, grid_grad, replicas, blockIdx.x & (gp.n_replicas - 1));
}

// grid_grad[level entries] += sum over replicas (fp32 accumulate, one rounding), for the replicated levels.
// Covers the half2 range [pair_begin, pair_end) of one replica (= the replicated levels of one level group).
__global__ void grid_replica_reduce_kernel(const __grid_constant__ GridParams gp, uint32_t F, const __half* __restrict__ replicas, __half* __restrict__ grid_grad,
	uint32_t pair_begin, uint32_t pair_end) {
	const uint32_t i = pair_begin + blockIdx.x * blockDim.x + threadIdx.x; // half2 index inside one replica
	const uint32_t pairs = gp.rep_entries * F / 2;
	if (i >= pair_end) return;
	float2 acc = make_float2(0.f, 0.f);
	for (uint32_t r0 = 0; r0 < gp.n_replicas; r0 += 8) { // 8 independent loads in flight (n_replicas is a multiple of 8)
		__half2 v[8];
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) v[k] = reinterpret_cast<const __half2*>(replicas)[(size_t)(r0 + k) * pairs + i];
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k) { const float2 f = __half22float2(v[k]); acc.x += f.x; acc.y += f.y; }
	}
	// locate the level this entry belongs to (a handful of coarse levels)
	const uint32_t entry = (i * 2) / F;
	for (uint32_t l = 0; l < gp.n_levels; ++l) {
		const GridLevel& lv = gp.level[l];
		if (lv.replicated && entry >= lv.rep_offset && entry < lv.rep_offset + lv.size) {
			__half2* dst = reinterpret_cast<__half2*>(grid_grad + (size_t)lv.offset * F) + (i - lv.rep_offset * F / 2);
			const float2 old = __half22float2(*dst);
			*dst = __floats2half2_rn(old.x + acc.x, old.y + acc.y);
			return;
		}
	}
}

// kernel_grid (dy_dx part, grid.h:171-211) + kernel_grid_backward_input (grid.h:322-349) in one pass: the reference stores
// dy/dx (fp32, D x E values per sample = 256 B at cfg #1) in the forward pass and contracts it with dL/dy in the backward pass;
// here the corner values are gathered again and contracted on the fly, nothing is stored.  Same fp32 arithmetic and
// summation order: per level and feature, dy/dx[d] = sum over the 2^(D-1) corner pairs along d of
// scale * prod_{other dims} w * (val_right - val_left); dL/dx[d] = sum_k dL/dy_k * dy_k/dx[d], k ascending.
// Thread = sample.
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }
__device__ __forceinline__ float to_f32(float v) { return v; }
// T = __half, or float for a stand-alone encoding created with fp32 precision (cpp_api.cu:156-165)
template <int D, int F, typename T = __half>
__global__ void __launch_bounds__(256) grid_backward_input_kernel(const __grid_constant__ GridParams gp, uint32_t B, const float* __restrict__ x,
	const T* __restrict__ grid, const T* __restrict__ dL_denc, float* __restrict__ dL_dx) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	float p[D];
	load_position<D>(x, i, p);
	float result[D];
#pragma unroll
	for (int d = 0; d < D; ++d) result[d] = 0.f;
	uint32_t n_levels = gp.n_levels;
	if (gp.max_level_gpu) { // dy/dx of the levels a sample does not see is zero (grid.h:69-92)
		const float max_level = (__ldg(gp.max_level_gpu + i) * (float)(gp.n_levels_total * F)) / (float)F;
		while (n_levels > 0 && (float)(n_levels - 1) >= max_level + 1e-3f) --n_levels;
	}
	for (uint32_t level = 0; level < n_levels; ++level) {
		const GridLevel& lv = gp.level[level];
		const T* table = grid + (size_t)lv.offset * F;
		float frac[D]; uint32_t cell[D];
#pragma unroll
		for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
		float dfrac[D]; // d(weight)/d(frac): 1 for Linear
#pragma unroll
		for (int d = 0; d < D; ++d) { dfrac[d] = gp.smoothstep ? smoothstep_derivative_dev(frac[d]) : 1.0f; if (gp.smoothstep) frac[d] = smoothstep_dev(frac[d]); }
		float dl[F];
#pragma unroll
		for (int f = 0; f < F; ++f) dl[f] = to_f32(dL_denc[(size_t)i * gp.padded_width + level * F + f]);
		float grads[F][D];
#pragma unroll
		for (int f = 0; f < F; ++f)
#pragma unroll
			for (int d = 0; d < D; ++d) grads[f][d] = 0.f;
#pragma unroll
		for (int gd = 0; gd < D; ++gd) {
#pragma unroll
			for (int idx = 0; idx < (1 << (D - 1)); ++idx) {
				float weight = lv.scale;
				uint32_t pg[D];
#pragma unroll
				for (int nd = 0; nd < D - 1; ++nd) {
					const int dim = nd >= gd ? nd + 1 : nd;
					if ((idx & (1 << nd)) == 0) { weight *= 1.0f - frac[dim]; pg[dim] = cell[dim]; }
					else { weight *= frac[dim]; pg[dim] = cell[dim] + 1u; }
				}
				pg[gd] = cell[gd];
				const uint32_t il = grid_index_dev<D>(lv, pg);
				pg[gd] = cell[gd] + 1u;
				const uint32_t ir = grid_index_dev<D>(lv, pg);
#pragma unroll
				for (int f = 0; f < F; ++f)
					grads[f][gd] += weight * (to_f32(table[(size_t)ir * F + f]) - to_f32(table[(size_t)il * F + f])) * dfrac[gd];
			}
		}
#pragma unroll
		for (int f = 0; f < F; ++f)
#pragma unroll
			for (int d = 0; d < D; ++d) result[d] += dl[f] * grads[f][d];
	}
#pragma unroll
	for (int d = 0; d < D; ++d) dL_dx[(size_t)i * D + d] = result[d];
}

// ---- fp32 precision (tcnn.Encoding(dtype=torch.float32); create_encoding<float>, cpp_api.cu:156-165) ------------------------
// The same kernel_grid / kernel_grid_backward arithmetic with T = float: fp32 table, fmaf interpolation in corner order,
// fp32 atomicAdd of weight * grad (grid.h:146-163, 248-252).  Thread = (sample, level); not a hot path of the trainer.
template <int D>
__device__ __forceinline__ void level_corner(const GridLevel& lv, const float (&frac)[D], const uint32_t (&cell)[D], int c, float& weight, uint32_t& idx) {
	weight = 1.0f;
	uint32_t pg[D];
#pragma unroll
	for (int d = 0; d < D; ++d) {
		if ((c >> d) & 1) { weight *= frac[d]; pg[d] = cell[d] + 1u; }
		else { weight *= 1.0f - frac[d]; pg[d] = cell[d]; }
	}
	idx = grid_index_dev<D>(lv, pg);
}
template <int D, int F>
__global__ void __launch_bounds__(256) grid_forward_f32_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const float* __restrict__ grid, float* __restrict__ encoded) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;
	float* out = encoded + (size_t)i * gp.padded_width;
	if (level == 0) for (uint32_t k = gp.n_features; k < gp.padded_width; ++k) out[k] = 0.0f; // padding (grid.h:749-759) and levels above max_level
	if (level >= gp.n_levels) return;
	const GridLevel& lv = gp.level[level];
	const float* table = grid + (size_t)lv.offset * F;
	float p[D];
	load_position<D>(x, i, p);
	float frac[D]; uint32_t cell[D];
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
	float result[F];
	if (gp.nearest) {
		const uint32_t idx = grid_index_dev<D>(lv, cell);
#pragma unroll
		for (int f = 0; f < F; ++f) result[f] = table[(size_t)idx * F + f];
	} else {
		if (gp.smoothstep) {
#pragma unroll
			for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
		}
#pragma unroll
		for (int f = 0; f < F; ++f) result[f] = 0.0f;
#pragma unroll
		for (int c = 0; c < (1 << D); ++c) {
			float weight; uint32_t idx;
			level_corner<D>(lv, frac, cell, c, weight, idx);
#pragma unroll
			for (int f = 0; f < F; ++f) result[f] = fmaf(weight, table[(size_t)idx * F + f], result[f]);
		}
	}
#pragma unroll
	for (int f = 0; f < F; ++f) out[level * F + f] = result[f];
}
template <int D, int F>
__global__ void __launch_bounds__(256) grid_backward_f32_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const float* __restrict__ dL_denc, float* __restrict__ grid_grad) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;
	if (level >= gp.n_levels) return; // set_max_level(<= 0): no active level, the launch still has one row of blocks
	const GridLevel& lv = gp.level[level];
	float* gtable = grid_grad + (size_t)lv.offset * F;
	float p[D];
	load_position<D>(x, i, p);
	float frac[D]; uint32_t cell[D];
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
	float g[F];
#pragma unroll
	for (int f = 0; f < F; ++f) g[f] = dL_denc[(size_t)i * gp.padded_width + level * F + f];
	if (gp.nearest) {
		const uint32_t idx = grid_index_dev<D>(lv, cell);
#pragma unroll
		for (int f = 0; f < F; ++f) atomicAdd(gtable + (size_t)idx * F + f, g[f]);
		return;
	}
	if (gp.smoothstep) {
#pragma unroll
		for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
	}
#pragma unroll
	for (int c = 0; c < (1 << D); ++c) {
		float weight; uint32_t idx;
		level_corner<D>(lv, frac, cell, c, weight, idx);
#pragma unroll
		for (int f = 0; f < F; ++f) atomicAdd(gtable + (size_t)idx * F + f, weight * g[f]);
	}
}

// ---- Nearest "interpolation" (grid.h:121-139, 279-282): the entry of the cell the shifted position falls into, weight 1 ----
// Same launch shapes as grid_forward_kernel / grid_backward_kernel (thread = one sample's 16-byte chunk of the encoded row).
template <int D, int F>
__global__ void __launch_bounds__(256) grid_nearest_forward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ grid, __half* __restrict__ encoded) {
	constexpr int LEVELS_PER_CHUNK = 8 / F;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t chunk = blockIdx.y;
	float p[D];
	load_position<D>(x, i, p);
	union { uint4 v; typename FeatVec<F>::type f[LEVELS_PER_CHUNK]; } out;
	out.v = make_uint4(0, 0, 0, 0);
#pragma unroll
	for (int s = 0; s < LEVELS_PER_CHUNK; ++s) {
		const uint32_t level = chunk * LEVELS_PER_CHUNK + s;
		if (level >= gp.n_levels) break;
		const GridLevel& lv = gp.level[level];
		float frac; uint32_t cell[D];
#pragma unroll
		for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac, cell[d]);
		const uint32_t idx = grid_index_dev<D>(lv, cell);
		out.f[s] = __ldg(reinterpret_cast<const typename FeatVec<F>::type*>(grid + ((size_t)lv.offset + idx) * F));
	}
	const size_t off = gp.image_subs ? wide_image_chunk(i, chunk, gp.image_subs) : ((size_t)i * gp.padded_width + chunk * 8) * 2;
	*reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(encoded) + off) = out.v;
}

template <int D, int F>
__global__ void __launch_bounds__(256) grid_nearest_backward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ dL_denc, __half* __restrict__ grid_grad, uint32_t chunk_begin) {
	constexpr int LEVELS_PER_CHUNK = 8 / F;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t chunk = chunk_begin + blockIdx.y;
	float p[D];
	load_position<D>(x, i, p);
	union { uint4 v; __half2 h2[4]; } g;
	g.v = __ldg(reinterpret_cast<const uint4*>(dL_denc + (size_t)i * gp.padded_width + chunk * 8));
#pragma unroll
	for (int s = 0; s < LEVELS_PER_CHUNK; ++s) {
		const uint32_t level = chunk * LEVELS_PER_CHUNK + s;
		if (level >= gp.n_levels) break;
		const GridLevel& lv = gp.level[level];
		float frac; uint32_t cell[D];
#pragma unroll
		for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac, cell[d]);
		__half* entry = grid_grad + ((size_t)lv.offset + grid_index_dev<D>(lv, cell)) * F;
#pragma unroll
		for (int f = 0; f < F / 2; ++f) red_add_h2(entry + 2 * f, g.h2[s * (F / 2) + f]); // (half)1.0f * grad = grad
	}
}


// stochastic_interpolation (grid.h:284-303): one draw per (sample, level) - random_val(1337, i + level * B), i.e. pcg32{1337}
// advanced by that index - picks cell or cell + 1 in EVERY dimension by comparing the same draw with the (interpolated)
// fraction; the whole gradient goes to that corner with weight 1.  The forward pass interpolates as usual.
__device__ __forceinline__ float pcg32_float_at(uint64_t state, uint64_t inc, uint64_t delta) {
	const uint32_t u = pcg32_output(pcg32_state_at(state, inc, delta));
	return __uint_as_float((u >> 9) | 0x3f800000u) - 1.0f;
}
template <int D, int F>
__global__ void __launch_bounds__(256) grid_stochastic_backward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ dL_denc, __half* __restrict__ grid_grad, uint32_t chunk_begin) {
	constexpr int LEVELS_PER_CHUNK = 8 / F;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t chunk = chunk_begin + blockIdx.y;
	float p[D];
	load_position<D>(x, i, p);
	union { uint4 v; __half2 h2[4]; } g;
	g.v = __ldg(reinterpret_cast<const uint4*>(dL_denc + (size_t)i * gp.padded_width + chunk * 8));
#pragma unroll
	for (int s = 0; s < LEVELS_PER_CHUNK; ++s) {
		const uint32_t level = chunk * LEVELS_PER_CHUNK + s;
		if (level >= gp.n_levels) break;
		const GridLevel& lv = gp.level[level];
		const float sample = pcg32_float_at(gp.rng_state, gp.rng_inc, (uint64_t)(uint32_t)(i + level * B));
		uint32_t cell[D];
#pragma unroll
		for (int d = 0; d < D; ++d) {
			float frac;
			pos_fract_dev(p[d], lv.scale, frac, cell[d]);
			if (gp.smoothstep) frac = smoothstep_dev(frac);
			if (!(sample >= frac)) cell[d] += 1u;
		}
		__half* entry = grid_grad + ((size_t)lv.offset + grid_index_dev<D>(lv, cell)) * F;
#pragma unroll
		for (int f = 0; f < F / 2; ++f) red_add_h2(entry + 2 * f, g.h2[s * (F / 2) + f]);
	}
}


// ---- general kernels: everything the specialised ones above leave out ---------------------------------------------
// One thread per (sample, level) like the reference's kernel_grid / kernel_grid_backward (grid.h:48-320), with every option as a
// run-time switch: any hash (grid_index_dev), Nearest / Linear / Smoothstep, stochastic_interpolation, the per-sample max level
// (grid.h:69-75, 236-243) and F = 1, whose gradients the reference accumulates in fp32 and casts afterwards (grad_t = float,
// grid.h:655-666, 877-885).  Chosen by the host for plans with a non-default hash, F = 1 or a per-sample max level.
template <int D, int F>
__global__ void __launch_bounds__(256) grid_general_forward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ grid, __half* __restrict__ encoded) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;                       // blockIdx.y runs over padded_width / F: the tail writes the zero padding
	__half result[F];
#pragma unroll
	for (int f = 0; f < F; ++f) result[f] = __float2half(0.f);
	bool active = level < gp.n_levels;
	if (active && gp.max_level_gpu) {
		const float max_level = (__ldg(gp.max_level_gpu + i) * (float)(gp.n_levels_total * F)) / (float)F;
		active = !((float)level >= max_level + 1e-3f);
	}
	if (active) {
		const GridLevel& lv = gp.level[level];
		float p[D], frac[D]; uint32_t cell[D];
		load_position<D>(x, i, p);
#pragma unroll
		for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
		const __half* table = grid + (size_t)lv.offset * F;
		if (gp.nearest) {
			const uint32_t idx = grid_index_dev<D>(lv, cell);
#pragma unroll
			for (int f = 0; f < F; ++f) result[f] = __ldg(table + (size_t)idx * F + f);
		} else {
			if (gp.smoothstep) {
#pragma unroll
				for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
			}
#pragma unroll
			for (int c = 0; c < (1 << D); ++c) {
				float weight = 1.0f; uint32_t pg[D];
#pragma unroll
				for (int d = 0; d < D; ++d) {
					weight *= (c & (1 << d)) == 0 ? 1.0f - frac[d] : frac[d];
					pg[d] = cell[d] + ((c >> d) & 1);
				}
				const uint32_t idx = grid_index_dev<D>(lv, pg);
				const __half wh = __float2half(weight);
#pragma unroll
				for (int f = 0; f < F; ++f) result[f] = __hfma(wh, __ldg(table + (size_t)idx * F + f), result[f]); // one rounding per corner, as __hfma2 per lane
			}
		}
	}
	const uint32_t col = level * F;
	const size_t off = gp.image_subs ? wide_image_chunk(i, col >> 3, gp.image_subs) + (col & 7u) * 2 : ((size_t)i * gp.padded_width + col) * 2;
	__half* dst = reinterpret_cast<__half*>(reinterpret_cast<uint8_t*>(encoded) + off);
#pragma unroll
	for (int f = 0; f < F; ++f) dst[f] = result[f];
}

// grad_f32 != nullptr (F = 1): fp32 atomics into a zeroed fp32 copy of the table gradient, cast by grid_cast_gradient_kernel.
template <int D, int F>
__global__ void __launch_bounds__(256) grid_general_backward_kernel(const __grid_constant__ GridParams gp, uint32_t B,
	const float* __restrict__ x, const __half* __restrict__ dL_denc, __half* __restrict__ grid_grad, float* __restrict__ grad_f32) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;
	if (level >= gp.n_levels) return;
	if (gp.max_level_gpu) {
		const float max_level = (__ldg(gp.max_level_gpu + i) * (float)(gp.n_levels_total * F)) / (float)F;
		if ((float)level > max_level + 1e-3f) return;
	}
	const GridLevel& lv = gp.level[level];
	float p[D], frac[D]; uint32_t cell[D];
	load_position<D>(x, i, p);
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
	if (gp.smoothstep && !gp.nearest) {
#pragma unroll
		for (int d = 0; d < D; ++d) frac[d] = smoothstep_dev(frac[d]);
	}
	__half g[F];
#pragma unroll
	for (int f = 0; f < F; ++f) g[f] = __ldg(dL_denc + (size_t)i * gp.padded_width + level * F + f);
	auto add = [&](const uint32_t (&pg)[D], float weight) {
		const size_t entry = ((size_t)lv.offset + grid_index_dev<D>(lv, pg)) * F;
		if constexpr (F == 1) {
			atomicAdd(grad_f32 + entry, weight * __half2float(g[0]));
		} else {
			const __half2 wh = __half2half2(__float2half(weight));
#pragma unroll
			for (int f = 0; f < F; f += 2) red_add_h2(grid_grad + entry + f, __hmul2(wh, __halves2half2(g[f], g[f + 1])));
		}
	};
	if (gp.nearest) { add(cell, 1.0f); return; }
	if (gp.stochastic) {
		const float sample = pcg32_float_at(gp.rng_state, gp.rng_inc, (uint64_t)(uint32_t)(i + level * B));
		uint32_t pg[D];
#pragma unroll
		for (int d = 0; d < D; ++d) pg[d] = sample >= frac[d] ? cell[d] : cell[d] + 1u;
		add(pg, 1.0f);
		return;
	}
#pragma unroll
	for (int c = 0; c < (1 << D); ++c) {
		float weight = 1.0f; uint32_t pg[D];
#pragma unroll
		for (int d = 0; d < D; ++d) {
			weight *= (c & (1 << d)) == 0 ? 1.0f - frac[d] : frac[d];
			pg[d] = cell[d] + ((c >> d) & 1);
		}
		add(pg, weight);
	}
}

// grad[i] = (half)tmp[i] (grid.h:877-885); accumulate: added to what grad holds (GradientMode::Accumulate)
__global__ void grid_cast_gradient_kernel(uint32_t n, const float* __restrict__ tmp, __half* __restrict__ grad, int accumulate) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	grad[i] = __float2half(accumulate ? __half2float(grad[i]) + tmp[i] : tmp[i]);
}

// ---- second order (encodings/grid.h:351-650) --------------------------------------------------------------------
// The backward pass of dL/dx = sum_k dL/dy_k * dy_k/dx (grid_backward_input_kernel): given dL_ddLdx = d(loss)/d(dL/dx), the
// gradients with respect to dL/dy, the table and x.  Needed when a loss contains the spatial gradient of the encoded field
// (eikonal / SDF terms) and autograd differentiates through the first backward pass (modules.py:120-160 of the reference).
__device__ __forceinline__ float smoothstep_2nd_derivative_dev(float v) { return 6.0f - 12.0f * v; }

// Interpolation weight, its first and second derivative with respect to the fractional coordinate (common_device.h:825-839).
template <int D>
__device__ __forceinline__ void level_weights(const GridParams& gp, const GridLevel& lv, const float (&p)[D], float (&w)[D], float (&dw)[D], float (&d2w)[D], uint32_t (&cell)[D]) {
#pragma unroll
	for (int d = 0; d < D; ++d) {
		float f;
		pos_fract_dev(p[d], lv.scale, f, cell[d]);
		w[d] = gp.smoothstep ? smoothstep_dev(f) : f;
		dw[d] = gp.smoothstep ? smoothstep_derivative_dev(f) : 1.0f;
		d2w[d] = gp.smoothstep ? smoothstep_2nd_derivative_dev(f) : 0.0f;
	}
}

// kernel_grid_backward_input_backward_dLdoutput (grid.h:624-650) and kernel_grid_backward_input_backward_input
// (grid.h:453-622) in one pass over the levels; thread = sample (no float atomics over levels / features as in the reference).
//   dL_ddLdy[i][k] = sum_d dy_k/dx_d * dL_ddLdx[i][d]                    (fp16 [B][padded_width], nullable)
//   dL_dx[i][d]    = Hessian of the interpolation contracted with dL/dy and dL_ddLdx: the diagonal part exists for
//                    Smoothstep only (d2w), the cross terms couple grad_dim with every other dim (fp32 [B][D], nullable)
template <int D, int F>
__global__ void __launch_bounds__(256) grid_backward_backward_input_kernel(const __grid_constant__ GridParams gp, uint32_t B, const float* __restrict__ x,
	const __half* __restrict__ grid, const __half* __restrict__ dL_denc, const float* __restrict__ dL_ddLdx, __half* __restrict__ dL_ddLdy, float* __restrict__ dL_dx) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	float p[D], v[D], result[D];
	load_position<D>(x, i, p);
#pragma unroll
	for (int d = 0; d < D; ++d) { v[d] = dL_ddLdx[(size_t)i * D + d]; result[d] = 0.f; }
	for (uint32_t level = 0; level < gp.n_levels; ++level) {
		const GridLevel& lv = gp.level[level];
		const __half* table = grid + (size_t)lv.offset * F;
		float w[D], dw[D], d2w[D]; uint32_t cell[D];
		level_weights<D>(gp, lv, p, w, dw, d2w, cell);
		float dl[F];
#pragma unroll
		for (int f = 0; f < F; ++f) dl[f] = dL_denc ? __half2float(dL_denc[(size_t)i * gp.padded_width + level * F + f]) : 0.f;
		// sum_f table[index][f] * dL/dy[f] * weight   (calc_dLdx, grid.h:529-540)
		auto contract = [&](const uint32_t (&pg)[D], float weight) {
			const uint32_t index = grid_index_dev<D>(lv, pg);
			float s = 0.f;
#pragma unroll
			for (int f = 0; f < F; ++f) s += __half2float(table[(size_t)index * F + f]) * dl[f] * weight;
			return s;
		};
		float dydx[F][D];
#pragma unroll
		for (int f = 0; f < F; ++f)
#pragma unroll
			for (int d = 0; d < D; ++d) dydx[f][d] = 0.f;
#pragma unroll
		for (int gd = 0; gd < D; ++gd) {
			float grad_out = 0.f;
#pragma unroll
			for (int idx = 0; idx < (1 << (D - 1)); ++idx) {
				// first derivative along gd (dy/dx, grid.h:171-211) and the diagonal of the Hessian share corners and the weight
				float weight = 1.0f; uint32_t pg[D];
#pragma unroll
				for (int nd = 0; nd < D - 1; ++nd) {
					const int dim = nd >= gd ? nd + 1 : nd;
					if ((idx & (1 << nd)) == 0) { weight *= 1.0f - w[dim]; pg[dim] = cell[dim]; }
					else { weight *= w[dim]; pg[dim] = cell[dim] + 1u; }
				}
				if (dL_ddLdy) {
					pg[gd] = cell[gd];
					const uint32_t il = grid_index_dev<D>(lv, pg);
					pg[gd] = cell[gd] + 1u;
					const uint32_t ir = grid_index_dev<D>(lv, pg);
#pragma unroll
					for (int f = 0; f < F; ++f)
						dydx[f][gd] += lv.scale * weight * (__half2float(table[(size_t)ir * F + f]) - __half2float(table[(size_t)il * F + f])) * dw[gd];
				}
				if (dL_dx) {
					if (gp.smoothstep) {
						const float wd = lv.scale * lv.scale * v[gd] * d2w[gd] * weight;
						pg[gd] = cell[gd]; grad_out += contract(pg, -wd);
						pg[gd] = cell[gd] + 1u; grad_out += contract(pg, wd);
					}
#pragma unroll
					for (int og = 0; og < D - 1; ++og) {
						const int rog = og >= gd ? og + 1 : og; // the other gradient dim
						float wo = lv.scale * lv.scale * v[rog] * dw[rog] * dw[gd];
						uint32_t pq[D];
#pragma unroll
						for (int nd = 0; nd < D - 1; ++nd) {
							const int dim = nd >= rog ? nd + 1 : nd;
							if ((idx & (1 << nd)) == 0) { if (dim != gd) wo *= 1.0f - w[dim]; else wo *= -1.0f; pq[dim] = cell[dim]; }
							else { if (dim != gd) wo *= w[dim]; pq[dim] = cell[dim] + 1u; }
						}
						pq[rog] = cell[rog]; grad_out += contract(pq, -wo);
						pq[rog] = cell[rog] + 1u; grad_out += contract(pq, wo);
					}
				}
			}
			result[gd] += grad_out;
		}
		if (dL_ddLdy) {
#pragma unroll
			for (int f = 0; f < F; ++f) {
				float r = 0.f;
#pragma unroll
				for (int d = 0; d < D; ++d) r += dydx[f][d] * v[d];
				dL_ddLdy[(size_t)i * gp.padded_width + level * F + f] = __float2half(r);
			}
		}
	}
	if (dL_ddLdy) for (uint32_t k = gp.n_features; k < gp.padded_width; ++k) dL_ddLdy[(size_t)i * gp.padded_width + k] = __float2half(0.f);
	if (dL_dx) {
#pragma unroll
		for (int d = 0; d < D; ++d) dL_dx[(size_t)i * D + d] = result[d];
	}
}

// kernel_grid_backward_input_backward_grid (grid.h:351-451): thread = sample x level; the two corners of every pair along
// grad_dim receive -/+ (half)weight * dL/dy with weight = scale * dL_ddLdx[gd] * dw[gd] * prod(other interpolation weights).
template <int D, int F>
__global__ void __launch_bounds__(256) grid_backward_backward_grid_kernel(const __grid_constant__ GridParams gp, uint32_t B, const float* __restrict__ x,
	const __half* __restrict__ dL_denc, const float* __restrict__ dL_ddLdx, __half* __restrict__ grid_grad) {
	static_assert(F % 2 == 0, "fp16 gradient scatter needs F >= 2");
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;
	if (level >= gp.n_levels) return; // set_max_level(<= 0): no active level, the launch still has one row of blocks
	const GridLevel& lv = gp.level[level];
	float p[D];
	load_position<D>(x, i, p);
	float w[D], dw[D], d2w[D]; uint32_t cell[D];
	level_weights<D>(gp, lv, p, w, dw, d2w, cell);
	__half2 g2[F / 2];
#pragma unroll
	for (int f = 0; f < F / 2; ++f) g2[f] = *reinterpret_cast<const __half2*>(dL_denc + (size_t)i * gp.padded_width + level * F + 2 * f);
	__half* gt = grid_grad + (size_t)lv.offset * F;
#pragma unroll
	for (int gd = 0; gd < D; ++gd) {
		const float grad_in = lv.scale * dL_ddLdx[(size_t)i * D + gd] * dw[gd];
#pragma unroll
		for (int idx = 0; idx < (1 << (D - 1)); ++idx) {
			float weight = grad_in; uint32_t pg[D];
#pragma unroll
			for (int nd = 0; nd < D - 1; ++nd) {
				const int dim = nd >= gd ? nd + 1 : nd;
				if ((idx & (1 << nd)) == 0) { weight *= 1.0f - w[dim]; pg[dim] = cell[dim]; }
				else { weight *= w[dim]; pg[dim] = cell[dim] + 1u; }
			}
#pragma unroll
			for (int side = 0; side < 2; ++side) {
				pg[gd] = cell[gd] + (uint32_t)side;
				const __half2 wh = __float2half2_rn(side ? weight : -weight);
				__half* dst = gt + (size_t)grid_index_dev<D>(lv, pg) * F;
#pragma unroll
				for (int f = 0; f < F / 2; ++f) red_add_h2(dst + 2 * f, __hmul2(wh, g2[f]));
			}
		}
	}
}

// Identity encoding in front of a stand-alone Network (encodings/identity.h:46-90, src/cpp_api.cu:151-153): fp32 [B][n_in] ->
// fp16 [B][width], columns >= n_in are ONE (identity.h:63), so that the first layer can learn a bias through them.
// Thread = sample x 8-column chunk.
__global__ void __launch_bounds__(256) identity_encode_kernel(uint32_t B, uint32_t n_in, uint32_t width, const float* __restrict__ x, __half* __restrict__ out,
	uint32_t image_subs = 0) {
	const uint32_t cpr = width / 8, q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= B * cpr) return;
	const uint32_t i = q / cpr, c0 = (q % cpr) * 8;
	union { uint4 v; __half h[8]; } o;
#pragma unroll
	for (uint32_t k = 0; k < 8; ++k) o.h[k] = __float2half(c0 + k < n_in ? __ldg(x + (size_t)i * n_in + c0 + k) : 1.0f);
	const size_t off = image_subs ? wide_image_chunk(i, c0 / 8, image_subs) : ((size_t)i * width + c0) * 2;
	*reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(out) + off) = o.v;
}
// identity_backward (identity.h:68-83): dL/dx = (float) dL/dy for the first n_in columns
__global__ void __launch_bounds__(256) identity_backward_kernel(uint32_t B, uint32_t n_in, uint32_t width, const __half* __restrict__ dL_dy, float* __restrict__ dL_dx) {
	const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= B * n_in) return;
	const uint32_t i = q / n_in, j = q % n_in;
	dL_dx[q] = __half2float(dL_dy[(size_t)i * width + j]);
}

// Parity helper: the per-corner entry indices exactly as the encode/scatter kernels compute them.
// out[(level * B + i) * 2^D + corner]   (same layout as oracle/ref_driver.cu dump_indices)
template <int D>
__global__ void grid_indices_kernel(const __grid_constant__ GridParams gp, uint32_t B, const float* __restrict__ x, uint32_t* __restrict__ out) {
	constexpr int NC = 1 << D;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= B) return;
	const uint32_t level = blockIdx.y;
	// (a diagnostic of the index function: every level of the table, whatever max_level says)
	const GridLevel& lv = gp.level[level];
	float p[D];
	load_position<D>(x, i, p);
	float frac[D]; uint32_t cell[D];
#pragma unroll
	for (int d = 0; d < D; ++d) pos_fract_dev(p[d], lv.scale, frac[d], cell[d]);
#pragma unroll
	for (int c = 0; c < NC; ++c) {
		uint32_t pg[D];
#pragma unroll
		for (int d = 0; d < D; ++d) pg[d] = cell[d] + ((c >> d) & 1u);
		out[((size_t)level * B + i) * NC + c] = grid_index_dev<D>(lv, pg);
	}
}

} // namespace b200
