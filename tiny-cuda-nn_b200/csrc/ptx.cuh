// Thin inline-PTX wrappers for sm_100a: mbarrier, proxy fences, tcgen05 (alloc / mma / commit / ld), red.f16x2.
// Hand-written for this project; bit layouts follow the PTX ISA "tcgen05" matrix/instruction descriptor tables.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile(
		"{\n\t"
		".reg .pred p;\n\t"
		"MBAR_WAIT:\n\t"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
		"@p bra MBAR_DONE;\n\t"
		"bra MBAR_WAIT;\n\t"
		"MBAR_DONE:\n\t"
		"}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

// 1-D bulk copy global -> shared through the TMA unit (SASS: UBLKCP), completion on an mbarrier.
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// 1-D bulk copy shared -> global through the TMA unit (bulk async-group completion).  The source must not be overwritten
// before wait_read has returned; the global writes are complete after wait_all (or at kernel end).
__device__ __forceinline__ void tma_bulk_s2g(void* dst_gmem, uint32_t src_smem_addr, uint32_t bytes) {
	asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(src_smem_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void tma_bulk_g2s_addr(uint32_t dst_smem_addr, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
		::"r"(dst_smem_addr), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// generic-proxy smem writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tcgen05 -------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// one full warp; writes the TMEM base address (lane 0, column c) to *dst_smem
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// Shared-memory matrix descriptor, 128-byte swizzle (layout_type 2), descriptor version 1 (Blackwell).
//   bits [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version | [61,64) layout
// K-major : rows of 128 B (64 halfs of K), 8-row groups SBO apart; LBO unused (1).
// MN-major: rows of 128 B (64 halfs of M/N) indexed by k, 8-k groups SBO apart, 64-wide M/N atoms LBO apart.
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
	uint64_t d = 0;
	d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
	d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
	d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
	d |= (uint64_t)1 << 46;
	d |= (uint64_t)2 << 61;
	return d;
}

// Instruction descriptor for kind::f16, fp16 A/B, fp32 accumulate.
//   [4,6) c_format=1(F32) | [7,10) a_format=0(F16) | [10,13) b_format=0 | 15 a_major | 16 b_major | [17,23) N>>3 | [24,29) M>>4
__host__ __device__ constexpr uint32_t make_idesc_f16(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major, uint32_t c_format_f32 = 1) {
	return (c_format_f32 << 4) | (a_mn_major << 15) | (b_mn_major << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
	asm volatile(
		"{\n\t"
		".reg .pred p;\n\t"
		"setp.ne.b32 p, %4, 0;\n\t"
		"tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
		"}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}

// all prior tcgen05.mma of this thread -> one arrival on the mbarrier when they have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns (lane = row of D).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x32.b32 "
		"{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
		"%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
		  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
		  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
		: "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x16.b32 "
		"{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
		: "r"(taddr) : "memory");
}
// fp16 accumulators (instruction descriptor c_format = F16): a 16-bit element sits in the low half of its 32-bit TMEM column;
// .pack::16b returns two adjacent columns per register (column 2i in the low half), i.e. the row as packed half2.
__device__ __forceinline__ void tmem_ld32_pack16(uint32_t taddr, uint32_t (&r)[32]) { // 64 columns
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "
		"{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
		"%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
		  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
		  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
		: "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16_pack16(uint32_t taddr, uint32_t (&r)[16]) { // 32 columns
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
		"{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
		: "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8_pack16(uint32_t taddr, uint32_t (&r)[8]) { // 16 columns
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
		: "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- shared memory by 32-bit address (explicit .shared space: STS/LDS instead of generic ST/LD) ----
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
	asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) { sts128(addr, v.x, v.y, v.z, v.w); }
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
	uint4 v;
	asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
	return v;
}
// packed fp16x2 helpers on raw 32-bit registers
__device__ __forceinline__ uint32_t h2_relu(uint32_t x) {
	const __half2 z = __floats2half2_rn(0.f, 0.f);
	__half2 r = __hmax2(*reinterpret_cast<__half2*>(&x), z);
	return *reinterpret_cast<uint32_t*>(&r);
}
// g * (a > 0 ? 1 : 0) per half: ReLU backward from the stored post-activation (common_device.h:243-248)
__device__ __forceinline__ uint32_t h2_mask_mul(uint32_t g, uint32_t act) {
	const __half2 z = __floats2half2_rn(0.f, 0.f);
	__half2 m = __hgt2(*reinterpret_cast<__half2*>(&act), z); // 1.0 / 0.0
	__half2 r = __hmul2(*reinterpret_cast<__half2*>(&g), m);
	return *reinterpret_cast<uint32_t*>(&r);
}

// Loads whose position in the instruction stream matters (software prefetch): ld.global.nc may otherwise be sunk to its
// first use by the compiler, which re-exposes the latency the prefetch was meant to hide.
__device__ __forceinline__ uint4 ldg128_here(const void* p) {
	uint4 v;
	asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}
__device__ __forceinline__ float ldg_f32_here(const float* p) {
	float v;
	asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
	return v;
}

// ---- misc ----------------------------------------------------------------------------------------
// fp16x2 reduction into global memory without return value (the op the reference scatter uses, vec.h:326-333)
__device__ __forceinline__ void red_add_h2(__half* addr, __half2 v) {
	asm volatile("red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;" ::"l"(addr), "r"(*reinterpret_cast<uint32_t*>(&v)) : "memory");
}
// 8- and 16-byte vector forms (SASS: REDG.E.ADD.F16x4 / F16x8), address aligned to the vector size
__device__ __forceinline__ void red_add_h2x2(__half* addr, uint32_t a, uint32_t b) {
	asm volatile("red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void red_add_h2x4(__half* addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
	asm volatile("red.relaxed.gpu.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// predicated forms: no branch around the reduction, so the scatter of several levels stays one schedulable block
__device__ __forceinline__ void red_add_h2_if(bool p, __half* addr, uint32_t v) {
	asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q red.relaxed.gpu.global.add.noftz.f16x2 [%0], %1;\n\t}" ::"l"(addr), "r"(v), "r"((uint32_t)p) : "memory");
}
__device__ __forceinline__ void red_add_h2x2_if(bool p, __half* addr, uint32_t a, uint32_t b) {
	asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\t@q red.relaxed.gpu.global.add.noftz.v2.f16x2 [%0], {%1, %2};\n\t}" ::"l"(addr), "r"(a), "r"(b), "r"((uint32_t)p) : "memory");
}
__device__ __forceinline__ uint32_t h2_bits(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
#ifdef TCNN_B200_PACK_RZ // development experiment (tests/dev/parity_probe.py): round the fp32 accumulators toward zero
	__half2 h = __halves2half2(__float2half_rz(a), __float2half_rz(b));
#else
	__half2 h = __floats2half2_rn(a, b);
#endif
	return *reinterpret_cast<uint32_t*>(&h);
}

// named barriers (producer/consumer hand-off between warpgroups of one CTA; id 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t n_threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n_threads) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t n_threads) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n_threads) : "memory"); }

// register re-balancing between warpgroups (all warps of the warpgroup execute it together)
template <uint32_t N> __device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <uint32_t N> __device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// ---- peer memory over NVLink / NVSwitch ---------------------------------------------------------------
// System-scope accesses for buffers that other GPUs read and write (symmetric allocations mapped into every rank).
__device__ __forceinline__ uint4 ld_sys_v4(const void* addr) {
	uint4 v;
	asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(addr) : "memory");
	return v;
}
__device__ __forceinline__ void st_sys_v4(void* addr, uint4 v) {
	asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint2 ld_sys_v2(const void* addr) {
	uint2 v;
	asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(addr) : "memory");
	return v;
}
__device__ __forceinline__ void st_sys_v2(void* addr, uint2 v) {
	asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* addr, uint32_t v) {
	asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* addr) {
	uint32_t v;
	asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
	return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
// counter += 1 with wrap to 0 after `wrap`; release/acquire at GPU scope (orders the CTA's earlier stores before it)
__device__ __forceinline__ uint32_t atom_inc_acq_rel_gpu(uint32_t* addr, uint32_t wrap) {
	uint32_t old;
	asm volatile("atom.acq_rel.gpu.global.inc.u32 %0, [%1], %2;" : "=r"(old) : "l"(addr), "r"(wrap) : "memory");
	return old;
}
__device__ __forceinline__ void st_release_sys(uint32_t* addr, uint32_t v) {
	asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* addr) {
	uint32_t v;
	asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
	return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
	uint64_t t;
	asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
	return t;
}
// NVSwitch multicast (NVLS): one load returns the sum over every rank's copy, one store lands in every rank's copy.
__device__ __forceinline__ uint4 multimem_ld_reduce_h2x4(const void* mc_addr) {
	uint4 v;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
		: "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_addr) : "memory");
	return v;
}
__device__ __forceinline__ uint2 multimem_ld_reduce_h2x2(const void* mc_addr) {
	uint2 v;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.f16x2 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(mc_addr) : "memory");
	return v;
}
__device__ __forceinline__ void multimem_st_v2(void* mc_addr, uint2 v) {
	asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc_addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void multimem_st_v4(void* mc_addr, uint4 v) {
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

} // namespace b200
