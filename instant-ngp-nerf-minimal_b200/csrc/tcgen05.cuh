// Thin inline-PTX layer over the sm_100a tensor-core path: TMEM allocation, tcgen05.mma (kind::f16, cta_group::1),
// tcgen05.commit -> mbarrier, tcgen05.ld, proxy fences, and the shared-memory / instruction descriptors.
// Operands use the NO-SWIZZLE canonical layouts: 8x8 fp16 core matrices stored as 8 rows of 16 bytes (128 contiguous
// bytes), so ONE shared-memory buffer serves both as a K-major operand (rows = M/N index) and as an MN-major operand
// (rows = K index) -- which is what lets the activations feed the next layer AND the weight-gradient GEMM untouched.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace ngp { namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
	uint32_t ok;
	asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
	             : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
	return ok != 0;
}
// non-blocking probe (try_wait may suspend the thread for a system-dependent time: wrong for a thread that polls several barriers)
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
	uint32_t ok;
	asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
	             : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
	return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	while (!mbar_try_wait(bar, parity)) { }
}

// ---- fences ------------------------------------------------------------------------------------------
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMEM --------------------------------------------------------------------------------------------
// one full warp; writes the base address (lane 0, column c) to *dst_smem
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t n_cols) {
	asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(dst_smem), "r"(n_cols) : "memory");
	asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t n_cols) {
	asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(n_cols) : "memory");
}

// 32 lanes x 16 consecutive 32-bit columns: thread t of warp w receives lane 32*(w%4)+t
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
	uint32_t r[16];
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	             : "r"(taddr) : "memory");
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
	#pragma unroll
	for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// the same load without the wait: several can be in flight, tmem_ld_wait() then covers all of them.  tmem_ld_pin() ties the
// destination registers to a point after the wait so that no use of them is scheduled above it.
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t* r) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
	               "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
	             : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_pin(uint32_t* r) {
	asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
	                  "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]) :: "memory");
}

// ---- wider / narrower TMEM loads and TMEM stores (32x32b: thread t of warp w owns lane 32*(w%4)+t, N consecutive columns) ----
__device__ __forceinline__ void tmem_ld8_nowait(uint32_t taddr, uint32_t* r) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t* r) {
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
	             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
	               "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
	             : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_pin8(uint32_t* r) { asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]) :: "memory"); }
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
	asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
	             :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
	asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
	             :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
	                "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- mbarrier arrive (one pending count) -------------------------------------------------------------
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
	asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" :: "r"(bar) : "memory");
}

// D[tmem] (+)= A[tmem] * B[smem]: the A operand comes from tensor memory -- M = 128 rows on the 128 lanes, the K = 16 fp16 values of
// one instruction in 8 consecutive 32-bit columns (two values per column, the even one in the low half).  Issued by ONE thread.
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
	             :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

// ---- descriptors -------------------------------------------------------------------------------------
// shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout type 0 = no swizzle [61,64)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
	return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46);
}

// instruction descriptor for kind::f16, fp16 inputs, fp32 accumulate (InstrDescriptor in the same header)
__host__ __device__ constexpr uint32_t instr_desc(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major) {
	return (1u << 4) | (a_mn_major << 15) | (b_mn_major << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
	asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
	             :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// all previously issued MMAs of this thread arrive on the mbarrier when they complete
__device__ __forceinline__ void mma_commit(uint32_t bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}

// An operand view over a core-matrix-tiled buffer whose rows hold F fp16 values:
//   element (row r, feature f) lives at (r/8)*16F + (f/8)*128 + (r%8)*16 + (f%8)*2 bytes.
struct Operand { uint32_t addr, lbo, sbo, kstep; };
// rows are the M/N index, features the K index
__device__ __forceinline__ Operand k_major(uint32_t buf, uint32_t F) { return {buf, 128u, 16u * F, 256u}; }
// rows are the K index, features the M/N index (first MN feature = f0, a multiple of 8)
__device__ __forceinline__ Operand mn_major(uint32_t buf, uint32_t F, uint32_t f0 = 0) { return {buf + (f0 >> 3) * 128u, 16u * F, 128u, 32u * F}; }

// K = 16*nk.  One thread.
__device__ __forceinline__ void gemm(uint32_t d_tmem, uint32_t idesc, Operand a, Operand b, uint32_t nk, uint32_t accumulate) {
	for (uint32_t k = 0; k < nk; ++k)
		mma_f16(d_tmem, smem_desc(a.addr + k * a.kstep, a.lbo, a.sbo), smem_desc(b.addr + k * b.kstep, b.lbo, b.sbo), idesc, (k > 0) ? 1u : accumulate);
}

// the same with A in tensor memory: a_tmem advances by 8 columns per K step of 16
__device__ __forceinline__ void gemm_ts(uint32_t d_tmem, uint32_t idesc, uint32_t a_tmem, Operand b, uint32_t nk, uint32_t accumulate) {
	for (uint32_t k = 0; k < nk; ++k)
		mma_f16_ts(d_tmem, a_tmem + k * 8u, smem_desc(b.addr + k * b.kstep, b.lbo, b.sbo), idesc, (k > 0) ? 1u : accumulate);
}

__device__ __forceinline__ uint32_t tile_off(uint32_t row, uint32_t chunk, uint32_t F) { return (row >> 3) * (16u * F) + chunk * 128u + (row & 7u) * 16u; }

}} // namespace ngp::tc
