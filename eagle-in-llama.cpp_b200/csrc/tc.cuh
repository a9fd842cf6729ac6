// tc.cuh — thin PTX wrappers for the Blackwell (sm_100a) async machinery used by mmq_tc.cu:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld / st / fences).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 tok;\n\tmbarrier.arrive.shared::cta.b64 tok, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 tok;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 tok, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    // suspend-time hint: the thread sleeps in hardware until the phase completes (or the limit passes) instead of
    // re-issuing the probe — polling warps otherwise eat ~20 % of the issue slots of the working warps (ncu, r1c)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity), "r"(0x989680u) : "memory");
    return ok != 0;
}
// non-blocking probe of a phase (no suspend window)
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) { }
}

// ---- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *m, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(m), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

// ---- tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem desc]   (kind::f16, A K-major in TMEM: lane = row, 2 halves per column)
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// four K = 16 steps issued back to back (one asm block: the issue slot is ~40 clk/MMA only when
// nothing sits between consecutive tcgen05.mma — measured in tools/ubench_mma.cu): A columns
// a_tmem + 8j, B descriptor b_desc + 32j bytes.  The first step overwrites D when acc_first == 0.
__device__ __forceinline__ void mma_f16_ts_x4(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc_first) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b32 a1, a2, a3;\n\t.reg .b64 b1, b2, b3;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\tsetp.eq.b32 q, 0, 0;\n\t"
                 "add.u32 a1, %1, 8;\n\tadd.u32 a2, %1, 16;\n\tadd.u32 a3, %1, 24;\n\t"
                 "add.u64 b1, %2, 2;\n\tadd.u64 b2, %2, 4;\n\tadd.u64 b3, %2, 6;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, q;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc_first) : "memory");
}
// eight K = 16 steps (two 64-element groups): A columns a_tmem + 8j, B = b_lo + 32j bytes (j < 4), b_hi + 32(j-4) bytes
__device__ __forceinline__ void mma_f16_ts_x8(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_lo, uint64_t b_hi, uint32_t idesc, uint32_t acc_first) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b32 a1, a2, a3, a4, a5, a6, a7;\n\t.reg .b64 b1, b2, b3, b5, b6, b7;\n\t"
                 "setp.ne.b32 p, %5, 0;\n\tsetp.eq.b32 q, 0, 0;\n\t"
                 "add.u32 a1, %1, 8;\n\tadd.u32 a2, %1, 16;\n\tadd.u32 a3, %1, 24;\n\tadd.u32 a4, %1, 32;\n\t"
                 "add.u32 a5, %1, 40;\n\tadd.u32 a6, %1, 48;\n\tadd.u32 a7, %1, 56;\n\t"
                 "add.u64 b1, %2, 2;\n\tadd.u64 b2, %2, 4;\n\tadd.u64 b3, %2, 6;\n\t"
                 "add.u64 b5, %3, 2;\n\tadd.u64 b6, %3, 4;\n\tadd.u64 b7, %3, 6;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %4, p;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a4], %3, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a5], b5, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a6], b6, %4, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a7], b7, %4, q;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_lo), "l"(b_hi), "r"(idesc), "r"(acc_first) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], eight K = 16 steps over two 64-element (128-byte) K-major boxes per operand:
// step j uses a_lo/b_lo + 32j bytes (j < 4) and a_hi/b_hi + 32(j-4) bytes
__device__ __forceinline__ void mma_f16_ss_x8(uint32_t d_tmem, uint64_t a_lo, uint64_t a_hi, uint64_t b_lo, uint64_t b_hi, uint32_t idesc, uint32_t acc_first) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b64 a1, a2, a3, a5, a6, a7, b1, b2, b3, b5, b6, b7;\n\t"
                 "setp.ne.b32 p, %6, 0;\n\tsetp.eq.b32 q, 0, 0;\n\t"
                 "add.u64 a1, %1, 2;\n\tadd.u64 a2, %1, 4;\n\tadd.u64 a3, %1, 6;\n\t"
                 "add.u64 a5, %2, 2;\n\tadd.u64 a6, %2, 4;\n\tadd.u64 a7, %2, 6;\n\t"
                 "add.u64 b1, %3, 2;\n\tadd.u64 b2, %3, 4;\n\tadd.u64 b3, %3, 6;\n\t"
                 "add.u64 b5, %4, 2;\n\tadd.u64 b6, %4, 4;\n\tadd.u64 b7, %4, 6;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %3, %5, p;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a1, b1, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a2, b2, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a3, b3, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %2, %4, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a5, b5, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a6, b6, %5, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], a7, b7, %5, q;\n\t}"
                 ::"r"(d_tmem), "l"(a_lo), "l"(a_hi), "l"(b_lo), "l"(b_hi), "r"(idesc), "r"(acc_first) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc], eight K = 16 steps: A columns a_tmem + 8j, B descriptor b_desc + j * 128 (2048 bytes:
// sixteen 128-byte rows of an MN-major operand)
__device__ __forceinline__ void mma_f16_ts_x8_rows(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc_first) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b32 a1, a2, a3, a4, a5, a6, a7;\n\t.reg .b64 b1, b2, b3, b4, b5, b6, b7;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\tsetp.eq.b32 q, 0, 0;\n\t"
                 "add.u32 a1, %1, 8;\n\tadd.u32 a2, %1, 16;\n\tadd.u32 a3, %1, 24;\n\tadd.u32 a4, %1, 32;\n\t"
                 "add.u32 a5, %1, 40;\n\tadd.u32 a6, %1, 48;\n\tadd.u32 a7, %1, 56;\n\t"
                 "add.u64 b1, %2, 128;\n\tadd.u64 b2, %2, 256;\n\tadd.u64 b3, %2, 384;\n\tadd.u64 b4, %2, 512;\n\t"
                 "add.u64 b5, %2, 640;\n\tadd.u64 b6, %2, 768;\n\tadd.u64 b7, %2, 896;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a4], b4, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a5], b5, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a6], b6, %3, q;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [a7], b7, %3, q;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc_first) : "memory");
}
template <int N> __device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ bool elect_one() {
    uint32_t e;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(e));
    return e != 0;
}
// arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// K-major, 128-byte swizzled smem operand (rows of 128 B, 8-row groups 1024 B apart)
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // start address, 16-byte units
    d |= (uint64_t)1 << 16;                           // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                 // stride byte offset: 8 rows x 128 B
    d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
    return d;
}
// MN-major, 128-byte swizzled smem operand: 64 MN elements contiguous per 128-byte row, 8 K rows per 1024-byte atom,
// the next 64 MN elements lbo_bytes away (canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units)
__device__ __forceinline__ uint64_t smem_desc_sw128_mn(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
#define TC_IDESC_B_MN (1u << 16)     // instruction-descriptor bit: B operand is MN-major
// instruction descriptor: f16 x f16 -> f32, both K-major, M = 128
__host__ __device__ __forceinline__ uint32_t idesc_f16_m128(int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
                   "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
                   "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}

}  // namespace tc

// host: 2-D f16 tensor map (defined in mmq_tc.cu next to the driver entry point); swizzle = 128 B
int eb_tc_map_2d_f16(CUtensorMap *m, const void *base, uint64_t d0, uint64_t d1, uint64_t stride1_bytes, uint32_t b0, uint32_t b1);
int eb_tc_ready();

