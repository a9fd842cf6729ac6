// mmq_tc.cu — K2 on 5th-generation tensor cores: tree-batched (T = 9..80) verify GEMM on Q4_K
// weights with tcgen05.mma (kind::f16, accumulators in TMEM), weights streamed from HBM exactly once
// by TMA.  Replaces mul_mat_q (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822); results follow
// ggml_vec_dot_q4_K_q8_K (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:6266): all integer sums are exact,
// only the order of the f32 additions over super-blocks differs.
//
// Exact integers on an f16 tensor pipe.  Per super-block (K = 256) and weight row r, column c:
//     y[r,c] += d[r]*dx[c] * SUM_k (sc[r,k/32]*q[r,k]) * a[c,k]  -  dmin[r]*dx[c] * SUM_j m[r,j]*bsum16[c,2j..2j+1]
// sc*q <= 63*15 = 945, a in [-127,127], m <= 63 and bsum16 in [-2032,2032] are all exactly
// representable in f16; products are exact and the f32 accumulation of integers stays exact below 2^24.
// So A = sc*q (f16, built by the dequant warps straight into TMEM), B = a as f16 (K-major, TMA 128B-swizzled
// tiles), D = f32 in TMEM, one fresh accumulator per super-block, scaled by d*dx in the epilogue warps.
//
// CTA = 128 weight rows, persistent over (row block, K split) items.  Warp roles (20 warps):
//   0      TMA producer of raw weight tiles  (128 rows x one super-block, 6-stage ring)
//   1      tcgen05.mma issuer (warp-uniform loop, one elected lane issues 8-9 MMAs per asm block)
//   2      TMA producer of the activation tiles (f16 quants, 16-sums)
//   3      TMEM allocator
//   4-11   dequant: smem raw blocks -> exact f16 integers -> TMEM A operand (lane = weight row)
//   12-19  epilogue: TMEM accumulators -> f32 registers (d*dx scaling), final store of the split-K plane
// Measured on B200 (tools/ubench*.cu, profiles/): tcgen05.mma costs ~165 clk each when issued from a divergent
// single-thread loop (R2UR traffic) and 40 clk (the N = 80 floor) when several are issued back to back from
// uniform registers; every mbarrier hand-off costs 100-300 clk of latency, so stages are coarse (half a
// super-block of A, a whole super-block of D) and the TMEM stores overlap the next group's arithmetic.
#include "common.cuh"
#include "tc.cuh"
#include <algorithm>
#include <cstdlib>
#include <cuda.h>

namespace {

constexpr int TC_THREADS = 640;
constexpr int S_W = 6;                          // weight ring depth
constexpr int W_STAGE = 128 * 144;              // Q4_K: 128 rows x 144 B
constexpr int B_BOX = 80 * 128;                 // one 128B-swizzled box: up to 80 columns x 64 f16
constexpr int B_STAGE = 4 * B_BOX;              // one super-block of activations
constexpr int OFF_W = 0;
constexpr int OFF_B = OFF_W + S_W * W_STAGE;    // 110592
constexpr int OFF_BS = OFF_B + 2 * B_STAGE;     // 192512
constexpr int OFF_HDR = OFF_BS + 2 * B_BOX;     // 212992
constexpr int OFF_DX = OFF_HDR + 4 * 128 * 8;   // 217088: per epilogue warp 2 x 40 floats
constexpr int OFF_BAR = OFF_DX + 8 * 80 * 4;    // 219648
constexpr int SMEM_BYTES = OFF_BAR + 512 + 1024;   // + barriers + alignment slack
// TMEM columns: D stage s at s*160 (main 80 | min 80), A half-super-block h at 320 + 64h, mins operand at 448 + 8*(i&1)
constexpr int D_STRIDE = 160, D_MIN = 80, A_OFF = 320, AM_OFF = 448;
// barrier indices
enum { W_FULL = 0, W_EMPTY = 6, B_FULL = 12, B_EMPTY = 14, BS_FULL = 16, BS_EMPTY = 18, A_FULL = 20, A_EMPTY = 22,
       D_FULL = 24, D_EMPTY = 26, HDR_FULL = 28, N_BARS = 32 };

// Q6_K variant (FMT = 1): stage = ql tile (128 rows x 128 B, 128B-swizzled) | qh tile (128 x 64 B, 64B-swizzled) | scales (128 x 16 B);
// one accumulator of 80 columns per stage (no mins term); A operand = two exact f16 addends per half-super-block
constexpr int S_W6 = 4;
constexpr int W6_QH = 128 * 128, W6_SC = W6_QH + 128 * 64, W6_STAGE = W6_SC + 128 * 16;    // 16384, 24576, 26624
static_assert(S_W6 * W6_STAGE <= S_W * W_STAGE, "Q6_K weight ring must fit the Q4_K one");
constexpr int D6_STRIDE = 80, A6_OFF = 192, A6_HALF = 128, A6_SECOND = 64;

struct TcArgs {
    const float *dT; float *out; int64_t N; int nbr, T, npad, n_rb, n_items, spp, early_trigger; long long *prof;
    uint32_t k_lo, k_hi, k_magic;    // 0x000f000f, 0x00f000f0, 0x64006400 passed as data: kept in registers so that (x & m) | c is ONE LOP3
    uint32_t k_q2;                   // 0x03000300
    const __half *w_d;               // Q6_K: per-block d plane [N][nbr]
    TpPush tp;                       // tensor parallel: mirror the planes into the peers' memory (n_peers = 0: off)
};

// optional in-kernel profile (EB_TC_PROF=1): cycles each role of CTA 0 spends waiting on its barriers
#define TCP_WAIT(slot, barrier, parity) do { if (PROF) { const long long t_ = clock64(); tc::mbar_wait(barrier, parity); pw[slot] += clock64() - t_; } else tc::mbar_wait(barrier, parity); } while (0)

__device__ __forceinline__ uint32_t and_or(uint32_t x, uint32_t m, uint32_t c) {
    uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(x), "r"(m), "r"(c)); return r;
}
// packed f32x2 arithmetic (sm_100): one issue slot for two FMAs
__device__ __forceinline__ unsigned long long pk2(float x, float y) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y)); return r; }
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) { unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) { unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t *>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2 *>(&u); }
// byte b (0..3) of word w as half2 (v, v), v in [0,255] exact
__device__ __forceinline__ __half2 byte_to_h2(uint32_t w, int b) {
    const uint32_t v = (w >> (8 * b)) & 0xffu;
    return __hsub2(u2h(v | (v << 16) | 0x64006400u), u2h(0x64006400u));
}

// one 64-element group (32 bytes: low nibbles = sub-block 2g, high nibbles = sub-block 2g+1) -> 32 TMEM columns
// of exact f16 integers sc*q.  Column 2w+e of a sub-block holds elements (4w+e, 4w+e+2): the activation plane
// h16 uses the same (0,2,1,3) order inside every group of four (quant_dev.cuh).
__device__ __forceinline__ void dequant_group(uint32_t (&r)[32], const uint4 &q0, const uint4 &q1, __half2 s_lo, __half2 s_hi,
                                              uint32_t k_lo, uint32_t k_hi, uint32_t k_magic) {
    const __half2 mul_lo = s_lo, add_lo = __hmul2(s_lo, __float2half2_rn(-1024.f));                       // (1024+q)*sc - 1024*sc
    const __half2 mul_hi = __hmul2(s_hi, __float2half2_rn(0.0625f)), add_hi = __hmul2(s_hi, __float2half2_rn(-64.f));   // (1024+16q)*sc/16 - 64*sc
    const uint32_t xs[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint32_t x = xs[w], z = x >> 8;                     // bytes (1,3) under the same masks as bytes (0,2)
        r[2 * w]          = h2u(__hfma2(u2h(and_or(x, k_lo, k_magic)), mul_lo, add_lo));
        r[2 * w + 1]      = h2u(__hfma2(u2h(and_or(z, k_lo, k_magic)), mul_lo, add_lo));
        r[16 + 2 * w]     = h2u(__hfma2(u2h(and_or(x, k_hi, k_magic)), mul_hi, add_hi));
        r[16 + 2 * w + 1] = h2u(__hfma2(u2h(and_or(z, k_hi, k_magic)), mul_hi, add_hi));
    }
}

// dequant role: raw Q4_K rows in smem -> TMEM A operand (lane = weight row).  MEMBER 0 handles groups 0 and 2
// (+ the header and the mins operand), MEMBER 1 groups 1 and 3; half-super-block h = groups 2h, 2h+1.
// Software-pipelined over the CTA's super-blocks (the role never needs to know which row block / K range a stage
// belongs to): every tcgen05.st is followed by the arithmetic of the NEXT group before its wait::st, so the TMEM
// store latency is never exposed — the second half's store overlaps loading + dequantising the next super-block.
template <int MEMBER, bool PROF>
__device__ __forceinline__ void dequant_q4_K(const TcArgs &p, uint8_t *smem, uint32_t bar0, uint32_t tmem, int q, int lane) {
    auto bar = [&](int i) { return bar0 + 8u * i; };
    const int row = 32 * q + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    float2 *hdr_sm = (float2 *)(smem + OFF_HDR);
    const uint32_t k_lo = p.k_lo, k_hi = p.k_hi, k_magic = p.k_magic;
    uint32_t n_total = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int sb0 = (item / p.n_rb) * p.spp;
        n_total += min(p.nbr, sb0 + p.spp) - sb0;
    }
    long long pw[4] = {0, 0, 0, 0}; const long long pt0 = PROF ? clock64() : 0;
    uint4 hd, qb0, qb1; uint32_t u1 = 0;
    uint32_t ra[32];
    // stage i: read this thread's row (header, its group of each half), release the stage, dequantise the first-half group
    auto fetch = [&](uint32_t i) {
        const uint32_t s = i % S_W;
        const uint8_t *rp = smem + OFF_W + s * W_STAGE + row * 144;
        hd = *(const uint4 *)rp;
        const uint4 qa0 = *(const uint4 *)(rp + 16 + 32 * MEMBER), qa1 = *(const uint4 *)(rp + 32 + 32 * MEMBER);
        qb0 = *(const uint4 *)(rp + 16 + 32 * (MEMBER + 2)); qb1 = *(const uint4 *)(rp + 32 + 32 * (MEMBER + 2));
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(bar(W_EMPTY + s));
        // 6-bit scales / mins (get_scale_min_k4, R/ggml/src/ggml-quants.c:631) as bytes of u0..u1 / u2..u3
        const uint32_t u0 = hd.y & 0x3f3f3f3fu;
        u1 = (hd.w & 0x0f0f0f0fu) | (((hd.y >> 6) & 0x03030303u) << 4);
        dequant_group(ra, qa0, qa1, byte_to_h2(u0, 2 * MEMBER), byte_to_h2(u0, 2 * MEMBER + 1), k_lo, k_hi, k_magic);
    };
    if (n_total > 0) {
        TCP_WAIT(0, bar(W_FULL + 0), 0);
        if (PROF && blockIdx.x == 0 && MEMBER == 0 && q == 0 && lane == 0) p.prof[64 + 3] = clock64();
        fetch(0);
    }
    for (uint32_t i = 0; i < n_total; i++) {
        // ---- first half-super-block (group MEMBER)
        TCP_WAIT(1, bar(A_EMPTY + 0), (i & 1) ^ 1);
        tc::fence_after();
        if (MEMBER == 0) {                                          // header for the epilogue, once per super-block
            const __half2 dm = u2h(hd.x);
            hdr_sm[(i & 3) * 128 + row] = make_float2(__low2float(dm), __high2float(dm));
        } else {                                                    // mins operand (the other member of the row pair: evens out the two)
            const uint32_t u2 = hd.z & 0x3f3f3f3fu, u3 = ((hd.w >> 4) & 0x0f0f0f0fu) | (((hd.z >> 6) & 0x03030303u) << 4);
            uint32_t mr[8];
#pragma unroll
            for (int j = 0; j < 8; j++) mr[j] = h2u(byte_to_h2(j < 4 ? u2 : u3, j & 3));
            tc::tmem_st8(lane_base + AM_OFF + 8 * (i & 1), mr);
        }
        tc::tmem_st32(lane_base + A_OFF + 32 * MEMBER, ra);
        // ---- second half (group MEMBER + 2): its arithmetic overlaps the TMEM store above
        uint32_t rb[32];
        dequant_group(rb, qb0, qb1, byte_to_h2(u1, 2 * MEMBER), byte_to_h2(u1, 2 * MEMBER + 1), k_lo, k_hi, k_magic);
        tc::wait_st();
        tc::fence_before();
        __syncwarp();
        if (lane == 0) {
            tc::mbar_arrive(bar(A_FULL + 0));
            if (MEMBER == 0) tc::mbar_arrive(bar(HDR_FULL + (i & 3)));
        }
        TCP_WAIT(2, bar(A_EMPTY + 1), (i & 1) ^ 1);
        tc::fence_after();
        tc::tmem_st32(lane_base + A_OFF + 64 + 32 * MEMBER, rb);
        // ---- next super-block's load + first group while that store is in flight (only if its weights have landed:
        // the hand-off of this half must not wait for HBM)
        const uint32_t nx = i + 1;
        const bool more = nx < n_total;
        const bool early = more && __all_sync(0xffffffffu, tc::mbar_test_wait(bar(W_FULL + nx % S_W), (nx / S_W) & 1));
        if (early) fetch(nx);
        tc::wait_st();
        tc::fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(bar(A_FULL + 1));
        if (more && !early) { TCP_WAIT(0, bar(W_FULL + nx % S_W), (nx / S_W) & 1); fetch(nx); }
    }
    if (PROF && blockIdx.x == 0 && q == 0 && lane == 0) {
        long long *o = p.prof + 16 + 8 * MEMBER;
        o[0] = clock64() - pt0; o[1] = pw[0]; o[2] = pw[1]; o[3] = pw[2]; o[4] = pw[3]; o[5] = n_total;
    }
}

// signed byte b of word w as half2 (v, v), v in [-128,127] exact
__device__ __forceinline__ __half2 sbyte_to_h2(uint32_t w, int b) {
    const uint32_t v = ((w >> (8 * b)) & 0xffu) ^ 0x80u;
    return __hsub2(u2h(v | (v << 16) | 0x64006400u), u2h(0x64806480u));     // (1024 + v + 128) - 1152
}
// Q6_K: weight = d * sc * (q - 32), q = ql | qh << 4, |sc * (q - 32)| <= 4096 does not fit f16 exactly, so the A operand is
// the sum of two exact f16 integers fed to two MMAs on the same accumulator:
//     A1 = sc * ql             (0 .. 15 * 128)          = (1024 + 16 ql) * (sc / 16) - 64 sc      nibble at mantissa bits 4-7
//     A2 = 16 sc * (qh - 2)    (multiples of 16 <= 4096) = (1024 + 256 qh) * (sc / 16) - 96 sc     2-bit field at mantissa bits 8-9
// (the bias product stays <= 96 * 128, inside the f16 range; a bias in the low mantissa bits would need 1024 * sc).
// One 32-element group G of a 128-element half (dequantize_row_q6_K, R/ggml/src/ggml-quants.c:1709): ql bytes 32 (G & 1) + l,
// nibble G >> 1; qh byte l, bits 2G..2G+1; K-steps 2G (l < 16) and 2G + 1.
template <int G>
__device__ __forceinline__ void dequant_group_q6(uint32_t (&r1)[16], uint32_t (&r2)[16], const uint4 &x0, const uint4 &x1, const uint4 &y0, const uint4 &y1,
                                                 __half2 sa, __half2 sb, uint32_t k_hi, uint32_t k_q2, uint32_t k_magic) {
    const uint32_t xs[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w}, ys[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
    const __half2 k16 = __float2half2_rn(0.0625f), k64 = __float2half2_rn(-64.f), k96 = __float2half2_rn(-96.f);
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const __half2 sc = w < 4 ? sa : sb;
        const __half2 m = __hmul2(sc, k16), c1 = __hmul2(sc, k64), c2 = __hmul2(sc, k96);
        const uint32_t x = xs[w], y = ys[w];
        const uint32_t a = G < 2 ? x << 4 : x, b = G < 2 ? x >> 4 : x >> 8;            // elements (4w, 4w+2) and (4w+1, 4w+3)
        const uint32_t a2 = y << (8 - 2 * G), b2 = y >> (2 * G);
        r1[2 * w]     = h2u(__hfma2(u2h(and_or(a, k_hi, k_magic)), m, c1));
        r1[2 * w + 1] = h2u(__hfma2(u2h(and_or(b, k_hi, k_magic)), m, c1));
        r2[2 * w]     = h2u(__hfma2(u2h(and_or(a2, k_q2, k_magic)), m, c2));
        r2[2 * w + 1] = h2u(__hfma2(u2h(and_or(b2, k_q2, k_magic)), m, c2));
    }
}
using tc::tmem_st16;
// dequant role for Q6_K: MEMBER m handles groups m and m + 2 of each half (same 32 ql bytes: low / high nibbles)
template <int MEMBER, bool PROF>
__device__ __forceinline__ void dequant_q6_K(const TcArgs &p, uint8_t *smem, uint32_t bar0, uint32_t tmem, int q, int lane) {
    auto bar = [&](int i) { return bar0 + 8u * i; };
    const int row = 32 * q + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    float2 *hdr_sm = (float2 *)(smem + OFF_HDR);
    const uint32_t k_hi = p.k_hi, k_magic = p.k_magic, k_q2 = p.k_q2;
    const int sw_ql = row & 7, sw_qh = (row >> 1) & 3;                  // TMA swizzle phases of this row
    uint32_t i = 0;
    long long pw[4] = {0, 0, 0, 0}; const long long pt0 = PROF ? clock64() : 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int ks = item / p.n_rb, rb = item % p.n_rb;
        const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
        const int64_t grow = (int64_t)rb * 128 + row;
        const __half *dp = p.w_d + grow * p.nbr;
        float dcur = (MEMBER == 0 && grow < p.N) ? __half2float(__ldg(dp + sb0)) : 0.f;
        for (int sb = sb0; sb < sb1; sb++, i++) {
            const uint32_t s = i % S_W6;
            TCP_WAIT(0, bar(W_FULL + s), (i / S_W6) & 1);
            const uint8_t *st = smem + OFF_W + s * W6_STAGE;
            const uint8_t *qlp = st + row * 128, *qhp = st + W6_QH + row * 64;
            const uint4 sc = *(const uint4 *)(st + W6_SC + row * 16);
            const float dnow = dcur;
            if (MEMBER == 0 && sb + 1 < sb1 && grow < p.N) dcur = __half2float(__ldg(dp + sb + 1));
            const uint32_t scw[4] = {sc.x, sc.y, sc.z, sc.w};
#pragma unroll
            for (int h = 0; h < 2; h++) {
                // this half's 32 ql bytes + 32 qh bytes (loaded per half: keeps the role inside its 96 registers)
                uint4 xl[2], yh[2];
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    xl[c] = *(const uint4 *)(qlp + (((4 * h + 2 * MEMBER + c) ^ sw_ql) << 4));
                    yh[c] = *(const uint4 *)(qhp + (((2 * h + c) ^ sw_qh) << 4));
                }
                if (h == 1) {                                       // last read of the stage
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(bar(W_EMPTY + s));
                }
                // group MEMBER: dequantise, store; group MEMBER + 2 is dequantised while those stores are in flight
                uint32_t r1a[16], r2a[16];
                dequant_group_q6<MEMBER>(r1a, r2a, xl[0], xl[1], yh[0], yh[1], sbyte_to_h2(scw[2 * h], 2 * MEMBER), sbyte_to_h2(scw[2 * h], 2 * MEMBER + 1),
                                         k_hi, k_q2, k_magic);
                TCP_WAIT(1 + h, bar(A_EMPTY + h), (i & 1) ^ 1);
                tc::fence_after();
                if (MEMBER == 0 && h == 0) hdr_sm[(i & 3) * 128 + row] = make_float2(dnow, 0.f);
                const uint32_t a = lane_base + A6_OFF + A6_HALF * h;
                tmem_st16(a + 16 * MEMBER, r1a); tmem_st16(a + A6_SECOND + 16 * MEMBER, r2a);
                uint32_t r1b[16], r2b[16];
                dequant_group_q6<MEMBER + 2>(r1b, r2b, xl[0], xl[1], yh[0], yh[1], sbyte_to_h2(scw[2 * h + 1], 2 * MEMBER), sbyte_to_h2(scw[2 * h + 1], 2 * MEMBER + 1),
                                             k_hi, k_q2, k_magic);
                tmem_st16(a + 16 * (MEMBER + 2), r1b); tmem_st16(a + A6_SECOND + 16 * (MEMBER + 2), r2b);
                tc::wait_st();
                tc::fence_before();
                __syncwarp();
                if (lane == 0) {
                    tc::mbar_arrive(bar(A_FULL + h));
                    if (MEMBER == 0 && h == 0) tc::mbar_arrive(bar(HDR_FULL + (i & 3)));
                }
            }
        }
    }
    if (PROF && blockIdx.x == 0 && q == 0 && lane == 0) {
        long long *o = p.prof + 16 + 8 * MEMBER;
        o[0] = clock64() - pt0; o[1] = pw[0]; o[2] = pw[1]; o[3] = pw[2]; o[4] = pw[3]; o[5] = i;
    }
}

// epilogue role: 8 warps; warp (q, hf) owns TMEM lanes 32q..32q+31 (weight rows) and columns [40hf, 40hf+40)
template <int FMT, bool PROF>
__device__ __forceinline__ void epilogue_role(const TcArgs &p, uint8_t *smem, uint32_t bar0, uint32_t tmem, int warp, int lane) {
    auto bar = [&](int i) { return bar0 + 8u * i; };
    const int q = warp & 3, hf = (warp - 12) >> 2, row = 32 * q + lane, c_base = 40 * hf;
    const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
    const float2 *hdr_sm = (const float2 *)(smem + OFF_HDR);
    float *dx_sm = (float *)(smem + OFF_DX) + (warp - 12) * 80;
    const bool tail = c_base + 32 < p.T;                                // columns [32,40) of this warp hold tokens
    pdl_wait();                                                         // dT is read and the output planes are written below
    uint32_t i = 0;
    long long pw[3] = {0, 0, 0}; const long long pt0 = PROF ? clock64() : 0;
    long long sec[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}; long long ck = 0;
#define EPI_CK(k) do { if (PROF) { const long long n_ = clock64(); sec[k] += n_ - ck; ck = n_; } } while (0)
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        const int ks = item / p.n_rb, rb = item % p.n_rb;
        const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
        unsigned long long acc2[20];                                    // 40 columns as f32 pairs
#pragma unroll
        for (int c = 0; c < 20; c++) acc2[c] = 0ull;
        // dx (activation scales) of a super-block: 40 columns, fetched one super-block ahead into two registers per
        // lane, staged through a per-warp smem strip and read back as broadcast float4
        const bool va = c_base + lane < p.T, vb = lane < 8 && c_base + 32 + lane < p.T;
        float dxa = va ? __ldg(p.dT + (int64_t)sb0 * EB_DT_PITCH + c_base + lane) : 0.f;
        float dxb = vb ? __ldg(p.dT + (int64_t)sb0 * EB_DT_PITCH + c_base + 32 + lane) : 0.f;
        for (int sb = sb0; sb < sb1; sb++, i++) {
            if (PROF) ck = clock64();
            float *dxs = dx_sm + (i & 1) * 40;
            dxs[lane] = dxa;
            if (lane < 8) dxs[32 + lane] = dxb;
            if (sb + 1 < sb1) {
                if (va) dxa = __ldg(p.dT + (int64_t)(sb + 1) * EB_DT_PITCH + c_base + lane);
                if (vb) dxb = __ldg(p.dT + (int64_t)(sb + 1) * EB_DT_PITCH + c_base + 32 + lane);
            }
            EPI_CK(0);
            TCP_WAIT(0, bar(HDR_FULL + (i & 3)), (i >> 2) & 1);
            EPI_CK(1);
            const float2 h = hdr_sm[(i & 3) * 128 + row];
            const unsigned long long hd2 = pk2(h.x, h.x), hm2 = pk2(-h.y, -h.y);
            const uint32_t s = i & 1;
            TCP_WAIT(1, bar(D_FULL + s), (i >> 1) & 1);
            EPI_CK(2);
            if (PROF && i == 0 && blockIdx.x == 0 && warp == 12 && lane == 0) p.prof[64 + 6] = clock64();
            tc::fence_after();
            EPI_CK(3);
            const long long te_ = PROF ? clock64() : 0;
            __syncwarp();
            const uint32_t t_main = lane_base + s * (FMT == 0 ? D_STRIDE : D6_STRIDE) + c_base, t_min = t_main + D_MIN;
            // two passes keep the live registers under the 120 granted by setmaxnreg: columns [0,16) then [16,40)
            auto fold4 = [&](const uint32_t *vm, const uint32_t *vn, int col0, int n4) {
#pragma unroll
                for (int e4 = 0; e4 < n4; e4++) {
                    const ulonglong2 dx = *(const ulonglong2 *)(dxs + col0 + 4 * e4);
                    const int k = 4 * e4, a = (col0 >> 1) + 2 * e4;
                    unsigned long long t0 = mul2(hd2, pk2(__uint_as_float(vm[k]), __uint_as_float(vm[k + 1])));
                    unsigned long long t1 = mul2(hd2, pk2(__uint_as_float(vm[k + 2]), __uint_as_float(vm[k + 3])));
                    if (FMT == 0) {
                        t0 = fma2(hm2, pk2(__uint_as_float(vn[k]), __uint_as_float(vn[k + 1])), t0);
                        t1 = fma2(hm2, pk2(__uint_as_float(vn[k + 2]), __uint_as_float(vn[k + 3])), t1);
                    }
                    acc2[a] = fma2(dx.x, t0, acc2[a]);
                    acc2[a + 1] = fma2(dx.y, t1, acc2[a + 1]);
                }
            };
            {
                uint32_t vm[16], vn[16];
                tc::tmem_ld16(t_main, vm);
                if (FMT == 0) tc::tmem_ld16(t_min, vn);
                tc::wait_ld();
                EPI_CK(4);
                fold4(vm, vn, 0, 4);
                EPI_CK(5);
            }
            {
                uint32_t vm[16], vn[16], wm[8], wn[8];
                tc::tmem_ld16(t_main + 16, vm);
                if (FMT == 0) tc::tmem_ld16(t_min + 16, vn);
                if (tail) { tc::tmem_ld8(t_main + 32, wm); if (FMT == 0) tc::tmem_ld8(t_min + 32, wn); }
                tc::wait_ld();
                EPI_CK(6);
                tc::fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(D_EMPTY + s));       // accumulators are in registers: hand the stage back
                EPI_CK(7);
                fold4(vm, vn, 16, 4);
                if (tail) fold4(wm, wn, 32, 2);
                EPI_CK(8);
            }
            if (PROF) pw[2] += clock64() - te_;
        }
        const long long tst_ = PROF ? clock64() : 0;
        // plane store: lanes = consecutive weight rows -> every warp store is one full 128-byte line; one running pointer,
        // warp-uniform column count (40 stores with per-store address arithmetic and branches cost ~3000 clk per item)
        const int64_t grow = (int64_t)rb * 128 + row;
        if (grow < p.N) {
            float *o = p.out + ((int64_t)ks * p.T + c_base) * p.N + grow;
            const int64_t ld = p.N;
            const int nc = min(40, p.T - c_base);
            if (nc == 40) {
#pragma unroll
                for (int c = 0; c < 20; c++) { o[0] = __uint_as_float((uint32_t)acc2[c]); o[ld] = __uint_as_float((uint32_t)(acc2[c] >> 32)); o += 2 * ld; }
            } else {
#pragma unroll
                for (int c = 0; c < 20; c++) {
                    if (2 * c < nc) o[0] = __uint_as_float((uint32_t)acc2[c]);
                    if (2 * c + 1 < nc) o[ld] = __uint_as_float((uint32_t)(acc2[c] >> 32));
                    o += 2 * ld;
                }
            }
        }
        // tensor parallel (common.cuh TpPush): the same values again, as {value, call number} pairs, into every peer's memory — posted
        // NVLink stores; the peers' consumer kernels take an element as arrived when its upper word is the call number
        if (p.tp.n_peers && grow < p.N) {
            const unsigned long long tag = (unsigned long long)(*(volatile const unsigned *)p.tp.seq + 1u) << 32;
            const int nc = min(40, p.T - c_base);
            for (int j = 0; j < p.tp.n_peers; j++) {
                unsigned long long *o = p.tp.peer[j] + ((int64_t)ks * p.T + c_base) * p.N + grow;
                const int64_t ld = p.N;
#pragma unroll
                for (int c = 0; c < 20; c++) {
                    if (2 * c < nc) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(o), "l"(tag | (acc2[c] & 0xffffffffull)) : "memory");
                    if (2 * c + 1 < nc) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(o + ld), "l"(tag | (acc2[c] >> 32)) : "memory");
                    o += 2 * ld;
                }
            }
        }
        if (PROF && blockIdx.x == 0 && warp == 12 && lane == 0) p.prof[89] += clock64() - tst_;
    }
    if (p.tp.n_peers && blockIdx.x == 0 && warp == 12 && lane == 0) {       // header: how many planes this rank pushes in this call
        const unsigned long long tag = (unsigned long long)(*(volatile const unsigned *)p.tp.seq + 1u) << 32;
        for (int j = 0; j < p.tp.n_peers; j++) asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p.tp.peer[j] + p.tp.hdr), "l"(tag | (unsigned)(p.n_items / p.n_rb)) : "memory");
    }
    if (PROF && blockIdx.x == 0 && warp == 12 && lane == 0) p.prof[64 + 7] = clock64();
    if (PROF && blockIdx.x == 0 && warp == 12 && lane == 0) { long long *o = p.prof + 32; o[0] = clock64() - pt0; o[1] = pw[0]; o[2] = pw[1]; o[3] = pw[2]; for (int k = 0; k < 9; k++) p.prof[80 + k] = sec[k]; }
#undef EPI_CK
}

// FMT 0 = Q4_K (tmW = the 144-byte blocks, tmBS = 16-element activation sums); FMT 1 = Q6_K (tmW = ql plane, tmQH / tmSC = qh / scales planes)
template <int FMT, bool PROF>
__global__ void __launch_bounds__(TC_THREADS, 1)
mmq_tc_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmBS,
              const __grid_constant__ CUtensorMap tmQH, const __grid_constant__ CUtensorMap tmSC, TcArgs p) {
    constexpr int NSW = FMT == 0 ? S_W : S_W6;
    extern __shared__ uint8_t smem_raw[];
    // pointer arithmetic on the __shared__ array (not an integer round-trip) keeps the address space: LDS/STS instead of generic LD/ST
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t sbase = tc::smem_u32(smem);
    const uint32_t bar0 = sbase + OFF_BAR;
    auto bar = [&](int i) { return bar0 + 8u * i; };
    volatile uint32_t *tmem_slot = (volatile uint32_t *)(smem + OFF_BAR + 8 * N_BARS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#define TCP_STAMP(k) do { if (PROF && blockIdx.x == 0) p.prof[64 + (k)] = clock64(); } while (0)
    if (threadIdx.x == 0) TCP_STAMP(0);

    if (threadIdx.x == 0) {
        for (int i = 0; i < NSW; i++) { tc::mbar_init(bar(W_FULL + i), 1); tc::mbar_init(bar(W_EMPTY + i), 8); }
        for (int i = 0; i < 2; i++) {
            tc::mbar_init(bar(B_FULL + i), 1); tc::mbar_init(bar(B_EMPTY + i), 1);
            tc::mbar_init(bar(BS_FULL + i), 1); tc::mbar_init(bar(BS_EMPTY + i), 1);
            tc::mbar_init(bar(A_FULL + i), 8); tc::mbar_init(bar(A_EMPTY + i), 1);
            tc::mbar_init(bar(D_FULL + i), 1); tc::mbar_init(bar(D_EMPTY + i), 8);
        }
        for (int i = 0; i < 4; i++) tc::mbar_init(bar(HDR_FULL + i), 4);
        tc::fence_barrier_init();
        tc::tma_prefetch_desc(&tmW); tc::tma_prefetch_desc(&tmB);
        if (FMT == 0) tc::tma_prefetch_desc(&tmBS); else { tc::tma_prefetch_desc(&tmQH); tc::tma_prefetch_desc(&tmSC); }
    }
    if (warp == 3) tc::tmem_alloc(tc::smem_u32((const void *)tmem_slot), 512);
    tc::fence_before();
    __syncthreads();
    tc::fence_after();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) TCP_STAMP(1);
    if (p.early_trigger) pdl_trigger();             // the next kernel may start its own prologue on SMs this grid has left

    if (warp < 4) {
        // ------------------------------------------------------------ weight producer (warp 0): the first tile goes out
        // before anything else (the register hand-over below costs several hundred cycles)
        auto issue_w = [&](uint32_t s, int sb, int rb) {
            if (FMT == 0) {
                tc::mbar_expect_tx(bar(W_FULL + s), W_STAGE);
                tc::tma_load_2d(sbase + OFF_W + s * W_STAGE, &tmW, sb * 144, rb * 128, bar(W_FULL + s));
            } else {
                const uint32_t dst = sbase + OFF_W + s * W6_STAGE;
                tc::mbar_expect_tx(bar(W_FULL + s), W6_STAGE);
                tc::tma_load_2d(dst, &tmW, sb * 128, rb * 128, bar(W_FULL + s));
                tc::tma_load_2d(dst + W6_QH, &tmQH, sb * 64, rb * 128, bar(W_FULL + s));
                tc::tma_load_2d(dst + W6_SC, &tmSC, sb * 16, rb * 128, bar(W_FULL + s));
            }
        };
        if (warp == 0 && lane == 0 && (int)blockIdx.x < p.n_items) {
            TCP_STAMP(2);
            issue_w(0, ((int)blockIdx.x / p.n_rb) * p.spp, (int)blockIdx.x % p.n_rb);
        }
        __syncwarp();
        tc::setmaxnreg_dec<40>();                   // hand registers to the epilogue warps
        if (warp == 0) {
            if (lane == 0) {
                uint32_t i = 0;
                long long pw[1] = {0}; const long long pt0 = PROF ? clock64() : 0;
                for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                    const int ks = item / p.n_rb, rb = item % p.n_rb;
                    const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                    for (int sb = sb0; sb < sb1; sb++, i++) {
                        if (i == 0) continue;           // already in flight
                        const uint32_t s = i % NSW, ph = (i / NSW) & 1;
                        TCP_WAIT(0, bar(W_EMPTY + s), ph ^ 1);
                        issue_w(s, sb, rb);
                    }
                }
                if (PROF && blockIdx.x == 0) { p.prof[0] = clock64() - pt0; p.prof[1] = pw[0]; }
            }
        } else if (warp == 2) {
            // ------------------------------------------------------------ activation producer
            if (lane == 0) {
                pdl_wait();                         // activations are the predecessor's output (the weight producer does not wait)
                uint32_t i = 0, nbx = 0;
                const uint32_t box_bytes = (uint32_t)p.npad * 128u;
                for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                    const int ks = item / p.n_rb;
                    const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                    for (int sb = sb0; sb < sb1; sb++, i++) {
                        if (FMT == 0 && (sb == sb0 || (sb & 3) == 0)) {
                            const uint32_t s2 = nbx & 1, ph2 = (nbx >> 1) & 1;
                            tc::mbar_wait(bar(BS_EMPTY + s2), ph2 ^ 1);
                            tc::mbar_expect_tx(bar(BS_FULL + s2), box_bytes);
                            tc::tma_load_2d(sbase + OFF_BS + s2 * B_BOX, &tmBS, 64 * (sb >> 2), 0, bar(BS_FULL + s2));
                            nbx++;
                        }
                        const uint32_t s = i & 1, ph = (i >> 1) & 1;
                        tc::mbar_wait(bar(B_EMPTY + s), ph ^ 1);
                        tc::mbar_expect_tx(bar(B_FULL + s), 4 * box_bytes);
#pragma unroll
                        for (int g = 0; g < 4; g++)
                            tc::tma_load_2d(sbase + OFF_B + s * B_STAGE + g * B_BOX, &tmB, sb * 256 + 64 * g, 0, bar(B_FULL + s));
                    }
                }
            }
        } else if (warp == 1) {
            // ------------------------------------------------------------ MMA issuer
            // The whole warp runs the loop so that every address stays in uniform registers; one elected lane issues.
            const bool leader = tc::elect_one();
            const uint32_t idesc = tc::idesc_f16_m128(p.npad);
            uint32_t i = 0, nbx = 0; int cur_bs = -1;
            long long pw[4] = {0, 0, 0, 0}; const long long pt0 = PROF ? clock64() : 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                const int ks = item / p.n_rb;
                const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                for (int sb = sb0; sb < sb1; sb++, i++) {
                    if (FMT == 0 && (sb == sb0 || (sb & 3) == 0)) {
                        if (cur_bs >= 0 && leader) tc::mma_commit(bar(BS_EMPTY + (cur_bs & 1)));
                        cur_bs = (int)nbx;
                        TCP_WAIT(0, bar(BS_FULL + (nbx & 1)), (nbx >> 1) & 1);
                        nbx++;
                    }
                    const uint32_t s = i & 1, ph = (i >> 1) & 1;
                    TCP_WAIT(1, bar(B_FULL + s), ph);
                    TCP_WAIT(2, bar(D_EMPTY + s), ph ^ 1);
                    const uint32_t d_main = tmem + s * (FMT == 0 ? D_STRIDE : D6_STRIDE), d_min = d_main + D_MIN;
                    const uint64_t bdesc = tc::smem_desc_sw128(sbase + OFF_B + s * B_STAGE);
                    const uint64_t mdesc = tc::smem_desc_sw128(sbase + OFF_BS + (cur_bs & 1) * B_BOX + 32 * (sb & 3));
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        TCP_WAIT(3, bar(A_FULL + h), i & 1);
                        if (i == 0 && h == 0 && lane == 0) TCP_STAMP(4);
                        tc::fence_after();
                        if (leader) {
                            const uint64_t b_lo = bdesc + (uint64_t)((2 * h) * (B_BOX >> 4)), b_hi = bdesc + (uint64_t)((2 * h + 1) * (B_BOX >> 4));
                            if (FMT == 0) {
                                tc::mma_f16_ts_x8(d_main, tmem + A_OFF + 64 * h, b_lo, b_hi, idesc, h != 0);
                                if (h == 0) tc::mma_f16_ts(d_min, tmem + AM_OFF + 8 * (i & 1), mdesc, idesc, 0);
                            } else {                                    // two exact addends of the weight, same activations, same accumulator
                                tc::mma_f16_ts_x8(d_main, tmem + A6_OFF + A6_HALF * h, b_lo, b_hi, idesc, h != 0);
                                tc::mma_f16_ts_x8(d_main, tmem + A6_OFF + A6_HALF * h + A6_SECOND, b_lo, b_hi, idesc, 1);
                            }
                            tc::mma_commit(bar(A_EMPTY + h));
                            if (h == 1) { tc::mma_commit(bar(B_EMPTY + s)); tc::mma_commit(bar(D_FULL + s)); if (i == 0) TCP_STAMP(5); }
                        }
                        __syncwarp();
                    }
                }
            }
            if (PROF && blockIdx.x == 0 && lane == 0) { long long *o = p.prof + 8; o[0] = clock64() - pt0; o[1] = pw[0]; o[2] = pw[1]; o[3] = pw[2]; o[4] = pw[3]; }
        }
        __syncwarp();
    } else if (warp < 12) {
        if (FMT == 0) {
            if (warp < 8) dequant_q4_K<0, PROF>(p, smem, bar0, tmem, warp & 3, lane);
            else dequant_q4_K<1, PROF>(p, smem, bar0, tmem, warp & 3, lane);
        } else {
            if (warp < 8) dequant_q6_K<0, PROF>(p, smem, bar0, tmem, warp & 3, lane);
            else dequant_q6_K<1, PROF>(p, smem, bar0, tmem, warp & 3, lane);
        }
    } else {
        tc::setmaxnreg_inc<120>();   // CTA pool: 128*40 + 256*96 + 256*120 = 60416 <= 640*96
        epilogue_role<FMT, PROF>(p, smem, bar0, tmem, warp, lane);
    }
    tc::fence_before();
    __syncthreads();
    if (warp == 3) tc::tmem_dealloc(tmem, 512);
    if (threadIdx.x == 96 + 0) TCP_STAMP(8);
#undef TCP_STAMP
}
// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

int make_map_2d(CUtensorMap *m, CUtensorMapDataType dt, int esize, const void *base, uint64_t d0, uint64_t d1, uint64_t stride1_bytes,
                uint32_t b0, uint32_t b1, CUtensorMapSwizzle sw) {
    const cuuint64_t gdim[2] = {d0, d1}, gstr[1] = {stride1_bytes};
    const cuuint32_t box[2] = {b0, b1}, estr[2] = {1, 1};
    (void)esize;
    const CUresult r = g_encode(m, dt, 2, const_cast<void *>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return eb_set_err(EB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d): dims %llu x %llu stride %llu box %u x %u", (int)r,
                                             (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)stride1_bytes, b0, b1);
    return EB_OK;
}

}  // namespace

int eb_tc_map_2d_f16(CUtensorMap *m, const void *base, uint64_t d0, uint64_t d1, uint64_t stride1_bytes, uint32_t b0, uint32_t b1) {
    if (!g_encode) return eb_set_err(EB_ERR_CUDA, "tensor maps unavailable");
    return make_map_2d(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, d0, d1, stride1_bytes, b0, b1, CU_TENSOR_MAP_SWIZZLE_128B);
}
int eb_tc_ready() { return g_encode != nullptr; }

int eb_mmq_tc_init() {
    if (!g_encode) {
        void *fn = nullptr; cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
            cudaGetLastError();
            return eb_set_err(EB_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
        }
        g_encode = (EncodeTiledFn)fn;
    }
    EB_CUDA(cudaFuncSetAttribute(mmq_tc_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    EB_CUDA(cudaFuncSetAttribute(mmq_tc_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    EB_CUDA(cudaFuncSetAttribute(mmq_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    EB_CUDA(cudaFuncSetAttribute(mmq_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    return EB_OK;
}

int eb_mmq_tc_supported(int type, int64_t N, int64_t K, int T) {
    static const int off = getenv("EB_NO_TC") ? 1 : 0;
    if (off || !g_encode) return 0;
    static const int off6 = getenv("EB_NO_TC_Q6") ? 1 : 0;
    // T <= 8 (draft head) stays on the dp4a / mma.sync kernels: measured 17.9 us vs 26.0 us for 22016 x 4096 at T = 8 — this
    // kernel is bound by its dequant role, not by N.  EB_TC_TMIN=1 routes small batches here (N = 16 MMAs) for experiments.
    static const int tmin = getenv("EB_TC_TMIN") ? atoi(getenv("EB_TC_TMIN")) : 9;
    static const int tmin6 = getenv("EB_TC_TMIN6") ? atoi(getenv("EB_TC_TMIN6")) : 9;
    return ((type == EB_Q4_K && T >= tmin) || (type == EB_Q6_K && !off6 && T >= tmin6)) && K % 256 == 0 && N >= 128 && T <= 80;
}

// split-K factor: planes <= 8, minimise rounds x (super-blocks per item + pipeline fill)
static int tc_pick_ksplit(int64_t N, int nbr) {
    const int n_rb = (int)cdiv(N, 128), G = eb_sm_count();
    static const int force = getenv("EB_TC_KS") ? atoi(getenv("EB_TC_KS")) : 0;     // tuning experiments
    if (force > 0) { const int ks = std::min(force, nbr), spp = (int)cdiv(nbr, ks); return (int)cdiv(nbr, spp); }
    int best = 1; double best_cost = 1e30;
    for (int ks = 1; ks <= 8 && ks <= nbr; ks++) {
        const int spp = (int)cdiv(nbr, ks), eff_ks = (int)cdiv(nbr, spp);
        if (eff_ks != ks) continue;
        const int rounds = (int)cdiv((int64_t)n_rb * ks, G);
        // per-item cost in super-block units.  Fit of the EB_TC_KS sweep of 22016 x 4096 (profiles/r2_summary.md):
        // t = 0.74 us + 0.835 us per super-block + 0.82 us per item, i.e. one item ~ one super-block
        static const double c_item = getenv("EB_TC_CITEM") ? atof(getenv("EB_TC_CITEM")) : 1.0;
        const double cost = rounds * (spp + c_item) + 0.25 * ks;
        if (cost < best_cost) { best_cost = cost; best = ks; }
    }
    return best;
}

// planes [ks][T][N] (unreduced split-K partials; ks returned)
int eb_mmq_tc_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st, const TpPush *push) {
    if (!eb_mmq_tc_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq_tc: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    if (!a.h16 || !a.dT || !a.bsum) return eb_set_err(EB_ERR_ARG, "mmq_tc: activations lack the f16 planes");
    const int nbr = (int)(K / 256);
    const int ks = tc_pick_ksplit(N, nbr);
    TcArgs p;
    p.dT = a.dT; p.out = planes; p.N = N; p.nbr = nbr; p.T = a.T; p.npad = (a.T + 15) / 16 * 16;
    p.k_lo = 0x000f000fu; p.k_hi = 0x00f000f0u; p.k_magic = 0x64006400u; p.k_q2 = 0x03000300u; p.w_d = nullptr;
    if (push) p.tp = *push; else { p.tp = TpPush{}; }
    p.n_rb = (int)cdiv(N, 128); p.spp = (int)cdiv(nbr, ks); p.n_items = p.n_rb * ks;
    static const int early = getenv("EB_TC_LATE_TRIGGER") ? 0 : 1;
    p.early_trigger = early;
    CUtensorMap tmW, tmB, tmBS, tmQH, tmSC;
    int rc;
    if (type == EB_Q4_K) {
        if ((rc = make_map_2d(&tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, W, (uint64_t)nbr * 144, (uint64_t)N, (uint64_t)nbr * 144, 144, 128, CU_TENSOR_MAP_SWIZZLE_NONE))) return rc;
        tmQH = tmW; tmSC = tmW;
    } else {                                             // Q6_K planes (common.cuh WPlanes): ql | qh | scales | d
        const uint8_t *base = (const uint8_t *)W; const uint64_t nb = (uint64_t)N * nbr;
        if ((rc = make_map_2d(&tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, (uint64_t)nbr * 128, (uint64_t)N, (uint64_t)nbr * 128, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
        if ((rc = make_map_2d(&tmQH, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base + nb * 128, (uint64_t)nbr * 64, (uint64_t)N, (uint64_t)nbr * 64, 64, 128, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
        if ((rc = make_map_2d(&tmSC, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base + nb * 192, (uint64_t)nbr * 16, (uint64_t)N, (uint64_t)nbr * 16, 16, 128, CU_TENSOR_MAP_SWIZZLE_NONE))) return rc;
        p.w_d = (const __half *)(base + nb * 208);
    }
    if ((rc = make_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, a.h16, (uint64_t)K, (uint64_t)a.T, (uint64_t)K * 2, 64, (uint32_t)p.npad, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = make_map_2d(&tmBS, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, a.bsum, (uint64_t)K / 16, (uint64_t)a.T, (uint64_t)K / 16 * 2, 64, (uint32_t)p.npad, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    static long long *prof_buf = nullptr; static const int want_prof = getenv("EB_TC_PROF") ? 1 : 0;
    if (want_prof && !prof_buf) { cudaMalloc((void **)&prof_buf, 512 * 8); cudaMemset(prof_buf, 0, 512 * 8); }
    p.prof = want_prof ? prof_buf : nullptr;
    const int grid = std::min(p.n_items, eb_sm_count());
    if (type == EB_Q4_K) {
        if (want_prof) mmq_tc_kernel<0, true><<<grid, TC_THREADS, SMEM_BYTES, st>>>(tmW, tmB, tmBS, tmQH, tmSC, p);
        else eb_launch_pdl_c(EB_PDL_GEMM, mmq_tc_kernel<0, false>, dim3(grid), dim3(TC_THREADS), SMEM_BYTES, st, tmW, tmB, tmBS, tmQH, tmSC, p);
    } else {
        if (want_prof) mmq_tc_kernel<1, true><<<grid, TC_THREADS, SMEM_BYTES, st>>>(tmW, tmB, tmBS, tmQH, tmSC, p);
        else eb_launch_pdl_c(EB_PDL_GEMM, mmq_tc_kernel<1, false>, dim3(grid), dim3(TC_THREADS), SMEM_BYTES, st, tmW, tmB, tmBS, tmQH, tmSC, p);
    }
    EB_CHECK_LAUNCH();
    *ks_out = ks;
    if (want_prof) {
        long long h[512]; cudaStreamSynchronize(st); cudaMemcpy(h, prof_buf, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[tc prof] N=%lld K=%lld T=%d ks=%d items=%d grid=%d | Wprod total %lld wait %lld | MMA total %lld bs %lld b %lld d_empty %lld a_full %lld | "
                "DQ0 total %lld w_full %lld a_empty0 %lld a_empty1 %lld - %lld n=%lld | DQ1 total %lld w_full %lld a_e0 %lld a_e1 %lld - %lld | EPI total %lld hdr %lld d_full %lld work %lld\n",
                (long long)N, (long long)K, a.T, ks, p.n_items, grid, h[0], h[1], h[8], h[9], h[10], h[11], h[12], h[16], h[17], h[18], h[19], h[20], h[21],
                h[24], h[25], h[26], h[27], h[28], h[32], h[33], h[34], h[35]);
        fprintf(stderr, "[tc epi sections] top %lld | hdr wait %lld | d_full wait %lld | fence_after %lld | ld1 %lld | fold1 %lld | ld2 %lld | arrive %lld | fold2 %lld | final stores %lld\n",
                h[80], h[81], h[82], h[83], h[84], h[85], h[86], h[87], h[88], h[89]);
        cudaMemsetAsync(prof_buf + 89, 0, 8, st);
        fprintf(stderr, "[tc timeline clk since entry] setup %lld | first TMA issue %lld | DQ sees W %lld | MMA sees A %lld | first D commit %lld | EPI sees D %lld | EPI done %lld | exit %lld\n",
                h[65] - h[64], h[66] - h[64], h[67] - h[64], h[68] - h[64], h[69] - h[64], h[70] - h[64], h[71] - h[64], h[72] - h[64]);
    }
    return EB_OK;
}
