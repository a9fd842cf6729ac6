// mmq_small.cu — draft-head GEMM: T <= 8 activation rows (one top-k level of the EAGLE tree) against Q4_K / Q6_K weights.
// Replaces mul_mat_vec_q / mul_mat_q at small batch (R/ggml/src/ggml-cuda/mmvq.cu, mmq.cuh) on the draft path of
// R/examples/speculative/speculative-eagle.cpp:505-520; per-block integer sums are exact (int8 mma.sync), so results differ
// from ggml_vec_dot_q4_K_q8_K / ggml_vec_dot_q6_K_q8_K (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:6266,7903) only by the order
// of the f32 additions over super-blocks.
//
// At T = 8 a GEMM is a pure weight stream (7 of them per level, 9-107 MB each) and every microsecond of fixed latency
// shows: the general kernel (mmq.cu) spends ~4.7 us per launch outside its streaming phase.  This kernel is built around
// the latency chain instead:
//   * persistent: one CTA per SM walks (32-row block, K split) items; the whole activation matrix (8 x K int8 + scales,
//     <= 116 KB for K = 11008) is staged in shared memory ONCE per CTA, so there is no per-super-block synchronisation;
//   * 16 warps = 2 row groups (16 rows, one m16n8k32 A tile) x 8 K groups: warp (rg, kg) owns super-blocks kg, kg+8, ..
//     of its rows, all warps' weight loads are in flight together, and the next super-block (also across items) is
//     requested before the current one is consumed;
//   * the first weight loads are issued BEFORE griddepcontrol.wait (weights are constants), so they overlap the
//     predecessor kernel's tail;
//   * the 8 K groups are summed through shared memory in fixed order (deterministic), one plane per K split.
#include "common.cuh"
#include <algorithm>
#include <cstdlib>
#include <type_traits>

namespace {

constexpr int SM_THREADS = 512, RG = 2, KG = 8;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void mma_u8s8_k32(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_s8s8_k16(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ void mma_f16_k16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// one super-block of activations in shared memory: 8 columns x [256 quants + pad] | d[8] | bsum16[8][16]
template <int FMT> struct Tile {
    static constexpr int PITCH = FMT == 0 ? 288 : 272;      // conflict-free 8-byte (Q4_K) / 4-byte (Q6_K) B-fragment loads
    static constexpr int D_OFF = 8 * PITCH, BS_OFF = D_OFF + 32;
    static constexpr int BYTES = (BS_OFF + (FMT == 0 ? 256 : 0) + 127) & ~127;
};

// weight registers of one (16-row group, super-block): two rows per lane (g, g + 8), the lane's 4 x 8 bytes of every row
struct W4 { int4 h0, h1; uint2 q0[4], q1[4]; };
struct W6 { uint32_t ql0[8], ql1[8], qh0[4], qh1[4]; int4 s0, s1; __half d0, d1; };

__device__ __forceinline__ void load_w(W4 &w, const void *W, int64_t r0, int64_t r1, int nbr, int64_t, int sb, int t) {
    const block_q4_K *b0 = (const block_q4_K *)W + r0 * nbr + sb, *b1 = (const block_q4_K *)W + r1 * nbr + sb;
    w.h0 = __ldg((const int4 *)b0); w.h1 = __ldg((const int4 *)b1);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        w.q0[j] = __ldg((const uint2 *)(b0->qs + j * 32 + 8 * t));
        w.q1[j] = __ldg((const uint2 *)(b1->qs + j * 32 + 8 * t));
    }
}
__device__ __forceinline__ void load_w(W6 &w, const void *W, int64_t r0, int64_t r1, int nbr, int64_t nb_total, int sb, int t) {
    const WPlanes p{(const uint8_t *)W, nb_total};
    const int64_t b0 = r0 * nbr + sb, b1 = r1 * nbr + sb;
    // ql: [n][c&1][u] words at n*64 + 32*(c&1) + 16u + 4t ; qh: [n][u] words at n*32 + 16u + 4t
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int off = (i >> 2) * 64 + ((i >> 1) & 1) * 32 + (i & 1) * 16 + 4 * t;
        w.ql0[i] = __ldg((const uint32_t *)(p.q6k_ql() + b0 * 128 + off));
        w.ql1[i] = __ldg((const uint32_t *)(p.q6k_ql() + b1 * 128 + off));
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int off = (i >> 1) * 32 + (i & 1) * 16 + 4 * t;
        w.qh0[i] = __ldg((const uint32_t *)(p.q6k_qh() + b0 * 64 + off));
        w.qh1[i] = __ldg((const uint32_t *)(p.q6k_qh() + b1 * 64 + off));
    }
    w.s0 = __ldg((const int4 *)(p.q6k_sc() + b0 * 16)); w.s1 = __ldg((const int4 *)(p.q6k_sc() + b1 * 16));
    w.d0 = p.q6k_d()[b0]; w.d1 = p.q6k_d()[b1];
}

// facc += one super-block.  Accumulator layout of m16n8: c0,c1 = row g, columns 2t, 2t+1; c2,c3 = row g + 8.
__device__ __forceinline__ void compute_sb(float (&facc)[4], const W4 &w, const uint8_t *st, int g, int t) {
    using TL = Tile<0>;
    // 6-bit scales / mins of both rows (get_scale_min_k4, R/ggml/src/ggml-quants.c:631): bytes of u[0..1] = scales, u[2..3] = mins
    uint32_t u0[4], u1[4];
    const uint32_t k1 = 0x3f3f3f3f, k2 = 0x0f0f0f0f, k3 = 0x03030303;
    u0[0] = w.h0.y & k1; u0[1] = (w.h0.w & k2) | (((w.h0.y >> 6) & k3) << 4); u0[2] = w.h0.z & k1; u0[3] = ((w.h0.w >> 4) & k2) | (((w.h0.z >> 6) & k3) << 4);
    u1[0] = w.h1.y & k1; u1[1] = (w.h1.w & k2) | (((w.h1.y >> 6) & k3) << 4); u1[2] = w.h1.z & k1; u1[3] = ((w.h1.w >> 4) & k2) | (((w.h1.z >> 6) & k3) << 4);
    int acc[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 4; j++) {
#pragma unroll
        for (int hi = 0; hi < 2; hi++) {
            const int s = 2 * j + hi;
            uint32_t af[4];
            af[0] = (w.q0[j].x >> (4 * hi)) & 0x0F0F0F0F; af[1] = (w.q1[j].x >> (4 * hi)) & 0x0F0F0F0F;
            af[2] = (w.q0[j].y >> (4 * hi)) & 0x0F0F0F0F; af[3] = (w.q1[j].y >> (4 * hi)) & 0x0F0F0F0F;
            const int sc0 = (u0[s >> 2] >> (8 * (s & 3))) & 0xff, sc1 = (u1[s >> 2] >> (8 * (s & 3))) & 0xff;
            const uint2 b = *(const uint2 *)(st + g * TL::PITCH + s * 32 + 8 * t);
            int c[4] = {0, 0, 0, 0};
            mma_u8s8_k32(c, af, b.x, b.y);
            acc[0] += sc0 * c[0]; acc[1] += sc0 * c[1]; acc[2] += sc1 * c[2]; acc[3] += sc1 * c[3];
        }
    }
    // min term: sum_j m_j * bsum32_j == [m_(k/2)] (16) x bsum16 (16) as one f16 MMA, exact
    uint32_t am[4];
    {
        const int m0a = (u0[2] >> (8 * t)) & 0xff, m0b = (u0[3] >> (8 * t)) & 0xff;
        const int m1a = (u1[2] >> (8 * t)) & 0xff, m1b = (u1[3] >> (8 * t)) & 0xff;
        const __half2 x0 = __half2half2(__int2half_rn(m0a)), x1 = __half2half2(__int2half_rn(m1a));
        const __half2 x2 = __half2half2(__int2half_rn(m0b)), x3 = __half2half2(__int2half_rn(m1b));
        am[0] = *(const uint32_t *)&x0; am[1] = *(const uint32_t *)&x1; am[2] = *(const uint32_t *)&x2; am[3] = *(const uint32_t *)&x3;
    }
    const __half2 dm0 = *(const __half2 *)&w.h0.x, dm1 = *(const __half2 *)&w.h1.x;
    const float d0 = __low2float(dm0), mn0 = __high2float(dm0), d1 = __low2float(dm1), mn1 = __high2float(dm1);
    const float *dx = (const float *)(st + TL::D_OFF);
    const uint32_t *bs = (const uint32_t *)(st + TL::BS_OFF + g * 32);
    float cm[4] = {0.f, 0.f, 0.f, 0.f};
    mma_f16_k16(cm, am, bs[t], bs[t + 4]);
    const float x0 = dx[2 * t], x1 = dx[2 * t + 1];
    facc[0] += (d0 * x0) * (float)acc[0] - (mn0 * x0) * cm[0];
    facc[1] += (d0 * x1) * (float)acc[1] - (mn0 * x1) * cm[1];
    facc[2] += (d1 * x0) * (float)acc[2] - (mn1 * x0) * cm[2];
    facc[3] += (d1 * x1) * (float)acc[3] - (mn1 * x1) * cm[3];
}
__device__ __forceinline__ void compute_sb(float (&facc)[4], const W6 &w, const uint8_t *st, int g, int t) {
    using TL = Tile<1>;
    int acc[4] = {0, 0, 0, 0};
    const int sc0w[4] = {w.s0.x, w.s0.y, w.s0.z, w.s0.w}, sc1w[4] = {w.s1.x, w.s1.y, w.s1.z, w.s1.w};
#pragma unroll
    for (int s = 0; s < 16; s++) {                          // sub-block s: n = s/8, c = (s%8)/2, u = s%2  (dequantize_row_q6_K, R/ggml/src/ggml-quants.c:1709)
        const int n = s >> 3, c = (s & 7) >> 1, u = s & 1;
        const int qi = n * 4 + (c & 1) * 2 + u, hi = n * 2 + u;
        const uint32_t a0 = __vsubss4(((w.ql0[qi] >> (4 * (c >> 1))) & 0x0F0F0F0Fu) | (((w.qh0[hi] >> (2 * c)) & 0x03030303u) << 4), 0x20202020u);
        const uint32_t a1 = __vsubss4(((w.ql1[qi] >> (4 * (c >> 1))) & 0x0F0F0F0Fu) | (((w.qh1[hi] >> (2 * c)) & 0x03030303u) << 4), 0x20202020u);
        const int sc0 = (int)(int8_t)((sc0w[s >> 2] >> (8 * (s & 3))) & 0xff), sc1 = (int)(int8_t)((sc1w[s >> 2] >> (8 * (s & 3))) & 0xff);
        const uint32_t b = *(const uint32_t *)(st + g * TL::PITCH + s * 16 + 4 * t);
        int cc[4] = {0, 0, 0, 0};
        mma_s8s8_k16(cc, a0, a1, b);
        acc[0] += sc0 * cc[0]; acc[1] += sc0 * cc[1]; acc[2] += sc1 * cc[2]; acc[3] += sc1 * cc[3];
    }
    const float d0 = __half2float(w.d0), d1 = __half2float(w.d1);
    const float *dx = (const float *)(st + TL::D_OFF);
    const float x0 = dx[2 * t], x1 = dx[2 * t + 1];
    facc[0] += (d0 * x0) * (float)acc[0]; facc[1] += (d0 * x1) * (float)acc[1];
    facc[2] += (d1 * x0) * (float)acc[2]; facc[3] += (d1 * x1) * (float)acc[3];
}

struct SmallArgs {
    const void *W; int64_t N, nb_total; int nbr, n_rb, n_items, spp; ActQ a; float *out; const int *gate;
};

template <int FMT>
__global__ void __launch_bounds__(SM_THREADS, 1) mmq_small_kernel(SmallArgs p) {
    using TL = Tile<FMT>;
    using WR = typename std::conditional<FMT == 0, W4, W6>::type;
    extern __shared__ __align__(128) uint8_t smem[];
    float4 *red = (float4 *)(smem + (size_t)p.nbr * TL::BYTES);             // [2][KG][RG][32]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int rg = warp & (RG - 1), kg = warp / RG;
    pdl_trigger();

    // this warp's walk over (item, super-block) pairs, one step ahead of the compute loop
    int p_item = blockIdx.x, p_sb = 0, p_sb1 = 0; int64_t p_r0 = 0;
    auto seek = [&](bool advance) {                         // position on the next pair that exists; false at the end
        if (advance) p_sb += KG;
        while (p_item < p.n_items) {
            if (!advance) {
                const int ksp = p_item / p.n_rb, rb = p_item % p.n_rb;
                const int sb0 = ksp * p.spp; p_sb1 = min(p.nbr, sb0 + p.spp); p_sb = sb0 + kg;
                const int64_t row0 = (int64_t)rb * (16 * RG) + rg * 16;
                p_r0 = row0 < p.N ? row0 : 0;               // N % 16 == 0; out-of-range groups read row 0 and never store
                advance = true;
            }
            if (p_sb < p_sb1) return true;
            p_item += gridDim.x; advance = false;
        }
        return false;
    };
    WR nxt;
    bool have = seek(false);
    if (have) load_w(nxt, p.W, p_r0 + g, p_r0 + g + 8, p.nbr, p.nb_total, p_sb, t);

    pdl_wait();                                             // activations are the predecessor's output; weights are constants
    if (gate_closed(p.gate)) return;                        // idle draft level: the (unconsumed) first weight loads are all that was issued
    {
        const ActQ &a = p.a;
        const int n_qs = p.nbr * 8 * 16;                    // 16-byte chunks of quants: [sb][col][16]
        for (int i = tid; i < n_qs; i += SM_THREADS) {
            const int sb = i >> 7, col = (i >> 4) & 7, ch = i & 15;
            uint8_t *dst = smem + (size_t)sb * TL::BYTES + col * TL::PITCH + ch * 16;
            if (col < a.T) cp_async16(dst, a.qs + (int64_t)col * a.K + sb * 256 + ch * 16);
            else *(int4 *)dst = make_int4(0, 0, 0, 0);
        }
        if (FMT == 0) {
            for (int i = tid; i < p.nbr * 16; i += SM_THREADS) {            // bsums: 2 x 16 B per (sb, column)
                const int sb = i >> 4, col = (i >> 1) & 7, ch = i & 1;
                uint8_t *dst = smem + (size_t)sb * TL::BYTES + TL::BS_OFF + col * 32 + ch * 16;
                if (col < a.T) cp_async16(dst, (const uint8_t *)(a.bsum + ((int64_t)col * p.nbr + sb) * 16) + ch * 16);
                else *(int4 *)dst = make_int4(0, 0, 0, 0);
            }
        }
        for (int i = tid; i < p.nbr * 8; i += SM_THREADS) {
            const int sb = i >> 3, col = i & 7;
            ((float *)(smem + (size_t)sb * TL::BYTES + TL::D_OFF))[col] = col < a.T ? a.d[(int64_t)col * p.nbr + sb] : 0.f;
        }
        asm volatile("cp.async.commit_group;");
        asm volatile("cp.async.wait_group 0;");
        __syncthreads();
    }

    int par = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, par ^= 1) {
        const int ksp = item / p.n_rb, rb = item % p.n_rb;
        const int sb0 = ksp * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
        float facc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int sb = sb0 + kg; sb < sb1; sb += KG) {
            const WR cur = nxt;
            have = seek(true);
            if (have) load_w(nxt, p.W, p_r0 + g, p_r0 + g + 8, p.nbr, p.nb_total, p_sb, t);
            compute_sb(facc, cur, smem + (size_t)sb * TL::BYTES, g, t);
        }
        // sum the K groups in fixed order; the buffer alternates so that one barrier per item is enough
        float4 *rd = red + par * (KG * RG * 32);
        rd[(kg * RG + rg) * 32 + lane] = make_float4(facc[0], facc[1], facc[2], facc[3]);
        __syncthreads();
        if (kg == 0) {
            float4 s = rd[rg * 32 + lane];
#pragma unroll
            for (int k = 1; k < KG; k++) { const float4 v = rd[(k * RG + rg) * 32 + lane]; s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
            const int64_t row = (int64_t)rb * (16 * RG) + rg * 16 + g;
            if (row < p.N) {                                // N % 16 == 0: row + 8 is in range too
                float *o = p.out + (int64_t)ksp * p.a.T * p.N;
                const int c0 = 2 * t, T = p.a.T;
                if (c0 < T) { o[(int64_t)c0 * p.N + row] = s.x; o[(int64_t)c0 * p.N + row + 8] = s.z; }
                if (c0 + 1 < T) { o[(int64_t)(c0 + 1) * p.N + row] = s.y; o[(int64_t)(c0 + 1) * p.N + row + 8] = s.w; }
            }
        }
    }
}

constexpr int RED_BYTES = 2 * KG * RG * 32 * 16;
constexpr int SMALL_SMEM_MAX = 200 * 1024;
template <int FMT> size_t small_smem(int nbr) { return (size_t)nbr * Tile<FMT>::BYTES + RED_BYTES; }

}  // namespace

int eb_mmq_small_supported(int type, int64_t N, int64_t K, int T) {
    static const bool off = getenv("EB_NO_MMQ_SMALL") != nullptr;
    if (off || (type != EB_Q4_K && type != EB_Q6_K) || T < 1 || T > 8 || N % 16 || K % 256) return 0;
    const int nbr = (int)(K / 256);
    return (type == EB_Q4_K ? small_smem<0>(nbr) : small_smem<1>(nbr)) <= (size_t)SMALL_SMEM_MAX;
}

// planes [ks][T][N] (ks returned; 1 for every shape of the 7B head)
int eb_mmq_small_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st) {
    if (!eb_mmq_small_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq_small: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    if (!a.kq) return eb_set_err(EB_ERR_ARG, "mmq_small: activations are not on the Q8_K grid");
    const int nbr = (int)(K / 256), G = eb_sm_count();
    const int n_rb = (int)cdiv(N, 16 * RG);
    // K split: rounds of the persistent grid x (serial super-blocks per warp + fixed cost per item)
    int best = 1; double best_cost = 1e30;
    for (int ks = 1; ks <= 8 && ks <= nbr; ks *= 2) {
        const int spp = (int)cdiv(nbr, ks);
        if ((int)cdiv(nbr, spp) != ks) continue;
        const double cost = (double)cdiv((int64_t)n_rb * ks, G) * (cdiv(spp, KG) + 2.0) + 0.1 * ks;
        if (cost < best_cost) { best_cost = cost; best = ks; }
    }
    SmallArgs p;
    p.W = W; p.N = N; p.nb_total = N * nbr; p.nbr = nbr; p.n_rb = n_rb; p.spp = (int)cdiv(nbr, best); p.n_items = n_rb * best; p.a = a; p.out = planes; p.gate = eb_gate();
    *ks_out = best;
    const int grid = std::min(p.n_items, G);
    if (type == EB_Q4_K) eb_launch_pdl_c(EB_PDL_GEMM, mmq_small_kernel<0>, dim3(grid), dim3(SM_THREADS), small_smem<0>(nbr), st, p);
    else eb_launch_pdl_c(EB_PDL_GEMM, mmq_small_kernel<1>, dim3(grid), dim3(SM_THREADS), small_smem<1>(nbr), st, p);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

int eb_mmq_small_init() {
    EB_CUDA(cudaFuncSetAttribute(mmq_small_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_MAX));
    EB_CUDA(cudaFuncSetAttribute(mmq_small_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMALL_SMEM_MAX));
    return EB_OK;
}
