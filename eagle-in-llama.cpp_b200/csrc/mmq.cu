// mmq.cu — K2: tree-batched (T = 9..128) verify GEMM on quantised weights, int8 tensor-core MMA.
// Replaces mul_mat_q (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822).  Weights are streamed from HBM
// exactly once; the activation tile (int8, CPU-backend quantisation grid) is staged through
// shared memory with cp.async and re-used by all warps of the CTA.  Integer sub-block sums are
// exact, so the result differs from ggml_vec_dot_q4_K_q8_K / q6_K_q8_K
// (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:6266,7903) only by f32 summation order.
//
// Tiling: CTA = 4 warps x 16 weight rows; grid = (N/64, ksplit); each CTA walks its super-blocks.
// Per super-block (K = 256):   Q4_K: 8 x mma.m16n8k32.u8.s8 per 8-column tile + 1 f16 MMA for the
// min term (m_j x bsum_j, exact in f16/f32);   Q6_K: 16 x mma.m16n8k16.s8.s8 per tile.
#include "common.cuh"
#include <cstdlib>

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

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

// shared-memory stage: per column  [qs: PITCH bytes][d: 4][bsum: 32]
template <int NT, int PITCH> struct ActTile {
    static constexpr int COLS = NT * 8;
    static constexpr int QS_BYTES = COLS * PITCH;
    static constexpr int D_OFF = QS_BYTES;                 // float [COLS]
    static constexpr int BS_OFF = D_OFF + COLS * 4;        // half [COLS][16]
    static constexpr int BYTES = (BS_OFF + COLS * 32 + 127) & ~127;
};

// all 128 threads cooperatively stage the activations of super-block sb
template <int NT, int PITCH>
__device__ __forceinline__ void load_act_tile(uint8_t *st, const ActQ &a, int nbr, int sb, int tid) {
    using AT = ActTile<NT, PITCH>;
    for (int i = tid; i < AT::COLS * 16; i += 128) {       // 16-byte chunks of qs
        const int col = i >> 4, ch = i & 15;
        uint8_t *dst = st + col * PITCH + ch * 16;
        if (col < a.T) cp_async16(dst, a.qs + (int64_t)col * a.K + sb * 256 + ch * 16);
        else *(int4 *)dst = make_int4(0, 0, 0, 0);
    }
    for (int i = tid; i < AT::COLS * 2; i += 128) {        // bsums: 2 x 16 B per column
        const int col = i >> 1, ch = i & 1;
        uint8_t *dst = st + AT::BS_OFF + col * 32 + ch * 16;
        if (col < a.T) cp_async16(dst, (const uint8_t *)(a.bsum + ((int64_t)col * nbr + sb) * 16) + ch * 16);
        else *(int4 *)dst = make_int4(0, 0, 0, 0);
    }
    for (int col = tid; col < AT::COLS; col += 128)
        ((float *)(st + AT::D_OFF))[col] = col < a.T ? a.d[(int64_t)col * nbr + sb] : 0.f;
}

__device__ __forceinline__ void store_out(float *out, int64_t N, int T, int64_t row, int col, float v0, float v1) {
    if (col < T) out[(int64_t)col * N + row] = v0;
    if (col + 1 < T) out[(int64_t)(col + 1) * N + row] = v1;
}

// ------------------------------------------------------------------ Q4_K
template <int NT>
__global__ void __launch_bounds__(128) mmq_q4_K(const block_q4_K *__restrict__ W, int64_t N, int nbr, ActQ a,
                                                float *__restrict__ out, int sb_per_split) {
    constexpr int PITCH = 288;                              // 256 + 32: conflict-free 8-byte B-fragment loads
    using AT = ActTile<NT, PITCH>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int64_t row0 = (int64_t)blockIdx.x * 64 + warp * 16;
    const bool active = row0 < N;                           // N % 16 == 0
    const int sb_begin = blockIdx.y * sb_per_split, sb_end = min(nbr, sb_begin + sb_per_split);
    out += (int64_t)blockIdx.y * a.T * N;                   // split-K partial plane

    const block_q4_K *w0 = W + (active ? row0 + g : 0) * nbr, *w1 = W + (active ? row0 + g + 8 : 0) * nbr;

    float facc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) { facc[nt][0] = facc[nt][1] = facc[nt][2] = facc[nt][3] = 0.f; }

    if (sb_begin < sb_end) { load_act_tile<NT, PITCH>(smem, a, nbr, sb_begin, tid); }
    cp_async_commit();

    for (int sb = sb_begin; sb < sb_end; sb++) {
        const int buf = (sb - sb_begin) & 1;
        uint8_t *st = smem + buf * AT::BYTES;
        // weights of this super-block: header + 4 x 8 bytes per row, two rows per lane
        int4 h0 = __ldg((const int4 *)&w0[sb]), h1 = __ldg((const int4 *)&w1[sb]);
        uint2 q0[4], q1[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            q0[j] = __ldg((const uint2 *)(w0[sb].qs + j * 32 + 8 * t));
            q1[j] = __ldg((const uint2 *)(w1[sb].qs + j * 32 + 8 * t));
        }
        if (sb + 1 < sb_end) load_act_tile<NT, PITCH>(smem + (buf ^ 1) * AT::BYTES, a, nbr, sb + 1, tid);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();

        // unpack the 6-bit scales/mins of both rows (kmask trick of ggml_vec_dot_q4_K_q8_K)
        uint32_t u0[4], u1[4];
        {
            const uint32_t k1 = 0x3f3f3f3f, k2 = 0x0f0f0f0f, k3 = 0x03030303;
            u0[0] = h0.y & k1; u0[1] = (h0.w & k2) | (((h0.y >> 6) & k3) << 4); u0[2] = h0.z & k1; u0[3] = ((h0.w >> 4) & k2) | (((h0.z >> 6) & k3) << 4);
            u1[0] = h1.y & k1; u1[1] = (h1.w & k2) | (((h1.y >> 6) & k3) << 4); u1[2] = h1.z & k1; u1[3] = ((h1.w >> 4) & k2) | (((h1.z >> 6) & k3) << 4);
        }   // bytes of u[0..1] = scales 0..7, bytes of u[2..3] = mins 0..7
        int acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) { acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0; }
#pragma unroll
        for (int j = 0; j < 4; j++) {
#pragma unroll
            for (int hi = 0; hi < 2; hi++) {
                const int s = 2 * j + hi;
                uint32_t af[4];
                af[0] = (q0[j].x >> (4 * hi)) & 0x0F0F0F0F; af[1] = (q1[j].x >> (4 * hi)) & 0x0F0F0F0F;
                af[2] = (q0[j].y >> (4 * hi)) & 0x0F0F0F0F; af[3] = (q1[j].y >> (4 * hi)) & 0x0F0F0F0F;
                const int sc0 = (u0[s >> 2] >> (8 * (s & 3))) & 0xff, sc1 = (u1[s >> 2] >> (8 * (s & 3))) & 0xff;
#pragma unroll
                for (int nt = 0; nt < NT; nt++) {
                    const uint2 b = *(const uint2 *)(st + (nt * 8 + g) * PITCH + s * 32 + 8 * t);
                    int c[4] = {0, 0, 0, 0};
                    mma_u8s8_k32(c, af, b.x, b.y);
                    acc[nt][0] += sc0 * c[0]; acc[nt][1] += sc0 * c[1];
                    acc[nt][2] += sc1 * c[2]; acc[nt][3] += sc1 * c[3];
                }
            }
        }
        // min term: sum_j m_j * bsum32_j  ==  [m_(k/2)] (16) x bsum16 (16) as one f16 MMA, exact
        uint32_t am[4];
        {
            const int m0a = (u0[2] >> (8 * t)) & 0xff, m0b = (u0[3] >> (8 * t)) & 0xff;
            const int m1a = (u1[2] >> (8 * t)) & 0xff, m1b = (u1[3] >> (8 * t)) & 0xff;
            const __half2 x0 = __half2half2(__int2half_rn(m0a)), x1 = __half2half2(__int2half_rn(m1a));
            const __half2 x2 = __half2half2(__int2half_rn(m0b)), x3 = __half2half2(__int2half_rn(m1b));
            am[0] = *(const uint32_t *)&x0; am[1] = *(const uint32_t *)&x1; am[2] = *(const uint32_t *)&x2; am[3] = *(const uint32_t *)&x3;
        }
        const __half2 dm0 = *(const __half2 *)&h0.x, dm1 = *(const __half2 *)&h1.x;
        const float d0 = __low2float(dm0), mn0 = __high2float(dm0), d1 = __low2float(dm1), mn1 = __high2float(dm1);
        const float *dx = (const float *)(st + AT::D_OFF);
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            const uint32_t *bs = (const uint32_t *)(st + AT::BS_OFF + (nt * 8 + g) * 32);
            float cm[4] = {0.f, 0.f, 0.f, 0.f};
            mma_f16_k16(cm, am, bs[t], bs[t + 4]);
            const float x0 = dx[nt * 8 + 2 * t], x1 = dx[nt * 8 + 2 * t + 1];
            facc[nt][0] += (d0 * x0) * (float)acc[nt][0] - (mn0 * x0) * cm[0];
            facc[nt][1] += (d0 * x1) * (float)acc[nt][1] - (mn0 * x1) * cm[1];
            facc[nt][2] += (d1 * x0) * (float)acc[nt][2] - (mn1 * x0) * cm[2];
            facc[nt][3] += (d1 * x1) * (float)acc[nt][3] - (mn1 * x1) * cm[3];
        }
        __syncthreads();                                    // stage `buf` may be overwritten next iteration
    }
    if (active) {
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            store_out(out, N, a.T, row0 + g, nt * 8 + 2 * t, facc[nt][0], facc[nt][1]);
            store_out(out, N, a.T, row0 + g + 8, nt * 8 + 2 * t, facc[nt][2], facc[nt][3]);
        }
    }
}

// ------------------------------------------------------------------ Q6_K
template <int NT>
__global__ void __launch_bounds__(128) mmq_q6_K(WPlanes w, int64_t N, int nbr, ActQ a, float *__restrict__ out, int sb_per_split) {
    constexpr int PITCH = 272;                              // 256 + 16: conflict-free 4-byte B-fragment loads
    using AT = ActTile<NT, PITCH>;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int64_t row0 = (int64_t)blockIdx.x * 64 + warp * 16;
    const bool active = row0 < N;
    const int sb_begin = blockIdx.y * sb_per_split, sb_end = min(nbr, sb_begin + sb_per_split);
    out += (int64_t)blockIdx.y * a.T * N;
    const int64_t r0 = active ? row0 + g : 0, r1 = active ? row0 + g + 8 : 0;

    float facc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) { facc[nt][0] = facc[nt][1] = facc[nt][2] = facc[nt][3] = 0.f; }
    if (sb_begin < sb_end) load_act_tile<NT, PITCH>(smem, a, nbr, sb_begin, tid);
    cp_async_commit();

    for (int sb = sb_begin; sb < sb_end; sb++) {
        const int buf = (sb - sb_begin) & 1;
        uint8_t *st = smem + buf * AT::BYTES;
        const int64_t b0 = r0 * nbr + sb, b1 = r1 * nbr + sb;
        // ql: [n][c&1][u] words at n*64 + 32*(c&1) + 16u + 4t ; qh: [n][u] words at n*32 + 16u + 4t
        uint32_t ql0[8], ql1[8], qh0[4], qh1[4];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int off = (i >> 2) * 64 + ((i >> 1) & 1) * 32 + (i & 1) * 16 + 4 * t;
            ql0[i] = __ldg((const uint32_t *)(w.q6k_ql() + b0 * 128 + off));
            ql1[i] = __ldg((const uint32_t *)(w.q6k_ql() + b1 * 128 + off));
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int off = (i >> 1) * 32 + (i & 1) * 16 + 4 * t;
            qh0[i] = __ldg((const uint32_t *)(w.q6k_qh() + b0 * 64 + off));
            qh1[i] = __ldg((const uint32_t *)(w.q6k_qh() + b1 * 64 + off));
        }
        const int4 s0 = __ldg((const int4 *)(w.q6k_sc() + b0 * 16)), s1 = __ldg((const int4 *)(w.q6k_sc() + b1 * 16));
        const float d0 = __half2float(w.q6k_d()[b0]), d1 = __half2float(w.q6k_d()[b1]);
        if (sb + 1 < sb_end) load_act_tile<NT, PITCH>(smem + (buf ^ 1) * AT::BYTES, a, nbr, sb + 1, tid);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();

        int acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) { acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0; }
        const int sc0w[4] = {s0.x, s0.y, s0.z, s0.w}, sc1w[4] = {s1.x, s1.y, s1.z, s1.w};
#pragma unroll
        for (int s = 0; s < 16; s++) {                      // sub-block s: n = s/8, c = (s%8)/2, u = s%2
            const int n = s >> 3, c = (s & 7) >> 1, u = s & 1;
            const int qi = n * 4 + (c & 1) * 2 + u, hi = n * 2 + u;
            const uint32_t a0 = __vsubss4(((ql0[qi] >> (4 * (c >> 1))) & 0x0F0F0F0Fu) | (((qh0[hi] >> (2 * c)) & 0x03030303u) << 4), 0x20202020u);
            const uint32_t a1 = __vsubss4(((ql1[qi] >> (4 * (c >> 1))) & 0x0F0F0F0Fu) | (((qh1[hi] >> (2 * c)) & 0x03030303u) << 4), 0x20202020u);
            const int sc0 = (int)(int8_t)((sc0w[s >> 2] >> (8 * (s & 3))) & 0xff), sc1 = (int)(int8_t)((sc1w[s >> 2] >> (8 * (s & 3))) & 0xff);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                const uint32_t b = *(const uint32_t *)(st + (nt * 8 + g) * PITCH + s * 16 + 4 * t);
                int cc[4] = {0, 0, 0, 0};
                mma_s8s8_k16(cc, a0, a1, b);
                acc[nt][0] += sc0 * cc[0]; acc[nt][1] += sc0 * cc[1];
                acc[nt][2] += sc1 * cc[2]; acc[nt][3] += sc1 * cc[3];
            }
        }
        const float *dx = (const float *)(st + AT::D_OFF);
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            const float x0 = dx[nt * 8 + 2 * t], x1 = dx[nt * 8 + 2 * t + 1];
            facc[nt][0] += (d0 * x0) * (float)acc[nt][0]; facc[nt][1] += (d0 * x1) * (float)acc[nt][1];
            facc[nt][2] += (d1 * x0) * (float)acc[nt][2]; facc[nt][3] += (d1 * x1) * (float)acc[nt][3];
        }
        __syncthreads();
    }
    if (active) {
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            store_out(out, N, a.T, row0 + g, nt * 8 + 2 * t, facc[nt][0], facc[nt][1]);
            store_out(out, N, a.T, row0 + g + 8, nt * 8 + 2 * t, facc[nt][2], facc[nt][3]);
        }
    }
}

// y[i] = sum_s partial[s][i]   (fixed order: deterministic)
__global__ void splitk_reduce_kernel(const float *__restrict__ partial, float *__restrict__ y, int64_t n, int ksplit) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    float4 s = *(const float4 *)(partial + i);
    for (int k = 1; k < ksplit; k++) { const float4 p = *(const float4 *)(partial + (int64_t)k * n + i); s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w; }
    *(float4 *)(y + i) = s;
}

// the same sum, one element per thread: n % 4 != 0 (ragged operator-layer shapes) leaves the planes unaligned for float4
__global__ void splitk_reduce_scalar_kernel(const float *__restrict__ partial, float *__restrict__ y, int64_t n, int ksplit) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = partial[i];
    for (int k = 1; k < ksplit; k++) s += partial[(int64_t)k * n + i];
    y[i] = s;
}

int eb_splitk_reduce(const float *partial, float *y, int64_t n, int ksplit, cudaStream_t st) {
    if (n % 4) {
        splitk_reduce_scalar_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(partial, y, n, ksplit);
        EB_CHECK_LAUNCH();
        return EB_OK;
    }
    splitk_reduce_kernel<<<(unsigned)cdiv(n / 4, 256), 256, 0, st>>>(partial, y, n, ksplit);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

int eb_mmq_supported(int type, int64_t N, int64_t K, int T) {
    return (type == EB_Q4_K || type == EB_Q6_K) && N % 16 == 0 && K % 256 == 0 && T > 1 && T <= 128;
}

int eb_mmq_pick_ksplit(int type, int64_t N, int64_t K, int T) {
    (void)type;
    const int nbr = (int)(K / 256);
    const int64_t row_blocks = cdiv(N, 64);
    // CTAs per SM that stream concurrently: 3 fit when the activation tile is 9 x 8 columns; with one 8-column tile (T <= 8,
    // draft head) 16 fit, and the kernel needs them all in flight to cover the HBM latency (each CTA walks its
    // super-blocks serially)
    static const int small_mult = getenv("EB_MMQ_SMALL_MULT") ? atoi(getenv("EB_MMQ_SMALL_MULT")) : 8;
    const int target = eb_sm_count() * (T <= 8 ? small_mult : 3);
    int ks = (int)cdiv(target, row_blocks);
    if (ks > nbr / 2) ks = nbr / 2;          // at least 2 super-blocks per CTA
    if (ks > 8) ks = 8;
    if (ks < 1) ks = 1;
    return ks;
}

template <int NT>
static int launch(int type, const void *W, int64_t N, int nbr, const ActQ &a, float *out, int ksplit, cudaStream_t st) {
    const int spp = (int)cdiv(nbr, ksplit);
    const dim3 grid((unsigned)cdiv(N, 64), (unsigned)ksplit);
    if (type == EB_Q4_K) {
        constexpr int smem = 2 * ActTile<NT, 288>::BYTES;
        mmq_q4_K<NT><<<grid, 128, smem, st>>>((const block_q4_K *)W, N, nbr, a, out, spp);
    } else {
        constexpr int smem = 2 * ActTile<NT, 272>::BYTES;
        WPlanes w{(const uint8_t *)W, N * nbr};
        mmq_q6_K<NT><<<grid, 128, smem, st>>>(w, N, nbr, a, out, spp);
    }
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// engine form: leave the split-K planes unreduced in `planes` ([ks][T][N]); the consumer kernel sums them
int eb_mmq_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st) {
    if (!eb_mmq_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    const int nbr = (int)(K / 256);
    int ksplit = eb_mmq_pick_ksplit(type, N, K, a.T);
    if (ksplit > nbr) ksplit = nbr;
    ksplit = (int)cdiv(nbr, cdiv(nbr, ksplit));
    *ks_out = ksplit;
    const int nt = (a.T + 7) / 8;
    if (nt <= 1) return launch<1>(type, W, N, nbr, a, planes, ksplit, st);
    if (nt <= 2) return launch<2>(type, W, N, nbr, a, planes, ksplit, st);
    if (nt <= 4) return launch<4>(type, W, N, nbr, a, planes, ksplit, st);
    if (nt <= 6) return launch<6>(type, W, N, nbr, a, planes, ksplit, st);
    if (nt <= 9) return launch<9>(type, W, N, nbr, a, planes, ksplit, st);
    if (nt <= 12) return launch<12>(type, W, N, nbr, a, planes, ksplit, st);
    return launch<16>(type, W, N, nbr, a, planes, ksplit, st);
}

// y [T][N]; partial: scratch of ksplit*T*N floats (only used when ksplit > 1)
int eb_mmq(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *y, float *partial, int ksplit, cudaStream_t st) {
    if (!eb_mmq_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    const int nbr = (int)(K / 256);
    if (ksplit < 1) ksplit = 1;
    if (ksplit > nbr) ksplit = nbr;
    ksplit = (int)cdiv(nbr, cdiv(nbr, ksplit));   // no empty splits
    float *out = ksplit > 1 ? partial : y;
    const int nt = (a.T + 7) / 8;
    int rc;
    if (nt <= 1) rc = launch<1>(type, W, N, nbr, a, out, ksplit, st);
    else if (nt <= 2) rc = launch<2>(type, W, N, nbr, a, out, ksplit, st);
    else if (nt <= 4) rc = launch<4>(type, W, N, nbr, a, out, ksplit, st);
    else if (nt <= 6) rc = launch<6>(type, W, N, nbr, a, out, ksplit, st);
    else if (nt <= 9) rc = launch<9>(type, W, N, nbr, a, out, ksplit, st);
    else if (nt <= 12) rc = launch<12>(type, W, N, nbr, a, out, ksplit, st);
    else rc = launch<16>(type, W, N, nbr, a, out, ksplit, st);
    if (rc) return rc;
    if (ksplit > 1) {
        const int64_t n = (int64_t)a.T * N;      // N % 16 == 0 so n % 4 == 0
        splitk_reduce_kernel<<<(unsigned)cdiv(n / 4, 256), 256, 0, st>>>(partial, y, n, ksplit);
        EB_CHECK_LAUNCH();
    }
    return EB_OK;
}

template <int NT> static int init_nt() {
    EB_CUDA(cudaFuncSetAttribute(mmq_q4_K<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * ActTile<NT, 288>::BYTES));
    EB_CUDA(cudaFuncSetAttribute(mmq_q6_K<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * ActTile<NT, 272>::BYTES));
    return EB_OK;
}
int eb_mmq_init() {   // dynamic-smem opt-in for every instantiation, once per process (never inside a graph capture)
    int rc;
    if ((rc = init_nt<1>()) || (rc = init_nt<2>()) || (rc = init_nt<4>()) || (rc = init_nt<6>()) || (rc = init_nt<9>()) || (rc = init_nt<12>()) || (rc = init_nt<16>())) return rc;
    return EB_OK;
}
