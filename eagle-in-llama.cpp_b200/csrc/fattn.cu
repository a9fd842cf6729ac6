// fattn.cu — K3: tree-masked flash attention for the verify batch (T <= 128 queries, D = 128,
// f16 KV cache, split-KV across CTAs).  Replaces flash_attn_ext_f16
// (R/ggml/src/ggml-cuda/fattn-mma-f16.cuh:806, launch_fattn fattn-common.cuh:676) and restates
// ggml_compute_forward_flash_attn_ext_f16 (R/ggml/src/ggml-cpu/ggml-cpu.c:10999-11212): Q is
// rounded to f16, S = K.Q accumulated in f32, masked cells are skipped, online softmax in f32.
// (The CPU keeps the V accumulator in f16; this kernel keeps it in f32 and rounds P to f16 for
// the P.V MMA — same order of error, checked against the oracle with the op tolerance.)
//
// One CTA = one head x one KV split x all T query rows (one warp per 16 rows), so every K/V byte
// is read from HBM once per layer.  The tree mask is either the ggml f16 mask tensor or the
// compact form (committed prefix + 128-bit ancestor set per row) kept in registers.
#include "common.cuh"
#include "spec.cuh"
#include "quant_dev.cuh"

#define FA_D 128
#define FA_TILE 64

__device__ __forceinline__ void fa_cp16(uint32_t saddr, const void *g) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g)); }
__device__ __forceinline__ void fa_zero16(uint32_t saddr) { asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(saddr), "r"(0)); }
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], uint32_t saddr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ void fa_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// tile row = one cell = 256 B = 16 chunks of 16 B; chunk index XOR-swizzled with the cell index
__device__ __forceinline__ uint32_t fa_off(int cell, int chunk) { return (uint32_t)(cell * 256 + (((chunk ^ cell) & 7) | (chunk & 8)) * 16); }

struct FaParams {
    const float *q; const __half *k, *v; int64_t kv_stride;
    const __half *mask; int64_t mask_stride;           // mode 1
    const uint32_t *tree_bits; int n_prefix;            // mode 0
    const FwdCtl *ctl; int n_ctx;                       // mode 0, device-resident lengths (engine)
    int T, H, n_kv, n_tiles, tiles_per_split, nsplit;
    float scale_log2;                                   // scale * log2(e)
    float *out; float *part_o; float *part_ml;          // partials when nsplit > 1
};

template <int MODE>
__global__ void __launch_bounds__(256) fattn_kernel(FaParams p) {
    extern __shared__ __align__(128) uint8_t smem[];    // [2 stages][K 16 KB | V 16 KB]
    const int tid = threadIdx.x, nthr = blockDim.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int h = blockIdx.x, split = blockIdx.y;
    if (MODE == 0 && p.ctl) {                           // lengths live on the device: no host round trip per step
        p.n_prefix = p.ctl->n_prefix;
        p.n_kv = min(p.n_ctx, p.n_prefix + p.ctl->n_tree);
        p.n_tiles = (p.n_kv + FA_TILE - 1) / FA_TILE;
        p.tiles_per_split = (p.n_tiles + p.nsplit - 1) / p.nsplit;
    }
    const int tile0 = split * p.tiles_per_split, tile1 = min(p.n_tiles, tile0 + p.tiles_per_split);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
    const __half *kh = p.k + (int64_t)h * FA_D, *vh = p.v + (int64_t)h * FA_D;

    auto load_tile = [&](int tile, int stage) {
        const uint32_t sk = sbase + stage * 32768, sv = sk + 16384;
        for (int i = tid; i < FA_TILE * 16; i += nthr) {
            const int cell = i >> 4, chunk = i & 15, c = tile * FA_TILE + cell;
            const uint32_t off = fa_off(cell, chunk);
            if (c < p.n_kv) {
                fa_cp16(sk + off, kh + (int64_t)c * p.kv_stride + chunk * 8);
                fa_cp16(sv + off, vh + (int64_t)c * p.kv_stride + chunk * 8);
            } else { fa_zero16(sk + off); fa_zero16(sv + off); }
        }
    };
    if (tile0 < tile1) load_tile(tile0, 0);
    asm volatile("cp.async.commit_group;");

    // Q fragments (f32 -> f16, GGML_FP32_TO_FP16 of ggml-cpu.c:11115), rows r0 = 16*warp + g, r1 = r0 + 8
    const int r0 = warp * 16 + g, r1 = r0 + 8;
    uint32_t qa[8][4];
#pragma unroll
    for (int ks = 0; ks < 8; ks++) {
        const int d0 = ks * 16 + 2 * t;
        float2 a = make_float2(0.f, 0.f), b = a, c = a, d = a;
        if (r0 < p.T) { const float *qr = p.q + ((int64_t)r0 * p.H + h) * FA_D; a = *(const float2 *)(qr + d0); c = *(const float2 *)(qr + d0 + 8); }
        if (r1 < p.T) { const float *qr = p.q + ((int64_t)r1 * p.H + h) * FA_D; b = *(const float2 *)(qr + d0); d = *(const float2 *)(qr + d0 + 8); }
        const __half2 ha = __floats2half2_rn(a.x, a.y), hb = __floats2half2_rn(b.x, b.y), hc = __floats2half2_rn(c.x, c.y), hd = __floats2half2_rn(d.x, d.y);
        qa[ks][0] = *(const uint32_t *)&ha; qa[ks][1] = *(const uint32_t *)&hb; qa[ks][2] = *(const uint32_t *)&hc; qa[ks][3] = *(const uint32_t *)&hd;
    }
    uint32_t tb0[4] = {0, 0, 0, 0}, tb1[4] = {0, 0, 0, 0};
    if (MODE == 0) {
        if (r0 < p.T) { const uint4 w = *(const uint4 *)(p.tree_bits + r0 * 4); tb0[0] = w.x; tb0[1] = w.y; tb0[2] = w.z; tb0[3] = w.w; }
        if (r1 < p.T) { const uint4 w = *(const uint4 *)(p.tree_bits + r1 * 4); tb1[0] = w.x; tb1[1] = w.y; tb1[2] = w.z; tb1[3] = w.w; }
    }
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; i++) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }

    for (int tile = tile0; tile < tile1; tile++) {
        const int stage = (tile - tile0) & 1;
        if (tile + 1 < tile1) load_tile(tile + 1, stage ^ 1);
        asm volatile("cp.async.commit_group;");
        asm volatile("cp.async.wait_group 1;");
        __syncthreads();
        const uint32_t sk = sbase + stage * 32768, sv = sk + 16384;
        if (warp * 16 < p.T) {      // warps beyond the last query tile only help with the loads

        // S = Q K^T  (16 rows x 64 cells)
        float s[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; nt++) { s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < 8; ks += 2) {
#pragma unroll
            for (int nt = 0; nt < 8; nt++) {
                uint32_t b[4];
                const int cell = nt * 8 + (lane & 7), chunk = 2 * ks + (lane >> 3);
                ldsm4(b, sk + fa_off(cell, chunk));
                fa_mma(s[nt], qa[ks], b[0], b[1]);
                fa_mma(s[nt], qa[ks + 1], b[2], b[3]);
            }
        }
        // scale + mask
        const int cbase = tile * FA_TILE;
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 8; nt++) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int c = cbase + nt * 8 + 2 * t + (e & 1);
                const int r = (e < 2) ? r0 : r1;
                float val = s[nt][e] * p.scale_log2;
                bool vis = c < p.n_kv && r < p.T;
                if (MODE == 0) {
                    if (vis && c >= p.n_prefix) {
                        const int j = c - p.n_prefix;
                        const uint32_t *tb = (e < 2) ? tb0 : tb1;
                        vis = j < 128 && ((tb[j >> 5] >> (j & 31)) & 1u);
                    }
                } else if (vis) {
                    const float mv = __half2float(p.mask[(int64_t)r * p.mask_stride + c]);
                    if (mv == -INFINITY) vis = false; else val += mv * 1.44269504f;
                }
                val = vis ? val : -INFINITY;
                s[nt][e] = val;
                if (e < 2) mx0 = fmaxf(mx0, val); else mx1 = fmaxf(mx1, val);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float base0 = mn0 == -INFINITY ? 0.f : mn0, base1 = mn1 == -INFINITY ? 0.f : mn1;
        const float corr0 = exp2f(m0 - base0), corr1 = exp2f(m1 - base1);   // m = -inf -> 0
        m0 = mn0; m1 = mn1;
        float rs0 = 0.f, rs1 = 0.f;
        // P is split into an f16 head and an f16 residual (two MMAs per tile) so that P.V carries
        // ~22 bits of P: activation quantisation downstream amplifies 1e-4 errors into logit noise
        uint32_t pa[4][4], pl[4][4];
#pragma unroll
        for (int nt = 0; nt < 8; nt++) {
            const float p0 = exp2f(s[nt][0] - base0), p1 = exp2f(s[nt][1] - base0);
            const float p2 = exp2f(s[nt][2] - base1), p3 = exp2f(s[nt][3] - base1);
            rs0 += p0 + p1; rs1 += p2 + p3;
            const __half2 h01 = __floats2half2_rn(p0, p1), h23 = __floats2half2_rn(p2, p3);
            const __half2 l01 = __floats2half2_rn(p0 - __low2float(h01), p1 - __high2float(h01));
            const __half2 l23 = __floats2half2_rn(p2 - __low2float(h23), p3 - __high2float(h23));
            pa[nt >> 1][(nt & 1) * 2 + 0] = *(const uint32_t *)&h01;
            pa[nt >> 1][(nt & 1) * 2 + 1] = *(const uint32_t *)&h23;
            pl[nt >> 1][(nt & 1) * 2 + 0] = *(const uint32_t *)&l01;
            pl[nt >> 1][(nt & 1) * 2 + 1] = *(const uint32_t *)&l23;
        }
        l0 = l0 * corr0 + rs0; l1 = l1 * corr1 + rs1;      // per-lane partial sums (reduced over t at the end)
#pragma unroll
        for (int nd = 0; nd < 16; nd++) { o[nd][0] *= corr0; o[nd][1] *= corr0; o[nd][2] *= corr1; o[nd][3] *= corr1; }
        // O += P V   (k = 64 cells in 4 steps of 16, n = 128 dims in 16 tiles of 8)
#pragma unroll
        for (int kk = 0; kk < 4; kk++) {
#pragma unroll
            for (int nd = 0; nd < 16; nd += 2) {
                uint32_t b[4];
                const int m = lane >> 3, cell = kk * 16 + (m & 1) * 8 + (lane & 7), chunk = nd + (m >> 1);
                ldsm4t(b, sv + fa_off(cell, chunk));
                fa_mma(o[nd], pa[kk], b[0], b[1]);
                fa_mma(o[nd + 1], pa[kk], b[2], b[3]);
                fa_mma(o[nd], pl[kk], b[0], b[1]);
                fa_mma(o[nd + 1], pl[kk], b[2], b[3]);
            }
        }
        }
        __syncthreads();
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
    float *dst0, *dst1;
    if (p.nsplit == 1) {
        dst0 = p.out + ((int64_t)r0 * p.H + h) * FA_D; dst1 = p.out + ((int64_t)r1 * p.H + h) * FA_D;
    } else {
        const int64_t pb = ((int64_t)h * p.nsplit + split) * p.T;
        dst0 = p.part_o + (pb + r0) * FA_D; dst1 = p.part_o + (pb + r1) * FA_D;
        if (t == 0) {
            if (r0 < p.T) { p.part_ml[(pb + r0) * 2] = m0; p.part_ml[(pb + r0) * 2 + 1] = l0; }
            if (r1 < p.T) { p.part_ml[(pb + r1) * 2] = m1; p.part_ml[(pb + r1) * 2 + 1] = l1; }
        }
    }
#pragma unroll
    for (int nd = 0; nd < 16; nd++) {
        if (r0 < p.T) *(float2 *)(dst0 + nd * 8 + 2 * t) = make_float2(o[nd][0] * inv0, o[nd][1] * inv0);
        if (r1 < p.T) *(float2 *)(dst1 + nd * 8 + 2 * t) = make_float2(o[nd][2] * inv1, o[nd][3] * inv1);
    }
}

// merge the KV splits: one warp per (row, head)
__global__ void fattn_combine_kernel(const float *__restrict__ part_o, const float *__restrict__ part_ml, float *__restrict__ out,
                                     int T, int H, int nsplit) {
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= T * H) return;
    const int r = w / H, h = w % H;
    float M = -INFINITY;
    for (int s = 0; s < nsplit; s++) M = fmaxf(M, part_ml[(((int64_t)h * nsplit + s) * T + r) * 2]);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f); float L = 0.f;
    for (int s = 0; s < nsplit; s++) {
        const int64_t pb = ((int64_t)h * nsplit + s) * T + r;
        const float m = part_ml[pb * 2], l = part_ml[pb * 2 + 1];
        if (l <= 0.f) continue;
        const float wgt = l * exp2f(m - M);
        const float4 v = *(const float4 *)(part_o + pb * FA_D + lane * 4);
        acc.x += wgt * v.x; acc.y += wgt * v.y; acc.z += wgt * v.z; acc.w += wgt * v.w; L += wgt;
    }
    const float inv = L > 0.f ? 1.f / L : 0.f;
    *(float4 *)(out + ((int64_t)r * H + h) * FA_D + lane * 4) = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
}

// merge the KV splits and quantise the result for the wo GEMM (Q8_K grid): one warp per (row, 256-value block = 2 heads),
// lane = (head-in-pair, 8 dims); the f32 attention output never goes to memory
__global__ void fattn_combine_quant_kernel(const float *__restrict__ part_o, const float *__restrict__ part_ml, int T, int H, int nsplit, ActQ a, const int *__restrict__ gate) {
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31, nb = H / 2;
    if (w >= T * nb) return;
    const int r = w / nb, b = w % nb, h = 2 * b + (lane >> 4), d0 = (lane & 15) * 8;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}; float L = 0.f;
    if (nsplit <= 4) {
        // every load of the (up to four) splits in flight at once: the rolled loops below cost two L2 round trips per split
        float2 ml[4]; float4 x[4], y[4];
#pragma unroll
        for (int s = 0; s < 4; s++) {
            ml[s] = make_float2(-INFINITY, 0.f);
            if (s < nsplit) {
                const int64_t pb = ((int64_t)h * nsplit + s) * T + r;
                ml[s] = *(const float2 *)(part_ml + pb * 2);
                x[s] = *(const float4 *)(part_o + pb * FA_D + d0); y[s] = *(const float4 *)(part_o + pb * FA_D + d0 + 4);
            }
        }
        float M = -INFINITY;
#pragma unroll
        for (int s = 0; s < 4; s++) if (s < nsplit) M = fmaxf(M, ml[s].x);
#pragma unroll
        for (int s = 0; s < 4; s++) {
            if (s < nsplit && ml[s].y > 0.f) {
                const float wgt = ml[s].y * exp2f(ml[s].x - M);
                acc[0] += wgt * x[s].x; acc[1] += wgt * x[s].y; acc[2] += wgt * x[s].z; acc[3] += wgt * x[s].w;
                acc[4] += wgt * y[s].x; acc[5] += wgt * y[s].y; acc[6] += wgt * y[s].z; acc[7] += wgt * y[s].w; L += wgt;
            }
        }
    } else {
        float M = -INFINITY;
        for (int s = 0; s < nsplit; s++) M = fmaxf(M, part_ml[(((int64_t)h * nsplit + s) * T + r) * 2]);
        for (int s = 0; s < nsplit; s++) {
            const int64_t pb = ((int64_t)h * nsplit + s) * T + r;
            const float m = part_ml[pb * 2], l = part_ml[pb * 2 + 1];
            if (l <= 0.f) continue;
            const float wgt = l * exp2f(m - M);
            const float4 x = *(const float4 *)(part_o + pb * FA_D + d0), y = *(const float4 *)(part_o + pb * FA_D + d0 + 4);
            acc[0] += wgt * x.x; acc[1] += wgt * x.y; acc[2] += wgt * x.z; acc[3] += wgt * x.w;
            acc[4] += wgt * y.x; acc[5] += wgt * y.y; acc[6] += wgt * y.z; acc[7] += wgt * y.w; L += wgt;
        }
    }
    const float inv = L > 0.f ? 1.f / L : 0.f;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = acc[j] * inv;
    const int64_t K = (int64_t)H * FA_D;
    quant_regs_q8K(v, a.qs + (int64_t)r * K + b * 256, a.d + (int64_t)r * nb + b, a.bsum + ((int64_t)r * nb + b) * 16, lane,
                   a.h16 ? a.h16 + (int64_t)r * K + b * 256 : nullptr, (a.dT && r < EB_DT_PITCH) ? a.dT + (int64_t)b * EB_DT_PITCH + r : nullptr);
}
void eb_fattn_combine_quant(const float *part_o, const float *part_ml, int T, int H, int nsplit, const ActQ &a, cudaStream_t st) {
    eb_launch_pdl_c(EB_PDL_COMBINE, fattn_combine_quant_kernel, dim3((unsigned)cdiv((int64_t)T * (H / 2), 8)), dim3(256), 0, st, part_o, part_ml, T, H, nsplit, a, eb_gate());
}
void eb_fattn_combine(const float *part_o, const float *part_ml, float *out, int T, int H, int nsplit, cudaStream_t st) {
    fattn_combine_kernel<<<(unsigned)cdiv((int64_t)T * H, 8), 256, 0, st>>>(part_o, part_ml, out, T, H, nsplit);
}
static int fa_pick_split(int H, int n_tiles) {
    int ns = (int)cdiv(2 * eb_sm_count(), H);
    if (ns > n_tiles) ns = n_tiles;
    if (ns < 1) ns = 1;
    return ns;
}
size_t eb_flash_attn_ws_size_impl(int T, int H, int D, int n_kv) {
    const int n_tiles = (int)cdiv(n_kv, FA_TILE);
    const int ns = fa_pick_split(H, n_tiles);
    return (size_t)H * ns * T * (D + 2) * 4 + 256;
}
int eb_flash_attn_impl(const float *q, const void *k, const void *v, int64_t kv_stride, const void *mask, int64_t mask_stride,
                       const uint32_t *tree_bits, int n_prefix, int T, int H, int D, int n_kv, float scale, float *out, void *ws, cudaStream_t st) {
    if (D != FA_D) return eb_set_err(EB_ERR_UNSUPPORTED, "flash_attn: head size %d (only 128)", D);
    if (T < 1 || T > 128) return eb_set_err(EB_ERR_ARG, "flash_attn: T=%d", T);
    if (!mask && !tree_bits) return eb_set_err(EB_ERR_ARG, "flash_attn: no mask");
    FaParams p;
    p.ctl = nullptr; p.n_ctx = n_kv;
    p.q = q; p.k = (const __half *)k; p.v = (const __half *)v; p.kv_stride = kv_stride;
    p.mask = (const __half *)mask; p.mask_stride = mask_stride; p.tree_bits = tree_bits; p.n_prefix = n_prefix;
    p.T = T; p.H = H;
    p.n_kv = mask ? n_kv : min(n_kv, n_prefix + T);       // compact form: cells past the tree are invisible
    p.n_tiles = (int)cdiv(p.n_kv, FA_TILE);
    // the split count must not depend on n_prefix so that the workspace bound holds
    p.nsplit = fa_pick_split(H, p.n_tiles);
    p.tiles_per_split = (int)cdiv(p.n_tiles, p.nsplit);
    p.nsplit = (int)cdiv(p.n_tiles, p.tiles_per_split);
    p.scale_log2 = scale * 1.44269504f;
    p.out = out; p.part_o = (float *)ws; p.part_ml = p.part_o + (size_t)H * p.nsplit * T * FA_D;
    if (p.nsplit > 1 && !ws) return eb_set_err(EB_ERR_ARG, "flash_attn: workspace required");
    const int nw = max(4, (T + 15) / 16);
    const dim3 grid(H, p.nsplit);
    if (mask) fattn_kernel<1><<<grid, nw * 32, 65536, st>>>(p); else fattn_kernel<0><<<grid, nw * 32, 65536, st>>>(p);
    EB_CHECK_LAUNCH();
    if (p.nsplit > 1) {
        fattn_combine_kernel<<<(unsigned)cdiv((int64_t)T * H, 8), 256, 0, st>>>(p.part_o, p.part_ml, out, T, H, p.nsplit);
        EB_CHECK_LAUNCH();
    }
    return EB_OK;
}

int eb_flash_attn_ctl(const float *q, const void *k, const void *v, int64_t kv_stride, const uint32_t *tree_bits, const FwdCtl *ctl,
                      int T, int H, int D, int n_ctx, float scale, float *out, void *ws, cudaStream_t st) {
    if (D != FA_D) return eb_set_err(EB_ERR_UNSUPPORTED, "flash_attn: head size %d (only 128)", D);
    if (T < 1 || T > 128) return eb_set_err(EB_ERR_ARG, "flash_attn: T=%d", T);
    FaParams p;
    p.q = q; p.k = (const __half *)k; p.v = (const __half *)v; p.kv_stride = kv_stride;
    p.mask = nullptr; p.mask_stride = 0; p.tree_bits = tree_bits; p.n_prefix = 0; p.ctl = ctl; p.n_ctx = n_ctx;
    p.T = T; p.H = H; p.n_kv = n_ctx; p.n_tiles = (int)cdiv(n_ctx, FA_TILE);
    p.nsplit = fa_pick_split(H, p.n_tiles); p.tiles_per_split = (int)cdiv(p.n_tiles, p.nsplit);
    p.scale_log2 = scale * 1.44269504f;
    p.out = out; p.part_o = (float *)ws; p.part_ml = p.part_o + (size_t)H * p.nsplit * T * FA_D;
    const int nw = max(4, (T + 15) / 16);
    fattn_kernel<0><<<dim3(H, p.nsplit), nw * 32, 65536, st>>>(p);
    EB_CHECK_LAUNCH();
    if (p.nsplit > 1) {
        fattn_combine_kernel<<<(unsigned)cdiv((int64_t)T * H, 8), 256, 0, st>>>(p.part_o, p.part_ml, out, T, H, p.nsplit);
        EB_CHECK_LAUNCH();
    } 
    return EB_OK;
}

int eb_fattn_init() {
    EB_CUDA(cudaFuncSetAttribute(fattn_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    EB_CUDA(cudaFuncSetAttribute(fattn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    return EB_OK;
}
