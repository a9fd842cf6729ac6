// mmq_tc.cu — K2 on 5th-generation tensor cores: tree-batched (T = 9..80) verify GEMM on Q4_K / Q6_K
// weights with tcgen05.mma (kind::f16, accumulators in TMEM), weights streamed from HBM exactly once
// by TMA.  Replaces mul_mat_q (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822); results follow
// ggml_vec_dot_q4_K_q8_K / q6_K_q8_K (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:6266,7903): all integer
// sums are exact, only the order of the f32 additions over super-blocks differs.
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
//   1      tcgen05.mma issuer (one thread)
//   2      TMA producer of the activation tiles (f16 quants, 16-sums)
//   3      TMEM allocator
//   4-11   dequant: smem raw blocks -> exact f16 integers -> TMEM A operand (lane = weight row)
//   12-19  epilogue: TMEM accumulators -> f32 registers (d*dx scaling), final store of the split-K plane
#include "common.cuh"
#include "tc.cuh"
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
constexpr int OFF_BAR = OFF_HDR + 4 * 128 * 8;  // 217088
constexpr int SMEM_BYTES = OFF_BAR + 512 + 1024;   // + barriers + alignment slack
// TMEM columns
constexpr int D_STRIDE = 160, D_MIN = 80, A_OFF = 320, AM_OFF = 448;
// barrier indices
enum { W_FULL = 0, W_EMPTY = 6, B_FULL = 12, B_EMPTY = 14, BS_FULL = 16, BS_EMPTY = 18, A_FULL = 20, A_EMPTY = 24,
       D_FULL = 28, D_EMPTY = 30, HDR_FULL = 32, N_BARS = 36 };

struct TcArgs {
    const float *dT; float *out; int64_t N; int nbr, T, npad, n_rb, n_items, spp;
};

__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t *>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2 *>(&u); }
// byte b (0..3) of word w as half2 (v, v), v in [0,255] exact
__device__ __forceinline__ __half2 byte_to_h2(uint32_t w, int b) {
    const uint32_t v = (w >> (8 * b)) & 0xffu;
    return __hsub2(u2h(v | (v << 16) | 0x64006400u), u2h(0x64006400u));
}

__global__ void __launch_bounds__(TC_THREADS, 1)
mmq_tc_q4_K(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmBS, TcArgs p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    const uint32_t sbase = tc::smem_u32(smem);
    const uint32_t bar0 = sbase + OFF_BAR;
    auto bar = [&](int i) { return bar0 + 8u * i; };
    volatile uint32_t *tmem_slot = (volatile uint32_t *)(smem + OFF_BAR + 8 * N_BARS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < S_W; i++) { tc::mbar_init(bar(W_FULL + i), 1); tc::mbar_init(bar(W_EMPTY + i), 8); }
        for (int i = 0; i < 2; i++) {
            tc::mbar_init(bar(B_FULL + i), 1); tc::mbar_init(bar(B_EMPTY + i), 1);
            tc::mbar_init(bar(BS_FULL + i), 1); tc::mbar_init(bar(BS_EMPTY + i), 1);
            tc::mbar_init(bar(D_FULL + i), 1); tc::mbar_init(bar(D_EMPTY + i), 8);
        }
        for (int i = 0; i < 4; i++) { tc::mbar_init(bar(A_FULL + i), 4); tc::mbar_init(bar(A_EMPTY + i), 1); tc::mbar_init(bar(HDR_FULL + i), 4); }
        tc::fence_barrier_init();
        tc::tma_prefetch_desc(&tmW); tc::tma_prefetch_desc(&tmB); tc::tma_prefetch_desc(&tmBS);
    }
    if (warp == 3) tc::tmem_alloc(tc::smem_u32((const void *)tmem_slot), 512);
    tc::fence_before();
    __syncthreads();
    tc::fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ---------------------------------------------------------------- weight producer
        if (lane == 0) {
            uint32_t i = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                const int ks = item / p.n_rb, rb = item % p.n_rb;
                const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                for (int sb = sb0; sb < sb1; sb++, i++) {
                    const uint32_t s = i % S_W, ph = (i / S_W) & 1;
                    tc::mbar_wait(bar(W_EMPTY + s), ph ^ 1);
                    tc::mbar_expect_tx(bar(W_FULL + s), W_STAGE);
                    tc::tma_load_2d(sbase + OFF_W + s * W_STAGE, &tmW, sb * 144, rb * 128, bar(W_FULL + s));
                }
            }
        }
        __syncwarp();
    } else if (warp == 2) {
        // ---------------------------------------------------------------- activation producer
        if (lane == 0) {
            uint32_t i = 0, nbx = 0;
            const uint32_t box_bytes = (uint32_t)p.npad * 128u;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                const int ks = item / p.n_rb;
                const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                for (int sb = sb0; sb < sb1; sb++, i++) {
                    if (sb == sb0 || (sb & 3) == 0) {
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
        __syncwarp();
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_f16_m128(p.npad);
            uint32_t i = 0, nbx = 0; int cur_bs = -1;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
                const int ks = item / p.n_rb;
                const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
                for (int sb = sb0; sb < sb1; sb++, i++) {
                    if (sb == sb0 || (sb & 3) == 0) {
                        if (cur_bs >= 0) tc::mma_commit(bar(BS_EMPTY + (cur_bs & 1)));
                        cur_bs = (int)nbx;
                        tc::mbar_wait(bar(BS_FULL + (nbx & 1)), (nbx >> 1) & 1);
                        nbx++;
                    }
                    const uint32_t s = i & 1, ph = (i >> 1) & 1;
                    tc::mbar_wait(bar(B_FULL + s), ph);
                    tc::mbar_wait(bar(D_EMPTY + s), ph ^ 1);
                    tc::fence_after();
                    const uint32_t d_main = tmem + s * D_STRIDE, d_min = d_main + D_MIN;
#pragma unroll
                    for (int g = 0; g < 4; g++) {
                        tc::mbar_wait(bar(A_FULL + g), i & 1);
                        tc::fence_after();
                        const uint32_t bsm = sbase + OFF_B + s * B_STAGE + g * B_BOX;
#pragma unroll
                        for (int j = 0; j < 4; j++)
                            tc::mma_f16_ts(d_main, tmem + A_OFF + 32 * g + 8 * j, tc::smem_desc_sw128(bsm + 32 * j), idesc, (g | j) != 0);
                        if (g == 0)
                            tc::mma_f16_ts(d_min, tmem + AM_OFF + 8 * (i & 1), tc::smem_desc_sw128(sbase + OFF_BS + (cur_bs & 1) * B_BOX + 32 * (sb & 3)), idesc, 0);
                        tc::mma_commit(bar(A_EMPTY + g));
                    }
                    tc::mma_commit(bar(B_EMPTY + s));
                    tc::mma_commit(bar(D_FULL + s));
                }
            }
        }
        __syncwarp();
    } else if (warp >= 4 && warp < 12) {
        // ---------------------------------------------------------------- dequant: raw Q4_K -> f16 integers in TMEM
        const int q = warp & 3, member = (warp - 4) >> 2, row = 32 * q + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
        float2 *hdr_sm = (float2 *)(smem + OFF_HDR);
        uint32_t i = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const int ks = item / p.n_rb;
            const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
            for (int sb = sb0; sb < sb1; sb++, i++) {
                const uint32_t s = i % S_W;
                tc::mbar_wait(bar(W_FULL + s), (i / S_W) & 1);
                const uint8_t *rp = smem + OFF_W + s * W_STAGE + row * 144;
                const uint4 hd = *(const uint4 *)rp;
                uint4 qv[2][2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    qv[h][0] = *(const uint4 *)(rp + 16 + 32 * (member + 2 * h));
                    qv[h][1] = *(const uint4 *)(rp + 32 + 32 * (member + 2 * h));
                }
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(W_EMPTY + s));
                // 6-bit scales / mins (get_scale_min_k4, R/ggml/src/ggml-quants.c:631) as bytes of u[0..1] / u[2..3]
                uint32_t u[4];
                u[0] = hd.y & 0x3f3f3f3fu; u[1] = (hd.w & 0x0f0f0f0fu) | (((hd.y >> 6) & 0x03030303u) << 4);
                u[2] = hd.z & 0x3f3f3f3fu; u[3] = ((hd.w >> 4) & 0x0f0f0f0fu) | (((hd.z >> 6) & 0x03030303u) << 4);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int g = member + 2 * h;                       // 64-element group: sub-blocks 2g (low nibbles), 2g+1 (high)
                    const __half2 s_lo = byte_to_h2(u[g >> 1], (2 * g) & 3), s_hi = byte_to_h2(u[g >> 1], (2 * g + 1) & 3);
                    const __half2 mul_lo = s_lo, add_lo = __hmul2(s_lo, __float2half2_rn(-1024.f));
                    const __half2 mul_hi = __hmul2(s_hi, __float2half2_rn(0.0625f)), add_hi = __hmul2(s_hi, __float2half2_rn(-64.f));
                    uint32_t r[32];
                    const uint32_t xs[8] = {qv[h][0].x, qv[h][0].y, qv[h][0].z, qv[h][0].w, qv[h][1].x, qv[h][1].y, qv[h][1].z, qv[h][1].w};
#pragma unroll
                    for (int w = 0; w < 8; w++) {
                        const uint32_t y = __byte_perm(xs[w], 0, 0x4140), z = __byte_perm(xs[w], 0, 0x4342);
                        r[2 * w]          = h2u(__hfma2(u2h((y & 0x000f000fu) | 0x64006400u), mul_lo, add_lo));
                        r[2 * w + 1]      = h2u(__hfma2(u2h((z & 0x000f000fu) | 0x64006400u), mul_lo, add_lo));
                        r[16 + 2 * w]     = h2u(__hfma2(u2h((y & 0x00f000f0u) | 0x64006400u), mul_hi, add_hi));
                        r[16 + 2 * w + 1] = h2u(__hfma2(u2h((z & 0x00f000f0u) | 0x64006400u), mul_hi, add_hi));
                    }
                    tc::mbar_wait(bar(A_EMPTY + g), (i & 1) ^ 1);
                    tc::fence_after();
                    if (g == 0) {                                        // member 0, first group: header + mins operand
                        const __half2 dm = u2h(hd.x);
                        hdr_sm[(i & 3) * 128 + row] = make_float2(__low2float(dm), __high2float(dm));
                        uint32_t mr[8];
#pragma unroll
                        for (int j = 0; j < 8; j++) mr[j] = h2u(byte_to_h2(u[2 + (j >> 2)], j & 3));
                        tc::tmem_st8(lane_base + AM_OFF + 8 * (i & 1), mr);
                    }
                    tc::tmem_st32(lane_base + A_OFF + 32 * g, r);
                    tc::wait_st();
                    tc::fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        tc::mbar_arrive(bar(A_FULL + g));
                        if (g == 0) tc::mbar_arrive(bar(HDR_FULL + (i & 3)));
                    }
                }
            }
        }
    } else if (warp >= 12) {
        // ---------------------------------------------------------------- epilogue
        const int q = warp & 3, hf = (warp - 12) >> 2, row = 32 * q + lane, c_base = 40 * hf;
        const uint32_t lane_base = tmem + ((uint32_t)(32 * q) << 16);
        const float2 *hdr_sm = (const float2 *)(smem + OFF_HDR);
        uint32_t i = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const int ks = item / p.n_rb, rb = item % p.n_rb;
            const int sb0 = ks * p.spp, sb1 = min(p.nbr, sb0 + p.spp);
            float acc[40];
#pragma unroll
            for (int c = 0; c < 40; c++) acc[c] = 0.f;
            for (int sb = sb0; sb < sb1; sb++, i++) {
                const float dxa = (c_base + lane < p.T) ? __ldg(p.dT + (int64_t)sb * EB_DT_PITCH + c_base + lane) : 0.f;
                const float dxb = (lane < 8 && c_base + 32 + lane < p.T) ? __ldg(p.dT + (int64_t)sb * EB_DT_PITCH + c_base + 32 + lane) : 0.f;
                tc::mbar_wait(bar(HDR_FULL + (i & 3)), (i >> 2) & 1);
                const float2 h = hdr_sm[(i & 3) * 128 + row];
                const uint32_t s = i & 1;
                tc::mbar_wait(bar(D_FULL + s), (i >> 1) & 1);
                tc::fence_after();
#pragma unroll
                for (int ch = 0; ch < 5; ch++) {
                    if (c_base + 8 * ch < p.npad) {
                        uint32_t vm[8], vn[8];
                        tc::tmem_ld8(lane_base + s * D_STRIDE + c_base + 8 * ch, vm);
                        tc::tmem_ld8(lane_base + s * D_STRIDE + D_MIN + c_base + 8 * ch, vn);
                        tc::wait_ld();
#pragma unroll
                        for (int e = 0; e < 8; e++) {
                            const int idx = 8 * ch + e;
                            const float dx = __shfl_sync(0xffffffffu, idx < 32 ? dxa : dxb, idx & 31);
                            acc[idx] += (h.x * dx) * __uint_as_float(vm[e]) - (h.y * dx) * __uint_as_float(vn[e]);
                        }
                    }
                }
                tc::fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(bar(D_EMPTY + s));
            }
            const int64_t grow = (int64_t)rb * 128 + row;
            if (grow < p.N) {
                float *o = p.out + (int64_t)ks * p.T * p.N + grow;
#pragma unroll
                for (int c = 0; c < 40; c++)
                    if (c_base + c < p.T) o[(int64_t)(c_base + c) * p.N] = acc[c];
            }
        }
    }
    tc::fence_before();
    __syncthreads();
    if (warp == 3) tc::tmem_dealloc(tmem, 512);
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

int eb_mmq_tc_init() {
    if (!g_encode) {
        void *fn = nullptr; cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
            cudaGetLastError();
            return eb_set_err(EB_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
        }
        g_encode = (EncodeTiledFn)fn;
    }
    EB_CUDA(cudaFuncSetAttribute(mmq_tc_q4_K, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    return EB_OK;
}

int eb_mmq_tc_supported(int type, int64_t N, int64_t K, int T) {
    static const int off = getenv("EB_NO_TC") ? 1 : 0;
    if (off || !g_encode) return 0;
    return type == EB_Q4_K && K % 256 == 0 && N >= 128 && T > 8 && T <= 80;
}

// split-K factor: planes <= 8, minimise rounds x (super-blocks per item + pipeline fill)
static int tc_pick_ksplit(int64_t N, int nbr) {
    const int n_rb = (int)cdiv(N, 128), G = eb_sm_count();
    int best = 1; double best_cost = 1e30;
    for (int ks = 1; ks <= 8 && ks <= nbr; ks++) {
        const int spp = (int)cdiv(nbr, ks), eff_ks = (int)cdiv(nbr, spp);
        if (eff_ks != ks) continue;
        const int rounds = (int)cdiv((int64_t)n_rb * ks, G);
        const double cost = rounds * (spp + 2.0) + 0.25 * ks;
        if (cost < best_cost) { best_cost = cost; best = ks; }
    }
    return best;
}

// planes [ks][T][N] (unreduced split-K partials; ks returned)
int eb_mmq_tc_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st) {
    if (!eb_mmq_tc_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq_tc: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    if (!a.h16 || !a.dT || !a.bsum) return eb_set_err(EB_ERR_ARG, "mmq_tc: activations lack the f16 planes");
    const int nbr = (int)(K / 256);
    const int ks = tc_pick_ksplit(N, nbr);
    TcArgs p;
    p.dT = a.dT; p.out = planes; p.N = N; p.nbr = nbr; p.T = a.T; p.npad = (a.T + 15) / 16 * 16;
    p.n_rb = (int)cdiv(N, 128); p.spp = (int)cdiv(nbr, ks); p.n_items = p.n_rb * ks;
    CUtensorMap tmW, tmB, tmBS;
    int rc;
    if ((rc = make_map_2d(&tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, W, (uint64_t)nbr * 144, (uint64_t)N, (uint64_t)nbr * 144, 144, 128, CU_TENSOR_MAP_SWIZZLE_NONE))) return rc;
    if ((rc = make_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, a.h16, (uint64_t)K, (uint64_t)a.T, (uint64_t)K * 2, 64, (uint32_t)p.npad, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = make_map_2d(&tmBS, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, a.bsum, (uint64_t)K / 16, (uint64_t)a.T, (uint64_t)K / 16 * 2, 64, (uint32_t)p.npad, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    const int grid = std::min(p.n_items, eb_sm_count());
    mmq_tc_q4_K<<<grid, TC_THREADS, SMEM_BYTES, st>>>(tmW, tmB, tmBS, p);
    EB_CHECK_LAUNCH();
    *ks_out = ks;
    return EB_OK;
}
