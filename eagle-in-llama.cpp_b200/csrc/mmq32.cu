// mmq32.cu — tree-batched GEMM for the 32-element block formats (Q4_0, Q8_0) on int8 tensor cores (mma.sync.m16n8k32),
// T = 2..128 columns.  Replaces mul_mat_q<GGML_TYPE_Q4_0 / Q8_0> (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822); oracle
// ggml_vec_dot_q4_0_q8_0 / ggml_vec_dot_q8_0_q8_0 (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:1912, 3663): per 32-block an exact
// int32 dot, scaled by d_w * d_x in f32.  Before this kernel those formats went through the matvec kernel 8 columns at a time
// (9 passes over the weights for a 65-row verify batch, activations re-read from L2 per block: 67 ms per Q8_0 7B verify step).
//
// CTA = 4 warps x 16 weight rows = 64 rows, all T columns (NT tiles of 8), a K range of the split.  Activations (Q8_0 grid: int8
// quants + f32 block scales) ride a double-buffered cp.async ring in shared memory, one chunk = 8 blocks = 256 values per column;
// weights go HBM -> registers (one chunk ahead) in the device planes layout (common.cuh WPlanes).  Bound: HBM; the scale
// application (4 I2F + 4 FMUL + 4 FFMA per MMA) keeps the issue slots at ~1.5x the HBM time for T = 65.
#include "common.cuh"
#include <algorithm>

namespace {

constexpr int KC = 8;                                     // blocks per chunk
template <int FMT> struct L32 {
    static constexpr int PITCH = FMT == 1 ? 288 : 272;    // bytes per activation row of a chunk: conflict-free 8-byte (Q8_0) / 4-byte (Q4_0) fragment loads
};

__device__ __forceinline__ void cp16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void mma_s8_k32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

struct WChunk { uint32_t a[KC][4]; __half d0[KC], d1[KC]; };

// A fragments of one chunk for rows r0 (= g) and r1 (= g + 8).  Q8_0: the lane takes 8 contiguous bytes per (row, block) —
// k-slots of the MMA are permuted consistently on the B side, the dot over the 32 values is the same.  Q4_0: 4 bytes per
// (row, block): low nibbles are values 4t..4t+3, high nibbles 16+4t.. (block_q4_0 layout), which is the MMA's own k order.
template <int FMT>
__device__ __forceinline__ void load_chunk(WChunk &w, const WPlanes &p, int64_t r0, int64_t r1, int nbr, int b0, int t) {
#pragma unroll
    for (int bb = 0; bb < KC; bb++) {
        const int64_t k0 = r0 * nbr + b0 + bb, k1 = r1 * nbr + b0 + bb;
        if (FMT == 1) {
            const uint2 x = __ldg((const uint2 *)(p.q80_qs() + k0 * 32 + 8 * t)), y = __ldg((const uint2 *)(p.q80_qs() + k1 * 32 + 8 * t));
            w.a[bb][0] = x.x; w.a[bb][1] = y.x; w.a[bb][2] = x.y; w.a[bb][3] = y.y;
            w.d0[bb] = p.q80_d()[k0]; w.d1[bb] = p.q80_d()[k1];
        } else {
            const uint32_t x = __ldg((const uint32_t *)(p.q40_qs() + k0 * 16 + 4 * t)), y = __ldg((const uint32_t *)(p.q40_qs() + k1 * 16 + 4 * t));
            w.a[bb][0] = __vsubss4(x & 0x0F0F0F0Fu, 0x08080808u); w.a[bb][1] = __vsubss4(y & 0x0F0F0F0Fu, 0x08080808u);
            w.a[bb][2] = __vsubss4((x >> 4) & 0x0F0F0F0Fu, 0x08080808u); w.a[bb][3] = __vsubss4((y >> 4) & 0x0F0F0F0Fu, 0x08080808u);
            w.d0[bb] = p.q40_d()[k0]; w.d1[bb] = p.q40_d()[k1];
        }
    }
}

template <int FMT, int NT>
__global__ void __launch_bounds__(128) mmq32_kernel(WPlanes wp, int64_t N, int nbr, ActQ a, float *__restrict__ out, int chunks_per_split) {
    constexpr int TP = NT * 8, PITCH = L32<FMT>::PITCH, STAGE = TP * PITCH + TP * KC * 4;
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int n_chunks = nbr / KC;
    const int c0 = blockIdx.y * chunks_per_split, c1 = min(n_chunks, c0 + chunks_per_split);
    const int64_t row0 = (int64_t)blockIdx.x * 64 + warp * 16 + g;
    const int64_t r0 = row0 < N ? row0 : 0, r1 = row0 + 8 < N ? row0 + 8 : 0;
    // rows >= T of the activation tiles are never loaded: zero them once in both stages
    for (int i = tid; i < 2 * STAGE / 16; i += 128) ((int4 *)smem)[i] = make_int4(0, 0, 0, 0);
    __syncthreads();
    auto stage_act = [&](int s, int c) {
        uint8_t *st = smem + s * STAGE;
        for (int i = tid; i < a.T * 16; i += 128) {                        // 16-byte pieces of the quants: [col][16]
            const int col = i >> 4, pc = i & 15;
            cp16(st + col * PITCH + pc * 16, a.qs + (int64_t)col * a.K + (int64_t)c * (KC * 32) + pc * 16);
        }
        for (int i = tid; i < a.T * 2; i += 128) {                         // block scales: KC floats = 2 pieces per column
            const int col = i >> 1, pc = i & 1;
            cp16(st + TP * PITCH + col * (KC * 4) + pc * 16, a.d + (int64_t)col * nbr + c * KC + pc * 4);
        }
        asm volatile("cp.async.commit_group;");
    };
    float acc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) { acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f; }
    WChunk nxt;
    if (c0 < c1) { stage_act(0, c0); load_chunk<FMT>(nxt, wp, r0, r1, nbr, c0 * KC, t); }
    for (int c = c0; c < c1; c++) {
        const int s = (c - c0) & 1;
        const WChunk cur = nxt;
        if (c + 1 < c1) { stage_act(s ^ 1, c + 1); load_chunk<FMT>(nxt, wp, r0, r1, nbr, (c + 1) * KC, t); asm volatile("cp.async.wait_group 1;"); }
        else asm volatile("cp.async.wait_group 0;");
        __syncthreads();
        const uint8_t *st = smem + s * STAGE;
        const float *dx = (const float *)(st + TP * PITCH);
#pragma unroll
        for (int bb = 0; bb < KC; bb++) {
            const float dw0 = __half2float(cur.d0[bb]), dw1 = __half2float(cur.d1[bb]);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                const uint8_t *br = st + (nt * 8 + g) * PITCH + bb * 32;
                uint32_t b0, b1;
                if (FMT == 1) { const uint2 v = *(const uint2 *)(br + 8 * t); b0 = v.x; b1 = v.y; }
                else { b0 = *(const uint32_t *)(br + 4 * t); b1 = *(const uint32_t *)(br + 16 + 4 * t); }
                int cc[4] = {0, 0, 0, 0};
                mma_s8_k32(cc, cur.a[bb][0], cur.a[bb][1], cur.a[bb][2], cur.a[bb][3], b0, b1);
                const float x0 = dx[(nt * 8 + 2 * t) * KC + bb], x1 = dx[(nt * 8 + 2 * t + 1) * KC + bb];
                acc[nt][0] += (float)cc[0] * (dw0 * x0); acc[nt][1] += (float)cc[1] * (dw0 * x1);
                acc[nt][2] += (float)cc[2] * (dw1 * x0); acc[nt][3] += (float)cc[3] * (dw1 * x1);
            }
        }
        __syncthreads();                                                    // the stage is overwritten two iterations later
    }
    float *o = out + (int64_t)blockIdx.y * a.T * N;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
        const int col = nt * 8 + 2 * t;
        if (col < a.T) { if (row0 < N) o[(int64_t)col * N + row0] = acc[nt][0]; if (row0 + 8 < N) o[(int64_t)col * N + row0 + 8] = acc[nt][2]; }
        if (col + 1 < a.T) { if (row0 < N) o[(int64_t)(col + 1) * N + row0] = acc[nt][1]; if (row0 + 8 < N) o[(int64_t)(col + 1) * N + row0 + 8] = acc[nt][3]; }
    }
}

template <int FMT, int NT> size_t smem32() { return 2 * ((size_t)NT * 8 * L32<FMT>::PITCH + (size_t)NT * 8 * KC * 4); }

template <int FMT, int NT>
int launch32(const WPlanes &wp, int64_t N, int nbr, const ActQ &a, float *planes, int ks, int cps, cudaStream_t st) {
    mmq32_kernel<FMT, NT><<<dim3((unsigned)cdiv(N, 64), ks), 128, smem32<FMT, NT>(), st>>>(wp, N, nbr, a, planes, cps);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
template <int FMT>
int dispatch32(const WPlanes &wp, int64_t N, int nbr, const ActQ &a, float *planes, int ks, int cps, cudaStream_t st) {
    const int nt = (a.T + 7) / 8;
    if (nt <= 1) return launch32<FMT, 1>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 2) return launch32<FMT, 2>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 4) return launch32<FMT, 4>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 9) return launch32<FMT, 9>(wp, N, nbr, a, planes, ks, cps, st);
    return launch32<FMT, 16>(wp, N, nbr, a, planes, ks, cps, st);
}

}  // namespace

int eb_mmq32_supported(int type, int64_t N, int64_t K, int T) {
    static const bool off = getenv("EB_NO_MMQ32") != nullptr;
    return !off && (type == EB_Q4_0 || type == EB_Q8_0) && K % 256 == 0 && N >= 1 && T >= 2 && T <= 128;
}

// planes [ks][T][N] (unreduced split-K partials; ks <= 8 returned)
int eb_mmq32_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st) {
    if (!eb_mmq32_supported(type, N, K, a.T)) return eb_set_err(EB_ERR_UNSUPPORTED, "mmq32: type %d N %lld K %lld T %d", type, (long long)N, (long long)K, a.T);
    if (a.kq) return eb_set_err(EB_ERR_ARG, "mmq32: activations are not on the Q8_0 grid");
    const int nbr = (int)(K / 32), n_chunks = nbr / KC;
    const int64_t rb = cdiv(N, 64);
    // enough CTAs to fill the machine ~3x over (each CTA walks its chunks serially), at least 2 chunks per split, at most 8 planes
    int ks = (int)std::min<int64_t>(8, std::max<int64_t>(1, cdiv(3 * (int64_t)eb_sm_count(), rb)));
    ks = std::max(1, std::min(ks, n_chunks / 2));
    int cps = (int)cdiv(n_chunks, ks); ks = (int)cdiv(n_chunks, cps);
    *ks_out = ks;
    WPlanes wp{(const uint8_t *)W, N * nbr};
    return type == EB_Q8_0 ? dispatch32<1>(wp, N, nbr, a, planes, ks, cps, st) : dispatch32<0>(wp, N, nbr, a, planes, ks, cps, st);
}

// dynamic-smem opt-in for every instantiation, once per process (never inside a graph capture)
template <int FMT, int NT> static int init32() { EB_CUDA(cudaFuncSetAttribute(mmq32_kernel<FMT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem32<FMT, NT>())); return EB_OK; }
int eb_mmq32_init() {
    int rc;
    if ((rc = init32<0, 1>()) || (rc = init32<0, 2>()) || (rc = init32<0, 4>()) || (rc = init32<0, 9>()) || (rc = init32<0, 16>())) return rc;
    if ((rc = init32<1, 1>()) || (rc = init32<1, 2>()) || (rc = init32<1, 4>()) || (rc = init32<1, 9>()) || (rc = init32<1, 16>())) return rc;
    return EB_OK;
}
