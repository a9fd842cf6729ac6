// mmq32.cu — tree-batched GEMM for the 32-element block formats (Q4_0, Q8_0) on int8 tensor cores (mma.sync.m16n8k32),
// T = 2..128 columns.  Replaces mul_mat_q<GGML_TYPE_Q4_0 / Q8_0> (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822); oracle
// ggml_vec_dot_q4_0_q8_0 / ggml_vec_dot_q8_0_q8_0 (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:1912, 3663): per 32-block an exact
// int32 dot, scaled by d_w * d_x in f32.  Before this kernel those formats went through the matvec kernel 8 columns at a time
// (9 passes over the weights for a 65-row verify batch, activations re-read from L2 per block: 67 ms per Q8_0 7B verify step).
//
// CTA = 4 warps x 16 weight rows = 64 rows, all T columns (NT tiles of 8), a K range of the split.  Activations (Q8_0 grid: int8
// quants + f32 block scales) ride a double-buffered cp.async ring in shared memory, one chunk = 8 blocks = 256 values per column;
// weights go HBM -> registers (one chunk ahead) in the device planes layout (common.cuh WPlanes).  Bound: HBM; the scale
// application is 2 FADD2 + 2 FMUL2 + 2 FFMA2 per MMA (magic-number accumulator init instead of I2F, packed f32x2 arithmetic).
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
__device__ __forceinline__ void cp4(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
// int8 MMA whose accumulator starts at the bit pattern of 1.5 * 2^23: the s32 result, read as f32, is 12582912 + dot exactly
// (|dot| <= 32 * 127 * 128 < 2^22, Q4_0's 16x-scaled nibbles included), so the int -> float conversion is one FADD instead of an I2F per output.
constexpr int MAGIC_I = 0x4B400000;
__device__ __forceinline__ void mma_s8_k32(uint64_t &c01, uint64_t &c23, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    int c0, c1, c2, c3;
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(c0), "=r"(c1), "=r"(c2), "=r"(c3) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(MAGIC_I));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c01) : "r"(c0), "r"(c1));
    asm("mov.b64 %0, {%1,%2};" : "=l"(c23) : "r"(c2), "r"(c3));
}
// packed f32x2 arithmetic (FADD2 / FMUL2 / FFMA2): two outputs per issue slot
__device__ __forceinline__ uint64_t pack2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) { uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ void fma2(uint64_t &acc, uint64_t a, uint64_t b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b)); }

struct WChunk { uint32_t a[KC][4]; uint4 d0, d1; };      // d0 / d1: the chunk's 8 f16 block scales of rows g / g + 8, one 16-byte load each

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
        } else {
            const uint32_t x = __ldg((const uint32_t *)(p.q40_qs() + k0 * 16 + 4 * t)), y = __ldg((const uint32_t *)(p.q40_qs() + k1 * 16 + 4 * t));
            // nibble n in the high half of a byte, top bit flipped, is the s8 value 16 * (n - 8): one LOP3 per register (plus a
            // shift for the low nibbles); the factor 16 is taken out of the block scale (exact: a power of two)
            w.a[bb][0] = ((x << 4) & 0xF0F0F0F0u) ^ 0x80808080u; w.a[bb][1] = ((y << 4) & 0xF0F0F0F0u) ^ 0x80808080u;
            w.a[bb][2] = (x & 0xF0F0F0F0u) ^ 0x80808080u; w.a[bb][3] = (y & 0xF0F0F0F0u) ^ 0x80808080u;
        }
    }
    const __half *dp = FMT == 1 ? p.q80_d() : p.q40_d();
    w.d0 = __ldg((const uint4 *)(dp + r0 * nbr + b0)); w.d1 = __ldg((const uint4 *)(dp + r1 * nbr + b0));
}
__device__ __forceinline__ float scale_of(const uint4 &v, int bb) {
    const uint32_t w = bb < 2 ? v.x : bb < 4 ? v.y : bb < 6 ? v.z : v.w;
    return __half2float(__ushort_as_half((unsigned short)((bb & 1) ? (w >> 16) : (w & 0xFFFFu))));
}

template <int FMT, int NT>
__global__ void __launch_bounds__(128) mmq32_kernel(WPlanes wp, int64_t N, int nbr, ActQ a, float *__restrict__ out, int chunks_per_split, const int *gate) {
    if (gate_closed(gate)) return;                       // idle draft level (common.cuh "launch gate"): the planes are not read either
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
        for (int i = tid; i < a.T * KC; i += 128) {                        // block scales, transposed to [block][col]: (x0, x1) is one 8-byte load
            const int col = i >> 3, bb = i & 7;
            cp4(st + TP * PITCH + (bb * TP + col) * 4, a.d + (int64_t)col * nbr + c * KC + bb);
        }
        asm volatile("cp.async.commit_group;");
    };
    uint64_t acc01[NT], acc23[NT];                                          // (row g: cols 2t, 2t+1), (row g + 8: same cols)
#pragma unroll
    for (int nt = 0; nt < NT; nt++) { acc01[nt] = acc23[nt] = pack2(0.f, 0.f); }
    const uint64_t neg_magic = pack2(-12582912.f, -12582912.f);
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
            constexpr float WS = FMT == 0 ? 0.0625f : 1.f;
            const float dw0f = scale_of(cur.d0, bb) * WS, dw1f = scale_of(cur.d1, bb) * WS;
            const uint64_t dw0 = pack2(dw0f, dw0f), dw1 = pack2(dw1f, dw1f);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
                const uint8_t *br = st + (nt * 8 + g) * PITCH + bb * 32;
                uint32_t b0, b1;
                if (FMT == 1) { const uint2 v = *(const uint2 *)(br + 8 * t); b0 = v.x; b1 = v.y; }
                else { b0 = *(const uint32_t *)(br + 4 * t); b1 = *(const uint32_t *)(br + 16 + 4 * t); }
                uint64_t c01, c23;
                mma_s8_k32(c01, c23, cur.a[bb][0], cur.a[bb][1], cur.a[bb][2], cur.a[bb][3], b0, b1);
                const uint64_t x01 = *(const uint64_t *)(dx + bb * TP + nt * 8 + 2 * t);
                fma2(acc01[nt], add2(c01, neg_magic), mul2(dw0, x01));
                fma2(acc23[nt], add2(c23, neg_magic), mul2(dw1, x01));
            }
        }
        __syncthreads();                                                    // the stage is overwritten two iterations later
    }
    float *o = out + (int64_t)blockIdx.y * a.T * N;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
        const int col = nt * 8 + 2 * t;
        float v0, v1, v2, v3;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(v0), "=f"(v1) : "l"(acc01[nt]));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(v2), "=f"(v3) : "l"(acc23[nt]));
        if (col < a.T) { if (row0 < N) o[(int64_t)col * N + row0] = v0; if (row0 + 8 < N) o[(int64_t)col * N + row0 + 8] = v2; }
        if (col + 1 < a.T) { if (row0 < N) o[(int64_t)(col + 1) * N + row0] = v1; if (row0 + 8 < N) o[(int64_t)(col + 1) * N + row0 + 8] = v3; }
    }
}

template <int FMT, int NT> size_t smem32() { return 2 * ((size_t)NT * 8 * L32<FMT>::PITCH + (size_t)NT * 8 * KC * 4); }

template <int FMT, int NT>
int launch32(const WPlanes &wp, int64_t N, int nbr, const ActQ &a, float *planes, int ks, int cps, cudaStream_t st) {
    mmq32_kernel<FMT, NT><<<dim3((unsigned)cdiv(N, 64), ks), 128, smem32<FMT, NT>(), st>>>(wp, N, nbr, a, planes, cps, eb_gate());
    EB_CHECK_LAUNCH();
    return EB_OK;
}
template <int FMT>
int dispatch32(const WPlanes &wp, int64_t N, int nbr, const ActQ &a, float *planes, int ks, int cps, cudaStream_t st) {
    const int nt = (a.T + 7) / 8;
    if (nt <= 1) return launch32<FMT, 1>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 2) return launch32<FMT, 2>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 3) return launch32<FMT, 3>(wp, N, nbr, a, planes, ks, cps, st);      // 16-node tree + root
    if (nt <= 4) return launch32<FMT, 4>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 5) return launch32<FMT, 5>(wp, N, nbr, a, planes, ks, cps, st);      // 32-node tree + root
    if (nt <= 7) return launch32<FMT, 7>(wp, N, nbr, a, planes, ks, cps, st);
    if (nt <= 9) return launch32<FMT, 9>(wp, N, nbr, a, planes, ks, cps, st);      // 64-node tree + root
    if (nt <= 12) return launch32<FMT, 12>(wp, N, nbr, a, planes, ks, cps, st);
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
    if (n_chunks >= 8) ks = std::max(ks, 2);            // measured: two K halves beat one pass even at 500 row blocks (lm_head 98.6 -> 83.7 us)
    ks = std::max(1, std::min(ks, n_chunks / 2));
    if (const char *e = getenv("EB_MMQ32_KS")) ks = std::max(1, std::min(std::min(8, n_chunks), atoi(e)));   // tuning sweeps
    int cps = (int)cdiv(n_chunks, ks); ks = (int)cdiv(n_chunks, cps);
    *ks_out = ks;
    WPlanes wp{(const uint8_t *)W, N * nbr};
    return type == EB_Q8_0 ? dispatch32<1>(wp, N, nbr, a, planes, ks, cps, st) : dispatch32<0>(wp, N, nbr, a, planes, ks, cps, st);
}

// dynamic-smem opt-in for every instantiation, once per process (never inside a graph capture)
template <int FMT, int NT> static int init32() { EB_CUDA(cudaFuncSetAttribute(mmq32_kernel<FMT, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem32<FMT, NT>())); return EB_OK; }
int eb_mmq32_init() {
    int rc;
    if ((rc = init32<0, 1>()) || (rc = init32<0, 2>()) || (rc = init32<0, 3>()) || (rc = init32<0, 4>()) || (rc = init32<0, 5>()) || (rc = init32<0, 7>()) ||
        (rc = init32<0, 9>()) || (rc = init32<0, 12>()) || (rc = init32<0, 16>())) return rc;
    if ((rc = init32<1, 1>()) || (rc = init32<1, 2>()) || (rc = init32<1, 3>()) || (rc = init32<1, 4>()) || (rc = init32<1, 5>()) || (rc = init32<1, 7>()) ||
        (rc = init32<1, 9>()) || (rc = init32<1, 12>()) || (rc = init32<1, 16>())) return rc;
    return EB_OK;
}
