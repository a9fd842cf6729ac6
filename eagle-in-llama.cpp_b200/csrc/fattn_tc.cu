// fattn_tc.cu — tree-masked flash attention on the 5th-gen tensor cores (tcgen05, accumulators in TMEM, K/V tiles by TMA).
// Same contract as fattn.cu (replaces flash_attn_ext_f16, R/ggml/src/ggml-cuda/fattn-mma-f16.cuh:806; CPU
// ggml_compute_forward_flash_attn_ext_f16, R/ggml/src/ggml-cpu/ggml-cpu.c:10999-11212) for the compact tree mask.
//
// One CTA = one head x one KV split; all T <= 128 query rows ride in the M = 128 dimension of the MMA, so every K/V byte
// is read from HBM exactly once per layer.  Per 128-cell tile:
//     S = Q K^T    tcgen05.mma SS   (Q, K K-major 128B-swizzled smem)           -> TMEM S[b]  (128 lanes x 128 f32 columns)
//     softmax      one thread per query row: tcgen05.ld S row -> registers, mask, running max (rescale of O only when the
//                  max grows by more than 2^8), P = exp2 as TWO f16 addends (hi = f16(P), lo = f16(P - hi): ~22 significant
//                  bits, i.e. the f32 P of the CPU kernel, R/ggml/src/ggml-cpu/ggml-cpu.c:11135-11170) -> tcgen05.st over S[b]
//                  (hi pairs in columns 0..63, lo pairs in columns 64..127)
//     O += P V     tcgen05.mma TS x 2 (P_hi, then P_lo on the same accumulator; V MN-major smem: the cache layout [cell][dim]
//                  needs no transpose).  Products f16 x f16 are exact, accumulation is f32: the only difference left against
//                  the reference is its f16 V accumulator (documented deviation)
// S is double-buffered so QK^T of tile i+2 and PV of tile i run on the tensor pipe while the softmax warps work on tile i+1.
// Roles: warp 0 TMA producer (3-stage K ring + 3-stage V ring, 32 KB tiles), warp 1 MMA issuer, warp 2 TMEM allocator,
// warps 4-7 softmax / correction / epilogue (warp 4+q owns TMEM lanes 32q..32q+31 = query rows).
#include "common.cuh"
#include "spec.cuh"
#include "tc.cuh"
#include <cstring>
#include <cstdlib>

namespace {
using namespace tc;

constexpr int FT_TILE = 128;
constexpr int FT_STAGES = 3;
constexpr uint32_t FT_BOX = 16384;                      // 128 rows x 64 f16
constexpr uint32_t FT_OFF_Q = 0;                        // 2 boxes
constexpr uint32_t FT_STAGE = 2 * FT_BOX;               // one K (or V) tile: dims 0-63 | dims 64-127
constexpr uint32_t FT_OFF_K = 2 * FT_BOX;               // K ring and V ring are independent: K(i) is free once S(i) is done,
constexpr uint32_t FT_OFF_V = FT_OFF_K + FT_STAGES * FT_STAGE;   // V(i) only after PV(i), so K runs further ahead
constexpr uint32_t FT_OFF_BAR = FT_OFF_V + FT_STAGES * FT_STAGE;
constexpr uint32_t FT_SMEM = FT_OFF_BAR + 256 + 1024;   // + alignment slack
enum { FB_Q = 0, FB_K_FULL = 1, FB_K_EMPTY = 4, FB_V_FULL = 7, FB_V_EMPTY = 10, FB_S_FULL = 13, FB_P_READY = 15, FB_O_DONE = 17, FB_COUNT = 18 };
constexpr uint32_t FT_COL_O = 256;                      // TMEM: S[0] 0..127, S[1] 128..255, O 256..383

struct FtArgs {
    const uint32_t *tree_bits; const FwdCtl *ctl;
    const uint32_t *mask_bits; const uint8_t *tile_kind; int mask_words;     // general 0 / -inf mask as bits (operator layer), or null
    int n_prefix, n_kv, n_ctx;
    int T, H, nsplit;
    float scale_log2;
    float *out, *part_o, *part_ml;
    const int *gate;        // launch gate (common.cuh): closed -> the CTA runs the zero-tile path
    long long *prof;        // EB_FA_PROF=1: per-section cycle totals of CTA (0,0)
};

__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_h2(float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *(const uint32_t *)&h; }

// visibility of the 32 cells [cb, cb+32) for one row: committed prefix = all visible, tree cell j = bit j of tb
__device__ __forceinline__ uint32_t vis_word(int cb, int n_prefix, const uint32_t (&tb)[4]) {
    const int lo = cb - n_prefix;
    if (lo <= -32) return 0xffffffffu;
    if (lo < 0) return ((1u << (-lo)) - 1u) | (tb[0] << (-lo));
    const int w = lo >> 5, s = lo & 31;
    const uint32_t a = w == 0 ? tb[0] : w == 1 ? tb[1] : w == 2 ? tb[2] : w == 3 ? tb[3] : 0u;
    const uint32_t b = w == 0 ? tb[1] : w == 1 ? tb[2] : w == 2 ? tb[3] : 0u;
    return __funnelshift_r(a, b, s);
}

__global__ void __launch_bounds__(256, 1)
fattn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, FtArgs p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = sbase + FT_OFF_BAR, slot = bars + FB_COUNT * 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.x, split = blockIdx.y;
    const int nsw = (p.T + 31) >> 5;                                         // softmax warps with at least one live row
    if (p.prof && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 128) p.prof[10] = clock64();

    if (threadIdx.x == 0) {
        mbar_init(bars + FB_Q * 8, 1);
        for (int s = 0; s < FT_STAGES; s++) {
            mbar_init(bars + (FB_K_FULL + s) * 8, 1); mbar_init(bars + (FB_K_EMPTY + s) * 8, 1);
            mbar_init(bars + (FB_V_FULL + s) * 8, 1); mbar_init(bars + (FB_V_EMPTY + s) * 8, 1);
        }
        for (int b = 0; b < 2; b++) { mbar_init(bars + (FB_S_FULL + b) * 8, 1); mbar_init(bars + (FB_P_READY + b) * 8, nsw); }
        mbar_init(bars + FB_O_DONE * 8, 1);
        fence_barrier_init();
        tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    }
    if (warp == 2) tmem_alloc(slot, 512);
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    // everything above overlapped the predecessor's tail (PDL); lengths, Q, K/V and the tree bits are earlier kernels' outputs
    pdl_trigger();
    pdl_wait();
    if (p.prof && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 128) p.prof[11] = clock64();
    int n_prefix = p.n_prefix, n_kv = p.n_kv;
    if (p.ctl) { n_prefix = p.ctl->n_prefix; n_kv = min(p.n_ctx, n_prefix + p.ctl->n_tree); }
    if (gate_closed(p.gate)) n_kv = 0;
    const int n_tiles = (n_kv + FT_TILE - 1) / FT_TILE, tps = (n_tiles + p.nsplit - 1) / p.nsplit;
    const int t0 = split * tps, n = max(0, min(n_tiles, t0 + tps) - t0);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (n > 0 && elect_one()) {
            mbar_expect_tx(bars + FB_Q * 8, 2 * FT_BOX);
            tma_load_2d(sbase + FT_OFF_Q, &tmQ, h * 128, 0, bars + FB_Q * 8);
            tma_load_2d(sbase + FT_OFF_Q + FT_BOX, &tmQ, h * 128 + 64, 0, bars + FB_Q * 8);
            for (int i = 0; i < n; i++) {
                const int s = i % FT_STAGES, u = i / FT_STAGES;
                const int cell0 = (t0 + i) * FT_TILE;
                if (u > 0) mbar_wait(bars + (FB_K_EMPTY + s) * 8, (u - 1) & 1);
                const uint32_t kf = bars + (FB_K_FULL + s) * 8, kd = sbase + FT_OFF_K + s * FT_STAGE;
                mbar_expect_tx(kf, FT_STAGE);
                tma_load_2d(kd, &tmK, h * 128, cell0, kf);
                tma_load_2d(kd + FT_BOX, &tmK, h * 128 + 64, cell0, kf);
                if (u > 0) mbar_wait(bars + (FB_V_EMPTY + s) * 8, (u - 1) & 1);
                const uint32_t vf = bars + (FB_V_FULL + s) * 8, vd = sbase + FT_OFF_V + s * FT_STAGE;
                mbar_expect_tx(vf, FT_STAGE);
                tma_load_2d(vd, &tmV, h * 128, cell0, vf);
                tma_load_2d(vd + FT_BOX, &tmV, h * 128 + 64, cell0, vf);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (warp-uniform loop, one elected lane issues)
        const uint32_t id_s = idesc_f16_m128(128), id_o = id_s | TC_IDESC_B_MN;
        const uint64_t q_lo = smem_desc_sw128(sbase + FT_OFF_Q), q_hi = smem_desc_sw128(sbase + FT_OFF_Q + FT_BOX);
        auto issue_s = [&](int i) {
            const int s = i % FT_STAGES;
            mbar_wait(bars + (FB_K_FULL + s) * 8, (i / FT_STAGES) & 1);
            fence_after();
            const uint32_t kb = sbase + FT_OFF_K + s * FT_STAGE;
            if (elect_one()) {
                mma_f16_ss_x8(tmem + (i & 1) * 128, q_lo, q_hi, smem_desc_sw128(kb), smem_desc_sw128(kb + FT_BOX), id_s, 0);
                mma_commit(bars + (FB_K_EMPTY + s) * 8);
                mma_commit(bars + (FB_S_FULL + (i & 1)) * 8);
            }
            __syncwarp();
        };
        if (n > 0) {
            mbar_wait(bars + FB_Q * 8, 0);
            issue_s(0);
            if (n > 1) issue_s(1);
            for (int i = 0; i < n; i++) {
                const int b = i & 1, s = i % FT_STAGES;
                mbar_wait(bars + (FB_V_FULL + s) * 8, (i / FT_STAGES) & 1);
                mbar_wait(bars + (FB_P_READY + b) * 8, (i >> 1) & 1);
                fence_after();
                const uint32_t vb = sbase + FT_OFF_V + s * FT_STAGE;
                if (elect_one()) {
                    mma_f16_ts_x8_rows(tmem + FT_COL_O, tmem + b * 128, smem_desc_sw128_mn(vb, FT_BOX), id_o, i > 0);
                    mma_f16_ts_x8_rows(tmem + FT_COL_O, tmem + b * 128 + 64, smem_desc_sw128_mn(vb, FT_BOX), id_o, 1);      // P_lo
                    mma_commit(bars + (FB_V_EMPTY + s) * 8);
                    mma_commit(bars + FB_O_DONE * 8);
                }
                __syncwarp();
                if (i + 2 < n) issue_s(i + 2);       // overwrites S[b] = P(i): tcgen05.mma of one thread execute in issue order
            }
        }
    } else if (warp >= 4 && (warp - 4) * 32 < p.T) {
        // ------------------------------------------------------------------ softmax / correction / epilogue: one thread per query row
        const int q = warp - 4, row = q * 32 + lane;
        const bool live = row < p.T;
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
        uint32_t tb[4] = {0, 0, 0, 0};
        if (live && p.tree_bits) { const uint4 w = *(const uint4 *)(p.tree_bits + row * 4); tb[0] = w.x; tb[1] = w.y; tb[2] = w.z; tb[3] = w.w; }
        const float c = p.scale_log2;
        float m = -INFINITY, l = 0.f;                                       // m in raw score units (before the scale)
        const bool prof = p.prof && blockIdx.x == 0 && blockIdx.y == 0 && warp == 4 && lane == 0;
        long long sec[6] = {0, 0, 0, 0, 0, 0}, ck = prof ? clock64() : 0; const long long t_start = ck;
#define FA_CK(k) do { if (prof) { const long long n_ = clock64(); sec[k] += n_ - ck; ck = n_; } } while (0)
        for (int i = 0; i < n; i++) {
            const int b = i & 1;
            mbar_wait(bars + (FB_S_FULL + b) * 8, (i >> 1) & 1);
            FA_CK(0);
            fence_after();
            const uint32_t ts = lane_base + b * 128;
            uint32_t sr[4][32];
            tmem_ld32(ts, sr[0]); tmem_ld32(ts + 32, sr[1]); tmem_ld32(ts + 64, sr[2]); tmem_ld32(ts + 96, sr[3]);
            wait_ld();
            FA_CK(1);
            const int cb0 = (t0 + i) * FT_TILE;
            if (p.mask_bits) {                                               // any 0 / -inf mask: one bit per (row, cell), tiles that hide nothing are flagged 0
                if (p.tile_kind[t0 + i]) {
                    uint4 w = make_uint4(0u, 0u, 0u, 0u);
                    if (live) w = *(const uint4 *)(p.mask_bits + (size_t)row * p.mask_words + (cb0 >> 5));
                    const uint32_t vws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                    for (int ch = 0; ch < 4; ch++) {
                        const uint32_t vw = vws[ch];
#pragma unroll
                        for (int k = 0; k < 32; k++) if (!((vw >> k) & 1u)) sr[ch][k] = 0xff800000u;   // -inf
                    }
                }
            } else if (cb0 + FT_TILE > n_prefix) {                           // tile touches the tree cells (or the cache tail)
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    const uint32_t vw = vis_word(cb0 + ch * 32, n_prefix, tb);
#pragma unroll
                    for (int k = 0; k < 32; k++) if (!((vw >> k) & 1u)) sr[ch][k] = 0xff800000u;   // -inf
                }
            }
            // 8 independent max chains (one dependent chain of 128 FMNMX costs ~600 clk on a single warp per scheduler)
            float tmx[8];
#pragma unroll
            for (int a = 0; a < 8; a++) tmx[a] = -INFINITY;
#pragma unroll
            for (int ch = 0; ch < 4; ch++)
#pragma unroll
                for (int k = 0; k < 32; k++) tmx[k & 7] = fmaxf(tmx[k & 7], __uint_as_float(sr[ch][k]));
            const float tm = fmaxf(fmaxf(fmaxf(tmx[0], tmx[1]), fmaxf(tmx[2], tmx[3])), fmaxf(fmaxf(tmx[4], tmx[5]), fmaxf(tmx[6], tmx[7])));
            FA_CK(2);
            // running max: O is rescaled only when the max moves by more than 2^8 (P stays within f16 range, the row sum is f32)
            const bool grow = m != -INFINITY && (tm - m) * c > 8.f;
            if (i > 0 && __any_sync(0xffffffffu, grow)) {
                mbar_wait(bars + FB_O_DONE * 8, (i - 1) & 1);
                fence_after();
                const float mn = fmaxf(m, tm);
                const float f = m == -INFINITY ? 1.f : ex2((m - mn) * c);
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    uint32_t o[32];
                    tmem_ld32(lane_base + FT_COL_O + ch * 32, o);
                    wait_ld();
#pragma unroll
                    for (int k = 0; k < 32; k++) o[k] = __float_as_uint(__uint_as_float(o[k]) * f);
                    tmem_st32(lane_base + FT_COL_O + ch * 32, o);
                }
                l *= f; m = mn;
            } else if (m == -INFINITY) {
                m = tm;
            }
            FA_CK(3);
            const float base = m == -INFINITY ? 0.f : m * c;
            float rs[4] = {0.f, 0.f, 0.f, 0.f};                                // independent partial row sums (ILP)
            // 32 cells at a time: their hi and lo pairs live together
            auto exp_chunk = [&](int ch, uint32_t (&ph)[16], uint32_t (&pl)[16]) {
#pragma unroll
                for (int k = 0; k < 32; k += 2) {
                    const float p0 = ex2(fmaf(__uint_as_float(sr[ch][k]), c, -base)), p1 = ex2(fmaf(__uint_as_float(sr[ch][k + 1]), c, -base));
                    rs[(k >> 1) & 3] += p0 + p1;
                    const __half2 hi = __floats2half2_rn(p0, p1);
                    const float2 hf2 = __half22float2(hi);
                    ph[k >> 1] = *(const uint32_t *)&hi;
                    pl[k >> 1] = pack_h2(p0 - hf2.x, p1 - hf2.y);
                }
            };
            {
                uint32_t ph0[16], pl0[16], ph1[16], pl1[16];
                exp_chunk(0, ph0, pl0); exp_chunk(1, ph1, pl1);
                // P(i) is written only after PV(i-1) has completed, for EVERY warp (not just the ones that rescaled above).  Measured
                // (tools/diag_attn.py, 4K cells, score std 3): with 16 PV MMAs per tile the softmax of a warp that did not rescale
                // can reach its tcgen05.st while PV(i-1) is still executing, and the outputs then differ from run to run (1e-2 ..
                // 3e-1 off); with this wait they are deterministic and within 4e-6 of float64.  Half of the exponentials are
                // computed before the wait so that it overlaps the tensor pipe.
                if (i > 0) { mbar_wait(bars + FB_O_DONE * 8, (i - 1) & 1); fence_after(); }
                tmem_st16(ts, ph0); tmem_st16(ts + 64, pl0); tmem_st16(ts + 16, ph1); tmem_st16(ts + 64 + 16, pl1);
            }
#pragma unroll
            for (int ch = 2; ch < 4; ch++) {
                uint32_t ph[16], pl[16];
                exp_chunk(ch, ph, pl);
                tmem_st16(ts + ch * 16, ph); tmem_st16(ts + 64 + ch * 16, pl);
            }
            l += (rs[0] + rs[1]) + (rs[2] + rs[3]);
            FA_CK(4);
            wait_st();
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + (FB_P_READY + b) * 8);
            FA_CK(5);
        }
        if (prof) { for (int k = 0; k < 6; k++) p.prof[k] = sec[k]; p.prof[6] = n; p.prof[7] = clock64() - t_start; }
        // epilogue: O / l -> output (or the split's partial + (max, sum) in the log2 domain)
        float *dst = nullptr;
        if (live) {
            if (p.nsplit == 1) dst = p.out + ((int64_t)row * p.H + h) * 128;
            else {
                const int64_t pb = ((int64_t)h * p.nsplit + split) * p.T + row;
                dst = p.part_o + pb * 128;
                p.part_ml[pb * 2] = m == -INFINITY ? -INFINITY : m * c; p.part_ml[pb * 2 + 1] = l;
            }
        }
        if (n > 0) { mbar_wait(bars + FB_O_DONE * 8, (n - 1) & 1); fence_after(); }
        if (prof) p.prof[8] = clock64() - t_start;
        const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll
        for (int ch = 0; ch < 4; ch++) {
            uint32_t o[32];
            if (n > 0) { tmem_ld32(lane_base + FT_COL_O + ch * 32, o); wait_ld(); }
#pragma unroll
            for (int k = 0; k < 32; k += 4) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n > 0) v = make_float4(__uint_as_float(o[k]) * inv, __uint_as_float(o[k + 1]) * inv, __uint_as_float(o[k + 2]) * inv, __uint_as_float(o[k + 3]) * inv);
                if (live) *(float4 *)(dst + ch * 32 + k) = v;
            }
        }
    }
    if (p.prof && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 128) p.prof[9] = clock64();
#undef FA_CK
    fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem, 512);
}

__global__ void q_to_f16_kernel(const float *__restrict__ q, __half *__restrict__ q16, int64_t n) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 v = *(const float4 *)(q + i);
    const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    *(uint2 *)(q16 + i) = make_uint2(*(const uint32_t *)&a, *(const uint32_t *)&b);
}

}  // namespace

// number of KV splits: one CTA per SM at most (the CTA owns the whole shared memory)
int eb_fattn_tc_nsplit(int H, int n_tiles) {
    int ns = std::max(1, eb_sm_count() / H);
    if (ns > n_tiles) ns = std::max(1, n_tiles);
    return ns;
}
int eb_fattn_tc_supported(int T, int D) {
    static const int off = getenv("EB_NO_TC_FA") ? 1 : 0;
    return !off && eb_tc_ready() && D == 128 && T >= 1 && T <= 128;
}
size_t eb_fattn_tc_ws_size(int T, int H, int n_kv) {
    const int ns = eb_fattn_tc_nsplit(H, (int)cdiv(n_kv, FT_TILE));
    return (size_t)H * ns * T * (128 + 2) * 4 + (size_t)T * H * 128 * 2 + 512;
}
void eb_fattn_combine(const float *part_o, const float *part_ml, float *out, int T, int H, int nsplit, cudaStream_t st);
void eb_fattn_combine_quant(const float *part_o, const float *part_ml, int T, int H, int nsplit, const ActQ &a, cudaStream_t st);

// q16: f16 copy of the (rotated) queries [T][H*128]; k, v: f16 cells with kv_stride elements between cells.
// ctl != null: lengths are read on the device (engine); otherwise n_prefix / n_kv are used.
int eb_fattn_tc(const __half *q16, const void *k, const void *v, int64_t kv_stride, int64_t n_cells, const uint32_t *tree_bits, const FwdCtl *ctl,
                int n_prefix, int n_kv, int T, int H, float scale, float *out, void *ws, cudaStream_t st, const ActQ *aq, int *quantized,
                const uint32_t *mask_bits, int mask_words, const uint8_t *tile_kind) {
    FtArgs p;
    if (quantized) *quantized = 0;
    p.tree_bits = tree_bits; p.ctl = ctl; p.n_prefix = n_prefix; p.n_kv = n_kv; p.n_ctx = (int)n_cells;
    p.mask_bits = mask_bits; p.mask_words = mask_words; p.tile_kind = tile_kind;
    if (!mask_bits && !tree_bits) return eb_set_err(EB_ERR_ARG, "flash_attn: no mask form given");
    p.T = T; p.H = H; p.scale_log2 = scale * 1.44269504f;
    const int n_tiles_max = (int)cdiv(n_cells, FT_TILE);     // the split count depends on the buffer size only (workspace bound)
    p.nsplit = eb_fattn_tc_nsplit(H, n_tiles_max);
    p.out = out; p.part_o = (float *)ws; p.part_ml = p.part_o + (size_t)H * p.nsplit * T * 128;
    if (p.nsplit > 1 && !ws) return eb_set_err(EB_ERR_ARG, "flash_attn: workspace required");
    static long long *prof_buf = nullptr; static const int want_prof = getenv("EB_FA_PROF") ? 1 : 0;
    if (want_prof && !prof_buf) { cudaMalloc((void **)&prof_buf, 64 * 8); cudaMemset(prof_buf, 0, 64 * 8); }
    p.prof = want_prof ? prof_buf : nullptr; p.gate = eb_gate();
    CUtensorMap tmQ, tmK, tmV;
    int rc;
    if ((rc = eb_tc_map_2d_f16(&tmQ, q16, (uint64_t)H * 128, (uint64_t)T, (uint64_t)H * 128 * 2, 64, 128))) return rc;
    if ((rc = eb_tc_map_2d_f16(&tmK, k, (uint64_t)H * 128, (uint64_t)n_cells, (uint64_t)kv_stride * 2, 64, 128))) return rc;
    if ((rc = eb_tc_map_2d_f16(&tmV, v, (uint64_t)H * 128, (uint64_t)n_cells, (uint64_t)kv_stride * 2, 64, 128))) return rc;
    eb_launch_pdl_c(EB_PDL_ATTN, fattn_tc_kernel, dim3(H, p.nsplit), dim3(256), FT_SMEM, st, tmQ, tmK, tmV, p);
    EB_CHECK_LAUNCH();
    if (want_prof) {
        long long h[16]; cudaStreamSynchronize(st); cudaMemcpy(h, prof_buf, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[fa prof] T=%d H=%d nsplit=%d tiles=%lld | softmax warp: s_full wait %lld | ld %lld | mask+max %lld | rescale %lld | exp+pack %lld | st+arrive %lld | loop %lld | +o_done %lld | setup %lld | exit-entry %lld\n",
                T, H, p.nsplit, h[6], h[0], h[1], h[2], h[3], h[4], h[5], h[7], h[8], h[11] - h[10], h[9] - h[10]);
    }
    if (p.nsplit > 1) {
        // the caller's wo GEMM wants the Q8_K grid: merge the splits and quantise in one pass, skipping the f32 output
#ifdef EB_TIMING_KNOBS
        static const char *skip_ = getenv("EB_SKIP");
#else
        const char *skip_ = nullptr;                                                     // the shipped library has no result-altering switch
#endif
        if (skip_ && strstr(skip_, "combine")) { if (quantized) *quantized = 1; }      // timing experiment only (-DEB_TIMING_KNOBS builds)
        else if (aq && aq->kq && aq->qs && H % 2 == 0 && aq->K == (int64_t)H * 128) { eb_fattn_combine_quant(p.part_o, p.part_ml, T, H, p.nsplit, *aq, st); if (quantized) *quantized = 1; }
        else eb_fattn_combine(p.part_o, p.part_ml, out, T, H, p.nsplit, st);
    }
    EB_CHECK_LAUNCH();
    return EB_OK;
}
// operator-layer entry: f32 queries are rounded to f16 first (GGML_FP32_TO_FP16 of ggml-cpu.c:11115), staged behind the partials in ws
int eb_fattn_tc_f32q(const float *q, const void *k, const void *v, int64_t kv_stride, const uint32_t *tree_bits, int n_prefix, int n_kv,
                     int T, int H, float scale, float *out, void *ws, cudaStream_t st) {
    if (!ws) return eb_set_err(EB_ERR_ARG, "flash_attn: workspace required");
    const int ns = eb_fattn_tc_nsplit(H, (int)cdiv(n_kv, FT_TILE));
    __half *q16 = (__half *)((uint8_t *)ws + (((size_t)H * ns * T * (128 + 2) * 4 + 255) & ~(size_t)255));
    const int64_t nq = (int64_t)T * H * 128;
    q_to_f16_kernel<<<(unsigned)cdiv(nq / 4, 256), 256, 0, st>>>(q, q16, nq);
    EB_CHECK_LAUNCH();
    return eb_fattn_tc(q16, k, v, kv_stride, n_kv, tree_bits, nullptr, n_prefix, std::min(n_kv, n_prefix + T), T, H, scale, out, ws, st, nullptr, nullptr, nullptr, 0, nullptr);
}
// ---- general 0 / -inf masks (the ggml f16 mask of GGML_OP_FLASH_ATTN_EXT, R/ggml/include/ggml.h:1778-1793) as bits
// layout of the bits buffer: uint32 [T][words] (words = 4 per 128-cell tile, bit = visible) | uint8 [n_tiles] tile kinds (0 = every
// row sees every cell of the tile, else 2) | int flag (1 = some mask value is neither 0 nor -inf: the bit form does not apply)
static int mask_words_of(int n_kv) { return 4 * (int)cdiv(n_kv, FT_TILE); }
size_t eb_mask_bits_size_impl(int T, int n_kv) { return (size_t)T * mask_words_of(n_kv) * 4 + (((size_t)cdiv(n_kv, FT_TILE) + 15) & ~(size_t)15) + 16; }
// one CTA per tile, one warp per row at a time: lane l reads cells l, l+32, l+64, l+96 (coalesced), four ballots = the four words
__global__ void __launch_bounds__(128) mask_to_bits_kernel(const __half *__restrict__ mask, int64_t mask_stride, int T, int n_kv, uint32_t *__restrict__ bits, int words,
                                                           uint8_t *__restrict__ kind, int *__restrict__ flag) {
    const int tile = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bool all = true, general = false;
    for (int r = warp; r < T; r += 4) {
        const __half *mr = mask + (size_t)r * mask_stride + (size_t)tile * FT_TILE;
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int c = tile * FT_TILE + 32 * k + lane;
            float v = -INFINITY;
            if (c < n_kv) v = __half2float(mr[32 * k + lane]);
            general |= !(v == 0.f || v == -INFINITY);
            w[k] = __ballot_sync(0xffffffffu, v == 0.f);
            all &= w[k] == 0xffffffffu;
        }
        if (lane == 0) *(uint4 *)(bits + (size_t)r * words + 4 * tile) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    const int every = __syncthreads_and(all ? 1 : 0);
    if (threadIdx.x == 0) kind[tile] = every ? 0 : 2;
    if (__any_sync(0xffffffffu, general) && lane == 0) atomicOr(flag, 1);
}
int eb_mask_to_bits_impl(const void *mask, int64_t mask_stride, int T, int n_kv, void *bits, cudaStream_t st) {
    const int words = mask_words_of(n_kv), n_tiles = (int)cdiv(n_kv, FT_TILE);
    uint8_t *kind = (uint8_t *)bits + (size_t)T * words * 4;
    int *flag = (int *)(kind + ((n_tiles + 15) & ~15));
    EB_CUDA(cudaMemsetAsync(flag, 0, 4, st));
    mask_to_bits_kernel<<<n_tiles, 128, 0, st>>>((const __half *)mask, mask_stride, T, n_kv, (uint32_t *)bits, words, kind, flag);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
const int *eb_mask_bits_flag(const void *bits, int T, int n_kv) {
    const int n_tiles = (int)cdiv(n_kv, FT_TILE);
    return (const int *)((const uint8_t *)bits + (size_t)T * mask_words_of(n_kv) * 4 + ((n_tiles + 15) & ~15));
}
// operator-layer entry with the bit mask: query rows in chunks of one 128-row tile (prompt-sized batches), f32 q -> f16 as above
int eb_fattn_tc_bits(const float *q, const void *k, const void *v, int64_t kv_stride, const void *bits, int T, int H, int n_kv, float scale, float *out, void *ws,
                     cudaStream_t st) {
    if (!ws) return eb_set_err(EB_ERR_ARG, "flash_attn: workspace required");
    const int words = mask_words_of(n_kv);
    const uint8_t *kind = (const uint8_t *)bits + (size_t)T * words * 4;
    const int ns = eb_fattn_tc_nsplit(H, (int)cdiv(n_kv, FT_TILE));
    for (int t0 = 0; t0 < T; t0 += 128) {
        const int tc = std::min(128, T - t0);
        __half *q16 = (__half *)((uint8_t *)ws + (((size_t)H * ns * tc * (128 + 2) * 4 + 255) & ~(size_t)255));
        const int64_t nq = (int64_t)tc * H * 128;
        q_to_f16_kernel<<<(unsigned)cdiv(nq / 4, 256), 256, 0, st>>>(q + (size_t)t0 * H * 128, q16, nq);
        EB_CHECK_LAUNCH();
        int rc = eb_fattn_tc(q16, k, v, kv_stride, n_kv, nullptr, nullptr, n_kv, n_kv, tc, H, scale, out + (size_t)t0 * H * 128, ws, st, nullptr, nullptr,
                             (const uint32_t *)bits + (size_t)t0 * words, words, kind);
        if (rc) return rc;
    }
    return EB_OK;
}
int eb_fattn_tc_init() {
    EB_CUDA(cudaFuncSetAttribute(fattn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FT_SMEM));
    return EB_OK;
}
