// common.cuh — shared device/host helpers for libeagle_b200.so (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/eagle_b200.h"

#define QK_K 256
#define QK 32

// ---- GGUF block layouts (R/ggml/src/ggml-common.h:161-327); all are only 2-byte aligned
struct __align__(2) block_q4_0 { __half d; uint8_t qs[16]; };                               // 18 B
struct __align__(2) block_q8_0 { __half d; int8_t qs[32]; };                                // 34 B
struct __align__(4) block_q4_K { __half d, dmin; uint8_t scales[12]; uint8_t qs[128]; };    // 144 B
struct __align__(2) block_q6_K { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; __half d; }; // 210 B
static_assert(sizeof(block_q4_0) == 18 && sizeof(block_q8_0) == 34, "block size");
static_assert(sizeof(block_q4_K) == 144 && sizeof(block_q6_K) == 210, "block size");

// ---- DEVICE layout of quantised weights.  Q4_K blocks (144 B) are 16-byte aligned and stay in
// GGUF order.  The 2-byte-aligned formats are split into aligned planes over the whole tensor
// (nb = total blocks) so that every load is a full-width aligned vector load; the total size is
// unchanged and the permutation is lossless (eb_repack / get_tensor invert it):
//   Q6_K: ql [nb][128] | qh [nb][64] | scales [nb][16] | d [nb] f16
//   Q4_0: qs [nb][16]  | d [nb] f16          Q8_0: qs [nb][32] | d [nb] f16
struct WPlanes {
    const uint8_t *base; int64_t nb; int64_t b0 = 0;
    __host__ __device__ void off(int64_t blk) { b0 += blk; }
    __host__ __device__ const uint8_t *q6k_ql() const { return base + b0 * 128; }
    __host__ __device__ const uint8_t *q6k_qh() const { return base + nb * 128 + b0 * 64; }
    __host__ __device__ const int8_t  *q6k_sc() const { return (const int8_t *)(base + nb * 192 + b0 * 16); }
    __host__ __device__ const __half  *q6k_d()  const { return (const __half *)(base + nb * 208) + b0; }
    __host__ __device__ const uint8_t *q40_qs() const { return base + b0 * 16; }
    __host__ __device__ const __half  *q40_d()  const { return (const __half *)(base + nb * 16) + b0; }
    __host__ __device__ const int8_t  *q80_qs() const { return (const int8_t *)(base + b0 * 32); }
    __host__ __device__ const __half  *q80_d()  const { return (const __half *)(base + nb * 32) + b0; }
};

// ---- quantised activations, device layout (planes instead of the CPU's AoS blocks):
//   qs    int8  [T][K]
//   d     f32   [T][K/B]      B = 256 (Q8_K grid) or 32 (Q8_0 grid; value is f16-rounded)
//   bsum  f16   [T][K/16]     Q8_K only: sums of 16 quants (|sum| <= 2032 is exact in f16)
//   h16   f16   [T][K]        Q8_K only: the same quants as f16 (exact) = UMMA B operand (mmq_tc.cu)
//   dT    f32   [K/256][80]   Q8_K only: d transposed (super-block major) for the GEMM epilogue
#define EB_DT_PITCH 80
struct ActQ {
    int8_t *qs; float *d; __half *bsum;
    int64_t K; int T; int kq;   // kq = 1: Q8_K grid, 0: Q8_0 grid
    __half *h16 = nullptr; float *dT = nullptr;
};

// ---- error handling
extern thread_local std::string g_eb_err;
int eb_set_err(int code, const char *fmt, ...);
#define EB_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return eb_set_err(EB_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)
#define EB_CHECK_LAUNCH() do { eb_count_launch(); cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) \
    return eb_set_err(EB_ERR_CUDA, "%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); } while (0)
void eb_count_launch();
long eb_launches_now(); void eb_launches_add(long n);   // kernels recorded in a CUDA graph are counted per replay
int eb_sm_count();
int eb_mmq_init(); int eb_fattn_init(); int eb_spec_init(); int eb_mmq32_init();
int eb_mmq32_supported(int type, int64_t N, int64_t K, int T);
int eb_mmq32_parts(int type, const void *W, int64_t N, int64_t K, const struct ActQ &a, float *planes, int *ks_out, cudaStream_t st);

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- programmatic dependent launch (PDL): a kernel launched through eb_launch_pdl may start while its predecessor in the
// stream is still running; it must execute pdl_wait() before it touches ANY global data written by earlier kernels (or
// writes data they read).  Everything before pdl_wait() — barrier init, TMEM allocation, tensor-map prefetch, the TMA
// stream of the (constant) weights — overlaps the predecessor's tail.  pdl_trigger() lets the successor start early.
// In a kernel launched the ordinary way both are no-ops.  EB_NO_PDL=1 turns the launch attribute off.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
enum { EB_PDL_GEMM = 1, EB_PDL_RESID = 2, EB_PDL_ROPE = 4, EB_PDL_ATTN = 8, EB_PDL_COMBINE = 16, EB_PDL_SWIGLU = 32, EB_PDL_CONCAT = 64, EB_PDL_TOPK = 128, EB_PDL_EXPAND = 256, EB_PDL_ALL = 511 };
int eb_pdl_enabled(int cls = EB_PDL_ALL);
void eb_pdl_scope(int mask);   // host side, per thread: kernel classes whose launches carry the PDL attribute (EB_PDL_ALL by default)
template <typename... KArgs, typename... Args>
static inline cudaError_t eb_launch_pdl_c(int cls, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = eb_pdl_enabled(cls) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- launch gate: the draft levels of a round are captured in a CUDA graph with fixed geometry, but the reference's draft
// loop stops as soon as no branch is drafting (R/examples/speculative/speculative-eagle.cpp:627-641).  A kernel launched
// while a gate is set (host side, eb_gate_set) receives a device pointer and returns right after pdl_wait() when the word
// it points to is 0 (the wait is kept so that the successor's wait still covers every earlier kernel).
const int *eb_gate(); void eb_gate_set(const int *g);
__device__ __forceinline__ bool gate_closed(const int *g) { return g != nullptr && *(const volatile int *)g == 0; }

// ---- device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// 2-byte-aligned 32-bit load (Q4_0/Q8_0/Q6_K blocks are not 4-byte aligned)
__device__ __forceinline__ uint32_t ld_u32_b2(const void *p) {
    const uint16_t *q = (const uint16_t *)p;
    return (uint32_t)q[0] | ((uint32_t)q[1] << 16);
}
// Q4_K 6-bit scale/min unpack (get_scale_min_k4, R/ggml/src/ggml-quants.c:631)
__device__ __forceinline__ void q4k_scale_min(int j, const uint8_t *q, int &sc, int &m) {
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

// ---- internal kernels' host launchers (defined across the .cu files)
int eb_quantize_act(const float *d_x, const ActQ &a, cudaStream_t st);                 // quant.cu
struct SpecState;
int eb_concat_quant(const float *x, const float *hid, int E, const ActQ &a, cudaStream_t st, const SpecState *spec = nullptr, const float *root_feat = nullptr, const float *node_feat = nullptr);
size_t eb_actq_bytes(int kq, int64_t K, int T);
ActQ   eb_actq_carve(void *ws, int kq, int64_t K, int T);
int eb_dequantize_impl(int type, const void *src, int64_t nb_total, int64_t blk0, float *dst, int64_t n, cudaStream_t st);
int eb_repack_impl(int type, void *gguf, void *dev, int64_t nb, int dir, cudaStream_t st);
int eb_synth_impl(int type, void *dst, int64_t n_elems, uint64_t seed, float scale, int norm_weight, cudaStream_t st);
int eb_mmvq(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *y, cudaStream_t st);   // mmvq.cu
int eb_mmq(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *y, float *partial, int ksplit, cudaStream_t st); // mmq.cu
int eb_mmq_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st);
int eb_fused_init();
int eb_mmq_tc_init();                                                                  // mmq_tc.cu (tcgen05 path)
int eb_mmq_tc_supported(int type, int64_t N, int64_t K, int T);
// Tensor-parallel hand-over fused into the GEMM (tp.cu, mmq_tc.cu, fused.cu): the epilogue of a K-split GEMM (wo, ffn_down) mirrors
// its split-K planes, as 8-byte {value, call number} pairs, straight into every peer's memory while it writes them locally; the
// consumer kernel (residual add + RMS norm + quantisation) adds the local planes and, rank by rank, the peers' planes as they arrive
// (an element whose upper word is the current call number has arrived).  No all-reduce kernel, no flag, no fence: the transfer
// overlaps the GEMM's tail.  seq = this rank's call counter of the slot set (last completed call).  Shards may differ in shape: each
// pushing rank also writes {its split-K plane count, call number} into a header word behind its planes (EB_TP_HDR elements in).
constexpr int EB_TP_MAX_PEERS = 7;
struct TpPush { unsigned long long *peer[EB_TP_MAX_PEERS]; int n_peers; int64_t hdr; const unsigned *seq; };
struct TpPull { const unsigned long long *src[EB_TP_MAX_PEERS]; int n_peers, my_pos; int64_t hdr, plane; unsigned *seq, *done, *err; };
int eb_mmq_tc_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st, const TpPush *push = nullptr);
int eb_splitk_reduce(const float *partial, float *y, int64_t n, int ksplit, cudaStream_t st);
int eb_mmq_supported(int type, int64_t N, int64_t K, int T);
// mmq_small.cu: T <= 8 (draft head)
int eb_mmq_small_init(); int eb_mmq_small_supported(int type, int64_t N, int64_t K, int T);
int eb_mmq_small_parts(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *planes, int *ks_out, cudaStream_t st);
int eb_mmq_pick_ksplit(int type, int64_t N, int64_t K, int T);
