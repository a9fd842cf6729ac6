// tp.cu — tensor-parallel communicator of the target's verify forward: a one-shot all-reduce over NVLink / NVSwitch PEER MEMORY,
// fused with the reduction of the GEMM's split-K planes.  Replaces the reference's row-split gather (cudaMemcpyPeerAsync of partial
// rows to the main device + events, R/ggml/src/ggml-cuda/ggml-cuda.cu:1530-1663) for the K-split wo / ffn_down outputs.
//
// One process per GPU.  Every rank allocates ONE communication buffer, exports it with cudaIpcGetMemHandle; the host program carries
// the 64-byte handles between the processes (bench.py / the tests use torch.distributed.all_gather_object) and each rank maps its
// peers' buffers with cudaIpcOpenMemHandle.  After that an all-reduce is a single kernel launch per rank, no host involvement, so it
// can sit inside a captured CUDA graph:
//
//   Flag-in-data ("LL") protocol: every 16-byte store carries two floats and two copies of the call number, {x, call, y, call}; each
//   8-byte half is written atomically, so a reader that sees the call number in both halves has the data that came with it.  No
//   fence, no separate flag, no acknowledgement: the latency of an exchange is one NVLink transit.
//   thread i of rank r  (1) sums this rank's ks split-K planes for its float4 of the [T][E] result and PUSHES it as two such stores into
//                           slot r of slot set p of EVERY rank (posted NVLink stores; own slot = a local store),
//                       (2) spins on the same float4 of every slot of its OWN memory until the call number shows (volatile loads past L1),
//                       (3) adds the slots IN RANK ORDER — every rank produces the same bits — and writes its private output.
//   That form doubles the bytes on the wire, which is the right trade at 2 ranks (13.2 us against 21.0 us for the flag form below,
//   65 x 5120 f32, device time from a graph replay) and the wrong one at 4 and 8 (34.9 / 79.1 us against 25.3 / 36.8 us: each rank
//   then sends 3 or 7 doubled copies).  With more than two ranks the kernel therefore uses plain float4 stores and one flag per
//   (rank, CTA slice): push, st.release.sys of the call number into flags[p][r][c] of every rank (a release store orders the CTA's
//   earlier writes behind it through the barrier before it), ld.acquire.sys spin on the own flags, rank-order sum.
//   Slot sets: the engine alternates p = 0 (after wo) and p = 1 (after ffn_down); the operator entry alternates p = 2, 3.  A rank can
//   only start its next call after finishing this one, and nobody finishes call n + 1 without every rank's data for it, so a set
//   written in call n + 2 is no longer being read from call n.  The call number never repeats within a set (32-bit counter; the
//   buffer starts zeroed and calls start at 1).
// A spin that sees no progress for ~2 s gives up and raises the error word (a hung GPU box is worse than a wrong number, and the host
// checks the word).
#include "common.cuh"
#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace {
constexpr int TP_MAX_RANKS = 8, TP_CTAS = 128, TP_THREADS = 512, TP_SETS = 4;

struct TpPeers { uint4 *ll[TP_MAX_RANKS]; unsigned *flags[TP_MAX_RANKS]; };   // per rank: data region ([TP_SETS][size][cap / 2] 16-byte LL elements, or
                                                                            // [TP_SETS][size][cap] floats in the flag form) and [TP_SETS][TP_MAX_RANKS][TP_CTAS] flags
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) { unsigned v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

__device__ __forceinline__ void st_ll(uint4 *p, float x, float y, unsigned f) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(x)), "r"(f), "r"(__float_as_uint(y)), "r"(f) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint4 *p) {
    uint4 v; asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); return v;
}

// seq: [TP_CTAS] call counters (own memory, one writer each); err: 1 word
__global__ void __launch_bounds__(TP_THREADS) tp_allreduce_kernel(const float *__restrict__ planes, int ks, int64_t n, float *__restrict__ out, int p,
                                                                  TpPeers peers, int rank, int size, int64_t cap, unsigned *seq, unsigned *err) {
    __shared__ unsigned s_call;
    const int c = blockIdx.x, tid = threadIdx.x;
    pdl_trigger();                                                      // the consumer's CTAs may take their places; they wait for this grid to finish
    const int64_t n4 = n >> 2, stride = (int64_t)gridDim.x * TP_THREADS, first = (int64_t)c * TP_THREADS + tid;
    const int64_t set0 = (int64_t)p * size * (cap >> 1);                // LL elements: two per float4
    pdl_wait();                                                         // the planes are the predecessor's output — and so may be seq[] (back-to-back calls)
    if (tid == 0) s_call = seq[c] + 1;
    __syncthreads();
    const unsigned call = s_call;
    // (1) this rank's partial: sum of its split-K planes in plane order (as the consumer kernels of the single-GPU path do), pushed
    for (int64_t i = first; i < n4; i += stride) {
        float4 a = *(const float4 *)(planes + 4 * i);
        for (int s = 1; s < ks; s++) { const float4 b = *(const float4 *)(planes + (int64_t)s * n + 4 * i); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        const int64_t e = set0 + (int64_t)rank * (cap >> 1) + 2 * i;
        for (int r = 0; r < size; r++) { st_ll(peers.ll[r] + e, a.x, a.y, call); st_ll(peers.ll[r] + e + 1, a.z, a.w, call); }
    }
    // (2) + (3) wait for the same float4 of every rank in this rank's own memory, add in rank order
    const uint4 *mine = peers.ll[rank] + set0;
    for (int64_t i = first; i < n4; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < size; r++) {
            const uint4 *q = mine + (int64_t)r * (cap >> 1) + 2 * i;
            uint4 u = ld_ll(q), v = ld_ll(q + 1);
            if (u.y != call || u.w != call || v.y != call || v.w != call) {
                const long long t0 = clock64();
                do {
                    u = ld_ll(q); v = ld_ll(q + 1);
                    if (clock64() - t0 > 4000000000ll) { atomicExch(err, 1u); break; }
                } while (u.y != call || u.w != call || v.y != call || v.w != call);
            }
            const float4 b = make_float4(__uint_as_float(u.x), __uint_as_float(u.z), __uint_as_float(v.x), __uint_as_float(v.z));
            if (r == 0) acc = b; else { acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w; }
        }
        *(float4 *)(out + 4 * i) = acc;
    }
    __syncthreads();
    if (tid == 0) seq[c] = call;
}

// more than two ranks: plain data + one flag per (rank, CTA slice)
__global__ void __launch_bounds__(TP_THREADS) tp_allreduce_flag_kernel(const float *__restrict__ planes, int ks, int64_t n, float *__restrict__ out, int p,
                                                                       TpPeers peers, int rank, int size, int64_t cap, unsigned *seq, unsigned *err) {
    __shared__ unsigned s_call;
    const int c = blockIdx.x, tid = threadIdx.x;
    pdl_trigger();
    const int64_t n4 = n >> 2, per = (n4 + gridDim.x - 1) / gridDim.x, lo = c * per, hi = min(n4, lo + per);
    const int64_t set0 = (int64_t)p * size * cap, slot = set0 + (int64_t)rank * cap;
    pdl_wait();
    if (tid == 0) s_call = seq[c] + 1;
    __syncthreads();
    const unsigned call = s_call;
    for (int64_t i = lo + tid; i < hi; i += TP_THREADS) {
        float4 a = *(const float4 *)(planes + 4 * i);
        for (int s = 1; s < ks; s++) { const float4 b = *(const float4 *)(planes + (int64_t)s * n + 4 * i); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        for (int r = 0; r < size; r++) *(float4 *)((float *)peers.ll[r] + slot + 4 * i) = a;
    }
    __syncthreads();
    if (tid < size) {
        st_release_sys(peers.flags[tid] + ((int64_t)p * TP_MAX_RANKS + rank) * TP_CTAS + c, call);
        const unsigned *f = peers.flags[rank] + ((int64_t)p * TP_MAX_RANKS + tid) * TP_CTAS + c;
        const long long t0 = clock64();
        while (ld_acquire_sys(f) != call) {
            if (clock64() - t0 > 4000000000ll) { atomicExch(err, 1u); break; }
        }
    }
    __syncthreads();
    const float *mine = (const float *)peers.ll[rank] + set0;
    for (int64_t i = lo + tid; i < hi; i += TP_THREADS) {
        float4 a = __ldcg((const float4 *)(mine + 4 * i));
        for (int r = 1; r < size; r++) { const float4 b = __ldcg((const float4 *)(mine + (int64_t)r * cap + 4 * i)); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        *(float4 *)(out + 4 * i) = a;
    }
    if (tid == 0) seq[c] = call;
}
}  // namespace

struct eb_tp {
    int rank = 0, size = 1; int64_t cap = 0;
    uint8_t *base = nullptr; size_t bytes = 0;
    uint4 *ll = nullptr; unsigned *seq = nullptr, *err = nullptr;
    TpPeers peers{}; void *opened[TP_MAX_RANKS] = {};
    unsigned op_calls = 0;                                  // operator-form calls (host side): slot sets 2 and 3 alternate
    uint8_t *peer_base[TP_MAX_RANKS] = {};                  // fused hand-over (TpPush / TpPull): every rank's buffer
};
static size_t tp_data_bytes(int64_t cap, int size) { return (size_t)TP_SETS * size * cap * 8; }      // LL: 8 bytes per float (the flag form uses half)
static size_t tp_flag_bytes() { return (size_t)TP_SETS * TP_MAX_RANKS * TP_CTAS * 4; }
// fused hand-over: [2 sets][size - 1 source ranks][8 planes][cap] 8-byte elements, then 2 x (seq, done) words
constexpr int TP_PLANES = 8;
static size_t tp_push_region(int64_t cap) { return (size_t)TP_PLANES * cap * 8 + 256; }      // planes + the header word
static size_t tp_push_bytes(int64_t cap, int size) { return (size_t)2 * (size - 1) * tp_push_region(cap); }

// max_elems: largest [T][E] result that will be reduced (floats, a multiple of 4)
extern "C" eb_tp *eb_tp_create(int rank, int size, int64_t max_elems) {
    if (size < 1 || size > TP_MAX_RANKS || rank < 0 || rank >= size || max_elems < 4 || (max_elems & 3)) { eb_set_err(EB_ERR_ARG, "tp_create: rank %d of %d, %lld elems", rank, size, (long long)max_elems); return nullptr; }
    eb_tp *t = new eb_tp(); t->rank = rank; t->size = size; t->cap = max_elems;
    t->bytes = tp_data_bytes(max_elems, size) + tp_flag_bytes() + TP_CTAS * 4 + 256 + tp_push_bytes(max_elems, size) + 256;
    if (cudaMalloc((void **)&t->base, t->bytes) != cudaSuccess || cudaMemset(t->base, 0, t->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        eb_set_err(EB_ERR_OOM, "tp_create: %s", cudaGetErrorString(cudaGetLastError())); delete t; return nullptr;
    }
    t->ll = (uint4 *)t->base;
    t->seq = (unsigned *)(t->base + tp_data_bytes(max_elems, size) + tp_flag_bytes()); t->err = t->seq + TP_CTAS;
    t->peers.ll[rank] = t->ll; t->peers.flags[rank] = (unsigned *)(t->base + tp_data_bytes(max_elems, size));
    t->peer_base[rank] = t->base;
    return t;
}
// 64-byte IPC handle of this rank's buffer (cudaIpcMemHandle_t), to be carried to the peers by the host program
extern "C" int eb_tp_handle(eb_tp *t, void *out64) {
    cudaIpcMemHandle_t h; EB_CUDA(cudaIpcGetMemHandle(&h, t->base));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t"); memcpy(out64, &h, 64);
    return EB_OK;
}
extern "C" int eb_tp_open_peer(eb_tp *t, int peer_rank, const void *handle64) {
    if (peer_rank < 0 || peer_rank >= t->size || peer_rank == t->rank) return eb_set_err(EB_ERR_ARG, "tp_open_peer: rank %d", peer_rank);
    cudaIpcMemHandle_t h; memcpy(&h, handle64, 64);
    void *p = nullptr;
    EB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    t->opened[peer_rank] = p;
    t->peers.ll[peer_rank] = (uint4 *)p; t->peers.flags[peer_rank] = (unsigned *)((uint8_t *)p + tp_data_bytes(t->cap, t->size));
    t->peer_base[peer_rank] = (uint8_t *)p;
    return EB_OK;
}
extern "C" void eb_tp_free(eb_tp *t) {
    if (!t) return;
    cudaDeviceSynchronize();
    for (void *p : t->opened) if (p) cudaIpcCloseMemHandle(p);
    cudaFree(t->base); delete t;
}
// 0 = healthy; 1 = a wait for a peer timed out (results since then are garbage)
extern "C" int eb_tp_error(eb_tp *t) {
    unsigned e = 0; if (cudaMemcpy(&e, t->err, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int)e;
}
// set: 0 / 1 = engine (after wo / after ffn_down); -1 = operator form (sets 2 / 3 by a host-side counter)
int eb_tp_allreduce_planes(eb_tp *t, const float *planes, int ks, int64_t n, float *out, cudaStream_t st, int set) {
    if (n > t->cap || (n & 3) || ks < 1 || !out) return eb_set_err(EB_ERR_ARG, "tp_allreduce: n=%lld (cap %lld) ks=%d out=%p", (long long)n, (long long)t->cap, ks, (void *)out);
    if (set > 1) return eb_set_err(EB_ERR_ARG, "tp_allreduce: set %d", set);
    for (int r = 0; r < t->size; r++) if (!t->peers.ll[r]) return eb_set_err(EB_ERR_ARG, "tp_allreduce: peer %d not opened", r);
    const int p = set >= 0 ? set : 2 + (int)(t->op_calls++ & 1);
    const int grid = std::min(TP_CTAS, eb_sm_count());
    static const int force = getenv("EB_TP_FORM") ? atoi(getenv("EB_TP_FORM")) : 0;          // experiments: 1 = flag-in-data, 2 = flag form, whatever the group size
    const bool ll = force ? force == 1 : t->size <= 2;
    if (ll) eb_launch_pdl_c(EB_PDL_RESID, tp_allreduce_kernel, dim3(grid), dim3(TP_THREADS), 0, st, planes, ks, n, out, p, t->peers, t->rank, t->size, t->cap, t->seq, t->err);
    else eb_launch_pdl_c(EB_PDL_RESID, tp_allreduce_flag_kernel, dim3(grid), dim3(TP_THREADS), 0, st, planes, ks, n, out, p, t->peers, t->rank, t->size, t->cap, t->seq, t->err);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
// operator form for tests and for hosts that drive their own layers: out[i] = sum over ranks of sum over planes
extern "C" int eb_tp_allreduce(eb_tp *t, const float *d_planes, int ks, int64_t n, float *d_out, void *stream) {
    if (!d_out) return eb_set_err(EB_ERR_ARG, "tp_allreduce: output required");
    return eb_tp_allreduce_planes(t, d_planes, ks, n, d_out, (cudaStream_t)stream, -1);
}

// device time of one all-reduce: `iters` calls captured in ONE CUDA graph (engine_sets != 0: slot sets 0 / 1 alternate as in the engine,
// else the operator's 2 / 3) and replayed, so that host launch overhead is not what is measured.
// Every rank must call this with the same arguments at the same time (the calls pair up across ranks).
extern "C" int eb_tp_bench(eb_tp *t, const float *d_planes, int ks, int64_t n, float *d_out, int engine_sets, int iters, float *us_per_call) {
    if (iters < 2 || (iters & 1)) return eb_set_err(EB_ERR_ARG, "tp_bench: iters must be even");
    cudaStream_t st; cudaEvent_t e0, e1; cudaGraph_t g = nullptr; cudaGraphExec_t ge = nullptr;
    EB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); EB_CUDA(cudaEventCreate(&e0)); EB_CUDA(cudaEventCreate(&e1));
    int rc = EB_OK;
    for (int i = 0; i < 2 && !rc; i++) rc = eb_tp_allreduce_planes(t, d_planes, ks, n, d_out, st, engine_sets ? (i & 1) : -1);     // warm both sets
    if (!rc) EB_CUDA(cudaStreamSynchronize(st));
    if (!rc) {
        EB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        for (int i = 0; i < iters && !rc; i++) rc = eb_tp_allreduce_planes(t, d_planes, ks, n, d_out, st, engine_sets ? (i & 1) : -1);
        EB_CUDA(cudaStreamEndCapture(st, &g));
    }
    if (!rc) {
        EB_CUDA(cudaGraphInstantiate(&ge, g, 0));
        EB_CUDA(cudaGraphLaunch(ge, st)); EB_CUDA(cudaStreamSynchronize(st));                  // warm replay
        EB_CUDA(cudaEventRecord(e0, st)); EB_CUDA(cudaGraphLaunch(ge, st)); EB_CUDA(cudaEventRecord(e1, st)); EB_CUDA(cudaStreamSynchronize(st));
        float ms = 0; EB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        *us_per_call = ms * 1e3f / iters;
    }
    if (ge) cudaGraphExecDestroy(ge);
    if (g) cudaGraphDestroy(g);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return rc;
}

// ---- fused hand-over (see common.cuh TpPush / TpPull)
static size_t tp_push_off(const eb_tp *t) { return tp_data_bytes(t->cap, t->size) + tp_flag_bytes() + TP_CTAS * 4 + 256; }
static int src_index(int src_rank, int dst_rank) { return src_rank < dst_rank ? src_rank : src_rank - 1; }      // position of src among dst's peers
// Off by default.  Measured on the 13B verify step (bench.py --config 4, same box, back to back): the hand-over fused into the GEMM
// epilogue 4889 us vs 4823 us with the separate all-reduce kernel at 2 ranks, 6755 vs 4568 us at 4 ranks.  The 8-byte remote stores
// (one per element per peer, 40-120 per lane per item) sit on the GEMM's critical path — its epilogue is already the role with the
// least slack — while the all-reduce kernel moves the same data with 16-byte coalesced stores from 65 536 threads
// that have nothing else to do; and either way every layer waits for the slowest rank twice.  EB_TP_FUSE=1 selects this path.
int eb_tp_fused_ok(const eb_tp *t, int ks, int64_t n) {
    static const int on = getenv("EB_TP_FUSE") ? 1 : 0;
    return on && t->size > 1 && ks <= TP_PLANES && n <= t->cap;
}
// where this rank's GEMM epilogue mirrors its planes (set 0 = after wo, 1 = after ffn_down)
TpPush eb_tp_push(const eb_tp *t, int set) {
    TpPush p{}; p.n_peers = t->size - 1; p.hdr = (int64_t)TP_PLANES * t->cap;
    unsigned *words = (unsigned *)(t->base + tp_push_off(t) + tp_push_bytes(t->cap, t->size));
    p.seq = words + 2 * set;
    int j = 0;
    for (int q = 0; q < t->size; q++) {
        if (q == t->rank) continue;
        p.peer[j++] = (unsigned long long *)(t->peer_base[q] + tp_push_off(t) + ((size_t)set * (t->size - 1) + src_index(t->rank, q)) * tp_push_region(t->cap));
    }
    return p;
}
// where the peers' planes arrive in this rank's memory; ks = planes per rank, plane = elements between planes
TpPull eb_tp_pull(const eb_tp *t, int set, int64_t plane) {
    TpPull p{}; p.n_peers = t->size - 1; p.my_pos = t->rank; p.hdr = (int64_t)TP_PLANES * t->cap; p.plane = plane;
    unsigned *words = (unsigned *)(t->base + tp_push_off(t) + tp_push_bytes(t->cap, t->size));
    p.seq = words + 2 * set; p.done = words + 2 * set + 1; p.err = t->err;
    int j = 0;
    for (int r = 0; r < t->size; r++) {
        if (r == t->rank) continue;
        p.src[j++] = (const unsigned long long *)(t->base + tp_push_off(t) + ((size_t)set * (t->size - 1) + src_index(r, t->rank)) * tp_push_region(t->cap));
    }
    return p;
}
