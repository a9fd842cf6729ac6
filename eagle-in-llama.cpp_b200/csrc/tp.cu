// tp.cu — tensor-parallel communicator of the target's verify forward: a one-shot all-reduce over NVLink / NVSwitch PEER MEMORY,
// fused with the reduction of the GEMM's split-K planes.  Replaces the reference's row-split gather (cudaMemcpyPeerAsync of partial
// rows to the main device + events, R/ggml/src/ggml-cuda/ggml-cuda.cu:1530-1663) for the K-split wo / ffn_down outputs.
//
// One process per GPU.  Every rank allocates ONE communication buffer, exports it with cudaIpcGetMemHandle; the host program carries
// the 64-byte handles between the processes (bench.py / the tests use torch.distributed.all_gather_object) and each rank maps its
// peers' buffers with cudaIpcOpenMemHandle.  After that an all-reduce is a single kernel launch per rank, no host involvement, so it
// can sit inside a captured CUDA graph:
//
//   CTA c of rank r  (1) sums this rank's ks split-K planes over its slice of the [T][E] result and PUSHES the partial into slot r of
//                        data[p] of EVERY rank (posted NVLink stores: no round trip; own slot = a local store),
//                    (2) publishes: __threadfence_system(), then st.release.sys of the call number into flags[p][r][c] of every rank,
//                    (3) waits until its own flags[p][*][c] carry that call number (ld.acquire.sys),
//                    (4) adds the size slots of slice c from its OWN memory, IN RANK ORDER, so all ranks produce the same bits, and
//                        writes its private output.
//   (A pull design — every rank reading its peers' partials with uncached loads — measured ~20 us per call at 2 GPUs: each load is an
//   NVLink round trip.  Pushing turns the transfer into posted writes that overlap each other.)
//   p = call number & 1: a rank can only start call n + 1 after finishing call n, and nobody can finish call n + 1 without every
//   rank's flag for it, so a slot written in call n + 2 is no longer being read from call n.
// Slices are matched CTA to CTA, so no grid-wide barrier is needed; all CTAs of a launch are co-resident (grid <= SM count), which
// the spin in (3) requires.  A spin that sees no progress for ~2 s gives up and raises the error word (a hung GPU box is worse than
// a wrong number, and the host checks the word).
#include "common.cuh"
#include <algorithm>
#include <cstring>

namespace {
constexpr int TP_MAX_RANKS = 8, TP_CTAS = 128, TP_THREADS = 512;

struct TpPeers { float *data[TP_MAX_RANKS]; unsigned *flags[TP_MAX_RANKS]; };

__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) { unsigned v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

// data: [2][size][cap] floats (slot r = rank r's partial); flags: [2][TP_MAX_RANKS][TP_CTAS]; seq: [TP_CTAS] call counters (own
// memory, one writer each); err: 1 word
__global__ void __launch_bounds__(TP_THREADS) tp_allreduce_kernel(const float *__restrict__ planes, int ks, int64_t n, float *__restrict__ out,
                                                                  TpPeers peers, int rank, int size, int64_t cap, unsigned *seq, unsigned *err) {
    __shared__ unsigned s_call;
    const int c = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) s_call = seq[c] + 1;
    __syncthreads();
    const unsigned call = s_call; const int p = call & 1;
    const int64_t n4 = n >> 2, per = (n4 + gridDim.x - 1) / gridDim.x, lo = c * per, hi = min(n4, lo + per);
    const int64_t slot = ((int64_t)p * size + rank) * cap;              // where rank `rank` writes in anybody's data
    // (1) this rank's partial: sum of its split-K planes in plane order (as the consumer kernels of the single-GPU path do), pushed
    for (int64_t i = lo + tid; i < hi; i += TP_THREADS) {
        float4 a = *(const float4 *)(planes + 4 * i);
        for (int s = 1; s < ks; s++) { const float4 b = *(const float4 *)(planes + (int64_t)s * n + 4 * i); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
        for (int r = 0; r < size; r++) *(float4 *)(peers.data[r] + slot + 4 * i) = a;
    }
    __syncthreads();
    // (2) publish slice c of this rank to everybody (own flag included: keeps step 3 uniform)
    if (tid < size) {
        __threadfence_system();
        st_release_sys(peers.flags[tid] + ((int64_t)p * TP_MAX_RANKS + rank) * TP_CTAS + c, call);
    }
    // (3) wait for slice c of every rank
    if (tid < size) {
        const unsigned *f = peers.flags[rank] + ((int64_t)p * TP_MAX_RANKS + tid) * TP_CTAS + c;
        const long long t0 = clock64();
        while (ld_acquire_sys(f) != call) {
            if (clock64() - t0 > 4000000000ll) { atomicExch(err, 1u); break; }
        }
    }
    __syncthreads();
    // (4) sum over ranks in rank order from local memory (written by the peers: read past L1): identical bits on every rank
    const float *mine = peers.data[rank] + (int64_t)p * size * cap;
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
    float *data = nullptr; unsigned *flags = nullptr, *seq = nullptr, *err = nullptr;
    TpPeers peers{}; void *opened[TP_MAX_RANKS] = {};
};
static size_t tp_data_bytes(int64_t cap, int size) { return (size_t)2 * size * cap * 4; }
static size_t tp_flag_bytes() { return (size_t)2 * TP_MAX_RANKS * TP_CTAS * 4; }
static void tp_carve(uint8_t *base, int64_t cap, int size, float **data, unsigned **flags) { *data = (float *)base; *flags = (unsigned *)(base + tp_data_bytes(cap, size)); }

// max_elems: largest [T][E] result that will be reduced (floats, a multiple of 4)
extern "C" eb_tp *eb_tp_create(int rank, int size, int64_t max_elems) {
    if (size < 1 || size > TP_MAX_RANKS || rank < 0 || rank >= size || max_elems < 4 || (max_elems & 3)) { eb_set_err(EB_ERR_ARG, "tp_create: rank %d of %d, %lld elems", rank, size, (long long)max_elems); return nullptr; }
    eb_tp *t = new eb_tp(); t->rank = rank; t->size = size; t->cap = max_elems;
    t->bytes = tp_data_bytes(max_elems, size) + tp_flag_bytes() + TP_CTAS * 4 + 256;
    if (cudaMalloc((void **)&t->base, t->bytes) != cudaSuccess || cudaMemset(t->base, 0, t->bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        eb_set_err(EB_ERR_OOM, "tp_create: %s", cudaGetErrorString(cudaGetLastError())); delete t; return nullptr;
    }
    tp_carve(t->base, t->cap, t->size, &t->data, &t->flags);
    t->seq = (unsigned *)(t->base + tp_data_bytes(max_elems, size) + tp_flag_bytes()); t->err = t->seq + TP_CTAS;
    t->peers.data[rank] = t->data; t->peers.flags[rank] = t->flags;
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
    tp_carve((uint8_t *)p, t->cap, t->size, &t->peers.data[peer_rank], &t->peers.flags[peer_rank]);
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
int eb_tp_allreduce_planes(eb_tp *t, const float *planes, int ks, int64_t n, float *out, cudaStream_t st) {
    if (n > t->cap || (n & 3) || ks < 1) return eb_set_err(EB_ERR_ARG, "tp_allreduce: n=%lld (cap %lld) ks=%d", (long long)n, (long long)t->cap, ks);
    for (int r = 0; r < t->size; r++) if (!t->peers.data[r]) return eb_set_err(EB_ERR_ARG, "tp_allreduce: peer %d not opened", r);
    const int grid = std::min(TP_CTAS, eb_sm_count());
    tp_allreduce_kernel<<<grid, TP_THREADS, 0, st>>>(planes, ks, n, out, t->peers, t->rank, t->size, t->cap, t->seq, t->err);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
// operator form for tests and for hosts that drive their own layers: out[i] = sum over ranks of sum over planes
extern "C" int eb_tp_allreduce(eb_tp *t, const float *d_planes, int ks, int64_t n, float *d_out, void *stream) {
    return eb_tp_allreduce_planes(t, d_planes, ks, n, d_out, (cudaStream_t)stream);
}
