// spec.cu — K11 + the fused device-side speculative-step kernels: top-k / argmax over the
// vocabulary, draft-tree expansion, greedy accept walk, KV compaction, feature hand-off.
// Nothing here returns to the host inside a round (the reference does a 32000-float D2H and a
// host partial_sort per node: R/src/llama.cpp:9637, R/src/llama-sampling.cpp:228-302).
#include "common.cuh"
#include <map>
#include <mutex>
#include "spec.cuh"

// ------------------------------------------------------------------ top-k / argmax
// One CTA (1024 threads) per logits row, staged in shared memory.  Each warp extracts the top k of its own 1/32 of the row
// with warp shuffles only (after a pick only the owning lane rescans its 1/32 of the warp's slice), then warp 0 merges the
// 32 sorted candidate lists.  No block-wide barrier inside the k rounds (the first version spent 2 per round: 22 us for
// 8 x 32000).  Ties resolve to the lower token id at every level, so the result equals a stable sort by (-logit, id);
// probs = softmax over the k kept logits (llama_sampler_softmax_impl, R/src/llama-sampling.cpp:130-150).
__device__ __forceinline__ void warp_argmax(float &bv, int &bi) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
}
// logits may be split-K planes [ks][rows][V] of the lm_head GEMM: they are summed (plane order, like splitk_reduce) while staging the row
__global__ void __launch_bounds__(1024) top_k_kernel(const float *__restrict__ logits, int ks, int64_t plane, int V, int k,
                                                     int32_t *__restrict__ ids, float *__restrict__ vals, float *__restrict__ probs, const int *__restrict__ gate) {
    extern __shared__ float row[];
    __shared__ float cv[32 * 16]; __shared__ int ci[32 * 16];
    __shared__ float kv[16]; __shared__ int ki[16];
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *x = logits + (int64_t)r * V;
    // stage the row: all loads of a thread in flight at once (a rolled scalar loop pays the L2 latency 31 times: 11 us)
    if ((V & 3) == 0 && ((uintptr_t)x & 15) == 0) {
        const float4 *x4 = (const float4 *)x; float4 *r4 = (float4 *)row;
        const int n4 = V >> 2;
        for (int i0 = 0; i0 < n4; i0 += 8 * 1024) {
            float4 t[8];
#pragma unroll
            for (int u = 0; u < 8; u++) { const int i = i0 + u * 1024 + tid; if (i < n4) t[u] = x4[i]; }
            for (int s = 1; s < ks; s++) {
                const float4 *p4 = (const float4 *)(x + s * plane);
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    const int i = i0 + u * 1024 + tid;
                    if (i < n4) { const float4 p = p4[i]; t[u].x += p.x; t[u].y += p.y; t[u].z += p.z; t[u].w += p.w; }
                }
            }
#pragma unroll
            for (int u = 0; u < 8; u++) { const int i = i0 + u * 1024 + tid; if (i < n4) r4[i] = t[u]; }
        }
    } else {
        for (int i = tid; i < V; i += 1024) { float v = x[i]; for (int s = 1; s < ks; s++) v += x[s * plane + i]; row[i] = v; }
    }
    __syncthreads();
    const int chunk = ((V + 31) / 32 + 31) & ~31, lo = warp * chunk, hi = min(V, lo + chunk);
    auto scan = [&](float &bv, int &bi) {
        bv = -INFINITY; bi = 0x7fffffff;
        for (int i = lo + lane; i < hi; i += 32) { const float v = row[i]; if (v > bv) { bv = v; bi = i; } }   // ascending i: first max kept
    };
    float mv; int mi;
    scan(mv, mi);
    for (int j = 0; j < k; j++) {
        float bv = mv; int bi = mi;
        warp_argmax(bv, bi);
        if (lane == 0) { cv[warp * 16 + j] = bv; ci[warp * 16 + j] = bi; }
        if (bi != 0x7fffffff && ((bi - lo) & 31) == lane) { row[bi] = -INFINITY; scan(mv, mi); }
        __syncwarp();
    }
    __syncthreads();
    if (warp == 0) {                                    // 32-way merge of the sorted candidate lists: lane = list
        int pnt = 0;
        for (int j = 0; j < k; j++) {
            float bv = pnt < k ? cv[lane * 16 + pnt] : -INFINITY; int bi = pnt < k ? ci[lane * 16 + pnt] : 0x7fffffff;
            const int mine = bi;
            warp_argmax(bv, bi);
            if (bi != 0x7fffffff && bi == mine) pnt++;
            if (lane == 0) { kv[j] = bv; ki[j] = bi; }
        }
    }
    __syncthreads();
    if (tid == 0) {
        const float mx = kv[0]; float sum = 0.f; float pr[16];
        for (int j = 0; j < k; j++) { pr[j] = expf(kv[j] - mx); sum += pr[j]; }
        for (int j = 0; j < k; j++) {
            ids[r * k + j] = ki[j]; vals[r * k + j] = kv[j];
            if (probs) probs[r * k + j] = pr[j] / sum;
        }
    }
}

// Draft-level form: grid (rows, TK_S).  A single CTA per row (above) reads its 128 KB x ks of planes at one SM's L2 bandwidth and
// leaves 140 SMs idle; here each CTA takes 1/TK_S of a row, writes its k sorted candidates to a scratch list, and the CTA that
// arrives last for a row (device counter) merges the TK_S lists.  Same comparator at every level (value, then lower id), so the
// result is identical to the single-CTA kernel and to the oracle.  scr: unsigned cnt[128] (zero before the first launch; every
// launch leaves it zero again; at a fixed place whatever `rows` is), then float v[rows][TK_S][16], int i[rows][TK_S][16].
constexpr int TK_S = 16, TK_THREADS = 256, TK_W = TK_THREADS / 32, TK_PER = 8;
size_t eb_top_k_scratch_bytes(int rows) { return 512 + (size_t)rows * TK_S * 16 * 8; }
// order-preserving float -> u32 key; 0 = "no candidate" (-inf is never picked, as in the single-CTA kernel)
__device__ __forceinline__ uint32_t tk_key(float v) {
    uint32_t u = __float_as_uint(v);
    if (u == 0x80000000u) u = 0u;                                   // -0 == +0: the tie goes to the lower id
    return v == -INFINITY ? 0u : (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float tk_val(uint32_t key) { return key == 0u ? -INFINITY : __uint_as_float((key & 0x80000000u) ? key ^ 0x80000000u : ~key); }
// Top k of the <= 32 x TK_PER candidates a warp holds in registers, by (value desc, id asc).  A candidate is one u64:
// (key << 32) | ~id, so "greater" is exactly the required order; 0 = none.  Each lane first sorts its own 8 (19 compare-exchanges,
// branch-free), then a round is two hardware warp reductions (redux.sync max of the head keys, then of ~id among the lanes
// that hold it) and a register shift in the winning lane — no shuffle trees, no rescans (those cost ~300 clk per round, three
// phases deep).  Lane j < k returns the j-th pick in (ov, oi).
__device__ __forceinline__ void tk_cex(uint64_t &a, uint64_t &b) { const uint64_t hi = a > b ? a : b, lo = a > b ? b : a; a = hi; b = lo; }
__device__ __forceinline__ void warp_topk_regs(uint64_t (&c)[TK_PER], int k, float &ov, int &oi) {
    static_assert(TK_PER == 8, "sorting network below is for 8 inputs");
    tk_cex(c[0], c[1]); tk_cex(c[2], c[3]); tk_cex(c[4], c[5]); tk_cex(c[6], c[7]);
    tk_cex(c[0], c[2]); tk_cex(c[1], c[3]); tk_cex(c[4], c[6]); tk_cex(c[5], c[7]);
    tk_cex(c[1], c[2]); tk_cex(c[5], c[6]); tk_cex(c[0], c[4]); tk_cex(c[3], c[7]);
    tk_cex(c[1], c[5]); tk_cex(c[2], c[6]);
    tk_cex(c[1], c[4]); tk_cex(c[3], c[6]);
    tk_cex(c[2], c[4]); tk_cex(c[3], c[5]);
    tk_cex(c[3], c[4]);
    ov = -INFINITY; oi = 0x7fffffff;
    const int lane = threadIdx.x & 31;
    for (int r = 0; r < k; r++) {
        const uint32_t hk = (uint32_t)(c[0] >> 32), hl = (uint32_t)c[0];
        const uint32_t mk = __reduce_max_sync(0xffffffffu, hk);
        if (mk == 0u) break;                                        // fewer candidates than k: the rest stay (-inf, 0x7fffffff)
        const uint32_t ml = __reduce_max_sync(0xffffffffu, hk == mk ? hl : 0u);
        if (lane == r) { ov = tk_val(mk); oi = (int)~ml; }
        if (hk == mk && hl == ml) {
#pragma unroll
            for (int j = 0; j + 1 < TK_PER; j++) c[j] = c[j + 1];
            c[TK_PER - 1] = 0ull;
        }
    }
}
__device__ __forceinline__ uint64_t tk_pack(float v, int id) { const uint32_t key = tk_key(v); return key == 0u ? 0ull : ((uint64_t)key << 32) | (uint32_t)~id; }
__global__ void __launch_bounds__(TK_THREADS) top_k_split_kernel(const float *__restrict__ logits, int ks, int64_t plane, int V, int k,
                                                                 int32_t *__restrict__ ids, float *__restrict__ vals, float *__restrict__ probs,
                                                                 const int *__restrict__ gate, float *scr_v, int *scr_i, unsigned *cnt) {
    __shared__ float cv[TK_S * 16]; __shared__ int ci[TK_S * 16];
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int r = blockIdx.x, sl = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (V + TK_S - 1) / TK_S;                          // slice length (<= TK_THREADS * TK_PER, checked by the launcher)
    const int base = sl * per, n = max(0, min(V, base + per) - base);
    const float *x = logits + (int64_t)r * V + base;
    // phase 1: every warp takes 1/8 of the slice, 8 values per lane straight from the planes (coalesced, all loads in flight)
    float val[TK_PER]; uint64_t c[TK_PER];
    const int wchunk = (n + TK_W - 1) / TK_W, lo = warp * wchunk, hi = min(n, lo + wchunk);
#pragma unroll
    for (int j = 0; j < TK_PER; j++) { const int i = lo + lane + 32 * j; val[j] = i < hi ? x[i] : -INFINITY; }
    for (int s = 1; s < ks; s++) {
#pragma unroll
        for (int j = 0; j < TK_PER; j++) { const int i = lo + lane + 32 * j; if (i < hi) val[j] += x[s * plane + i]; }
    }
#pragma unroll
    for (int j = 0; j < TK_PER; j++) c[j] = tk_pack(val[j], base + lo + lane + 32 * j);      // -inf (also every i >= hi) packs to "none"
    float ov; int oi;
    warp_topk_regs(c, k, ov, oi);
    if (lane < 16) { cv[warp * 16 + lane] = ov; ci[warp * 16 + lane] = oi; }
    __syncthreads();
    if (warp != 0) return;
    // phase 2: warp 0 merges the 8 warp lists (128 slots, 4 per lane)
    auto take = [&](const float *lv, const int *li, int slots) {
#pragma unroll
        for (int j = 0; j < TK_PER; j++) {
            const int q = lane + 32 * j; const bool ok = q < slots && (q & 15) < k && li[q] != 0x7fffffff;
            c[j] = ok ? tk_pack(lv[q], li[q]) : 0ull;
        }
    };
    take(cv, ci, TK_W * 16);
    warp_topk_regs(c, k, ov, oi);
    float *gv = scr_v + ((int64_t)r * TK_S + sl) * 16; int *gi = scr_i + ((int64_t)r * TK_S + sl) * 16;
    if (lane < 16) { __stcg(gv + lane, ov); __stcg(gi + lane, oi); }
    __threadfence();
    __syncwarp();
    int last = 0;
    if (lane == 0) last = atomicAdd(cnt + r, 1u) == TK_S - 1;
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    // phase 3: the slice that arrives last merges the TK_S slice lists of the row (256 slots, 8 per lane)
    __threadfence();
    for (int i = lane; i < TK_S * 16; i += 32) { cv[i] = __ldcg(scr_v + (int64_t)r * TK_S * 16 + i); ci[i] = __ldcg(scr_i + (int64_t)r * TK_S * 16 + i); }
    __syncwarp();
    take(cv, ci, TK_S * 16);
    warp_topk_regs(c, k, ov, oi);
    if (lane == 0) cnt[r] = 0;
    // probs = softmax over the k kept logits, summed in list order like the single-CTA kernel (llama_sampler_softmax_impl)
    const float mx = __shfl_sync(0xffffffffu, ov, 0);
    const float pe = lane < k ? expf(ov - mx) : 0.f;
    float sum = 0.f;
    for (int j = 0; j < k; j++) sum += __shfl_sync(0xffffffffu, pe, j);
    if (lane < k) {
        ids[r * k + lane] = oi; vals[r * k + lane] = ov;
        if (probs) probs[r * k + lane] = pe / sum;
    }
}
int eb_top_k_impl(const float *logits, int rows, int V, int k, int32_t *ids, float *vals, float *probs, cudaStream_t st, int ks, int64_t plane, void *scratch) {
    if (k < 1 || k > 16 || V > 56000) return eb_set_err(EB_ERR_ARG, "top_k: k=%d V=%d", k, V);
    if (rows == 0) return EB_OK;
    if (ks > 1 && (plane & 3)) return eb_set_err(EB_ERR_ARG, "top_k: plane stride %lld", (long long)plane);
    static const bool no_split = getenv("EB_NO_TOPK_SPLIT") != nullptr;
    if (scratch && !no_split && V >= TK_S * 64 && V <= TK_S * TK_THREADS * TK_PER && rows <= 128) {
        unsigned *cnt = (unsigned *)scratch; float *sv = (float *)((char *)scratch + 512); int *si = (int *)(sv + (size_t)rows * TK_S * 16);
        eb_launch_pdl_c(EB_PDL_TOPK, top_k_split_kernel, dim3(rows, TK_S), dim3(TK_THREADS), 0, st, logits, ks, plane, V, k, ids, vals, probs, eb_gate(), sv, si, cnt);
        EB_CHECK_LAUNCH();
        return EB_OK;
    }
    eb_launch_pdl_c(EB_PDL_TOPK, top_k_kernel, dim3(rows), dim3(1024), (size_t)V * 4, st, logits, ks, plane, V, k, ids, vals, probs, eb_gate());
    EB_CHECK_LAUNCH();
    return EB_OK;
}

__global__ void __launch_bounds__(1024) argmax_kernel(const float *__restrict__ logits, int ks, int64_t plane, int V, int32_t *__restrict__ ids) {
    __shared__ float wv[32]; __shared__ int wi[32];
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *x = logits + (int64_t)r * V;
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int i0 = tid; i0 < V; i0 += 8 * 1024) {        // 8 loads in flight per thread
        float t[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int i = i0 + u * 1024; t[u] = i < V ? x[i] : -INFINITY; }
        for (int s = 1; s < ks; s++) {                  // split-K planes of the lm_head, summed in plane order
#pragma unroll
            for (int u = 0; u < 8; u++) { const int i = i0 + u * 1024; if (i < V) t[u] += x[s * plane + i]; }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) if (t[u] > bv) { bv = t[u]; bi = i0 + u * 1024; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { wv[warp] = bv; wi[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        bv = wv[lane]; bi = wi[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) ids[r] = bi == 0x7fffffff ? 0 : bi;
    }
}
// GGML_OP_ARGSORT (ggml_argsort / ggml_top_k, R/ggml/include/ggml.h:1761-1776; CUDA: R/ggml/src/ggml-cuda/argsort.cu): per row the
// indices that sort the row, ascending or descending.  One CTA per row, bitonic network over the next power of two (<= 1024) in shared
// memory; padding sorts last; equal values keep index order (the comparison falls back to the index), so the result is deterministic.
__global__ void argsort_kernel(const float *__restrict__ x, int n, int n_pad, int desc, int32_t *__restrict__ ids) {
    extern __shared__ int32_t idx[];
    const float *row = x + (int64_t)blockIdx.x * n;
    const int t = threadIdx.x;
    idx[t] = t;
    __syncthreads();
    // a sorts before b ?
    auto before = [&](int a, int b) {
        if (a >= n) return false;
        if (b >= n) return true;
        const float va = row[a], vb = row[b];
        if (va != vb) return desc ? va > vb : va < vb;
        return a < b;
    };
    for (int k = 2; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int p = t ^ j;
            if (p > t) {
                const int a = idx[t], b = idx[p];
                const bool up = (t & k) == 0;                 // this pair sorts "forward"
                if (up ? before(b, a) : before(a, b)) { idx[t] = b; idx[p] = a; }
            }
            __syncthreads();
        }
    }
    if (t < n) ids[(int64_t)blockIdx.x * n + t] = idx[t];
}
int eb_argsort_impl(const float *x, int rows, int n, int descending, int32_t *ids, cudaStream_t st) {
    if (rows == 0) return EB_OK;
    if (n < 1 || n > 1024) return eb_set_err(EB_ERR_UNSUPPORTED, "argsort: row length %d (1..1024)", n);
    int n_pad = 1; while (n_pad < n) n_pad <<= 1;
    argsort_kernel<<<rows, n_pad, (size_t)n_pad * 4, st>>>(x, n, n_pad, descending, ids);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_argmax_impl(const float *logits, int rows, int V, int32_t *ids, cudaStream_t st, int ks, int64_t plane) {
    if (rows == 0) return EB_OK;
    argmax_kernel<<<rows, 1024, 0, st>>>(logits, ks, plane, V, ids);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ tree state kernels
__device__ void set_bits_from_parent(SpecState *s, int node) {
    uint32_t *b = s->tree_bits + node * 4;
    const int par = s->parent[node];
    if (par >= 0) { const uint32_t *pb = s->tree_bits + par * 4; b[0] = pb[0]; b[1] = pb[1]; b[2] = pb[2]; b[3] = pb[3]; }
    else { b[0] = b[1] = b[2] = b[3] = 0; }
    b[node >> 5] |= 1u << (node & 31);
}

// Start of a round: node 0 = last accepted token (R/examples/speculative/speculative-eagle.cpp:516-517),
// branch 0 drafting, first draft batch = {root} with the target's feature as hidden input.
__global__ void round_begin_kernel(SpecState *s, SpecCfg c) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    for (int i = 0; i < c.max_nodes; i++) {            // dummy rows: attend only themselves
        s->tokens[i] = 0; s->parent[i] = -1; s->pos[i] = s->n_past; s->depth[i] = 0;
        uint32_t *b = s->tree_bits + i * 4; b[0] = b[1] = b[2] = b[3] = 0; b[i >> 5] = 1u << (i & 31);
    }
    s->tokens[0] = s->root_token; s->n_nodes = 1; s->n_seq_cur = 1; s->cur_depth = 0; s->done = 0;
    for (int b = 0; b < 16; b++) { s->br_drafting[b] = 0; s->br_last_node[b] = 0; s->br_row[b] = 0; }
    s->br_drafting[0] = 1;
    s->dft_n = 1; s->dft_node_base = 0; s->gate_fwd = 1; s->gate_lm = 1;
    for (int r = 0; r < c.n_branch; r++) {
        s->dft_tokens[r] = r == 0 ? s->root_token : 0; s->dft_pos[r] = s->n_past; s->dft_src[r] = -1;
        uint32_t *b = s->dft_bits + r * 4; b[0] = b[1] = b[2] = b[3] = 0; b[r >> 5] = 1u << (r & 31);   // bit of cell node_base + r
    }
    s->ctl_head.n_prefix = s->n_past; s->ctl_head.node_base = 0; s->ctl_head.n_tree = c.n_branch;
    s->ctl_tgt.n_prefix = s->n_past; s->ctl_tgt.node_base = 0; s->ctl_tgt.n_tree = c.max_nodes;
}

// hidden input rows of the draft batch: root <- target feature, others <- the head's own
// result_norm of the parent node (F5: R/src/llama.cpp:10170, R/src/llama-context.cpp:83-88)
__global__ void gather_hidden_kernel(const SpecState *s, const float *root_feat, const float *node_feat, float *hidden, int E, int n_branch) {
    const int r = blockIdx.y; const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= E || r >= n_branch) return;
    const int src = s->dft_src[r];
    const float *p = src < 0 ? root_feat : node_feat + (int64_t)src * E;
    *(float4 *)(hidden + (int64_t)r * E + i) = *(const float4 *)(p + i);
}
// after a head forward: feat rows -> per-node feature store
__global__ void scatter_feat_kernel(const SpecState *s, const float *feat, float *node_feat, int E, int n_branch) {
    const int r = blockIdx.y; const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= E || r >= s->dft_n) return;
    const int node = s->dft_node_base + r;
    *(float4 *)(node_feat + (int64_t)node * E + i) = *(const float4 *)(feat + (int64_t)r * E + i);
}

// One depth level of the reference's tree expansion (speculative-eagle.cpp:527-642), on device.
// cand_ids/cand_p: [n_branch][top_k] sorted candidates of the rows of the last draft batch.
// The walk is serial (one thread), but every access to the state would be a dependent global round trip: the CTA copies
// the state and the candidate lists into shared memory, thread 0 works there, the CTA copies the state back.
__device__ void tree_expand_serial(SpecState *s, const SpecCfg &c, const int32_t *cand_ids, const float *cand_p);
__global__ void __launch_bounds__(256) tree_expand_kernel(SpecState *gs, SpecCfg c, const int32_t *g_ids, const float *g_p) {
    __shared__ __align__(16) uint32_t sm_state[(sizeof(SpecState) + 3) / 4];
    __shared__ int32_t sm_ids[256]; __shared__ float sm_p[256];
    pdl_trigger(); pdl_wait();
    constexpr int NW = (int)(sizeof(SpecState) / 4);
    static_assert(sizeof(SpecState) % 4 == 0, "SpecState is copied word-wise");
    for (int i = threadIdx.x; i < NW; i += blockDim.x) sm_state[i] = ((const uint32_t *)gs)[i];
    for (int i = threadIdx.x; i < c.n_branch * c.top_k; i += blockDim.x) { sm_ids[i] = g_ids[i]; sm_p[i] = g_p[i]; }
    __syncthreads();
    if (threadIdx.x == 0) tree_expand_serial((SpecState *)sm_state, c, sm_ids, sm_p);
    __syncthreads();
    for (int i = threadIdx.x; i < NW; i += blockDim.x) ((uint32_t *)gs)[i] = sm_state[i];
}
__device__ void tree_expand_serial(SpecState *s, const SpecCfg &c, const int32_t *cand_ids, const float *cand_p) {
    if (s->done) {                                     // nothing left to draft: park the dummy rows behind the real nodes
        const int base = s->n_nodes;
        s->dft_n = 0; s->dft_node_base = base; s->gate_fwd = 0; s->gate_lm = 0;
        for (int r = 0; r < c.n_branch; r++) {
            s->dft_tokens[r] = 0; s->dft_pos[r] = s->n_past; s->dft_src[r] = -1;
            uint32_t *b = s->dft_bits + r * 4; b[0] = b[1] = b[2] = b[3] = 0;
            const int cell = base + r; if (cell < 128) b[cell >> 5] = 1u << (cell & 31);
        }
        s->ctl_head.node_base = base; s->ctl_head.n_tree = min(128, base + c.n_branch);
        return;
    }
    const int n_seq_dft = c.n_branch;
    int skip[16];
    for (int b = 0; b < 16; b++) skip[b] = 0;
    const int first_new = s->n_nodes;
    int n_new = 0;
    int new_tok[16], new_src[16], new_br[16];
    for (int b = 0; b < n_seq_dft; b++) {
        if (!s->br_drafting[b] || skip[b]) continue;
        const int32_t *ids = cand_ids + s->br_row[b] * c.top_k; const float *pr = cand_p + s->br_row[b] * c.top_k;
        int sa[16]; int nsa = 0; sa[nsa++] = b;
        for (int f = 1; f < 8 && f < c.top_k; f++) {                                  // :556-597
            if (s->n_seq_cur < n_seq_dft && pr[f] > c.p_split) {
                const int nb = s->n_seq_cur++;
                s->br_drafting[nb] = 1; skip[nb] = 1; s->br_last_node[nb] = s->br_last_node[b]; s->br_row[nb] = s->br_row[b];
                sa[nsa++] = nb;
            } else break;
        }
        for (int is = 0; is < nsa; is++) {                                            // :599-625
            const int br = sa[is];
            if (s->n_nodes + n_new >= c.max_nodes) { s->br_drafting[br] = 0; continue; }   // capacity guard (not in the reference)
            new_tok[n_new] = ids[is]; new_src[n_new] = s->br_last_node[br]; new_br[n_new] = br; n_new++;
            if (s->n_nodes + n_new > c.n_draft) s->br_drafting[br] = 0;               // batch_tgt.n_tokens > n_draft
        }
        // chain driver's confidence stop (R/common/speculative.cpp:283): the token just drafted stays, the branch that drew it stops
        if (c.p_min > 0.f && pr[0] < c.p_min) s->br_drafting[b] = 0;
    }
    for (int i = 0; i < n_new; i++) {
        const int node = first_new + i, par = new_src[i];
        s->tokens[node] = new_tok[i]; s->parent[node] = par; s->depth[node] = s->depth[par] + 1; s->pos[node] = s->n_past + s->depth[node];
        set_bits_from_parent(s, node);
        s->br_last_node[new_br[i]] = node; s->br_row[new_br[i]] = i;
    }
    s->n_nodes = first_new + n_new;
    s->cur_depth++;
    // next draft batch = the new nodes (every node is run through the head so that it owns a KV cell)
    s->dft_n = n_new; s->dft_node_base = first_new;
    for (int r = 0; r < c.n_branch; r++) {
        const bool live = r < n_new; const int node = first_new + r;
        s->dft_tokens[r] = live ? new_tok[r] : 0;
        s->dft_pos[r] = live ? s->pos[node] : s->n_past;
        s->dft_src[r] = live ? new_src[r] : -1;
        uint32_t *b = s->dft_bits + r * 4;
        if (live) { const uint32_t *nb = s->tree_bits + node * 4; b[0] = nb[0]; b[1] = nb[1]; b[2] = nb[2]; b[3] = nb[3]; }
        else { b[0] = b[1] = b[2] = b[3] = 0; const int cell = first_new + r; if (cell < 128) b[cell >> 5] = 1u << (cell & 31); }
    }
    s->ctl_head.node_base = first_new; s->ctl_head.n_tree = min(128, first_new + c.n_branch);
    int any = 0;
    for (int b = 0; b < n_seq_dft; b++) any |= s->br_drafting[b];
    if (n_new == 0 || !any || s->n_nodes > c.n_draft) s->done = 1;                    // :627-641
    s->gate_fwd = n_new > 0; s->gate_lm = !s->done;      // the new nodes still need their draft-KV cells; their logits only if drafting goes on
}

// Greedy accept walk (speculative-eagle.cpp:395-449) + state update.  best[t] = target argmax at
// node t.  fixed_accept >= 0: accept exactly that many draft nodes down first children (bench).
__global__ void accept_kernel(SpecState *s, SpecCfg c, const int32_t *best) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int cur = 0, n = 0;
    s->path[n++] = 0;
    for (;;) {
        int next = -1;
        for (int ch = cur + 1; ch < s->n_nodes; ch++) {
            if (s->parent[ch] != cur) continue;
            if (c.fixed_accept >= 0 ? true : s->tokens[ch] == best[cur]) { next = ch; break; }
        }
        if (next < 0 || (c.fixed_accept >= 0 && n - 1 >= c.fixed_accept)) break;
        cur = next; s->path[n++] = cur;
    }
    s->n_accept = n; s->bonus = best[cur]; s->last_node = cur;
}

// KV compaction, phase 1: gather accepted cells to a staging buffer; phase 2: write them back at
// the contiguous tail.  (Reference: metadata-only seq_keep/seq_cp + lazy defrag,
// R/src/llama-kv-cache.cpp:368-499, R/src/llama.cpp:1384-1450; here the rows move immediately.)
__global__ void kv_gather_kernel(const SpecState *s, const __half *kc, const __half *vc, __half *stage, int64_t layer_stride, int E, int max_acc) {
    const int i = blockIdx.y, l = blockIdx.z;
    if (i >= s->n_accept || s->path[i] == i) return;
    const int64_t src = (int64_t)l * layer_stride + (int64_t)(s->n_past + s->path[i]) * E;
    const int64_t dst = (((int64_t)l * max_acc + i) * 2) * E;
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) * 8; e < E; e += gridDim.x * blockDim.x * 8) {
        *(int4 *)(stage + dst + e) = *(const int4 *)(kc + src + e);
        *(int4 *)(stage + dst + E + e) = *(const int4 *)(vc + src + e);
    }
}
__global__ void kv_scatter_kernel(const SpecState *s, __half *kc, __half *vc, const __half *stage, int64_t layer_stride, int E, int max_acc) {
    const int i = blockIdx.y, l = blockIdx.z;
    if (i >= s->n_accept || s->path[i] == i) return;
    const int64_t dstc = (int64_t)l * layer_stride + (int64_t)(s->n_past + i) * E;
    const int64_t src = (((int64_t)l * max_acc + i) * 2) * E;
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) * 8; e < E; e += gridDim.x * blockDim.x * 8) {
        *(int4 *)(kc + dstc + e) = *(const int4 *)(stage + src + e);
        *(int4 *)(vc + dstc + e) = *(const int4 *)(stage + src + E + e);
    }
}
// end of round: commit, hand the target feature of the last accepted node to the head, fill the
// host-visible result record
__global__ void round_end_kernel(SpecState *s, const float *tgt_feat, float *root_feat, int E, eb_round_result *res) {
    const int last = s->last_node;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) root_feat[i] = tgt_feat[(int64_t)last * E + i];
    if (blockIdx.x == 0) {                                  // one thread per tree slot (a single thread walking 128 x 3 words took 19.5 us)
        const int n_accept = s->n_accept;
        if (threadIdx.x < 128) { const int i = threadIdx.x; res->tokens[i] = s->tokens[i]; res->parent[i] = s->parent[i]; res->path[i] = i < n_accept ? s->path[i] : -1; }
        __syncthreads();
        if (threadIdx.x == 0) {
            res->n_nodes = s->n_nodes; res->n_accept = n_accept; res->bonus = s->bonus;
            s->n_past += n_accept; s->root_token = s->bonus;    // commit: nobody else in this launch reads n_past / root_token
        }
    }
}

int eb_spec_round_begin(SpecState *s, const SpecCfg &c, cudaStream_t st) { round_begin_kernel<<<1, 32, 0, st>>>(s, c); EB_CHECK_LAUNCH(); return EB_OK; }
int eb_spec_gather_hidden(const SpecState *s, const float *root_feat, const float *node_feat, float *hidden, int E, int n_branch, cudaStream_t st) {
    gather_hidden_kernel<<<dim3((unsigned)cdiv(E / 4, 256), n_branch), 256, 0, st>>>(s, root_feat, node_feat, hidden, E, n_branch); EB_CHECK_LAUNCH(); return EB_OK;
}
int eb_spec_scatter_feat(const SpecState *s, const float *feat, float *node_feat, int E, int n_branch, cudaStream_t st) {
    scatter_feat_kernel<<<dim3((unsigned)cdiv(E / 4, 256), n_branch), 256, 0, st>>>(s, feat, node_feat, E, n_branch); EB_CHECK_LAUNCH(); return EB_OK;
}
int eb_spec_expand(SpecState *s, const SpecCfg &c, const int32_t *cand_ids, const float *cand_p, cudaStream_t st) {
    if (c.n_branch * c.top_k > 256) return eb_set_err(EB_ERR_ARG, "expand: n_branch * top_k > 256");
    eb_launch_pdl_c(EB_PDL_EXPAND, tree_expand_kernel, dim3(1), dim3(256), 0, st, s, c, cand_ids, cand_p); EB_CHECK_LAUNCH(); return EB_OK;
}
int eb_spec_accept(SpecState *s, const SpecCfg &c, const int32_t *best, cudaStream_t st) { accept_kernel<<<1, 32, 0, st>>>(s, c, best); EB_CHECK_LAUNCH(); return EB_OK; }
int eb_spec_compact(SpecState *s, void *kc, void *vc, void *stage, int64_t layer_stride, int E, int L, int max_acc, cudaStream_t st) {
    const dim3 grid(2, max_acc, L);
    kv_gather_kernel<<<grid, 256, 0, st>>>(s, (const __half *)kc, (const __half *)vc, (__half *)stage, layer_stride, E, max_acc); EB_CHECK_LAUNCH();
    kv_scatter_kernel<<<grid, 256, 0, st>>>(s, (__half *)kc, (__half *)vc, (const __half *)stage, layer_stride, E, max_acc); EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_spec_round_end(SpecState *s, const float *tgt_feat, float *root_feat, int E, eb_round_result *res, cudaStream_t st) {
    round_end_kernel<<<4, 256, 0, st>>>(s, tgt_feat, root_feat, E, res); EB_CHECK_LAUNCH();
    return EB_OK;
}

// scratch of the operator-layer eb_top_k (sessions own theirs): one per stream, so that calls on different streams do not share the
// slice lists and arrival counters (they used to: ADVICE round 1).  Created on first use, kept until the process exits.
static void *g_topk_scr = nullptr;      // the NULL stream's
static std::mutex g_topk_mu;
static std::map<cudaStream_t, void *> g_topk_by_stream;
void *eb_top_k_default_scratch(cudaStream_t st) {
    if (!st) return g_topk_scr;
    std::lock_guard<std::mutex> lk(g_topk_mu);
    auto it = g_topk_by_stream.find(st);
    if (it != g_topk_by_stream.end()) return it->second;
    void *p = nullptr;
    if (cudaMalloc(&p, eb_top_k_scratch_bytes(128)) != cudaSuccess || cudaMemset(p, 0, eb_top_k_scratch_bytes(128)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    g_topk_by_stream[st] = p;
    return p;
}

// ---- eb_top_k_noalloc: the 16-slice top-k with its scratch in a file-scope device array, for callers that cannot allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc); no
// counterpart in the reference.  The array is part of the module image: every GPU that loads the library holds these 66 048 bytes
// until the process exits.  In HOST code the name `g_topk_static` is only a stand-in for the symbol, not a device address (a kernel
// given it would fault, or quietly use host memory where the GPU can reach it): its device address is taken once per device, with
// that device current, by cudaGetSymbolAddress in eb_spec_init() and kept in g_topk_static_ptr[device].
#define EB_TOPK_NOALLOC_BYTES (512 + (size_t)EB_TOP_K_NOALLOC_MAX_ROWS * TK_S * 16 * 8)      // arrival counters | slice lists (vals, ids)
static_assert(EB_TOP_K_NOALLOC_MAX_ROWS * 4 <= 512, "one arrival counter per row in the first 512 bytes");
static __device__ __align__(256) unsigned char g_topk_static[EB_TOPK_NOALLOC_BYTES];
constexpr int EB_MAX_DEVICES = 64;
static void *g_topk_static_ptr[EB_MAX_DEVICES] = {};

int eb_spec_init() {
    EB_CUDA(cudaFuncSetAttribute(top_k_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 56000 * 4));
    if (!g_topk_scr) { EB_CUDA(cudaMalloc(&g_topk_scr, eb_top_k_scratch_bytes(128))); EB_CUDA(cudaMemset(g_topk_scr, 0, eb_top_k_scratch_bytes(128))); }
    int dev = -1;
    EB_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < EB_MAX_DEVICES && !g_topk_static_ptr[dev]) {
        void *ptr = nullptr;
        EB_CUDA(cudaGetSymbolAddress(&ptr, g_topk_static));          // device address of the array on THIS device
        EB_CUDA(cudaMemset(ptr, 0, EB_TOPK_NOALLOC_BYTES));
        g_topk_static_ptr[dev] = ptr;
    }
    return EB_OK;
}

// Allocation-free entry point: inputs and outputs are the caller's device buffers; nothing here calls cudaMalloc / cudaFree or an
// allocating library routine (the kernel launch and a 512-byte cudaMemsetAsync are the only CUDA calls, both legal under stream
// capture).  Every call on a device shares g_topk_static, which keeps the previous call's slice lists: the lists are fully
// rewritten before they are read (each slice writes its 16 slots, the last arriver reads them), the arrival counters must start at
// zero and are cleared on the stream before the launch.  Two calls that overlap on the GPU (different streams, or concurrent graph
// branches) race on the array: serialise them.
extern "C" int eb_top_k_noalloc(const float *d_logits, int rows, int V, int k, int32_t *d_ids, float *d_vals, float *d_probs, void *stream) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= EB_MAX_DEVICES || !g_topk_static_ptr[dev])
        return eb_set_err(EB_ERR_CUDA, "top_k_noalloc: call eb_device_init(device) for the current device first (this entry point sets nothing up itself)");
    if (!d_logits || !d_ids || !d_vals) return eb_set_err(EB_ERR_ARG, "top_k_noalloc: null buffer");
    if (rows < 0 || rows > EB_TOP_K_NOALLOC_MAX_ROWS)
        return eb_set_err(EB_ERR_ARG, "top_k_noalloc: rows=%d exceeds the static scratch (EB_TOP_K_NOALLOC_MAX_ROWS = %d): use eb_top_k", rows, EB_TOP_K_NOALLOC_MAX_ROWS);
    if (k < 1 || k > 16) return eb_set_err(EB_ERR_ARG, "top_k_noalloc: k=%d (1..16)", k);
    if (V < TK_S * 64 || V > TK_S * TK_THREADS * TK_PER)
        return eb_set_err(EB_ERR_ARG, "top_k_noalloc: V=%d outside the 16-slice kernel's range [%d, %d]: use eb_top_k", V, TK_S * 64, TK_S * TK_THREADS * TK_PER);
    if (rows == 0) return EB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char *scr = (unsigned char *)g_topk_static_ptr[dev];
    unsigned *cnt = (unsigned *)scr; float *sv = (float *)(scr + 512); int *si = (int *)(sv + (size_t)rows * TK_S * 16);
    EB_CUDA(cudaMemsetAsync(cnt, 0, 512, st));                       // counters start at zero whatever an earlier (aborted) call left
    top_k_split_kernel<<<dim3(rows, TK_S), dim3(TK_THREADS), 0, st>>>(d_logits, 1, 0, V, k, d_ids, d_vals, d_probs, nullptr, sv, si, cnt);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
