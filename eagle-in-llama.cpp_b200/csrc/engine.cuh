// engine.cuh — model / session containers of the speculative-step engine.
#pragma once
#include <vector>
#include "common.cuh"
#include "spec.cuh"

struct DTensor { int type = -1; void *data = nullptr; int64_t rows = 0, cols = 0; size_t bytes = 0; bool owned = false; };
// rows of wq|wk(|wv) and gate|up stored back to back (Q4_K keeps the GGUF row-major block layout, so a row-wise
// concatenation is one bigger matrix): one GEMM launch instead of three / two
struct LayerFuse { int n_qkv = 1; bool gate_up = false; };

struct eb_model {
    int E, H, F, V, L, is_eagle; float eps, freq_base;
    // tensor-parallel shard (eb_model_create_tp): H = heads held by this rank, D = head size, Ea = H * D = rows of wq/wk/wv and
    // columns of wo held here, F = ffn rows held here.  Unsharded models: D = E / H, Ea = E.
    int D = 0, Ea = 0;
    DTensor g[EB_T_GLOBAL_COUNT];
    std::vector<std::vector<DTensor>> layers;
    std::vector<LayerFuse> fuse; std::vector<void *> fused_bufs; bool finalized = false;
};

int eb_model_alloc_tensor(eb_model *m, int layer, int which, int type, void **d_ptr, size_t *bytes);

struct KVCache { __half *k = nullptr, *v = nullptr; int64_t layer_stride = 0; int L = 0, E = 0, n_ctx = 0; };

// activation buffers for one forward of up to Tmax rows
struct Workspace {
    int Tmax = 0;
    float *x, *cur, *q, *k, *v, *att, *gate, *up, *cat, *hidden, *feat, *logits, *partial;
    float *pq, *pk, *pv, *po, *pgate, *pup;      // split-K planes [<=8][T][N], summed by the consumer kernel
    void *actq2;                                  // second activation-grid buffer
    void *actq; size_t actq_bytes; void *fa_ws; size_t fa_bytes; size_t partial_bytes;
    int32_t *tokens, *pos; uint32_t *bits; FwdCtl *ctl;
    float *rope_tab;                              // [T][D/2] (cos, sin)
    __half *q16;                                  // f16 copy of the rotated queries (TMA source of the tcgen05 attention)
};

struct eb_tp;      // tp.cu: peer-memory communicator of a tensor-parallel group
int eb_tp_allreduce_planes(eb_tp *tp, const float *planes, int ks, int64_t n, float *out, cudaStream_t st, int set);
int eb_tp_fused_ok(const eb_tp *tp, int ks, int64_t n);
TpPush eb_tp_push(const eb_tp *tp, int set);
TpPull eb_tp_pull(const eb_tp *tp, int set, int64_t plane);

struct ActTap { int8_t *buf = nullptr; size_t cap = 0, off = 0; int sites = 0; };

struct eb_session {
    ActTap tap;
    eb_model *tgt, *head; eb_session_params p; SpecCfg cfg;
    KVCache kv_t, kv_h; Workspace ws_t, ws_h;
    SpecState *state; float *root_feat, *node_feat; int32_t *cand_ids; float *cand_vals, *cand_p; void *topk_scr; int32_t *best;
    __half *stage; int max_acc;
    eb_round_result *d_res, *h_res;
    cudaStream_t st; cudaEvent_t ev0, ev1;
    cudaGraphExec_t graph_round = nullptr, graph_verify = nullptr; int graph_verify_T = 0;
    long graph_round_kernels = 0, graph_verify_kernels = 0;
    int last_T[2] = {0, 0};
    eb_tp *tp = nullptr;   // tensor-parallel group of the target forward (eb_session_set_tp), or null
    float *tp_buf = nullptr;
    int n_past_h = 0;      // host mirror of SpecState.n_past (advanced from the round result): no extra D2H per round
};

// ops implemented in the other translation units
int eb_rms_norm_impl(const float *x, const float *w, float *y, int64_t n, int64_t rows, float eps, cudaStream_t st);
int eb_rope_norm_impl(const float *src, float *dst, const int32_t *pos, int D, int n_head, int T, float freq_base, cudaStream_t st);
int eb_add_impl(const float *a, const float *b, float *y, int64_t n, int64_t nb, cudaStream_t st);
int eb_mul_impl(const float *a, const float *b, float *y, int64_t n, int64_t nb, cudaStream_t st);
int eb_silu_impl(const float *x, float *y, int64_t n, cudaStream_t st);
int eb_scale_impl(const float *x, float *y, int64_t n, float s, cudaStream_t st);
int eb_relu_impl(const float *x, float *y, int64_t n, cudaStream_t st);
int eb_bias_relu_impl(const float *x, const float *b, float *y, int64_t n, int nb, cudaStream_t st);
int eb_cpy_f32_f16_impl(const float *x, void *y, int64_t n, cudaStream_t st);
int eb_copy_strided_impl(int src_type, int dst_type, const void *src, void *dst, const int64_t *ne, const int64_t *nb_src, const int64_t *dne, const int64_t *nb_dst, cudaStream_t st);
int eb_mul_mat_dense_impl(int a_type, const void *a, int64_t a_nb1, int64_t a_nb2, int64_t a_batches, const float *x, int64_t x_nb1, int64_t x_nb2,
                          float *y, int64_t N, int64_t K, int64_t T, int64_t B, cudaStream_t st);
int eb_soft_max_ext_impl(const float *x, const void *mask, int mask_type, int64_t mask_stride, float *y, int64_t n, int64_t rows, int64_t rows_per_batch, float scale, cudaStream_t st);
int eb_concat2_impl(const float *a, const float *b, float *y, int E, int T, cudaStream_t st);
int eb_swiglu_impl(const float *gate, const float *up, float *y, int64_t n, cudaStream_t st);
int eb_get_rows_dev(int type, const void *W, int64_t n_rows, int64_t K, const int32_t *d_ids, int n, float *out, cudaStream_t st);
size_t eb_flash_attn_ws_size_impl(int T, int H, int D, int n_kv);
int eb_flash_attn_impl(const float *q, const void *k, const void *v, int64_t kv_stride, const void *mask, int64_t mask_stride,
                       const uint32_t *tree_bits, int n_prefix, int T, int H, int D, int n_kv, float scale, float *out, void *ws, cudaStream_t st);
// engine forms: cell base / prefix length are read from device memory (CUDA-graph friendly)
int eb_rope_kv_store_ctl(float *q, const float *k, const float *v, const int32_t *pos, void *kc, void *vc, const FwdCtl *ctl,
                         int D, int n_head, int T, float freq_base, cudaStream_t st);
int eb_flash_attn_ctl(const float *q, const void *k, const void *v, int64_t kv_stride, const uint32_t *tree_bits, const FwdCtl *ctl,
                      int T, int H, int D, int n_ctx, float scale, float *out, void *ws, cudaStream_t st);
int eb_resid_norm_quant(float *x, const float *part, int ks, const float *w, float eps, int n, int T, const ActQ &aq, float *feat, cudaStream_t st, const float *bias = nullptr,
                        const SpecState *spec = nullptr, float *node_feat = nullptr, const TpPull *pull = nullptr);
struct PlaneRef { const float *p; int ks; int64_t ld; };   // split-K planes [ks][T][ld] of a (possibly fused) GEMM, first column of interest at p
int eb_rope_table(const int32_t *pos, int D, int T, float freq_base, float *tab, cudaStream_t st);
int eb_rope_kv_store_fused(PlaneRef qp, PlaneRef kp, PlaneRef vp, float *q, void *q16, const float *tab, void *kc, void *vc,
                           const FwdCtl *ctl, int D, int n_head, int T, cudaStream_t st, const int32_t *pos = nullptr, float freq_base = 10000.f);
// fattn_tc.cu: tcgen05 tree attention
int eb_fattn_tc_init(); int eb_fattn_tc_supported(int T, int D); size_t eb_fattn_tc_ws_size(int T, int H, int n_kv);
int eb_fattn_tc(const __half *q16, const void *k, const void *v, int64_t kv_stride, int64_t n_cells, const uint32_t *tree_bits, const FwdCtl *ctl,
                int n_prefix, int n_kv, int T, int H, float scale, float *out, void *ws, cudaStream_t st, const ActQ *aq, int *quantized,
                const uint32_t *mask_bits = nullptr, int mask_words = 0, const uint8_t *tile_kind = nullptr);
size_t eb_mask_bits_size_impl(int T, int n_kv);
int eb_mask_to_bits_impl(const void *mask, int64_t mask_stride, int T, int n_kv, void *bits, cudaStream_t st);
const int *eb_mask_bits_flag(const void *bits, int T, int n_kv);
int eb_fattn_tc_bits(const float *q, const void *k, const void *v, int64_t kv_stride, const void *bits, int T, int H, int n_kv, float scale, float *out, void *ws,
                     cudaStream_t st);
int eb_fattn_tc_f32q(const float *q, const void *k, const void *v, int64_t kv_stride, const uint32_t *tree_bits, int n_prefix, int n_kv,
                     int T, int H, float scale, float *out, void *ws, cudaStream_t st);
int eb_swiglu_quant(PlaneRef gate, PlaneRef up, int n, int T, const ActQ &aq, cudaStream_t st);
int eb_mul_mat_impl(int type, const void *W, int64_t N, int64_t K, const float *x, int T, float *y, void *actq_ws, float *partial, cudaStream_t st, bool reuse_act = false);
