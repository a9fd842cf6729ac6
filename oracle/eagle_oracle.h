/*
 * eagle_oracle.h — CPU restatement of the reference's EAGLE speculative-step hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (eagle-in-llama.cpp_b200/,
 * include/, bench.py's GPU arm) may call into this file; only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Parity pin: every arithmetic function here is checked in tests/test_oracle_vs_ref.py against
 * the reference's own ggml-cpu code compiled from /root/reference (oracle/_ref/libggml-ref.so,
 * recipe oracle/Makefile) and against the committed fixtures under tests/golden/.  The
 * EAGLE-specific composition (tree expansion with head features, accept walk, KV compaction)
 * has no test in the reference: that part is "parity unpinned by the reference" and pinned only
 * by the rules written in DESIGN.md.
 *
 * All paths below are relative to R = /root/reference/llama.cpp/.
 */
#ifndef EAGLE_ORACLE_H
#define EAGLE_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* tensor types: numeric values of enum ggml_type, R/ggml/include/ggml.h:351-390 */
enum { EO_F32 = 0, EO_F16 = 1, EO_Q4_0 = 2, EO_Q8_0 = 8, EO_Q4_K = 12, EO_Q6_K = 14 };

uint16_t eo_f32_to_f16(float f);
float    eo_f16_to_f32(uint16_t h);

size_t eo_row_size(int type, int64_t k);          /* bytes of one row of k elements   */
int    eo_block_elems(int type);                  /* 32 / 256 / 1                      */

/* R/ggml/src/ggml-quants.c:255 (q4_0) 349 (q8_0) 1280 (q4_K) 1690 (q6_K) */
void eo_dequantize_row(int type, const void *src, float *dst, int64_t k);

/* activation quantisers: R/ggml/src/ggml-quants.c:2479 (q8_K, iscale=-127/max) and
 * R/ggml/src/ggml-cpu/ggml-cpu-quants.c:738 (q8_0, AVX2 branch: id = 127/amax, round-to-even) */
void eo_quantize_row_q8_K(const float *x, void *y, int64_t k);
void eo_quantize_row_q8_0(const float *x, void *y, int64_t k);

/* one dot product, weights row (type) x quantised activations; scalar branches of
 * R/ggml/src/ggml-cpu/ggml-cpu-quants.c:1912,3663,6266,7903 */
float eo_vec_dot(int type, int64_t k, const void *w_row, const void *act_q);

/* Y[t*N + n] = sum_k W[n][k] X[t*K + k]   — R/ggml/src/ggml-cpu/ggml-cpu.c:7526 */
void eo_mul_mat(int type, const void *W, int64_t N, int64_t K, const float *X, int64_t T, float *Y);

/* R/ggml/src/ggml-cpu/ggml-cpu.c:7098 (double accumulation, no weight) */
void eo_rms_norm(const float *x, float *y, int64_t n, int64_t rows, float eps);
/* R/ggml/src/ggml-cpu/ggml-cpu.c:9449, mode NORM: x is [T][n_head][D], rotate pairs (2i,2i+1) */
void eo_rope_norm(float *x, const int32_t *pos, int D, int n_head, int T, float freq_base);
/* R/ggml/src/ggml-cpu/ggml-cpu.c:8547 */
void eo_get_rows(int type, const void *W, int64_t K, const int32_t *ids, int n, float *out);

/* R/ggml/src/ggml-cpu/ggml-cpu.c:10999-11212.  q [T][H][D] f32, k/v: f16 cells [n_kv][H][D],
 * vis[T][n_kv] != 0 means mask value 0, == 0 means -inf (cell skipped).  out [T][H][D]. */
void eo_flash_attn(const float *q, const uint16_t *k, const uint16_t *v, const uint8_t *vis,
                   int T, int H, int D, int n_kv, float scale, float *out);

/* checker switch for the GPU kernel's documented deviation: 1 = keep the V accumulator in f32
 * instead of the reference's f16 (default 0 = reference arithmetic) */
void eo_set_fattn_acc_f32(int on);

/* KQ mask rule of R/src/llama-context.cpp:161-193: cell i visible to query j iff the cell holds
 * the query's first seq id and cell_pos[i] <= pos[j].  seq sets are 64-bit bitsets. */
void eo_build_kq_mask(const int32_t *cell_pos, const uint64_t *cell_seq, int n_kv,
                      const int32_t *q_pos, const int32_t *q_seq0, int T, uint8_t *vis);

/* top-k by logit, descending; ties broken towards the lower token id (the reference's
 * std::partial_sort is unstable: R/src/llama-sampling.cpp:228-302; ties are flagged by tests).
 * probs = softmax over the k kept logits (sampler chain top_k -> dist). */
void eo_top_k(const float *logits, int V, int k, int32_t *ids, float *vals, float *probs);
int  eo_argmax(const float *logits, int V);

/* ---- model container + forward (build_llama R/src/llama.cpp:1647, build_eagle :1839) ---- */
typedef struct eo_model eo_model;
enum { EO_T_TOK_EMBD = 0, EO_T_OUTPUT_NORM, EO_T_OUTPUT, EO_T_FC, EO_T_FC_BIAS, EO_T_GLOBAL_COUNT };
enum { EO_L_ATTN_NORM = 0, EO_L_WQ, EO_L_WK, EO_L_WV, EO_L_WO, EO_L_FFN_NORM, EO_L_GATE, EO_L_UP,
       EO_L_DOWN, EO_L_COUNT };

eo_model *eo_model_new(int E, int H, int F, int V, int L, float eps, float freq_base, int is_eagle);
void      eo_model_free(eo_model *m);
void      eo_model_set_global(eo_model *m, int which, int type, const void *data);
void      eo_model_set_layer(eo_model *m, int layer, int which, int type, const void *data);
/* eagle heads use the target's output matrix as lm_head (R/src/llama.cpp:2016) */
void      eo_model_set_lm_head_from(eo_model *m, const eo_model *target);

typedef struct eo_kv eo_kv;  /* per-layer f16 K and V, [n_ctx][E] */
eo_kv    *eo_kv_new(int L, int E, int n_ctx);
void      eo_kv_free(eo_kv *kv);
uint16_t *eo_kv_k(eo_kv *kv, int layer);
uint16_t *eo_kv_v(eo_kv *kv, int layer);

/* Forward T tokens.  The new K/V rows are stored at cells [kv_head, kv_head+T); attention runs
 * over cells [0, n_kv) with visibility vis[T][n_kv].  hidden_in [T][E] is only read by eagle
 * heads.  Outputs: logits [T][V], feat [T][E] = result_norm (either may be NULL). */
/* tests only: append the int8 activation quants of every quantisation site of eo_forward to buf (NULL = off) */
void eo_set_tap(int8_t *buf, size_t cap);
size_t eo_tap_bytes(void);
void eo_forward(const eo_model *m, eo_kv *kv, const int32_t *tokens, const int32_t *pos,
                const float *hidden_in, int T, int kv_head, int n_kv, const uint8_t *vis,
                float *logits, float *feat);

/* ---- speculative step pieces (R/examples/speculative/speculative-eagle.cpp) ---- */
/* ancestor-or-self closure from parent[] (parent[0] = -1): vis_tree[T][T] */
void eo_tree_vis(const int32_t *parent, int T, uint8_t *vis_tree);
/* greedy accept walk (:395-449): best[t] = argmax of target logits at node t.  Starting at the
 * root, follow the child whose token equals best[cur]; children are tried in node order.
 * Returns n_acc = number of accepted nodes including the root; path[] holds their node ids;
 * *bonus = best[last accepted node]. */
int  eo_accept_walk(const int32_t *tokens, const int32_t *parent, int T, const int32_t *best,
                    int32_t *path, int32_t *bonus);
/* KV compaction (:459-472 + defrag R/src/llama.cpp:1384-1450): move cells kv_head+path[i] to
 * kv_head+i for every layer. */
void eo_kv_compact(eo_kv *kv, int kv_head, const int32_t *path, int n_acc);

#ifdef __cplusplus
}
#endif
#endif
