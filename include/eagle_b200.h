/*
 * eagle_b200.h — C-ABI of libeagle_b200.so: the B200 (sm_100a) implementation of the EAGLE
 * speculative-decoding hot path of mkjsym/EAGLE-in-llama.cpp.
 *
 * Plain pointers and sizes only; no torch / ggml types.  Two layers:
 *   (1) operator entry points — one per ggml op on the path (what a ggml backend's
 *       graph_compute dispatches to; INTEGRATION.md shows the reference-side stub that binds
 *       them behind the reference's ggml_backend_i vtable);
 *   (2) the fused speculative-step engine (model upload, tree-verify forward, draft-head
 *       forward + tree expansion, accept walk, KV compaction) with no host round trip inside
 *       a round.
 * Pointers named d_* are device pointers, h_* are host pointers.  `stream` is a cudaStream_t
 * passed as void* (NULL = the library's own stream).  Every function returns 0 on success or a
 * negative eb_status; eb_last_error() gives the message.  There is no CPU fallback: every entry
 * point fails with EB_ERR_CUDA when no sm_100 device is usable.
 *
 * R = /root/reference/llama.cpp/ in the citations below.
 */
#ifndef EAGLE_B200_H
#define EAGLE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EB_API __attribute__((visibility("default")))

typedef enum { EB_OK = 0, EB_ERR_ARG = -1, EB_ERR_CUDA = -2, EB_ERR_UNSUPPORTED = -3, EB_ERR_OOM = -4 } eb_status;

/* numeric values of enum ggml_type, R/ggml/include/ggml.h:351-390 */
typedef enum { EB_F32 = 0, EB_F16 = 1, EB_Q4_0 = 2, EB_Q8_0 = 8, EB_Q4_K = 12, EB_Q6_K = 14 } eb_type;

EB_API const char *eb_last_error(void);
EB_API int  eb_version(void);
EB_API int  eb_device_init(int device);                 /* selects the device, checks sm_100 */
EB_API int  eb_device_sm_count(void);
/* ggml_backend_device_i::get_memory, R/ggml/src/ggml-backend-impl.h:143 (CUDA: ggml-cuda.cu:2944) */
EB_API int  eb_device_memory(size_t *free_bytes, size_t *total_bytes);
EB_API size_t eb_row_size(int type, int64_t k);         /* ggml_row_size, R/ggml/src/ggml.c:1183 */

/* ---- device memory (ggml_backend_buffer_i alloc/set/get, R/ggml/src/ggml-backend-impl.h:41-72) */
EB_API int eb_malloc(void **d_ptr, size_t bytes);
EB_API int eb_free(void *d_ptr);
EB_API int eb_memset(void *d_ptr, int value, size_t bytes, void *stream);
EB_API int eb_upload(void *d_dst, const void *h_src, size_t bytes, void *stream);    /* set_tensor */
EB_API int eb_download(void *h_dst, const void *d_src, size_t bytes, void *stream);  /* get_tensor */
EB_API int eb_synchronize(void *stream);
/* One stream per execution context and stream-ordered copies / events between them: what ggml_backend_i asks of an
 * asynchronous backend (R/ggml/src/ggml-backend-impl.h:87-124; CUDA backend: set/get_tensor_async
 * R/ggml/src/ggml-cuda/ggml-cuda.cu:2390-2440, event_record / event_wait :2790-2815, event_new / free / synchronize
 * :3161-3198).  Streams are non-blocking (no implicit ordering with the NULL stream). */
EB_API int eb_malloc_host(void **h_ptr, size_t bytes);   /* pinned: the host buffer type of the device (ggml-cuda.cu:1048-1108) */
EB_API int eb_free_host(void *h_ptr);
EB_API int eb_stream_create(void **stream);
EB_API int eb_stream_destroy(void *stream);
EB_API int eb_device_synchronize(void);
EB_API int eb_upload_async(void *d_dst, const void *h_src, size_t bytes, void *stream);
EB_API int eb_download_async(void *h_dst, const void *d_src, size_t bytes, void *stream);
EB_API int eb_event_create(void **event);
EB_API int eb_event_destroy(void *event);
EB_API int eb_event_record(void *event, void *stream);
EB_API int eb_stream_wait_event(void *stream, void *event);
EB_API int eb_event_synchronize(void *event);
/* device-to-device copy: GGML_OP_CPY / CONT / DUP of a contiguous same-type tensor and buffer_i::cpy_tensor
 * (R/ggml/src/ggml-cuda/cpy.cu:526, ggml-cuda.cu:603) */
EB_API int eb_copy(void *d_dst, const void *d_src, size_t bytes, void *stream);

/* ---- operator layer ------------------------------------------------------------------ */
/* Device layout of quantised weights.  Q4_K blocks (16-byte aligned) stay in GGUF order; the
 * 2-byte-aligned GGUF formats (Q4_0 18 B, Q8_0 34 B, Q6_K 210 B blocks, R/ggml/src/ggml-common.h:
 * 161-327) are split into aligned planes over the whole tensor so that every kernel load is a
 * full-width aligned vector load.  Same byte count, lossless; eb_unpack restores the GGUF bytes
 * (this is what a backend's set_tensor / get_tensor do: R/ggml/src/ggml-backend-impl.h:50-52).
 * All operator entry points below take weights in DEVICE layout. */
EB_API int eb_repack(int type, const void *d_gguf, void *d_dev, int64_t n_elems, void *stream);
EB_API int eb_unpack(int type, const void *d_dev, void *d_gguf, int64_t n_elems, void *stream);
/* from_float of the CPU mul_mat (quantize_row_q8_K R/ggml/src/ggml-quants.c:2479 when kq = 1,
 * quantize_row_q8_0 R/ggml/src/ggml-cpu/ggml-cpu-quants.c:738 when kq = 0), output as planes:
 * qs int8 [T][K]; d f32 [T][K/256 | K/32]; bsum f16 [T][K/16] (kq = 1 only).  Replaces
 * quantize_q8_1 R/ggml/src/ggml-cuda/quantize.cu:4. */
EB_API int eb_quantize_act_planes(int kq, const float *d_x, int64_t K, int T, int8_t *d_qs, float *d_d,
                                  void *d_bsum_f16, void *stream);
/* to_float of a quantised tensor: replaces dequantize_block_* R/ggml/src/ggml-cuda/convert.cu:71,190,255
 * (CPU: dequantize_row_* R/ggml/src/ggml-quants.c:255,349,1280,1690).  Bit-exact. */
EB_API int eb_dequantize(int type, const void *d_src, float *d_dst, int64_t n_elems, void *stream);

/* GGML_OP_MUL_MAT: d_y[t*N + n] = sum_k W[n][k] * d_x[t*K + k].  Replaces mul_mat_vec_q
 * (R/ggml/src/ggml-cuda/mmvq.cu:49-200, T <= 8) and mul_mat_q (R/ggml/src/ggml-cuda/mmq.cuh:2502-2822,
 * T > 8).  Activations are quantised on the CPU backend's grid (Q8_K for K-quants, Q8_0 for
 * Q4_0/Q8_0: R/ggml/src/ggml-cpu/ggml-cpu.c:254-419,7588-7623) so integer dots equal the oracle's.
 * d_ws: scratch of at least eb_mul_mat_ws_size() bytes. */
EB_API size_t eb_mul_mat_ws_size(int type, int64_t N, int64_t K, int64_t T);
EB_API int eb_mul_mat(int type, const void *d_W, int64_t N, int64_t K, const float *d_x, int64_t T,
                      float *d_y, void *d_ws, void *stream);
/* the same with the activation quantisation shared between consecutive calls on one workspace: reuse_activation != 0 says that the
 * previous eb_mul_mat / eb_mul_mat_ex call quantised this very d_x (same K, same T <= 128, same grid: K-quants share Q8_K, Q4_0 / Q8_0
 * share Q8_0) and nothing else used d_ws since.  wq / wk / wv and ffn_gate / ffn_up read one tensor each: the reference's CPU path
 * quantises src1 once per MUL_MAT node too, into the same values (R/ggml/src/ggml-cpu/ggml-cpu.c:7590-7620), so results are unchanged. */
EB_API int eb_mul_mat_ex(int w_type, const void *d_W, int64_t N, int64_t K, const float *d_x, int64_t T, float *d_y, void *d_ws,
                         int reuse_activation, void *stream);

/* GGML_OP_RMS_NORM (+ optional GGML_OP_MUL by a weight row): R/ggml/src/ggml-cuda/norm.cu:108 */
EB_API int eb_rms_norm(const float *d_x, const float *d_weight_or_null, float *d_y, int64_t n,
                       int64_t rows, float eps, void *stream);
/* GGML_OP_ROPE, mode NORM, in place on [T][n_head][D]: R/ggml/src/ggml-cuda/rope.cu:41 */
EB_API int eb_rope_norm(float *d_x, const int32_t *d_pos, int D, int n_head, int T, float freq_base, void *stream);
/* out of place: dst = rope(src); dst may equal src */
EB_API int eb_rope_norm_to(const float *d_src, float *d_dst, const int32_t *d_pos, int D, int n_head, int T, float freq_base, void *stream);
/* GGML_OP_GET_ROWS with dequantisation: R/ggml/src/ggml-cuda/getrows.cu:5-148 */
EB_API int eb_get_rows_n(int type, const void *d_W, int64_t n_rows, int64_t K, const int32_t *d_ids, int n, float *d_out, void *stream);
/* elementwise: R/ggml/src/ggml-cuda/binbcast.cu:26, unary.cu:45,201.  b is broadcast over rows
 * when nb < n (nb must divide n). */
EB_API int eb_add(const float *d_a, const float *d_b, float *d_y, int64_t n, int64_t nb, void *stream);
EB_API int eb_mul(const float *d_a, const float *d_b, float *d_y, int64_t n, int64_t nb, void *stream);
EB_API int eb_silu(const float *d_x, float *d_y, int64_t n, void *stream);
/* GGML_OP_SCALE: y = x * s (R/ggml/src/ggml-cuda/scale.cu) */
EB_API int eb_scale(const float *d_x, float *d_y, int64_t n, float s, void *stream);
EB_API int eb_relu(const float *d_x, float *d_y, int64_t n, void *stream);
/* GGML_OP_CONCAT along dim 0 of two f32 [E][T] tensors -> [2E][T] (build_eagle's [embd | hidden],
 * R/src/llama.cpp:1893; R/ggml/src/ggml-cuda/concat.cu:4) */
EB_API int eb_concat2(const float *d_a, const float *d_b, float *d_y, int E, int T, void *stream);
/* GGML_OP_CPY f32 -> f16 (KV store), R/ggml/src/ggml-cuda/cpy.cu:35 */
EB_API int eb_cpy_f32_f16(const float *d_x, void *d_y_f16, int64_t n, void *stream);

/* ---- FA-off attention block (llm_build_kqv without flash attention, R/src/llama.cpp:762-812 — the reference's default unless -fa;
 * off the engine's path, which always runs the tree flash-attention kernel).  Strides are in BYTES, as ggml's nb[]. */
/* GGML_OP_CPY / CONT / DUP over arbitrary 4-D strides, f32 -> f32 | f16, f16 -> f16 (R/ggml/src/ggml-cuda/cpy.cu:526;
 * CPU: ggml_compute_forward_dup, R/ggml/src/ggml-cpu/ggml-cpu.c:4113): element i of the source's logical order goes to
 * element i of the destination's logical order (ne / dne may differ in shape, not in count). */
EB_API int eb_copy_strided(int src_type, int dst_type, const void *d_src, void *d_dst, const int64_t ne[4], const int64_t nb_src[4],
                           const int64_t dne[4], const int64_t nb_dst[4], void *stream);
/* GGML_OP_MUL_MAT with an f16 / f32 src0 and a batch dimension (KQ = K x Q and KQV = V^T x softmax): y[b][t][n] = sum_k
 * a[b / (B / a_batches)][n][k] * x[b][t][k]; x is rounded to f16 first when a is f16, products accumulate in f32, as the CPU's
 * vec_dot_f16 path does (R/ggml/src/ggml-cpu/ggml-cpu.c:7588-7623).  a: rows nb1 apart, batches nb2 apart; y contiguous [B][T][N]. */
EB_API int eb_mul_mat_dense(int a_type, const void *d_a, int64_t a_nb1, int64_t a_nb2, int64_t a_batches, const float *d_x, int64_t x_nb1,
                            int64_t x_nb2, float *d_y, int64_t N, int64_t K, int64_t T, int64_t B, void *stream);
/* GGML_OP_SOFT_MAX (ggml_soft_max_ext with max_bias = 0, R/ggml/src/ggml-cuda/softmax.cu:13; CPU R/ggml/src/ggml-cpu/ggml-cpu.c:9042):
 * y = softmax(scale * x + mask) over rows of n; mask (f16 or f32, rows mask_stride elements apart, or NULL) row = row % rows_per_batch */
EB_API int eb_soft_max_ext(const float *d_x, const void *d_mask, int mask_type, int64_t mask_stride, float *d_y, int64_t n, int64_t rows,
                           int64_t rows_per_batch, float scale, void *stream);

/* GGML_OP_FLASH_ATTN_EXT: replaces flash_attn_ext_f16 R/ggml/src/ggml-cuda/fattn-mma-f16.cuh:806
 * (CPU: R/ggml/src/ggml-cpu/ggml-cpu.c:10999).  q f32 [T][H][D]; k, v f16 cells, row stride
 * kv_stride elements between cells, head h at offset h*D; out f32 [T][H][D].  D must be 128.
 * Mask, one of:  d_mask_f16 [T_pad][n_kv] (ggml layout, R/ggml/include/ggml.h:1778-1793), or
 * the compact tree form: cells [0,n_prefix) visible to all rows, cell n_prefix+j visible to row t
 * iff bit j of d_tree_bits[t*4 .. t*4+3] (128 bits per row) is set; other cells invisible. */
EB_API size_t eb_flash_attn_ws_size(int T, int H, int D, int n_kv);
EB_API int eb_flash_attn(const float *d_q, const void *d_k, const void *d_v, int64_t kv_stride,
                         const void *d_mask_f16, int64_t mask_stride,
                         const uint32_t *d_tree_bits, int n_prefix,
                         int T, int H, int D, int n_kv, float scale, float *d_out, void *d_ws, void *stream);

/* The same operator on the tcgen05 kernel for ANY mask whose values are 0 or -inf (what llama_set_inputs writes,
 * R/src/llama-context.cpp:136-201): the f16 mask is first turned into one bit per (row, cell) plus a per-tile summary
 * (eb_mask_to_bits, once per mask: every layer of a graph shares it), then eb_flash_attn_bits runs the attention for T of
 * any size (128-row chunks).  *h_general (optional, host) = 1 when some mask value is neither 0 nor -inf (ALiBi slopes, the
 * random masks of R/tests/test-backend-ops.cpp): the bit form does not apply and the caller uses eb_flash_attn.  Reading
 * it synchronises the stream once.  d_bits: eb_mask_bits_size(T, n_kv) bytes.  kv_stride must be a multiple of 8. */
EB_API size_t eb_mask_bits_size(int T, int n_kv);
EB_API int eb_mask_to_bits(const void *d_mask_f16, int64_t mask_stride, int T, int n_kv, void *d_bits, int *h_general, void *stream);
EB_API int eb_flash_attn_bits_supported(int T, int H, int D, int64_t kv_stride);
EB_API int eb_flash_attn_bits(const float *d_q, const void *d_k, const void *d_v, int64_t kv_stride, const void *d_bits,
                              int T, int H, int D, int n_kv, float scale, float *d_out, void *d_ws, void *stream);

/* GGML_OP_ARGMAX / top-k (R/ggml/src/ggml-cuda/argmax.cu:8, argsort.cu:11; host sampler
 * R/src/llama-sampling.cpp:228-302): per row top-k by logit, ties to the lower id; probs =
 * softmax over the k kept logits. k <= 16.  eb_top_k keeps one scratch list per stream (created on first use): calls on
 * different streams may overlap, calls on one stream are ordered anyway (sessions own their scratch). */
EB_API int eb_top_k(const float *d_logits, int rows, int V, int k, int32_t *d_ids, float *d_vals, float *d_probs, void *stream);
EB_API int eb_argmax(const float *d_logits, int rows, int V, int32_t *d_ids, void *stream);
/* GGML_OP_ARGSORT (ggml_argsort, and ggml_top_k = its first k columns; R/ggml/src/ggml-cuda/argsort.cu:11): per row of n <= 1024 values
 * the indices in ascending (descending != 0: descending) order of value; equal values keep index order. */
EB_API int eb_argsort(const float *d_x, int rows, int n, int descending, int32_t *d_ids, void *stream);
/* eb_top_k without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same results as
 * eb_top_k (the 16-slice kernel); its scratch is a file-scope `static __device__` array of EB_TOP_K_NOALLOC_MAX_ROWS rows
 * (66 048 bytes, held by every GPU that loads the library until exit).  Limits: rows <= EB_TOP_K_NOALLOC_MAX_ROWS, 1024 <= V <=
 * 32768, k <= 16; anything else is refused with EB_ERR_ARG.  The device must have been set up with eb_device_init() beforehand
 * (this call sets nothing up and calls no cudaMalloc / cudaFree).  All calls on a device share the array: calls that overlap on
 * the GPU (different streams) race — serialise them. */
#define EB_TOP_K_NOALLOC_MAX_ROWS 32
EB_API int eb_top_k_noalloc(const float *d_logits, int rows, int V, int k, int32_t *d_ids, float *d_vals, float *d_probs, void *stream);

/* ---- engine layer -------------------------------------------------------------------- */
typedef struct eb_model eb_model;      /* weights of one GGUF model resident in HBM */
typedef struct eb_session eb_session;  /* target + head + both KV caches + tree state */

/* tensor slots (R/src/llama-arch.cpp:1301-1326 names in comments) */
enum { EB_T_TOK_EMBD = 0 /*token_embd*/, EB_T_OUTPUT_NORM /*output_norm*/, EB_T_OUTPUT /*output*/,
       EB_T_FC /*fc.weight*/, EB_T_FC_BIAS /*fc.bias*/, EB_T_GLOBAL_COUNT };
enum { EB_L_ATTN_NORM = 0, EB_L_WQ, EB_L_WK, EB_L_WV, EB_L_WO, EB_L_FFN_NORM, EB_L_GATE, EB_L_UP,
       EB_L_DOWN, EB_L_COUNT };

/* hparams: R/src/llama-model.cpp:410-487.  is_eagle selects build_eagle (R/src/llama.cpp:1839). */
EB_API eb_model *eb_model_create(int E, int H, int F, int V, int L, float eps, float freq_base, int is_eagle);
EB_API void      eb_model_free(eb_model *m);
/* upload one tensor in its GGUF block layout (host pointer); layer = -1 for globals */
EB_API int eb_model_set_tensor(eb_model *m, int layer, int which, int type, const void *h_data);
/* fill one tensor with synthetic random blocks directly on the device (bench fixtures) */
EB_API int eb_model_synth_tensor(eb_model *m, int layer, int which, int type, uint64_t seed, float scale);
/* read a tensor back in GGUF block layout (get_tensor) */
EB_API int eb_model_get_tensor(const eb_model *m, int layer, int which, void *h_data, size_t bytes);
EB_API int eb_model_share_lm_head(eb_model *head, const eb_model *target);   /* R/src/llama.cpp:2016 */
EB_API size_t eb_model_weight_bytes(const eb_model *m);   /* bytes one forward streams */
/* hp = {E, H, F, V, L, is_eagle}, hf = {rms eps, rope freq base} */
EB_API int eb_model_get_hparams(const eb_model *m, int32_t *hp, float *hf);
/* GGUF file -> resident model.  Replaces, for the tensors of this path, gguf_init_from_file (R/ggml/src/gguf.cpp) +
 * llama_model_loader's per-tensor ggml_backend_tensor_set (R/src/llama-model-loader.cpp): the header is parsed here (GGUF v2 / v3;
 * arch "llama" = target, "eagle" = draft head, tensor names of R/src/llama-arch.cpp:1301-1326), the data section streams through two
 * page-locked buffers into the final allocations (plane-layout types through a device staging area + repack on the stream).
 * An eagle file without output.weight is completed with eb_model_share_lm_head.  seconds / bytes (optional) report the upload.
 * eb_gguf_probe reads the header only (no device needed): tensors of this path and their bytes. */
EB_API int eb_model_load_gguf(const char *path, eb_model **out, double *seconds, uint64_t *bytes);
EB_API int eb_gguf_probe(const char *path, int32_t *hp, float *hf, int32_t *n_tensors_used, uint64_t *bytes_used);

/* ---- tensor parallelism of the target's verify forward (BASELINE config 4).  Replaces the reference's --split-mode row
 * (ggml_backend_cuda_split_buffer_type / ggml_cuda_op_mul_mat, R/ggml/src/ggml-cuda/ggml-cuda.cu:735-748,1359-1668: weight rows
 * split across GPUs, partial rows gathered to the main device by cudaMemcpyPeerAsync) with N-split wq|wk|wv|gate|up + K-split
 * wo|down and ONE peer-memory all-reduce kernel per K-split output (csrc/tp.cu), one process per GPU.
 * eb_model_create_tp: this rank's shard (n_head_local heads of size D = 128, F_local ffn rows, both * may be uneven across ranks,
 * multiples of 256 elements); tensors are set with eb_model_set_tensor in the shard's own shapes: wq/wk/wv [n_head_local*D][E],
 * wo [E][n_head_local*D], gate/up [F_local][E], down [E][F_local]; embeddings, norms and output are replicated.
 * eb_tp_create / _handle / _open_peer: the communicator; the host program carries the 64-byte handles between the processes.
 * eb_session_set_tp attaches it to a session whose target is a shard model; every rank must then make the same forward calls. */
typedef struct eb_tp eb_tp;
EB_API eb_model *eb_model_create_tp(int E, int D, int n_head_local, int F_local, int V, int L, float eps, float freq_base);
EB_API eb_tp *eb_tp_create(int rank, int size, int64_t max_elems);
EB_API int  eb_tp_handle(eb_tp *tp, void *out64);
EB_API int  eb_tp_open_peer(eb_tp *tp, int peer_rank, const void *handle64);
EB_API void eb_tp_free(eb_tp *tp);
EB_API int  eb_tp_error(eb_tp *tp);      /* 1 after a wait for a peer timed out */
/* out[i] = sum over ranks of sum over the ks planes [ks][n] of every rank; identical bits on all ranks; n % 4 == 0 */
EB_API int  eb_tp_allreduce(eb_tp *tp, const float *d_planes, int ks, int64_t n, float *d_out, void *stream);
/* device time of one all-reduce, measured over `iters` (even) calls replayed from one CUDA graph (host launch overhead excluded);
 * engine_sets != 0 alternates the engine's slot sets, else the operator's.  Collective: all ranks call it together. */
EB_API int  eb_tp_bench(eb_tp *tp, const float *d_planes, int ks, int64_t n, float *d_out, int engine_sets, int iters, float *us_per_call);

typedef struct {
    int n_ctx;          /* KV cells per cache */
    int max_nodes;      /* verify batch capacity incl. root (<= 128) */
    int n_branch;       /* -np: max simultaneous draft branches (<= 16) */
    int n_draft;        /* --draft-max: stop a branch once the verify batch exceeds it */
    int max_depth;      /* draft levels captured in the graph */
    int top_k;          /* draft candidates per node (reference: 10, R/common/speculative.cpp:47) */
    float p_split;      /* --draft-p-split, R/common/common.h:183 */
    int fixed_accept;   /* >= 0: accept exactly this many draft nodes per round (bench mode) */
    int use_graph;      /* capture the round in a CUDA graph */
    float p_min;        /* --draft-p-min (R/common/speculative.cpp:283, default 0.75 R/common/common.h:184): a branch stops drafting after a
                         * token whose draft probability (softmax over the kept top-k) is below p_min; 0 = off, as in the tree driver */
} eb_session_params;

EB_API eb_session *eb_session_create(eb_model *target, eb_model *head, const eb_session_params *p);
EB_API void        eb_session_free(eb_session *s);

/* One forward of `model` (0 = target, 1 = head) over T tokens with tree structure parent[]
 * (parent[i] < i, -1 = attends the committed prefix only), appended at the cache tail without
 * committing.  Host in / host out; used by parity tests and by prefill.  h_hidden [T][E] only
 * for the head.  h_logits [T][V] / h_feat [T][E] may be NULL. */
EB_API int eb_session_forward(eb_session *s, int model, const int32_t *h_tokens, const int32_t *h_pos,
                              const int32_t *h_parent, const float *h_hidden, int T,
                              float *h_logits, float *h_feat);
/* The same forward driven by a tree-shaped llama_batch (R/include/llama.h:247-257; built by common_batch_add,
 * R/common/common.cpp:1784, with the seq-id sets of R/examples/speculative/speculative-eagle.cpp:556-625): token i goes to cell
 * n_past + i, and token j sees it iff token i's seq-id set contains token j's first seq id and pos[i] <= pos[j] — the rule of
 * the reference's KQ-mask builder, R/src/llama-context.cpp:163-172.  eb_tree_bits_from_batch is that rule alone (pure host code:
 * bits_out[j*4 ..] = 128-bit visibility set of row j, the compact mask form of eb_flash_attn). */
EB_API int eb_tree_bits_from_batch(int n_tokens, const int32_t *h_pos, const int32_t *h_n_seq_id, const int32_t *const *h_seq_id, uint32_t *bits_out);
EB_API int eb_session_forward_batch(eb_session *s, int model, const int32_t *h_tokens, const int32_t *h_pos, const int32_t *h_n_seq_id,
                                    const int32_t *const *h_seq_id, const float *h_hidden, int T, float *h_logits, float *h_feat);
/* Activation tap — the counterpart of the reference's per-node eval callback (llama_context_params.cb_eval ->
 * ggml_backend_sched_set_eval_callback, R/src/llama.cpp:9562; demo R/examples/eval-callback/eval-callback.cpp:161): while a tap
 * buffer is set, every eb_session_forward / _forward_batch appends the int8 activation plane [T][K] of each quantisation site
 * (per layer: input of wq|wk|wv, of wo, of gate|up, of ffn_down; then the lm_head input; the head's fc input first) to d_buf.
 * d_buf = NULL turns it off.  eb_session_tap_state reports bytes / sites written since set_tap.  Not used inside eb_session_round. */
EB_API int eb_session_set_tp(eb_session *s, eb_tp *tp);
EB_API int eb_session_set_tap(eb_session *s, void *d_buf, size_t bytes);
EB_API int eb_session_tap_state(const eb_session *s, size_t *bytes_written, int *n_sites);
/* commit the first n of the T tokens of the last forward of `model` (chain prefill) */
EB_API int eb_session_commit(eb_session *s, int model, int n);
EB_API int eb_session_n_past(const eb_session *s, int model);
/* Prompt prefill of any length — llama_decode_init + llama_decode_initial of the chain driver (R/examples/speculative-simple/
 * speculative-eagle.cpp:150,194; R/src/llama.cpp:10512,10256), SURVEY §8(f) N1: the tokens go through the target in chunks of
 * one verify tile and through the draft head with the previous token's target feature as hidden input; both caches are committed.
 * h_logits_last [V] / h_feat_last [E]: logits and feature (result_norm) of the LAST prompt token (either may be NULL). */
EB_API int eb_session_prefill(eb_session *s, const int32_t *h_tokens, int n, float *h_logits_last, float *h_feat_last);
/* KV cache rows as raw f16 [n_cells][E] of one layer (which: 0 = K, 1 = V; model: 0 = target, 1 = head): what the reference's
 * session save / load does through ggml_backend_tensor_get / _set on the cache tensors (R/src/llama-context.cpp:1229-2190).
 * eb_session_set_n_past declares how many leading cells are the committed prefix (after writing a cache). */
EB_API int eb_session_kv_write(eb_session *s, int model, int layer, int which, int cell0, int n_cells, const void *h_f16);
EB_API int eb_session_kv_read(eb_session *s, int model, int layer, int which, int cell0, int n_cells, void *h_f16);
EB_API int eb_session_set_n_past(eb_session *s, int n_past);
/* fill both KV caches with n_past cells of synthetic N(0,1) K/V (bench: 4K context) */
EB_API int eb_session_synth_kv(eb_session *s, int n_past, uint64_t seed);
/* start decoding: last committed token id + the target feature (result_norm) that produced it */
EB_API int eb_session_set_root(eb_session *s, int32_t id_last, const float *h_feat_or_null);

typedef struct {
    int n_nodes;        /* verify batch size incl. root */
    int n_accept;       /* accepted nodes incl. root */
    int32_t bonus;      /* target argmax at the last accepted node = next round's root */
    int32_t tokens[128];   /* tree tokens (node order) */
    int32_t parent[128];
    int32_t path[128];     /* accepted node ids */
    float   gpu_ms;        /* device time of the round (CUDA events) */
} eb_round_result;

/* One speculative round entirely on the device: draft-head forward(s) + top-k tree expansion,
 * tree-verify forward of the target, greedy accept walk, KV compaction of both caches, feature
 * hand-off.  Only h_out crosses the bus (one small D2H). */
EB_API int eb_session_round(eb_session *s, eb_round_result *h_out);
/* verify forward alone on the current tree (micro-bench: device time in *gpu_ms) */
/* The draft half alone: what common_speculative_gen_draft hands to its caller (R/common/speculative.cpp:137-299).  Runs the head
 * forwards, top-k and expansion of the session's configuration from the current root and returns the drafted tree (node 0 = the root
 * token id_last, h_parent[0] = -1); with n_branch = 1 this is the reference's chain draft, p_min stop included.  Nothing is
 * committed; a following eb_session_round() drafts again from the same state. */
EB_API int eb_session_gen_draft(eb_session *s, int32_t *h_tokens, int32_t *h_parent, int cap, int *n_nodes);
EB_API int eb_session_verify_only(eb_session *s, int T, float *gpu_ms);
/* micro-bench of one hot kernel class (GEMM for T > 8, matvec for T <= 8): average device ms per
 * launch over `iters` launches cycling through n_bufs distinct weight buffers (HBM-resident,
 * larger than L2 in total), CUDA events on the launching stream */
EB_API int eb_bench_mul_mat_kernel(int type, int64_t N, int64_t K, int T, int n_bufs, int iters, float *ms_per_launch);
/* same for GGML_OP_FLASH_ATTN_EXT with the compact tree mask (chain tree over the last T cells; n_bufs distinct K/V caches) */
EB_API int eb_bench_flash_attn(int T, int H, int D, int n_kv, int n_bufs, int iters, float *ms_per_call);
/* launches of this library's kernels since the last call (bench "gpu_launches") */
EB_API long eb_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif
