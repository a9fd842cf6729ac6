// ggml-b200.cpp — the reference-side binding: a ggml backend ("B200") whose graph_compute
// dispatches the operators of the EAGLE speculative-step path to libeagle_b200.so.
//
// This is the one translation unit INTEGRATION.md §A describes.  It is compiled against the
// reference's own headers where they lie (R/ggml/include, R/ggml/src/ggml-backend-impl.h; nothing
// is copied into this repo) and speaks only the C-ABI of include/eagle_b200.h to the kernels.
// It exports `ggml_backend_init` (GGML_BACKEND_DL_IMPL, R/ggml/src/ggml-backend-impl.h:222-251),
// so the reference loads it with `ggml_backend_load(path)` / GGML_BACKEND_PATH
// (R/ggml/src/ggml-backend-reg.cpp:220-263,577-581) exactly like libggml-cuda.so; ggml-base symbols
// (ggml_nbytes, ggml_backend_buffer_init, ...) are left undefined and resolved by the host process.
//
// Vtables implemented (all of R/ggml/src/ggml-backend-impl.h): ggml_backend_reg_i,
// ggml_backend_device_i, ggml_backend_buffer_type_i, ggml_backend_buffer_i, ggml_backend_i.
// Counterparts in the backend this replaces: R/ggml/src/ggml-cuda/ggml-cuda.cu:480-700 (buffers),
// :2096-2400 (compute_forward), :2693 (graph_compute), :2938-3306 (device), :3310-3420 (reg).
//
// Quantised weights live in the library's DEVICE layout (eagle_b200.h: Q4_K verbatim, Q6_K / Q4_0 /
// Q8_0 as aligned planes): set_tensor repacks, get_tensor unpacks — bit-exact GGUF bytes both ways.
// No CPU fallback: ops this file does not list are refused in supports_op, so ggml_backend_sched
// keeps them on whatever other backend the host has; a failing launch fails the graph.
#define GGML_BACKEND_DL
#define GGML_BACKEND_SHARED
#define GGML_BACKEND_BUILD
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"

#include "eagle_b200.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#define B200_NAME "B200"

static void b200_log_err(const char *what) { fprintf(stderr, "ggml-b200: %s: %s\n", what, eb_last_error()); }
#define B200_OK(call) do { if ((call) != EB_OK) { b200_log_err(#call); GGML_ABORT("ggml-b200: device call failed"); } } while (0)

static bool b200_repacked(ggml_type t) { return t == GGML_TYPE_Q6_K || t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0; }
static bool b200_weight_type(ggml_type t) { return t == GGML_TYPE_Q4_K || b200_repacked(t); }

// ---------------------------------------------------------------------------------- buffer
struct b200_buffer_ctx { void *base = nullptr; size_t size = 0; };

static void b200_buf_free(ggml_backend_buffer_t b) { auto *c = (b200_buffer_ctx *)b->context; eb_free(c->base); delete c; }
static void *b200_buf_base(ggml_backend_buffer_t b) { return ((b200_buffer_ctx *)b->context)->base; }
static void b200_buf_init_tensor(ggml_backend_buffer_t, ggml_tensor *t) {
    // a strided / partial view of a plane-repacked weight cannot be addressed; the path never builds one
    if (t->view_src && b200_repacked(t->type) && !(t->view_offs == 0 && ggml_nbytes(t) == ggml_nbytes(t->view_src)))
        GGML_ABORT("ggml-b200: partial views of %s tensors are not supported", ggml_type_name(t->type));
}
// The buffer interface is blocking and knows no backend: compute runs on per-backend non-blocking streams, so these calls first
// wait for the whole device (ggml_backend_cuda_buffer_set_tensor does the same with cudaStreamPerThread + synchronize, ggml-cuda.cu:571-590)
static void b200_buf_memset_tensor(ggml_backend_buffer_t, ggml_tensor *t, uint8_t v, size_t off, size_t size) {
    B200_OK(eb_device_synchronize());
    B200_OK(eb_memset((char *)t->data + off, v, size, nullptr)); B200_OK(eb_synchronize(nullptr));
}
// host -> device, blocking (ggml_backend_cuda_buffer_set_tensor, ggml-cuda.cu:571)
// GGUF-layout staging area of the plane-layout types: one grow-only device allocation for the process (a model load is ~100
// set_tensor calls of such tensors; a cudaMalloc / cudaFree pair per call synchronises the device twice each time)
static std::mutex g_stage_mu;
static void *g_stage = nullptr; static size_t g_stage_size = 0;
static void *b200_stage(size_t need) {
    if (need > g_stage_size) { if (g_stage) eb_free(g_stage); g_stage = nullptr; g_stage_size = 0; if (eb_malloc(&g_stage, need) != EB_OK) return nullptr; g_stage_size = need; }
    return g_stage;
}
static void b200_buf_set_tensor(ggml_backend_buffer_t, ggml_tensor *t, const void *data, size_t off, size_t size) {
    B200_OK(eb_device_synchronize());
    if (!b200_repacked(t->type)) { B200_OK(eb_upload((char *)t->data + off, data, size, nullptr)); return; }
    const size_t total = ggml_nbytes(t);
    std::lock_guard<std::mutex> lk(g_stage_mu);
    void *stage = b200_stage(total);
    if (!stage) { b200_log_err("set_tensor: staging"); GGML_ABORT("ggml-b200: out of device memory"); }
    if (off != 0 || size != total) B200_OK(eb_unpack((int)t->type, t->data, stage, ggml_nelements(t), nullptr));   // read-modify-write
    B200_OK(eb_upload((char *)stage + off, data, size, nullptr));
    B200_OK(eb_repack((int)t->type, stage, t->data, ggml_nelements(t), nullptr));
    B200_OK(eb_synchronize(nullptr));
}
static void b200_buf_get_tensor(ggml_backend_buffer_t, const ggml_tensor *t, void *data, size_t off, size_t size) {
    B200_OK(eb_device_synchronize());
    if (!b200_repacked(t->type)) { B200_OK(eb_download(data, (const char *)t->data + off, size, nullptr)); return; }
    std::lock_guard<std::mutex> lk(g_stage_mu);
    void *stage = b200_stage(ggml_nbytes(t));
    if (!stage) { b200_log_err("get_tensor: staging"); GGML_ABORT("ggml-b200: out of device memory"); }
    B200_OK(eb_unpack((int)t->type, t->data, stage, ggml_nelements(t), nullptr));
    B200_OK(eb_download(data, (const char *)stage + off, size, nullptr));
}
static ggml_backend_buffer_type_t b200_buffer_type();
static bool b200_buf_cpy_tensor(ggml_backend_buffer_t, const ggml_tensor *src, ggml_tensor *dst) {
    if (!src->buffer || src->buffer->buft != b200_buffer_type() || !ggml_is_contiguous(src) || !ggml_is_contiguous(dst)) return false;
    B200_OK(eb_device_synchronize());
    B200_OK(eb_copy(dst->data, src->data, ggml_nbytes(src), nullptr)); B200_OK(eb_synchronize(nullptr));
    return true;
}
static void b200_buf_clear(ggml_backend_buffer_t b, uint8_t v) {
    auto *c = (b200_buffer_ctx *)b->context; B200_OK(eb_device_synchronize()); B200_OK(eb_memset(c->base, v, c->size, nullptr)); B200_OK(eb_synchronize(nullptr));
}
static const ggml_backend_buffer_i b200_buffer_iface = {
    b200_buf_free, b200_buf_base, b200_buf_init_tensor, b200_buf_memset_tensor, b200_buf_set_tensor, b200_buf_get_tensor,
    b200_buf_cpy_tensor, b200_buf_clear, /* reset */ nullptr,
};

// ---------------------------------------------------------------------------------- buffer type
static const char *b200_buft_name(ggml_backend_buffer_type_t) { return B200_NAME; }
static ggml_backend_buffer_t b200_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    auto *c = new b200_buffer_ctx(); c->size = size;
    if (eb_malloc(&c->base, size) != EB_OK) { b200_log_err("alloc_buffer"); delete c; return nullptr; }    // NULL on OOM, as the contract asks
    return ggml_backend_buffer_init(buft, b200_buffer_iface, c, size);
}
static size_t b200_buft_alignment(ggml_backend_buffer_type_t) { return 256; }     // TMA boxes and 16-byte vector loads
static size_t b200_buft_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor *t) { return (ggml_nbytes(t) + 255) & ~(size_t)255; }
static bool b200_buft_is_host(ggml_backend_buffer_type_t) { return false; }

// ---------------------------------------------------------------------------------- host (pinned) buffer type
// What libllama asks a GPU device for so that inputs and the logits / embeddings output buffer live in page-locked memory
// (llama_output_reserve, R/src/llama-context.cpp:884-958; ggml_backend_cuda_host_buffer_type, ggml-cuda.cu:1048-1108): a CPU buffer over
// cudaMallocHost memory, so stream-ordered copies to and from it are real DMA transfers.  Falls back to a plain CPU buffer when pinning fails.
static const char *b200_host_buft_name(ggml_backend_buffer_type_t) { return B200_NAME "_Host"; }
static void b200_host_buf_free(ggml_backend_buffer_t b) { eb_free_host(b->context); }
static ggml_backend_buffer_t b200_host_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    void *p = nullptr;
    if (getenv("GGML_B200_NO_PINNED") || eb_malloc_host(&p, size) != EB_OK || !p) return ggml_backend_buft_alloc_buffer(ggml_backend_cpu_buffer_type(), size);
    ggml_backend_buffer_t b = ggml_backend_cpu_buffer_from_ptr(p, size);
    b->buft = buft; b->iface.free_buffer = b200_host_buf_free;
    return b;
}
static ggml_backend_dev_t b200_the_device();
static ggml_backend_buffer_type_t b200_host_buffer_type() {
    static ggml_backend_buffer_type t = {
        { b200_host_buft_name, b200_host_buft_alloc, ggml_backend_cpu_buffer_type()->iface.get_alignment, /* get_max_size */ nullptr,
          ggml_backend_cpu_buffer_type()->iface.get_alloc_size, ggml_backend_cpu_buffer_type()->iface.is_host },
        b200_the_device(), nullptr };
    return &t;
}

// ---------------------------------------------------------------------------------- backend
// one ggml_backend_t = one stream (ggml_backend_cuda_context::stream(), R/ggml/src/ggml-cuda/common.cuh): graphs, async copies and
// events of one backend are ordered on it; two backends of the same device overlap
struct b200_backend_ctx {
    void *stream = nullptr;
    void *ws = nullptr; size_t ws_size = 0;      // scratch of eb_mul_mat / eb_flash_attn, grown on demand
    void *scratch(size_t need) {
        if (need > ws_size) { if (ws) { eb_synchronize(stream); eb_free(ws); } ws = nullptr; ws_size = 0; aq_src = nullptr; if (eb_malloc(&ws, need) != EB_OK) return nullptr; ws_size = need; }
        return ws;
    }
    // FLASH_ATTN_EXT: the mask of a graph in bit form (every layer uses the same KQ_mask tensor), valid for one graph_compute
    void *bits = nullptr; size_t bits_size = 0;
    const void *mask_data = nullptr; int mask_T = 0, mask_kv = 0; bool mask_bits_ok = false;
    // MUL_MAT: the activation planes left in the scratch by the previous MUL_MAT (wq / wk / wv and ffn_gate / ffn_up read one tensor
    // each).  Valid until a node writes into that tensor, another op takes the scratch, or the graph ends.
    const char *aq_src = nullptr; size_t aq_bytes = 0; int64_t aq_K = 0, aq_T = 0; int aq_kq = -1; void *aq_ws = nullptr;
    void aq_drop() { aq_src = nullptr; }
};

static const char *b200_be_name(ggml_backend_t) { return B200_NAME; }
static void b200_be_free(ggml_backend_t be) {
    auto *c = (b200_backend_ctx *)be->context;
    eb_synchronize(c->stream);
    if (c->ws) eb_free(c->ws);
    if (c->bits) eb_free(c->bits);
    eb_stream_destroy(c->stream);
    delete c; delete be;
}
static void b200_be_sync(ggml_backend_t be) { B200_OK(eb_synchronize(((b200_backend_ctx *)be->context)->stream)); }
// stream-ordered copies (ggml_backend_cuda_set_tensor_async / get_tensor_async, R/ggml/src/ggml-cuda/ggml-cuda.cu:2390-2410); weights in
// plane layout go through the blocking repack path
static void b200_be_set_tensor_async(ggml_backend_t be, ggml_tensor *t, const void *data, size_t off, size_t size) {
    ggml_backend_buffer_t buf = t->view_src ? t->view_src->buffer : t->buffer;
    if (b200_repacked(t->type)) { buf->iface.set_tensor(buf, t, data, off, size); return; }
    B200_OK(eb_upload_async((char *)t->data + off, data, size, ((b200_backend_ctx *)be->context)->stream));
}
static void b200_be_get_tensor_async(ggml_backend_t be, const ggml_tensor *t, void *data, size_t off, size_t size) {
    ggml_backend_buffer_t buf = t->view_src ? t->view_src->buffer : t->buffer;
    if (b200_repacked(t->type)) { buf->iface.get_tensor(buf, t, data, off, size); return; }
    B200_OK(eb_download_async(data, (const char *)t->data + off, size, ((b200_backend_ctx *)be->context)->stream));
}
static ggml_backend_buffer_type_t b200_buffer_type();
static bool b200_be_cpy_tensor_async(ggml_backend_t src_be, ggml_backend_t dst_be, const ggml_tensor *src, ggml_tensor *dst) {
    // both ends on this device type, contiguous, not in plane layout: a device copy on the destination's stream after the source's work
    if (src_be->iface.get_name != b200_be_name || dst_be->iface.get_name != b200_be_name) return false;
    ggml_backend_buffer_t sb = src->view_src ? src->view_src->buffer : src->buffer, db = dst->view_src ? dst->view_src->buffer : dst->buffer;
    if (!sb || !db || sb->buft != b200_buffer_type() || db->buft != b200_buffer_type()) return false;
    if (!ggml_is_contiguous(src) || !ggml_is_contiguous(dst) || ggml_nbytes(src) != ggml_nbytes(dst)) return false;
    auto *sc = (b200_backend_ctx *)src_be->context; auto *dc = (b200_backend_ctx *)dst_be->context;
    if (sc != dc) {
        void *ev = nullptr;
        B200_OK(eb_event_create(&ev)); B200_OK(eb_event_record(ev, sc->stream)); B200_OK(eb_stream_wait_event(dc->stream, ev)); B200_OK(eb_event_destroy(ev));
    }
    B200_OK(eb_copy(dst->data, src->data, ggml_nbytes(src), dc->stream));
    return true;
}
// events (ggml_backend_cuda_event_record / event_wait, ggml-cuda.cu:2790-2815)
static void b200_be_event_record(ggml_backend_t be, ggml_backend_event_t ev) { B200_OK(eb_event_record(ev->context, ((b200_backend_ctx *)be->context)->stream)); }
static void b200_be_event_wait(ggml_backend_t be, ggml_backend_event_t ev) { B200_OK(eb_stream_wait_event(((b200_backend_ctx *)be->context)->stream, ev->context)); }

static float op_f32(const ggml_tensor *t, int i) { float v; memcpy(&v, (const int32_t *)t->op_params + i, 4); return v; }
static bool same_shape_contig(const ggml_tensor *a, const ggml_tensor *b) { return ggml_are_same_shape(a, b) && ggml_is_contiguous(a) && ggml_is_contiguous(b); }

// q = permute(reshape(Q, D, H, T), 0, 2, 1, 3): logical [D, T, H], memory [T][H][D] (llm_build_kqv, R/src/llama.cpp:724)
static bool fa_q_layout_ok(const ggml_tensor *q) {
    return q->type == GGML_TYPE_F32 && q->ne[3] == 1 && q->nb[0] == 4 && q->nb[2] == (size_t)q->ne[0] * 4 && q->nb[1] == (size_t)q->ne[0] * q->ne[2] * 4;
}
// k/v = view_3d(cache, D, n_kv, H, row, D*2): cell stride nb[1], head h at h*D elements (R/src/llama.cpp:736-758)
static bool fa_kv_layout_ok(const ggml_tensor *k, const ggml_tensor *q) {
    return k->type == GGML_TYPE_F16 && k->ne[0] == q->ne[0] && k->ne[2] == q->ne[2] && k->ne[3] == 1 && k->nb[0] == 2 && k->nb[2] == (size_t)k->ne[0] * 2 && k->nb[1] % 2 == 0 &&
           k->nb[1] >= (size_t)k->ne[0] * k->ne[2] * 2;
}

// supports_op (ggml_backend_cuda_device_supports_op, ggml-cuda.cu:2979): exactly the cases graph_compute handles
static bool b200_dev_supports_op(ggml_backend_dev_t, const ggml_tensor *op) {
    const ggml_tensor *a = op->src[0], *b = op->src[1];
    switch (op->op) {
    case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE: return true;
    case GGML_OP_MUL_MAT:
        if (a->type == GGML_TYPE_F16 || a->type == GGML_TYPE_F32)      // KQ / KQV of the FA-off attention: strided views, batched over heads
            return b->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && a->nb[0] == ggml_type_size(a->type) && b->nb[0] == 4 && a->ne[3] == 1 && b->ne[3] == 1 &&
                   a->ne[2] >= 1 && b->ne[2] % a->ne[2] == 0 && ggml_is_contiguous(op);
        return b200_weight_type(a->type) && ggml_is_contiguous(a) && a->ne[2] == 1 && a->ne[3] == 1 && b->type == GGML_TYPE_F32 && ggml_is_contiguous(b) &&
               b->ne[2] == 1 && b->ne[3] == 1 && a->ne[0] % (a->type == GGML_TYPE_Q4_K || a->type == GGML_TYPE_Q6_K ? 256 : 128) == 0 && op->type == GGML_TYPE_F32;   // 128: the Q8_0-grid quantiser works on 128-value groups
    case GGML_OP_FLASH_ATTN_EXT: {
        const ggml_tensor *v = op->src[2], *m = op->src[3];
        float max_bias = op_f32(op, 1), softcap = op_f32(op, 2);
        return a->ne[0] == 128 && fa_q_layout_ok(a) && fa_kv_layout_ok(b, a) && fa_kv_layout_ok(v, a) && b->nb[1] == v->nb[1] && b->ne[1] == v->ne[1] && m && m->type == GGML_TYPE_F16 &&
               ggml_is_contiguous(m) && m->ne[0] == b->ne[1] && m->ne[1] >= a->ne[1] && max_bias == 0.0f && softcap == 0.0f;
    }
    case GGML_OP_GET_ROWS:
        return (b200_weight_type(a->type) || a->type == GGML_TYPE_F16 || a->type == GGML_TYPE_F32) && ggml_is_contiguous(a) && a->ne[2] == 1 && b->type == GGML_TYPE_I32 && ggml_is_contiguous(b) && b->ne[1] == 1;
    case GGML_OP_RMS_NORM: return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && a->ne[0] % 4 == 0 && a->ne[0] <= 16384;
    case GGML_OP_ROPE: {
        const int32_t *p = (const int32_t *)op->op_params;     // R/ggml/src/ggml.c:3532: n_dims, mode, freq_base, freq_scale, ext_factor, attn_factor
        return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && a->ne[3] == 1 && p[1] == a->ne[0] && p[2] == 0 /* NORM */ && !op->src[2] && op_f32(op, 6) == 1.0f &&
               op_f32(op, 7) == 0.0f && op_f32(op, 8) == 1.0f && b->type == GGML_TYPE_I32;
    }
    case GGML_OP_ADD: case GGML_OP_MUL:   // same shape, or b = one row broadcast over the rows of a (norm weight, fc bias)
        return a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && ggml_is_contiguous(b) &&
               (ggml_are_same_shape(a, b) || (b->ne[0] == a->ne[0] && ggml_nelements(b) == b->ne[0]));
    case GGML_OP_UNARY: return (ggml_get_unary_op(op) == GGML_UNARY_OP_SILU || ggml_get_unary_op(op) == GGML_UNARY_OP_RELU) && a->type == GGML_TYPE_F32 && ggml_is_contiguous(a);
    case GGML_OP_CONCAT:    // [E, T] | [E, T] along dim 0 (build_eagle)
        return ((const int32_t *)op->op_params)[0] == 0 && a->type == GGML_TYPE_F32 && same_shape_contig(a, b) && a->ne[2] == 1 && a->ne[3] == 1;
    case GGML_OP_CPY: case GGML_OP_DUP: case GGML_OP_CONT: {    // any strides (transposed V store, merge of the permuted KQV)
        const ggml_tensor *d = op->op == GGML_OP_CPY ? op->src[1] : op;
        if (ggml_nelements(a) != ggml_nelements(d)) return false;
        if (ggml_is_contiguous(a) && ggml_is_contiguous(d) && a->type == d->type && !ggml_is_quantized(a->type)) return true;
        return (a->type == GGML_TYPE_F32 && (d->type == GGML_TYPE_F16 || d->type == GGML_TYPE_F32)) || (a->type == GGML_TYPE_F16 && d->type == GGML_TYPE_F16);
    }
    case GGML_OP_SOFT_MAX: {    // ggml_soft_max_ext(kq, mask, scale, max_bias = 0)
        const ggml_tensor *m = op->src[1];
        return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && a->ne[0] <= 48000 && op_f32(op, 1) == 0.0f &&
               (!m || ((m->type == GGML_TYPE_F16 || m->type == GGML_TYPE_F32) && m->ne[0] == a->ne[0] && m->ne[1] >= a->ne[1] && m->ne[2] == 1 && m->ne[3] == 1 &&
                       m->nb[0] == ggml_type_size(m->type)));
    }
    case GGML_OP_ARGMAX: return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a);
    case GGML_OP_ARGSORT: return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && a->ne[0] <= 1024;     // ggml_top_k = argsort + view
    case GGML_OP_SCALE: return a->type == GGML_TYPE_F32 && ggml_is_contiguous(a) && op->type == GGML_TYPE_F32;
    default: return false;
    }
}

static int b200_compute_node(b200_backend_ctx *ctx, ggml_tensor *d) {
    const ggml_tensor *a = d->src[0], *b = d->src[1];
    void *st = ctx->stream;
    switch (d->op) {
    case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE: return EB_OK;
    case GGML_OP_MUL_MAT: {     // a = weights [K, N], b = activations f32 [K, T] -> d f32 [N, T]
        if (a->type == GGML_TYPE_F16 || a->type == GGML_TYPE_F32)
            return eb_mul_mat_dense((int)a->type, a->data, (int64_t)a->nb[1], (int64_t)a->nb[2], a->ne[2], (const float *)b->data, (int64_t)b->nb[1], (int64_t)b->nb[2],
                                    (float *)d->data, a->ne[1], a->ne[0], b->ne[1], b->ne[2], st);
        void *ws = ctx->scratch(eb_mul_mat_ws_size((int)a->type, a->ne[1], a->ne[0], b->ne[1]));
        if (!ws) return EB_ERR_OOM;
        const int kq = a->type == GGML_TYPE_Q4_K || a->type == GGML_TYPE_Q6_K;
        const int64_t K = a->ne[0], T = b->ne[1];
        const bool reuse = T <= 128 && ctx->aq_src == (const char *)b->data && ctx->aq_K == K && ctx->aq_T == T && ctx->aq_kq == kq && ctx->aq_ws == ws;
        const int rc = eb_mul_mat_ex((int)a->type, a->data, a->ne[1], K, (const float *)b->data, T, (float *)d->data, ws, reuse ? 1 : 0, st);
        ctx->aq_src = T <= 128 && rc == EB_OK ? (const char *)b->data : nullptr; ctx->aq_bytes = (size_t)T * K * 4; ctx->aq_K = K; ctx->aq_T = T; ctx->aq_kq = kq; ctx->aq_ws = ws;
        return rc;
    }
    case GGML_OP_FLASH_ATTN_EXT: {
        const ggml_tensor *v = d->src[2], *m = d->src[3];
        const int D = (int)a->ne[0], T = (int)a->ne[1], H = (int)a->ne[2], n_kv = (int)b->ne[1];
        const int64_t kv_stride = (int64_t)(b->nb[1] / 2);
        void *ws = ctx->scratch(eb_flash_attn_ws_size(T, H, D, n_kv));
        if (!ws) return EB_ERR_OOM;
        ctx->aq_drop();                                             // the attention workspace overlays the activation planes
        // llama_set_inputs writes 0 / -inf masks (R/src/llama-context.cpp:136-201): those run on the tcgen05 kernel through the mask's
        // bit form, built once per graph (all layers share KQ_mask); anything else (ALiBi, test masks) takes the generic kernel
        static const bool generic_only = getenv("GGML_B200_FA_GENERIC") != nullptr;      // A/B switch: always take the f16-mask kernel
        if (!generic_only && eb_flash_attn_bits_supported(T, H, D, kv_stride) && m->nb[1] == (size_t)m->ne[0] * 2) {
            if (ctx->mask_data != m->data || ctx->mask_T != T || ctx->mask_kv != n_kv) {
                const size_t need = eb_mask_bits_size(T, n_kv);
                if (need > ctx->bits_size) {
                    if (ctx->bits) { eb_synchronize(st); eb_free(ctx->bits); }
                    ctx->bits = nullptr; ctx->bits_size = 0;
                    if (eb_malloc(&ctx->bits, need) != EB_OK) return EB_ERR_OOM;
                    ctx->bits_size = need;
                }
                int general = 1;
                int rc = eb_mask_to_bits(m->data, (int64_t)(m->nb[1] / 2), T, n_kv, ctx->bits, &general, st); if (rc) return rc;
                ctx->mask_data = m->data; ctx->mask_T = T; ctx->mask_kv = n_kv; ctx->mask_bits_ok = general == 0;
            }
            if (ctx->mask_bits_ok)
                return eb_flash_attn_bits((const float *)a->data, b->data, v->data, kv_stride, ctx->bits, T, H, D, n_kv, op_f32(d, 0), (float *)d->data, ws, st);
        }
        return eb_flash_attn((const float *)a->data, b->data, v->data, kv_stride, m->data, (int64_t)(m->nb[1] / 2), nullptr, 0,
                             T, H, D, n_kv, op_f32(d, 0), (float *)d->data, ws, st);
    }
    case GGML_OP_GET_ROWS: return eb_get_rows_n((int)a->type, a->data, a->ne[1], a->ne[0], (const int32_t *)b->data, (int)b->ne[0], (float *)d->data, st);
    case GGML_OP_RMS_NORM: return eb_rms_norm((const float *)a->data, nullptr, (float *)d->data, a->ne[0], ggml_nrows(a), op_f32(d, 0), st);
    case GGML_OP_ROPE:          // ggml's dst may or may not alias src0 (ggml-alloc inplace): the kernel reads src and writes dst
        return eb_rope_norm_to((const float *)a->data, (float *)d->data, (const int32_t *)b->data, (int)a->ne[0], (int)a->ne[1], (int)a->ne[2], op_f32(d, 5), st);
    case GGML_OP_ADD: return eb_add((const float *)a->data, (const float *)b->data, (float *)d->data, ggml_nelements(d), ggml_nelements(b), st);
    case GGML_OP_MUL: return eb_mul((const float *)a->data, (const float *)b->data, (float *)d->data, ggml_nelements(d), ggml_nelements(b), st);
    case GGML_OP_UNARY:
        return ggml_get_unary_op(d) == GGML_UNARY_OP_SILU ? eb_silu((const float *)a->data, (float *)d->data, ggml_nelements(d), st)
                                                          : eb_relu((const float *)a->data, (float *)d->data, ggml_nelements(d), st);
    case GGML_OP_CONCAT: return eb_concat2((const float *)a->data, (const float *)b->data, (float *)d->data, (int)a->ne[0], (int)a->ne[1], st);
    case GGML_OP_CPY: case GGML_OP_DUP: case GGML_OP_CONT: {
        const ggml_tensor *dst = d->op == GGML_OP_CPY ? d->src[1] : d;     // ggml_cpy: the node is a view of src[1]
        if (ggml_is_contiguous(a) && ggml_is_contiguous(dst)) {
            if (a->type == GGML_TYPE_F32 && dst->type == GGML_TYPE_F16) return eb_cpy_f32_f16((const float *)a->data, dst->data, ggml_nelements(a), st);   // KV store, mask cast
            if (a->type == dst->type) return eb_copy(dst->data, a->data, ggml_nbytes(a), st);
        }
        const int64_t nbs[4] = {(int64_t)a->nb[0], (int64_t)a->nb[1], (int64_t)a->nb[2], (int64_t)a->nb[3]};
        const int64_t nbd[4] = {(int64_t)dst->nb[0], (int64_t)dst->nb[1], (int64_t)dst->nb[2], (int64_t)dst->nb[3]};
        return eb_copy_strided((int)a->type, (int)dst->type, a->data, dst->data, a->ne, nbs, dst->ne, nbd, st);
    }
    case GGML_OP_SOFT_MAX: {
        const ggml_tensor *m = d->src[1];
        return eb_soft_max_ext((const float *)a->data, m ? m->data : nullptr, m ? (int)m->type : 0, m ? (int64_t)(m->nb[1] / ggml_type_size(m->type)) : 0,
                               (float *)d->data, a->ne[0], ggml_nrows(a), a->ne[1], op_f32(d, 0), st);
    }
    case GGML_OP_ARGMAX: return eb_argmax((const float *)a->data, (int)ggml_nrows(a), (int)a->ne[0], (int32_t *)d->data, st);
    case GGML_OP_SCALE: return eb_scale((const float *)a->data, (float *)d->data, ggml_nelements(d), op_f32(d, 0), st);
    case GGML_OP_ARGSORT: return eb_argsort((const float *)a->data, (int)ggml_nrows(a), (int)a->ne[0], ((const int32_t *)d->op_params)[0] == GGML_SORT_ORDER_DESC, (int32_t *)d->data, st);
    default: return EB_ERR_UNSUPPORTED;
    }
}

// graph_compute (ggml_backend_cuda_graph_compute, ggml-cuda.cu:2693): asynchronous on the backend's stream;
// a failing op fails the graph (-> llama_decode returns -3, R/src/llama.cpp:9610-9617)
static ggml_status b200_be_graph_compute(ggml_backend_t be, ggml_cgraph *g) {
    auto *ctx = (b200_backend_ctx *)be->context;
    const int n = ggml_graph_n_nodes(g);
    ctx->mask_data = nullptr;                    // the mask's contents change between graphs even when its address does not
    ctx->aq_drop();
    static const bool timing = getenv("GGML_B200_TIMING") != nullptr;
    const int64_t t0 = timing ? ggml_time_us() : 0;
    for (int i = 0; i < n; i++) {
        ggml_tensor *d = ggml_graph_node(g, i);
        if (ggml_is_empty(d)) continue;
        // a node that writes into the tensor whose quantisation is cached ends the cache (views and reshapes compute nothing)
        if (ctx->aq_src && d->op != GGML_OP_MUL_MAT && d->op != GGML_OP_NONE && d->op != GGML_OP_RESHAPE && d->op != GGML_OP_VIEW && d->op != GGML_OP_PERMUTE &&
            d->op != GGML_OP_TRANSPOSE && d->data) {
            const char *w0 = (const char *)d->data, *w1 = w0 + ggml_nbytes(d);
            if (w0 < ctx->aq_src + ctx->aq_bytes && ctx->aq_src < w1) ctx->aq_drop();
        }
        const int rc = b200_compute_node(ctx, d);
        if (rc == EB_ERR_OOM) { b200_log_err(ggml_op_name(d->op)); return GGML_STATUS_ALLOC_FAILED; }
        if (rc != EB_OK) { fprintf(stderr, "ggml-b200: %s (%s): %s\n", ggml_op_name(d->op), d->name, rc == EB_ERR_UNSUPPORTED ? "unsupported op" : eb_last_error()); return GGML_STATUS_FAILED; }
    }
    ctx->aq_drop();
    if (timing) fprintf(stderr, "ggml-b200: graph_compute issued %d nodes in %lld us (host side; the stream runs on)\n", n, (long long)(ggml_time_us() - t0));
    return GGML_STATUS_SUCCESS;
}

static const ggml_backend_i b200_backend_iface = {
    b200_be_name, b200_be_free, b200_be_set_tensor_async, b200_be_get_tensor_async, b200_be_cpy_tensor_async, b200_be_sync,
    nullptr, nullptr, nullptr, nullptr, b200_be_graph_compute, b200_be_event_record, b200_be_event_wait,
};
static ggml_guid_t b200_guid() { static ggml_guid g = {0xb2, 0x00, 0xea, 0x61, 0x1e, 0x5b, 0x10, 0x0a, 0x73, 0xc9, 0x05, 0x7a, 0x4d, 0x11, 0x20, 0x26}; return &g; }

// ---------------------------------------------------------------------------------- device
struct b200_device_ctx { int device = 0; std::string name, desc; };
static const char *b200_dev_name(ggml_backend_dev_t d) { return ((b200_device_ctx *)d->context)->name.c_str(); }
static const char *b200_dev_desc(ggml_backend_dev_t d) { return ((b200_device_ctx *)d->context)->desc.c_str(); }
static void b200_dev_memory(ggml_backend_dev_t, size_t *free, size_t *total) { if (eb_device_memory(free, total) != EB_OK) { *free = 0; *total = 0; } }
static enum ggml_backend_dev_type b200_dev_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }
static void b200_dev_props(ggml_backend_dev_t d, ggml_backend_dev_props *p) {
    p->name = b200_dev_name(d); p->description = b200_dev_desc(d); p->type = GGML_BACKEND_DEVICE_TYPE_GPU;
    b200_dev_memory(d, &p->memory_free, &p->memory_total);
    p->caps = { /* async */ true, /* host_buffer */ true, /* buffer_from_host_ptr */ false, /* events */ true };
}
static ggml_backend_t b200_dev_init_backend(ggml_backend_dev_t dev, const char *) {
    if (eb_device_init(((b200_device_ctx *)dev->context)->device) != EB_OK) { b200_log_err("init_backend"); return nullptr; }   // refuses anything that is not sm_100
    auto *c = new b200_backend_ctx();
    if (eb_stream_create(&c->stream) != EB_OK) { b200_log_err("init_backend: stream"); delete c; return nullptr; }
    return new ggml_backend{ b200_guid(), b200_backend_iface, dev, c };
}
static ggml_backend_buffer_type_t b200_dev_buffer_type(ggml_backend_dev_t) { return b200_buffer_type(); }
static ggml_backend_buffer_type_t b200_dev_host_buffer_type(ggml_backend_dev_t) { return b200_host_buffer_type(); }
// tensors in the pinned host type are computed on by the CPU backend (it is a host buffer), not by this device
static bool b200_dev_supports_buft(ggml_backend_dev_t, ggml_backend_buffer_type_t buft) { return buft == b200_buffer_type(); }
static bool b200_dev_offload_op(ggml_backend_dev_t, const ggml_tensor *) { return false; }    // weights are resident: nothing is offloaded per op

// events (ggml_backend_cuda_device_event_new / free / synchronize, ggml-cuda.cu:3161-3198)
static ggml_backend_event_t b200_dev_event_new(ggml_backend_dev_t dev) {
    void *ev = nullptr;
    if (eb_event_create(&ev) != EB_OK) { b200_log_err("event_new"); return nullptr; }
    return new ggml_backend_event{ dev, ev };
}
static void b200_dev_event_free(ggml_backend_dev_t, ggml_backend_event_t ev) { eb_event_destroy(ev->context); delete ev; }
static void b200_dev_event_sync(ggml_backend_dev_t, ggml_backend_event_t ev) { B200_OK(eb_event_synchronize(ev->context)); }

static const ggml_backend_device_i b200_device_iface = {
    b200_dev_name, b200_dev_desc, b200_dev_memory, b200_dev_type, b200_dev_props, b200_dev_init_backend, b200_dev_buffer_type,
    b200_dev_host_buffer_type, /* buffer_from_host_ptr */ nullptr, b200_dev_supports_op, b200_dev_supports_buft, b200_dev_offload_op,
    b200_dev_event_new, b200_dev_event_free, b200_dev_event_sync,
};

// ---------------------------------------------------------------------------------- reg
static const char *b200_reg_name(ggml_backend_reg_t) { return B200_NAME; }
static ggml_backend_dev_t b200_reg_device(ggml_backend_reg_t reg, size_t i);
static size_t b200_reg_count(ggml_backend_reg_t) { return eb_device_init(0) == EB_OK ? 1 : 0; }    // one replica per process
// optional entry points by name (R/src/llama.cpp:9086-9095 asks for the CPU thread-pool setters: NULL here);
// the fused speculative-step engine is reachable through the same hook
static void *b200_reg_proc(ggml_backend_reg_t, const char *name) {
    static const struct { const char *n; void *p; } tab[] = {
        {"eb_model_create", (void *)eb_model_create}, {"eb_model_set_tensor", (void *)eb_model_set_tensor}, {"eb_model_share_lm_head", (void *)eb_model_share_lm_head},
        {"eb_model_free", (void *)eb_model_free}, {"eb_session_create", (void *)eb_session_create}, {"eb_session_free", (void *)eb_session_free},
        {"eb_session_forward", (void *)eb_session_forward}, {"eb_session_commit", (void *)eb_session_commit}, {"eb_session_set_root", (void *)eb_session_set_root},
        {"eb_session_round", (void *)eb_session_round}, {"eb_last_error", (void *)eb_last_error},
    };
    for (auto &e : tab) if (!strcmp(e.n, name)) return e.p;
    return nullptr;
}
static const ggml_backend_reg_i b200_reg_iface = { b200_reg_name, b200_reg_count, b200_reg_device, b200_reg_proc };

static ggml_backend_reg_t b200_reg() {
    static ggml_backend_reg reg = { GGML_BACKEND_API_VERSION, b200_reg_iface, nullptr };
    return &reg;
}
static ggml_backend_dev_t b200_the_device() { return b200_reg_device(b200_reg(), 0); }
static ggml_backend_dev_t b200_reg_device(ggml_backend_reg_t reg, size_t i) {
    static b200_device_ctx dctx;
    static ggml_backend_device dev = { b200_device_iface, reg, &dctx };
    static std::once_flag once;
    std::call_once(once, [] { dctx.device = 0; dctx.name = B200_NAME "0"; dctx.desc = "NVIDIA B200 (sm_100a) via libeagle_b200"; });
    GGML_ASSERT(i == 0);
    return &dev;
}
static ggml_backend_buffer_type_t b200_buffer_type() {
    static ggml_backend_buffer_type buft = {
        { b200_buft_name, b200_buft_alloc, b200_buft_alignment, /* get_max_size */ nullptr, b200_buft_alloc_size, b200_buft_is_host },
        b200_reg_device(b200_reg(), 0), nullptr,
    };
    return &buft;
}

extern "C" { GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void); }
ggml_backend_reg_t ggml_backend_b200_reg(void) { return b200_reg(); }
static int b200_score() { return eb_device_init(0) == EB_OK ? 100 : 0; }     // 0 = "not supported on this system": the loader skips us
GGML_BACKEND_DL_IMPL(ggml_backend_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(b200_score)
