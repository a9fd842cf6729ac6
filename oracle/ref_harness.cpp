// ref_harness.cpp — thin C-ABI shim over the REFERENCE's own ggml + ggml-cpu code.
//
// TEST INFRASTRUCTURE ONLY.  Compiled by oracle/Makefile against the headers and sources where
// they lie under /root/reference/llama.cpp/ggml (nothing is copied into this repo); the output
// goes to oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).  It exists to
// (1) pin oracle/eagle_oracle.c against the real reference arithmetic and (2) serve as the
// "reference" CPU baseline in bench.py (cpu_baseline.kind = "reference", --impl reference).
//
// Every op below is executed by the reference's ggml-cpu backend (ggml_backend_cpu_init +
// ggml_backend_graph_compute); this file only builds the same ggml graphs that
// R/src/llama.cpp build_llama (:1647) / build_eagle (:1839) build.
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#define RH_API extern "C" __attribute__((visibility("default")))

static ggml_backend_t g_cpu = nullptr;
static int g_threads = 1;

// ggml's f16 lookup tables are filled by the first ggml_init (R/ggml/src/ggml.c ggml_init)
__attribute__((constructor)) static void rh_init() {
    ggml_init_params p = {1024, nullptr, false};
    ggml_free(ggml_init(p));
    ggml_cpu_init();
}

static ggml_backend_t cpu() {
    if (!g_cpu) { g_cpu = ggml_backend_cpu_init(); ggml_backend_cpu_set_n_threads(g_cpu, g_threads); }
    return g_cpu;
}

RH_API void ref_set_threads(int n) { g_threads = n > 0 ? n : 1; if (g_cpu) ggml_backend_cpu_set_n_threads(g_cpu, g_threads); }
RH_API int  ref_get_threads(void) { return g_threads; }
RH_API size_t ref_row_size(int type, int64_t k) { return ggml_row_size((ggml_type)type, k); }

RH_API void ref_dequantize_row(int type, const void *src, float *dst, int64_t k) {
    ggml_get_type_traits((ggml_type)type)->to_float(src, dst, k);
}
// weights quantiser used by llama-quantize (ggml_quantize_chunk -> quantize_q4_K etc.)
RH_API size_t ref_quantize(int type, const float *src, void *dst, int64_t nrows, int64_t n_per_row) {
    return ggml_quantize_chunk((ggml_type)type, src, dst, 0, nrows, n_per_row, nullptr);
}
// activation quantiser the CPU mul_mat uses (type_traits_cpu[...].from_float)
RH_API void ref_quantize_act(int type, const float *x, void *y, int64_t k) {
    ggml_get_type_traits_cpu((ggml_type)type)->from_float(x, y, k);
}
RH_API float ref_vec_dot(int type, int64_t k, const void *w, const void *a) {
    float s = 0;
    ggml_get_type_traits_cpu((ggml_type)type)->vec_dot((int)k, &s, 0, w, 0, a, 0, 1);
    return s;
}

struct Ctx {
    ggml_context *c;
    explicit Ctx(size_t bytes) { ggml_init_params p = {bytes, nullptr, false}; c = ggml_init(p); }
    ~Ctx() { ggml_free(c); }
};

static void run(ggml_context *c, ggml_tensor *out) {
    ggml_cgraph *gf = ggml_new_graph_custom(c, 16384, false);
    ggml_build_forward_expand(gf, out);
    ggml_backend_graph_compute(cpu(), gf);
}

// tensor that aliases caller memory (weights are never copied)
static ggml_tensor *alias2d(ggml_context *c, int type, int64_t ne0, int64_t ne1, const void *data) {
    ggml_init_params p = {ggml_tensor_overhead() * 2, nullptr, true};
    (void)p;
    ggml_tensor *t = ggml_new_tensor_2d(c, (ggml_type)type, ne0, 1);   // tiny allocation, then re-point
    t->ne[1] = ne1;
    t->nb[2] = t->nb[1] * ne1; t->nb[3] = t->nb[2];
    t->data = const_cast<void *>(data);
    return t;
}

RH_API void ref_mul_mat(int type, const void *W, int64_t N, int64_t K, const float *X, int64_t T, float *Y) {
    Ctx cx((size_t)T * (K + N) * 4 + (64u << 20));
    ggml_tensor *w = alias2d(cx.c, type, K, N, W);
    ggml_tensor *x = alias2d(cx.c, GGML_TYPE_F32, K, T, X);
    ggml_tensor *y = ggml_mul_mat(cx.c, w, x);
    run(cx.c, y);
    memcpy(Y, y->data, (size_t)T * N * 4);
}

RH_API void ref_rms_norm(const float *x, float *y, int64_t n, int64_t rows, float eps) {
    Ctx cx((size_t)rows * n * 8 + (16u << 20));
    ggml_tensor *a = alias2d(cx.c, GGML_TYPE_F32, n, rows, x);
    ggml_tensor *r = ggml_rms_norm(cx.c, a, eps);
    run(cx.c, r);
    memcpy(y, r->data, (size_t)rows * n * 4);
}

RH_API void ref_rope_norm(float *x, const int32_t *pos, int D, int n_head, int T, float freq_base) {
    Ctx cx((size_t)T * n_head * D * 8 + (16u << 20));
    ggml_tensor *a = ggml_new_tensor_3d(cx.c, GGML_TYPE_F32, D, n_head, T);
    memcpy(a->data, x, ggml_nbytes(a));
    ggml_tensor *p = ggml_new_tensor_1d(cx.c, GGML_TYPE_I32, T);
    memcpy(p->data, pos, (size_t)T * 4);
    ggml_tensor *r = ggml_rope_ext(cx.c, a, p, nullptr, D, 0 /*LLAMA_ROPE_TYPE_NORM*/, 0, freq_base, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
    run(cx.c, r);
    memcpy(x, r->data, ggml_nbytes(a));
}

RH_API void ref_get_rows(int type, const void *W, int64_t K, int64_t n_rows, const int32_t *ids, int n, float *out) {
    Ctx cx((size_t)n * K * 4 + (16u << 20));
    ggml_tensor *w = alias2d(cx.c, type, K, n_rows, W);
    ggml_tensor *i = ggml_new_tensor_1d(cx.c, GGML_TYPE_I32, n);
    memcpy(i->data, ids, (size_t)n * 4);
    ggml_tensor *r = ggml_get_rows(cx.c, w, i);
    run(cx.c, r);
    memcpy(out, r->data, (size_t)n * K * 4);
}

static ggml_tensor *make_mask(ggml_context *c, const uint8_t *vis, int T, int n_kv) {
    const int Tp = GGML_PAD(T, GGML_KQ_MASK_PAD);
    ggml_tensor *m32 = ggml_new_tensor_2d(c, GGML_TYPE_F32, n_kv, Tp);
    float *d = (float *)m32->data;
    for (int j = 0; j < Tp; j++) for (int i = 0; i < n_kv; i++)
        d[(size_t)j * n_kv + i] = (j < T && vis[(size_t)j * n_kv + i]) ? 0.0f : -INFINITY;
    return ggml_cast(c, m32, GGML_TYPE_F16);               // R/src/llama.cpp:1488
}

// q [T][H][D] f32; k,v f16 cells [n_kv][H][D]; vis [T][n_kv]; out [T][H][D]
RH_API void ref_flash_attn(const float *q, const uint16_t *k, const uint16_t *v, const uint8_t *vis,
                           int T, int H, int D, int n_kv, float scale, float *out) {
    Ctx cx((size_t)T * H * D * 16 + (size_t)n_kv * (T + 64) * 8 + (32u << 20));
    ggml_context *c = cx.c;
    ggml_tensor *qc = ggml_new_tensor_3d(c, GGML_TYPE_F32, D, H, T);
    memcpy(qc->data, q, ggml_nbytes(qc));
    ggml_tensor *qp = ggml_permute(c, qc, 0, 2, 1, 3);
    const size_t row = (size_t)H * D * 2;
    ggml_tensor *kc = alias2d(c, GGML_TYPE_F16, (int64_t)H * D, n_kv, k);
    ggml_tensor *vc = alias2d(c, GGML_TYPE_F16, (int64_t)H * D, n_kv, v);
    ggml_tensor *kv3 = ggml_view_3d(c, kc, D, n_kv, H, row, (size_t)D * 2, 0);
    ggml_tensor *vv3 = ggml_view_3d(c, vc, D, n_kv, H, row, (size_t)D * 2, 0);
    ggml_tensor *mask = make_mask(c, vis, T, n_kv);
    ggml_tensor *r = ggml_flash_attn_ext(c, qp, kv3, vv3, mask, scale, 0.0f, 0.0f);
    ggml_flash_attn_ext_set_prec(r, GGML_PREC_F32);
    run(c, r);
    memcpy(out, r->data, (size_t)T * H * D * 4);
}

// ---------------------------------------------------------------- model forward
enum { T_TOK_EMBD = 0, T_OUTPUT_NORM, T_OUTPUT, T_FC, T_FC_BIAS, T_GLOBAL_COUNT };
enum { L_ATTN_NORM = 0, L_WQ, L_WK, L_WV, L_WO, L_FFN_NORM, L_GATE, L_UP, L_DOWN, L_COUNT };
struct RTensor { int type = 0; const void *data = nullptr; };
struct RModel {
    int E, H, F, V, L, is_eagle; float eps, freq_base;
    RTensor g[T_GLOBAL_COUNT];
    std::vector<std::vector<RTensor>> layers;
};
struct RKV { int L, E, n_ctx; std::vector<std::vector<uint16_t>> k, v; };

RH_API RModel *ref_model_new(int E, int H, int F, int V, int L, float eps, float freq_base, int is_eagle) {
    RModel *m = new RModel();
    m->E = E; m->H = H; m->F = F; m->V = V; m->L = L; m->eps = eps; m->freq_base = freq_base; m->is_eagle = is_eagle;
    m->layers.assign(L, std::vector<RTensor>(L_COUNT));
    return m;
}
RH_API void ref_model_free(RModel *m) { delete m; }
RH_API void ref_model_set_global(RModel *m, int which, int type, const void *data) { m->g[which].type = type; m->g[which].data = data; }
RH_API void ref_model_set_layer(RModel *m, int l, int which, int type, const void *data) { m->layers[l][which].type = type; m->layers[l][which].data = data; }
RH_API void ref_model_set_lm_head_from(RModel *m, const RModel *t) { m->g[T_OUTPUT] = t->g[T_OUTPUT]; }
RH_API RKV *ref_kv_new(int L, int E, int n_ctx) {
    RKV *kv = new RKV(); kv->L = L; kv->E = E; kv->n_ctx = n_ctx;
    kv->k.assign(L, std::vector<uint16_t>((size_t)n_ctx * E, 0)); kv->v = kv->k;
    return kv;
}
RH_API void ref_kv_free(RKV *kv) { delete kv; }
RH_API uint16_t *ref_kv_k(RKV *kv, int l) { return kv->k[l].data(); }
RH_API uint16_t *ref_kv_v(RKV *kv, int l) { return kv->v[l].data(); }

static ggml_tensor *norm(ggml_context *c, ggml_tensor *x, const RTensor &w, int E, float eps) {   // llm_build_norm :329-365
    ggml_tensor *r = ggml_rms_norm(c, x, eps);
    if (w.data) r = ggml_mul(c, r, alias2d(c, GGML_TYPE_F32, E, 1, w.data));
    return r;
}

// Same graph as build_llama / build_eagle with flash attention on (llm_build_kqv :758-761),
// f16 KV cache, all T rows as outputs.
RH_API void ref_forward(const RModel *m, RKV *kv, const int32_t *tokens, const int32_t *pos,
                        const float *hidden_in, int T, int kv_head, int n_kv, const uint8_t *vis,
                        float *logits, float *feat) {
    const int E = m->E, H = m->H, F = m->F, D = E / H, V = m->V;
    size_t act = (size_t)T * (size_t)(8 * E + 3 * F) * 4 * (size_t)(m->L + 2) + (size_t)T * V * 4 + (size_t)n_kv * (T + 64) * 8;
    Ctx cx(act + (256u << 20));
    ggml_context *c = cx.c;
    ggml_cgraph *gf = ggml_new_graph_custom(c, 32768, false);

    ggml_tensor *tok = ggml_new_tensor_1d(c, GGML_TYPE_I32, T); memcpy(tok->data, tokens, (size_t)T * 4);
    ggml_tensor *ps = ggml_new_tensor_1d(c, GGML_TYPE_I32, T);  memcpy(ps->data, pos, (size_t)T * 4);
    ggml_tensor *mask = make_mask(c, vis, T, n_kv);

    ggml_tensor *inpL = ggml_get_rows(c, alias2d(c, m->g[T_TOK_EMBD].type, E, V, m->g[T_TOK_EMBD].data), tok);
    if (m->is_eagle) {
        ggml_tensor *hid = ggml_new_tensor_2d(c, GGML_TYPE_F32, E, T); memcpy(hid->data, hidden_in, (size_t)T * E * 4);
        ggml_tensor *cat = ggml_concat(c, inpL, hid, 0);
        ggml_tensor *r = ggml_mul_mat(c, alias2d(c, m->g[T_FC].type, 2 * E, E, m->g[T_FC].data), cat);
        if (m->g[T_FC_BIAS].data) r = ggml_add(c, r, alias2d(c, GGML_TYPE_F32, E, 1, m->g[T_FC_BIAS].data));
        inpL = ggml_relu(c, r);
    }
    const float kq_scale = 1.0f / sqrtf((float)D);
    for (int l = 0; l < m->L; l++) {
        const std::vector<RTensor> &w = m->layers[l];
        ggml_tensor *inpSA = inpL;
        ggml_tensor *cur = norm(c, inpL, w[L_ATTN_NORM], E, m->eps);
        ggml_tensor *Q = ggml_mul_mat(c, alias2d(c, w[L_WQ].type, E, E, w[L_WQ].data), cur);
        ggml_tensor *K = ggml_mul_mat(c, alias2d(c, w[L_WK].type, E, E, w[L_WK].data), cur);
        ggml_tensor *Vc = ggml_mul_mat(c, alias2d(c, w[L_WV].type, E, E, w[L_WV].data), cur);
        Q = ggml_rope_ext(c, ggml_reshape_3d(c, Q, D, H, T), ps, nullptr, D, 0, 0, m->freq_base, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
        K = ggml_rope_ext(c, ggml_reshape_3d(c, K, D, H, T), ps, nullptr, D, 0, 0, m->freq_base, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
        // llm_build_kv_store :228-270
        ggml_tensor *kc = alias2d(c, GGML_TYPE_F16, E, kv->n_ctx, kv->k[l].data());
        ggml_tensor *vc = alias2d(c, GGML_TYPE_F16, E, kv->n_ctx, kv->v[l].data());
        const size_t row = (size_t)E * 2;
        ggml_tensor *kview = ggml_view_1d(c, kc, (int64_t)T * E, row * kv_head);
        ggml_tensor *vview = ggml_view_1d(c, vc, (int64_t)T * E, row * kv_head);
        ggml_build_forward_expand(gf, ggml_cpy(c, ggml_reshape_1d(c, K, (int64_t)T * E), kview));
        ggml_build_forward_expand(gf, ggml_cpy(c, ggml_reshape_1d(c, Vc, (int64_t)T * E), vview));
        // llm_build_kqv :706-828 (flash-attn branch)
        ggml_tensor *q = ggml_permute(c, Q, 0, 2, 1, 3);
        ggml_tensor *k3 = ggml_view_3d(c, kc, D, n_kv, H, row, (size_t)D * 2, 0);
        ggml_tensor *v3 = ggml_view_3d(c, vc, D, n_kv, H, row, (size_t)D * 2, 0);
        cur = ggml_flash_attn_ext(c, q, k3, v3, mask, kq_scale, 0.0f, 0.0f);
        ggml_flash_attn_ext_set_prec(cur, GGML_PREC_F32);
        cur = ggml_reshape_2d(c, cur, E, T);
        cur = ggml_mul_mat(c, alias2d(c, w[L_WO].type, E, E, w[L_WO].data), cur);
        ggml_tensor *ffn_inp = ggml_add(c, cur, inpSA);
        cur = norm(c, ffn_inp, w[L_FFN_NORM], E, m->eps);
        // llm_build_ffn :456-581, LLM_FFN_SILU + LLM_FFN_PAR
        ggml_tensor *up = ggml_mul_mat(c, alias2d(c, w[L_UP].type, E, F, w[L_UP].data), cur);
        ggml_tensor *gt = ggml_mul_mat(c, alias2d(c, w[L_GATE].type, E, F, w[L_GATE].data), cur);
        gt = ggml_silu(c, gt);
        cur = ggml_mul(c, gt, up);
        cur = ggml_mul_mat(c, alias2d(c, w[L_DOWN].type, F, E, w[L_DOWN].data), cur);
        inpL = ggml_add(c, cur, ffn_inp);
    }
    ggml_tensor *res_norm = norm(c, inpL, m->g[T_OUTPUT_NORM], E, m->eps);
    ggml_build_forward_expand(gf, res_norm);
    ggml_tensor *out = nullptr;
    if (logits) {
        out = ggml_mul_mat(c, alias2d(c, m->g[T_OUTPUT].type, E, V, m->g[T_OUTPUT].data), res_norm);
        ggml_build_forward_expand(gf, out);
    }
    ggml_backend_graph_compute(cpu(), gf);
    if (feat) memcpy(feat, res_norm->data, (size_t)T * E * 4);
    if (logits) memcpy(logits, out->data, (size_t)T * V * 4);
}
