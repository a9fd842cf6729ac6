// ggml_graph_harness.cpp — TEST INFRASTRUCTURE ONLY (built by eagle-in-llama.cpp_b200/ggml_backend/Makefile
// into oracle/_ref/, against the reference's headers where they lie).
//
// Proves the drop-in: it loads libggml-b200.so through the reference's own plugin loader
// (ggml_backend_load, R/ggml/src/ggml-backend-reg.cpp:220), builds the SAME ggml graph as
// build_llama / build_eagle (R/src/llama.cpp:1647-1811, 1839-2019) with backend-allocated tensors
// (ggml_backend_alloc_ctx_tensors + ggml_gallocr, the way libllama does), and runs it either on the
// reference's ggml-cpu backend or on the plugin.  tests/test_gpu_ggml_backend.py compares the two —
// the same comparison R/tests/test-backend-ops.cpp makes per op.
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#define GH_API extern "C" __attribute__((visibility("default")))

enum { T_TOK_EMBD = 0, T_OUTPUT_NORM, T_OUTPUT, T_FC, T_FC_BIAS, T_GLOBAL_COUNT };
enum { L_ATTN_NORM = 0, L_WQ, L_WK, L_WV, L_WO, L_FFN_NORM, L_GATE, L_UP, L_DOWN, L_COUNT };
struct GTensor { int type = 0; const void *data = nullptr; };
struct GModel {
    int E, H, F, V, L, is_eagle; float eps, freq_base;
    GTensor g[T_GLOBAL_COUNT];
    std::vector<std::vector<GTensor>> layers;
};

static ggml_backend_reg_t g_reg = nullptr;
static ggml_backend_t g_be[2] = {nullptr, nullptr};      // 0 = reference ggml-cpu, 1 = the loaded plugin
static char g_err[512] = "";

GH_API const char *gh_last_error(void) { return g_err; }
GH_API int gh_load_plugin(const char *path) {
    if (g_reg) return 0;
    g_reg = ggml_backend_load(path);
    if (!g_reg) { snprintf(g_err, sizeof g_err, "ggml_backend_load(%s) failed", path); return -1; }
    if (ggml_backend_reg_dev_count(g_reg) < 1) { snprintf(g_err, sizeof g_err, "plugin reports no device"); return -2; }
    return 0;
}
GH_API const char *gh_plugin_name(void) { return g_reg ? ggml_backend_reg_name(g_reg) : ""; }
static ggml_backend_t backend(int which) {
    if (!g_be[which]) {
        if (which == 0) { g_be[0] = ggml_backend_cpu_init(); ggml_backend_cpu_set_n_threads(g_be[0], 8); }
        else if (g_reg) g_be[1] = ggml_backend_dev_init(ggml_backend_reg_dev_get(g_reg, 0), nullptr);
    }
    return g_be[which];
}

GH_API GModel *gh_model_new(int E, int H, int F, int V, int L, float eps, float freq_base, int is_eagle) {
    GModel *m = new GModel();
    m->E = E; m->H = H; m->F = F; m->V = V; m->L = L; m->eps = eps; m->freq_base = freq_base; m->is_eagle = is_eagle;
    m->layers.assign(L, std::vector<GTensor>(L_COUNT));
    return m;
}
GH_API void gh_model_free(GModel *m) { delete m; }
GH_API void gh_model_set_global(GModel *m, int which, int type, const void *data) { m->g[which] = {type, data}; }
GH_API void gh_model_set_layer(GModel *m, int l, int which, int type, const void *data) { m->layers[l][which] = {type, data}; }

// set_tensor -> get_tensor round trip of one weight tensor through the plugin's buffer (GGUF bytes must come back)
GH_API int gh_roundtrip(int which, int type, int64_t K, int64_t N, const void *in, void *out) {
    ggml_backend_t be = backend(which); if (!be) { snprintf(g_err, sizeof g_err, "no backend %d", which); return -1; }
    ggml_init_params p = {ggml_tensor_overhead() * 4, nullptr, true};
    ggml_context *c = ggml_init(p);
    ggml_tensor *t = ggml_new_tensor_2d(c, (ggml_type)type, K, N);
    ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors(c, be);
    if (!buf) { ggml_free(c); snprintf(g_err, sizeof g_err, "alloc failed"); return -2; }
    const size_t nb = ggml_nbytes(t), half = (nb / 2) & ~(size_t)1;
    ggml_backend_tensor_set(t, in, 0, nb);
    if (half) ggml_backend_tensor_set(t, (const char *)in + half, half, nb - half);     // partial overwrite with the same bytes
    ggml_backend_tensor_get(t, out, 0, nb);
    ggml_backend_buffer_free(buf); ggml_free(c);
    return 0;
}

struct Built { ggml_tensor *tok, *pos, *mask32, *hid, *res_norm, *logits; std::vector<ggml_tensor *> kc, vc; };

static ggml_tensor *wt(ggml_context *cw, std::vector<std::pair<ggml_tensor *, const void *>> &ups, const GTensor &g, int64_t ne0, int64_t ne1) {
    ggml_tensor *t = ggml_new_tensor_2d(cw, (ggml_type)g.type, ne0, ne1);
    ups.push_back({t, g.data});
    return t;
}
static ggml_tensor *norm(ggml_context *c, ggml_tensor *x, ggml_tensor *w, float eps) {     // llm_build_norm R/src/llama.cpp:329-365
    ggml_tensor *r = ggml_rms_norm(c, x, eps);
    return w ? ggml_mul(c, r, w) : r;
}

// One forward of the model over T tree tokens; K/V caches are host arrays [L][n_ctx][E] f16 that are
// uploaded before and downloaded after (so the caller sees the stored cells).  vis [T][n_kv] 0/1.
GH_API int gh_forward(int which, const GModel *m, uint16_t *kv_k, uint16_t *kv_v, int n_ctx, const int32_t *tokens, const int32_t *pos,
                      const float *hidden_in, int T, int kv_head, int n_kv, const uint8_t *vis, float *logits, float *feat, int *n_nodes_out,
                      int use_sched, int *n_splits_out) {
    ggml_backend_t be = backend(which); if (!be) { snprintf(g_err, sizeof g_err, "no backend %d", which); return -1; }
    const bool fa_off = (use_sched & 2) != 0;      // bit 1: build the attention block without flash attention (V cache transposed)
    use_sched &= 1;
    const int E = m->E, H = m->H, F = m->F, D = E / H, V = m->V, L = m->L;
    ggml_init_params pw = {ggml_tensor_overhead() * (size_t)(16 + 16 * L), nullptr, true};
    ggml_init_params pc = {ggml_tensor_overhead() * (size_t)(256 + 64 * L) + ggml_graph_overhead_custom(8192, false), nullptr, true};
    ggml_context *cw = ggml_init(pw), *c = ggml_init(pc);
    std::vector<std::pair<ggml_tensor *, const void *>> ups;
    Built b{};

    // ---- weights + KV cache: static tensors in a backend buffer (llama_model_loader / llama_kv_cache_init)
    ggml_tensor *tok_embd = wt(cw, ups, m->g[T_TOK_EMBD], E, V);
    ggml_tensor *out_norm = m->g[T_OUTPUT_NORM].data ? wt(cw, ups, m->g[T_OUTPUT_NORM], E, 1) : nullptr;
    ggml_tensor *output = logits ? wt(cw, ups, m->g[T_OUTPUT], E, V) : nullptr;
    ggml_tensor *fc = nullptr, *fc_b = nullptr;
    if (m->is_eagle) { fc = wt(cw, ups, m->g[T_FC], 2 * E, E); if (m->g[T_FC_BIAS].data) fc_b = wt(cw, ups, m->g[T_FC_BIAS], E, 1); }
    struct LW { ggml_tensor *an, *wq, *wk, *wv, *wo, *fn, *gate, *up, *down; };
    std::vector<LW> lw(L);
    for (int l = 0; l < L; l++) {
        const std::vector<GTensor> &w = m->layers[l];
        lw[l].an = w[L_ATTN_NORM].data ? wt(cw, ups, w[L_ATTN_NORM], E, 1) : nullptr;
        lw[l].wq = wt(cw, ups, w[L_WQ], E, E); lw[l].wk = wt(cw, ups, w[L_WK], E, E); lw[l].wv = wt(cw, ups, w[L_WV], E, E); lw[l].wo = wt(cw, ups, w[L_WO], E, E);
        lw[l].fn = w[L_FFN_NORM].data ? wt(cw, ups, w[L_FFN_NORM], E, 1) : nullptr;
        lw[l].gate = wt(cw, ups, w[L_GATE], E, F); lw[l].up = wt(cw, ups, w[L_UP], E, F); lw[l].down = wt(cw, ups, w[L_DOWN], F, E);
        b.kc.push_back(ggml_new_tensor_1d(cw, GGML_TYPE_F16, (int64_t)n_ctx * E));
        b.vc.push_back(ggml_new_tensor_1d(cw, GGML_TYPE_F16, (int64_t)n_ctx * E));
    }
    ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors(cw, be);
    if (!wbuf) { snprintf(g_err, sizeof g_err, "weight buffer allocation failed"); ggml_free(cw); ggml_free(c); return -2; }
    for (auto &u : ups) ggml_backend_tensor_set(u.first, u.second, 0, ggml_nbytes(u.first));
    for (int l = 0; l < L; l++) {
        ggml_backend_tensor_set(b.kc[l], kv_k + (size_t)l * n_ctx * E, 0, (size_t)n_ctx * E * 2);
        ggml_backend_tensor_set(b.vc[l], kv_v + (size_t)l * n_ctx * E, 0, (size_t)n_ctx * E * 2);
    }

    // ---- the graph
    ggml_cgraph *gf = ggml_new_graph_custom(c, 8192, false);
    const int Tp = GGML_PAD(T, GGML_KQ_MASK_PAD);
    b.tok = ggml_new_tensor_1d(c, GGML_TYPE_I32, T); ggml_set_input(b.tok);
    b.pos = ggml_new_tensor_1d(c, GGML_TYPE_I32, T); ggml_set_input(b.pos);
    b.mask32 = ggml_new_tensor_2d(c, GGML_TYPE_F32, n_kv, Tp); ggml_set_input(b.mask32);
    ggml_tensor *mask = ggml_cast(c, b.mask32, GGML_TYPE_F16);                            // R/src/llama.cpp:1488
    ggml_tensor *inpL = ggml_get_rows(c, tok_embd, b.tok);
    if (m->is_eagle) {                                                                    // build_eagle :1885-1900
        b.hid = ggml_new_tensor_2d(c, GGML_TYPE_F32, E, T); ggml_set_input(b.hid);
        ggml_tensor *r = ggml_mul_mat(c, fc, ggml_concat(c, inpL, b.hid, 0));
        if (fc_b) r = ggml_add(c, r, fc_b);
        inpL = ggml_relu(c, r);
    }
    const float kq_scale = 1.0f / sqrtf((float)D);
    const size_t row = (size_t)E * 2;
    for (int l = 0; l < L; l++) {
        ggml_tensor *inpSA = inpL;
        ggml_tensor *cur = norm(c, inpL, lw[l].an, m->eps);
        ggml_tensor *Q = ggml_mul_mat(c, lw[l].wq, cur), *K = ggml_mul_mat(c, lw[l].wk, cur), *Vc = ggml_mul_mat(c, lw[l].wv, cur);
        Q = ggml_rope_ext(c, ggml_reshape_3d(c, Q, D, H, T), b.pos, nullptr, D, 0, 0, m->freq_base, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
        K = ggml_rope_ext(c, ggml_reshape_3d(c, K, D, H, T), b.pos, nullptr, D, 0, 0, m->freq_base, 1.0f, 0.0f, 1.0f, 32.0f, 1.0f);
        ggml_build_forward_expand(gf, ggml_cpy(c, ggml_reshape_1d(c, K, (int64_t)T * E), ggml_view_1d(c, b.kc[l], (int64_t)T * E, row * kv_head)));    // llm_build_kv_store :228-270
        ggml_tensor *q = ggml_permute(c, Q, 0, 2, 1, 3);
        ggml_tensor *k3 = ggml_view_3d(c, b.kc[l], D, n_kv, H, row, (size_t)D * 2, 0);
        if (!fa_off) {                                                                    // llm_build_kqv :706-761, flash-attn branch
            ggml_build_forward_expand(gf, ggml_cpy(c, ggml_reshape_1d(c, Vc, (int64_t)T * E), ggml_view_1d(c, b.vc[l], (int64_t)T * E, row * kv_head)));
            ggml_tensor *v3 = ggml_view_3d(c, b.vc[l], D, n_kv, H, row, (size_t)D * 2, 0);
            cur = ggml_flash_attn_ext(c, q, k3, v3, mask, kq_scale, 0.0f, 0.0f);
            ggml_flash_attn_ext_set_prec(cur, GGML_PREC_F32);
            cur = ggml_reshape_2d(c, cur, E, T);
        } else {                                                                          // :762-812 and the transposed V store :258-264
            ggml_build_forward_expand(gf, ggml_cpy(c, ggml_transpose(c, Vc), ggml_view_2d(c, b.vc[l], T, E, (size_t)n_ctx * 2, (size_t)kv_head * 2)));
            ggml_tensor *kq = ggml_mul_mat(c, k3, q);
            ggml_mul_mat_set_prec(kq, GGML_PREC_F32);
            kq = ggml_soft_max_ext(c, kq, mask, kq_scale, 0.0f);
            ggml_tensor *v3 = ggml_view_3d(c, b.vc[l], n_kv, D, H, (size_t)n_ctx * 2, (size_t)n_ctx * D * 2, 0);
            ggml_tensor *kqv = ggml_mul_mat(c, v3, kq);
            cur = ggml_cont_2d(c, ggml_permute(c, kqv, 0, 2, 1, 3), E, T);
        }
        cur = ggml_mul_mat(c, lw[l].wo, cur);
        ggml_tensor *ffn_inp = ggml_add(c, cur, inpSA);
        cur = norm(c, ffn_inp, lw[l].fn, m->eps);
        ggml_tensor *up = ggml_mul_mat(c, lw[l].up, cur), *gt = ggml_silu(c, ggml_mul_mat(c, lw[l].gate, cur));   // llm_build_ffn :456-581
        cur = ggml_mul_mat(c, lw[l].down, ggml_mul(c, gt, up));
        inpL = ggml_add(c, cur, ffn_inp);
    }
    b.res_norm = norm(c, inpL, out_norm, m->eps);
    // use_sched: an op the plugin does not claim (NEG, twice: exact identity, and not an in-place view like ggml_clamp) sits between the norm and the lm_head, so ggml_backend_sched has to cut the graph
    // into plugin | CPU | plugin splits and move tensors across the backends in both directions
    if (use_sched) b.res_norm = ggml_neg(c, ggml_neg(c, b.res_norm));
    ggml_set_output(b.res_norm);
    ggml_build_forward_expand(gf, b.res_norm);
    if (logits) { b.logits = ggml_mul_mat(c, output, b.res_norm); ggml_set_output(b.logits); ggml_build_forward_expand(gf, b.logits); }

    ggml_backend_dev_t dev = ggml_backend_get_device(be);
    int rc = 0;
    ggml_gallocr_t ga = nullptr; ggml_backend_sched_t sched = nullptr;
    auto set_inputs = [&]() {
        std::vector<float> m32((size_t)n_kv * Tp);
        for (int j = 0; j < Tp; j++) for (int i = 0; i < n_kv; i++) m32[(size_t)j * n_kv + i] = (j < T && vis[(size_t)j * n_kv + i]) ? 0.0f : -INFINITY;
        ggml_backend_tensor_set(b.tok, tokens, 0, (size_t)T * 4);
        ggml_backend_tensor_set(b.pos, pos, 0, (size_t)T * 4);
        ggml_backend_tensor_set(b.mask32, m32.data(), 0, m32.size() * 4);
        if (b.hid) ggml_backend_tensor_set(b.hid, hidden_in, 0, (size_t)T * E * 4);
    };
    if (!use_sched) {
        // ---- every node must be claimed by the backend (this is what ggml_backend_sched asks before assigning a split)
        for (int i = 0; i < ggml_graph_n_nodes(gf) && !rc; i++)
            if (!ggml_backend_dev_supports_op(dev, ggml_graph_node(gf, i))) {
                snprintf(g_err, sizeof g_err, "backend %s refuses node %d: %s", ggml_backend_name(be), i, ggml_op_name(ggml_graph_node(gf, i)->op)); rc = -3;
            }
        ga = ggml_gallocr_new(ggml_backend_get_default_buffer_type(be));
        if (!rc && !ggml_gallocr_alloc_graph(ga, gf)) { snprintf(g_err, sizeof g_err, "graph allocation failed"); rc = -4; }
        if (!rc) {
            set_inputs();
            const ggml_status st = ggml_backend_graph_compute(be, gf);
            if (st != GGML_STATUS_SUCCESS) { snprintf(g_err, sizeof g_err, "graph_compute -> %s", ggml_status_to_string(st)); rc = -5; }
        }
    } else {
        // ---- the way libllama runs it: ggml_backend_sched over {this backend, CPU} (R/src/llama.cpp:11700-11760, R/ggml/src/ggml-backend.cpp:1448)
        ggml_backend_t bes[2] = {be, backend(0)};
        const int nb = which == 0 ? 1 : 2;
        sched = ggml_backend_sched_new(which == 0 ? bes + 1 : bes, nullptr, nb, 8192, false);
        if (!sched || !ggml_backend_sched_alloc_graph(sched, gf)) { snprintf(g_err, sizeof g_err, "sched allocation failed"); rc = -4; }
        if (!rc) {
            set_inputs();
            const ggml_status st = ggml_backend_sched_graph_compute(sched, gf);
            if (st != GGML_STATUS_SUCCESS) { snprintf(g_err, sizeof g_err, "sched_graph_compute -> %s", ggml_status_to_string(st)); rc = -5; }
            if (n_splits_out) *n_splits_out = ggml_backend_sched_get_n_splits(sched);
            // every MUL_MAT and the attention must have been placed on the plugin, the NEGs on the CPU
            for (int i = 0; i < ggml_graph_n_nodes(gf) && !rc && which == 1; i++) {
                ggml_tensor *nd = ggml_graph_node(gf, i);
                const bool on_plugin = ggml_backend_sched_get_tensor_backend(sched, nd) == be;
                if ((nd->op == GGML_OP_MUL_MAT || nd->op == GGML_OP_FLASH_ATTN_EXT) && !on_plugin) { snprintf(g_err, sizeof g_err, "node %d (%s) was not given to the plugin", i, ggml_op_name(nd->op)); rc = -6; }
                if (nd->op == GGML_OP_UNARY && ggml_get_unary_op(nd) == GGML_UNARY_OP_NEG && on_plugin) { snprintf(g_err, sizeof g_err, "NEG was given to the plugin"); rc = -6; }
            }
        }
    }
    if (!rc) {
        if (feat) ggml_backend_tensor_get(b.res_norm, feat, 0, (size_t)T * E * 4);
        if (logits) ggml_backend_tensor_get(b.logits, logits, 0, (size_t)T * V * 4);
        for (int l = 0; l < L; l++) {
            ggml_backend_tensor_get(b.kc[l], kv_k + (size_t)l * n_ctx * E, 0, (size_t)n_ctx * E * 2);
            ggml_backend_tensor_get(b.vc[l], kv_v + (size_t)l * n_ctx * E, 0, (size_t)n_ctx * E * 2);
        }
        if (n_nodes_out) *n_nodes_out = ggml_graph_n_nodes(gf);
    }
    if (ga) ggml_gallocr_free(ga);
    if (sched) ggml_backend_sched_free(sched);
    ggml_backend_buffer_free(wbuf); ggml_free(c); ggml_free(cw);
    return rc;
}

// Two ggml_backend_t of the plugin's device = two streams, chained the way ggml_backend_sched chains pipeline copies
// (R/ggml/src/ggml-backend.cpp: event_record after a split, event_wait + tensor_copy_async before the next): backend A uploads x
// (set_tensor_async) and runs a long chain y = (chain + 1) * x, records an event; backend B waits for it, takes y with
// cpy_tensor_async, computes z = y * y and hands it back with get_tensor_async.  Without stream ordering B would read y before A wrote it.
GH_API int gh_two_backends(int64_t n, int chain, const float *x_host, float *z_host, int *caps_out) {
    if (!g_reg) { snprintf(g_err, sizeof g_err, "plugin not loaded"); return -1; }
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(g_reg, 0);
    ggml_backend_dev_props props; ggml_backend_dev_get_props(dev, &props);
    if (caps_out) *caps_out = (props.caps.async ? 1 : 0) | (props.caps.events ? 2 : 0);
    if (!props.caps.async || !props.caps.events) { snprintf(g_err, sizeof g_err, "device reports async %d events %d", props.caps.async, props.caps.events); return -2; }
    ggml_backend_t A = ggml_backend_dev_init(dev, nullptr), B = ggml_backend_dev_init(dev, nullptr);
    ggml_backend_event_t ev = ggml_backend_event_new(dev);
    if (!A || !B || !ev || A == B) { snprintf(g_err, sizeof g_err, "could not create two backends + an event"); return -3; }
    ggml_init_params p = {ggml_tensor_overhead() * (size_t)(chain + 8) + ggml_graph_overhead() * 2, nullptr, true};
    ggml_context *ca = ggml_init(p), *cb = ggml_init(p);
    ggml_tensor *x = ggml_new_tensor_1d(ca, GGML_TYPE_F32, n), *y = x;
    for (int i = 0; i < chain; i++) y = ggml_add(ca, y, x);
    ggml_cgraph *ga = ggml_new_graph(ca); ggml_build_forward_expand(ga, y);
    ggml_tensor *y2 = ggml_new_tensor_1d(cb, GGML_TYPE_F32, n), *z = ggml_mul(cb, y2, y2);
    ggml_cgraph *gb = ggml_new_graph(cb); ggml_build_forward_expand(gb, z);
    ggml_backend_buffer_t ba = ggml_backend_alloc_ctx_tensors(ca, A), bb = ggml_backend_alloc_ctx_tensors(cb, B);
    int rc = 0;
    if (!ba || !bb) { snprintf(g_err, sizeof g_err, "alloc failed"); rc = -4; }
    if (!rc) {
        ggml_backend_tensor_set_async(A, x, x_host, 0, (size_t)n * 4);
        if (ggml_backend_graph_compute_async(A, ga) != GGML_STATUS_SUCCESS) { snprintf(g_err, sizeof g_err, "graph A failed"); rc = -5; }
        ggml_backend_event_record(ev, A);
        ggml_backend_event_wait(B, ev);
        ggml_backend_tensor_copy_async(A, B, y, y2);
        if (!rc && ggml_backend_graph_compute_async(B, gb) != GGML_STATUS_SUCCESS) { snprintf(g_err, sizeof g_err, "graph B failed"); rc = -6; }
        ggml_backend_tensor_get_async(B, z, z_host, 0, (size_t)n * 4);
        ggml_backend_synchronize(B);
        ggml_backend_event_synchronize(ev);
        ggml_backend_synchronize(A);
    }
    if (ba) ggml_backend_buffer_free(ba);
    if (bb) ggml_backend_buffer_free(bb);
    ggml_backend_event_free(ev);
    ggml_backend_free(A); ggml_backend_free(B);
    ggml_free(ca); ggml_free(cb);
    return rc;
}
