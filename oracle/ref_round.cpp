// ref_round.cpp — TEST INFRASTRUCTURE: a thin C harness over the REFERENCE's own libllama + common (compiled from the
// sources where they lie by oracle/Makefile into oracle/_ref/libllama-ref.so / libcommon-ref.so).  Nothing here restates
// reference logic: every function below only marshals arguments into reference calls, so that tests/ can pin
//   * llama_decode on a tree-shaped llama_batch (R/src/llama.cpp:11916 -> llama_decode_impl :9486), on ggml-cpu or on any
//     backend plug-in loaded with ggml_backend_load (R/ggml/src/ggml-backend-reg.cpp:393) — the drop-in check,
//   * the KQ mask libllama builds for that batch (llama_set_inputs_initial, R/src/llama-context.cpp:475, rule :136-201),
//   * the sampler chains of the drivers (top-k 10 draft chain R/common/speculative.cpp:47-58; greedy recipe top-k 1;
//     llama_sampler_top_k_impl R/src/llama-sampling.cpp:228-302) and the chain accept walk
//     common_sampler_sample_and_accept_n (R/common/sampling.cpp:423-458),
//   * the KV-cache sequence operations (R/src/llama-kv-cache.cpp:368,431,478) + defrag (R/src/llama.cpp:10941).
// Only tests/, __graft_entry__.smoke() and bench.py's CPU arm load this library.
#include "llama.h"
#include "llama-context.h"
#include "llama-kv-cache.h"
#include "common.h"
#include "sampling.h"
#include "speculative.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"

#include <cstring>
#include <string>
#include <vector>

struct rr_tap_rec { std::string weight; int64_t K, T; std::vector<float> x; };
struct rr_ctx {
    llama_context *ctx = nullptr;
    bool tap_on = false; std::vector<rr_tap_rec> taps;   // f32 src1 of every MUL_MAT over a quantised weight, in graph order
    std::vector<float> mask;        // KQ mask of the last decode, captured before the first node runs
    int64_t mask_ne0 = 0, mask_ne1 = 0;
    bool want_mask = false;
};

static bool rr_eval_cb(struct ggml_tensor *t, bool ask, void *user) {
    (void)t;
    rr_ctx *r = (rr_ctx *)user;
    if (ask && r->want_mask && r->ctx && r->ctx->inp_KQ_mask) {
        ggml_tensor *m = r->ctx->inp_KQ_mask;
        r->mask.resize(ggml_nelements(m));
        ggml_backend_tensor_get(m, r->mask.data(), 0, ggml_nbytes(m));
        r->mask_ne0 = m->ne[0]; r->mask_ne1 = m->ne[1];
        r->want_mask = false;
    }
    // tap: the f32 activation operand of every MUL_MAT on a quantised weight (what the CPU backend's from_float quantises,
    // R/ggml/src/ggml-cpu/ggml-cpu.c:7588-7623), read right after the node ran
    if (r->tap_on && t->op == GGML_OP_MUL_MAT && ggml_is_quantized(t->src[0]->type) && t->src[1]->type == GGML_TYPE_F32) {
        if (ask) return true;
        const ggml_tensor *x = t->src[1];
        rr_tap_rec rec; rec.weight = t->src[0]->name; rec.K = x->ne[0]; rec.T = x->ne[1];
        rec.x.resize((size_t)rec.K * rec.T);
        ggml_backend_tensor_get(x, rec.x.data(), 0, rec.x.size() * sizeof(float));
        r->taps.push_back(std::move(rec));
        return true;
    }
    return false;                   // we never need any other node's data: the scheduler runs the graph as usual
}

static void rr_log_quiet(ggml_log_level level, const char *text, void *) {
    if (level >= GGML_LOG_LEVEL_ERROR) fputs(text, stderr);
}

extern "C" {

// loads a ggml backend plug-in through the reference's loader; returns the number of registered devices (or -1)
int rr_load_backend(const char *path) {
    llama_log_set(rr_log_quiet, nullptr);
    if (path && *path && !ggml_backend_load(path)) return -1;
    return (int)ggml_backend_dev_count();
}
int rr_dev_count(void) { return (int)ggml_backend_dev_count(); }
const char *rr_dev_name(int i) { return ggml_backend_dev_name(ggml_backend_dev_get(i)); }

void *rr_model_load(const char *path, int n_gpu_layers) {
    llama_log_set(rr_log_quiet, nullptr);
    llama_backend_init();
    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = n_gpu_layers;
    mp.use_mmap = false;
    return llama_model_load_from_file(path, mp);
}
void rr_model_free(void *m) { llama_model_free((llama_model *)m); }
int rr_model_n_vocab(void *m) { return llama_vocab_n_tokens(llama_model_get_vocab((llama_model *)m)); }
// ggml type of a tensor of the loaded model (by GGUF name), -1 if absent: pins llama-quantize's type mixture
int rr_model_tensor_type(void *m, const char *name) {
    for (const auto &it : ((llama_model *)m)->tensors_by_name) if (it.first == name) return (int)it.second->type;
    return -1;
}

// embeddings = true + pooling NONE: the fork only allocates logits when `embeddings` is on (R/src/llama-context.cpp:898-902)
void *rr_ctx_new(void *model, int n_ctx, int n_batch, int n_seq_max, int n_threads, int flash_attn, int type_kv) {
    rr_ctx *r = new rr_ctx();
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = n_ctx; cp.n_batch = n_batch; cp.n_ubatch = n_batch; cp.n_seq_max = n_seq_max;
    cp.n_threads = n_threads; cp.n_threads_batch = n_threads;
    cp.flash_attn = flash_attn != 0;
    cp.embeddings = true; cp.pooling_type = LLAMA_POOLING_TYPE_NONE;
    cp.type_k = (ggml_type)type_kv; cp.type_v = (ggml_type)type_kv;
    cp.defrag_thold = -1.0f;        // defrag only when the test asks for it
    cp.cb_eval = rr_eval_cb; cp.cb_eval_user_data = r;
    r->ctx = llama_init_from_model((llama_model *)model, cp);
    if (!r->ctx) { delete r; return nullptr; }
    return r;
}
void rr_ctx_free(void *c) { rr_ctx *r = (rr_ctx *)c; if (r) { llama_free(r->ctx); delete r; } }

// llama_decode on a (tree-shaped) batch: token, pos, seq-id sets (seq ids: row i at seq[i*max_seq ..], n_seq[i] valid); logits
// requested for every token; copies the [n][V] logits out (row i = batch token i).  Returns llama_decode's status.
int rr_decode(void *c, int n, const int32_t *tok, const int32_t *pos, const int32_t *n_seq, const int32_t *seq, int max_seq, float *logits_out) {
    rr_ctx *r = (rr_ctx *)c;
    llama_batch b = llama_batch_init(n, 0, max_seq);
    for (int i = 0; i < n; i++) {
        std::vector<llama_seq_id> ids(seq + (size_t)i * max_seq, seq + (size_t)i * max_seq + n_seq[i]);
        common_batch_add(b, tok[i], pos[i], ids, true);
    }
    r->want_mask = true;
    const int rc = llama_decode(r->ctx, b);
    if (rc == 0 && logits_out) {
        const int V = llama_vocab_n_tokens(llama_model_get_vocab(llama_get_model(r->ctx)));
        for (int i = 0; i < n; i++) memcpy(logits_out + (size_t)i * V, llama_get_logits_ith(r->ctx, i), (size_t)V * sizeof(float));
    }
    llama_batch_free(b);
    return rc;
}
// activation tap for the NEXT rr_decode calls (on = 1 clears the list)
void rr_tap_enable(void *c, int on) { rr_ctx *r = (rr_ctx *)c; r->tap_on = on != 0; if (on) r->taps.clear(); }
int rr_tap_count(void *c) { return (int)((rr_ctx *)c)->taps.size(); }
int rr_tap_info(void *c, int i, char *weight_name, int cap, int64_t *K, int64_t *T) {
    const rr_tap_rec &t = ((rr_ctx *)c)->taps[i];
    snprintf(weight_name, cap, "%s", t.weight.c_str()); *K = t.K; *T = t.T;
    return 0;
}
int rr_tap_data(void *c, int i, float *out) { const rr_tap_rec &t = ((rr_ctx *)c)->taps[i]; memcpy(out, t.x.data(), t.x.size() * sizeof(float)); return 0; }

// the KQ mask libllama built for the last rr_decode: f32 [ne1 = pad(n_tokens, 64)][ne0 = n_kv]; returns element count
int64_t rr_kq_mask(void *c, float *out, int64_t cap, int64_t *ne0, int64_t *ne1) {
    rr_ctx *r = (rr_ctx *)c;
    *ne0 = r->mask_ne0; *ne1 = r->mask_ne1;
    const int64_t n = (int64_t)r->mask.size();
    if (out && cap >= n) memcpy(out, r->mask.data(), n * sizeof(float));
    return n;
}
// overwrite the logits of the last decode (rows 0..n_rows-1): lets the sampler tests feed arbitrary logits to the reference's code
int rr_set_logits(void *c, const float *logits, int n_rows) {
    rr_ctx *r = (rr_ctx *)c;
    const int V = llama_vocab_n_tokens(llama_model_get_vocab(llama_get_model(r->ctx)));
    for (int i = 0; i < n_rows; i++) memcpy((float *)llama_get_logits_ith(r->ctx, i), logits + (size_t)i * V, (size_t)V * sizeof(float));
    return 0;
}

// the draft sampler of common_speculative_init (R/common/speculative.cpp:47-58: samplers = {TOP_K}, top_k = k) applied to
// logits row idx exactly as common_speculative_gen_draft does (:257-269): candidates sorted, p = softmax over the kept k
int rr_draft_candidates(void *c, int idx, int k, int32_t *ids, float *logits, float *probs) {
    rr_ctx *r = (rr_ctx *)c;
    common_params_sampling params;
    params.no_perf = false;
    params.top_k = k;
    params.samplers = { COMMON_SAMPLER_TYPE_TOP_K };
    common_sampler *smpl = common_sampler_init(llama_get_model(r->ctx), params);
    common_sampler_sample(smpl, r->ctx, idx, true);
    const llama_token_data_array *cur = common_sampler_get_candidates(smpl);
    const int n = (int)cur->size;
    for (int i = 0; i < n && i < k; i++) { ids[i] = cur->data[i].id; logits[i] = cur->data[i].logit; probs[i] = cur->data[i].p; }
    common_sampler_free(smpl);
    return n;
}
// the documented greedy recipe (--sampling-seq k --top-k 1 --temp 0, R/examples/speculative-simple/README.md:6,10) through
// common_sampler_sample_and_accept_n (R/common/sampling.cpp:423-458): rows 0..n_draft of the last decode are the chain
// [id_last, draft...]; returns the accepted ids (1..n_draft+1)
int rr_accept_n(void *c, const int32_t *draft, int n_draft, int32_t *out_ids) {
    rr_ctx *r = (rr_ctx *)c;
    common_params_sampling params;
    params.top_k = 1; params.temp = 0.0f;
    params.samplers = { COMMON_SAMPLER_TYPE_TOP_K };
    common_sampler *smpl = common_sampler_init(llama_get_model(r->ctx), params);
    llama_tokens d(draft, draft + n_draft);
    const std::vector<llama_token> ids = common_sampler_sample_and_accept_n(smpl, r->ctx, d);
    for (size_t i = 0; i < ids.size(); i++) out_ids[i] = ids[i];
    common_sampler_free(smpl);
    return (int)ids.size();
}

// ---- KV cache sequence operations (public API) + metadata / data read-back (internals)
void rr_kv_seq_rm(void *c, int seq, int p0, int p1) { llama_kv_cache_seq_rm(((rr_ctx *)c)->ctx, seq, p0, p1); }
void rr_kv_seq_cp(void *c, int src, int dst, int p0, int p1) { llama_kv_cache_seq_cp(((rr_ctx *)c)->ctx, src, dst, p0, p1); }
void rr_kv_seq_keep(void *c, int seq) { llama_kv_cache_seq_keep(((rr_ctx *)c)->ctx, seq); }
void rr_kv_clear(void *c) { llama_kv_cache_clear(((rr_ctx *)c)->ctx); }
void rr_kv_defrag(void *c) { llama_kv_cache_defrag(((rr_ctx *)c)->ctx); llama_kv_cache_update(((rr_ctx *)c)->ctx); }
// per cell: pos (-1 = empty) and the seq-id set as a bit mask; returns the cache size
int rr_kv_cells(void *c, int32_t *pos, uint64_t *seq_mask, int cap) {
    const llama_kv_cache &kv = ((rr_ctx *)c)->ctx->kv_self;
    for (uint32_t i = 0; i < kv.size && (int)i < cap; i++) {
        pos[i] = kv.cells[i].pos;
        uint64_t m = 0;
        for (llama_seq_id s : kv.cells[i].seq_id) if (s >= 0 && s < 64) m |= 1ull << s;
        seq_mask[i] = m;
    }
    return (int)kv.size;
}
// K (which = 0) or V (which = 1) row of one cell of one layer as raw f16 (flash-attention layout: V not transposed)
int rr_kv_read_row(void *c, int layer, int which, int cell, uint16_t *out, int n_embd) {
    const llama_kv_cache &kv = ((rr_ctx *)c)->ctx->kv_self;
    ggml_tensor *t = which ? kv.v_l[layer] : kv.k_l[layer];
    if (t->type != GGML_TYPE_F16 || (which && kv.v_trans)) return -1;
    ggml_backend_tensor_get(t, out, (size_t)cell * n_embd * 2, (size_t)n_embd * 2);
    return 0;
}

// synthetic committed prefix (full-size parity runs and the timed CPU arm: a 4K-token CPU prefill of a 7B model would take minutes):
// raw f16 rows [n][n_embd] written into one layer's K / V cache tensor from cell0, and the cell metadata of a prefix of n cells at
// positions 0..n-1 that belongs to sequences 0..n_seq-1 (what the cache looks like after the driver's seq_cp 0 -> s, speculative-eagle.cpp:648-652)
int rr_kv_write_rows(void *c, int layer, int which, int cell0, int n, const uint16_t *rows, int n_embd) {
    const llama_kv_cache &kv = ((rr_ctx *)c)->ctx->kv_self;
    ggml_tensor *t = which ? kv.v_l[layer] : kv.k_l[layer];
    if (t->type != GGML_TYPE_F16 || (which && kv.v_trans)) return -1;
    ggml_backend_tensor_set(t, rows, (size_t)cell0 * n_embd * 2, (size_t)n * n_embd * 2);
    return 0;
}
int rr_kv_set_prefix(void *c, int n, int n_seq) {
    llama_kv_cache &kv = ((rr_ctx *)c)->ctx->kv_self;
    if ((uint32_t)n > kv.size) return -1;
    for (uint32_t i = 0; i < kv.size; i++) { kv.cells[i].pos = -1; kv.cells[i].seq_id.clear(); kv.cells[i].delta = 0; }
    for (int i = 0; i < n; i++) { kv.cells[i].pos = i; for (int s = 0; s < n_seq; s++) kv.cells[i].seq_id.insert(s); }
    kv.head = n; kv.used = n;
    return 0;
}

// time one llama_decode of a tree batch (the CPU arm of bench.py): returns microseconds (ggml_time_us), < 0 on failure
int64_t rr_time_decode(void *c, int n, const int32_t *tok, const int32_t *pos, const int32_t *n_seq, const int32_t *seq, int max_seq) {
    rr_ctx *r = (rr_ctx *)c;
    llama_batch b = llama_batch_init(n, 0, max_seq);
    for (int i = 0; i < n; i++) {
        std::vector<llama_seq_id> ids(seq + (size_t)i * max_seq, seq + (size_t)i * max_seq + n_seq[i]);
        common_batch_add(b, tok[i], pos[i], ids, true);
    }
    r->want_mask = false;
    const int64_t t0 = ggml_time_us();
    const int rc = llama_decode(r->ctx, b);
    llama_synchronize(r->ctx);
    const int64_t t1 = ggml_time_us();
    llama_batch_free(b);
    return rc == 0 ? t1 - t0 : -1;
}

}  // extern "C"
