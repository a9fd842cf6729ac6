// gguf_load.cu — GGUF file -> resident eb_model (the on-disk format in front of the path).
// Replaces, for the tensors of this path, the reference's loader walk (gguf_init_from_file R/ggml/src/gguf.cpp, llama_model_loader
// R/src/llama-model-loader.cpp: one ggml_backend_tensor_set per tensor from an mmap or a bounce buffer) with a streaming upload:
// the file is read in 32 MB pieces into two page-locked buffers, each piece goes to the device with cudaMemcpyAsync while the next
// one is being read, Q4_K / f16 / f32 tensors land directly in their final allocation, plane-layout types (Q6_K, Q4_0, Q8_0) in one of
// two device staging areas and are repacked on the stream.  The format itself is parsed here (GGUF v2 / v3, little endian:
// header, key/value section with all 13 value types, tensor infos, aligned data section); arch "llama" (target) and "eagle" (draft
// head, R/src/llama-arch.cpp:1301-1326) with the tensor names of R/src/llama-arch.cpp.  No tokenizer data is interpreted.
#include "common.cuh"
#include "engine.cuh"
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

namespace {

struct GTensorInfo { std::string name; int n_dims = 0; uint64_t dims[4] = {1, 1, 1, 1}; int type = -1; uint64_t offset = 0; };
struct GFile {
    std::string arch; uint32_t alignment = 32, version = 0;
    std::map<std::string, double> num;               // every scalar key (integers and floats fit a double exactly at these magnitudes)
    std::vector<GTensorInfo> tensors; uint64_t data_off = 0, file_size = 0;
};

struct Cursor {
    const uint8_t *p, *end; bool ok = true;
    template <typename T> T get() { T v{}; if (p + sizeof(T) > end) { ok = false; return v; } memcpy(&v, p, sizeof(T)); p += sizeof(T); return v; }
    std::string str() { const uint64_t n = get<uint64_t>(); if (!ok || n > (uint64_t)(end - p)) { ok = false; return ""; } std::string s((const char *)p, n); p += n; return s; }
    void skip(uint64_t n) { if (n > (uint64_t)(end - p)) ok = false; else p += n; }
};
// GGUF value types (R/ggml/include/gguf.h): 0 u8, 1 i8, 2 u16, 3 i16, 4 u32, 5 i32, 6 f32, 7 bool, 8 string, 9 array, 10 u64, 11 i64, 12 f64
size_t scalar_size(uint32_t t) { static const size_t s[13] = {1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8}; return t < 13 ? s[t] : 0; }
bool read_scalar(Cursor &c, uint32_t t, double *out) {
    switch (t) {
        case 0: *out = c.get<uint8_t>(); return true;   case 1: *out = c.get<int8_t>(); return true;
        case 2: *out = c.get<uint16_t>(); return true;  case 3: *out = c.get<int16_t>(); return true;
        case 4: *out = c.get<uint32_t>(); return true;  case 5: *out = c.get<int32_t>(); return true;
        case 6: *out = c.get<float>(); return true;     case 7: *out = c.get<uint8_t>(); return true;
        case 10: *out = (double)c.get<uint64_t>(); return true; case 11: *out = (double)c.get<int64_t>(); return true;
        case 12: *out = c.get<double>(); return true;
    }
    return false;
}

int parse(const uint8_t *base, uint64_t size, GFile &g) {
    Cursor c{base, base + size};
    if (size < 24 || memcmp(base, "GGUF", 4)) return eb_set_err(EB_ERR_ARG, "gguf: bad magic");
    c.skip(4);
    g.version = c.get<uint32_t>();
    if (g.version < 2 || g.version > 3) return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: version %u", g.version);
    const uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
    if (n_tensors > (1u << 20) || n_kv > (1u << 20)) return eb_set_err(EB_ERR_ARG, "gguf: implausible counts");
    for (uint64_t i = 0; i < n_kv && c.ok; i++) {
        const std::string key = c.str(); const uint32_t t = c.get<uint32_t>();
        if (t == 8) { const std::string v = c.str(); if (key == "general.architecture") g.arch = v; }
        else if (t == 9) {
            const uint32_t et = c.get<uint32_t>(); const uint64_t n = c.get<uint64_t>();
            if (et == 8) { for (uint64_t k = 0; k < n && c.ok; k++) c.str(); }
            else if (scalar_size(et)) c.skip(n * scalar_size(et));
            else return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: array of type %u in key %s", et, key.c_str());
        } else {
            double v; if (!read_scalar(c, t, &v)) return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: value type %u in key %s", t, key.c_str());
            g.num[key] = v;
        }
    }
    if (g.num.count("general.alignment")) g.alignment = (uint32_t)g.num["general.alignment"];
    if (!g.alignment || (g.alignment & (g.alignment - 1))) return eb_set_err(EB_ERR_ARG, "gguf: alignment %u", g.alignment);
    g.tensors.resize(n_tensors);
    for (auto &t : g.tensors) {
        t.name = c.str(); t.n_dims = (int)c.get<uint32_t>();
        if (!c.ok || t.n_dims < 1 || t.n_dims > 4) return eb_set_err(EB_ERR_ARG, "gguf: tensor info");
        for (int d = 0; d < t.n_dims; d++) t.dims[d] = c.get<uint64_t>();
        t.type = (int)c.get<uint32_t>(); t.offset = c.get<uint64_t>();
    }
    if (!c.ok) return eb_set_err(EB_ERR_ARG, "gguf: truncated header");
    const uint64_t pos = (uint64_t)(c.p - base);
    g.data_off = (pos + g.alignment - 1) / g.alignment * g.alignment; g.file_size = size;
    return EB_OK;
}

// tensor name -> (layer, which); false for tensors this path does not use (rope_freqs, tokenizer tables ...)
bool slot_of(const std::string &n, int *layer, int *which) {
    static const struct { const char *n; int w; } G[] = {{"token_embd.weight", EB_T_TOK_EMBD}, {"output_norm.weight", EB_T_OUTPUT_NORM}, {"output.weight", EB_T_OUTPUT},
                                                          {"fc.weight", EB_T_FC}, {"fc.bias", EB_T_FC_BIAS}};
    for (auto &e : G) if (n == e.n) { *layer = -1; *which = e.w; return true; }
    static const struct { const char *n; int w; } Ls[] = {{"attn_norm", EB_L_ATTN_NORM}, {"attn_q", EB_L_WQ}, {"attn_k", EB_L_WK}, {"attn_v", EB_L_WV}, {"attn_output", EB_L_WO},
                                                           {"ffn_norm", EB_L_FFN_NORM}, {"ffn_gate", EB_L_GATE}, {"ffn_up", EB_L_UP}, {"ffn_down", EB_L_DOWN}};
    int l = -1, used = 0; char mid[64];
    if (sscanf(n.c_str(), "blk.%d.%63[a-z_].weight%n", &l, mid, &used) == 2 && used == (int)n.size())
        for (auto &e : Ls) if (!strcmp(mid, e.n)) { *layer = l; *which = e.w; return true; }
    return false;
}
double key(const GFile &g, const char *suffix, double dflt) {
    auto it = g.num.find(g.arch + "." + suffix);
    return it == g.num.end() ? dflt : it->second;
}
int hparams_of(const GFile &g, int hp[6], float hf[2]) {
    if (g.arch != "llama" && g.arch != "eagle") return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: architecture '%s' (this path: llama, eagle)", g.arch.c_str());
    hp[0] = (int)key(g, "embedding_length", 0); hp[1] = (int)key(g, "attention.head_count", 0); hp[2] = (int)key(g, "feed_forward_length", 0);
    hp[3] = (int)key(g, "vocab_size", 0); hp[4] = (int)key(g, "block_count", 0); hp[5] = g.arch == "eagle";
    if (!hp[3]) for (auto &t : g.tensors) if (t.name == "token_embd.weight") hp[3] = (int)t.dims[1];
    if (key(g, "attention.head_count_kv", hp[1]) != hp[1]) return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: grouped-query attention is outside this path");
    hf[0] = (float)key(g, "attention.layer_norm_rms_epsilon", 1e-5); hf[1] = (float)key(g, "rope.freq_base", 10000.0);
    if (hp[0] <= 0 || hp[1] <= 0 || hp[2] <= 0 || hp[3] <= 0 || hp[4] <= 0) return eb_set_err(EB_ERR_ARG, "gguf: missing hyper-parameters (E %d H %d F %d V %d L %d)", hp[0], hp[1], hp[2], hp[3], hp[4]);
    return EB_OK;
}

// one piece of the file into page-locked memory with several threads (one pread stream copies ~7 GB/s out of the page cache; the
// H2D copy of the previous piece runs meanwhile)
bool read_parallel(int fd, uint8_t *dst, size_t n, uint64_t foff) {
    constexpr int NT = 4;
    auto part = [&](size_t lo, size_t hi) {
        while (lo < hi) { const ssize_t r = pread(fd, dst + lo, hi - lo, (off_t)(foff + lo)); if (r <= 0) return false; lo += (size_t)r; }
        return true;
    };
    if (n < (4u << 20)) return part(0, n);
    bool ok[NT]; std::thread th[NT - 1];
    const size_t per = (n / NT + 4095) & ~(size_t)4095;
    for (int i = 1; i < NT; i++) th[i - 1] = std::thread([&, i] { ok[i] = part(std::min(n, i * per), std::min(n, (i + 1) * per)); });
    ok[0] = part(0, std::min(n, per));
    bool all = ok[0];
    for (int i = 1; i < NT; i++) { th[i - 1].join(); all = all && ok[i]; }
    return all;
}

struct Mapped {
    int fd = -1; const uint8_t *base = nullptr; uint64_t size = 0;
    int open_file(const char *path) {
        fd = open(path, O_RDONLY);
        if (fd < 0) return eb_set_err(EB_ERR_ARG, "gguf: cannot open %s", path);
        struct stat st; if (fstat(fd, &st) || st.st_size < 24) return eb_set_err(EB_ERR_ARG, "gguf: %s is not a GGUF file", path);
        size = (uint64_t)st.st_size;
        void *m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return eb_set_err(EB_ERR_ARG, "gguf: mmap(%s) failed", path);
        base = (const uint8_t *)m;
        return EB_OK;
    }
    ~Mapped() { if (base) munmap((void *)base, size); if (fd >= 0) close(fd); }
};
}  // namespace

// header only, no device needed: hp = {E, H, F, V, L, is_eagle}, hf = {rms eps, rope freq base}
extern "C" int eb_gguf_probe(const char *path, int32_t *hp, float *hf, int32_t *n_tensors_used, uint64_t *bytes_used) {
    Mapped f; int rc = f.open_file(path); if (rc) return rc;
    GFile g; if ((rc = parse(f.base, f.size, g))) return rc;
    int h[6]; float q[2]; if ((rc = hparams_of(g, h, q))) return rc;
    for (int i = 0; i < 6; i++) hp[i] = h[i];
    hf[0] = q[0]; hf[1] = q[1];
    int n = 0; uint64_t b = 0;
    for (auto &t : g.tensors) {
        int l, w; if (!slot_of(t.name, &l, &w)) continue;
        const size_t rs = eb_row_size(t.type, (int64_t)t.dims[0]);
        if (!rs) return eb_set_err(EB_ERR_UNSUPPORTED, "gguf: tensor %s has type %d", t.name.c_str(), t.type);
        const uint64_t nb = rs * t.dims[1] * t.dims[2] * t.dims[3];
        if (g.data_off + t.offset + nb > g.file_size) return eb_set_err(EB_ERR_ARG, "gguf: tensor %s runs past the end of the file", t.name.c_str());
        n++; b += nb;
    }
    if (n_tensors_used) *n_tensors_used = n;
    if (bytes_used) *bytes_used = b;
    return EB_OK;
}

extern "C" int eb_model_load_gguf(const char *path, eb_model **out, double *seconds, uint64_t *bytes) {
    if (!out) return eb_set_err(EB_ERR_ARG, "load_gguf: out == NULL");
    *out = nullptr;
    int dev_n = 0;
    if (cudaGetDeviceCount(&dev_n) != cudaSuccess || dev_n == 0) { cudaGetLastError(); return eb_set_err(EB_ERR_CUDA, "no CUDA device: this library has no CPU fallback"); }
    const auto t0 = std::chrono::steady_clock::now();
    Mapped f; int rc = f.open_file(path); if (rc) return rc;
    GFile g; if ((rc = parse(f.base, f.size, g))) return rc;
    int hp[6]; float hf[2]; if ((rc = hparams_of(g, hp, hf))) return rc;
    eb_model *m = eb_model_create(hp[0], hp[1], hp[2], hp[3], hp[4], hf[0], hf[1], hp[5]);
    if (!m) return EB_ERR_ARG;
    // tensors in file order
    std::vector<const GTensorInfo *> order;
    for (auto &t : g.tensors) { int l, w; if (slot_of(t.name, &l, &w)) order.push_back(&t); }
    std::sort(order.begin(), order.end(), [](const GTensorInfo *a, const GTensorInfo *b) { return a->offset < b->offset; });
    constexpr size_t CH = 32u << 20;
    uint8_t *pin[2] = {nullptr, nullptr}; void *dstage[2] = {nullptr, nullptr}; cudaEvent_t ev_pin[2] = {}, ev_stage[2] = {};
    cudaStream_t st = nullptr;
    size_t max_planar = 0; uint64_t total = 0;
    for (auto *t : order) {
        const size_t nb = eb_row_size(t->type, (int64_t)t->dims[0]) * t->dims[1];
        if (t->type == EB_Q4_0 || t->type == EB_Q8_0 || t->type == EB_Q6_K) max_planar = std::max(max_planar, nb);
    }
    auto cleanup = [&](int code) {
        if (st) cudaStreamSynchronize(st);
        for (int i = 0; i < 2; i++) { if (pin[i]) cudaFreeHost(pin[i]); if (dstage[i]) cudaFree(dstage[i]); if (ev_pin[i]) cudaEventDestroy(ev_pin[i]); if (ev_stage[i]) cudaEventDestroy(ev_stage[i]); }
        if (st) cudaStreamDestroy(st);
        if (code) { eb_model_free(m); }
        return code;
    };
#define LG(call) do { if ((call) != cudaSuccess) return cleanup(eb_set_err(EB_ERR_CUDA, "load_gguf: %s: %s", #call, cudaGetErrorString(cudaGetLastError()))); } while (0)
    LG(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) {
        LG(cudaMallocHost((void **)&pin[i], CH)); LG(cudaEventCreateWithFlags(&ev_pin[i], cudaEventDisableTiming)); LG(cudaEventCreateWithFlags(&ev_stage[i], cudaEventDisableTiming));
        if (max_planar) LG(cudaMalloc(&dstage[i], max_planar + 256));
    }
    int pk = 0, sk = 0; bool pin_used[2] = {false, false}, stage_used[2] = {false, false};
    for (auto *t : order) {
        int layer, which; slot_of(t->name, &layer, &which);
        if (layer >= hp[4]) return cleanup(eb_set_err(EB_ERR_ARG, "gguf: tensor %s beyond block_count %d", t->name.c_str(), hp[4]));
        const size_t nb = eb_row_size(t->type, (int64_t)t->dims[0]) * t->dims[1];
        if (!nb || t->dims[2] != 1 || t->dims[3] != 1) return cleanup(eb_set_err(EB_ERR_UNSUPPORTED, "gguf: tensor %s: type %d, %d dims", t->name.c_str(), t->type, t->n_dims));
        const uint64_t foff = g.data_off + t->offset;
        if (foff + nb > g.file_size) return cleanup(eb_set_err(EB_ERR_ARG, "gguf: tensor %s runs past the end of the file", t->name.c_str()));
        void *final_ptr = nullptr; size_t final_bytes = 0;
        if ((rc = eb_model_alloc_tensor(m, layer, which, t->type, &final_ptr, &final_bytes))) return cleanup(rc);
        if (final_bytes != nb) return cleanup(eb_set_err(EB_ERR_ARG, "gguf: tensor %s is %llu x %llu, the model wants %zu bytes", t->name.c_str(),
                                                         (unsigned long long)t->dims[0], (unsigned long long)t->dims[1], final_bytes));
        const bool planar = t->type == EB_Q4_0 || t->type == EB_Q8_0 || t->type == EB_Q6_K;
        uint8_t *dst = (uint8_t *)final_ptr;
        if (planar) {
            if (stage_used[sk]) LG(cudaEventSynchronize(ev_stage[sk]));        // the repack that last read this staging area has finished
            dst = (uint8_t *)dstage[sk];
        }
        for (size_t off = 0; off < nb; off += CH) {
            const size_t n = std::min(CH, nb - off);
            if (pin_used[pk]) LG(cudaEventSynchronize(ev_pin[pk]));            // the copy out of this pinned buffer has finished
            if (!read_parallel(f.fd, pin[pk], n, foff + off)) return cleanup(eb_set_err(EB_ERR_ARG, "gguf: read error in tensor %s", t->name.c_str()));
            LG(cudaMemcpyAsync(dst + off, pin[pk], n, cudaMemcpyHostToDevice, st));
            LG(cudaEventRecord(ev_pin[pk], st)); pin_used[pk] = true; pk ^= 1;
        }
        if (planar) {
            const int be = t->type == EB_Q6_K ? 256 : 32;
            if ((rc = eb_repack_impl(t->type, dstage[sk], final_ptr, (int64_t)(t->dims[0] * t->dims[1] / be), 0, st))) return cleanup(rc);
            LG(cudaEventRecord(ev_stage[sk], st)); stage_used[sk] = true; sk ^= 1;
        }
        total += nb;
    }
    LG(cudaStreamSynchronize(st));
#undef LG
    cleanup(0);
    *out = m;
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (bytes) *bytes = total;
    return EB_OK;
}
