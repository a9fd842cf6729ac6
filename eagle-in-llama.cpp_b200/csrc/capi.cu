// capi.cu — the operator-layer C-ABI (include/eagle_b200.h) over the kernels, plus device
// bring-up, error reporting and launch accounting.
#include "engine.cuh"
#include <atomic>
#include <cstdarg>
#include <vector>

thread_local std::string g_eb_err;
static std::atomic<long> g_launches{0};
static int g_sm_count = 0;
static int g_device = -1;

int eb_set_err(int code, const char *fmt, ...) {
    char buf[1024]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    g_eb_err = buf;
    return code;
}
void eb_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
long eb_launches_now() { return g_launches.load(); }
void eb_launches_add(long n) { g_launches.fetch_add(n); }
int eb_sm_count() { return g_sm_count > 0 ? g_sm_count : 148; }

static thread_local const int *g_gate = nullptr;
const int *eb_gate() { return g_gate; }
void eb_gate_set(const int *g) { g_gate = g; }

static thread_local int g_pdl_scope = EB_PDL_ALL;
void eb_pdl_scope(int mask) { g_pdl_scope = mask; }
int eb_pdl_enabled(int cls) { static const int on = getenv("EB_NO_PDL") ? 0 : 1; return on && (g_pdl_scope & cls) != 0; }

extern "C" {

const char *eb_last_error(void) { return g_eb_err.c_str(); }
int eb_version(void) { return 100; }
long eb_launch_count(int reset) { return reset ? g_launches.exchange(0) : g_launches.load(); }

int eb_device_init(int device) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return eb_set_err(EB_ERR_CUDA, "no CUDA device: this library has no CPU fallback"); }
    if (device < 0 || device >= n) return eb_set_err(EB_ERR_ARG, "device %d of %d", device, n);
    cudaDeviceProp prop;
    EB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return eb_set_err(EB_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    EB_CUDA(cudaSetDevice(device));
    g_sm_count = prop.multiProcessorCount; g_device = device;
    int rc;
    if ((rc = eb_mmq_init()) || (rc = eb_mmq32_init()) || (rc = eb_mmq_small_init()) || (rc = eb_mmq_tc_init()) || (rc = eb_fattn_tc_init()) || (rc = eb_fattn_init()) || (rc = eb_spec_init()) || (rc = eb_fused_init())) return rc;
    return EB_OK;
}
int eb_device_sm_count(void) { return g_sm_count; }
int eb_device_memory(size_t *free_bytes, size_t *total_bytes) {
    if (g_device < 0) { int rc = eb_device_init(0); if (rc) return rc; }
    size_t f = 0, t = 0; EB_CUDA(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f; if (total_bytes) *total_bytes = t;
    return EB_OK;
}

size_t eb_row_size(int type, int64_t k) {
    switch (type) {
        case EB_F32: return (size_t)k * 4;
        case EB_F16: return (size_t)k * 2;
        case EB_Q4_0: return k % 32 ? 0 : (size_t)(k / 32) * 18;
        case EB_Q8_0: return k % 32 ? 0 : (size_t)(k / 32) * 34;
        case EB_Q4_K: return k % 256 ? 0 : (size_t)(k / 256) * 144;
        case EB_Q6_K: return k % 256 ? 0 : (size_t)(k / 256) * 210;
    }
    return 0;
}

#define NEED_DEV() do { if (g_device < 0) { int rc_ = eb_device_init(0); if (rc_) return rc_; } } while (0)
static inline cudaStream_t S(void *s) { return (cudaStream_t)s; }

int eb_malloc(void **p, size_t bytes) { NEED_DEV(); if (cudaMalloc(p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return eb_set_err(EB_ERR_OOM, "cudaMalloc(%zu)", bytes); } return EB_OK; }
int eb_free(void *p) { EB_CUDA(cudaFree(p)); return EB_OK; }
int eb_memset(void *p, int v, size_t bytes, void *st) { EB_CUDA(cudaMemsetAsync(p, v, bytes, S(st))); return EB_OK; }
int eb_upload(void *d, const void *h, size_t bytes, void *st) { NEED_DEV(); EB_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, S(st))); EB_CUDA(cudaStreamSynchronize(S(st))); return EB_OK; }
int eb_download(void *h, const void *d, size_t bytes, void *st) { EB_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, S(st))); EB_CUDA(cudaStreamSynchronize(S(st))); return EB_OK; }
int eb_copy(void *d, const void *s, size_t bytes, void *st) { if (bytes && d != s) EB_CUDA(cudaMemcpyAsync(d, s, bytes, cudaMemcpyDeviceToDevice, S(st))); return EB_OK; }
int eb_synchronize(void *st) { EB_CUDA(cudaStreamSynchronize(S(st))); return EB_OK; }

// GGUF block layout <-> device layout (aligned planes), see common.cuh WPlanes
int eb_repack(int type, const void *d_gguf, void *d_dev, int64_t n_elems, void *st) {
    NEED_DEV();
    if (type == EB_Q4_0 || type == EB_Q8_0 || type == EB_Q6_K) return eb_repack_impl(type, const_cast<void *>(d_gguf), d_dev, n_elems / (type == EB_Q6_K ? 256 : 32), 0, S(st));
    EB_CUDA(cudaMemcpyAsync(d_dev, d_gguf, eb_row_size(type, n_elems), cudaMemcpyDeviceToDevice, S(st)));
    return EB_OK;
}
int eb_unpack(int type, const void *d_dev, void *d_gguf, int64_t n_elems, void *st) {
    NEED_DEV();
    if (type == EB_Q4_0 || type == EB_Q8_0 || type == EB_Q6_K) return eb_repack_impl(type, d_gguf, const_cast<void *>(d_dev), n_elems / (type == EB_Q6_K ? 256 : 32), 1, S(st));
    EB_CUDA(cudaMemcpyAsync(d_gguf, d_dev, eb_row_size(type, n_elems), cudaMemcpyDeviceToDevice, S(st)));
    return EB_OK;
}

int eb_dequantize(int type, const void *d_src, float *d_dst, int64_t n_elems, void *st) {
    NEED_DEV();
    const int be = (type == EB_Q4_K || type == EB_Q6_K) ? 256 : (type == EB_Q4_0 || type == EB_Q8_0) ? 32 : 1;
    return eb_dequantize_impl(type, d_src, n_elems / be, 0, d_dst, n_elems, S(st));
}

size_t eb_mul_mat_ws_size(int type, int64_t N, int64_t K, int64_t T) {
    (void)type;
    return 2 * eb_actq_bytes(1, K, (int)T) + (size_t)8 * T * N * 4 + 1024;
}
int eb_mul_mat(int type, const void *d_W, int64_t N, int64_t K, const float *d_x, int64_t T, float *d_y, void *d_ws, void *st) {
    return eb_mul_mat_ex(type, d_W, N, K, d_x, T, d_y, d_ws, 0, st);
}
// reuse_activation != 0: the previous eb_mul_mat / eb_mul_mat_ex call on this workspace quantised the SAME d_x (same K, T <= 128, same
// activation grid: K-quants share Q8_K, Q4_0 / Q8_0 share Q8_0) and nothing else touched the workspace since — the planes are reused
int eb_mul_mat_ex(int type, const void *d_W, int64_t N, int64_t K, const float *d_x, int64_t T, float *d_y, void *d_ws, int reuse_activation, void *st) {
    NEED_DEV();
    if (T < 1) return EB_OK;
    if (!d_ws) return eb_set_err(EB_ERR_ARG, "mul_mat: workspace required");
    if (reuse_activation && T > 128) return eb_set_err(EB_ERR_ARG, "mul_mat_ex: activation reuse needs T <= 128");
    const size_t aq = 2 * eb_actq_bytes(1, K, (int)T);
    // batches beyond one verify tile are processed in column chunks of 128
    for (int64_t t0 = 0; t0 < T; t0 += 128) {
        const int tc = (int)std::min<int64_t>(128, T - t0);
        int rc = eb_mul_mat_impl(type, d_W, N, K, d_x + t0 * K, tc, d_y + t0 * N, d_ws, (float *)((char *)d_ws + ((aq + 255) & ~(size_t)255)), S(st), reuse_activation != 0);
        if (rc) return rc;
    }
    return EB_OK;
}
// activation quantiser exposed for the parity tests: planes as documented in common.cuh
int eb_quantize_act_planes(int kq, const float *d_x, int64_t K, int T, int8_t *d_qs, float *d_d, void *d_bsum_f16, void *st) {
    NEED_DEV();
    ActQ a; a.qs = d_qs; a.d = d_d; a.bsum = (__half *)d_bsum_f16; a.K = K; a.T = T; a.kq = kq;
    return eb_quantize_act(d_x, a, S(st));
}

int eb_rms_norm(const float *x, const float *w, float *y, int64_t n, int64_t rows, float eps, void *st) { NEED_DEV(); return eb_rms_norm_impl(x, w, y, n, rows, eps, S(st)); }
int eb_rope_norm(float *x, const int32_t *pos, int D, int n_head, int T, float freq_base, void *st) { NEED_DEV(); return eb_rope_norm_impl(x, x, pos, D, n_head, T, freq_base, S(st)); }
int eb_rope_norm_to(const float *src, float *dst, const int32_t *pos, int D, int n_head, int T, float freq_base, void *st) { NEED_DEV(); return eb_rope_norm_impl(src, dst, pos, D, n_head, T, freq_base, S(st)); }
int eb_get_rows_n(int type, const void *W, int64_t n_rows, int64_t K, const int32_t *ids, int n, float *out, void *st) { NEED_DEV(); return eb_get_rows_dev(type, W, n_rows, K, ids, n, out, S(st)); }
int eb_add(const float *a, const float *b, float *y, int64_t n, int64_t nb, void *st) { NEED_DEV(); return eb_add_impl(a, b, y, n, nb, S(st)); }
int eb_mul(const float *a, const float *b, float *y, int64_t n, int64_t nb, void *st) { NEED_DEV(); return eb_mul_impl(a, b, y, n, nb, S(st)); }
int eb_scale(const float *x, float *y, int64_t n, float s, void *st) { NEED_DEV(); return eb_scale_impl(x, y, n, s, S(st)); }
int eb_silu(const float *x, float *y, int64_t n, void *st) { NEED_DEV(); return eb_silu_impl(x, y, n, S(st)); }
int eb_relu(const float *x, float *y, int64_t n, void *st) { NEED_DEV(); return eb_relu_impl(x, y, n, S(st)); }
int eb_copy_strided(int src_type, int dst_type, const void *src, void *dst, const int64_t *ne, const int64_t *nb_src, const int64_t *dne, const int64_t *nb_dst, void *st) {
    NEED_DEV(); return eb_copy_strided_impl(src_type, dst_type, src, dst, ne, nb_src, dne, nb_dst, S(st));
}
int eb_mul_mat_dense(int a_type, const void *a, int64_t a_nb1, int64_t a_nb2, int64_t a_batches, const float *x, int64_t x_nb1, int64_t x_nb2,
                     float *y, int64_t N, int64_t K, int64_t T, int64_t B, void *st) {
    NEED_DEV(); return eb_mul_mat_dense_impl(a_type, a, a_nb1, a_nb2, a_batches, x, x_nb1, x_nb2, y, N, K, T, B, S(st));
}
int eb_soft_max_ext(const float *x, const void *mask, int mask_type, int64_t mask_stride, float *y, int64_t n, int64_t rows, int64_t rows_per_batch, float scale, void *st) {
    NEED_DEV(); return eb_soft_max_ext_impl(x, mask, mask_type, mask_stride, y, n, rows, rows_per_batch, scale, S(st));
}
int eb_concat2(const float *a, const float *b, float *y, int E, int T, void *st) { NEED_DEV(); return eb_concat2_impl(a, b, y, E, T, S(st)); }
int eb_cpy_f32_f16(const float *x, void *y, int64_t n, void *st) { NEED_DEV(); return eb_cpy_f32_f16_impl(x, y, n, S(st)); }

size_t eb_flash_attn_ws_size(int T, int H, int D, int n_kv) { T = std::min(T, 128); return std::max(eb_flash_attn_ws_size_impl(T, H, D, n_kv), eb_fattn_tc_ws_size(T, H, n_kv)); }
int eb_flash_attn(const float *q, const void *k, const void *v, int64_t kv_stride, const void *mask, int64_t mask_stride,
                  const uint32_t *tree_bits, int n_prefix, int T, int H, int D, int n_kv, float scale, float *out, void *ws, void *st) {
    NEED_DEV();
    if (!mask && tree_bits && eb_fattn_tc_supported(T, D) && kv_stride % 8 == 0)      // compact tree mask: tcgen05 kernel
        return eb_fattn_tc_f32q(q, k, v, kv_stride, tree_bits, n_prefix, n_kv, T, H, scale, out, ws, S(st));
    if (mask && T > 128) {
        // prompt-sized batches (llama_decode_impl_init, R/src/llama.cpp:10512: T = 512): query rows in chunks of one verify tile,
        // each with its own mask rows; the chunks run back to back on the stream and share the workspace
        for (int t0 = 0; t0 < T; t0 += 128) {
            const int tc = std::min(128, T - t0);
            int rc = eb_flash_attn_impl(q + (size_t)t0 * H * D, k, v, kv_stride, (const __half *)mask + (size_t)t0 * mask_stride, mask_stride, nullptr, 0,
                                        tc, H, D, n_kv, scale, out + (size_t)t0 * H * D, ws, S(st));
            if (rc) return rc;
        }
        return EB_OK;
    }
    return eb_flash_attn_impl(q, k, v, kv_stride, mask, mask_stride, tree_bits, n_prefix, T, H, D, n_kv, scale, out, ws, S(st));
}
// the ggml mask in bit form (any mask whose values are 0 or -inf): lets the tcgen05 kernel serve GGML_OP_FLASH_ATTN_EXT
size_t eb_mask_bits_size(int T, int n_kv) { return eb_mask_bits_size_impl(T, n_kv); }
int eb_mask_to_bits(const void *mask, int64_t mask_stride, int T, int n_kv, void *bits, int *h_general, void *st) {
    NEED_DEV();
    if (T < 1 || n_kv < 1 || !mask || !bits) return eb_set_err(EB_ERR_ARG, "mask_to_bits: T %d n_kv %d", T, n_kv);
    int rc = eb_mask_to_bits_impl(mask, mask_stride, T, n_kv, bits, S(st)); if (rc) return rc;
    if (h_general) {                                    // the caller decides on the host which kernel runs: one 4-byte read per mask
        EB_CUDA(cudaMemcpyAsync(h_general, eb_mask_bits_flag(bits, T, n_kv), 4, cudaMemcpyDeviceToHost, S(st)));
        EB_CUDA(cudaStreamSynchronize(S(st)));
    }
    return EB_OK;
}
int eb_flash_attn_bits_supported(int T, int H, int D, int64_t kv_stride) { (void)H; return T >= 1 && eb_fattn_tc_supported(std::min(T, 128), D) && kv_stride % 8 == 0; }
int eb_flash_attn_bits(const float *q, const void *k, const void *v, int64_t kv_stride, const void *bits, int T, int H, int D, int n_kv, float scale, float *out,
                       void *ws, void *st) {
    NEED_DEV();
    if (!eb_flash_attn_bits_supported(T, H, D, kv_stride)) return eb_set_err(EB_ERR_UNSUPPORTED, "flash_attn_bits: T %d D %d kv_stride %lld", T, D, (long long)kv_stride);
    return eb_fattn_tc_bits(q, k, v, kv_stride, bits, T, H, n_kv, scale, out, ws, S(st));
}
// streams and events for hosts that own several execution contexts (one ggml_backend_t = one stream)
// pinned host memory (ggml_backend_cuda_host_buffer_type, ggml-cuda.cu:1048-1108): inputs and the logits buffer of a host program
int eb_malloc_host(void **p, size_t bytes) { NEED_DEV(); if (cudaMallocHost(p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); *p = nullptr; return eb_set_err(EB_ERR_OOM, "cudaMallocHost(%zu)", bytes); } return EB_OK; }
int eb_free_host(void *p) { if (p) EB_CUDA(cudaFreeHost(p)); return EB_OK; }
int eb_stream_create(void **st) { NEED_DEV(); cudaStream_t s; EB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking)); *st = (void *)s; return EB_OK; }
int eb_stream_destroy(void *st) { if (st) EB_CUDA(cudaStreamDestroy(S(st))); return EB_OK; }
int eb_device_synchronize(void) { NEED_DEV(); EB_CUDA(cudaDeviceSynchronize()); return EB_OK; }
int eb_upload_async(void *d, const void *h, size_t bytes, void *st) { NEED_DEV(); EB_CUDA(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, S(st))); return EB_OK; }
int eb_download_async(void *h, const void *d, size_t bytes, void *st) { NEED_DEV(); EB_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, S(st))); return EB_OK; }
int eb_event_create(void **ev) { NEED_DEV(); cudaEvent_t e; EB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); *ev = (void *)e; return EB_OK; }
int eb_event_destroy(void *ev) { if (ev) EB_CUDA(cudaEventDestroy((cudaEvent_t)ev)); return EB_OK; }
int eb_event_record(void *ev, void *st) { EB_CUDA(cudaEventRecord((cudaEvent_t)ev, S(st))); return EB_OK; }
int eb_stream_wait_event(void *st, void *ev) { EB_CUDA(cudaStreamWaitEvent(S(st), (cudaEvent_t)ev, 0)); return EB_OK; }
int eb_event_synchronize(void *ev) { EB_CUDA(cudaEventSynchronize((cudaEvent_t)ev)); return EB_OK; }

int eb_top_k(const float *logits, int rows, int V, int k, int32_t *ids, float *vals, float *probs, void *st) { NEED_DEV(); return eb_top_k_impl(logits, rows, V, k, ids, vals, probs, S(st), 1, 0, rows <= 128 ? eb_top_k_default_scratch(S(st)) : nullptr); }
int eb_argsort(const float *x, int rows, int n, int descending, int32_t *ids, void *st) { NEED_DEV(); return eb_argsort_impl(x, rows, n, descending, ids, S(st)); }
int eb_argmax(const float *logits, int rows, int V, int32_t *ids, void *st) { NEED_DEV(); return eb_argmax_impl(logits, rows, V, ids, S(st)); }


// Micro-bench of ONE hot kernel launch class, timed with CUDA events on the launching stream:
// the GEMM (T > 8) or matvec (T <= 8) over `n_bufs` distinct weight buffers in rotation (so that
// weights always come from HBM, not L2).  Activations are pre-quantised outside the timed region.
// Returns the average milliseconds per launch in *ms_per_launch.
int eb_bench_mul_mat_kernel(int type, int64_t N, int64_t K, int T, int n_bufs, int iters, float *ms_per_launch) {
    NEED_DEV();
    if (n_bufs < 1 || iters < 1 || T < 1 || T > 128) return eb_set_err(EB_ERR_ARG, "bench: bad arguments");
    const size_t wbytes = eb_row_size(type, K) * N;
    if (!wbytes) return eb_set_err(EB_ERR_ARG, "bench: type/K");
    std::vector<void *> w(n_bufs, nullptr);
    float *x = nullptr, *y = nullptr, *partial = nullptr; void *aq = nullptr;
    cudaStream_t st; cudaEvent_t e0, e1;
    EB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); EB_CUDA(cudaEventCreate(&e0)); EB_CUDA(cudaEventCreate(&e1));
    int rc = EB_OK;
    const int kq = (type == EB_Q4_K || type == EB_Q6_K);
    for (int i = 0; i < n_bufs && !rc; i++) {
        if (cudaMalloc(&w[i], wbytes + 256) != cudaSuccess) { rc = eb_set_err(EB_ERR_OOM, "bench: weights"); break; }
        void *tmp = w[i];
        if (type == EB_Q6_K || type == EB_Q4_0 || type == EB_Q8_0) {     // synth writes GGUF order; planes need a repack
            if (cudaMalloc(&tmp, wbytes) != cudaSuccess) { rc = eb_set_err(EB_ERR_OOM, "bench: tmp"); break; }
        }
        rc = eb_synth_impl(type, tmp, N * K, 1234 + i, 0.02f, 0, st);
        if (!rc && tmp != w[i]) { rc = eb_repack_impl(type, tmp, w[i], N * K / (kq ? 256 : 32), 0, st); cudaStreamSynchronize(st); cudaFree(tmp); }
    }
    if (!rc && (cudaMalloc((void **)&x, (size_t)T * K * 4) || cudaMalloc((void **)&y, (size_t)T * N * 4) ||
                cudaMalloc((void **)&partial, (size_t)8 * T * N * 4) || cudaMalloc(&aq, eb_actq_bytes(kq, K, T) + 256)))
        rc = eb_set_err(EB_ERR_OOM, "bench: activations");
    if (!rc) rc = eb_synth_impl(EB_F32, x, (int64_t)T * K, 99, 1.0f, 0, st);
    ActQ a = eb_actq_carve(aq, kq, K, T);
    if (!rc) rc = eb_quantize_act(x, a, st);
    const bool use_mmq = eb_mmq_supported(type, N, K, T);
    const int ks = use_mmq ? eb_mmq_pick_ksplit(type, N, K, T) : 1;
    const bool use_tc = eb_mmq_tc_supported(type, N, K, T);
    const bool use_small = T > 1 && eb_mmq_small_supported(type, N, K, T);
    auto one = [&](int i) { int k2; return use_small ? eb_mmq_small_parts(type, w[i % n_bufs], N, K, a, partial, &k2, st) : use_tc ? eb_mmq_tc_parts(type, w[i % n_bufs], N, K, a, partial, &k2, st) : use_mmq ? eb_mmq(type, w[i % n_bufs], N, K, a, y, partial, ks, st) : eb_mmq32_supported(type, N, K, T) ? eb_mmq32_parts(type, w[i % n_bufs], N, K, a, partial, &k2, st) : eb_mmvq(type, w[i % n_bufs], N, K, a, y, st); };
    for (int i = 0; i < 3 && !rc; i++) rc = one(i);
    if (!rc) {
        // the timed launches are replayed from a CUDA graph: host-side work per launch (tensor-map
        // encoding, argument marshalling) stays outside the measurement, as in the engine's graph
        cudaStreamSynchronize(st);
        cudaGraph_t g = nullptr; cudaGraphExec_t ge = nullptr;
        const long before = eb_launches_now();
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench: capture");
        for (int i = 0; i < iters && !rc; i++) rc = one(i);
        if (cudaStreamEndCapture(st, &g) != cudaSuccess && !rc) rc = eb_set_err(EB_ERR_CUDA, "bench: end capture");
        (void)before;
        if (!rc && cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench: instantiate");
        if (!rc) {
            cudaGraphLaunch(ge, st);                       // warm replay
            cudaStreamSynchronize(st);
            cudaEventRecord(e0, st);
            cudaGraphLaunch(ge, st);
            cudaEventRecord(e1, st);
            if (cudaStreamSynchronize(st) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench: %s", cudaGetErrorString(cudaGetLastError()));
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            if (ms_per_launch) *ms_per_launch = ms / iters;
        }
        if (ge) cudaGraphExecDestroy(ge);
        if (g) cudaGraphDestroy(g);
    }
    for (void *p : w) cudaFree(p);
    cudaFree(x); cudaFree(y); cudaFree(partial); cudaFree(aq);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return rc;
}

// Micro-bench of the FLASH_ATTN_EXT operator with the compact tree mask (BASELINE config 5): `iters` calls replayed from a CUDA
// graph, cycling through n_bufs distinct K/V caches (so K/V come from HBM), CUDA events on the launching stream.  The tree is a
// chain over the last T cells; the first n_kv - T cells are the committed prefix.
int eb_bench_flash_attn(int T, int H, int D, int n_kv, int n_bufs, int iters, float *ms_per_call) {
    NEED_DEV();
    if (T < 1 || T > 128 || n_kv <= T || n_bufs < 1 || iters < 1 || D != 128) return eb_set_err(EB_ERR_ARG, "bench_flash_attn: bad arguments");
    const int64_t E = (int64_t)H * D;
    std::vector<void *> kc(n_bufs, nullptr), vc(n_bufs, nullptr);
    float *q = nullptr, *out = nullptr; void *ws = nullptr; uint32_t *bits = nullptr;
    cudaStream_t st; cudaEvent_t e0, e1;
    EB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)); EB_CUDA(cudaEventCreate(&e0)); EB_CUDA(cudaEventCreate(&e1));
    int rc = EB_OK;
    for (int i = 0; i < n_bufs && !rc; i++) {
        if (cudaMalloc(&kc[i], (size_t)n_kv * E * 2) != cudaSuccess || cudaMalloc(&vc[i], (size_t)n_kv * E * 2) != cudaSuccess) { rc = eb_set_err(EB_ERR_OOM, "bench_flash_attn: KV"); break; }
        rc = eb_synth_impl(EB_F16, kc[i], (int64_t)n_kv * E, 500 + 2 * i, 1.0f, 0, st);
        if (!rc) rc = eb_synth_impl(EB_F16, vc[i], (int64_t)n_kv * E, 501 + 2 * i, 1.0f, 0, st);
    }
    if (!rc && (cudaMalloc((void **)&q, (size_t)T * E * 4) || cudaMalloc((void **)&out, (size_t)T * E * 4) || cudaMalloc(&ws, eb_flash_attn_ws_size(T, H, D, n_kv)) ||
                cudaMalloc((void **)&bits, (size_t)T * 16)))
        rc = eb_set_err(EB_ERR_OOM, "bench_flash_attn: activations");
    if (!rc) rc = eb_synth_impl(EB_F32, q, (int64_t)T * E, 7, 1.0f, 0, st);
    if (!rc) {
        std::vector<uint32_t> hb((size_t)T * 4, 0u);
        for (int t = 0; t < T; t++) for (int j = 0; j <= t; j++) hb[t * 4 + (j >> 5)] |= 1u << (j & 31);
        if (cudaMemcpyAsync(bits, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench_flash_attn: bits");
    }
    auto one = [&](int i) { return eb_flash_attn(q, kc[i % n_bufs], vc[i % n_bufs], E, nullptr, 0, bits, n_kv - T, T, H, D, n_kv, 1.0f / sqrtf((float)D), out, ws, st); };
    for (int i = 0; i < 3 && !rc; i++) rc = one(i);
    if (!rc) {
        cudaStreamSynchronize(st);
        cudaGraph_t g = nullptr; cudaGraphExec_t ge = nullptr;
        const long before = eb_launches_now();
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench_flash_attn: capture");
        for (int i = 0; i < iters && !rc; i++) rc = one(i);
        if (cudaStreamEndCapture(st, &g) != cudaSuccess && !rc) rc = eb_set_err(EB_ERR_CUDA, "bench_flash_attn: end capture");
        eb_launches_add(before - eb_launches_now());
        if (!rc && cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench_flash_attn: instantiate");
        if (!rc) {
            cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
            cudaEventRecord(e0, st); cudaGraphLaunch(ge, st); cudaEventRecord(e1, st);
            if (cudaStreamSynchronize(st) != cudaSuccess) rc = eb_set_err(EB_ERR_CUDA, "bench_flash_attn: %s", cudaGetErrorString(cudaGetLastError()));
            float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
            if (ms_per_call) *ms_per_call = ms / iters;
        }
        if (ge) cudaGraphExecDestroy(ge);
        if (g) cudaGraphDestroy(g);
    }
    for (void *p : kc) cudaFree(p);
    for (void *p : vc) cudaFree(p);
    cudaFree(q); cudaFree(out); cudaFree(ws); cudaFree(bits);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return rc;
}

}  // extern "C"
