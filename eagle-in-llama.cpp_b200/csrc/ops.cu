// ops.cu — glue operators of the path (K4-K9): RMS-norm (+weight, + fused activation
// quantisation), RoPE (+ fused f32->f16 KV store), get_rows, add / mul / silu / relu, SwiGLU.
// All are HBM/L2-bound element-wise work: coalesced 128-bit accesses, one pass over the data.
#include "common.cuh"
#include "spec.cuh"

// ------------------------------------------------------------------ RMS norm
// ggml_compute_forward_rms_norm_f32 (R/ggml/src/ggml-cpu/ggml-cpu.c:7098): sum of f32 squares
// accumulated in double, mean -> f32, scale = 1/sqrtf(mean + eps); optional weight (GGML_OP_MUL).
// One CTA (256 threads) per row.  If `aq.qs` is set the normalised row is also quantised on the
// activation grid (fusion of RMS_NORM -> MUL -> from_float of the following MUL_MATs).
__global__ void __launch_bounds__(256) rms_norm_kernel(const float *__restrict__ x, const float *__restrict__ w, float *__restrict__ y,
                                                      int n, float eps) {
    extern __shared__ float row[];
    __shared__ double red[8];
    __shared__ float s_scale;
    const int r = blockIdx.x, tid = threadIdx.x;
    const float *xr = x + (int64_t)r * n;
    double acc = 0.0;
    for (int i = tid * 4; i < n; i += 1024) {
        const float4 v = *(const float4 *)(xr + i);
        *(float4 *)(row + i) = v;
        acc += (double)(v.x * v.x); acc += (double)(v.y * v.y); acc += (double)(v.z * v.z); acc += (double)(v.w * v.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((tid & 31) == 0) red[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        double s = 0; for (int i = 0; i < 8; i++) s += red[i];
        const float mean = (float)(s / n);
        s_scale = 1.0f / sqrtf(mean + eps);
    }
    __syncthreads();
    const float sc = s_scale;
    float *yr = y + (int64_t)r * n;
    for (int i = tid * 4; i < n; i += 1024) {
        float4 v = *(float4 *)(row + i);
        v.x = __fmul_rn(v.x, sc); v.y = __fmul_rn(v.y, sc); v.z = __fmul_rn(v.z, sc); v.w = __fmul_rn(v.w, sc);
        if (w) { const float4 g = *(const float4 *)(w + i); v.x = __fmul_rn(v.x, g.x); v.y = __fmul_rn(v.y, g.y); v.z = __fmul_rn(v.z, g.z); v.w = __fmul_rn(v.w, g.w); }
        *(float4 *)(yr + i) = v;
    }
}
int eb_rms_norm_impl(const float *x, const float *w, float *y, int64_t n, int64_t rows, float eps, cudaStream_t st) {
    if (n % 4 || n > 16384) return eb_set_err(EB_ERR_ARG, "rms_norm: n=%lld", (long long)n);
    if (rows == 0) return EB_OK;
    rms_norm_kernel<<<(unsigned)rows, 256, (size_t)n * 4, st>>>(x, w, y, (int)n, eps);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ RoPE (NORM) [+ KV store]
// ggml_compute_forward_rope_f32 (R/ggml/src/ggml-cpu/ggml-cpu.c:9449) with ggml_rope_cache_init
// (:9376): theta_0 = pos, theta_{i+1} = theta_i * theta_scale (sequential f32 products, restated
// exactly), pairs (2i, 2i+1).  One CTA per token: the first D/2 threads build the token's cos/sin
// cache in shared memory (like the reference's per-row cache), then all threads rotate the
// n_head * D/2 pairs of the row.  src may equal dst (in place) or not (ggml's ROPE node is not
// always in place).
__global__ void __launch_bounds__(256) rope_norm_kernel(const float *__restrict__ src, float *__restrict__ dst, const int32_t *__restrict__ pos, int D, int n_head, float theta_scale) {
    __shared__ float2 cs[128];
    const int t = blockIdx.x, half = D / 2;
    for (int i = threadIdx.x; i < half; i += blockDim.x) {
        float theta = (float)pos[t];
        for (int j = 0; j < i; j++) theta = __fmul_rn(theta, theta_scale);
        cs[i] = make_float2(cosf(theta), sinf(theta));
    }
    __syncthreads();
    const int64_t row = (int64_t)t * n_head * D;
    for (int idx = threadIdx.x; idx < n_head * half; idx += blockDim.x) {
        const int h = idx / half, i = idx - h * half;
        const float2 k = cs[i];
        const int64_t o = row + (int64_t)h * D + 2 * i;
        const float2 v = *(const float2 *)(src + o);
        *(float2 *)(dst + o) = make_float2(__fsub_rn(__fmul_rn(v.x, k.x), __fmul_rn(v.y, k.y)), __fadd_rn(__fmul_rn(v.x, k.y), __fmul_rn(v.y, k.x)));
    }
}
int eb_rope_norm_impl(const float *src, float *dst, const int32_t *pos, int D, int n_head, int T, float freq_base, cudaStream_t st) {
    if (T == 0) return EB_OK;
    if (D > 256 || (D & 1)) return eb_set_err(EB_ERR_ARG, "rope: D=%d", D);
    const float theta_scale = powf(freq_base, -2.0f / D);
    rope_norm_kernel<<<T, 256, 0, st>>>(src, dst, pos, D, n_head, theta_scale);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// Fused engine form: q,k,v f32 [T][E] -> rotated q (f32, in place), rotated k -> f16 cache rows,
// v -> f16 cache rows (llm_build_kv_store R/src/llama.cpp:228-270 + ggml_cpy f32->f16).
// The reference rounds the rotated K to f16 in the cache and keeps Q in f32 until flash-attn
// converts it (ggml-cpu.c:11115) — the attention kernel does that conversion.
__global__ void rope_kv_store_kernel(float *__restrict__ q, const float *__restrict__ k, const float *__restrict__ v,
                                     const int32_t *__restrict__ pos, __half *__restrict__ kc, __half *__restrict__ vc,
                                     const FwdCtl *__restrict__ ctl, int D, int n_head, int T, float theta_scale) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (ctl) { const int64_t cell0 = (int64_t)(ctl->n_prefix + ctl->node_base) * n_head * D; kc += cell0; vc += cell0; }
    const int half = D / 2;
    if (idx >= T * half) return;
    const int t = idx / half, i = idx % half;
    float theta = (float)pos[t];
    for (int j = 0; j < i; j++) theta = __fmul_rn(theta, theta_scale);
    const float c = cosf(theta), s = sinf(theta);
    const int64_t base = (int64_t)t * n_head * D + 2 * i;
    for (int h = 0; h < n_head; h++) {
        const int64_t o = base + (int64_t)h * D;
        const float2 a = *(float2 *)(q + o);
        *(float2 *)(q + o) = make_float2(__fsub_rn(__fmul_rn(a.x, c), __fmul_rn(a.y, s)), __fadd_rn(__fmul_rn(a.x, s), __fmul_rn(a.y, c)));
        const float2 b = *(const float2 *)(k + o);
        *(__half2 *)(kc + o) = __floats2half2_rn(__fsub_rn(__fmul_rn(b.x, c), __fmul_rn(b.y, s)), __fadd_rn(__fmul_rn(b.x, s), __fmul_rn(b.y, c)));
        const float2 w = *(const float2 *)(v + o);
        *(__half2 *)(vc + o) = __floats2half2_rn(w.x, w.y);
    }
}
int eb_rope_kv_store_impl(float *q, const float *k, const float *v, const int32_t *pos, void *kc, void *vc,
                          int D, int n_head, int T, float freq_base, cudaStream_t st) {
    if (T == 0) return EB_OK;
    const float theta_scale = powf(freq_base, -2.0f / D);
    rope_kv_store_kernel<<<cdiv(T * (D / 2), 64), 64, 0, st>>>(q, k, v, pos, (__half *)kc, (__half *)vc, nullptr, D, n_head, T, theta_scale);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

int eb_rope_kv_store_ctl(float *q, const float *k, const float *v, const int32_t *pos, void *kc, void *vc, const FwdCtl *ctl,
                         int D, int n_head, int T, float freq_base, cudaStream_t st) {
    if (T == 0) return EB_OK;
    const float theta_scale = powf(freq_base, -2.0f / D);
    rope_kv_store_kernel<<<cdiv(T * (D / 2), 64), 64, 0, st>>>(q, k, v, pos, (__half *)kc, (__half *)vc, ctl, D, n_head, T, theta_scale);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ element-wise
__global__ void add_kernel(const float *a, const float *b, float *y, int64_t n, int64_t nb) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __fadd_rn(a[i], b[nb == n ? i : i % nb]);
}
__global__ void mul_kernel(const float *a, const float *b, float *y, int64_t n, int64_t nb) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __fmul_rn(a[i], b[nb == n ? i : i % nb]);
}
// ggml_silu_f32: x/(1+expf(-x))  (R/ggml/src/ggml-cpu/ggml-cpu.c:1838)
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__global__ void silu_kernel(const float *x, float *y, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = silu_f(x[i]);
}
__global__ void scale_kernel(const float *x, float *y, int64_t n, float s) {      // GGML_OP_SCALE: ggml_vec_scale_f32, R/ggml/src/ggml-cpu/ggml-cpu.c
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __fmul_rn(x[i], s);
}
__global__ void relu_kernel(const float *x, float *y, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = x[i] > 0.f ? x[i] : 0.f;
}
// y = silu(gate) * up   (llm_build_ffn LLM_FFN_SILU + LLM_FFN_PAR, R/src/llama.cpp:456-581)
__global__ void swiglu_kernel(const float *__restrict__ gate, const float *__restrict__ up, float *__restrict__ y, int64_t n) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const float4 g = *(const float4 *)(gate + i), u = *(const float4 *)(up + i);
    *(float4 *)(y + i) = make_float4(__fmul_rn(silu_f(g.x), u.x), __fmul_rn(silu_f(g.y), u.y), __fmul_rn(silu_f(g.z), u.z), __fmul_rn(silu_f(g.w), u.w));
}
// y = relu(x + bias)   (llm_build_fc, R/src/llama.cpp:367-420)
__global__ void bias_relu_kernel(const float *__restrict__ x, const float *__restrict__ b, float *__restrict__ y, int64_t n, int nb) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float r = b ? __fadd_rn(x[i], b[i % nb]) : x[i];
    y[i] = r > 0.f ? r : 0.f;
}
__global__ void cpy_f32_f16_kernel(const float *x, __half *y, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __float2half_rn(x[i]);
}
// [T][E] embedding rows and [T][E] hidden rows -> [T][2E] (ggml_concat dim 0, embedding first)
__global__ void concat2_kernel(const float *__restrict__ a, const float *__restrict__ b, float *__restrict__ y, int E, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t t = i / (2 * E); const int c = (int)(i % (2 * E));
    y[i] = c < E ? a[t * E + c] : b[t * E + c - E];
}

#define EW_LAUNCH(kernel, n, ...) do { if ((n) > 0) { kernel<<<(unsigned)cdiv((n), 256), 256, 0, st>>>(__VA_ARGS__); EB_CHECK_LAUNCH(); } return EB_OK; } while (0)
int eb_add_impl(const float *a, const float *b, float *y, int64_t n, int64_t nb, cudaStream_t st) { EW_LAUNCH(add_kernel, n, a, b, y, n, nb); }
int eb_mul_impl(const float *a, const float *b, float *y, int64_t n, int64_t nb, cudaStream_t st) { EW_LAUNCH(mul_kernel, n, a, b, y, n, nb); }
int eb_silu_impl(const float *x, float *y, int64_t n, cudaStream_t st) { EW_LAUNCH(silu_kernel, n, x, y, n); }
int eb_scale_impl(const float *x, float *y, int64_t n, float s, cudaStream_t st) { EW_LAUNCH(scale_kernel, n, x, y, n, s); }
int eb_relu_impl(const float *x, float *y, int64_t n, cudaStream_t st) { EW_LAUNCH(relu_kernel, n, x, y, n); }
int eb_bias_relu_impl(const float *x, const float *b, float *y, int64_t n, int nb, cudaStream_t st) { EW_LAUNCH(bias_relu_kernel, n, x, b, y, n, nb); }
int eb_cpy_f32_f16_impl(const float *x, void *y, int64_t n, cudaStream_t st) { EW_LAUNCH(cpy_f32_f16_kernel, n, x, (__half *)y, n); }
int eb_concat2_impl(const float *a, const float *b, float *y, int E, int T, cudaStream_t st) { const int64_t n = (int64_t)T * 2 * E; EW_LAUNCH(concat2_kernel, n, a, b, y, E, n); }
int eb_swiglu_impl(const float *gate, const float *up, float *y, int64_t n, cudaStream_t st) {
    if (n % 4) return eb_set_err(EB_ERR_ARG, "swiglu: n %% 4");
    if (n) { swiglu_kernel<<<(unsigned)cdiv(n / 4, 256), 256, 0, st>>>(gate, up, y, n); EB_CHECK_LAUNCH(); }
    return EB_OK;
}

// ------------------------------------------------------------------ get_rows (K7)
// ggml_compute_forward_get_rows_q (R/ggml/src/ggml-cpu/ggml-cpu.c:8547): dequantise row ids[i].
int eb_get_rows_impl(int type, const void *W, int64_t n_rows, int64_t K, const int32_t *h_ids_or_null, const int32_t *d_ids, int n, float *out, cudaStream_t st);

__global__ void get_rows_f32_kernel(const float *__restrict__ W, int64_t K, const int32_t *__restrict__ ids, float *__restrict__ out) {
    const int r = blockIdx.y; const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < K) out[(int64_t)r * K + i] = W[(int64_t)ids[r] * K + i];
}
__global__ void get_rows_f16_kernel(const __half *__restrict__ W, int64_t K, const int32_t *__restrict__ ids, float *__restrict__ out) {
    const int r = blockIdx.y; const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < K) out[(int64_t)r * K + i] = __half2float(W[(int64_t)ids[r] * K + i]);
}
// quantised: one CTA of 64 threads per (block, row); same arithmetic as the dequant kernels
__global__ void get_rows_q_kernel(int type, const uint8_t *__restrict__ W, int64_t nb_total, int nbr, const int32_t *__restrict__ ids, float *__restrict__ out,
                                  const int *__restrict__ gate) {
    if (gate_closed(gate)) return;
    const int r = blockIdx.y, b = blockIdx.x, tid = threadIdx.x;
    const int64_t blk = (int64_t)ids[r] * nbr + b;
    WPlanes w{W, nb_total};
    if (type == EB_Q4_K) {
        const block_q4_K *x = (const block_q4_K *)W + blk;
        float *y = out + ((int64_t)r * nbr + b) * 256;
        const int j = tid / 16, l0 = (tid % 16) * 2;
        const float d = __half2float(x->d), mn = __half2float(x->dmin);
        int sc1, m1, sc2, m2;
        q4k_scale_min(2 * j, x->scales, sc1, m1); q4k_scale_min(2 * j + 1, x->scales, sc2, m2);
        const float d1 = __fmul_rn(d, (float)sc1), mm1 = __fmul_rn(mn, (float)m1), d2 = __fmul_rn(d, (float)sc2), mm2 = __fmul_rn(mn, (float)m2);
        for (int l = l0; l < l0 + 2; l++) {
            const int q = x->qs[j * 32 + l];
            y[j * 64 + l] = __fsub_rn(__fmul_rn(d1, (float)(q & 0xF)), mm1);
            y[j * 64 + l + 32] = __fsub_rn(__fmul_rn(d2, (float)(q >> 4)), mm2);
        }
    } else if (type == EB_Q6_K) {
        float *y = out + ((int64_t)r * nbr + b) * 256;
        const int n = tid / 32, l = tid % 32, is = l / 16;
        const float d = __half2float(w.q6k_d()[blk]);
        const uint8_t *ql = w.q6k_ql() + blk * 128 + 64 * n, *qh = w.q6k_qh() + blk * 64 + 32 * n;
        const int8_t *sc = w.q6k_sc() + blk * 16 + 8 * n;
        const int q1 = (int8_t)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32, q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
        const int q3 = (int8_t)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32, q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
        y[128 * n + l] = __fmul_rn(__fmul_rn(d, (float)sc[is]), (float)q1); y[128 * n + l + 32] = __fmul_rn(__fmul_rn(d, (float)sc[is + 2]), (float)q2);
        y[128 * n + l + 64] = __fmul_rn(__fmul_rn(d, (float)sc[is + 4]), (float)q3); y[128 * n + l + 96] = __fmul_rn(__fmul_rn(d, (float)sc[is + 6]), (float)q4);
    } else if (type == EB_Q4_0) {
        if (tid >= 16) return;
        float *y = out + ((int64_t)r * nbr + b) * 32;
        const float d = __half2float(w.q40_d()[blk]); const int q = w.q40_qs()[blk * 16 + tid];
        y[tid] = __fmul_rn((float)((q & 0xF) - 8), d); y[tid + 16] = __fmul_rn((float)((q >> 4) - 8), d);
    } else if (type == EB_Q8_0) {
        if (tid >= 32) return;
        out[((int64_t)r * nbr + b) * 32 + tid] = __fmul_rn((float)w.q80_qs()[blk * 32 + tid], __half2float(w.q80_d()[blk]));
    }
}
int eb_get_rows_dev(int type, const void *W, int64_t n_rows, int64_t K, const int32_t *d_ids, int n, float *out, cudaStream_t st) {
    if (n == 0) return EB_OK;
    if (type == EB_F32) get_rows_f32_kernel<<<dim3((unsigned)cdiv(K, 256), n), 256, 0, st>>>((const float *)W, K, d_ids, out);
    else if (type == EB_F16) get_rows_f16_kernel<<<dim3((unsigned)cdiv(K, 256), n), 256, 0, st>>>((const __half *)W, K, d_ids, out);
    else {
        const int be = (type == EB_Q4_K || type == EB_Q6_K) ? 256 : 32;
        if (K % be) return eb_set_err(EB_ERR_ARG, "get_rows: K");
        const int nbr = (int)(K / be);
        get_rows_q_kernel<<<dim3(nbr, n), 64, 0, st>>>(type, (const uint8_t *)W, n_rows * nbr, nbr, d_ids, out, eb_gate());
    }
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ FA-off operator surface (llm_build_kqv without flash attention,
// R/src/llama.cpp:762-812; the reference's default unless -fa is given).  Off the engine's hot path (the engine always runs the tree
// flash-attention kernel); here so that the ggml plugin keeps the whole attention block on the device for FA-off graphs too.
struct Strides4 { int64_t ne[4]; int64_t sa[4]; int64_t sd[4]; };     // element counts, source / destination strides in BYTES

// GGML_OP_CPY / CONT / DUP over arbitrary 4-D strides (R/ggml/src/ggml-cuda/cpy.cu:526: ggml_cuda_cpy); f32 -> f32 | f16, f16 -> f16.
// The element index runs over the SOURCE's logical order; ggml_cpy into a differently shaped destination of the same size walks the
// destination in its own logical order (ggml_compute_forward_dup_f32), so the destination is addressed through its own ne/strides.
template <typename TS, typename TD>
__global__ void strided_copy_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, Strides4 s, int64_t dne0, int64_t dne1, int64_t dne2, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t i0 = i % s.ne[0], i1 = (i / s.ne[0]) % s.ne[1], i2 = (i / (s.ne[0] * s.ne[1])) % s.ne[2], i3 = i / (s.ne[0] * s.ne[1] * s.ne[2]);
    const int64_t d0 = i % dne0, d1 = (i / dne0) % dne1, d2 = (i / (dne0 * dne1)) % dne2, d3 = i / (dne0 * dne1 * dne2);
    const TS v = *(const TS *)(src + i0 * s.sa[0] + i1 * s.sa[1] + i2 * s.sa[2] + i3 * s.sa[3]);
    *(TD *)(dst + d0 * s.sd[0] + d1 * s.sd[1] + d2 * s.sd[2] + d3 * s.sd[3]) = (TD)v;
}
int eb_copy_strided_impl(int src_type, int dst_type, const void *src, void *dst, const int64_t *ne, const int64_t *nb_src, const int64_t *dne, const int64_t *nb_dst, cudaStream_t st) {
    Strides4 s; int64_t n = 1, dn = 1;
    for (int k = 0; k < 4; k++) { s.ne[k] = ne[k]; s.sa[k] = nb_src[k]; s.sd[k] = nb_dst[k]; n *= ne[k]; dn *= dne[k]; }
    if (n != dn) return eb_set_err(EB_ERR_ARG, "copy_strided: %lld source elements, %lld destination elements", (long long)n, (long long)dn);
    if (n == 0) return EB_OK;
    const unsigned grid = (unsigned)cdiv(n, 256);
    if (src_type == EB_F32 && dst_type == EB_F32) strided_copy_kernel<float, float><<<grid, 256, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, s, dne[0], dne[1], dne[2], n);
    else if (src_type == EB_F32 && dst_type == EB_F16) strided_copy_kernel<float, __half><<<grid, 256, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, s, dne[0], dne[1], dne[2], n);
    else if (src_type == EB_F16 && dst_type == EB_F16) strided_copy_kernel<__half, __half><<<grid, 256, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, s, dne[0], dne[1], dne[2], n);
    else return eb_set_err(EB_ERR_UNSUPPORTED, "copy_strided: types %d -> %d", src_type, dst_type);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// GGML_OP_MUL_MAT with an f16 (or f32) src0 and batch dimension: y[b][t][n] = sum_k a[b / r][n][k] * x[b][t][k] (KQ and KQV of the
// FA-off attention).  The CPU converts x to f16 first when a is f16 (vec_dot_type, R/ggml/src/ggml-cpu/ggml-cpu.c:7588-7623) and
// accumulates the f16 products in f32 (ggml_vec_dot_f16); restated here.  One warp per output element, lanes stride over k.
template <typename TA>
__global__ void __launch_bounds__(256) mul_mat_dense_kernel(const uint8_t *__restrict__ a, const uint8_t *__restrict__ x, float *__restrict__ y,
                                                          int64_t N, int64_t K, int64_t T, int64_t B, int64_t bcast,
                                                          int64_t a_nb1, int64_t a_nb2, int64_t x_nb1, int64_t x_nb2) {
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); const int lane = threadIdx.x & 31;
    if (w >= N * T * B) return;
    const int64_t n = w % N, t = (w / N) % T, b = w / (N * T);
    const TA *ar = (const TA *)(a + n * a_nb1 + (b / bcast) * a_nb2);
    const float *xr = (const float *)(x + t * x_nb1 + b * x_nb2);
    float acc = 0.f;
    for (int64_t k = lane; k < K; k += 32) {
        if (sizeof(TA) == 2) acc += __half2float(((const __half *)ar)[k]) * __half2float(__float2half_rn(xr[k]));
        else acc += ((const float *)ar)[k] * xr[k];
    }
    acc = warp_sum(acc);
    if (lane == 0) y[w] = acc;
}
int eb_mul_mat_dense_impl(int a_type, const void *a, int64_t a_nb1, int64_t a_nb2, int64_t a_batches, const float *x, int64_t x_nb1, int64_t x_nb2,
                          float *y, int64_t N, int64_t K, int64_t T, int64_t B, cudaStream_t st) {
    if (a_batches < 1 || B % a_batches) return eb_set_err(EB_ERR_ARG, "mul_mat_dense: %lld batches of src0 for %lld of src1", (long long)a_batches, (long long)B);
    const int64_t n = N * T * B;
    if (n == 0) return EB_OK;
    const unsigned grid = (unsigned)cdiv(n, 8);
    if (a_type == EB_F16) mul_mat_dense_kernel<__half><<<grid, 256, 0, st>>>((const uint8_t *)a, (const uint8_t *)x, y, N, K, T, B, B / a_batches, a_nb1, a_nb2, x_nb1, x_nb2);
    else if (a_type == EB_F32) mul_mat_dense_kernel<float><<<grid, 256, 0, st>>>((const uint8_t *)a, (const uint8_t *)x, y, N, K, T, B, B / a_batches, a_nb1, a_nb2, x_nb1, x_nb2);
    else return eb_set_err(EB_ERR_UNSUPPORTED, "mul_mat_dense: src0 type %d", a_type);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// GGML_OP_SOFT_MAX with mask and scale (ggml_soft_max_ext, max_bias = 0): y = softmax(scale * x + mask) over rows of n values;
// mask row = (row % rows_per_batch) of an f16 or f32 [n][>= rows_per_batch] tensor, shared by all heads
// (ggml_compute_forward_soft_max_f32, R/ggml/src/ggml-cpu/ggml-cpu.c:9042: max, expf, sum in double, scale by 1/sum).
__global__ void __launch_bounds__(256) soft_max_ext_kernel(const float *__restrict__ x, const void *__restrict__ mask, int mask_f16, int64_t mask_stride,
                                                         float *__restrict__ y, int n, int64_t rows_per_batch, float scale) {
    extern __shared__ float row[];
    __shared__ float red_f[8]; __shared__ double red_d[8];
    const int64_t r = blockIdx.x; const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *xr = x + r * n;
    const int64_t mr = r % rows_per_batch;
    float mx = -INFINITY;
    for (int i = tid; i < n; i += 256) {
        float v = xr[i] * scale;
        if (mask) v += mask_f16 ? __half2float(((const __half *)mask)[mr * mask_stride + i]) : ((const float *)mask)[mr * mask_stride + i];
        row[i] = v; mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    if (lane == 0) red_f[warp] = mx;
    __syncthreads();
    mx = red_f[0];
    for (int i = 1; i < 8; i++) mx = fmaxf(mx, red_f[i]);
    double sum = 0.0;
    for (int i = tid; i < n; i += 256) { const float e = mx == -INFINITY ? 0.f : expf(row[i] - mx); row[i] = e; sum += (double)e; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) red_d[warp] = sum;
    __syncthreads();
    sum = 0.0;
    for (int i = 0; i < 8; i++) sum += red_d[i];
    const float inv = (float)(1.0 / sum);
    float *yr = y + r * n;
    for (int i = tid; i < n; i += 256) yr[i] = row[i] * inv;
}
int eb_soft_max_ext_impl(const float *x, const void *mask, int mask_type, int64_t mask_stride, float *y, int64_t n, int64_t rows, int64_t rows_per_batch, float scale, cudaStream_t st) {
    if (n < 1 || n > 48000) return eb_set_err(EB_ERR_ARG, "soft_max: n=%lld", (long long)n);
    if (mask && mask_type != EB_F16 && mask_type != EB_F32) return eb_set_err(EB_ERR_UNSUPPORTED, "soft_max: mask type %d", mask_type);
    if (rows == 0) return EB_OK;
    static bool attr_set = false;
    if (!attr_set) { EB_CUDA(cudaFuncSetAttribute(soft_max_ext_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 48000 * 4)); attr_set = true; }
    soft_max_ext_kernel<<<(unsigned)rows, 256, (size_t)n * 4, st>>>(x, mask, mask_type == EB_F16, mask_stride, y, (int)n, rows_per_batch > 0 ? rows_per_batch : rows, scale);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
