/*
 * eagle_oracle.c — CPU restatement of the reference's EAGLE speculative-step hot path.
 * TEST INFRASTRUCTURE ONLY (see eagle_oracle.h).  Plain C, scalar, single-threaded unless
 * compiled with -fopenmp (rows of a mul_mat are independent, so threading does not change
 * results).  Every function cites the reference file:line it restates; R = /root/reference/llama.cpp/.
 */
#include "eagle_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define QK_K 256
#define QK 32

/* ------------------------------------------------------------------ fp16 */
/* IEEE binary16 <-> binary32, round-to-nearest-even: what GGML_FP32_TO_FP16 does on x86 with
 * F16C (_cvtss_sh(x, 0)), R/ggml/src/ggml-impl.h:328-340. */
float eo_f16_to_f32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ff, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            do { e++; man <<= 1; } while (!(man & 0x400));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ff) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}

uint16_t eo_f32_to_f16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000;
    uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) return (uint16_t)(sign | 0x7c00 | ((ax > 0x7f800000u) ? 0x200 | ((ax >> 13) & 0x3ff) : 0));
    if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00);            /* overflows after rounding */
    if (ax < 0x33000001u) return (uint16_t)sign;                         /* < 2^-25 (ties to even -> 0) */
    int e = (int)(ax >> 23) - 127;
    uint32_t man = (ax & 0x7fffff) | 0x800000;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;
    uint32_t half = 1u << (shift - 1), rem = man & ((1u << shift) - 1), q = man >> shift;
    if (rem > half || (rem == half && (q & 1))) q++;
    if (e < -14) return (uint16_t)(sign | q);                            /* subnormal (q may carry into exp) */
    return (uint16_t)(sign | (((uint32_t)(e + 15) << 10) + (q - 0x400)));
}

/* ------------------------------------------------------------------ block formats */
/* R/ggml/src/ggml-common.h:161-327 */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[QK / 2]; } blk_q4_0;                       /* 18 B */
typedef struct { uint16_t d; int8_t qs[QK]; } blk_q8_0;                            /* 34 B */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[QK_K / 2]; } blk_q4_K; /* 144 B */
typedef struct { uint8_t ql[QK_K / 2]; uint8_t qh[QK_K / 4]; int8_t scales[QK_K / 16]; uint16_t d; } blk_q6_K; /* 210 B */
typedef struct { float d; int8_t qs[QK_K]; int16_t bsums[QK_K / 16]; } blk_q8_K;    /* 292 B */
#pragma pack(pop)

int eo_block_elems(int type) {
    switch (type) {
        case EO_Q4_0: case EO_Q8_0: return QK;
        case EO_Q4_K: case EO_Q6_K: return QK_K;
        default: return 1;
    }
}

size_t eo_row_size(int type, int64_t k) {
    switch (type) {
        case EO_F32: return (size_t)k * 4;
        case EO_F16: return (size_t)k * 2;
        case EO_Q4_0: return (size_t)(k / QK) * sizeof(blk_q4_0);
        case EO_Q8_0: return (size_t)(k / QK) * sizeof(blk_q8_0);
        case EO_Q4_K: return (size_t)(k / QK_K) * sizeof(blk_q4_K);
        case EO_Q6_K: return (size_t)(k / QK_K) * sizeof(blk_q6_K);
    }
    return 0;
}

/* R/ggml/src/ggml-quants.c:631 */
static inline void get_scale_min_k4(int j, const uint8_t *q, uint8_t *d, uint8_t *m) {
    if (j < 4) { *d = q[j] & 63; *m = q[j + 4] & 63; }
    else {
        *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        *m = (q[j + 4] >> 4) | ((q[j - 0] >> 6) << 4);
    }
}

void eo_dequantize_row(int type, const void *src, float *y, int64_t k) {
    switch (type) {
    case EO_F32: memcpy(y, src, (size_t)k * 4); return;
    case EO_F16: { const uint16_t *x = src; for (int64_t i = 0; i < k; i++) y[i] = eo_f16_to_f32(x[i]); return; }
    case EO_Q4_0: {                                   /* ggml-quants.c:255 */
        const blk_q4_0 *x = src;
        for (int64_t i = 0; i < k / QK; i++) {
            const float d = eo_f16_to_f32(x[i].d);
            for (int j = 0; j < QK / 2; j++) {
                y[i * QK + j] = ((x[i].qs[j] & 0x0F) - 8) * d;
                y[i * QK + j + QK / 2] = ((x[i].qs[j] >> 4) - 8) * d;
            }
        }
        return; }
    case EO_Q8_0: {                                   /* ggml-quants.c:349 */
        const blk_q8_0 *x = src;
        for (int64_t i = 0; i < k / QK; i++) {
            const float d = eo_f16_to_f32(x[i].d);
            for (int j = 0; j < QK; j++) y[i * QK + j] = x[i].qs[j] * d;
        }
        return; }
    case EO_Q4_K: {                                   /* ggml-quants.c:1280 */
        const blk_q4_K *x = src;
        for (int64_t i = 0; i < k / QK_K; i++) {
            const uint8_t *q = x[i].qs;
            const float d = eo_f16_to_f32(x[i].d), min = eo_f16_to_f32(x[i].dmin);
            int is = 0; uint8_t sc, m;
            for (int j = 0; j < QK_K; j += 64) {
                get_scale_min_k4(is + 0, x[i].scales, &sc, &m);
                const float d1 = d * sc, m1 = min * m;
                get_scale_min_k4(is + 1, x[i].scales, &sc, &m);
                const float d2 = d * sc, m2 = min * m;
                for (int l = 0; l < 32; ++l) *y++ = d1 * (q[l] & 0xF) - m1;
                for (int l = 0; l < 32; ++l) *y++ = d2 * (q[l] >> 4) - m2;
                q += 32; is += 2;
            }
        }
        return; }
    case EO_Q6_K: {                                   /* ggml-quants.c:1690 */
        const blk_q6_K *x = src;
        for (int64_t i = 0; i < k / QK_K; i++) {
            const float d = eo_f16_to_f32(x[i].d);
            const uint8_t *ql = x[i].ql, *qh = x[i].qh; const int8_t *sc = x[i].scales;
            for (int n = 0; n < QK_K; n += 128) {
                for (int l = 0; l < 32; ++l) {
                    int is = l / 16;
                    const int8_t q1 = (int8_t)((ql[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                    const int8_t q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                    const int8_t q3 = (int8_t)((ql[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                    const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                    y[l + 0] = d * sc[is + 0] * q1;
                    y[l + 32] = d * sc[is + 2] * q2;
                    y[l + 64] = d * sc[is + 4] * q3;
                    y[l + 96] = d * sc[is + 6] * q4;
                }
                y += 128; ql += 64; qh += 32; sc += 8;
            }
        }
        return; }
    }
}

/* ------------------------------------------------------------------ activation quantisers */
static inline int nearest_int(float fval) {            /* ggml-quants.c:559 */
    float val = fval + 12582912.f;
    int i; memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

void eo_quantize_row_q8_K(const float *x, void *vy, int64_t k) {   /* ggml-quants.c:2479 */
    blk_q8_K *y = vy;
    for (int64_t i = 0; i < k / QK_K; i++) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; ++j) { float ax = fabsf(x[j]); if (ax > amax) { amax = ax; max = x[j]; } }
        if (!amax) { y[i].d = 0; memset(y[i].qs, 0, QK_K); memset(y[i].bsums, 0, sizeof(y[i].bsums)); x += QK_K; continue; }
        const float iscale = -127.f / max;
        for (int j = 0; j < QK_K; ++j) { int v = nearest_int(iscale * x[j]); y[i].qs[j] = (int8_t)(v < 127 ? v : 127); }
        for (int j = 0; j < QK_K / 16; ++j) {
            int sum = 0;
            for (int ii = 0; ii < 16; ++ii) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1 / iscale;
        x += QK_K;
    }
}

void eo_quantize_row_q8_0(const float *x, void *vy, int64_t k) {   /* ggml-cpu-quants.c:738, AVX2 branch 795-860 */
    blk_q8_0 *y = vy;
    for (int64_t i = 0; i < k / QK; i++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) { float a = fabsf(x[i * QK + j]); if (a > amax) amax = a; }
        const float d = amax / 127.f;
        y[i].d = eo_f32_to_f16(d);
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        for (int j = 0; j < QK; ++j) y[i].qs[j] = (int8_t)nearbyintf(x[i * QK + j] * id);   /* round-to-even */
    }
}

/* ------------------------------------------------------------------ dot products */
/* scalar tails/branches of ggml-cpu-quants.c: q4_0 :2373-2389, q8_0 :3990-4000 */
static float vec_dot_q4_0_q8_0(int64_t n, const blk_q4_0 *x, const blk_q8_0 *y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < n / QK; ++ib) {
        int sumi0 = 0, sumi1 = 0;
        for (int j = 0; j < QK / 2; ++j) {
            const int v0 = (x[ib].qs[j] & 0x0F) - 8, v1 = (x[ib].qs[j] >> 4) - 8;
            sumi0 += v0 * y[ib].qs[j];
            sumi1 += v1 * y[ib].qs[j + QK / 2];
        }
        sumf += (sumi0 + sumi1) * eo_f16_to_f32(x[ib].d) * eo_f16_to_f32(y[ib].d);
    }
    return sumf;
}

static float vec_dot_q8_0_q8_0(int64_t n, const blk_q8_0 *x, const blk_q8_0 *y) {
    float sumf = 0;
    for (int64_t ib = 0; ib < n / QK; ++ib) {
        int sumi = 0;
        for (int j = 0; j < QK; j++) sumi += x[ib].qs[j] * y[ib].qs[j];
        sumf += sumi * (eo_f16_to_f32(x[ib].d) * eo_f16_to_f32(y[ib].d));
    }
    return sumf;
}

/* ggml-cpu-quants.c:7021-7077 (generic branch of ggml_vec_dot_q4_K_q8_K) */
static float vec_dot_q4_K_q8_K(int64_t n, const blk_q4_K *x, const blk_q8_K *y) {
    static const uint32_t kmask1 = 0x3f3f3f3f, kmask2 = 0x0f0f0f0f, kmask3 = 0x03030303;
    uint32_t utmp[4];
    const uint8_t *scales = (const uint8_t *)&utmp[0], *mins = (const uint8_t *)&utmp[2];
    int8_t aux8[QK_K]; int16_t aux16[8]; float sums[8]; int32_t aux32[8];
    memset(sums, 0, sizeof(sums));
    float sumf = 0;
    for (int64_t i = 0; i < n / QK_K; ++i) {
        const uint8_t *q4 = x[i].qs; const int8_t *q8 = y[i].qs;
        memset(aux32, 0, sizeof(aux32));
        int8_t *a = aux8;
        for (int j = 0; j < QK_K / 64; ++j) {
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] & 0xF);
            a += 32;
            for (int l = 0; l < 32; ++l) a[l] = (int8_t)(q4[l] >> 4);
            a += 32; q4 += 32;
        }
        memcpy(utmp, x[i].scales, 12);
        utmp[3] = ((utmp[2] >> 4) & kmask2) | (((utmp[1] >> 6) & kmask3) << 4);
        const uint32_t uaux = utmp[1] & kmask1;
        utmp[1] = (utmp[2] & kmask2) | (((utmp[0] >> 6) & kmask3) << 4);
        utmp[2] = uaux; utmp[0] &= kmask1;
        int sumi = 0;
        for (int j = 0; j < QK_K / 16; ++j) sumi += y[i].bsums[j] * mins[j / 2];
        a = aux8; int is = 0;
        for (int j = 0; j < QK_K / 32; ++j) {
            int32_t scale = scales[is++];
            for (int r = 0; r < 4; r++) {
                for (int l = 0; l < 8; ++l) aux16[l] = q8[l] * a[l];
                for (int l = 0; l < 8; ++l) aux32[l] += scale * aux16[l];
                q8 += 8; a += 8;
            }
        }
        const float d = eo_f16_to_f32(x[i].d) * y[i].d;
        for (int l = 0; l < 8; ++l) sums[l] += d * aux32[l];
        const float dmin = eo_f16_to_f32(x[i].dmin) * y[i].d;
        sumf -= dmin * sumi;
    }
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    return sumf;
}

/* ggml-cpu-quants.c:8560-8610 (generic branch of ggml_vec_dot_q6_K_q8_K) */
static float vec_dot_q6_K_q8_K(int64_t n, const blk_q6_K *x, const blk_q8_K *y) {
    int8_t aux8[QK_K]; int16_t aux16[8]; float sums[8]; int32_t aux32[8];
    memset(sums, 0, sizeof(sums));
    float sumf = 0;
    for (int64_t i = 0; i < n / QK_K; ++i) {
        const uint8_t *q4 = x[i].ql, *qh = x[i].qh; const int8_t *q8 = y[i].qs;
        memset(aux32, 0, sizeof(aux32));
        int8_t *a = aux8;
        for (int j = 0; j < QK_K; j += 128) {
            for (int l = 0; l < 32; ++l) {
                a[l + 0] = (int8_t)((q4[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                a[l + 32] = (int8_t)((q4[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                a[l + 64] = (int8_t)((q4[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                a[l + 96] = (int8_t)((q4[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
            }
            a += 128; q4 += 64; qh += 32;
        }
        a = aux8; int is = 0;
        for (int j = 0; j < QK_K / 16; ++j) {
            int scale = x[i].scales[is++];
            for (int r = 0; r < 2; r++) {
                for (int l = 0; l < 8; ++l) aux16[l] = q8[l] * a[l];
                for (int l = 0; l < 8; ++l) aux32[l] += scale * aux16[l];
                q8 += 8; a += 8;
            }
        }
        const float d = eo_f16_to_f32(x[i].d) * y[i].d;
        for (int l = 0; l < 8; ++l) sums[l] += d * aux32[l];
    }
    for (int l = 0; l < 8; ++l) sumf += sums[l];
    return sumf;
}

float eo_vec_dot(int type, int64_t k, const void *w, const void *a) {
    switch (type) {
        case EO_Q4_0: return vec_dot_q4_0_q8_0(k, w, a);
        case EO_Q8_0: return vec_dot_q8_0_q8_0(k, w, a);
        case EO_Q4_K: return vec_dot_q4_K_q8_K(k, w, a);
        case EO_Q6_K: return vec_dot_q6_K_q8_K(k, w, a);
    }
    return NAN;
}

/* ggml_compute_forward_mul_mat, ggml-cpu.c:7526: src1 rows are converted to the weight type's
 * vec_dot_type (Q8_0 for Q4_0/Q8_0, Q8_K for Q4_K/Q6_K; table ggml-cpu.c:254-419), then every
 * output is one vec_dot.  F32/F16 weights: plain f32 dot (f16 weights: activations rounded to
 * f16 first, as vec_dot_type F16 does). */
void eo_mul_mat(int type, const void *W, int64_t N, int64_t K, const float *X, int64_t T, float *Y) {
    const size_t rs = eo_row_size(type, K);
    if (type == EO_F32 || type == EO_F16) {
        for (int64_t t = 0; t < T; t++) {
            #pragma omp parallel for schedule(static)
            for (int64_t n = 0; n < N; n++) {
                double acc = 0;
                if (type == EO_F32) { const float *w = (const float *)W + n * K; for (int64_t k = 0; k < K; k++) acc += (double)w[k] * X[t * K + k]; }
                else { const uint16_t *w = (const uint16_t *)W + n * K; for (int64_t k = 0; k < K; k++) acc += (double)eo_f16_to_f32(w[k]) * eo_f16_to_f32(eo_f32_to_f16(X[t * K + k])); }
                Y[t * N + n] = (float)acc;
            }
        }
        return;
    }
    const int kq = (type == EO_Q4_K || type == EO_Q6_K);
    const size_t as = kq ? (size_t)(K / QK_K) * sizeof(blk_q8_K) : (size_t)(K / QK) * sizeof(blk_q8_0);
    char *aq = malloc(as * (size_t)T);
    for (int64_t t = 0; t < T; t++) {
        if (kq) eo_quantize_row_q8_K(X + t * K, aq + t * as, K);
        else eo_quantize_row_q8_0(X + t * K, aq + t * as, K);
    }
    #pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++)
        for (int64_t t = 0; t < T; t++)
            Y[t * N + n] = eo_vec_dot(type, K, (const char *)W + n * rs, aq + t * as);
    free(aq);
}

/* ------------------------------------------------------------------ glue ops */
void eo_rms_norm(const float *x, float *y, int64_t n, int64_t rows, float eps) {  /* ggml-cpu.c:7098 */
    for (int64_t r = 0; r < rows; r++) {
        double sum = 0.0;
        for (int64_t i = 0; i < n; i++) sum += (double)(x[r * n + i] * x[r * n + i]);
        const float mean = (float)(sum / n);
        const float scale = 1.0f / sqrtf(mean + eps);
        for (int64_t i = 0; i < n; i++) y[r * n + i] = x[r * n + i] * scale;
    }
}

void eo_rope_norm(float *x, const int32_t *pos, int D, int n_head, int T, float freq_base) { /* ggml-cpu.c:9449, 9376 */
    const float theta_scale = powf(freq_base, -2.0f / D);
    for (int t = 0; t < T; t++) {
        float theta = (float)pos[t];
        for (int i0 = 0; i0 < D; i0 += 2) {
            const float c = cosf(theta), s = sinf(theta);
            for (int h = 0; h < n_head; h++) {
                float *p = x + ((size_t)t * n_head + h) * D + i0;
                const float x0 = p[0], x1 = p[1];
                p[0] = x0 * c - x1 * s;
                p[1] = x0 * s + x1 * c;
            }
            theta *= theta_scale;
        }
    }
}

void eo_get_rows(int type, const void *W, int64_t K, const int32_t *ids, int n, float *out) { /* ggml-cpu.c:8547 */
    const size_t rs = eo_row_size(type, K);
    for (int i = 0; i < n; i++) eo_dequantize_row(type, (const char *)W + (size_t)ids[i] * rs, out + (size_t)i * K, K);
}

/* ggml-cpu.c:10999-11212 with K,V f16: Q row -> f16, s = dot(K,Q) in f32, online softmax with an
 * f16 VKQ accumulator (ggml_vec_scale_f16 / ggml_vec_mad_f16 round to f16 after every op). */
static int g_fa_acc_f32 = 0;
/* 0 (default): the reference's f16 VKQ accumulator.  1: f32 accumulator — NOT the reference's
 * arithmetic; it is the checker for the GPU kernel's documented deviation:
 * everything else in the forward stays the reference's. */
void eo_set_fattn_acc_f32(int on) { g_fa_acc_f32 = on; }

static void flash_attn_f32acc(const float *q, const uint16_t *k, const uint16_t *v, const uint8_t *vis,
                              int T, int H, int D, int n_kv, float scale, float *out) {
    #pragma omp parallel for collapse(2) schedule(static)
    for (int t = 0; t < T; t++) for (int h = 0; h < H; h++) {
        float qf[512], acc[512];
        const float *pq = q + ((size_t)t * H + h) * D;
        for (int d = 0; d < D; d++) { qf[d] = eo_f16_to_f32(eo_f32_to_f16(pq[d])); acc[d] = 0; }
        float S = 0.0f, M = -INFINITY;
        for (int ic = 0; ic < n_kv; ic++) {
            if (!vis[(size_t)t * n_kv + ic]) continue;
            const uint16_t *kr = k + ((size_t)ic * H + h) * D, *vr = v + ((size_t)ic * H + h) * D;
            float s = 0; for (int d = 0; d < D; d++) s += eo_f16_to_f32(kr[d]) * qf[d];
            s = s * scale;
            const float Mold = M; float ms = 1.0f, vs = 1.0f;
            if (s > M) { M = s; ms = expf(Mold - M); for (int d = 0; d < D; d++) acc[d] *= ms; }
            else vs = expf(s - M);
            for (int d = 0; d < D; d++) acc[d] += eo_f16_to_f32(vr[d]) * vs;
            S = S * ms + vs;
        }
        const float S_inv = 1.0f / S;
        float *po = out + ((size_t)t * H + h) * D;
        for (int d = 0; d < D; d++) po[d] = acc[d] * S_inv;
    }
}

void eo_flash_attn(const float *q, const uint16_t *k, const uint16_t *v, const uint8_t *vis,
                   int T, int H, int D, int n_kv, float scale, float *out) {
    if (g_fa_acc_f32) { flash_attn_f32acc(q, k, v, vis, T, H, D, n_kv, scale, out); return; }
    #pragma omp parallel for collapse(2) schedule(static)
    for (int t = 0; t < T; t++) for (int h = 0; h < H; h++) {
        float qf[512]; uint16_t acc[512];
        const float *pq = q + ((size_t)t * H + h) * D;
        for (int d = 0; d < D; d++) { qf[d] = eo_f16_to_f32(eo_f32_to_f16(pq[d])); acc[d] = 0; }
        float S = 0.0f, M = -INFINITY;
        for (int ic = 0; ic < n_kv; ic++) {
            if (!vis[(size_t)t * n_kv + ic]) continue;
            const uint16_t *kr = k + ((size_t)ic * H + h) * D, *vr = v + ((size_t)ic * H + h) * D;
            float s = 0; for (int d = 0; d < D; d++) s += eo_f16_to_f32(kr[d]) * qf[d];
            s = s * scale;
            const float Mold = M; float ms = 1.0f, vs = 1.0f;
            if (s > M) { M = s; ms = expf(Mold - M); for (int d = 0; d < D; d++) acc[d] = eo_f32_to_f16(eo_f16_to_f32(acc[d]) * ms); }
            else vs = expf(s - M);
            for (int d = 0; d < D; d++) acc[d] = eo_f32_to_f16(eo_f16_to_f32(acc[d]) + eo_f16_to_f32(vr[d]) * vs);
            S = S * ms + vs;
        }
        const float S_inv = 1.0f / S;
        float *po = out + ((size_t)t * H + h) * D;
        for (int d = 0; d < D; d++) po[d] = eo_f16_to_f32(acc[d]) * S_inv;
    }
}

void eo_build_kq_mask(const int32_t *cell_pos, const uint64_t *cell_seq, int n_kv,
                      const int32_t *q_pos, const int32_t *q_seq0, int T, uint8_t *vis) { /* llama-context.cpp:161-193 */
    for (int j = 0; j < T; j++)
        for (int i = 0; i < n_kv; i++)
            vis[(size_t)j * n_kv + i] = ((cell_seq[i] >> q_seq0[j]) & 1) && cell_pos[i] <= q_pos[j];
}

void eo_top_k(const float *logits, int V, int k, int32_t *ids, float *vals, float *probs) { /* llama-sampling.cpp:228-302 */
    int n = 0;
    for (int i = 0; i < V; i++) {
        int p = n;
        while (p > 0 && logits[i] > vals[p - 1]) p--;      /* strict >: earlier (lower) id wins ties */
        if (p >= k) continue;
        int end = n < k ? n : k - 1;
        for (int j = end; j > p; j--) { vals[j] = vals[j - 1]; ids[j] = ids[j - 1]; }
        vals[p] = logits[i]; ids[p] = i;
        if (n < k) n++;
    }
    float mx = vals[0], sum = 0;                            /* llama_sampler_softmax_impl :130-150 */
    for (int i = 0; i < n; i++) { probs[i] = expf(vals[i] - mx); sum += probs[i]; }
    for (int i = 0; i < n; i++) probs[i] /= sum;
}

int eo_argmax(const float *logits, int V) {
    int b = 0;
    for (int i = 1; i < V; i++) if (logits[i] > logits[b]) b = i;
    return b;
}

/* ------------------------------------------------------------------ model forward */
typedef struct { int type; const void *data; } eo_tensor;
struct eo_model {
    int E, H, F, V, L, is_eagle; float eps, freq_base;
    eo_tensor g[EO_T_GLOBAL_COUNT];
    eo_tensor (*layers)[EO_L_COUNT];
};
struct eo_kv { int L, E, n_ctx; uint16_t **k, **v; };

eo_model *eo_model_new(int E, int H, int F, int V, int L, float eps, float freq_base, int is_eagle) {
    eo_model *m = calloc(1, sizeof(*m));
    m->E = E; m->H = H; m->F = F; m->V = V; m->L = L; m->eps = eps; m->freq_base = freq_base; m->is_eagle = is_eagle;
    m->layers = calloc((size_t)L, sizeof(*m->layers));
    return m;
}
void eo_model_free(eo_model *m) { if (m) { free(m->layers); free(m); } }
void eo_model_set_global(eo_model *m, int which, int type, const void *data) { m->g[which].type = type; m->g[which].data = data; }
void eo_model_set_layer(eo_model *m, int layer, int which, int type, const void *data) { m->layers[layer][which].type = type; m->layers[layer][which].data = data; }
void eo_model_set_lm_head_from(eo_model *m, const eo_model *t) { m->g[EO_T_OUTPUT] = t->g[EO_T_OUTPUT]; }

eo_kv *eo_kv_new(int L, int E, int n_ctx) {
    eo_kv *kv = calloc(1, sizeof(*kv));
    kv->L = L; kv->E = E; kv->n_ctx = n_ctx;
    kv->k = calloc((size_t)L, sizeof(void *)); kv->v = calloc((size_t)L, sizeof(void *));
    for (int l = 0; l < L; l++) { kv->k[l] = calloc((size_t)n_ctx * E, 2); kv->v[l] = calloc((size_t)n_ctx * E, 2); }
    return kv;
}
void eo_kv_free(eo_kv *kv) { if (!kv) return; for (int l = 0; l < kv->L; l++) { free(kv->k[l]); free(kv->v[l]); } free(kv->k); free(kv->v); free(kv); }
uint16_t *eo_kv_k(eo_kv *kv, int l) { return kv->k[l]; }
uint16_t *eo_kv_v(eo_kv *kv, int l) { return kv->v[l]; }

static void mm(const eo_tensor *w, int64_t N, int64_t K, const float *x, int T, float *y) { eo_mul_mat(w->type, w->data, N, K, x, T, y); }

/* Activation tap (checker for rounding flips, tests only): when set, eo_forward appends the int8 quants [T][K] of every
 * activation-quantisation site, on the grid the following mul_mat uses (from_float of its vec_dot_type, ggml-cpu.c:254-419,
 * 7588-7623), in the order: [fc input] ; per layer wq|wk|wv input, wo input, gate|up input, ffn_down input ; lm_head input. */
static int8_t *g_tap; static size_t g_tap_cap, g_tap_off;
void eo_set_tap(int8_t *buf, size_t cap) { g_tap = buf; g_tap_cap = buf ? cap : 0; g_tap_off = 0; }
size_t eo_tap_bytes(void) { return g_tap_off; }
static void tap_site(int wtype, const float *x, int64_t K, int T) {
    if (!g_tap || g_tap_off + (size_t)T * K > g_tap_cap) return;
    const int kq = (wtype == EO_Q4_K || wtype == EO_Q6_K);
    if (!kq && wtype != EO_Q4_0 && wtype != EO_Q8_0) { g_tap_off += (size_t)T * K; return; }
    for (int t = 0; t < T; t++) {
        int8_t *dst = g_tap + g_tap_off + (size_t)t * K;
        if (kq) {
            blk_q8_K *b = malloc((size_t)(K / QK_K) * sizeof(blk_q8_K));
            eo_quantize_row_q8_K(x + (size_t)t * K, b, K);
            for (int64_t i = 0; i < K / QK_K; i++) memcpy(dst + i * QK_K, b[i].qs, QK_K);
            free(b);
        } else {
            blk_q8_0 *b = malloc((size_t)(K / QK) * sizeof(blk_q8_0));
            eo_quantize_row_q8_0(x + (size_t)t * K, b, K);
            for (int64_t i = 0; i < K / QK; i++) memcpy(dst + i * QK, b[i].qs, QK);
            free(b);
        }
    }
    g_tap_off += (size_t)T * K;
}

/* build_llama R/src/llama.cpp:1647-1811 and build_eagle :1839-2019 (FA on, f16 KV, no biases
 * except fc_bias, rope type NORM, kq_scale 1/sqrt(D)).  Differences of the eagle head, all the
 * reference's behaviour: input = relu(fc([embd ; hidden]) + fc_bias) (:1863-1869, :367-420);
 * attn_norm and output_norm have no weight (tensors not loaded, R/src/llama-model.cpp:1594,1608);
 * lm_head is the target's output matrix (:2016). */
void eo_forward(const eo_model *m, eo_kv *kv, const int32_t *tokens, const int32_t *pos,
                const float *hidden_in, int T, int kv_head, int n_kv, const uint8_t *vis,
                float *logits, float *feat) {
    const int E = m->E, H = m->H, F = m->F, D = E / H;
    const size_t TE = (size_t)T * E;
    float *x = malloc(TE * 4), *cur = malloc(TE * 4), *q = malloc(TE * 4), *k = malloc(TE * 4), *v = malloc(TE * 4);
    float *att = malloc(TE * 4), *gate = malloc((size_t)T * F * 4), *up = malloc((size_t)T * F * 4);

    eo_get_rows(m->g[EO_T_TOK_EMBD].type, m->g[EO_T_TOK_EMBD].data, E, tokens, T, x);
    if (m->is_eagle) {
        float *cat = malloc(TE * 2 * 4);
        for (int t = 0; t < T; t++) {                       /* ggml_concat(inpL, hidd, 0): embedding first */
            memcpy(cat + (size_t)t * 2 * E, x + (size_t)t * E, (size_t)E * 4);
            memcpy(cat + (size_t)t * 2 * E + E, hidden_in + (size_t)t * E, (size_t)E * 4);
        }
        tap_site(m->g[EO_T_FC].type, cat, 2 * E, T);
        mm(&m->g[EO_T_FC], E, 2 * E, cat, T, x);
        const float *b = m->g[EO_T_FC_BIAS].data;
        for (int t = 0; t < T; t++) for (int i = 0; i < E; i++) {
            float r = x[(size_t)t * E + i] + (b ? b[i] : 0.0f);
            x[(size_t)t * E + i] = r > 0.f ? r : 0.f;       /* ggml_relu: (x > 0) ? x : 0, ggml-cpu.c:6547 */
        }
        free(cat);
    }
    for (int l = 0; l < m->L; l++) {
        const eo_tensor *w = m->layers[l];
        eo_rms_norm(x, cur, E, T, m->eps);
        if (w[EO_L_ATTN_NORM].data) { const float *g = w[EO_L_ATTN_NORM].data; for (size_t i = 0; i < TE; i++) cur[i] *= g[i % E]; }
        tap_site(w[EO_L_WQ].type, cur, E, T);
        mm(&w[EO_L_WQ], E, E, cur, T, q); mm(&w[EO_L_WK], E, E, cur, T, k); mm(&w[EO_L_WV], E, E, cur, T, v);
        eo_rope_norm(q, pos, D, H, T, m->freq_base);
        eo_rope_norm(k, pos, D, H, T, m->freq_base);
        for (size_t i = 0; i < TE; i++) {                   /* llm_build_kv_store :228-270, f32 -> f16 */
            kv->k[l][(size_t)kv_head * E + i] = eo_f32_to_f16(k[i]);
            kv->v[l][(size_t)kv_head * E + i] = eo_f32_to_f16(v[i]);
        }
        eo_flash_attn(q, kv->k[l], kv->v[l], vis, T, H, D, n_kv, 1.0f / sqrtf((float)D), att);
        tap_site(w[EO_L_WO].type, att, E, T);
        mm(&w[EO_L_WO], E, E, att, T, cur);
        for (size_t i = 0; i < TE; i++) x[i] = cur[i] + x[i];          /* ffn_inp = cur + inpSA */
        eo_rms_norm(x, cur, E, T, m->eps);
        if (w[EO_L_FFN_NORM].data) { const float *g = w[EO_L_FFN_NORM].data; for (size_t i = 0; i < TE; i++) cur[i] *= g[i % E]; }
        tap_site(w[EO_L_UP].type, cur, E, T);
        mm(&w[EO_L_UP], F, E, cur, T, up); mm(&w[EO_L_GATE], F, E, cur, T, gate);
        for (size_t i = 0; i < (size_t)T * F; i++) {        /* ggml_silu_f32: x/(1+expf(-x)), ggml-cpu.c:1838; then mul */
            const float g = gate[i];
            gate[i] = (g / (1.0f + expf(-g))) * up[i];
        }
        tap_site(w[EO_L_DOWN].type, gate, F, T);
        mm(&w[EO_L_DOWN], E, F, gate, T, cur);
        for (size_t i = 0; i < TE; i++) x[i] = cur[i] + x[i];
    }
    eo_rms_norm(x, cur, E, T, m->eps);
    if (m->g[EO_T_OUTPUT_NORM].data) { const float *g = m->g[EO_T_OUTPUT_NORM].data; for (size_t i = 0; i < TE; i++) cur[i] *= g[i % E]; }
    if (feat) memcpy(feat, cur, TE * 4);                    /* "result_norm" */
    tap_site(m->g[EO_T_OUTPUT].type, cur, E, T);
    if (logits) mm(&m->g[EO_T_OUTPUT], m->V, E, cur, T, logits);
    free(x); free(cur); free(q); free(k); free(v); free(att); free(gate); free(up);
}

/* ------------------------------------------------------------------ speculative step pieces */
void eo_tree_vis(const int32_t *parent, int T, uint8_t *vt) {
    memset(vt, 0, (size_t)T * T);
    for (int t = 0; t < T; t++) for (int a = t; a >= 0; a = parent[a]) vt[(size_t)t * T + a] = 1;
}

/* R/examples/speculative/speculative-eagle.cpp:395-449 (greedy): at each depth the target's
 * argmax at the last kept node is compared with every active branch's token at that depth;
 * the first matching branch (lowest index = earliest created child) is kept. */
int eo_accept_walk(const int32_t *tokens, const int32_t *parent, int T, const int32_t *best,
                   int32_t *path, int32_t *bonus) {
    int cur = 0, n = 0;
    path[n++] = 0;
    for (;;) {
        int next = -1;
        for (int c = cur + 1; c < T; c++) if (parent[c] == cur && tokens[c] == best[cur]) { next = c; break; }
        if (next < 0) break;
        cur = next; path[n++] = cur;
    }
    *bonus = best[cur];
    return n;
}

void eo_kv_compact(eo_kv *kv, int kv_head, const int32_t *path, int n_acc) {
    const size_t rb = (size_t)kv->E * 2;
    for (int l = 0; l < kv->L; l++)
        for (int i = 0; i < n_acc; i++) {                   /* path[i] >= i, so ascending order is safe */
            if (path[i] == i) continue;
            memcpy((char *)kv->k[l] + (size_t)(kv_head + i) * rb, (char *)kv->k[l] + (size_t)(kv_head + path[i]) * rb, rb);
            memcpy((char *)kv->v[l] + (size_t)(kv_head + i) * rb, (char *)kv->v[l] + (size_t)(kv_head + path[i]) * rb, rb);
        }
}
