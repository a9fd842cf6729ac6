// quant.cu — dequantisation (bit-exact to the reference), activation quantisers on the CPU
// backend's grids, and synthetic-tensor generators for bench fixtures.
#include "common.cuh"
#include "spec.cuh"
#include "quant_dev.cuh"

// ------------------------------------------------------------------ dequantize (K13)
// One thread per 4 output floats would be faster; these kernels exist for parity tests and
// get_rows, so they favour a direct restatement.  R/ggml/src/ggml-quants.c:255,349,1280,1690.
__global__ void dequant_q4_0_kernel(WPlanes w, float *__restrict__ y, int64_t nb) {
    int64_t i = (int64_t)blockIdx.x * (blockDim.x / 16) + threadIdx.x / 16;
    if (i >= nb) return;
    const int j = threadIdx.x % 16;
    const float d = __half2float(w.q40_d()[i]);
    const int q = w.q40_qs()[i * 16 + j];
    y[i * 32 + j] = __fmul_rn((float)((q & 0xF) - 8), d);
    y[i * 32 + j + 16] = __fmul_rn((float)((q >> 4) - 8), d);
}
__global__ void dequant_q8_0_kernel(WPlanes w, float *__restrict__ y, int64_t nb) {
    int64_t i = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (i >= nb) return;
    const int j = threadIdx.x % 32;
    y[i * 32 + j] = __fmul_rn((float)w.q80_qs()[i * 32 + j], __half2float(w.q80_d()[i]));
}
__global__ void dequant_q4_K_kernel(const block_q4_K *__restrict__ x, float *__restrict__ y, int64_t nb) {
    const int64_t i = blockIdx.x;           // 1 block of 64 threads per super-block, 4 values each
    if (i >= nb) return;
    const int tid = threadIdx.x, j = tid / 16, l0 = (tid % 16) * 2;   // j: 64-element group, l0: 2 bytes
    const float d = __half2float(x[i].d), mn = __half2float(x[i].dmin);
    int sc1, m1, sc2, m2;
    q4k_scale_min(2 * j, x[i].scales, sc1, m1);
    q4k_scale_min(2 * j + 1, x[i].scales, sc2, m2);
    const float d1 = __fmul_rn(d, (float)sc1), mm1 = __fmul_rn(mn, (float)m1);
    const float d2 = __fmul_rn(d, (float)sc2), mm2 = __fmul_rn(mn, (float)m2);
    float *yy = y + i * QK_K + j * 64;
#pragma unroll
    for (int l = l0; l < l0 + 2; l++) {
        const int q = x[i].qs[j * 32 + l];
        yy[l] = __fsub_rn(__fmul_rn(d1, (float)(q & 0xF)), mm1);
        yy[l + 32] = __fsub_rn(__fmul_rn(d2, (float)(q >> 4)), mm2);
    }
}
__global__ void dequant_q6_K_kernel(WPlanes w, float *__restrict__ y, int64_t nb) {
    const int64_t i = blockIdx.x;           // 64 threads: n = tid/32 half, l = tid%32
    if (i >= nb) return;
    const int n = threadIdx.x / 32, l = threadIdx.x % 32, is = l / 16;
    const float d = __half2float(w.q6k_d()[i]);
    const uint8_t *ql = w.q6k_ql() + i * 128 + 64 * n, *qh = w.q6k_qh() + i * 64 + 32 * n;
    const int8_t *sc = w.q6k_sc() + i * 16 + 8 * n;
    const int q1 = (int8_t)((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
    const int q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
    const int q3 = (int8_t)((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
    const int q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
    float *yy = y + i * QK_K + 128 * n;
    yy[l] = __fmul_rn(__fmul_rn(d, (float)sc[is + 0]), (float)q1);
    yy[l + 32] = __fmul_rn(__fmul_rn(d, (float)sc[is + 2]), (float)q2);
    yy[l + 64] = __fmul_rn(__fmul_rn(d, (float)sc[is + 4]), (float)q3);
    yy[l + 96] = __fmul_rn(__fmul_rn(d, (float)sc[is + 6]), (float)q4);
}
__global__ void dequant_f16_kernel(const __half *__restrict__ x, float *__restrict__ y, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __half2float(x[i]);
}

// src is in DEVICE layout (see WPlanes); blk0 = first block to convert, n = element count.
int eb_dequantize_impl(int type, const void *src, int64_t nb_total, int64_t blk0, float *dst, int64_t n, cudaStream_t st) {
    if (n == 0) return EB_OK;
    WPlanes w{(const uint8_t *)src, nb_total};
    switch (type) {
        case EB_F32: EB_CUDA(cudaMemcpyAsync(dst, (const float *)src + blk0, n * 4, cudaMemcpyDeviceToDevice, st)); return EB_OK;
        case EB_F16: dequant_f16_kernel<<<cdiv(n, 256), 256, 0, st>>>((const __half *)src + blk0, dst, n); break;
        case EB_Q4_0: if (n % 32) return eb_set_err(EB_ERR_ARG, "q4_0: n %% 32"); w.off(blk0); dequant_q4_0_kernel<<<cdiv(n / 32, 16), 256, 0, st>>>(w, dst, n / 32); break;
        case EB_Q8_0: if (n % 32) return eb_set_err(EB_ERR_ARG, "q8_0: n %% 32"); w.off(blk0); dequant_q8_0_kernel<<<cdiv(n / 32, 8), 256, 0, st>>>(w, dst, n / 32); break;
        case EB_Q4_K: if (n % 256) return eb_set_err(EB_ERR_ARG, "q4_K: n %% 256"); dequant_q4_K_kernel<<<n / 256, 64, 0, st>>>((const block_q4_K *)src + blk0, dst, n / 256); break;
        case EB_Q6_K: if (n % 256) return eb_set_err(EB_ERR_ARG, "q6_K: n %% 256"); w.off(blk0); dequant_q6_K_kernel<<<n / 256, 64, 0, st>>>(w, dst, n / 256); break;
        default: return eb_set_err(EB_ERR_UNSUPPORTED, "dequantize: type %d", type);
    }
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ GGUF <-> device layout
// Lossless permutation of the block bytes into aligned planes (Q4_K is already 16-B aligned and
// stays as is).  dir = 0: GGUF -> device, 1: device -> GGUF.  One thread per block.
__global__ void repack_kernel(int type, uint8_t *gguf, uint8_t *dev, int64_t nb, int dir) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb) return;
    WPlanes w{dev, nb};
    auto mv = [&](uint8_t *g, uint8_t *d, int n) { for (int k = 0; k < n; k++) { if (dir == 0) d[k] = g[k]; else g[k] = d[k]; } };
    if (type == EB_Q4_0) { uint8_t *g = gguf + i * 18; mv(g, (uint8_t *)(w.q40_d() + i), 2); mv(g + 2, (uint8_t *)w.q40_qs() + i * 16, 16); }
    else if (type == EB_Q8_0) { uint8_t *g = gguf + i * 34; mv(g, (uint8_t *)(w.q80_d() + i), 2); mv(g + 2, (uint8_t *)w.q80_qs() + i * 32, 32); }
    else if (type == EB_Q6_K) { uint8_t *g = gguf + i * 210; mv(g, (uint8_t *)w.q6k_ql() + i * 128, 128); mv(g + 128, (uint8_t *)w.q6k_qh() + i * 64, 64);
                                mv(g + 192, (uint8_t *)w.q6k_sc() + i * 16, 16); mv(g + 208, (uint8_t *)(w.q6k_d() + i), 2); }
}
int eb_repack_impl(int type, void *gguf, void *dev, int64_t nb, int dir, cudaStream_t st) {
    if (type == EB_Q4_0 || type == EB_Q8_0 || type == EB_Q6_K) {
        repack_kernel<<<cdiv(nb, 128), 128, 0, st>>>(type, (uint8_t *)gguf, (uint8_t *)dev, nb, dir);
        EB_CHECK_LAUNCH();
        return EB_OK;
    }
    return eb_set_err(EB_ERR_ARG, "repack: type %d has no device layout", type);
}

// ------------------------------------------------------------------ activation quantisers (K12)
__global__ void quant_act_kernel(const float *__restrict__ x, ActQ a) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (a.kq) {
        const int64_t nb = a.K / 256, total = nb * a.T;
        if (warp >= total) return;
        const int64_t t = warp / nb, b = warp % nb;
        quant_block_q8K(x + warp * 256, a.qs + warp * 256, a.d + warp, a.bsum + warp * 16, lane,
                        a.h16 ? a.h16 + warp * 256 : nullptr, (a.dT && t < EB_DT_PITCH) ? a.dT + b * EB_DT_PITCH + t : nullptr);
    } else {
        const int64_t total = a.K / 128 * a.T;
        if (warp >= total) return;
        quant_block_q8_0x4(x + warp * 128, a.qs + warp * 128, a.d + warp * 4, lane);
    }
}

// EAGLE head input: row t = [token embedding (E) | parent feature (E)] (ggml_concat, R/src/llama.cpp:1880) quantised for the fc
// GEMM without materialising the concatenation: quantisation block b comes from `x` (b < E / block) or from `hid`
__global__ void concat_quant_kernel(const float *__restrict__ x, const float *__restrict__ hid, int E, ActQ a,
                                    const SpecState *__restrict__ spec, const float *__restrict__ root_feat, const float *__restrict__ node_feat,
                                    const int *__restrict__ gate) {
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int per = a.kq ? 256 : 128;
    const int64_t nb = a.K / per;
    if (warp >= nb * a.T) return;
    const int64_t t = warp / nb, b = warp % nb;
    const int64_t col = b * per;
    const float *src;
    if (col < E) src = x + t * E + col;
    else if (spec) {                                        // hidden row of draft row t: root <- target feature, others <- the head's own
        const int par = spec->dft_src[t];                   // result_norm of the parent node (R/src/llama.cpp:10170, R/src/llama-context.cpp:83-88)
        src = (par < 0 ? root_feat : node_feat + (int64_t)par * E) + (col - E);
    } else src = hid + t * E + (col - E);
    if (a.kq) quant_block_q8K(src, a.qs + warp * 256, a.d + warp, a.bsum + warp * 16, lane,
                              a.h16 ? a.h16 + warp * 256 : nullptr, (a.dT && t < EB_DT_PITCH) ? a.dT + b * EB_DT_PITCH + t : nullptr);
    else quant_block_q8_0x4(src, a.qs + warp * 128, a.d + warp * 4, lane);
}
int eb_concat_quant(const float *x, const float *hid, int E, const ActQ &a, cudaStream_t st, const SpecState *spec, const float *root_feat, const float *node_feat) {
    const int per = a.kq ? 256 : 128;
    if (a.K != 2 * (int64_t)E || E % per) return eb_set_err(EB_ERR_ARG, "concat_quant: E=%d K=%lld", E, (long long)a.K);
    const int64_t warps = a.K / per * a.T;
    if (warps == 0) return EB_OK;
    eb_launch_pdl_c(EB_PDL_CONCAT, concat_quant_kernel, dim3((unsigned)cdiv(warps, 8)), dim3(256), 0, st, x, hid, E, a, spec, root_feat, node_feat, eb_gate());
    EB_CHECK_LAUNCH();
    return EB_OK;
}

size_t eb_actq_bytes(int kq, int64_t K, int T) {
    size_t qs = (size_t)T * K, d = (size_t)T * (K / (kq ? 256 : 32)) * 4, bs = kq ? (size_t)T * (K / 16) * 2 : 0;
    size_t h16 = kq ? (size_t)T * K * 2 : 0, dT = kq ? (size_t)(K / 256) * EB_DT_PITCH * 4 : 0;
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    return al(qs) + al(d) + al(bs) + al(h16) + al(dT);
}
ActQ eb_actq_carve(void *ws, int kq, int64_t K, int T) {
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    ActQ a; char *p = (char *)ws;
    a.qs = (int8_t *)p; p += al((size_t)T * K);
    a.d = (float *)p; p += al((size_t)T * (K / (kq ? 256 : 32)) * 4);
    a.bsum = kq ? (__half *)p : nullptr; if (kq) p += al((size_t)T * (K / 16) * 2);
    a.h16 = kq ? (__half *)p : nullptr; if (kq) p += al((size_t)T * K * 2);
    a.dT = kq ? (float *)p : nullptr;
    a.K = K; a.T = T; a.kq = kq;
    return a;
}
int eb_quantize_act(const float *d_x, const ActQ &a, cudaStream_t st) {
    if (a.K % (a.kq ? 256 : 128)) return eb_set_err(EB_ERR_ARG, "quantize_act: K=%lld not a multiple of %d", (long long)a.K, a.kq ? 256 : 128);
    const int64_t warps = a.kq ? a.K / 256 * a.T : a.K / 128 * a.T;
    if (warps == 0) return EB_OK;
    quant_act_kernel<<<cdiv(warps, 8), 256, 0, st>>>(d_x, a);
    EB_CHECK_LAUNCH();
    return EB_OK;
}

// ------------------------------------------------------------------ synthetic tensors (bench fixtures)
__device__ __forceinline__ uint64_t splitmix(uint64_t &s) {
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ float u01(uint64_t &s) { return (splitmix(s) >> 40) * (1.0f / 16777216.0f); }
__device__ __forceinline__ float gauss(uint64_t &s) {
    const float u1 = fmaxf(u01(s), 1e-7f), u2 = u01(s);
    return sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
}
// Random but well-formed blocks whose dequantised values are roughly zero-mean with std ~scale.
// One thread per block.
__global__ void synth_kernel(int type, void *dst, int64_t nb, uint64_t seed, float scale) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb) return;
    uint64_t s = seed * 0x100000001b3ull + (uint64_t)i * 0x9e3779b97f4a7c15ull;
    splitmix(s);
    if (type == EB_Q4_K) {
        block_q4_K *b = (block_q4_K *)dst + i;
        // sub-block scale sc in [24,63], min m chosen so that values are centred: d*sc*(q-7.5)
        const float amp = scale * (2.2f + 0.6f * u01(s));         // block half-range
        const float d = amp / (7.5f * 48.f);
        b->d = __float2half_rn(d); b->dmin = __float2half_rn(d * 7.5f);
        uint8_t sc[8], m[8];
        for (int j = 0; j < 8; j++) { sc[j] = 24 + (splitmix(s) % 40); m[j] = sc[j]; }
        for (int j = 0; j < 4; j++) {
            b->scales[j] = (sc[j] & 63) | ((sc[j + 4] >> 4) << 6);
            b->scales[j + 4] = (m[j] & 63) | ((m[j + 4] >> 4) << 6);
            b->scales[j + 8] = (sc[j + 4] & 0xF) | ((m[j + 4] & 0xF) << 4);
        }
        for (int j = 0; j < 128; j += 8) { uint64_t r = splitmix(s); for (int k = 0; k < 8; k++) b->qs[j + k] = (uint8_t)(r >> (8 * k)); }
    } else if (type == EB_Q6_K) {
        block_q6_K *b = (block_q6_K *)dst + i;
        const float amp = scale * (2.2f + 0.6f * u01(s));
        b->d = __float2half_rn(amp / (32.f * 96.f));
        for (int j = 0; j < 16; j++) b->scales[j] = (int8_t)(64 + (splitmix(s) % 64));
        for (int j = 0; j < 128; j += 8) { uint64_t r = splitmix(s); for (int k = 0; k < 8; k++) b->ql[j + k] = (uint8_t)(r >> (8 * k)); }
        for (int j = 0; j < 64; j += 8) { uint64_t r = splitmix(s); for (int k = 0; k < 8; k++) b->qh[j + k] = (uint8_t)(r >> (8 * k)); }
    } else if (type == EB_Q4_0) {
        block_q4_0 *b = (block_q4_0 *)dst + i;
        b->d = __float2half_rn(scale * (2.2f + 0.6f * u01(s)) / 8.f);
        for (int j = 0; j < 16; j += 8) { uint64_t r = splitmix(s); for (int k = 0; k < 8; k++) b->qs[j + k] = (uint8_t)(r >> (8 * k)); }
    } else if (type == EB_Q8_0) {
        block_q8_0 *b = (block_q8_0 *)dst + i;
        b->d = __float2half_rn(scale * (2.2f + 0.6f * u01(s)) / 127.f);
        for (int j = 0; j < 32; j += 8) { uint64_t r = splitmix(s); for (int k = 0; k < 8; k++) b->qs[j + k] = (int8_t)(r >> (8 * k)); }
    } else if (type == EB_F32) {
        ((float *)dst)[i] = scale * gauss(s);
    } else if (type == EB_F16) {
        ((__half *)dst)[i] = __float2half_rn(scale * gauss(s));
    }
}
// norm weights: 1 + N(0, scale^2)
__global__ void synth_norm_kernel(float *dst, int64_t n, uint64_t seed, float scale) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t s = seed * 0x100000001b3ull + (uint64_t)i * 0x9e3779b97f4a7c15ull;
    splitmix(s);
    dst[i] = 1.f + scale * gauss(s);
}
int eb_synth_impl(int type, void *dst, int64_t n_elems, uint64_t seed, float scale, int norm_weight, cudaStream_t st) {
    if (norm_weight) { synth_norm_kernel<<<cdiv(n_elems, 256), 256, 0, st>>>((float *)dst, n_elems, seed, scale); EB_CHECK_LAUNCH(); return EB_OK; }
    int be = (type == EB_Q4_K || type == EB_Q6_K) ? 256 : (type == EB_Q4_0 || type == EB_Q8_0) ? 32 : 1;
    const int64_t nb = n_elems / be;
    synth_kernel<<<cdiv(nb, 256), 256, 0, st>>>(type, dst, nb, seed, scale);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
