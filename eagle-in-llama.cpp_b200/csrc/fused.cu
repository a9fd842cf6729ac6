// fused.cu — engine-only fusions of the glue between the GEMMs and attention.  Every kernel reads
// the split-K partial planes of the producing GEMM directly (fixed summation order -> results
// are deterministic), so no stand-alone reduce / add / norm / quantise launches remain:
//   resid_norm_quant : x += sum_s part[s];  cur = rms_norm(x) * w;  cur -> activation grid
//                      (ADD -> RMS_NORM -> MUL -> from_float of R/src/llama.cpp:1763-1790)
//   rope_kv_store    : q,k,v = sum_s part[s]; RoPE(q,k); K,V -> f16 cache rows (:1710-1757, :228-270)
//   swiglu_quant     : silu(sum gate) * sum up -> activation grid (llm_build_ffn :456-581)
// All are one pass over [T][E|F] f32 data that lives in L2; coalesced float4 accesses.
#include "common.cuh"
#include "quant_dev.cuh"
#include "spec.cuh"
#include "engine.cuh"

struct Parts { const float *p; int ks; int64_t plane; int64_t ld; };   // plane = T*ld elements between split planes, ld = row pitch (a fused GEMM's N)
static Parts parts_of(const PlaneRef &r, int T) { return Parts{r.p, r.ks, (int64_t)T * r.ld, r.ld}; }

// sum of the split-K planes in plane order.  The loads of four planes are issued together: a rolled loop pays one L2 round
// trip per plane (5 planes = 3 us in a kernel whose whole job is 65 rows)
__device__ __forceinline__ float4 sum_parts(const Parts &s, int64_t off) {
    const float *p = s.p + off;
    float4 b[4];
#pragma unroll
    for (int k = 0; k < 4; k++) if (k == 0 || k < s.ks) b[k] = *(const float4 *)(p + k * s.plane);
    float4 a = b[0];
#pragma unroll
    for (int k = 1; k < 4; k++) if (k < s.ks) { a.x += b[k].x; a.y += b[k].y; a.z += b[k].z; a.w += b[k].w; }
    if (s.ks > 4) {
#pragma unroll
        for (int k = 0; k < 4; k++) if (k + 4 < s.ks) b[k] = *(const float4 *)(p + (k + 4) * s.plane);
#pragma unroll
        for (int k = 0; k < 4; k++) if (k + 4 < s.ks) { a.x += b[k].x; a.y += b[k].y; a.z += b[k].z; a.w += b[k].w; }
    }
    return a;
}

// tensor parallel (common.cuh TpPull): eight consecutive elements of the planes one peer pushed into this rank's memory, summed in
// plane order; an element has arrived when its upper word is the call number (the spin gives up after ~2 s and raises the error word)
__device__ __forceinline__ void pull_sum8(const unsigned long long *src, int64_t hdr, int64_t plane, int64_t off, unsigned call, unsigned *err, float4 &s0, float4 &s1) {
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int ks = 0;
    {                                                   // the peer's plane count of this call
        const long long t0 = clock64();
        for (;;) {
            unsigned long long h; asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(h) : "l"(src + hdr) : "memory");
            if ((unsigned)(h >> 32) == call) { ks = (int)(unsigned)h; break; }
            if (clock64() - t0 > 4000000000ll) { atomicExch(err, 1u); break; }
        }
    }
    for (int k = 0; k < ks; k++) {
        const unsigned long long *q = src + (int64_t)k * plane + off;
        uint4 u[4];
        long long t0 = 0; bool timing = false;
        for (;;) {
#pragma unroll
            for (int i = 0; i < 4; i++) asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(u[i].x), "=r"(u[i].y), "=r"(u[i].z), "=r"(u[i].w) : "l"(q + 2 * i) : "memory");
            bool ok = true;
#pragma unroll
            for (int i = 0; i < 4; i++) ok = ok && u[i].y == call && u[i].w == call;
            if (ok) break;
            if (!timing) { timing = true; t0 = clock64(); }
            else if (clock64() - t0 > 4000000000ll) { atomicExch(err, 1u); break; }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float a = __uint_as_float(u[i].x), b = __uint_as_float(u[i].z);
            if (k == 0) { v[2 * i] = a; v[2 * i + 1] = b; } else { v[2 * i] += a; v[2 * i + 1] += b; }
        }
    }
    s0 = make_float4(v[0], v[1], v[2], v[3]); s1 = make_float4(v[4], v[5], v[6], v[7]);
}

// one CTA (256 threads) per token row
// bias != null: x = relu(sum_s part[s] + bias) instead of x += sum_s part[s]  (the EAGLE head's fc + bias + ReLU, R/src/llama.cpp:1886-1890)
__global__ void __launch_bounds__(256) resid_norm_quant_kernel(float *__restrict__ x, Parts add, const float *__restrict__ w, float eps,
                                                              int n, ActQ aq, float *__restrict__ feat, const float *__restrict__ bias) {
    extern __shared__ float row[];
    __shared__ double red[8];
    __shared__ float s_scale;
    const int r = blockIdx.x, tid = threadIdx.x;
    float *xr = x + (int64_t)r * n;
    double acc = 0.0;
    for (int i = tid * 4; i < n; i += 1024) {
        float4 v = bias ? make_float4(0.f, 0.f, 0.f, 0.f) : *(const float4 *)(xr + i);
        if (add.p) {                                       // ffn_inp = cur + inpSA (ggml_add(cur, inpSA))
            const float4 c = sum_parts(add, (int64_t)r * n + i);
            if (bias) {
                const float4 b = *(const float4 *)(bias + i);
                v = make_float4(fmaxf(__fadd_rn(c.x, b.x), 0.f), fmaxf(__fadd_rn(c.y, b.y), 0.f), fmaxf(__fadd_rn(c.z, b.z), 0.f), fmaxf(__fadd_rn(c.w, b.w), 0.f));
            } else {
                v.x = __fadd_rn(c.x, v.x); v.y = __fadd_rn(c.y, v.y); v.z = __fadd_rn(c.z, v.z); v.w = __fadd_rn(c.w, v.w);
            }
            *(float4 *)(xr + i) = v;
        }
        *(float4 *)(row + i) = v;
        acc += (double)(v.x * v.x); acc += (double)(v.y * v.y); acc += (double)(v.z * v.z); acc += (double)(v.w * v.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((tid & 31) == 0) red[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        double s = 0; for (int i = 0; i < 8; i++) s += red[i];
        s_scale = 1.0f / sqrtf((float)(s / n) + eps);
    }
    __syncthreads();
    const float sc = s_scale;
    for (int i = tid * 4; i < n; i += 1024) {
        float4 v = *(float4 *)(row + i);
        v.x = __fmul_rn(v.x, sc); v.y = __fmul_rn(v.y, sc); v.z = __fmul_rn(v.z, sc); v.w = __fmul_rn(v.w, sc);
        if (w) { const float4 g = *(const float4 *)(w + i); v.x = __fmul_rn(v.x, g.x); v.y = __fmul_rn(v.y, g.y); v.z = __fmul_rn(v.z, g.z); v.w = __fmul_rn(v.w, g.w); }
        *(float4 *)(row + i) = v;
        if (feat) *(float4 *)(feat + (int64_t)r * n + i) = v;
    }
    __syncthreads();
    if (aq.qs) quant_row_smem(row, aq, r, tid, 256);
}

// Q8_K-grid rows of n = 8 * blockDim.x <= 8192 values: thread t owns elements 8t..8t+7 in registers for the whole kernel
// (warp w = quantisation block w), one __syncthreads for the row sum, no shared-memory staging of the row
template <int MAXT>
__global__ void __launch_bounds__(MAXT) resid_norm_quant_regs_kernel(float *__restrict__ x, Parts add, const float *__restrict__ w, float eps,
                                                                    int n, ActQ a, float *__restrict__ feat, const float *__restrict__ bias,
                                                                    const SpecState *__restrict__ spec, float *__restrict__ node_feat, const int *__restrict__ gate, TpPull pull) {
    __shared__ double red[32];
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    const int64_t off = (int64_t)r * n + tid * 8;
    float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
    if (!bias) { v0 = *(const float4 *)(x + off); v1 = *(const float4 *)(x + off + 4); }
    if (add.p) {
        float4 c0 = sum_parts(add, off), c1 = sum_parts(add, off + 4);
        if (pull.n_peers) {
            // tensor parallel: this rank's planes (above) and every peer's, added IN RANK ORDER so that all ranks produce the same bits
            const unsigned call = *(volatile const unsigned *)pull.seq + 1u;
            float4 t0 = c0, t1 = c1;
            for (int rr = 0; rr <= pull.n_peers; rr++) {
                float4 s0 = c0, s1 = c1;
                if (rr != pull.my_pos) pull_sum8(pull.src[rr < pull.my_pos ? rr : rr - 1], pull.hdr, pull.plane, off, call, pull.err, s0, s1);
                if (rr == 0) { t0 = s0; t1 = s1; }
                else { t0.x += s0.x; t0.y += s0.y; t0.z += s0.z; t0.w += s0.w; t1.x += s1.x; t1.y += s1.y; t1.z += s1.z; t1.w += s1.w; }
            }
            c0 = t0; c1 = t1;
            // the last CTA to get here closes the call (the next GEMM on this slot set is a whole layer away)
            __syncthreads();
            if (tid == 0 && atomicAdd(pull.done, 1u) == gridDim.x - 1) { *pull.done = 0u; *(volatile unsigned *)pull.seq = call; }
        }
        if (bias) {
            const float4 b0 = *(const float4 *)(bias + tid * 8), b1 = *(const float4 *)(bias + tid * 8 + 4);
            v0 = make_float4(fmaxf(__fadd_rn(c0.x, b0.x), 0.f), fmaxf(__fadd_rn(c0.y, b0.y), 0.f), fmaxf(__fadd_rn(c0.z, b0.z), 0.f), fmaxf(__fadd_rn(c0.w, b0.w), 0.f));
            v1 = make_float4(fmaxf(__fadd_rn(c1.x, b1.x), 0.f), fmaxf(__fadd_rn(c1.y, b1.y), 0.f), fmaxf(__fadd_rn(c1.z, b1.z), 0.f), fmaxf(__fadd_rn(c1.w, b1.w), 0.f));
        } else {
            v0.x = __fadd_rn(c0.x, v0.x); v0.y = __fadd_rn(c0.y, v0.y); v0.z = __fadd_rn(c0.z, v0.z); v0.w = __fadd_rn(c0.w, v0.w);
            v1.x = __fadd_rn(c1.x, v1.x); v1.y = __fadd_rn(c1.y, v1.y); v1.z = __fadd_rn(c1.z, v1.z); v1.w = __fadd_rn(c1.w, v1.w);
        }
        *(float4 *)(x + off) = v0; *(float4 *)(x + off + 4) = v1;
    }
    float4 g0 = make_float4(1.f, 1.f, 1.f, 1.f), g1 = g0;
    if (w) { g0 = *(const float4 *)(w + tid * 8); g1 = *(const float4 *)(w + tid * 8 + 4); }
    double acc = (double)(v0.x * v0.x); acc += (double)(v0.y * v0.y); acc += (double)(v0.z * v0.z); acc += (double)(v0.w * v0.w);
    acc += (double)(v1.x * v1.x); acc += (double)(v1.y * v1.y); acc += (double)(v1.z * v1.z); acc += (double)(v1.w * v1.w);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    double s = 0;
    for (int i = 0; i < nw; i++) s += red[i];
    const float sc = 1.0f / sqrtf((float)(s / n) + eps);
    float v[8] = {__fmul_rn(v0.x, sc), __fmul_rn(v0.y, sc), __fmul_rn(v0.z, sc), __fmul_rn(v0.w, sc),
                  __fmul_rn(v1.x, sc), __fmul_rn(v1.y, sc), __fmul_rn(v1.z, sc), __fmul_rn(v1.w, sc)};
    if (w) {
        v[0] = __fmul_rn(v[0], g0.x); v[1] = __fmul_rn(v[1], g0.y); v[2] = __fmul_rn(v[2], g0.z); v[3] = __fmul_rn(v[3], g0.w);
        v[4] = __fmul_rn(v[4], g1.x); v[5] = __fmul_rn(v[5], g1.y); v[6] = __fmul_rn(v[6], g1.z); v[7] = __fmul_rn(v[7], g1.w);
    }
    if (feat) { *(float4 *)(feat + off) = make_float4(v[0], v[1], v[2], v[3]); *(float4 *)(feat + off + 4) = make_float4(v[4], v[5], v[6], v[7]); }
    if (spec && r < spec->dft_n) {                          // draft level: the row is also the node's feature (input of its children)
        float *nf = node_feat + (int64_t)(spec->dft_node_base + r) * n + tid * 8;
        *(float4 *)nf = make_float4(v[0], v[1], v[2], v[3]); *(float4 *)(nf + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (a.qs) {
        const int nb = n / 256;
        quant_regs_q8K(v, a.qs + (int64_t)r * n + warp * 256, a.d + (int64_t)r * nb + warp, a.bsum + ((int64_t)r * nb + warp) * 16, lane,
                       a.h16 ? a.h16 + (int64_t)r * n + warp * 256 : nullptr, (a.dT && r < EB_DT_PITCH) ? a.dT + (int64_t)warp * EB_DT_PITCH + r : nullptr);
    }
}

// cos/sin of every (token, rotation pair), once per forward instead of once per layer: the reference's sequential theta
// recurrence (theta *= theta_scale, R/ggml/src/ggml-cpu/ggml-cpu.c rope_yarn / ggml_rope_cache_init)
__global__ void rope_table_kernel(const int32_t *__restrict__ pos, float theta_scale, int half, float2 *__restrict__ tab) {
    const int t = blockIdx.x, i = threadIdx.x;
    if (i >= half) return;
    float theta = (float)pos[t];
    for (int j = 0; j < i; j++) theta = __fmul_rn(theta, theta_scale);
    tab[t * half + i] = make_float2(cosf(theta), sinf(theta));
}
// grid (T, E/2048): q, k, v = sums of the split-K planes of the (possibly fused) QKV GEMM; RoPE on q and k; K/V rows go
// straight to their cache cells; q additionally as f16 for the tcgen05 attention
__global__ void __launch_bounds__(512) rope_kv_store_fused_kernel(Parts qp, Parts kp, Parts vp, float *__restrict__ q, __half *__restrict__ q16,
                                                                 const float2 *__restrict__ tab, __half *__restrict__ kc, __half *__restrict__ vc,
                                                                 const FwdCtl *__restrict__ ctl, int D, int n_head,
                                                                 const int32_t *__restrict__ pos, float theta_scale, const int *__restrict__ gate) {
    pdl_trigger(); pdl_wait();
    if (gate_closed(gate)) return;
    const int t = blockIdx.x, tid = threadIdx.x, half = D / 2;
    const int64_t cell0 = ctl ? (int64_t)(ctl->n_prefix + ctl->node_base) * n_head * D : 0;
    const int E = n_head * D;
    // float4 = two rotation pairs
    for (int e = (blockIdx.y * blockDim.x + tid) * 4; e < E; e += gridDim.y * blockDim.x * 4) {
        const int i = (e % D) / 2;
        float4 cs;                                                   // (cos0, sin0, cos1, sin1)
        if (tab) cs = *(const float4 *)(tab + t * half + i);
        else {                                                       // single-layer model (the EAGLE head): no table pass, same recurrence
            float theta = (float)pos[t];
            for (int j = 0; j < i; j++) theta = __fmul_rn(theta, theta_scale);
            const float th1 = __fmul_rn(theta, theta_scale);
            cs = make_float4(cosf(theta), sinf(theta), cosf(th1), sinf(th1));
        }
        const float c0 = cs.x, s0 = cs.y, c1 = cs.z, s1 = cs.w;
        const int64_t o = (int64_t)t * E + e;
        const float4 a = sum_parts(qp, t * qp.ld + e), b = sum_parts(kp, t * kp.ld + e), v = sum_parts(vp, t * vp.ld + e);
        const float4 qr = make_float4(__fsub_rn(__fmul_rn(a.x, c0), __fmul_rn(a.y, s0)), __fadd_rn(__fmul_rn(a.x, s0), __fmul_rn(a.y, c0)),
                                      __fsub_rn(__fmul_rn(a.z, c1), __fmul_rn(a.w, s1)), __fadd_rn(__fmul_rn(a.z, s1), __fmul_rn(a.w, c1)));
        *(float4 *)(q + o) = qr;
        if (q16) {                                          // the attention kernel's f16 queries (GGML_FP32_TO_FP16 of ggml-cpu.c:11115)
            const __half2 q0 = __floats2half2_rn(qr.x, qr.y), q1 = __floats2half2_rn(qr.z, qr.w);
            *(uint2 *)(q16 + o) = make_uint2(*(const uint32_t *)&q0, *(const uint32_t *)&q1);
        }
        const __half2 k0 = __floats2half2_rn(__fsub_rn(__fmul_rn(b.x, c0), __fmul_rn(b.y, s0)), __fadd_rn(__fmul_rn(b.x, s0), __fmul_rn(b.y, c0)));
        const __half2 k1 = __floats2half2_rn(__fsub_rn(__fmul_rn(b.z, c1), __fmul_rn(b.w, s1)), __fadd_rn(__fmul_rn(b.z, s1), __fmul_rn(b.w, c1)));
        const __half2 v0 = __floats2half2_rn(v.x, v.y), v1 = __floats2half2_rn(v.z, v.w);
        *(uint2 *)(kc + cell0 + o) = make_uint2(*(const uint32_t *)&k0, *(const uint32_t *)&k1);
        *(uint2 *)(vc + cell0 + o) = make_uint2(*(const uint32_t *)&v0, *(const uint32_t *)&v1);
    }
}

__device__ __forceinline__ float silu_ref(float x) { return x / (1.0f + expf(-x)); }   // ggml_silu_f32
// one warp per quantisation block (256 values on the Q8_K grid, 4 x 32 on the Q8_0 grid): the blocks of a row are
// independent, so the grid is (blocks, T) instead of one CTA per row — 65 x 43 warps for the 7B ffn instead of 65 CTAs
__device__ __forceinline__ float4 swiglu4(const Parts &gate, const Parts &up, int r, int i) {
    const float4 g = sum_parts(gate, r * gate.ld + i), u = sum_parts(up, r * up.ld + i);
    return make_float4(__fmul_rn(silu_ref(g.x), u.x), __fmul_rn(silu_ref(g.y), u.y), __fmul_rn(silu_ref(g.z), u.z), __fmul_rn(silu_ref(g.w), u.w));
}
__global__ void __launch_bounds__(256) swiglu_quant_kernel(Parts gate, Parts up, int n, ActQ a, const int *__restrict__ lgate) {
    pdl_trigger(); pdl_wait();
    if (gate_closed(lgate)) return;
    const int r = blockIdx.y, lane = threadIdx.x & 31, b = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int64_t row0 = (int64_t)r * n;
    if (a.kq) {
        const int nb = n / 256;
        if (b >= nb) return;
        const float4 x = swiglu4(gate, up, r, b * 256 + lane * 8), y = swiglu4(gate, up, r, b * 256 + lane * 8 + 4);
        const float v[8] = {x.x, x.y, x.z, x.w, y.x, y.y, y.z, y.w};
        quant_regs_q8K(v, a.qs + row0 + b * 256, a.d + (int64_t)r * nb + b, a.bsum + ((int64_t)r * nb + b) * 16, lane,
                       a.h16 ? a.h16 + row0 + b * 256 : nullptr, (a.dT && r < EB_DT_PITCH) ? a.dT + (int64_t)b * EB_DT_PITCH + r : nullptr);
    } else {
        const int ng = n / 128;
        if (b >= ng) return;
        quant_regs_q8_0x4(swiglu4(gate, up, r, b * 128 + lane * 4), a.qs + row0 + b * 128, a.d + (int64_t)r * (n / 32) + b * 4, lane);
    }
}

// plain row quantisation of an f32 [T][K] tensor that is already final (attention output)
int eb_resid_norm_quant(float *x, const float *part, int ks, const float *w, float eps, int n, int T, const ActQ &aq, float *feat, cudaStream_t st, const float *bias,
                        const SpecState *spec, float *node_feat, const TpPull *pull) {
    TpPull pl{}; if (pull) pl = *pull;
    if (bias && !part) return eb_set_err(EB_ERR_ARG, "resid_norm_quant: bias without planes");
    if (n % 4 || n > 16384) return eb_set_err(EB_ERR_ARG, "resid_norm_quant: n=%d", n);
    if (T == 0) return EB_OK;
    Parts p{part, ks, (int64_t)T * n, n};
    if (n % 256 == 0 && n <= 8192 && (!aq.qs || (aq.kq && aq.K == n)))
        if (n <= 4096) eb_launch_pdl_c(EB_PDL_RESID, resid_norm_quant_regs_kernel<512>, dim3(T), dim3(n / 8), 0, st, x, p, w, eps, n, aq, feat, bias, spec, node_feat, eb_gate(), pl);
        else eb_launch_pdl_c(EB_PDL_RESID, resid_norm_quant_regs_kernel<1024>, dim3(T), dim3(n / 8), 0, st, x, p, w, eps, n, aq, feat, bias, spec, node_feat, eb_gate(), pl);
    else {
        if (pl.n_peers) return eb_set_err(EB_ERR_UNSUPPORTED, "resid_norm_quant: the tensor-parallel hand-over needs the register kernel");
        if (spec) return eb_set_err(EB_ERR_UNSUPPORTED, "resid_norm_quant: node-feature scatter needs the register kernel (n %% 256 == 0, n <= 8192, K-quant grid)");
        resid_norm_quant_kernel<<<T, 256, (size_t)n * 4, st>>>(x, p, w, eps, n, aq, feat, bias);
    }
    EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_rope_table(const int32_t *pos, int D, int T, float freq_base, float *tab, cudaStream_t st) {
    if (T == 0) return EB_OK;
    if (D > 256 || D % 4) return eb_set_err(EB_ERR_ARG, "rope: D=%d", D);
    rope_table_kernel<<<T, 128, 0, st>>>(pos, powf(freq_base, -2.0f / D), D / 2, (float2 *)tab);
    EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_rope_kv_store_fused(PlaneRef qp, PlaneRef kp, PlaneRef vp, float *q, void *q16, const float *tab, void *kc, void *vc,
                           const FwdCtl *ctl, int D, int n_head, int T, cudaStream_t st, const int32_t *pos, float freq_base) {
    if (!tab && !pos) return eb_set_err(EB_ERR_ARG, "rope: neither table nor positions");
    if (T == 0) return EB_OK;
    if (D > 256 || D % 4) return eb_set_err(EB_ERR_ARG, "rope: D=%d", D);
    const int E = n_head * D;
    const int threads = 512, gy = (int)cdiv(E / 4, threads);
    eb_launch_pdl_c(EB_PDL_ROPE, rope_kv_store_fused_kernel, dim3(T, gy), dim3(threads), 0, st, parts_of(qp, T), parts_of(kp, T), parts_of(vp, T), q, (__half *)q16, (const float2 *)tab,
                  (__half *)kc, (__half *)vc, ctl, D, n_head, pos, powf(freq_base, -2.0f / D), eb_gate());
    EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_swiglu_quant(PlaneRef gate, PlaneRef up, int n, int T, const ActQ &aq, cudaStream_t st) {
    if (T == 0) return EB_OK;
    const int per = aq.kq ? 256 : 128;
    if (n % per || aq.K != n) return eb_set_err(EB_ERR_ARG, "swiglu: n=%d", n);
    eb_launch_pdl_c(EB_PDL_SWIGLU, swiglu_quant_kernel, dim3((unsigned)cdiv(n / per, 8), T), dim3(256), 0, st, parts_of(gate, T), parts_of(up, T), n, aq, eb_gate());
    EB_CHECK_LAUNCH();
    return EB_OK;
}
int eb_fused_init() {
    EB_CUDA(cudaFuncSetAttribute(resid_norm_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    return EB_OK;
}
