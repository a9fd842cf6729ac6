// mmvq.cu — K1: dequant-dot matvec for the draft head and small batches (T <= 8 per pass).
// Replaces mul_mat_vec_q (R/ggml/src/ggml-cuda/mmvq.cu:49-200).  One warp per weight row,
// 128-bit aligned loads of the weight planes straight from HBM (each byte is read once),
// int8 dp4a against activations quantised on the CPU backend's grid, warp-shuffle reduction.
// Integer sums are identical to ggml_vec_dot_*_q8_* (R/ggml/src/ggml-cpu/ggml-cpu-quants.c:1912,
// 3663,6266,7903); only the order of the f32 additions differs.
#include "common.cuh"

__device__ __forceinline__ int4 ldg_stream(const void *p) {      // weights: read once, keep out of L1
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ int dot16(const int4 &w, const int4 &a, int acc) {
    acc = __dp4a(w.x, a.x, acc); acc = __dp4a(w.y, a.y, acc); acc = __dp4a(w.z, a.z, acc); return __dp4a(w.w, a.w, acc);
}
#define AND4(v, m) make_int4((v).x & (m), (v).y & (m), (v).z & (m), (v).w & (m))
#define SHR4(v, s) make_int4((int)((uint32_t)(v).x >> (s)), (int)((uint32_t)(v).y >> (s)), (int)((uint32_t)(v).z >> (s)), (int)((uint32_t)(v).w >> (s)))

// 6-bit scale / min of sub-block j from the three scale words of a Q4_K header held in registers (q4k_scale_min on bytes would put
// the header into local memory for the dynamic index)
__device__ __forceinline__ void q4k_scale_min_reg(int j, uint32_t w0, uint32_t w1, uint32_t w2, int &sc, int &m) {
    const int sh = 8 * (j & 3);
    const uint32_t b0 = (w0 >> sh) & 0xFFu, b1 = (w1 >> sh) & 0xFFu, b2 = (w2 >> sh) & 0xFFu;
    if (j < 4) { sc = (int)(b0 & 63u); m = (int)(b1 & 63u); }
    else { sc = (int)((b2 & 0xFu) | ((b0 >> 6) << 4)); m = (int)((b2 >> 4) | ((b1 >> 6) << 4)); }
}

// K-quant rows: 8 lanes per super-block, 4 super-blocks per warp step, and 4 steps (16 super-blocks = one 4096-wide row) requested
// before the first is used, so every lane has eight 16-byte weight loads in flight instead of two
template <int NC>
__global__ void __launch_bounds__(128) mmvq_q4_K(const block_q4_K *__restrict__ W, int64_t N, int nbr, ActQ a, int t0, float *__restrict__ y, const int *gate) {
    if (gate_closed(gate)) return;
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= N) return;
    const block_q4_K *wr = W + row * nbr;
    const int g = lane >> 3, c = lane & 7, j = c >> 1;
    float acc[NC];
#pragma unroll
    for (int t = 0; t < NC; t++) acc[t] = 0.f;
    for (int sbb = 0; sbb < nbr; sbb += 16) {
        int4 hdr[4], q[4];
        int4 alo1[4], ahi1[4]; float dx1[4]; __half2 bs1[4];          // NC == 1: the column's operands, requested with the weights
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int sb = sbb + 4 * u + g;
            hdr[u] = make_int4(0, 0, 0, 0); q[u] = make_int4(0, 0, 0, 0);
            if (sb < nbr) { hdr[u] = ldg_stream(&wr[sb]); q[u] = ldg_stream((const uint8_t *)&wr[sb] + 16 + 16 * c); }
            if constexpr (NC == 1) {
                const int sbc = min(sb, nbr - 1);                      // clamped: the loads need no predicate and can issue at once
                const int8_t *ap = a.qs + (int64_t)t0 * a.K + sbc * 256 + j * 64 + (c & 1) * 16;
                alo1[u] = *(const int4 *)ap; ahi1[u] = *(const int4 *)(ap + 32);
                dx1[u] = a.d[(int64_t)t0 * nbr + sbc];
                bs1[u] = *(const __half2 *)(a.bsum + ((int64_t)t0 * nbr + sbc) * 16 + 2 * c);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int sb = sbb + 4 * u + g;
            if (sb < nbr) {
                const __half2 dm = *(const __half2 *)&hdr[u].x;
                const float dw = __low2float(dm), dmin = __high2float(dm);
                int sc_lo, m_lo, sc_hi, m_hi, sc_c, m_c;
                q4k_scale_min_reg(2 * j, (uint32_t)hdr[u].y, (uint32_t)hdr[u].z, (uint32_t)hdr[u].w, sc_lo, m_lo);
                q4k_scale_min_reg(2 * j + 1, (uint32_t)hdr[u].y, (uint32_t)hdr[u].z, (uint32_t)hdr[u].w, sc_hi, m_hi);
                q4k_scale_min_reg(c, (uint32_t)hdr[u].y, (uint32_t)hdr[u].z, (uint32_t)hdr[u].w, sc_c, m_c);
                const int4 qlo = AND4(q[u], 0x0F0F0F0F), qhi = AND4(SHR4(q[u], 4), 0x0F0F0F0F);
                if constexpr (NC == 1) {
                    const int isum = dot16(qlo, alo1[u], 0) * sc_lo + dot16(qhi, ahi1[u], 0) * sc_hi;
                    const float bsum = __low2float(bs1[u]) + __high2float(bs1[u]);
                    acc[0] += (dw * dx1[u]) * (float)isum - (dmin * dx1[u]) * ((float)m_c * bsum);
                } else {
#pragma unroll
                    for (int t = 0; t < NC; t++) {
                        const int8_t *ap = a.qs + (int64_t)(t0 + t) * a.K + sb * 256 + j * 64 + (c & 1) * 16;
                        const int4 alo = *(const int4 *)ap, ahi = *(const int4 *)(ap + 32);
                        const int isum = dot16(qlo, alo, 0) * sc_lo + dot16(qhi, ahi, 0) * sc_hi;
                        const float dx = a.d[(int64_t)(t0 + t) * nbr + sb];
                        const __half2 bs = *(const __half2 *)(a.bsum + ((int64_t)(t0 + t) * nbr + sb) * 16 + 2 * c);
                        const float bsum = __low2float(bs) + __high2float(bs);
                        acc[t] += (dw * dx) * (float)isum - (dmin * dx) * ((float)m_c * bsum);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < NC; t++) { const float s = warp_sum(acc[t]); if (lane == 0) y[(int64_t)(t0 + t) * N + row] = s; }
}

template <int NC>
__global__ void __launch_bounds__(128) mmvq_q6_K(WPlanes w, int64_t N, int nbr, ActQ a, int t0, float *__restrict__ y, const int *gate) {
    if (gate_closed(gate)) return;
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= N) return;
    const int g = lane >> 3, c = lane & 7, n = c >> 2, cc = c & 3;
    const int shift = (cc >= 2) ? 2 : 0;
    float acc[NC];
#pragma unroll
    for (int t = 0; t < NC; t++) acc[t] = 0.f;
    for (int sbb = 0; sbb < nbr; sbb += 16) {
        int4 ql[4], qh[4]; int sc_lo[4], sc_hi[4]; float dw[4];
        int4 alo1[4], ahi1[4]; float dx1[4];                          // NC == 1: the column's operands, requested with the weights
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int sb = sbb + 4 * u + g;
            ql[u] = make_int4(0, 0, 0, 0); qh[u] = make_int4(0, 0, 0, 0); sc_lo[u] = sc_hi[u] = 0; dw[u] = 0.f;
            if constexpr (NC == 1) {
                const int sbc = min(sb, nbr - 1);
                const int8_t *ap = a.qs + (int64_t)t0 * a.K + sbc * 256 + n * 128 + 16 * cc;
                alo1[u] = *(const int4 *)ap; ahi1[u] = *(const int4 *)(ap + 64);
                dx1[u] = a.d[(int64_t)t0 * nbr + sbc];
            }
            if (sb < nbr) {
                const int64_t blk = row * nbr + sb;
                ql[u] = ldg_stream(w.q6k_ql() + blk * 128 + 16 * c);
                qh[u] = ldg_stream(w.q6k_qh() + blk * 64 + n * 32 + 16 * (cc & 1));
                const int8_t *sc = w.q6k_sc() + blk * 16 + 8 * n;
                sc_lo[u] = sc[cc]; sc_hi[u] = sc[4 + cc];
                dw[u] = __half2float(w.q6k_d()[blk]);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int sb = sbb + 4 * u + g;
            if (sb < nbr) {
                int4 qlo, qhi;
#define Q6(dst, lw, hw, sh, hsh) dst = __vsubss4((((uint32_t)(lw) >> (sh)) & 0x0F0F0F0Fu) | ((((uint32_t)(hw) >> (hsh)) & 0x03030303u) << 4), 0x20202020u)
                Q6(qlo.x, ql[u].x, qh[u].x, 0, shift); Q6(qlo.y, ql[u].y, qh[u].y, 0, shift); Q6(qlo.z, ql[u].z, qh[u].z, 0, shift); Q6(qlo.w, ql[u].w, qh[u].w, 0, shift);
                Q6(qhi.x, ql[u].x, qh[u].x, 4, shift + 4); Q6(qhi.y, ql[u].y, qh[u].y, 4, shift + 4); Q6(qhi.z, ql[u].z, qh[u].z, 4, shift + 4); Q6(qhi.w, ql[u].w, qh[u].w, 4, shift + 4);
#undef Q6
                if constexpr (NC == 1) {
                    const int isum = dot16(qlo, alo1[u], 0) * sc_lo[u] + dot16(qhi, ahi1[u], 0) * sc_hi[u];
                    acc[0] += (dw[u] * dx1[u]) * (float)isum;
                } else {
#pragma unroll
                    for (int t = 0; t < NC; t++) {
                        const int8_t *ap = a.qs + (int64_t)(t0 + t) * a.K + sb * 256 + n * 128 + 16 * cc;
                        const int4 alo = *(const int4 *)ap, ahi = *(const int4 *)(ap + 64);
                        const int isum = dot16(qlo, alo, 0) * sc_lo[u] + dot16(qhi, ahi, 0) * sc_hi[u];
                        acc[t] += (dw[u] * a.d[(int64_t)(t0 + t) * nbr + sb]) * (float)isum;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int t = 0; t < NC; t++) { const float s = warp_sum(acc[t]); if (lane == 0) y[(int64_t)(t0 + t) * N + row] = s; }
}

template <int NC>
__global__ void __launch_bounds__(128) mmvq_q4_0(WPlanes w, int64_t N, int nbr, ActQ a, int t0, float *__restrict__ y, const int *gate) {
    if (gate_closed(gate)) return;
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= N) return;
    float acc[NC];
#pragma unroll
    for (int t = 0; t < NC; t++) acc[t] = 0.f;
#pragma unroll 4
    for (int b = lane; b < nbr; b += 32) {
        const int64_t blk = row * nbr + b;
        const int4 q = ldg_stream(w.q40_qs() + blk * 16);
        const float dw = __half2float(w.q40_d()[blk]);
        int4 qlo = AND4(q, 0x0F0F0F0F), qhi = AND4(SHR4(q, 4), 0x0F0F0F0F);
        qlo = make_int4(__vsubss4(qlo.x, 0x08080808u), __vsubss4(qlo.y, 0x08080808u), __vsubss4(qlo.z, 0x08080808u), __vsubss4(qlo.w, 0x08080808u));
        qhi = make_int4(__vsubss4(qhi.x, 0x08080808u), __vsubss4(qhi.y, 0x08080808u), __vsubss4(qhi.z, 0x08080808u), __vsubss4(qhi.w, 0x08080808u));
#pragma unroll
        for (int t = 0; t < NC; t++) {
            const int8_t *ap = a.qs + (int64_t)(t0 + t) * a.K + b * 32;
            const int isum = dot16(qhi, *(const int4 *)(ap + 16), dot16(qlo, *(const int4 *)ap, 0));
            acc[t] += (float)isum * dw * a.d[(int64_t)(t0 + t) * nbr + b];
        }
    }
#pragma unroll
    for (int t = 0; t < NC; t++) { const float s = warp_sum(acc[t]); if (lane == 0) y[(int64_t)(t0 + t) * N + row] = s; }
}

template <int NC>
__global__ void __launch_bounds__(128) mmvq_q8_0(WPlanes w, int64_t N, int nbr, ActQ a, int t0, float *__restrict__ y, const int *gate) {
    if (gate_closed(gate)) return;
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= N) return;
    float acc[NC];
#pragma unroll
    for (int t = 0; t < NC; t++) acc[t] = 0.f;
#pragma unroll 4
    for (int hb = lane; hb < nbr * 2; hb += 32) {          // half blocks of 16 quants
        const int b = hb >> 1;
        const int64_t blk = row * nbr + b;
        const int4 q = ldg_stream(w.q80_qs() + blk * 32 + (hb & 1) * 16);
        const float dw = __half2float(w.q80_d()[blk]);
#pragma unroll
        for (int t = 0; t < NC; t++) {
            const int8_t *ap = a.qs + (int64_t)(t0 + t) * a.K + hb * 16;
            acc[t] += (float)dot16(q, *(const int4 *)ap, 0) * (dw * a.d[(int64_t)(t0 + t) * nbr + b]);
        }
    }
#pragma unroll
    for (int t = 0; t < NC; t++) { const float s = warp_sum(acc[t]); if (lane == 0) y[(int64_t)(t0 + t) * N + row] = s; }
}

template <int NC>
static void launch_nc(int type, const void *W, int64_t N, int64_t K, const ActQ &a, int t0, float *y, cudaStream_t st) {
    const dim3 grid((unsigned)cdiv(N, 4)), block(128);
    const int be = (type == EB_Q4_K || type == EB_Q6_K) ? 256 : 32;
    const int nbr = (int)(K / be);
    WPlanes w{(const uint8_t *)W, N * nbr};
    switch (type) {
        case EB_Q4_K: mmvq_q4_K<NC><<<grid, block, 0, st>>>((const block_q4_K *)W, N, nbr, a, t0, y, eb_gate()); break;
        case EB_Q6_K: mmvq_q6_K<NC><<<grid, block, 0, st>>>(w, N, nbr, a, t0, y, eb_gate()); break;
        case EB_Q4_0: mmvq_q4_0<NC><<<grid, block, 0, st>>>(w, N, nbr, a, t0, y, eb_gate()); break;
        case EB_Q8_0: mmvq_q8_0<NC><<<grid, block, 0, st>>>(w, N, nbr, a, t0, y, eb_gate()); break;
    }
}

int eb_mmvq(int type, const void *W, int64_t N, int64_t K, const ActQ &a, float *y, cudaStream_t st) {
    if (type != EB_Q4_K && type != EB_Q6_K && type != EB_Q4_0 && type != EB_Q8_0) return eb_set_err(EB_ERR_UNSUPPORTED, "mmvq: type %d", type);
    int t0 = 0;
    while (t0 < a.T) {
        const int rem = a.T - t0;
        if (rem >= 8) { launch_nc<8>(type, W, N, K, a, t0, y, st); t0 += 8; }
        else if (rem >= 4) { launch_nc<4>(type, W, N, K, a, t0, y, st); t0 += 4; }
        else if (rem >= 2) { launch_nc<2>(type, W, N, K, a, t0, y, st); t0 += 2; }
        else { launch_nc<1>(type, W, N, K, a, t0, y, st); t0 += 1; }
        EB_CHECK_LAUNCH();
    }
    return EB_OK;
}
