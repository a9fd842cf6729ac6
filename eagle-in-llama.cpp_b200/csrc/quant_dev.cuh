// quant_dev.cuh — device-side activation quantisers shared by quant.cu and the fused kernels.
#pragma once
#include "common.cuh"

// value-in-registers forms: lane l of a warp owns elements 8l..8l+7 of a 256-block (Q8_K) or
// elements 4l..4l+3 of four consecutive 32-blocks (Q8_0).
// Q8_K grid (quantize_row_q8_K_ref, R/ggml/src/ggml-quants.c:2479): one warp per 256-block,
// lane l owns elements 8l..8l+7.  "max" is the signed value of the FIRST element reaching the
// largest |x| (strict > in the reference scan), so ties resolve to the lowest index.
__device__ __forceinline__ void quant_regs_q8K(const float (&v)[8], int8_t *__restrict__ qs,
                                               float *__restrict__ dout, __half *__restrict__ bsum, int lane,
                                               __half *__restrict__ h16 = nullptr, float *__restrict__ dT = nullptr) {
    float amax = 0.f, mx = 0.f; int idx = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) { const float ax = fabsf(v[j]); if (ax > amax) { amax = ax; mx = v[j]; idx = lane * 8 + j; } }
    if (amax == 0.f) idx = 1 << 20;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o), om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; mx = om; idx = oi; }
    }
    int q[8];
    if (amax == 0.f) {
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = 0;
        if (lane == 0) { *dout = 0.f; if (dT) *dT = 0.f; }
    } else {
        const float iscale = __fdiv_rn(-127.f, mx);
#pragma unroll
        for (int j = 0; j < 8; j++) q[j] = min(127, __float2int_rn(__fmul_rn(iscale, v[j])));
        if (lane == 0) { const float dd = __fdiv_rn(1.f, iscale); *dout = dd; if (dT) *dT = dd; }
    }
    if (h16) {                                                // exact: |q| <= 127.  K order inside each group of 4 is
        // (0,2,1,3): the tc GEMM's dequant extracts nibbles of bytes (0,2) and (1,3) of a word with one mask each
        const __half2 a = __floats2half2_rn((float)q[0], (float)q[2]), b = __floats2half2_rn((float)q[1], (float)q[3]);
        const __half2 c = __floats2half2_rn((float)q[4], (float)q[6]), e = __floats2half2_rn((float)q[5], (float)q[7]);
        ((uint4 *)h16)[lane] = make_uint4(*(const uint32_t *)&a, *(const uint32_t *)&b, *(const uint32_t *)&c, *(const uint32_t *)&e);
    }
    uint32_t w0 = (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((uint32_t)(q[3] & 0xff) << 24);
    uint32_t w1 = (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((uint32_t)(q[7] & 0xff) << 24);
    ((uint2 *)qs)[lane] = make_uint2(w0, w1);
    int s = q[0] + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + q[7];
    s += __shfl_xor_sync(0xffffffffu, s, 1);                  // 16 elements = 2 lanes
    if ((lane & 1) == 0) bsum[lane >> 1] = __int2half_rn(s);  // |s| <= 2032: exact
}
// Q8_0 grid (quantize_row_q8_0 AVX2 branch, R/ggml/src/ggml-cpu/ggml-cpu-quants.c:808-860):
// 8 lanes per 32-block, 4 elements each.
__device__ __forceinline__ void quant_regs_q8_0x4(const float4 a, int8_t *__restrict__ qs,
                                                  float *__restrict__ dout, int lane) {
    float amax = fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(a.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d = __fdiv_rn(amax, 127.f);
    const float id = amax != 0.f ? __fdiv_rn(127.f, amax) : 0.f;
    const int q0 = __float2int_rn(__fmul_rn(a.x, id)), q1 = __float2int_rn(__fmul_rn(a.y, id));
    const int q2 = __float2int_rn(__fmul_rn(a.z, id)), q3 = __float2int_rn(__fmul_rn(a.w, id));
    ((uint32_t *)qs)[lane] = (q0 & 0xff) | ((q1 & 0xff) << 8) | ((q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
    if ((lane & 7) == 0) dout[lane >> 3] = __half2float(__float2half_rn(d));   // stored as f16 in block_q8_0
}


__device__ __forceinline__ void quant_block_q8K(const float *__restrict__ x, int8_t *__restrict__ qs,
                                                float *__restrict__ dout, __half *__restrict__ bsum, int lane,
                                                __half *__restrict__ h16 = nullptr, float *__restrict__ dT = nullptr) {
    float v[8];
    const float4 a = ((const float4 *)x)[lane * 2], b = ((const float4 *)x)[lane * 2 + 1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    quant_regs_q8K(v, qs, dout, bsum, lane, h16, dT);
}
__device__ __forceinline__ void quant_block_q8_0x4(const float *__restrict__ x, int8_t *__restrict__ qs,
                                                   float *__restrict__ dout, int lane) {
    quant_regs_q8_0x4(((const float4 *)x)[lane], qs, dout, lane);
}
// quantise one row held in shared memory (optionally times a weight row): all warps of the CTA
__device__ __forceinline__ void quant_row_smem(const float *row, const ActQ &a, int t, int tid, int nthreads) {
    const int lane = tid & 31, warp = tid >> 5, nw = nthreads >> 5;
    if (a.kq) {
        const int nb = (int)(a.K / 256);
        for (int b = warp; b < nb; b += nw)
            quant_block_q8K(row + b * 256, a.qs + (int64_t)t * a.K + b * 256, a.d + (int64_t)t * nb + b, a.bsum + ((int64_t)t * nb + b) * 16, lane,
                            a.h16 ? a.h16 + (int64_t)t * a.K + b * 256 : nullptr, (a.dT && t < EB_DT_PITCH) ? a.dT + (int64_t)b * EB_DT_PITCH + t : nullptr);
    } else {
        const int ng = (int)(a.K / 128);
        for (int g = warp; g < ng; g += nw)
            quant_block_q8_0x4(row + g * 128, a.qs + (int64_t)t * a.K + g * 128, a.d + (int64_t)t * (a.K / 32) + g * 4, lane);
    }
}
