// ubench_mma.cu — tcgen05.mma issue/throughput micro-benchmark (kind::f16, M = 128, K = 16):
// TS (A in TMEM) vs SS (A in smem), N in {16..256}.  Operands are zeros; only timing matters.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../eagle-in-llama.cpp_b200/csrc/tc.cuh"
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void mma_f16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(128) k_mma(int mode, int n, int reps, int kadv, long long *cyc) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint32_t tmem_slot; __shared__ uint64_t barmem;
    const uint32_t sb = tc::smem_u32(smem), bar = tc::smem_u32(&barmem);
    for (int i = threadIdx.x; i < 65536 / 4; i += 128) ((uint32_t *)smem)[i] = 0;
    if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
    if (threadIdx.x < 32) tc::tmem_alloc(tc::smem_u32(&tmem_slot), 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc::fence_before(); __syncthreads(); tc::fence_after();
    const uint32_t tmem = tmem_slot;
    if (mode == 4 || mode == 5) {          // 8 MMAs per asm block, descriptors precomputed
        if (threadIdx.x < 32) {
            const uint32_t idesc = tc::idesc_f16_m128(n);
            uint32_t elected;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
            const uint64_t b0 = tc::smem_desc_sw128(sb + 32768), b1 = tc::smem_desc_sw128(sb + 32768 + 32), b2 = tc::smem_desc_sw128(sb + 32768 + 64), b3 = tc::smem_desc_sw128(sb + 32768 + 96);
            const uint64_t a0 = tc::smem_desc_sw128(sb), a1 = tc::smem_desc_sw128(sb + 32), a2 = tc::smem_desc_sw128(sb + 64), a3 = tc::smem_desc_sw128(sb + 96);
            const uint32_t ta = tmem + 256;
            long long t0 = clock64();
            for (int r = 0; r < reps; r += 8) {
                if (elected) {
                    if (mode == 4)
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %3, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %4, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %5, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %6, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %3, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %4, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %5, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %6, %2, p;\n\t}"
                            ::"r"(tmem), "r"(ta), "r"(idesc), "l"(b0), "l"(b1), "l"(b2), "l"(b3) : "memory");
                    else
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %7, %3, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], %8, %4, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %9, %5, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], %10, %6, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %7, %3, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], %8, %4, %2, p;\n\t"
                            "tcgen05.mma.cta_group::1.kind::f16 [%0], %9, %5, %2, p;\n\t" "tcgen05.mma.cta_group::1.kind::f16 [%0], %10, %6, %2, p;\n\t}"
                            ::"r"(tmem), "r"(ta), "r"(idesc), "l"(b0), "l"(b1), "l"(b2), "l"(b3), "l"(a0), "l"(a1), "l"(a2), "l"(a3) : "memory");
                }
                __syncwarp();
            }
            long long t1 = clock64();
            if (elected) tc::mma_commit(bar);
            tc::mbar_wait(bar, 0);
            long long t2 = clock64();
            if (blockIdx.x == 0 && threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t0; }
        }
    } else if (mode >= 2) {                       // warp-uniform issue loop, one elected lane issues
        if (threadIdx.x < 32) {
            const uint32_t idesc = tc::idesc_f16_m128(n);
            uint32_t elected;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
            long long t0 = clock64();
            for (int r = 0; r < reps; r++) {
                const uint32_t ko = 32 * (r & 3);
                if (elected) {
                    if (mode == 2) tc::mma_f16_ts(tmem, tmem + 256 + 8 * (r & 15), tc::smem_desc_sw128(sb + 32768 + ko), idesc, 1);
                    else mma_f16_ss(tmem, tc::smem_desc_sw128(sb + ko), tc::smem_desc_sw128(sb + 32768 + ko), idesc, 1);
                }
                __syncwarp();
            }
            long long t1 = clock64();
            if (elected) tc::mma_commit(bar);
            tc::mbar_wait(bar, 0);
            long long t2 = clock64();
            if (blockIdx.x == 0 && threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t0; }
        }
    } else if (threadIdx.x == 0) {
        const uint32_t idesc = tc::idesc_f16_m128(n);
        long long t0 = clock64();
        for (int r = 0; r < reps; r++) {
            const uint32_t ko = 32 * (r % kadv);
            if (mode == 0) tc::mma_f16_ts(tmem, tmem + 256 + 8 * (r % 16), tc::smem_desc_sw128(sb + 32768 + ko), idesc, 1);
            else mma_f16_ss(tmem, tc::smem_desc_sw128(sb + ko), tc::smem_desc_sw128(sb + 32768 + ko), idesc, 1);
        }
        long long t1 = clock64();
        tc::mma_commit(bar);
        tc::mbar_wait(bar, 0);
        long long t2 = clock64();
        if (blockIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t0; }
    }
    tc::fence_before(); __syncthreads();
    if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

int main() {
    long long *cyc; CK(cudaMalloc(&cyc, 64)); long long h[2];
    CK(cudaFuncSetAttribute(k_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
    for (int mode = 2; mode < 6; mode++)
        for (int n : {16, 64, 80, 128, 256})
            for (int reps : {16, 256}) {
                k_mma<<<148, 128, 70000>>>(mode, n, reps, 4, cyc); CK(cudaDeviceSynchronize());
                CK(cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost));
                printf("%s N=%3d reps=%3d: issue %.1f clk/mma, complete %.1f clk/mma (%.0f MAC/clk/SM)\n", mode == 0 ? "TS" : mode == 1 ? "SS" : mode == 2 ? "TS-uniform" : mode == 3 ? "SS-uniform" : mode == 4 ? "TS-block8" : "SS-block8", n, reps,
                       (double)h[0] / reps, (double)h[1] / reps, 128.0 * n * 16 * reps / h[1]);
            }
    return 0;
}
