// ubench.cu — pipe-rate micro-benchmarks that decide the GEMM design on B200 (sm_100a):
// legacy mma.sync int8 / f16 rates, IMAD / I2F rates, tcgen05.ld (TMEM read) bandwidth.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench tools/ubench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); return 1; } } while (0)

__global__ void k_imma(int iters, int *out, long long *cyc) {
    int c[8][4]; uint32_t a[4] = {threadIdx.x, 2, 3, 4}, b0 = 5, b1 = 6;
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_hmma(int iters, float *out, long long *cyc) {
    float c[8][4]; uint32_t a[4] = {threadIdx.x, 2, 3, 4}, b0 = 5, b1 = 6;
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_imad(int iters, int *out, long long *cyc) {
    int c[8], a = threadIdx.x + 3, b = threadIdx.x * 7 + 1;
#pragma unroll
    for (int i = 0; i < 8; i++) c[i] = i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("mad.lo.s32 %0, %1, %2, %0;" : "+r"(c[i]) : "r"(a), "r"(b));
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void k_i2f(int iters, float *out, long long *cyc) {
    int c[8]; float f[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { c[i] = threadIdx.x + i; f[i] = 0.f; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) { float t; asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(t) : "r"(c[i])); f[i] += t; c[i] ^= it; }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// TMEM read bandwidth: 4 warps, each reads 32 lanes x 16 columns per instruction
__global__ void __launch_bounds__(128) k_tmem_ld(int iters, uint32_t *out, long long *cyc) {
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t base = tmem_base + ((uint32_t)(warp * 32) << 16);
    uint32_t acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int cc = 0; cc < 4; cc++) {
            uint32_t r[16];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                           "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                         : "r"(base + cc * 16 + (it & 7) * 64));
            asm volatile("tcgen05.wait::ld.sync.aligned;");
#pragma unroll
            for (int i = 0; i < 16; i++) acc ^= r[i];
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base));
}

int main(int argc, char **argv) {
    const int which = argc > 1 ? atoi(argv[1]) : 0;
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    void *out; long long *cyc; CK(cudaMalloc(&out, 1 << 24)); CK(cudaMalloc(&cyc, nsm * 8 * 8));
    long long h[4096];
    const int iters = 2000;
    if (which == 0) {
        for (int warps : {4, 8, 16, 32}) {
            k_imma<<<nsm, warps * 32>>>(iters, (int *)out, cyc); CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h, cyc, nsm * 8, cudaMemcpyDeviceToHost));
            printf("IMMA.16832.u8s8  warps/SM=%2d  %.0f MAC/clk/SM\n", warps, (double)warps * iters * 8 * 4096 / h[0]);
            k_hmma<<<nsm, warps * 32>>>(iters, (float *)out, cyc); CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h, cyc, nsm * 8, cudaMemcpyDeviceToHost));
            printf("HMMA.16816.f32   warps/SM=%2d  %.0f MAC/clk/SM\n", warps, (double)warps * iters * 8 * 2048 / h[0]);
            k_imad<<<nsm, warps * 32>>>(iters, (int *)out, cyc); CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h, cyc, nsm * 8, cudaMemcpyDeviceToHost));
            printf("IMAD             warps/SM=%2d  %.1f lanes/clk/SM\n", warps, (double)warps * iters * 8 * 32 / h[0]);
            k_i2f<<<nsm, warps * 32>>>(iters, (float *)out, cyc); CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h, cyc, nsm * 8, cudaMemcpyDeviceToHost));
            printf("I2F+FADD+XOR     warps/SM=%2d  %.1f lanes/clk/SM (triples)\n", warps, (double)warps * iters * 8 * 32 / h[0]);
        }
    } else {
        k_tmem_ld<<<nsm, 128>>>(iters, (uint32_t *)out, cyc); CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h, cyc, nsm * 8, cudaMemcpyDeviceToHost));
        printf("tcgen05.ld 32x32b.x16 (+wait each): %.1f B/clk/SM (4 warps)\n", (double)iters * 4 * 4 * 32 * 16 * 4 / h[0]);
    }
    return 0;
}
