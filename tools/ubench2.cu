// ubench2.cu — issue rates of the CUDA-core instructions the tc GEMM's dequant / epilogue warps use.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
template <int OP> __global__ void k(int iters, uint32_t *out, long long *cyc) {
    uint32_t c[8]; float f[8]; uint32_t a = threadIdx.x * 3 + 1, b = threadIdx.x + 7;
    float fa = 1.0001f + threadIdx.x, fb = 0.5f;
#pragma unroll
    for (int i = 0; i < 8; i++) { c[i] = i + threadIdx.x; f[i] = i; }
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f[i]) : "f"(fa), "f"(fb));
            if (OP == 1) asm volatile("fma.rn.f16x2 %0, %1, %2, %0;" : "+r"(c[i]) : "r"(a), "r"(b));
            if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(c[i]) : "r"(a), "r"(b));
            if (OP == 3) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(c[i]) : "r"(a), "r"(b));
            if (OP == 4) asm volatile("shf.r.clamp.b32 %0, %0, %1, %2;" : "+r"(c[i]) : "r"(a), "r"(b));
            if (OP == 5) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(fb));
            if (OP == 6) asm volatile("shfl.sync.idx.b32 %0, %0, %1, 0x1f, 0xffffffff;" : "+r"(c[i]) : "r"(b & 31));
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i] + __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    uint32_t *out; long long *cyc; CK(cudaMalloc(&out, 1 << 24)); CK(cudaMalloc(&cyc, 4096)); long long h[1];
    const char *names[] = {"FFMA", "HFMA2", "LOP3", "PRMT", "SHF", "FMUL", "SHFL"};
    const int iters = 2000, warps = 16;
#define RUN(OP) k<OP><<<148, warps * 32>>>(iters, out, cyc); CK(cudaDeviceSynchronize()); CK(cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost)); \
    printf("%-6s %.1f lanes/clk/SM\n", names[OP], (double)warps * iters * 8 * 32 / h[0]);
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    return 0;
}
