// does MUFU.EX2 cost scale with the number of active lanes of a warp?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float *out, int active, long long *clk) {
    const int lane = threadIdx.x & 31;
    float a[8];
    for (int i = 0; i < 8; i++) a[i] = -0.001f * (lane + i + 1);
    __syncthreads();
    const long long t0 = clock64();
    if (lane < active) {
        for (int it = 0; it < 512; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
        }
    }
    __syncwarp();
    const long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
// interleaved-lane variant: lanes with (lane % stride) == 0 active
__global__ void k2(float *out, int stride, long long *clk) {
    const int lane = threadIdx.x & 31;
    float a[8];
    for (int i = 0; i < 8; i++) a[i] = -0.001f * (lane + i + 1);
    __syncthreads();
    const long long t0 = clock64();
    if (lane % stride == 0) {
        for (int it = 0; it < 512; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
        }
    }
    __syncwarp();
    const long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
int main() {
    float *o; long long *c, h; cudaMalloc(&o, 4096 * 4); cudaMalloc(&c, 8);
    for (int warps : {1, 4}) {
        for (int act : {32, 24, 16, 8, 4, 1}) {
            k<<<1, 32 * warps>>>(o, act, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
            printf("warps %d (one per scheduler) lanes 0..%2d active: %.2f clk per warp-level EX2\n", warps, act - 1, (double)h / 4096);
        }
        for (int st : {2, 4, 8}) {
            k2<<<1, 32 * warps>>>(o, st, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
            printf("warps %d every %d-th lane active: %.2f clk per warp-level EX2\n", warps, st, (double)h / 4096);
        }
    }
    return 0;
}
