// ubench_tmem.cu — TMEM traffic under the tc GEMM's access pattern: tcgen05.ld bandwidth (8 warps, x32),
// tcgen05.mma (TS, N = 80) rate, commit latency, and both running concurrently.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../eagle-in-llama.cpp_b200/csrc/tc.cuh"
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s line %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// mode bit0: MMA warp active, bit1: loader warps active, bit2: storer warps (tcgen05.st x32) active
__global__ void __launch_bounds__(640, 1) k(int mode, int reps, long long *out) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint32_t tmem_slot; __shared__ uint64_t barmem[2];
    const uint32_t sb = tc::smem_u32(smem), bar = tc::smem_u32(&barmem[0]);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 65536 / 4; i += 640) ((uint32_t *)smem)[i] = 0;
    if (threadIdx.x == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
    if (warp == 3) tc::tmem_alloc(tc::smem_u32(&tmem_slot), 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc::fence_before(); __syncthreads(); tc::fence_after();
    const uint32_t tmem = tmem_slot;
    if (warp == 1 && (mode & 1)) {
        const bool leader = tc::elect_one();
        const uint32_t idesc = tc::idesc_f16_m128(80);
        const uint64_t bdesc = tc::smem_desc_sw128(sb);
        long long t0 = clock64();
        for (int r = 0; r < reps; r++) {
            if (leader) tc::mma_f16_ts_x8(tmem + 160 * (r & 1), tmem + 320 + 64 * (r & 1), bdesc, bdesc + 640, idesc, 1);
            __syncwarp();
        }
        long long t1 = clock64();
        if (leader) tc::mma_commit(bar);
        tc::mbar_wait(bar, 0);
        long long t2 = clock64();
        if (blockIdx.x == 0 && lane == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    } else if (warp >= 12 && (mode & 2)) {
        const uint32_t lane_base = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + 40 * ((warp - 12) >> 2);
        uint32_t acc = 0;
        long long t0 = clock64();
        for (int r = 0; r < reps; r++) {
            uint32_t v[32], w[32];
            tc::tmem_ld32(lane_base + 160 * (r & 1), v); tc::tmem_ld32(lane_base + 160 * (r & 1) + 80, w);
            tc::wait_ld();
#pragma unroll
            for (int i = 0; i < 32; i++) acc ^= v[i] + w[i];
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0 && warp == 12) { out[2] = t1 - t0; out[3] = acc; }
    } else if (warp >= 4 && warp < 12 && (mode & 4)) {
        const uint32_t lane_base = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + 320 + 32 * ((warp - 4) >> 2);
        uint32_t v[32];
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = i + lane;
        long long t0 = clock64();
        for (int r = 0; r < reps; r++) {
            tc::tmem_st32(lane_base + 64 * (r & 1), v);
            tc::wait_st();
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0 && warp == 4) out[4] = t1 - t0;
    }
    tc::fence_before(); __syncthreads();
    if (warp == 3) tc::tmem_dealloc(tmem, 512);
}
int main() {
    long long *out; CK(cudaMalloc(&out, 64)); long long h[8];
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
    const int reps = 64;
    for (int mode : {1, 2, 4, 3, 5, 6, 7}) {
        CK(cudaMemset(out, 0, 64));
        k<<<148, 640, 70000>>>(mode, reps, out); CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h, out, 64, cudaMemcpyDeviceToHost));
        printf("mode %d:", mode);
        if (mode & 1) printf("  MMA x8 block: issue %.0f clk, complete %.0f clk per 8 MMAs;", (double)h[0] / reps, (double)h[1] / reps);
        if (mode & 2) printf("  LD 8 warps x (2 x32): %.0f clk per round = %.0f B/clk/SM;", (double)h[2] / reps, 8.0 * 2 * 32 * 32 * 4 * reps / h[2]);
        if (mode & 4) printf("  ST 8 warps x32: %.0f clk per round = %.0f B/clk/SM;", (double)h[4] / reps, 8.0 * 32 * 32 * 4 * reps / h[4]);
        printf("\n");
    }
    return 0;
}
