// chain of dependent tiny kernels inside a CUDA graph: cost of one launch boundary, with and without programmatic dependent launch
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_plain(float *x) { if (threadIdx.x == 0 && blockIdx.x == 0) x[0] = x[0] + 1.f; }
__global__ void k_pdl(float *x) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) x[0] = x[0] + 1.f;
}
// PDL, trigger late (after the work)
__global__ void k_pdl_late(float *x) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) x[0] = x[0] + 1.f;
    asm volatile("griddepcontrol.launch_dependents;");
}
template <typename K> float run(K kern, bool pdl, int grid, int block, size_t smem, float *d, int n) {
    cudaStream_t st; cudaStreamCreate(&st);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < n; i++) {
        cudaLaunchConfig_t c = {}; c.gridDim = dim3(grid); c.blockDim = dim3(block); c.dynamicSmemBytes = smem; c.stream = st;
        cudaLaunchAttribute a; a.id = cudaLaunchAttributeProgrammaticStreamSerialization; a.val.programmaticStreamSerializationAllowed = 1;
        if (pdl) { c.attrs = &a; c.numAttrs = 1; }
        cudaLaunchKernelEx(&c, kern, d);
    }
    cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; i++) cudaGraphLaunch(ge, st);
    cudaEventRecord(e0, st);
    for (int i = 0; i < 10; i++) cudaGraphLaunch(ge, st);
    cudaEventRecord(e1, st); cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) printf("error\n");
    return ms * 1000.f / (10 * n);
}
int main() {
    float *d; cudaMalloc(&d, 4); cudaMemset(d, 0, 4);
    const int n = 200;
    struct { int g, b; size_t s; } cfg[] = {{1, 32, 0}, {65, 512, 0}, {148, 256, 0}, {148, 640, 200 * 1024}, {2080, 256, 0}};
    for (auto c : cfg) {
        printf("grid %4d x %3d thr, smem %6zu: plain %.2f us | pdl early-trigger %.2f us | pdl late-trigger %.2f us\n", c.g, c.b, c.s,
               run(k_plain, false, c.g, c.b, c.s, d, n), run(k_pdl, true, c.g, c.b, c.s, d, n), run(k_pdl_late, true, c.g, c.b, c.s, d, n));
    }
    return 0;
}
