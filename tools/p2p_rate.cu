// Dev tool: NVLink peer-memory bandwidth from a kernel, by transfer size and direction (single process, GPUs 0 and 1).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/p2p_rate.bin tools/p2p_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512) copy_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += 2 * stride) {
        const size_t i2 = i + stride;
        const float4 a = __ldcg(src + i);
        float4 b = a;
        if (i2 < n4) b = __ldcg(src + i2);
        __stcg(dst + i, a);
        if (i2 < n4) __stcg(dst + i2, b);
    }
}
int main() {
    int nd = 0; cudaGetDeviceCount(&nd);
    if (nd < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t maxb = (size_t)256 << 20;
    float4 *l0, *l1, *r1;
    cudaSetDevice(1); cudaMalloc(&r1, maxb); cudaMemset(r1, 1, maxb);
    cudaSetDevice(0); cudaDeviceEnablePeerAccess(1, 0); cudaMalloc(&l0, maxb); cudaMalloc(&l1, maxb); cudaMemset(l0, 1, maxb);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const char *names[3] = {"local -> local", "remote read (peer -> local)", "remote write (local -> peer)"};
    for (size_t bytes = (size_t)1 << 20; bytes <= maxb; bytes *= 4)
        for (int mode = 0; mode < 3; ++mode)
            for (int blocks : {144, 592}) {
                const float4 *src = mode == 1 ? r1 : l0; float4 *dst = mode == 2 ? r1 : l1;
                for (int w = 0; w < 3; ++w) copy_kernel<<<blocks, 512>>>(src, dst, bytes / 16);
                cudaEventRecord(e0);
                const int reps = 10;
                for (int w = 0; w < reps; ++w) copy_kernel<<<blocks, 512>>>(src, dst, bytes / 16);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                printf("%6.0f MB  %-30s %3d blocks: %7.1f us  %7.1f GB/s\n", bytes / 1048576.0, names[mode], blocks, ms * 1e3 / reps, bytes / (ms * 1e-3 / reps) / 1e9);
            }
    return 0;
}
