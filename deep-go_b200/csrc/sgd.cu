// SGD:step (optimizer.lua:24-25): p <- p + (-rate)*g on the flat Torch7-order vector.  HBM-bound:
// 12 B per parameter (read p, read g, write p); 128-bit accesses, grid sized to the SM count.
#include "common.cuh"

__global__ void __launch_bounds__(256) sgd_kernel(float *__restrict__ p, const float *__restrict__ g, size_t n, float neg_rate) {
    dg_pdl_launch_dependents();
    dg_pdl_wait();
    const size_t n4 = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    float4 *p4 = reinterpret_cast<float4 *>(p);
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = p4[i];
        const float4 b = g4[i];
        a.x = fmaf(neg_rate, b.x, a.x); a.y = fmaf(neg_rate, b.y, a.y);
        a.z = fmaf(neg_rate, b.z, a.z); a.w = fmaf(neg_rate, b.w, a.w);
        p4[i] = a;
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = fmaf(neg_rate, g[i], p[i]);
}

// same update with the rate read from device memory, so a captured CUDA graph of the step can be replayed while
// the host keeps decaying the rate (optimizer.lua:26)
__global__ void __launch_bounds__(256) sgd_dev_rate_kernel(float *__restrict__ p, const float *__restrict__ g, size_t n, const float *__restrict__ rate) {
    dg_pdl_launch_dependents();
    dg_pdl_wait();
    const float neg_rate = -*rate;
    const size_t n4 = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    float4 *p4 = reinterpret_cast<float4 *>(p);
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 a = p4[i];
        const float4 b = g4[i];
        a.x = fmaf(neg_rate, b.x, a.x); a.y = fmaf(neg_rate, b.y, a.y);
        a.z = fmaf(neg_rate, b.z, a.z); a.w = fmaf(neg_rate, b.w, a.w);
        p4[i] = a;
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = fmaf(neg_rate, g[i], p[i]);
}

__global__ void __launch_bounds__(256) fill_kernel(float *__restrict__ p, size_t n, float v) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = v;
}

int dg_launch_sgd(dg_ctx *ctx, float *p, const float *g, size_t n, float rate) {
    if (n == 0) return 0;
    size_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks == 0) blocks = 1;
    { dg_prof_scope _ps(ctx, DG_K_SGD);
    DG_CUDA(dg_launch_pdl(sgd_kernel, dim3((unsigned)blocks), dim3(256), 0, ctx->stream, p, g, n, -rate));
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
int dg_launch_sgd_dev_rate(dg_ctx *ctx, float *p, const float *g, size_t n, const float *rate_dev) {
    if (n == 0) return 0;
    size_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks == 0) blocks = 1;
    { dg_prof_scope _ps(ctx, DG_K_SGD);
    DG_CUDA(dg_launch_pdl(sgd_dev_rate_kernel, dim3((unsigned)blocks), dim3(256), 0, ctx->stream, p, g, n, rate_dev));
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
int dg_launch_fill(dg_ctx *ctx, float *p, size_t n, float v) {
    if (n == 0) return 0;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    { dg_prof_scope _ps(ctx, DG_K_OTHER);
    fill_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(p, n, v);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
