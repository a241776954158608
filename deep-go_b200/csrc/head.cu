// Row-wise log-softmax over the 361 points + ClassNLLCriterion forward/backward + top-1
// (experiments.lua:150-151, :45; train.lua:8-9,29-37), fused into one warp-primitive kernel.
#include "common.cuh"

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, s));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    return v;
}

// one warp per position; h = ReLU'd last-layer map (the reference keeps the ReLU before LogSoftMax,
// experiments.lua:147).  dz = scale*(softmax - onehot) masked by h > 0 (ReLU backward).
__global__ void __launch_bounds__(128) logsoftmax_nll_kernel(const float *__restrict__ h, const int32_t *__restrict__ targets,
                                                             float *__restrict__ logp, float *__restrict__ loss_n,
                                                             float *__restrict__ dz, int32_t *__restrict__ argmax, int n, float scale) {
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const float *hr = h + (size_t)row * DG_POINTS;
    float v[12];
    float mx = -INFINITY;
    int best = 0;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        const int p = j * 32 + lane;
        v[j] = p < DG_POINTS ? hr[p] : -INFINITY;
        if (v[j] > mx) { mx = v[j]; best = p; }
    }
    // first index of the maximum
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const float om = __shfl_xor_sync(0xffffffffu, mx, s);
        const int ob = __shfl_xor_sync(0xffffffffu, best, s);
        if (om > mx || (om == mx && ob < best)) { mx = om; best = ob; }
    }
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < 12; ++j) se += (j * 32 + lane < DG_POINTS) ? expf(v[j] - mx) : 0.f;
    se = warp_sum(se);
    const float lse = mx + logf(se);
    const int t = targets ? targets[row] : -1;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        const int p = j * 32 + lane;
        if (p < DG_POINTS) {
            const float lp = v[j] - lse;
            logp[(size_t)row * DG_POINTS + p] = lp;
            if (dz) {
                float g = scale * (expf(lp) - (p == t ? 1.f : 0.f));
                dz[(size_t)row * DG_POINTS + p] = v[j] > 0.f ? g : 0.f;
            }
            if (p == t && loss_n) loss_n[row] = -lp;
        }
    }
    if (lane == 0 && argmax) argmax[row] = best;
}

// deterministic single-block reduction: mean NLL (sizeAverage) and top-1 error count
__global__ void __launch_bounds__(256) loss_reduce_kernel(const float *__restrict__ loss_n, const int32_t *__restrict__ argmax,
                                                          const int32_t *__restrict__ targets, int n, float *__restrict__ loss_out,
                                                          int *__restrict__ errors_out) {
    __shared__ float sl[256];
    __shared__ int se[256];
    float s = 0.f;
    int e = 0;
    for (int i = threadIdx.x; i < n; i += 256) {
        s += loss_n[i];
        if (argmax && targets) e += argmax[i] != targets[i];
    }
    sl[threadIdx.x] = s;
    se[threadIdx.x] = e;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) { sl[threadIdx.x] += sl[threadIdx.x + w]; se[threadIdx.x] += se[threadIdx.x + w]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (loss_out) *loss_out = sl[0] / (float)n;
        if (errors_out) *errors_out = se[0];
    }
}

int dg_launch_logsoftmax_nll(dg_ctx *ctx, const float *h_last, const int32_t *targets, float *logp, float *loss_n,
                             float *dz_last, int32_t *argmax, int n, float scale) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_HEAD);
    logsoftmax_nll_kernel<<<(n + 3) / 4, 128, 0, ctx->stream>>>(h_last, targets, logp, loss_n, dz_last, argmax, n, scale);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
int dg_launch_loss_reduce(dg_ctx *ctx, const float *loss_n, const int32_t *argmax, const int32_t *targets, int n,
                          float *loss_out, int *errors_out) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_OTHER);
    loss_reduce_kernel<<<1, 256, 0, ctx->stream>>>(loss_n, argmax, targets, n, loss_out, errors_out);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
