// fp32 CUDA-core convolution path (DG_PREC_FP32): the reference's CUDA arithmetic type
// (train.lua:99 casts to float; cunn SpatialConvolutionMM is fp32).  NCHW, as Torch7 stores it.
// This path exists for reference-precision parity and as an on-device cross-check of the
// tcgen05 path; it is not the fast path.
#include "common.cuh"

// out[n][o][p] = act( sum_{i,t} in[n][i][p+off(t)] * w(o,i,t) + b[o] + a[o][p] ) [* (mask[n][o][p] > 0)]
//   forward  : w(o,i,t) = W[(o*cin + i)*KK + t]                (experiments.lua:138, cross-correlation)
//   transposed (input gradient): w(o,i,t) = W[(i*cout + o)*KK + KK-1-t]
template <int K, int CO_T, int CI_T>
__global__ void __launch_bounds__(384) conv_simt_kernel(const float *__restrict__ in, const float *__restrict__ W,
                                                        const float *__restrict__ b, const float *__restrict__ a,
                                                        const float *__restrict__ mask, float *__restrict__ out,
                                                        int cin, int cout, int transposed, int relu) {
    constexpr int KK = K * K, PAD = K / 2, PW = DG_BOARD + 2 * PAD, PP = PW * PW;
    __shared__ float xs[CI_T][PP];
    __shared__ float ws[CO_T][CI_T][KK];
    const int n = blockIdx.x, o0 = blockIdx.y * CO_T, p = threadIdx.x;
    const int px = p / DG_BOARD, py = p - px * DG_BOARD;
    for (int i = threadIdx.x; i < CI_T * PP; i += blockDim.x) (&xs[0][0])[i] = 0.f;
    float acc[CO_T];
#pragma unroll
    for (int o = 0; o < CO_T; ++o) acc[o] = 0.f;
    for (int i0 = 0; i0 < cin; i0 += CI_T) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < CI_T * DG_POINTS; idx += blockDim.x) {
            const int i = idx / DG_POINTS, q = idx - i * DG_POINTS;
            const int qx = q / DG_BOARD, qy = q - qx * DG_BOARD;
            xs[i][(qx + PAD) * PW + qy + PAD] = (i0 + i < cin) ? in[((size_t)n * cin + i0 + i) * DG_POINTS + q] : 0.f;
        }
        for (int idx = threadIdx.x; idx < CO_T * CI_T * KK; idx += blockDim.x) {
            const int o = idx / (CI_T * KK), r = idx - o * (CI_T * KK), i = r / KK, t = r - i * KK;
            float v = 0.f;
            if (o0 + o < cout && i0 + i < cin)
                v = transposed ? W[((size_t)(i0 + i) * cout + o0 + o) * KK + (KK - 1 - t)] : W[((size_t)(o0 + o) * cin + i0 + i) * KK + t];
            ws[o][i][t] = v;
        }
        __syncthreads();
        if (p < DG_POINTS) {
#pragma unroll 1
            for (int i = 0; i < CI_T; ++i) {
#pragma unroll
                for (int ky = 0; ky < K; ++ky)
#pragma unroll
                    for (int kx = 0; kx < K; ++kx) {
                        const float v = xs[i][(px + ky) * PW + py + kx];
#pragma unroll
                        for (int o = 0; o < CO_T; ++o) acc[o] = fmaf(v, ws[o][i][ky * K + kx], acc[o]);
                    }
            }
        }
    }
    if (p < DG_POINTS) {
#pragma unroll
        for (int o = 0; o < CO_T; ++o) {
            if (o0 + o >= cout) break;
            float v = acc[o];
            if (b) v += b[o0 + o];
            if (a) v += a[(size_t)(o0 + o) * DG_POINTS + p];
            if (relu) v = v > 0.f ? v : 0.f;
            const size_t oi = ((size_t)n * cout + o0 + o) * DG_POINTS + p;
            if (mask) v = mask[oi] > 0.f ? v : 0.f;
            out[oi] = v;
        }
    }
}

int dg_simt_conv(dg_ctx *ctx, const float *in, const float *W, const float *b, const float *a, const float *mask,
                 float *out, int n, int cin, int cout, int k, int transposed, int relu) {
    if (n <= 0) return 0;
    constexpr int CO_T = 8, CI_T = 8;
    dim3 grid(n, (cout + CO_T - 1) / CO_T);
    if (k == 3) { dg_prof_scope _ps(ctx, transposed ? DG_K_DGRAD : DG_K_CONV); conv_simt_kernel<3, CO_T, CI_T><<<grid, 384, 0, ctx->stream>>>(in, W, b, a, mask, out, cin, cout, transposed, relu); }
    else if (k == 5) { dg_prof_scope _ps(ctx, transposed ? DG_K_DGRAD : DG_K_CONV_FIRST); conv_simt_kernel<5, CO_T, CI_T><<<grid, 384, 0, ctx->stream>>>(in, W, b, a, mask, out, cin, cout, transposed, relu); }
    else DG_REQUIRE(false, "conv kernel size %d not supported (3 or 5)", k);
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// gW[o][i][t] += sum_{n,p} dz[n][o][p] * x[n][i][p+off(t)]       (SpatialConvolutionMM accGradParameters)
template <int K, int CI_T>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(const float *__restrict__ x, const float *__restrict__ dz,
                                                         float *__restrict__ gW, int n, int cin, int cout) {
    constexpr int KK = K * K, PAD = K / 2, PW = DG_BOARD + 2 * PAD, PP = PW * PW;
    __shared__ float xs[CI_T][PP];
    __shared__ float ds[DG_POINTS];
    __shared__ float red[CI_T * KK];
    const int o = blockIdx.x, i0 = blockIdx.y * CI_T;
    for (int i = threadIdx.x; i < CI_T * PP; i += blockDim.x) (&xs[0][0])[i] = 0.f;
    for (int i = threadIdx.x; i < CI_T * KK; i += blockDim.x) red[i] = 0.f;
    float acc[CI_T][KK];
#pragma unroll
    for (int i = 0; i < CI_T; ++i)
#pragma unroll
        for (int t = 0; t < KK; ++t) acc[i][t] = 0.f;
    for (int s = 0; s < n; ++s) {
        __syncthreads();
        for (int q = threadIdx.x; q < DG_POINTS; q += blockDim.x) ds[q] = dz[((size_t)s * cout + o) * DG_POINTS + q];
        for (int idx = threadIdx.x; idx < CI_T * DG_POINTS; idx += blockDim.x) {
            const int i = idx / DG_POINTS, q = idx - i * DG_POINTS;
            const int qx = q / DG_BOARD, qy = q - qx * DG_BOARD;
            xs[i][(qx + PAD) * PW + qy + PAD] = (i0 + i < cin) ? x[((size_t)s * cin + i0 + i) * DG_POINTS + q] : 0.f;
        }
        __syncthreads();
        for (int p = threadIdx.x; p < DG_POINTS; p += blockDim.x) {
            const float d = ds[p];
            if (d == 0.f) continue;
            const int px = p / DG_BOARD, py = p - px * DG_BOARD;
#pragma unroll
            for (int i = 0; i < CI_T; ++i)
#pragma unroll
                for (int ky = 0; ky < K; ++ky)
#pragma unroll
                    for (int kx = 0; kx < K; ++kx) acc[i][ky * K + kx] = fmaf(d, xs[i][(px + ky) * PW + py + kx], acc[i][ky * K + kx]);
        }
    }
#pragma unroll
    for (int i = 0; i < CI_T; ++i)
#pragma unroll
        for (int t = 0; t < KK; ++t) {
            float v = acc[i][t];
#pragma unroll
            for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
            if ((threadIdx.x & 31) == 0) atomicAdd(&red[i * KK + t], v);
        }
    __syncthreads();
    for (int idx = threadIdx.x; idx < CI_T * KK; idx += blockDim.x) {
        const int i = idx / KK, t = idx - i * KK;
        if (i0 + i < cin) gW[((size_t)o * cin + i0 + i) * KK + t] += red[idx];
    }
}

int dg_simt_wgrad(dg_ctx *ctx, const float *x, const float *dz, float *gW, int n, int cin, int cout, int k) {
    if (n <= 0) return 0;
    if (k == 3) {
        constexpr int CI_T = 4;
        { dg_prof_scope _ps(ctx, DG_K_WGRAD);
        wgrad_simt_kernel<3, CI_T><<<dim3(cout, (cin + CI_T - 1) / CI_T), 256, 0, ctx->stream>>>(x, dz, gW, n, cin, cout);
        }
    } else if (k == 5) {
        constexpr int CI_T = 2;
        { dg_prof_scope _ps(ctx, DG_K_WGRAD);
        wgrad_simt_kernel<5, CI_T><<<dim3(cout, (cin + CI_T - 1) / CI_T), 256, 0, ctx->stream>>>(x, dz, gW, n, cin, cout);
        }
    } else DG_REQUIRE(false, "conv kernel size %d not supported (3 or 5)", k);
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// ga[o][p] += sum_n dz[n][o][p]  (nn.Add gradBias) ; gb[o] += sum_{n,p} dz[n][o][p]  (conv gradBias)
__global__ void __launch_bounds__(384) bias_grad_simt_kernel(const float *__restrict__ dz, float *__restrict__ ga,
                                                             float *__restrict__ gb, int n, int cout) {
    __shared__ float red[12];
    const int o = blockIdx.x, p = threadIdx.x;
    float s = 0.f;
    if (p < DG_POINTS) {
        for (int i = 0; i < n; ++i) s += dz[((size_t)i * cout + o) * DG_POINTS + p];
        ga[(size_t)o * DG_POINTS + p] += s;
    }
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sh);
    if ((p & 31) == 0) red[p >> 5] = s;
    __syncthreads();
    if (p == 0) {
        float t = 0.f;
        for (int w = 0; w < 12; ++w) t += red[w];
        gb[o] += t;
    }
}

int dg_simt_bias_grad(dg_ctx *ctx, const float *dz, float *ga, float *gb, int n, int cout) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_BIAS_GRAD);
    bias_grad_simt_kernel<<<cout, 384, 0, ctx->stream>>>(dz, ga, gb, n, cout);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
