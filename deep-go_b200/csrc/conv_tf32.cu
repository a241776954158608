// DG_PREC_TF32 path: the conv stack of experiments.lua:133-153 on the tensor cores in TF32 -- the tensor-core type closest to the
// reference's CUDA arithmetic (train.lua:99-100 casts to float, cunn's SpatialConvolutionMM is fp32): fp32 storage everywhere,
// operands read by tcgen05.mma kind::tf32 (10-bit mantissa inside the multiplier), fp32 accumulation in TMEM.
//
// Same implicit-GEMM scheme as the bf16 path (conv_umma.cu), with 4-byte elements: "blocked image" per position
// [C/4][R][4] fp32, R rows of 4 channels (16 B), row index over the zero-padded board; a k x k tap is the same image at a row
// offset, i.e. a different descriptor start address.  No-swizzle canonical layout: core matrix = 8 rows x 16 B, an MMA takes
// K = 8 channels = two 16-byte chunks.  A 128-channel fp32 image is 221 KB and does not fit beside a weight ring, so the image
// goes through shared memory in PANELS of `kp` channels (ring of image buffers), the accumulators collecting all panels and taps;
// layers run one launch each (no chaining: that needs the whole image resident).
//   forward / input gradient: D[row][co] += A(image panel, K-major) x B(weights [tap][ci/4][co][4], K-major)
//   weight gradient:          D[co][ci]  += A(dZ panel, MN-major)   x B(X panel shifted by the tap, MN-major), K = rows
// The 1-channel last layer, log-softmax and NLL stay on the fp32 CUDA-core kernels (conv_simt.cu, head.cu): 0.15 % of the FLOPs.
#include "model.cuh"
#include "umma.cuh"
#include "../../include/deepgo.h"
#include <algorithm>
#include <cstdlib>

using namespace umma;

#define T32_R3 432
#define T32_R5 600
#define T32_MAX_STAGES 8

__host__ __device__ __forceinline__ uint32_t make_idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
    uint32_t d = 0;
    d |= 1u << 4;                       // D format f32
    d |= 2u << 7;                       // A format tf32
    d |= 2u << 10;                      // B format tf32
    d |= (uint32_t)(a_mn_major & 1) << 15;
    d |= (uint32_t)(b_mn_major & 1) << 16;
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// The tensor core reads the upper 19 bits of an fp32 operand, i.e. it TRUNCATES to TF32.  Everything an MMA will read (stored
// activations, stored dZ, weight operands) is therefore rounded to nearest TF32 when it is written: half the error, no bias.
__device__ __forceinline__ float tf32_rna(float x) { uint32_t u; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x)); return __uint_as_float(u); }

struct T32Conv {
    const float *in;        // [n][cin/4][R_in][4]
    const float *w;         // [taps][cin/4][cout][4]   (input gradient: taps flipped, in/out channels swapped)
    const float *bias;      // mode 0: [cout/4][R_out][4] fp32, b[c] + a[c][x][y]
    const float *mask;      // mode 1: [n][cout/4][R_mask][4], output kept where mask > 0
    float *out;             // [n][cout/4][R_out][4]
    int n, cin, cout;
    int R_in, pitch_in, pad_in, ksize, first, ntiles;
    int R_out, pitch_out, pad_out;
    int R_mask, pitch_mask, pad_mask;
    int nb, kp, nstages, nimg, mode;
};

// 384 threads: warp 0 producer (bulk copies); warps 1, 10, 11 issue the MMAs, each for its own accumulator tile(s) (the tensor pipe's
// queue is shallow: with a single issuing thread every barrier wait and commit shows up as a gap, see conv_umma.cu); warps 2..9
// epilogue (TMEM lane quarter = warp % 4, two warps per quarter taking alternate 32-column blocks)
#define T32_MMA_WARPS 3
__global__ void __launch_bounds__(384, 1) conv_tf32_kernel(const __grid_constant__ T32Conv p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t w_full[T32_MAX_STAGES], w_empty[T32_MAX_STAGES], img_full[2], img_empty[2], acc_full, acc_empty;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks_kp = p.kp >> 2, nkp = p.cin / p.kp, nnb = p.cout / p.nb, taps = p.ksize * p.ksize;
    const uint32_t panel_bytes = (uint32_t)chunks_kp * p.R_in * 16;
    const uint32_t panel_pad = (panel_bytes + 1023) & ~1023u;
    const uint32_t stage_bytes = (uint32_t)chunks_kp * p.nb * 16;
    uint8_t *ring = smem + (size_t)p.nimg * panel_pad;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], T32_MMA_WARPS); }
        for (int i = 0; i < 2; ++i) { mbar_init(&img_full[i], 1); mbar_init(&img_empty[i], T32_MMA_WARPS); }
        mbar_init(&acc_full, T32_MMA_WARPS);
        mbar_init(&acc_empty, 8);
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(&tmem_base_s, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        uint32_t s = 0, ph = 0, ib = 0, iph = 0;
        const size_t tap_stride = (size_t)(p.cin >> 2) * p.cout * 16;
        for (int pos = blockIdx.x; pos < p.n; pos += gridDim.x)
            for (int nbk = 0; nbk < nnb; ++nbk)
                for (int k = 0; k < nkp; ++k) {
                    mbar_wait(&img_empty[ib], iph ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&img_full[ib], panel_bytes);
                        const uint8_t *src = (const uint8_t *)p.in + ((size_t)pos * (p.cin >> 2) + (size_t)k * chunks_kp) * p.R_in * 16;
                        for (int c = 0; c < chunks_kp; ++c)
                            bulk_g2s(smem + (size_t)ib * panel_pad + (size_t)c * p.R_in * 16, src + (size_t)c * p.R_in * 16, (uint32_t)p.R_in * 16, &img_full[ib]);
                    }
                    __syncwarp();
                    if (++ib == (uint32_t)p.nimg) { ib = 0; iph ^= 1; }
                    for (int t = 0; t < taps; ++t) {
                        mbar_wait(&w_empty[s], ph ^ 1);
                        if (elect_one()) {
                            mbar_arrive_expect_tx(&w_full[s], stage_bytes);
                            const uint8_t *wsrc = (const uint8_t *)p.w + (size_t)t * tap_stride + ((size_t)k * chunks_kp * p.cout + (size_t)nbk * p.nb) * 16;
                            if (nnb == 1) bulk_g2s(ring + (size_t)s * stage_bytes, wsrc, stage_bytes, &w_full[s]);      // one N block: the stage is one contiguous run
                            else
                                for (int c = 0; c < chunks_kp; ++c)
                                    bulk_g2s(ring + (size_t)s * stage_bytes + (size_t)c * p.nb * 16, wsrc + (size_t)c * p.cout * 16, (uint32_t)p.nb * 16, &w_full[s]);
                        }
                        __syncwarp();
                        if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
                    }
                }
    } else if (warp == 1 || warp >= 10) {
        const int mma_id = warp == 1 ? 0 : warp - 9;
        const uint32_t idesc = make_idesc_tf32(128, p.nb, 0, 0);
        const uint32_t hi = desc_hi(128);
        const int ksteps = p.kp >> 3;
        uint32_t s = 0, ph = 0, ib = 0, iph = 0, a = 0;
        for (int pos = blockIdx.x; pos < p.n; pos += gridDim.x)
            for (int nbk = 0; nbk < nnb; ++nbk, ++a) {
                mbar_wait(&acc_empty, (a & 1) ^ 1);
                tc_fence_after();
                for (int k = 0; k < nkp; ++k) {
                    mbar_wait(&img_full[ib], iph);
                    tc_fence_after();
                    const uint32_t a_lo_k = desc_lo(smem_u32(smem) + ib * panel_pad, (uint32_t)p.R_in * 16) + (uint32_t)(p.first - p.pad_in * p.pitch_in - p.pad_in);
                    for (int ky = 0; ky < p.ksize; ++ky)
                        for (int kx = 0; kx < p.ksize; ++kx) {
                            mbar_wait(&w_full[s], ph);
                            tc_fence_after();
                            if (elect_one()) {
                                const uint32_t b_lo_s = desc_lo(smem_u32(ring) + s * stage_bytes, (uint32_t)p.nb * 16);
                                for (int tile = mma_id; tile < p.ntiles; tile += T32_MMA_WARPS) {
                                    uint32_t a_lo = a_lo_k + (uint32_t)(ky * p.pitch_in + kx) + (uint32_t)tile * 128, b_lo = b_lo_s;
                                    for (int ks = 0; ks < ksteps; ++ks, a_lo += 2u * p.R_in, b_lo += 2u * p.nb)
                                        mma_tf32(tmem_base + (uint32_t)tile * p.nb, desc_join(a_lo, hi), desc_join(b_lo, hi), idesc, (uint32_t)((k | ky | kx | ks) != 0));
                                }
                                mma_commit(&w_empty[s]);
                            }
                            __syncwarp();
                            if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
                        }
                    if (elect_one()) mma_commit(&img_empty[ib]);
                    __syncwarp();
                    if (++ib == (uint32_t)p.nimg) { ib = 0; iph ^= 1; }
                }
                if (elect_one()) mma_commit(&acc_full);
                __syncwarp();
            }
    } else {
        const int q = warp & 3, half = (warp - 2) >> 2;
        const uint32_t tm_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int chunks_out = p.cout >> 2;
        uint32_t a = 0;
        for (int pos = blockIdx.x; pos < p.n; pos += gridDim.x)
            for (int nbk = 0; nbk < nnb; ++nbk, ++a) {
                mbar_wait(&acc_full, a & 1);
                tc_fence_after();
                for (int tile = 0; tile < p.ntiles; ++tile) {
                    const int rho = p.first + tile * 128 + q * 32 + lane;
                    const int xp = rho / p.pitch_in, yp = rho - xp * p.pitch_in;
                    const int x = xp - p.pad_in, y = yp - p.pad_in;
                    const bool valid = x >= 0 && x < DG_BOARD && y >= 0 && y < DG_BOARD;
                    const int rho_out = (x + p.pad_out) * p.pitch_out + y + p.pad_out;
                    const int rho_m = (x + p.pad_mask) * p.pitch_mask + y + p.pad_mask;
                    for (int c0 = half * 32; c0 < p.nb; c0 += 64) {          // nb is a multiple of 32
                        const int chunk0 = (nbk * p.nb + c0) >> 2;
                        // the eight bias (or mask) lines of the block are requested first, under the TMEM load: the drain is a chain of
                        // L2 round trips otherwise
                        float4 bm[8];
                        if (valid) {
                            const float4 *src = p.mode == 0 ? reinterpret_cast<const float4 *>(p.bias) + (size_t)chunk0 * p.R_out + rho_out
                                                            : reinterpret_cast<const float4 *>(p.mask) + ((size_t)pos * chunks_out + chunk0) * p.R_mask + rho_m;
                            const size_t stride = p.mode == 0 ? p.R_out : p.R_mask;
#pragma unroll
                            for (int g = 0; g < 8; ++g) bm[g] = __ldg(src + (size_t)g * stride);
                        }
                        uint32_t r[32];
                        tmem_ld32(tm_lane + (uint32_t)tile * p.nb + c0, r);
                        tmem_ld_wait();
                        if (!valid) continue;
                        float4 *dst = reinterpret_cast<float4 *>(p.out) + ((size_t)pos * chunks_out + chunk0) * p.R_out + rho_out;
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            float4 v = make_float4(__uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]), __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]));
                            const float4 b = bm[g];
                            if (p.mode == 0) {
                                v.x = fmaxf(v.x + b.x, 0.f); v.y = fmaxf(v.y + b.y, 0.f); v.z = fmaxf(v.z + b.z, 0.f); v.w = fmaxf(v.w + b.w, 0.f);
                            } else {
                                v.x = b.x > 0.f ? v.x : 0.f; v.y = b.y > 0.f ? v.y : 0.f; v.z = b.z > 0.f ? v.z : 0.f; v.w = b.w > 0.f ? v.w : 0.f;
                            }
                            v.x = tf32_rna(v.x); v.y = tf32_rna(v.y); v.z = tf32_rna(v.z); v.w = tf32_rna(v.w);
                            dst[(size_t)g * p.R_out] = v;
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty);
            }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------------
// weight gradient.  D[co][ci] = sum over rows of dZ[row][co] * X[row + tap][ci] needs both operands with the contraction index
// (rows) along the 16-byte lines' OTHER axis: "MN-major" descriptors.  Measured on B200: tcgen05.mma kind::tf32 with MN-major
// operands in the no-swizzle layout leaves the accumulator at ZERO whatever LBO / SBO say (the same kernel with the K-major bit
// produces numbers), and a K-major copy of the images would need tap shifts of 4 bytes where descriptor addresses move in steps
// of 16.  So the weight gradient of this path runs on the bf16 tensor-core kernel of conv_umma.cu with both fp32 images split
// into bf16 hi + lo parts, three products per tap -- hi*hi + hi*lo + lo*hi, 16 significant bits per factor, fp32 accumulation --
// which is MORE accurate than TF32's 11 bits; the price is three launches per layer.
// ---------------------------------------------------------------------------------------------
// fp32 blocked [n][C/4][R][4] -> bf16 blocked hi and lo [n][C/8][R][8]: v = hi + lo + O(2^-17 |v|); pad rows are zero in, zero out
__global__ void __launch_bounds__(256) t32_split_bf16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo,
                                                             size_t n_chunk8_rows, int R) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunk8_rows; i += (size_t)gridDim.x * blockDim.x) {
        const size_t c8 = i / R, row = i - c8 * R;                       // c8 runs over (position, chunk of 8 channels)
        const float4 a = __ldg(reinterpret_cast<const float4 *>(src) + (2 * c8) * R + row), b = __ldg(reinterpret_cast<const float4 *>(src) + (2 * c8 + 1) * R + row);
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        __align__(16) __nv_bfloat16 h[8], l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { h[e] = __float2bfloat16(v[e]); l[e] = __float2bfloat16(v[e] - __bfloat162float(h[e])); }
        reinterpret_cast<uint4 *>(hi)[i] = *reinterpret_cast<const uint4 *>(h);
        reinterpret_cast<uint4 *>(lo)[i] = *reinterpret_cast<const uint4 *>(l);
    }
}

// ---------------------------------------------------------------------------------------------
// layout converters, bias gradients, operand packing
// ---------------------------------------------------------------------------------------------
// fp32 NCHW [n][c_true][361] -> blocked [n][c_pad/4][R][4] (channels >= c_true zero; pad rows stay zero from allocation)
__global__ void __launch_bounds__(384) t32_nchw_to_blocked_kernel(const float *__restrict__ src, float *__restrict__ dst, int c_true, int c_pad, int R, int pitch, int pad) {
    const int n = blockIdx.x, pt = threadIdx.x;
    if (pt >= DG_POINTS) return;
    const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
    const int rho = (x + pad) * pitch + y + pad;
    for (int c = blockIdx.y; c < (c_pad >> 2); c += gridDim.y) {
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = c * 4 + e < c_true ? src[((size_t)n * c_true + c * 4 + e) * DG_POINTS + pt] : 0.f;
        reinterpret_cast<float4 *>(dst)[((size_t)n * (c_pad >> 2) + c) * R + rho] = make_float4(tf32_rna(v[0]), tf32_rna(v[1]), tf32_rna(v[2]), tf32_rna(v[3]));
    }
}
__global__ void __launch_bounds__(384) t32_blocked_to_nchw_kernel(const float *__restrict__ src, float *__restrict__ dst, int C, int R, int pitch, int pad) {
    const int n = blockIdx.x, pt = threadIdx.x;
    if (pt >= DG_POINTS) return;
    const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
    const int rho = (x + pad) * pitch + y + pad;
    for (int c = blockIdx.y; c < (C >> 2); c += gridDim.y) {
        const float4 v = reinterpret_cast<const float4 *>(src)[((size_t)n * (C >> 2) + c) * R + rho];
        float *o = dst + ((size_t)n * C + c * 4) * DG_POINTS + pt;
        o[0] = v.x; o[DG_POINTS] = v.y; o[2 * DG_POINTS] = v.z; o[3 * DG_POINTS] = v.w;
    }
}
// ga[c][p] += sum_n dz[n][c][rho(p)] ; gb[c] += sum_{n,p} dz   (nn.Add gradBias, conv gradBias) from a blocked fp32 dZ image
__global__ void __launch_bounds__(384) t32_bias_grad_kernel(const float *__restrict__ dz, float *__restrict__ ga, float *__restrict__ gb, int n, int cout,
                                                            int R, int pitch, int pad, int groups) {
    __shared__ float red[12][4];
    const int chunk = blockIdx.x, grp = blockIdx.y, p = threadIdx.x;
    float s[4] = {0.f, 0.f, 0.f, 0.f};
    if (p < DG_POINTS) {
        const int x = p / DG_BOARD, y = p - x * DG_BOARD;
        const int rho = (x + pad) * pitch + y + pad;
        for (int i = grp; i < n; i += groups) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(dz) + ((size_t)i * (cout >> 2) + chunk) * R + rho);
            s[0] += v.x; s[1] += v.y; s[2] += v.z; s[3] += v.w;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) atomicAdd(&ga[(size_t)(chunk * 4 + e) * DG_POINTS + p], s[e]);
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float v = s[e];
#pragma unroll
        for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
        if ((p & 31) == 0) red[p >> 5][e] = v;
    }
    __syncthreads();
    if (p < 4) {
        float t = 0.f;
        for (int w = 0; w < 12; ++w) t += red[w][p];
        atomicAdd(&gb[chunk * 4 + p], t);
    }
}
// master fp32 (Torch7 order) -> operands of one conv layer: wf[t][ci/4][co][4] = W[co][ci][t] (ci >= cin_true -> 0),
// wd[t][co/4][ci][4] = W[co][ci][KK-1-t] (may be NULL), ab[c/4][rho][4] = b[c] + a[c][x][y] in the 3x3 output geometry
struct T32Pack { const float *W, *b, *a; float *wf, *wd, *ab; int cin_true, cin, cout, ksize; };
__global__ void __launch_bounds__(256) t32_pack_kernel(const T32Pack *__restrict__ table, int R_out, int pitch_out, int pad_out) {
    const T32Pack L = table[blockIdx.y];
    const int KK = L.ksize * L.ksize, nwf = KK * L.cin * L.cout, stride = gridDim.x * blockDim.x;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nwf; idx += stride) {
        const int e = idx & 3, co = (idx >> 2) % L.cout, c4 = (idx / (4 * L.cout)) % (L.cin >> 2), t = idx / (4 * L.cout * (L.cin >> 2));
        const int ci = c4 * 4 + e;
        L.wf[idx] = ci < L.cin_true ? tf32_rna(L.W[((size_t)co * L.cin_true + ci) * KK + t]) : 0.f;
    }
    if (L.wd)
        for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nwf; idx += stride) {
            const int e = idx & 3, ci = (idx >> 2) % L.cin, c4 = (idx / (4 * L.cin)) % (L.cout >> 2), t = idx / (4 * L.cin * (L.cout >> 2));
            const int co = c4 * 4 + e;
            L.wd[idx] = ci < L.cin_true ? tf32_rna(L.W[((size_t)co * L.cin_true + ci) * KK + (KK - 1 - t)]) : 0.f;
        }
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < L.cout * DG_POINTS; idx += stride) {
        const int c = idx / DG_POINTS, pt = idx - c * DG_POINTS;
        const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
        L.ab[((size_t)(c >> 2) * R_out + (x + pad_out) * pitch_out + y + pad_out) * 4 + (c & 3)] = L.b[c] + L.a[idx];
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct T32Geom { int pitch, pad, rows, first, ntiles; };
static T32Geom t32_geom(int k) {
    T32Geom g;
    g.pad = k / 2; g.pitch = DG_BOARD + g.pad; g.rows = k == 5 ? T32_R5 : T32_R3;
    g.first = g.pad * g.pitch + g.pad; g.ntiles = k == 5 ? 4 : 3;
    return g;
}

struct Tf32Path {
    T32Geom g_first, g3;
    int cin0_pad = 48;                     // 37 input planes padded to a multiple of 16 (N block of the first layer's weight gradient)
    float *x0 = nullptr;                   // [B][12][R_first][4]
    std::vector<float *> h;                // h[l], l = 1..L-1: [B][C/4][R3][4]
    std::vector<float *> dzl;              // dZ of conv layer l: l = 0 in the first layer's geometry, l >= 1 in the 3x3 geometry
    std::vector<float *> wf, wd, ab;       // per conv layer 0..L-2
    float *gacc = nullptr;                 // weight-gradient accumulation blocks [tap][cin/4][cout][4] of all conv layers
    std::vector<size_t> gacc_off;
    size_t gacc_floats = 0;
    std::vector<__nv_bfloat16 *> dz_hi, dz_lo, x_hi, x_lo;      // bf16 hi / lo splits of every conv layer's dZ and X images
    T32Pack *pack_table = nullptr;
    FinalizeLayer *fin_table = nullptr;
    int num_sms = 148;
};

int dg_tf32_create(dg_model *m) {
    DG_REQUIRE(m->C % 32 == 0 && m->C >= 32 && m->C <= 256, "DG_PREC_TF32 needs channelSize %% 32 == 0 and <= 256 (got %d)", m->C);
    Tf32Path *b = new Tf32Path();
    m->t32 = b;
    b->g_first = t32_geom(m->k_first);
    b->g3 = t32_geom(3);
    cudaDeviceProp prop;
    DG_CUDA(cudaGetDeviceProperties(&prop, m->ctx->dev));
    b->num_sms = prop.multiProcessorCount;
    const size_t B = m->max_batch;
    const int L = m->L, C = m->C;
    cudaStream_t st = m->ctx->stream;
    auto alloc0 = [&](float **p, size_t floats) -> int {
        DG_CUDA(cudaMalloc((void **)p, floats * 4));
        DG_CUDA(cudaMemsetAsync(*p, 0, floats * 4, st));
        return 0;
    };
    if (int rc = alloc0(&b->x0, B * b->cin0_pad * b->g_first.rows)) return rc;
    b->h.assign(L, nullptr);
    for (int l = 1; l < L; ++l) if (int rc = alloc0(&b->h[l], B * (size_t)C * T32_R3)) return rc;
    b->dzl.assign(L - 1, nullptr);
    for (int l = 0; l < L - 1; ++l) if (int rc = alloc0(&b->dzl[l], B * (size_t)C * (l == 0 ? b->g_first.rows : T32_R3))) return rc;
    b->wf.assign(L - 1, nullptr); b->wd.assign(L - 1, nullptr); b->ab.assign(L - 1, nullptr);
    b->gacc_off.assign(L - 1, 0);
    std::vector<T32Pack> tab(L - 1);
    std::vector<FinalizeLayer> ft(L - 1);
    for (int l = 0; l < L - 1; ++l) {
        const LayerInfo &li = m->layers[l];
        const int cin_pad = l == 0 ? b->cin0_pad : li.cin;
        const size_t nw = (size_t)li.k * li.k * cin_pad * li.cout;
        if (int rc = alloc0(&b->wf[l], nw)) return rc;
        if (l > 0) if (int rc = alloc0(&b->wd[l], nw)) return rc;
        if (int rc = alloc0(&b->ab[l], (size_t)li.cout * T32_R3)) return rc;
        b->gacc_off[l] = b->gacc_floats; b->gacc_floats += nw;
        tab[l] = T32Pack{m->params + li.offW, m->params + li.offB, m->params + li.offA, b->wf[l], b->wd[l], b->ab[l], li.cin, cin_pad, li.cout, li.k};
    }
    DG_CUDA(cudaMalloc((void **)&b->gacc, b->gacc_floats * 4));
    b->dz_hi.assign(L - 1, nullptr); b->dz_lo.assign(L - 1, nullptr); b->x_hi.assign(L - 1, nullptr); b->x_lo.assign(L - 1, nullptr);
    for (int l = 0; l < L - 1; ++l) {
        const size_t R = l == 0 ? b->g_first.rows : T32_R3;
        const size_t dzb = B * (size_t)C * R * 2, xb = B * (size_t)(l == 0 ? b->cin0_pad : C) * R * 2;
        DG_CUDA(cudaMalloc((void **)&b->dz_hi[l], dzb)); DG_CUDA(cudaMalloc((void **)&b->dz_lo[l], dzb));
        DG_CUDA(cudaMalloc((void **)&b->x_hi[l], xb)); DG_CUDA(cudaMalloc((void **)&b->x_lo[l], xb));
    }
    for (int l = 0; l < L - 1; ++l) {
        const LayerInfo &li = m->layers[l];
        FinalizeLayer f{};
        f.gacc = b->gacc + b->gacc_off[l]; f.gW = m->grads + li.offW; f.cin_true = li.cin; f.cin = l == 0 ? b->cin0_pad : li.cin; f.cout = li.cout; f.ksize = li.k;
        f.gab = nullptr; f.ga = nullptr; f.gb = nullptr; f.pitch = f.pad = f.first = 0;
        ft[l] = f;
    }
    DG_CUDA(cudaMalloc((void **)&b->pack_table, tab.size() * sizeof(T32Pack)));
    DG_CUDA(cudaMemcpyAsync(b->pack_table, tab.data(), tab.size() * sizeof(T32Pack), cudaMemcpyHostToDevice, st));
    DG_CUDA(cudaMalloc((void **)&b->fin_table, ft.size() * sizeof(FinalizeLayer)));
    DG_CUDA(cudaMemcpyAsync(b->fin_table, ft.data(), ft.size() * sizeof(FinalizeLayer), cudaMemcpyHostToDevice, st));
    DG_CUDA(cudaStreamSynchronize(st));
    DG_CUDA(cudaFuncSetAttribute(conv_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
    return 0;
}

void dg_tf32_destroy(dg_model *m) {
    Tf32Path *b = m->t32;
    if (!b) return;
    cudaFree(b->x0);
    for (float *p : b->h) if (p) cudaFree(p);
    for (float *p : b->dzl) if (p) cudaFree(p);
    for (float *p : b->wf) if (p) cudaFree(p);
    for (float *p : b->wd) if (p) cudaFree(p);
    for (float *p : b->ab) if (p) cudaFree(p);
    cudaFree(b->gacc); cudaFree(b->pack_table); cudaFree(b->fin_table);
    for (auto *v : {&b->dz_hi, &b->dz_lo, &b->x_hi, &b->x_lo}) for (__nv_bfloat16 *p : *v) if (p) cudaFree(p);
    delete b;
    m->t32 = nullptr;
}

int dg_tf32_params_changed(dg_model *m) {
    Tf32Path *b = m->t32;
    dg_ctx *c = m->ctx;
    { dg_prof_scope _ps(c, DG_K_PACK);
    t32_pack_kernel<<<dim3(b->num_sms, m->L - 1), 256, 0, c->stream>>>(b->pack_table, T32_R3, b->g3.pitch, b->g3.pad); }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

static int t32_launch_conv(dg_model *m, T32Conv cp) {
    dg_ctx *c = m->ctx;
    cp.nb = cp.cout <= 128 ? cp.cout : (cp.cout % 96 == 0 ? 96 : (cp.cout % 128 == 0 ? 128 : 64));
    DG_REQUIRE(cp.cout % cp.nb == 0 && cp.nb % 32 == 0 && cp.ntiles * cp.nb <= 512, "tf32 path: unsupported channel count %d", cp.cout);
    cp.kp = cp.cin <= 48 ? cp.cin : 32;                          // channels per image panel
    DG_REQUIRE(cp.cin % cp.kp == 0 && cp.kp % 8 == 0, "tf32 path: unsupported input channel count %d", cp.cin);
    const size_t panel = (((size_t)(cp.kp / 4) * cp.R_in * 16) + 1023) & ~(size_t)1023, stage = (size_t)cp.kp * cp.nb * 4;
    const size_t budget = 227 * 1024 - 2048;
    cp.nimg = (cp.cin / cp.kp > 1 && 2 * panel + 3 * stage <= budget) ? 2 : 1;
    DG_REQUIRE(cp.nimg * panel + 2 * stage <= budget, "tf32 path: image panel (%zu B) + weight ring do not fit shared memory", panel);
    cp.nstages = (int)std::min<size_t>((budget - cp.nimg * panel) / stage, T32_MAX_STAGES);
    const size_t smem = cp.nimg * panel + cp.nstages * stage;
    const int grid = std::min(cp.n, m->t32->num_sms);
    { dg_prof_scope _ps(c, cp.mode ? DG_K_DGRAD : (cp.ksize == 5 ? DG_K_CONV_FIRST : DG_K_CONV));
    conv_tf32_kernel<<<grid, 384, smem, c->stream>>>(cp); }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// conv layers 0 .. L-2 on the tensor cores; leaves h_{L-1} in NCHW fp32 in m->act[L-1] for the fp32 head (api.cu)
int dg_tf32_forward(dg_model *m) {
    Tf32Path *b = m->t32;
    dg_ctx *c = m->ctx;
    t32_nchw_to_blocked_kernel<<<dim3(m->n, 4), 384, 0, c->stream>>>(m->act[0], b->x0, m->in_planes, b->cin0_pad, b->g_first.rows, b->g_first.pitch, b->g_first.pad);
    c->launches++;
    for (int l = 0; l < m->L - 1; ++l) {
        const T32Geom &gi = l == 0 ? b->g_first : b->g3;
        T32Conv cp{};
        cp.in = l == 0 ? b->x0 : b->h[l]; cp.w = b->wf[l]; cp.bias = b->ab[l]; cp.out = b->h[l + 1];
        cp.n = m->n; cp.cin = l == 0 ? b->cin0_pad : m->layers[l].cin; cp.cout = m->layers[l].cout;
        cp.R_in = gi.rows; cp.pitch_in = gi.pitch; cp.pad_in = gi.pad; cp.ksize = m->layers[l].k; cp.first = gi.first; cp.ntiles = gi.ntiles;
        cp.R_out = T32_R3; cp.pitch_out = b->g3.pitch; cp.pad_out = b->g3.pad;
        cp.mode = 0;
        if (int rc = t32_launch_conv(m, cp)) return rc;
    }
    t32_blocked_to_nchw_kernel<<<dim3(m->n, 8), 384, 0, c->stream>>>(b->h[m->L - 1], m->act[m->L - 1], m->C, T32_R3, b->g3.pitch, b->g3.pad);
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// dz_top: dZ of conv layer L-2 in NCHW fp32 (the input gradient of the head layer, already masked by its ReLU).  Accumulates the
// gradients of conv layers L-2 .. 0 into m->grads.
int dg_tf32_backward(dg_model *m, const float *dz_top) {
    Tf32Path *b = m->t32;
    dg_ctx *c = m->ctx;
    const int L = m->L;
    {
        const T32Geom &go = (L - 2) == 0 ? b->g_first : b->g3;
        t32_nchw_to_blocked_kernel<<<dim3(m->n, 8), 384, 0, c->stream>>>(dz_top, b->dzl[L - 2], m->C, m->C, go.rows, go.pitch, go.pad);
        c->launches++;
    }
    DG_CUDA(cudaMemsetAsync(b->gacc, 0, b->gacc_floats * 4, c->stream));
    for (int l = L - 2; l >= 0; --l) {
        const LayerInfo &li = m->layers[l];
        const T32Geom &g = l == 0 ? b->g_first : b->g3;
        const int cin_pad = l == 0 ? b->cin0_pad : li.cin;
        {   // nn.Add gradBias + conv gradBias
            const int groups = m->n >= 64 ? 32 : (m->n >= 8 ? 8 : 1);
            dg_prof_scope _ps(c, DG_K_BIAS_GRAD);
            t32_bias_grad_kernel<<<dim3(li.cout / 4, groups), 384, 0, c->stream>>>(b->dzl[l], m->grads + li.offA, m->grads + li.offB, m->n, li.cout, g.rows, g.pitch, g.pad, groups);
            c->launches++;
        }
        if (l > 0) {   // input gradient through the conv, masked by the ReLU of the layer below (h_l > 0)
            const T32Geom &go = (l - 1) == 0 ? b->g_first : b->g3;
            T32Conv cp{};
            cp.in = b->dzl[l]; cp.w = b->wd[l]; cp.mask = b->h[l]; cp.out = b->dzl[l - 1];
            cp.n = m->n; cp.cin = li.cout; cp.cout = li.cin;
            cp.R_in = T32_R3; cp.pitch_in = b->g3.pitch; cp.pad_in = b->g3.pad; cp.ksize = 3; cp.first = b->g3.first; cp.ntiles = b->g3.ntiles;
            cp.R_out = go.rows; cp.pitch_out = go.pitch; cp.pad_out = go.pad;
            cp.R_mask = T32_R3; cp.pitch_mask = b->g3.pitch; cp.pad_mask = b->g3.pad;
            cp.mode = 1;
            if (int rc = t32_launch_conv(m, cp)) return rc;
        }
    }
    // ---- weight gradients: bf16 hi / lo splits of every layer's dZ and X, then hi*hi + hi*lo + lo*hi as "layers" of persistent launches
    // of the bf16 weight-gradient kernel (15 products = 5 layers per launch for the 3x3 layers; the first layer on its own)
    for (int l = L - 2; l >= 0; --l) {
        const LayerInfo &li = m->layers[l];
        const T32Geom &g = l == 0 ? b->g_first : b->g3;
        const int cin_pad = l == 0 ? b->cin0_pad : li.cin;
        t32_split_bf16_kernel<<<2 * b->num_sms, 256, 0, c->stream>>>(b->dzl[l], b->dz_hi[l], b->dz_lo[l], (size_t)m->n * (li.cout / 8) * g.rows, g.rows);
        t32_split_bf16_kernel<<<2 * b->num_sms, 256, 0, c->stream>>>(l == 0 ? b->x0 : b->h[l], b->x_hi[l], b->x_lo[l], (size_t)m->n * (cin_pad / 8) * g.rows, g.rows);
        c->launches += 2;
    }
    for (int l = L - 2; l >= 1;) {
        const int nlay = std::min(l, 5);
        const void *dz[15], *x[15];
        float *ga[15];
        int k = 0;
        for (int j = 0; j < nlay; ++j) {
            const int lj = l - j;
            const void *dzs[3] = {b->dz_hi[lj], b->dz_hi[lj], b->dz_lo[lj]}, *xs[3] = {b->x_hi[lj], b->x_lo[lj], b->x_hi[lj]};
            for (int q = 0; q < 3; ++q, ++k) { dz[k] = dzs[q]; x[k] = xs[q]; ga[k] = b->gacc + b->gacc_off[lj]; }
        }
        if (int rc = dg_bf16_wgrad_multi(c, b->num_sms, k, dz, x, ga, m->n, m->C, m->C, 3)) return rc;
        l -= nlay;
    }
    {
        const void *dz[3] = {b->dz_hi[0], b->dz_hi[0], b->dz_lo[0]}, *x[3] = {b->x_hi[0], b->x_lo[0], b->x_hi[0]};
        float *ga[3] = {b->gacc, b->gacc, b->gacc};
        if (int rc = dg_bf16_wgrad_multi(c, b->num_sms, 3, dz, x, ga, m->n, b->cin0_pad, m->layers[0].cout, m->layers[0].k)) return rc;
    }
    if (int rc = dg_launch_wgrad_finalize(c, b->fin_table, L - 1, c->stream)) return rc;
    DG_CUDA(cudaGetLastError());
    return 0;
}
