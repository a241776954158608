// Feature-plane encoder: preprocess() of dataloader.lua:50-92 with the D4 symmetry hook (:41-44).
// One thread per board point; raw planes are read byte-wise (coalesced per plane), each of the 37
// output planes is written coalesced (fp32) or as one ballot word per warp (bit-planes).
#include "common.cuh"

// inverse D4 action: output point (x',y') <- source point (x,y) with T_s(x,y) = (x',y').
// bits: b0 transpose, b1 flip x, b2 flip y applied in that order.
__device__ __forceinline__ int dg_src_point(int p, int sym) {
    int x = p / DG_BOARD, y = p - x * DG_BOARD;
    if (sym & 2) x = DG_BOARD - 1 - x;
    if (sym & 4) y = DG_BOARD - 1 - y;
    if (sym & 1) { int t = x; x = y; y = t; }
    return x * DG_BOARD + y;
}

// 37 plane values of one point as a bit mask (bit c = plane c, 0-based channel order of dataloader.lua:6-14)
__device__ __forceinline__ unsigned long long dg_point_planes(const uint8_t *__restrict__ rec, int src, int player, int rank) {
    const int stone = rec[0 * DG_POINTS + src];
    const int lib = rec[1 * DG_POINTS + src];
    const int la = rec[(2 + player - 1) * DG_POINTS + src];
    const int kills = rec[(4 + player - 1) * DG_POINTS + src];
    const int age = rec[6 * DG_POINTS + src];
    const int ladder = rec[(7 + player - 1) * DG_POINTS + src];
    unsigned long long m = 0;
    m |= (unsigned long long)(stone == 0) << 0;                       // :57
    m |= (unsigned long long)(stone == player) << 1;                  // :58
    m |= (unsigned long long)(stone == 3 - player) << 2;              // :59
    if (lib >= 1 && lib <= 3) m |= 1ull << (3 + lib - 1);             // :68
    m |= (unsigned long long)(lib >= 4) << 6;                         // :69
    m |= (unsigned long long)(stone == 0 && la == 0) << 7;            // :73
    if (la >= 1 && la <= 5) m |= 1ull << (7 + la);                    // :74
    m |= (unsigned long long)(la >= 6) << 13;                         // :75
    if (kills >= 1 && kills <= 6) m |= 1ull << (14 + kills - 1);      // :78
    m |= (unsigned long long)(kills >= 7) << 20;                      // :79
    if (age >= 1 && age <= 5) m |= 1ull << (21 + age - 1);            // :82
    m |= (unsigned long long)(ladder >= 1) << 26;                     // :85
    m |= 1ull << (27 + rank);                                         // :87 (plane 27 itself stays 0)
    return m;
}

__global__ void __launch_bounds__(384) encode_f32_kernel(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ player,
                                                        const uint8_t *__restrict__ rank, const uint8_t *__restrict__ sym,
                                                        float *__restrict__ out) {
    const int n = blockIdx.x, p = threadIdx.x;
    if (p >= DG_POINTS) return;
    const int s = sym ? sym[n] : 0;
    const unsigned long long m = dg_point_planes(raw + (size_t)n * DG_RAW_PLANES * DG_POINTS, dg_src_point(p, s), player[n], rank[n]);
    float *o = out + (size_t)n * DG_PLANES * DG_POINTS + p;
#pragma unroll
    for (int c = 0; c < DG_PLANES; ++c) o[c * DG_POINTS] = (float)((m >> c) & 1ull);
}

__global__ void __launch_bounds__(384) encode_bits_kernel(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ player,
                                                         const uint8_t *__restrict__ rank, const uint8_t *__restrict__ sym,
                                                         uint32_t *__restrict__ out) {
    const int n = blockIdx.x, p = threadIdx.x;
    const int s = sym ? sym[n] : 0;
    unsigned long long m = 0;
    if (p < DG_POINTS) m = dg_point_planes(raw + (size_t)n * DG_RAW_PLANES * DG_POINTS, dg_src_point(p, s), player[n], rank[n]);
    const int warp = p >> 5, lane = p & 31;
    uint32_t *o = out + (size_t)n * DG_PLANES * 12 + warp;
#pragma unroll
    for (int c = 0; c < DG_PLANES; ++c) {
        const unsigned w = __ballot_sync(0xffffffffu, (m >> c) & 1ull);
        if (lane == 0) o[c * 12] = w;
    }
}

__global__ void labels_kernel(const uint8_t *__restrict__ x, const uint8_t *__restrict__ y, const uint8_t *__restrict__ sym,
                              int n, int32_t *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    dg_pdl_wait();
    if (i >= n) return;
    int a = x[i] - 1, b = y[i] - 1;
    const int s = sym ? sym[i] : 0;
    if (s & 1) { int t = a; a = b; b = t; }
    if (s & 2) a = DG_BOARD - 1 - a;
    if (s & 4) b = DG_BOARD - 1 - b;
    out[i] = DG_BOARD * a + b;
}

int dg_launch_encode_f32(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym, int n, float *out) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_ENCODE);
    encode_f32_kernel<<<n, 384, 0, ctx->stream>>>(raw, player, rank, sym, out);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
int dg_launch_encode_bits(dg_ctx *ctx, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym, int n, uint32_t *out) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_ENCODE);
    encode_bits_kernel<<<n, 384, 0, ctx->stream>>>(raw, player, rank, sym, out);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
int dg_launch_labels(dg_ctx *ctx, const uint8_t *x, const uint8_t *y, const uint8_t *sym, int n, int32_t *out) {
    if (n <= 0) return 0;
    { dg_prof_scope _ps(ctx, DG_K_OTHER);
    labels_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(x, y, sym, n, out);
    }
    ctx->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
