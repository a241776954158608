// Dev tool: per-SM throughput of cp.async.bulk global->shared as a function of the copy size (timing only).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/bulk_rate.bin tools/bulk_rate.cu
// Every CTA streams its own slice of a buffer (L2-resident or HBM-sized) through a 3-stage ring of 64 KB stages; one
// elected thread issues `pieces` copies per stage, a second warp waits for the stage and frees it at once.
#include "../deep-go_b200/csrc/umma.cuh"
#include <cstdio>
#include <cstdlib>
using namespace umma;

__global__ void __launch_bounds__(64, 1) bulk_rate_kernel(const uint8_t *buf, size_t per_cta, int stage_bytes, int pieces, int iters, long long *out, int share) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full[3], empty[3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { for (int s = 0; s < 3; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); } fence_barrier_init(); }
    __syncthreads();
    const uint8_t *src = buf + (size_t)(blockIdx.x / share) * per_cta;   // `share` neighbouring CTAs stream the same bytes
    const int piece = stage_bytes / pieces;
    const long long t0 = clock64();
    long long issue = 0;
    if (warp == 0) {
        uint32_t s = 0, ph = 0;
        size_t off = 0;
        for (int i = 0; i < iters; ++i) {
            mbar_wait(&empty[s], ph ^ 1);
            const long long a = clock64();
            if (elect_one()) {
                mbar_arrive_expect_tx(&full[s], stage_bytes);
                for (int c = 0; c < pieces; ++c) bulk_g2s(smem + (size_t)s * stage_bytes + (size_t)c * piece, src + off + (size_t)c * piece, piece, &full[s]);
            }
            __syncwarp();
            issue += clock64() - a;
            off += stage_bytes; if (off + stage_bytes > per_cta) off = 0;
            if (++s == 3) { s = 0; ph ^= 1; }
        }
    } else {
        uint32_t s = 0, ph = 0;
        for (int i = 0; i < iters; ++i) {
            mbar_wait(&full[s], ph);
            if (lane == 0) mbar_arrive(&empty[s]);
            __syncwarp();
            if (++s == 3) { s = 0; ph ^= 1; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) { out[blockIdx.x * 2] = clock64() - t0; out[blockIdx.x * 2 + 1] = issue; }
}

int main() {
    const int grid = 148, stage = 64 * 1024, iters = 400;
    uint8_t *buf; long long *d;
    const size_t big = (size_t)2 << 30;
    cudaMalloc(&buf, big); cudaMemset(buf, 1, big); cudaMalloc(&d, grid * 16);
    cudaFuncSetAttribute(bulk_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * stage);
    for (int share = 1; share <= 3; share += 2)
    for (int resident = 1; resident >= 0; --resident)
        for (int pieces = (share == 1 ? 1 : 32); pieces <= (share == 1 ? 64 : 32); pieces *= 2) {
            const size_t per_cta = resident ? (size_t)256 * 1024 : big / grid / stage * stage;   // 37 MB total stays in L2; 2 GB does not
            for (int rep = 0; rep < 2; ++rep) bulk_rate_kernel<<<grid, 64, 3 * stage>>>(buf, per_cta, stage, pieces, iters, d, share);
            long long h[grid * 2];
            if (cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
            double t = 0, is = 0; for (int i = 0; i < grid; ++i) { t += (double)h[2 * i] / grid; is += (double)h[2 * i + 1] / grid; }
            printf("share %d %s  %2d copies of %5d B per 64 KB stage: %6.1f B/clk per SM (%7.0f cycles per stage, %6.0f of them issuing)\n", share, resident ? "L2 " : "HBM", pieces, stage / pieces,
                   (double)stage * iters / t, t / iters, is / iters);
        }
    return 0;
}
