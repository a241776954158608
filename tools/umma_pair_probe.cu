// Dev tool: checks the cta_group::2 (CTA pair) conventions before the conv chain relies on them:
//   each CTA supplies its own 128 rows of A and HALF of the N rows of B (K-major, no swizzle, blocked [chunk][row][8]);
//   the leader issues tcgen05.mma.cta_group::2 with M = 256; each CTA's TMEM then holds D_r = A_r x [B_0 ; B_1]^T (128 x N);
//   completion is multicast to a barrier in both CTAs; the peer signals "my operands have landed" with a remote arrive.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_pair_probe tools/umma_pair_probe.cu
#include "../deep-go_b200/csrc/umma.cuh"
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
using namespace umma;

constexpr int ROWS_A = 192, K = 64, N = 128, SHIFT = 21;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) pair_kernel(const uint8_t *a_img, const uint8_t *b_img, float *d_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar_load, bar_peer, bar_mma;
    __shared__ uint32_t tmem_base;
    const uint32_t rank = cluster_ctarank();
    const uint32_t a_bytes = ROWS_A * K * 2, bh_bytes = (N / 2) * K * 2;
    uint8_t *sa = smem, *sb = smem + 32768;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar_load, 1); mbar_init(&bar_peer, 1); mbar_init(&bar_mma, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc2(&tmem_base, 256); tmem_relinquish2(); }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tm = tmem_base;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar_load, a_bytes + bh_bytes);
        bulk_g2s(sa, a_img + (size_t)rank * a_bytes, a_bytes, &bar_load);          // this CTA's A image
        bulk_g2s(sb, b_img + (size_t)rank * bh_bytes, bh_bytes, &bar_load);        // this CTA's half of B (rows 64*rank ..)
        mbar_wait(&bar_load, 0);
        if (rank == 1) mbar_arrive_remote(mapa_u32(smem_u32(&bar_peer), 0));       // tell the leader
    }
    if (rank == 0 && threadIdx.x == 0) {
        mbar_wait_cluster(&bar_peer, 0);
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(256, N, 0, 0);
        for (int k = 0; k < K / 16; ++k) {
            const uint64_t ad = make_desc(smem_u32(sa) + SHIFT * 16 + k * 2 * ROWS_A * 16, ROWS_A * 16, 128);
            const uint64_t bd = make_desc(smem_u32(sb) + k * 2 * (N / 2) * 16, (N / 2) * 16, 128);
            mma_bf16_pair(tm, ad, bd, idesc, k > 0);
        }
        mma_commit_pair(&bar_mma, 3);
    }
    __syncthreads();
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) d_out[((size_t)rank * 128 + threadIdx.x) * N + c0 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 0) tmem_dealloc2(tm, 256);
}

static float bf(uint16_t v) { uint32_t u = (uint32_t)v << 16; float f; memcpy(&f, &u, 4); return f; }
static uint16_t tobf(float f) { __nv_bfloat16 b = __float2bfloat16(f); uint16_t u; memcpy(&u, &b, 2); return u; }

int main() {
    // A: two images (one per CTA) [K/8][ROWS_A][8]; B: [2 halves][K/8][N/2][8]
    std::vector<uint16_t> ia((size_t)2 * ROWS_A * K), ib((size_t)N * K);
    std::vector<float> fa(ia.size()), fb(ib.size());
    srand(7);
    for (int r = 0; r < 2; ++r)
        for (int row = 0; row < ROWS_A; ++row)
            for (int c = 0; c < K; ++c) {
                const float v = (float)((rand() % 17) - 8) / 8.f;
                fa[((size_t)r * ROWS_A + row) * K + c] = v;
                ia[(size_t)r * ROWS_A * K + ((size_t)(c / 8) * ROWS_A + row) * 8 + c % 8] = tobf(v);
            }
    for (int n = 0; n < N; ++n)
        for (int c = 0; c < K; ++c) {
            const float v = (float)((rand() % 13) - 6) / 4.f;
            fb[(size_t)n * K + c] = v;
            const int h = n / (N / 2), nn = n % (N / 2);
            ib[(size_t)h * (N / 2) * K + ((size_t)(c / 8) * (N / 2) + nn) * 8 + c % 8] = tobf(v);
        }
    uint8_t *da, *db; float *dd;
    cudaMalloc(&da, ia.size() * 2); cudaMalloc(&db, ib.size() * 2); cudaMalloc(&dd, 2 * 128 * N * 4);
    cudaMemcpy(da, ia.data(), ia.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, ib.data(), ib.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dd, 0xff, 2 * 128 * N * 4);
    cudaFuncSetAttribute(pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    pair_kernel<<<2, 128, 65536>>>(da, db, dd);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("pair probe: CUDA ERROR %s\n", cudaGetErrorString(e)); return 2; }
    std::vector<float> got((size_t)2 * 128 * N);
    cudaMemcpy(got.data(), dd, got.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0; double maxerr = 0;
    for (int r = 0; r < 2; ++r)
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < N; ++n) {
                float s = 0;
                for (int k = 0; k < K; ++k) s += bf(tobf(fa[((size_t)r * ROWS_A + SHIFT + m) * K + k])) * bf(tobf(fb[(size_t)n * K + k]));
                const double d = fabs((double)s - got[((size_t)r * 128 + m) * N + n]);
                if (!(d <= 1e-3)) bad++;
                if (d > maxerr) maxerr = d;
            }
    printf("pair probe (cta_group::2, M=256, N=%d, B split in halves, A row shift %d): %s maxerr=%g bad=%d\n", N, SHIFT, bad ? "FAIL" : "PASS", maxerr, bad);
    return bad != 0;
}
