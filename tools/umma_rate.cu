// Dev tool: tcgen05.mma issue-rate microbenchmark (timing only; operands are whatever is in shared memory).
// Answers: how many cycles does one cta_group::1 MMA take for a given N / layout / accumulator pattern?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_rate tools/umma_rate.cu
#include "../deep-go_b200/csrc/umma.cuh"
#include <cstdio>
#include <vector>
using namespace umma;
__device__ __forceinline__ uint32_t b_lo_dummy(uint32_t x) { return x; }

struct RCfg { int N, nmma, nacc, a_mn, b_mn, layout, a_lbo, a_sbo, b_lbo, b_sbo, a_stride, b_stride, a_span, b_span, a_off, group, gap; };

__global__ void __launch_bounds__(128) rate_kernel(RCfg c, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar, bar2;
    __shared__ uint32_t tmem_base;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&bar2, 1 << 20); fence_barrier_init(); }
    if (threadIdx.x < 32) { tmem_alloc(&tmem_base, 512); tmem_relinquish(); }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x < 32) {
        // whole warp converged; one elected lane issues (canonical pattern)
        const uint32_t idesc = make_idesc_bf16(128, c.N, c.a_mn, c.b_mn);
        const uint32_t sa = smem_u32(smem), sb = sa + 100 * 1024;
        uint32_t hi_a = desc_hi(c.a_sbo) | ((uint32_t)c.layout << 29), hi_b = desc_hi(c.b_sbo) | ((uint32_t)c.layout << 29);
        uint32_t alo[8], blo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { alo[j] = desc_lo(sa + c.a_off + (j % c.a_span) * c.a_stride, c.a_lbo); blo[j] = desc_lo(sb + (j % c.b_span) * c.b_stride, c.b_lbo); }
        const long long t0 = clock64();
        if (elect_one()) {
            for (int i = 0; i < c.nmma; i += 8) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    mma_bf16(tmem_base + (j % c.nacc) * c.N, desc_join(alo[j], hi_a), desc_join(b_lo_dummy(blo[j]), hi_b), idesc, (i | (j >= c.nacc)) != 0);
                if (c.group && ((i + 8) % c.group) == 0 && i + 8 < c.nmma) {
                    // stage boundary as in the conv kernel: commit to a barrier nobody waits on, then `gap` cycles of scalar work
                    mma_commit(&bar2);
                    const long long g0 = clock64();
                    while (clock64() - g0 < c.gap) { }
                }
            }
            mma_commit(&bar);
        }
        __syncwarp();
        const long long t1 = clock64();
        mbar_wait(&bar, 0);
        const long long t2 = clock64();
        if (threadIdx.x == 0) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem_base, 512);
}

static void run(const char *name, RCfg c, int grid) {
    long long *d;
    cudaMalloc(&d, grid * 16);
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int rep = 0; rep < 2; ++rep) rate_kernel<<<grid, 128, 200 * 1024>>>(c, d);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<long long> h(grid * 2);
    cudaMemcpy(h.data(), d, grid * 16, cudaMemcpyDeviceToHost);
    double issue = 0, done = 0;
    for (int i = 0; i < grid; ++i) { issue += h[2 * i]; done += h[2 * i + 1]; }
    printf("%-58s grid %3d: %7.1f cyc/MMA issue, %7.1f cyc/MMA complete  (%s)\n", name, grid, issue / grid / c.nmma, done / grid / c.nmma,
           e == cudaSuccess ? "ok" : cudaGetErrorString(e));
    cudaFree(d);
}

int main(int argc, char **argv) {
    const int R = 432;
    if (argc > 1) {
        // weight-gradient operand geometry: MN-major, chunk stride (SBO) = rows per stage x 16 B, with and without one row of padding
        for (int grid : {1, 148})
            for (int a_rows : {128, 129, 132, 136})
                for (int b_rows : {136, 137, 140, 132}) {
                    RCfg m = {128, 512, 3, 1, 1, 0, 128, a_rows * 16, 128, b_rows * 16, 256, 256, 8, 8};
                    char nm[96]; snprintf(nm, sizeof nm, "MN-major N=128 3 acc, A chunk stride %d B, B chunk stride %d B", a_rows * 16, b_rows * 16);
                    run(nm, m, grid);
                }
        return 0;
    }
    // K-major no-swizzle blocked image: LBO = R*16 (chunk stride), SBO = 128, k-step = 2 chunks
    RCfg k128 = {128, 512, 1, 0, 0, 0, R * 16, 128, 128 * 16, 128, 2 * R * 16, 2 * 128 * 16, 4, 4};
    for (int grid : {1, 148}) {
        run("K-major noswz N=128 1 acc", k128, grid);
        RCfg c = k128; c.nacc = 3; run("K-major noswz N=128 3 acc round-robin", c, grid);
        c = k128; c.N = 256; c.b_lbo = 256 * 16; c.b_stride = 2 * 256 * 16; run("K-major noswz N=256 1 acc", c, grid);
        c = k128; c.N = 64; c.b_lbo = 64 * 16; c.b_stride = 2 * 64 * 16; run("K-major noswz N=64 1 acc", c, grid);
        c = k128; c.a_stride = 0; c.b_stride = 0; run("K-major noswz N=128 same operands every MMA", c, grid);
        // 128B swizzle K-major (canonical GEMM layout): rows of 128 B, SBO = 1024, k-step = 32 B inside the atom
        RCfg s = {128, 512, 1, 0, 0, 2, 16, 1024, 16, 1024, 32, 32, 4, 4};
        run("K-major SW128 N=128 1 acc", s, grid);
        s.N = 256; run("K-major SW128 N=256 1 acc", s, grid);
        // MN-major no-swizzle (weight gradient): LBO = 128 (K groups), SBO = chunk stride, k-step = 256 B
        RCfg m = {128, 512, 1, 1, 1, 0, 128, 64 * 16, 128, 72 * 16, 256, 256, 4, 4};
        run("MN-major noswz N=128 1 acc", m, grid);
        m.nacc = 3; run("MN-major noswz N=128 3 acc round-robin", m, grid);
        c = k128; c.a_off = 16; run("K-major noswz N=128, A start +1 row (misaligned)", c, grid);
        c = k128; c.a_off = 21 * 16; run("K-major noswz N=128, A start +21 rows", c, grid);
        m = RCfg{128, 512, 1, 1, 1, 0, 128, 64 * 16, 128, 72 * 16, 256, 256, 4, 4, 16}; run("MN-major noswz N=128, A start +1 row", m, grid);
        s = RCfg{128, 512, 1, 0, 0, 2, 16, 1024, 16, 1024, 32, 32, 4, 4, 128}; run("K-major SW128 N=128, A start +1 row (128 B)", s, grid);
        s.a_off = 21 * 128; run("K-major SW128 N=128, A start +21 rows", s, grid);
        s = RCfg{128, 512, 1, 1, 1, 2, 8192, 1024, 8192, 1024, 2048, 2048, 4, 4, 128}; run("MN-major SW128 N=128, A start +1 row", s, grid);
        // stage boundaries: a commit every `group` MMAs, followed by `gap` cycles in which the issuing thread does something else
        for (int gap : {0, 100, 200, 400, 800}) {
            RCfg g = k128; g.nmma = 480; g.nacc = 3; g.group = 24; g.gap = gap;
            char nm[96]; snprintf(nm, sizeof nm, "K-major N=128: commit every 24 MMAs + %d-cycle gap", gap);
            run(nm, g, grid);
        }
        { RCfg g = k128; g.nmma = 480; g.nacc = 3; g.group = 8; g.gap = 0; run("K-major N=128: commit every 8 MMAs, no gap", g, grid); }
    }
    return 0;
}
