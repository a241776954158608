// Stand-alone hardware probe (dev tool, not part of the library): checks, on a real B200, the
// descriptor conventions the conv kernels rely on -- no-swizzle K-major and MN-major smem
// descriptors over the blocked [chunk][row][8] bf16 image, arbitrary row shifts of the start
// address (the 3x3 taps), the instruction descriptor, and the TMEM lane/column mapping.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o gpurun_out/umma_probe tools/umma_probe.cu
#include "../deep-go_b200/csrc/umma.cuh"
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

using namespace umma;

struct Cfg {
    int layout, base_off_mode;   // layout_type (0 none, 2 = 128B swizzle); 1 = set descriptor base_offset from the start address
    int N;                 // MMA N (<=256), M fixed 128
    int ksteps;            // MMAs issued (K = 16 each)
    int a_mn, b_mn;        // 1 = MN-major
    uint32_t a_off, a_lbo, a_sbo, a_step;   // bytes
    uint32_t b_off, b_lbo, b_sbo, b_step;
    uint32_t a_bytes, b_bytes;              // image sizes to copy
};

__global__ void __launch_bounds__(128) probe_kernel(const uint8_t *__restrict__ a_img, const uint8_t *__restrict__ b_img, float *__restrict__ d_out, Cfg c) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar_load, bar_mma;
    __shared__ uint32_t tmem_base;
    uint8_t *sa = smem, *sb = smem + ((c.a_bytes + 1023) / 1024) * 1024;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(&tmem_base, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_base;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar_load, c.a_bytes + c.b_bytes);
        bulk_g2s(sa, a_img, c.a_bytes, &bar_load);
        bulk_g2s(sb, b_img, c.b_bytes, &bar_load);
        mbar_wait(&bar_load, 0);
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(128, c.N, c.a_mn, c.b_mn);
        for (int k = 0; k < c.ksteps; ++k) {
            uint64_t ad = make_desc(smem_u32(sa) + c.a_off + k * c.a_step, c.a_lbo, c.a_sbo);
            uint64_t bd = make_desc(smem_u32(sb) + c.b_off + k * c.b_step, c.b_lbo, c.b_sbo);
            ad |= (uint64_t)c.layout << 61; bd |= (uint64_t)c.layout << 61;
            if (c.base_off_mode) {
                ad |= (uint64_t)(((smem_u32(sa) + c.a_off + k * c.a_step) >> 7) & 7) << 49;
                bd |= (uint64_t)(((smem_u32(sb) + c.b_off + k * c.b_step) >> 7) & 7) << 49;
            }
            mma_bf16(tm, ad, bd, idesc, k > 0);
        }
        mma_commit(&bar_mma);
    }
    __syncthreads();
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    for (int c0 = 0; c0 + 32 <= c.N; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + c0, r);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) d_out[(size_t)threadIdx.x * c.N + c0 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 256);
}

static float bf(uint16_t v) { uint32_t u = (uint32_t)v << 16; float f; memcpy(&f, &u, 4); return f; }
static uint16_t tobf(float f) { __nv_bfloat16 b = __float2bfloat16(f); uint16_t u; memcpy(&u, &b, 2); return u; }

// blocked image: element (row r, channel c) at ((c/8)*rows + r)*8 + c%8
static int run_case(const char *name, int N, int rows_a, int ch_a, int rows_b, int ch_b, int a_mn, int b_mn, int shift, int K) {
    // logical problem: D[m][n] = sum_k A(m,k) B(n,k)
    //   K-major  A: A(m,k) = imgA[row = shift+m][ch = k]        (conv fwd: rows = positions, K = channels)
    //   MN-major A: A(m,k) = imgA[row = shift+k][ch = m]        (wgrad: K = rows, M = channels)
    std::vector<uint16_t> ia((size_t)rows_a * ch_a), ib((size_t)rows_b * ch_b);
    srand(1234);
    for (auto &v : ia) v = tobf((float)((rand() % 17) - 8) / 8.f);
    for (auto &v : ib) v = tobf((float)((rand() % 13) - 6) / 4.f);
    auto A = [&](int r, int c) { return bf(ia[((size_t)(c / 8) * rows_a + r) * 8 + c % 8]); };
    auto B = [&](int r, int c) { return bf(ib[((size_t)(c / 8) * rows_b + r) * 8 + c % 8]); };
    Cfg c{};
    c.N = N; c.ksteps = K / 16; c.a_mn = a_mn; c.b_mn = b_mn;
    c.a_bytes = (uint32_t)ia.size() * 2; c.b_bytes = (uint32_t)ib.size() * 2;
    if (!a_mn) { c.a_off = shift * 16; c.a_lbo = rows_a * 16; c.a_sbo = 128; c.a_step = 2 * rows_a * 16; }
    else       { c.a_off = shift * 16; c.a_lbo = 128; c.a_sbo = rows_a * 16; c.a_step = 256; }
    if (!b_mn) { c.b_off = 0; c.b_lbo = rows_b * 16; c.b_sbo = 128; c.b_step = 2 * rows_b * 16; }
    else       { c.b_off = 0; c.b_lbo = 128; c.b_sbo = rows_b * 16; c.b_step = 256; }
    std::vector<float> ref((size_t)128 * N, 0.f), got((size_t)128 * N, -1.f);
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            float s = 0;
            for (int k = 0; k < K; ++k) {
                const float a = a_mn ? A(shift + k, m) : A(shift + m, k);
                const float b = b_mn ? B(k, n) : B(n, k);
                s += a * b;
            }
            ref[(size_t)m * N + n] = s;
        }
    uint8_t *da, *db; float *dd;
    cudaMalloc(&da, c.a_bytes); cudaMalloc(&db, c.b_bytes); cudaMalloc(&dd, got.size() * 4);
    cudaMemcpy(da, ia.data(), c.a_bytes, cudaMemcpyHostToDevice);
    cudaMemcpy(db, ib.data(), c.b_bytes, cudaMemcpyHostToDevice);
    cudaMemset(dd, 0xff, got.size() * 4);
    const size_t smem = ((c.a_bytes + 1023) / 1024) * 1024 + c.b_bytes + 1024;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    probe_kernel<<<1, 128, smem>>>(da, db, dd, c);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-40s CUDA ERROR %s\n", name, cudaGetErrorString(e)); return 2; }
    cudaMemcpy(got.data(), dd, got.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int bad = 0;
    for (size_t i = 0; i < ref.size(); ++i) { double d = fabs((double)ref[i] - got[i]); if (!(d <= 1e-3)) bad++; if (d > maxerr) maxerr = d; }
    printf("%-40s %s  maxerr=%g bad=%d/%zu  (ref[5]=%g got[5]=%g)\n", name, bad ? "FAIL" : "PASS", maxerr, bad, ref.size(), ref[5], got[5]);
    cudaFree(da); cudaFree(db); cudaFree(dd);
    return bad != 0;
}

// 128B-swizzled panel image: element (row r, channel c) of a 64-channel panel P = c/64 at
//   ((P*rows + r)*8 + (((c%64)/8) ^ (r & 7)))*8 + c%8   -- rows of 128 B, 16-byte chunks XOR-ed with the row index
static int run_sw128(const char *name, int N, int rows_a, int ch_a, int rows_b, int ch_b, int a_mn, int b_mn, int shift, int K, int base_off_mode) {
    std::vector<uint16_t> ia((size_t)rows_a * ch_a), ib((size_t)rows_b * ch_b);
    srand(4321);
    auto idx = [](int rows, int r, int c) { return ((size_t)((c / 64) * rows + r) * 8 + (((c % 64) / 8) ^ (r & 7))) * 8 + c % 8; };
    std::vector<float> fa((size_t)rows_a * ch_a), fb((size_t)rows_b * ch_b);
    for (int r = 0; r < rows_a; ++r) for (int c = 0; c < ch_a; ++c) { float v = (float)((rand() % 17) - 8) / 8.f; fa[(size_t)r * ch_a + c] = v; ia[idx(rows_a, r, c)] = tobf(v); }
    for (int r = 0; r < rows_b; ++r) for (int c = 0; c < ch_b; ++c) { float v = (float)((rand() % 13) - 6) / 4.f; fb[(size_t)r * ch_b + c] = v; ib[idx(rows_b, r, c)] = tobf(v); }
    Cfg c{};
    c.layout = 2; c.base_off_mode = base_off_mode;
    c.N = N; c.ksteps = K / 16; c.a_mn = a_mn; c.b_mn = b_mn;
    c.a_bytes = (uint32_t)ia.size() * 2; c.b_bytes = (uint32_t)ib.size() * 2;
    // K-major: rows = M/N index, K = channels: SBO = 1024 (8 rows), k-step = 32 B inside the 128 B row (64 channels per panel)
    // MN-major: rows = K index, M/N = channels: LBO = panel stride (next 64 channels), SBO = 1024 (8 K rows), k-step = 16 rows = 2048 B
    if (!a_mn) { c.a_off = shift * 128; c.a_lbo = 16; c.a_sbo = 1024; c.a_step = 32; }
    else       { c.a_off = shift * 128; c.a_lbo = rows_a * 128; c.a_sbo = 1024; c.a_step = 2048; }
    if (!b_mn) { c.b_off = 0; c.b_lbo = 16; c.b_sbo = 1024; c.b_step = 32; }
    else       { c.b_off = 0; c.b_lbo = rows_b * 128; c.b_sbo = 1024; c.b_step = 2048; }
    std::vector<float> ref((size_t)128 * N, 0.f), got((size_t)128 * N, -1.f);
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            float s = 0;
            for (int k = 0; k < K; ++k) {
                const float a = a_mn ? fa[(size_t)(shift + k) * ch_a + m] : fa[(size_t)(shift + m) * ch_a + k];
                const float b = b_mn ? fb[(size_t)k * ch_b + n] : fb[(size_t)n * ch_b + k];
                s += a * b;
            }
            ref[(size_t)m * N + n] = s;
        }
    uint8_t *da, *db; float *dd;
    cudaMalloc(&da, c.a_bytes); cudaMalloc(&db, c.b_bytes); cudaMalloc(&dd, got.size() * 4);
    cudaMemcpy(da, ia.data(), c.a_bytes, cudaMemcpyHostToDevice);
    cudaMemcpy(db, ib.data(), c.b_bytes, cudaMemcpyHostToDevice);
    cudaMemset(dd, 0xff, got.size() * 4);
    const size_t smem = ((c.a_bytes + 1023) / 1024) * 1024 + c.b_bytes + 1024;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    probe_kernel<<<1, 128, smem>>>(da, db, dd, c);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-52s CUDA ERROR %s\n", name, cudaGetErrorString(e)); return 2; }
    cudaMemcpy(got.data(), dd, got.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int bad = 0;
    for (size_t i = 0; i < ref.size(); ++i) { double d = fabs((double)ref[i] - got[i]); if (!(d <= 1e-3)) bad++; if (d > maxerr) maxerr = d; }
    printf("%-52s %s  maxerr=%g bad=%d/%zu\n", name, bad ? "FAIL" : "PASS", maxerr, bad, ref.size());
    cudaFree(da); cudaFree(db); cudaFree(dd);
    return bad != 0;
}

int main() {
    int fails = 0;
    for (int bo = 0; bo < 2; ++bo) {
        printf("--- 128B swizzle, base_offset %s\n", bo ? "set from start address" : "left 0");
        run_sw128("SW128 Kmajor shift0 N128 K64", 128, 192, 64, 128, 64, 0, 0, 0, 64, bo);
        run_sw128("SW128 Kmajor shift1 N128 K64", 128, 192, 64, 128, 64, 0, 0, 1, 64, bo);
        run_sw128("SW128 Kmajor shift8 N128 K64", 128, 192, 64, 128, 64, 0, 0, 8, 64, bo);
        run_sw128("SW128 Kmajor shift21 N128 K64", 128, 192, 64, 128, 64, 0, 0, 21, 64, bo);
        run_sw128("SW128 MNmajor A,B shift0 N128 K64rows ch128", 128, 96, 128, 96, 128, 1, 1, 0, 64, bo);
        run_sw128("SW128 MNmajor A,B shift5 N128 K64rows ch128", 128, 96, 128, 96, 128, 1, 1, 5, 64, bo);
        run_sw128("SW128 MNmajor A shift21 N64 ch128/64", 64, 96, 128, 96, 64, 1, 1, 21, 64, bo);
    }
    // conv forward shape: A = position image (rows x 64ch) K-major, B = weights (N rows x 64 ch) K-major
    fails += run_case("Kmajor shift0 N128 K64", 128, 192, 64, 128, 64, 0, 0, 0, 64);
    fails += run_case("Kmajor shift1 N128 K64", 128, 192, 64, 128, 64, 0, 0, 1, 64);
    fails += run_case("Kmajor shift21 N128 K64", 128, 192, 64, 128, 64, 0, 0, 21, 64);
    fails += run_case("Kmajor shift42 N64 K128", 64, 192, 128, 64, 128, 0, 0, 42, 128);
    fails += run_case("Kmajor shift3 N192 K32", 192, 160, 32, 192, 32, 0, 0, 3, 32);
    fails += run_case("Kmajor shift7 N16 K48", 16, 160, 48, 16, 48, 0, 0, 7, 48);
    // wgrad shape: A = dZ image (rows x 128 ch) MN-major (M = channel), B = X image (rows x N ch) MN-major, K = rows
    fails += run_case("MNmajor A,B shift0 N128 K64rows", 128, 96, 128, 96, 128, 1, 1, 0, 64);
    fails += run_case("MNmajor A,B shift5 N128 K64rows", 128, 96, 128, 96, 128, 1, 1, 5, 64);
    fails += run_case("MNmajor A shift21, B Kmajor N64", 64, 96, 128, 64, 64, 1, 0, 21, 64);
    printf("probe: %d failing case(s)\n", fails);
    return fails ? 1 : 0;
}
