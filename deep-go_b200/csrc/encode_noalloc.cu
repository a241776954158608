// dg_encode_noalloc: the feature-plane encoder for callers that must not allocate device memory.
//
// Who needs it.  dg_encode stages its host buffers through a grow-only device buffer that it cudaMalloc's on
// demand.  Some users of this library cannot afford that call:
//   * loader threads that encode small batches while a training step is in flight on the same GPU -- cudaMalloc
//     and cudaFree synchronise the whole device, so one allocation in a loader stalls the step it runs beside;
//   * processes where another framework's caching allocator has already reserved (nearly) all of HBM, so a
//     cudaMalloc from this library simply fails.
// For them the staging memory lives in the library image itself: a fixed, file-scope __device__ array.  Every GPU
// that loads the library carries these DG_NOALLOC_SCRATCH_BYTES (about 3.7 MB) until the process exits, which is
// why the limit is a modest 64 positions per call.
//
// Rules for callers
//   * n <= DG_NOALLOC_MAX_POSITIONS (64); larger batches are rejected with an error, not truncated.
//   * The first call on a device looks up the device address of the array (cudaGetSymbolAddress) and caches it;
//     make that first call before any stream capture starts.
//   * All calls on a device share the one array.  The entry point synchronises before it returns, so calls made
//     one after another are safe; two calls that are in flight on the same GPU at the same time (two threads, two
//     contexts) overwrite each other's staging data -- serialise them yourself.
//   * The array keeps the previous call's bytes.  Nothing in it has to start at zero: each call copies in exactly
//     the n records it will read and the kernel writes every output element of those n positions before they are
//     copied back, so stale data beyond position n is never read.
//   * No cudaMalloc, cudaFree or other allocating routine is called on this path.
#include "common.cuh"
#include "../../include/deepgo.h"
#include <mutex>

#define DG_NOALLOC_MAX 64
// layout inside the array: raw records | player,rank,sym bytes | output planes (fp32 is the larger format)
#define NA_RAW_BYTES ((size_t)DG_NOALLOC_MAX * DG_RAW_PLANES * DG_POINTS)
#define NA_META_OFF ((NA_RAW_BYTES + 255) / 256 * 256)
#define NA_OUT_OFF ((NA_META_OFF + 3 * DG_NOALLOC_MAX + 255) / 256 * 256)
#define NA_OUT_BYTES ((size_t)DG_NOALLOC_MAX * DG_PLANES * DG_POINTS * 4)
#define DG_NOALLOC_SCRATCH_BYTES (NA_OUT_OFF + NA_OUT_BYTES)

static __device__ __align__(256) unsigned char g_noalloc_scratch[DG_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address; the address each device
// uses is fetched once with cudaGetSymbolAddress and kept here.
static unsigned char *g_scratch_addr[64] = {nullptr};
static std::mutex g_scratch_mu;

static int scratch_for_device(int dev, unsigned char **out) {
    DG_REQUIRE(dev >= 0 && dev < 64, "dg_encode_noalloc: device index %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        DG_CUDA(cudaSetDevice(dev));                                   // the lookup is per device: make it current first
        DG_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        DG_REQUIRE(p != nullptr, "dg_encode_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[dev] = (unsigned char *)p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

// same plane definitions as encode.cu (dataloader.lua:50-92); kept local so this file stands alone
__device__ __forceinline__ int na_src_point(int p, int sym) {
    int x = p / DG_BOARD, y = p - x * DG_BOARD;
    if (sym & 2) x = DG_BOARD - 1 - x;
    if (sym & 4) y = DG_BOARD - 1 - y;
    if (sym & 1) { int t = x; x = y; y = t; }
    return x * DG_BOARD + y;
}
__device__ __forceinline__ unsigned long long na_point_planes(const uint8_t *__restrict__ rec, int src, int player, int rank) {
    const int stone = rec[src], lib = rec[DG_POINTS + src], la = rec[(2 + player - 1) * DG_POINTS + src];
    const int kills = rec[(4 + player - 1) * DG_POINTS + src], age = rec[6 * DG_POINTS + src], ladder = rec[(7 + player - 1) * DG_POINTS + src];
    unsigned long long m = 0;
    m |= (unsigned long long)(stone == 0) << 0;
    m |= (unsigned long long)(stone == player) << 1;
    m |= (unsigned long long)(stone == 3 - player) << 2;
    if (lib >= 1 && lib <= 3) m |= 1ull << (3 + lib - 1);
    m |= (unsigned long long)(lib >= 4) << 6;
    m |= (unsigned long long)(stone == 0 && la == 0) << 7;
    if (la >= 1 && la <= 5) m |= 1ull << (7 + la);
    m |= (unsigned long long)(la >= 6) << 13;
    if (kills >= 1 && kills <= 6) m |= 1ull << (14 + kills - 1);
    m |= (unsigned long long)(kills >= 7) << 20;
    if (age >= 1 && age <= 5) m |= 1ull << (21 + age - 1);
    m |= (unsigned long long)(ladder >= 1) << 26;
    m |= 1ull << (27 + rank);
    return m;
}

// One block per position.  `scratch` is the device address of g_noalloc_scratch handed over by the host; records and
// per-position bytes are read from its front, planes are written to its tail (fmt 0: fp32 NCHW, fmt 1: bit-planes).
__global__ void __launch_bounds__(384) encode_noalloc_kernel(unsigned char *__restrict__ scratch, int use_sym, int fmt) {
    const int n = blockIdx.x, p = threadIdx.x;
    const uint8_t *raw = scratch + (size_t)n * DG_RAW_PLANES * DG_POINTS;
    const uint8_t *meta = scratch + NA_META_OFF;
    const int player = meta[n], rank = meta[DG_NOALLOC_MAX + n], sym = use_sym ? meta[2 * DG_NOALLOC_MAX + n] : 0;
    unsigned long long m = 0;
    if (p < DG_POINTS) m = na_point_planes(raw, na_src_point(p, sym), player, rank);
    if (fmt == 0) {
        if (p >= DG_POINTS) return;
        float *o = reinterpret_cast<float *>(scratch + NA_OUT_OFF) + (size_t)n * DG_PLANES * DG_POINTS + p;
#pragma unroll
        for (int c = 0; c < DG_PLANES; ++c) o[c * DG_POINTS] = (float)((m >> c) & 1ull);
    } else {
        uint32_t *o = reinterpret_cast<uint32_t *>(scratch + NA_OUT_OFF) + (size_t)n * DG_PLANES * 12 + (p >> 5);
#pragma unroll
        for (int c = 0; c < DG_PLANES; ++c) {
            const unsigned w = __ballot_sync(0xffffffffu, (m >> c) & 1ull);
            if ((p & 31) == 0) o[c * 12] = w;
        }
    }
}

extern "C" int dg_encode_noalloc(dg_ctx *c, const uint8_t *raw, const uint8_t *player, const uint8_t *rank, const uint8_t *sym,
                                 int n, void *out, int outfmt) {
    DG_REQUIRE(c, "dg_encode_noalloc: ctx is NULL");
    DG_REQUIRE(n >= 0, "dg_encode_noalloc: n < 0");
    DG_REQUIRE(n <= DG_NOALLOC_MAX, "dg_encode_noalloc: %d positions exceed the fixed staging area (at most %d per call; use dg_encode for larger batches)",
               n, DG_NOALLOC_MAX);
    if (n == 0) return 0;
    DG_REQUIRE(raw && player && rank && out, "dg_encode_noalloc: NULL buffer");
    DG_REQUIRE(outfmt == DG_FMT_F32_NCHW || outfmt == DG_FMT_BITPLANES, "dg_encode_noalloc: unknown output format %d", outfmt);
    for (int i = 0; i < n; ++i) {
        DG_REQUIRE(player[i] == 1 || player[i] == 2, "record %d: player %d not in {1,2}", i, player[i]);
        DG_REQUIRE(rank[i] >= 1 && rank[i] <= 9, "record %d: rank %d outside 1..9 (dataloader.lua:87 would index out of range)", i, rank[i]);
        DG_REQUIRE(!sym || sym[i] < 8, "record %d: symmetry id %d outside 0..7", i, sym[i]);
    }
    unsigned char *d = nullptr;
    if (int rc = scratch_for_device(c->dev, &d)) return rc;
    DG_CUDA(cudaSetDevice(c->dev));
    DG_CUDA(cudaMemcpyAsync(d, raw, (size_t)n * DG_RAW_PLANES * DG_POINTS, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + NA_META_OFF, player, n, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + NA_META_OFF + DG_NOALLOC_MAX, rank, n, cudaMemcpyHostToDevice, c->stream));
    if (sym) DG_CUDA(cudaMemcpyAsync(d + NA_META_OFF + 2 * DG_NOALLOC_MAX, sym, n, cudaMemcpyHostToDevice, c->stream));
    encode_noalloc_kernel<<<n, 384, 0, c->stream>>>(d, sym != nullptr, outfmt == DG_FMT_F32_NCHW ? 0 : 1);
    c->launches++;
    DG_CUDA(cudaGetLastError());
    const size_t out_b = outfmt == DG_FMT_F32_NCHW ? (size_t)n * DG_PLANES * DG_POINTS * 4 : (size_t)n * DG_PLANES * 12 * 4;
    DG_CUDA(cudaMemcpyAsync(out, d + NA_OUT_OFF, out_b, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
