// DG_PREC_BF16 path: the conv stack of experiments.lua:133-153 as implicit-GEMM tcgen05 kernels.
//
// Activation layout in HBM ("blocked image"): per position  [C/8][R][8] bf16, R rows of 8 channels
// (16 B).  Row index rho = x'*pitch + y' over a zero-padded board whose right pad column(s) double
// as the next board row's left pad (pitch = 19 + pad): a k x k tap (ky,kx) is then the SAME image
// read at row offset (ky-pad)*pitch + (kx-pad).  With the no-swizzle UMMA canonical layout
// (8 rows x 16 B core matrices) that is just a different descriptor start address, so one
// bulk-copied image in shared memory feeds all k*k taps: no im2col, no re-fetch per tap.
//   forward / input-gradient:  D[rho][co] += A(image rows, K-major) x B(weights [ci/8][co][8], K-major)
//   weight-gradient:           D[co][ci]  += A(dZ image, MN-major)  x B(X image shifted, MN-major)
// Accumulators live in TMEM; the epilogue fuses bias + per-point bias + ReLU (forward) or the
// ReLU mask (backward) and writes bf16 back in the same blocked layout.
#include "model.cuh"
#include "umma.cuh"
#include "../../include/deepgo.h"
#include <cstdio>
#include <cstdlib>
#include <algorithm>

using namespace umma;
typedef __nv_bfloat16 bf16;

#define DG_R3 432      // image rows, 3x3 geometry: pitch 20, pad 1 (rows 0..420 used, MMA tiles read up to 425)
#define DG_R5 600      // image rows, 5x5 geometry: pitch 21, pad 2 (rows 0..484 used, MMA tiles read up to 599)

struct Geom { int pitch, pad, rows, first, ntiles; };
static Geom geom_for(int k) {
    Geom g;
    g.pad = k / 2; g.pitch = DG_BOARD + g.pad; g.rows = k == 5 ? DG_R5 : DG_R3;
    g.first = g.pad * g.pitch + g.pad;
    g.ntiles = k == 5 ? 4 : 3;
    return g;
}

// ---------------------------------------------------------------------------------------------
// conv forward / input gradient
// ---------------------------------------------------------------------------------------------
#define CONV_MAX_CHAIN 16
struct ConvParams {
    const bf16 *in;                    // [n][cin/8][R_in][8]  input image of the first layer of the chain
    // A chain is `nchain` consecutive layers run by one kernel with the activation image resident in shared memory:
    // layer j's epilogue writes its output rows back over the image (all MMAs of layer j are complete by then) and
    // layer j+1 starts from there -- no launch, no image reload, the weight ring keeps streaming across layers.
    // nchain > 1 needs cin == cout <= 128 (one N block, one K panel) and the 3x3 geometry in and out.
    const bf16 *w[CONV_MAX_CHAIN];     // [taps][cin/8][cout][8]
    bf16 *out[CONV_MAX_CHAIN];         // [n][cout/8][R_out][8]   every layer's output also goes to HBM (backward needs it)
    const bf16 *bias[CONV_MAX_CHAIN];  // mode 0: [cout/8][R_out][8] bf16, b[c] + a[c][x][y] (16 bytes per row and channel chunk)
    const bf16 *mask[CONV_MAX_CHAIN];  // mode 1: [n][cout/8][R_mask][8], output kept where mask > 0
    int nchain;
    int n, cin, cout;
    unsigned long long *tl; int tl_id;
    // first layer fed from raw records: the MMA warps build the input image in shared memory themselves (and store it to enc_out for
    // the weight gradient) instead of waiting for a bulk copy of an image a separate encode kernel wrote
    const uint8_t *enc_raw, *enc_player, *enc_rank, *enc_sym;
    bf16 *enc_out;
    int R_in, pitch_in, pad_in, ksize, first, ntiles;
    int R_out, pitch_out, pad_out;     // geometry of the LAST layer's HBM output (intermediate layers: same as the input)
    int R_mask, pitch_mask, pad_mask;
    int nb;               // N block (<= 128, multiple of 32); cout % nb == 0
    int kp;               // channels per weight ring stage (multiple of 16); cin % kp == 0
    int nstages;
    int mode;
    long long *dbg;       // optional per-CTA phase timestamps (DG_TIMELINE=1), 8 slots per CTA
};

__device__ __forceinline__ int dg_src_point_d(int p, int sym) {
    int x = p / DG_BOARD, y = p - x * DG_BOARD;
    if (sym & 2) x = DG_BOARD - 1 - x;
    if (sym & 4) y = DG_BOARD - 1 - y;
    if (sym & 1) { int t = x; x = y; y = t; }
    return x * DG_BOARD + y;
}

// preprocess() of one point (dataloader.lua:50-92): the 37 one-hot planes (+ zero padding to 48) as a bit mask, plane i = bit i
__device__ __forceinline__ unsigned long long dg_plane_bits(const uint8_t *__restrict__ rec, int src, int pl, int rk) {
    const int stone = rec[src], lib = rec[DG_POINTS + src], la = rec[(2 + pl - 1) * DG_POINTS + src];
    const int kills = rec[(4 + pl - 1) * DG_POINTS + src], age = rec[6 * DG_POINTS + src], ladder = rec[(7 + pl - 1) * DG_POINTS + src];
    unsigned long long m = 0;
    m |= (unsigned long long)(stone == 0) << 0;
    m |= (unsigned long long)(stone == pl) << 1;
    m |= (unsigned long long)(stone == 3 - pl) << 2;
    if (lib >= 1 && lib <= 3) m |= 1ull << (3 + lib - 1);
    m |= (unsigned long long)(lib >= 4) << 6;
    m |= (unsigned long long)(stone == 0 && la == 0) << 7;
    if (la >= 1 && la <= 5) m |= 1ull << (7 + la);
    m |= (unsigned long long)(la >= 6) << 13;
    if (kills >= 1 && kills <= 6) m |= 1ull << (14 + kills - 1);
    m |= (unsigned long long)(kills >= 7) << 20;
    if (age >= 1 && age <= 5) m |= 1ull << (21 + age - 1);
    m |= (unsigned long long)(ladder >= 1) << 26;
    m |= 1ull << (27 + rk);
    return m;
}
// 8 planes of that mask as 8 bf16 (1.0 = 0x3F80)
__device__ __forceinline__ uint4 dg_plane_chunk(unsigned long long m, int c) {
    const uint32_t bits = (uint32_t)(m >> (c * 8)) & 0xFFu;
    uint4 o;
    o.x = ((bits & 1) ? 0x3F80u : 0u) | ((bits & 2) ? 0x3F800000u : 0u);
    o.y = ((bits & 4) ? 0x3F80u : 0u) | ((bits & 8) ? 0x3F800000u : 0u);
    o.z = ((bits & 16) ? 0x3F80u : 0u) | ((bits & 32) ? 0x3F800000u : 0u);
    o.w = ((bits & 64) ? 0x3F80u : 0u) | ((bits & 128) ? 0x3F800000u : 0u);
    return o;
}

#define DBG_STAMP(slot) do { if (p.dbg) p.dbg[blockIdx.x * 8 + (slot)] = clock64(); } while (0)
__device__ __forceinline__ long long dg_globaltimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define CHAIN_STAMP(l, k) do { if (p.dbg && blockIdx.x == 0) p.dbg[148 * 10 + (l) * 4 + (k)] = clock64(); } while (0)
#define CONV_BIAS_RING 6
#define CONV_MAX_STAGES 8
#define CONV_MAX_KPANELS 4

// PAIR = true: two CTAs of a cluster (two SMs of a TPC) work as one cta_group::2 unit.  Each CTA holds ITS position's image
// (its 128-row halves of the M = 256 MMA) and HALF of every weight stage (64 of the 128 output channels); the leader CTA
// issues the MMAs for both, the tensor cores share the weight halves.  Per SM that halves the weight bytes written into and
// read out of shared memory -- the bandwidth that bounds the single-CTA kernel (87 instead of 65 cycles per MMA).
#define CONV_MMA_WARPS 3
template <bool PAIR, int MODE>      // MODE 0: forward (bias + ReLU), 1: input gradient (ReLU mask of the layer's stored input)
__global__ void __launch_bounds__(384, 1) conv_umma_kernel(const __grid_constant__ ConvParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t w_full[CONV_MAX_STAGES], w_empty[CONV_MAX_STAGES], w_peer[CONV_MAX_STAGES];
    __shared__ uint64_t img_full[CONV_MAX_KPANELS], img_peer, img_empty, acc_full, acc_ready, acc_ready01;
    __shared__ uint32_t tmem_base_s;
    __shared__ uint32_t row_s[4][256], mask_s[MODE == 1 ? 8 : 1][256];   // per-thread tables of the epilogue warps (see there)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const int chunks_in = p.cin >> 3;
    const uint32_t img_bytes = (uint32_t)chunks_in * p.R_in * 16;
    const int nb_loc = PAIR ? p.nb >> 1 : p.nb;                 // weight rows (output channels) this CTA keeps per stage
    const uint32_t stage_bytes = (uint32_t)p.kp * nb_loc * 2;
    uint8_t *img = smem;
    uint8_t *ring = smem + ((img_bytes + 1023) & ~1023u);
    const int taps = p.ksize * p.ksize;
    const int nkp = p.cin / p.kp;
    const int nnb = p.cout / p.nb;
    const int chunks_kp = p.kp >> 3;
    // positions: one per CTA; a pair walks two neighbouring positions together (the second may not exist when n is odd)
    // Chains with three accumulator tiles rotate them through the four 128-column TMEM blocks: tile t of the a-th layer lives in block
    // (t + 3a) mod 4, so the NEXT layer's tile 0 owns a block that is free while this layer's tiles are still being drained.  Its MMAs
    // need the output rows of tiles 0 and 1 only (21-row halo), so the warp issuing tile 0 starts as soon as those two are drained
    // and runs up to the depth of the weight ring ahead -- under the drain of tile 2.
    const bool rotate = !PAIR && p.ntiles == 3 && p.nchain > 1 && p.nb * 4 <= 512 && p.cout == p.nb;
    const int pos_first = PAIR ? 2 * (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int pos_step = PAIR ? (int)(gridDim.x & ~1u) : (int)gridDim.x;

    dg_pdl_launch_dependents();
    DG_TL_IN(p.tl, p.tl_id);
    if (threadIdx.x == 0) {
        DBG_STAMP(0);
        for (int s = 0; s < p.nstages; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], CONV_MMA_WARPS); mbar_init(&w_peer[s], 1); }
        for (int k = 0; k < nkp; ++k) mbar_init(&img_full[k], 1);
        mbar_init(&img_peer, 1);
        mbar_init(&img_empty, CONV_MMA_WARPS);    // every MMA-issuing warp commits its own MMAs
        mbar_init(&acc_full, CONV_MMA_WARPS);
        mbar_init(&acc_ready, PAIR ? 16 : 8);
        mbar_init(&acc_ready01, 8);
        fence_barrier_init();
    }
    if (warp == 1) {
        if (PAIR) { tmem_alloc2(&tmem_base_s, 512); tmem_relinquish2(); }
        else { tmem_alloc(&tmem_base_s, 512); tmem_relinquish(); }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    dg_pdl_wait();                 // everything above overlapped the previous kernel's tail; its output is visible from here on
    DG_TL_GO(p.tl, p.tl_id);
    if (threadIdx.x == 0) { DBG_STAMP(1); if (p.dbg) p.dbg[148 * 8 + blockIdx.x * 2] = dg_globaltimer(); }

    if (warp == 0) {
        // ===== producer warp (converged; one elected lane issues the bulk copies) =====
        // Stage / phase bookkeeping uses counters only: this warp and the MMA warp are single-issue scalar code,
        // an integer division per stage costs more than the stage's MMAs.
        uint32_t s = 0, ph = 0, j = 0;
        const uint32_t panel_bytes = (uint32_t)chunks_kp * p.R_in * 16;
        const size_t tap_stride = (size_t)chunks_in * p.cout * 16, kp_stride = (size_t)chunks_kp * p.cout * 16;
        for (int pb = pos_first; pb < p.n; pb += pos_step, ++j) {
            const int pos = pb + (int)rank;
            mbar_wait(&img_empty, (j & 1) ^ 1);
            if (pos < p.n && !p.enc_raw && elect_one()) {
                const uint8_t *src = (const uint8_t *)p.in + (size_t)pos * img_bytes;
                for (int k = 0; k < nkp; ++k) {
                    mbar_arrive_expect_tx(&img_full[k], panel_bytes);
                    for (int c = 0; c < chunks_kp; ++c) {
                        const size_t off = (size_t)(k * chunks_kp + c) * p.R_in * 16;
                        bulk_g2s(img + off, src + off, (uint32_t)p.R_in * 16, &img_full[k]);
                    }
                }
            }
            __syncwarp();
            for (int l = 0; l < p.nchain; ++l)
                for (int nbk = 0; nbk < nnb; ++nbk) {
                    const uint8_t *w_k = (const uint8_t *)p.w[l] + ((size_t)nbk * p.nb + (size_t)rank * nb_loc) * 16;
                    for (int k = 0; k < nkp; ++k, w_k += kp_stride) {
                        const uint8_t *wsrc = w_k;
                        for (int t = 0; t < taps; ++t, wsrc += tap_stride) {
                            mbar_wait(&w_empty[s], ph ^ 1);
                            if (elect_one()) {
                                mbar_arrive_expect_tx(&w_full[s], stage_bytes);
                                uint8_t *dst = ring + (size_t)s * stage_bytes;
                                // W[t][k*chunks_kp + c][nbk*nb (+ rank*nb/2) .. ][8]
                                if (nnb == 1 && !PAIR) {
                                    bulk_g2s(dst, wsrc, stage_bytes, &w_full[s]);
                                } else {
                                    for (int c = 0; c < chunks_kp; ++c)
                                        bulk_g2s(dst + (size_t)c * nb_loc * 16, wsrc + (size_t)c * p.cout * 16, (uint32_t)nb_loc * 16, &w_full[s]);
                                }
                            }
                            __syncwarp();
                            if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
                        }
                    }
                }
        }
    } else if (warp == 1 && PAIR && rank == 1) {
        // ===== relay warp of the second CTA: tells the leader when this CTA's image and weight halves have landed =====
        const uint32_t r_img = mapa_u32(smem_u32(&img_peer), 0);
        uint32_t s = 0, ph = 0, j = 0;
        for (int pb = pos_first; pb < p.n; pb += pos_step, ++j) {
            const bool has_pos = pb + 1 < p.n;
            for (int l = 0; l < p.nchain; ++l) {
                if (l == 0) {
                    if (has_pos) for (int k = 0; k < nkp; ++k) mbar_wait(&img_full[k], j & 1);
                    if (elect_one()) mbar_arrive_remote(r_img);
                    __syncwarp();
                }
                for (int t = 0; t < taps * nkp * nnb; ++t) {
                    mbar_wait(&w_full[s], ph);
                    if (elect_one()) mbar_arrive_remote(mapa_u32(smem_u32(&w_peer[s]), 0));
                    __syncwarp();
                    if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (PAIR && rank == 1 && warp >= 10) {
        // second CTA of a pair: the leader issues all MMAs
    } else if (warp == 1 || warp >= 10) {
        // ===== three MMA warps (each converged, so descriptors stay in uniform registers; one elected lane issues) =====
        // The tensor pipe's queue is shallow (measured, tools/umma_rate.cu: a gap in the issue stream is ~90% exposed), so
        // whatever an issuing thread does between MMAs -- barrier waits, commits, counters -- idles the pipe.  Each of the
        // three warps owns one accumulator tile (tile % 3) and walks every weight stage for it: while one warp does its
        // per-stage bookkeeping the other two keep the pipe fed, and every accumulator still sees its MMAs from a single
        // thread in program order (bit-reproducible, no cross-thread hazard on a TMEM tile).
        const int mma_id = warp == 1 ? 0 : warp - 9;
        const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, p.nb, 0, 0);
        const uint32_t img_a = smem_u32(img), ring_a = smem_u32(ring);
        const uint32_t hi = desc_hi(128);
        const uint32_t a_lo_first = desc_lo(img_a, (uint32_t)p.R_in * 16) + (uint32_t)(p.first - p.pad_in * p.pitch_in - p.pad_in);
        const uint32_t b_lo_ring = desc_lo(ring_a, (uint32_t)nb_loc * 16);
        const uint32_t a_kstep = 2u * p.R_in, b_kstep = 2u * nb_loc, stage_units = stage_bytes >> 4;
        const uint32_t a_panel = (uint32_t)chunks_kp * p.R_in;
        const int ksteps = p.kp >> 4;
        uint32_t s = 0, ph = 0, j = 0, a = 0;
        uint32_t b_lo_s = b_lo_ring;
        long long wait_cycles = 0;
        auto wait_stage = [&](uint32_t st, uint32_t phase) {
            mbar_wait(&w_full[st], phase);
            if (PAIR) mbar_wait_cluster(&w_peer[st], phase);
            tc_fence_after();
        };
        for (int pb = pos_first; pb < p.n; pb += pos_step, ++j) {
            if (!PAIR && p.enc_raw) {
                // on-device encode fused into the first layer: these 96 threads turn the raw record of the position into the blocked bf16
                // image (dataloader.lua:50-92 preprocess + the symmetry of transform()), in shared memory and in HBM
                const int tix = mma_id * 32 + lane;
                mbar_wait(&img_empty, (j & 1) ^ 1);                          // the previous position's MMAs no longer read the image
                const uint8_t *rec = p.enc_raw + (size_t)pb * DG_RAW_PLANES * DG_POINTS;
                const int pl = p.enc_player[pb], rk = p.enc_rank[pb], sy = p.enc_sym ? p.enc_sym[pb] : 0;
                unsigned long long pm[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {                                // all loads first: 4 points x 6 bytes in flight
                    const int pt = tix + 96 * k;
                    pm[k] = pt < DG_POINTS ? dg_plane_bits(rec, dg_src_point_d(pt, sy), pl, rk) : 0ull;
                }
                const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
                for (uint32_t i = tix; i < img_bytes / 16; i += 96) *reinterpret_cast<uint4 *>(img + (size_t)i * 16) = z4;   // pads stay zero
                asm volatile("bar.sync 1, 96;" ::: "memory");
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int pt = tix + 96 * k;
                    if (pt < DG_POINTS) {
                        const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
                        const int rho = (x + p.pad_in) * p.pitch_in + y + p.pad_in;
#pragma unroll
                        for (int c = 0; c < 6; ++c) {
                            const uint4 o = dg_plane_chunk(pm[k], c);
                            *reinterpret_cast<uint4 *>(img + ((size_t)c * p.R_in + rho) * 16) = o;
                            *reinterpret_cast<uint4 *>(p.enc_out + (((size_t)pb * 6 + c) * p.R_in + rho) * 8) = o;
                        }
                    }
                }
                fence_proxy_async();                                         // generic-proxy st.shared -> visible to the tensor core's async proxy
                asm volatile("bar.sync 1, 96;" ::: "memory");
            }
            for (int l = 0; l < p.nchain; ++l)
                for (int nbk = 0; nbk < nnb; ++nbk, ++a) {
                    // the epilogue warps have drained the previous accumulators and written the previous layer's output rows into
                    // the shared-memory image (chains); the first MMA of every tile overwrites its accumulator
                    if (PAIR) mbar_wait_cluster(&acc_ready, a & 1);
                    else if (rotate && mma_id == 0) mbar_wait(&acc_ready01, a & 1);      // tile 0 may start before tile 2 of the previous layer is drained
                    else mbar_wait(&acc_ready, a & 1);
                    tc_fence_after();
                    if (lane == 0 && j == 0 && mma_id == 0) CHAIN_STAMP(l, 0);
                    uint32_t a_lo_k = a_lo_first;
                    for (int k = 0; k < nkp; ++k, a_lo_k += a_panel) {
                        if (l == 0 && nbk == 0 && !(!PAIR && p.enc_raw)) {
                            mbar_wait(&img_full[k], j & 1);
                            if (PAIR && k == 0) mbar_wait_cluster(&img_peer, j & 1);
                            if (j == 0 && lane == 0 && mma_id == 0) DBG_STAMP(2 + (k > 0));
                        }
                        uint32_t a_lo_ky = a_lo_k;                    // + ky*pitch + kx rows, in 16-byte units
                        for (int ky = 0; ky < p.ksize; ++ky, a_lo_ky += p.pitch_in) {
                            for (int kx = 0; kx < p.ksize; ++kx) {
                                const long long tw0 = p.dbg ? clock64() : 0;
                                const bool first_stage = (k | ky | kx) == 0;      // the first MMA of a tile overwrites its accumulator
                                wait_stage(s, ph);
                                if (p.dbg) wait_cycles += clock64() - tw0;
                                if (elect_one()) {
                                    for (int tile = mma_id; tile < p.ntiles; tile += CONV_MMA_WARPS) {
                                        uint32_t a_lo = a_lo_ky + kx + (uint32_t)tile * 128, b_lo = b_lo_s;
                                        const uint32_t d_tmem = tmem_base + (rotate ? (uint32_t)((tile + 3 * a) & 3) : (uint32_t)tile) * p.nb;
#pragma unroll 4
                                        for (int ks = 0; ks < ksteps; ++ks, a_lo += a_kstep, b_lo += b_kstep) {
                                            const uint32_t accumulate = first_stage ? (uint32_t)(ks != 0) : 1u;
                                            if (PAIR) mma_bf16_pair(d_tmem, desc_join(a_lo, hi), desc_join(b_lo, hi), idesc, accumulate);
                                            else mma_bf16(d_tmem, desc_join(a_lo, hi), desc_join(b_lo, hi), idesc, accumulate);
                                        }
                                    }
                                    if (PAIR) mma_commit_pair(&w_empty[s], 3); else mma_commit(&w_empty[s]);
                                }
                                __syncwarp();
                                b_lo_s += stage_units;
                                if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; b_lo_s = b_lo_ring; }
                            }
                        }
                    }
                    if (elect_one()) { if (PAIR) mma_commit_pair(&acc_full, 3); else mma_commit(&acc_full); }
                    __syncwarp();
                    if (lane == 0 && j == 0 && mma_id == 0) CHAIN_STAMP(l, 1);
                    if (a == 0 && lane == 0 && mma_id == 0) { DBG_STAMP(4); if (p.dbg) p.dbg[blockIdx.x * 8 + 7] = wait_cycles; }
                }
            if (elect_one()) { if (PAIR) mma_commit_pair(&img_empty, 3); else mma_commit(&img_empty); }
            __syncwarp();
        }
    } else {
        // ===== epilogue warps 2..9: TMEM -> registers -> ReLU or mask -> bf16 -> HBM (+ shared-memory image in a chain) =====
        // warp w may only touch TMEM lanes 32*(w%4)..+31; two warps share a lane quarter and split the 32-column blocks.
        // An "item" is one 32-column block of one accumulator tile.  The item loops are ROLLED on purpose: fully unrolled, this
        // branch was ~100 KB of SASS and every layer streamed it through the instruction caches (the drain ran at a quarter
        // of the speed it has now).  What varies per item and cannot be recomputed cheaply lives in two small per-thread
        // tables in shared memory.
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int et = (int)threadIdx.x - 64;                      // 0..255: this thread's column in the per-thread tables
        const int chunks_out = p.cout >> 3;
        const int nblk32 = p.nb >> 5;
        const int my_blocks = (nblk32 - half + 1) >> 1;           // blocks half, half+2, ...
        const int items = p.ntiles * my_blocks;                    // <= 8 (ntiles * nb <= 512 TMEM columns)
        const int rho0 = p.first + q * 32 + lane;                  // this thread's image row in tile 0 (+128 per tile)
        const uint32_t tm_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        // rows are the same for every layer and position: bit 31 = a board point, bits 12..23 = row in the geometry of the
        // launch's last output, bits 0..11 = row in the mask geometry
        for (int tile = 0; tile < p.ntiles; ++tile) {
            const int rho = rho0 + tile * 128;
            const int xp = rho / p.pitch_in, yp = rho - xp * p.pitch_in;
            const int x = xp - p.pad_in, y = yp - p.pad_in;
            const bool on_board = x >= 0 && x < DG_BOARD && y >= 0 && y < DG_BOARD;
            const uint32_t r_last = (uint32_t)((x + p.pad_out) * p.pitch_out + y + p.pad_out);
            const uint32_t r_mask = (uint32_t)((x + p.pad_mask) * p.pitch_mask + y + p.pad_mask);
            row_s[tile][et] = on_board ? (0x80000000u | (r_last << 12) | r_mask) : 0u;
        }
        bool has_pos_flag = true;                                  // false: this CTA of a pair has no position (n odd); nothing is loaded or stored
        auto advance = [&](int &tile, int &blk) { if (++blk == my_blocks) { blk = 0; ++tile; } };
        auto row_of = [&](int tile, int blk, int l, int &rho, int &rho_out, int &rho_m, bool &valid, int &c0) {
            const uint32_t ri = row_s[tile][et];
            c0 = (blk * 2 + half) * 32;
            rho = rho0 + tile * 128;
            valid = has_pos_flag && (ri >> 31);
            // intermediate layers of a chain keep the input geometry; only the last layer's HBM output may differ
            rho_out = l == p.nchain - 1 ? (int)((ri >> 12) & 0xFFFu) : rho;
            rho_m = (int)(ri & 0xFFFu);
        };
        // one item's bias (32 channels of one row) of layer l, N block nbk, into registers
        auto load_bias_item = [&](int tile, int blk, int l, int nbk, uint4 (&bq)[4]) {
            int rho, rho_out, rho_m, c0; bool valid;
            row_of(tile, blk, l, rho, rho_out, rho_m, valid, c0);
            const int R_b = l == p.nchain - 1 ? p.R_out : p.R_in;
            const uint4 *src = reinterpret_cast<const uint4 *>(p.bias[l]) + (size_t)((nbk * p.nb + c0) >> 3) * R_b + rho_out;
#pragma unroll
            for (int g = 0; g < 4; ++g, src += R_b) {
                bq[g] = make_uint4(0u, 0u, 0u, 0u);
                if (valid) bq[g] = __ldg(src);
            }
        };
        // hand-over to the MMA warp: in a pair the barrier lives in the leader CTA (the second CTA arrives remotely)
        const uint32_t r_acc_ready = PAIR ? mapa_u32(smem_u32(&acc_ready), 0) : 0u;
        auto arrive_ready = [&]() {
            if (lane == 0) { if (PAIR && rank == 1) mbar_arrive_remote(r_acc_ready); else mbar_arrive(&acc_ready); }
        };
        const uint32_t img_s = smem_u32(img);
        const uint32_t one2 = 0x3F803F80u;                         // bf16x2 {1, 1}
        uint32_t a = 0;
        for (int pb = pos_first; pb < p.n; pb += pos_step) {
            const int pos = pb + (int)rank;
            const bool has_pos = pos < p.n;                        // the second CTA of the last pair may have no position
            has_pos_flag = has_pos;
            for (int l = 0; l < p.nchain; ++l)
                for (int nbk = 0; nbk < nnb; ++nbk, ++a) {
                    const bool chain_more = l + 1 < p.nchain;      // this layer's output is the next layer's shared-memory image
                    if (l == 0 || nnb > 1) {
                        // first layer of a position (or a multi-N-block layer): the accumulators are free (the first MMA of a tile
                        // overwrites it); inside a chain the hand-over happens at the end of the previous layer's drain
                        tc_fence_before();
                        __syncwarp();
                        arrive_ready();
                        if (rotate && lane == 0) mbar_arrive(&acc_ready01);
                    }
                    // backward: while the MMAs run, the ReLU mask of each item (stored input of the layer > 0) is compressed to one
                    // 32-bit word, laid out so that the drain expands it with byte permutes: element (chunk g, pair j, half h)
                    // sits at the top of byte 2*(j&1)+h, shifted down by k = 2*g + (j>>1)
                    if (MODE == 1) {
                        int tile = 0, blk = 0;
                        for (int item = 0; item < items; ++item) {
                            int rho, rho_out, rho_m, c0; bool valid;
                            row_of(tile, blk, l, rho, rho_out, rho_m, valid, c0);
                            uint32_t mbits = 0;
                            if (valid) {
                                const uint4 *src = reinterpret_cast<const uint4 *>(p.mask[l]) + ((size_t)pos * chunks_out + ((nbk * p.nb + c0) >> 3)) * p.R_mask + rho_m;
                                uint4 mk[4];
#pragma unroll
                                for (int g = 0; g < 4; ++g) mk[g] = __ldg(src + (size_t)g * p.R_mask);
#pragma unroll
                                for (int g = 0; g < 4; ++g) {
                                    const uint32_t mw[4] = {mk[g].x, mk[g].y, mk[g].z, mk[g].w};
#pragma unroll
                                    for (int j = 0; j < 4; ++j) {
                                        const int k = 2 * g + (j >> 1);
                                        // bf16 > 0  <=>  sign bit clear and not zero
                                        const uint32_t lo = mw[j] & 0xFFFFu, hi = mw[j] >> 16;
                                        if (lo != 0 && (lo & 0x8000u) == 0) mbits |= 1u << (8 * (2 * (j & 1)) + 7 - k);
                                        if (hi != 0 && (hi & 0x8000u) == 0) mbits |= 1u << (8 * (2 * (j & 1) + 1) + 7 - k);
                                    }
                                }
                            }
                            mask_s[item][et] = mbits;
                            advance(tile, blk);
                        }
                    }
                    // forward: the layer's bias (bf16, b[c] + a[c][x][y]) of the first CONV_BIAS_RING items is fetched while the MMAs
                    // run; a slot is refilled (items beyond the ring) right after the drain has used it.  Loads inside the drain
                    // queue behind its stores, so the ring is as deep as the registers allow.
                    uint4 bq[MODE == 0 ? CONV_BIAS_RING : 1][4];
                    int ftile = 0, fblk = 0;                       // fetch cursor
                    if (MODE == 0) {
#pragma unroll
                        for (int u = 0; u < CONV_BIAS_RING; ++u)
                            if (u < items) { load_bias_item(ftile, fblk, l, nbk, bq[u]); advance(ftile, fblk); }
                    }
                    mbar_wait(&acc_full, a & 1);
                    tc_fence_after();
                    if (a == 0 && threadIdx.x == 64) DBG_STAMP(5);
                    if (threadIdx.x == 64 && pb == pos_first) CHAIN_STAMP(l, 2);
                    const int R_o = chain_more ? p.R_in : p.R_out;
                    uint4 *out_pos = reinterpret_cast<uint4 *>(p.out[l]) + (size_t)pos * chunks_out * R_o;
                    int tile = 0, blk = 0;                         // drain cursor
                    for (int item0 = 0; item0 < items; item0 += CONV_BIAS_RING) {
#pragma unroll
                        for (int u = 0; u < CONV_BIAS_RING; ++u) {   // unrolled by the ring depth: static register indices
                            const int item = item0 + u;
                            if (item < items) {
                                int rho, rho_out, rho_m, c0; bool valid;
                                row_of(tile, blk, l, rho, rho_out, rho_m, valid, c0);
                                const int chunk0 = (nbk * p.nb + c0) >> 3;
                                uint32_t r[32];
                                tmem_ld32(tm_lane + (rotate ? (uint32_t)((tile + 3 * a) & 3) : (uint32_t)tile) * p.nb + c0, r);
                                const uint32_t mbits = MODE == 1 ? mask_s[item][et] : 0u;
                                tmem_ld_wait();
                                if (valid) {
                                    uint4 *og = out_pos + (size_t)chunk0 * R_o + rho_out;
                                    uint32_t sg = img_s + (uint32_t)(chunk0 * p.R_in + rho) * 16u;
#pragma unroll
                                    for (int g = 0; g < 4; ++g, og += R_o, sg += (uint32_t)p.R_in * 16u) {
                                        uint32_t w[4];
#pragma unroll
                                        for (int j = 0; j < 4; ++j) {
                                            const __nv_bfloat162 t = __floats2bfloat162_rn(__uint_as_float(r[g * 8 + 2 * j]), __uint_as_float(r[g * 8 + 2 * j + 1]));
                                            w[j] = *reinterpret_cast<const uint32_t *>(&t);
                                        }
                                        if (MODE == 0) {
                                            // relu(acc + bias) on bf16 pairs: one packed fma per pair
                                            const uint4 bb = bq[MODE == 0 ? u : 0][g];
                                            const uint32_t bw[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
                                            for (int j = 0; j < 4; ++j) asm("fma.rn.relu.bf16x2 %0, %1, %2, %3;" : "=r"(w[j]) : "r"(w[j]), "r"(one2), "r"(bw[j]));
                                        } else {
                                            // each byte permute replicates two mask bits (byte tops) over a bf16 pair
#pragma unroll
                                            for (int hg = 0; hg < 2; ++hg) {
                                                const uint32_t x = mbits << (2 * g + hg);
                                                uint32_t m0, m1;
                                                asm("prmt.b32 %0, %1, %1, 0x9988;" : "=r"(m0) : "r"(x));
                                                asm("prmt.b32 %0, %1, %1, 0xBBAA;" : "=r"(m1) : "r"(x));
                                                w[2 * hg] &= m0;
                                                w[2 * hg + 1] &= m1;
                                            }
                                        }
                                        *og = make_uint4(w[0], w[1], w[2], w[3]);
                                        if (chain_more) asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sg), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
                                    }
                                }
                                if (MODE == 0 && item + CONV_BIAS_RING < items) { load_bias_item(ftile, fblk, l, nbk, bq[MODE == 0 ? u : 0]); advance(ftile, fblk); }
                                advance(tile, blk);
                                if (rotate && chain_more && tile == 2 && blk == 0) {
                                    // tiles 0 and 1 are drained: the next layer's tile 0 has its input rows and a free accumulator block
                                    fence_proxy_async();
                                    tc_fence_before();
                                    __syncwarp();
                                    if (lane == 0) mbar_arrive(&acc_ready01);
                                }
                            }
                        }
                    }
                    if (chain_more) {
                        // next layer: the accumulators are free and its input image is complete once every epilogue warp gets here
                        fence_proxy_async();                       // generic-proxy st.shared -> visible to the tensor core's async proxy
                        tc_fence_before();
                        __syncwarp();
                        arrive_ready();
                    } else {
                        tc_fence_before();
                        __syncwarp();
                    }
                    if (a == 0 && threadIdx.x == 64) DBG_STAMP(6);
                    if (threadIdx.x == 64 && pb == pos_first) CHAIN_STAMP(l, 3);
                }
        }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();          // a pair: neither CTA may leave while the other still uses its memory
    if (warp == 1) { if (PAIR) tmem_dealloc2(tmem_base, 512); else tmem_dealloc(tmem_base, 512); }
    DG_TL_OUT(p.tl, p.tl_id);
    if (threadIdx.x == 0 && p.dbg) p.dbg[148 * 8 + blockIdx.x * 2 + 1] = dg_globaltimer();
}

// ---------------------------------------------------------------------------------------------
// weight gradient: D[co][ci] per tap, split-K over (position, 64-row panel) units, partial tiles to
// scratch, then a deterministic reduction into the Torch7-order gradient vector.
// ---------------------------------------------------------------------------------------------
#define WG_MAX_LAYERS 16
struct WgradParams {
    // A launch covers `nl` layers of identical shape: every CTA keeps its (kernel row ky, split) role and walks the layers,
    // so launch, prologue and the tail of the L2 reductions are paid once and the panel ring never drains between layers.
    const bf16 *dz[WG_MAX_LAYERS];   // [n][cout/8][R][8]   (A, MN-major: M = cout channel)
    const bf16 *x[WG_MAX_LAYERS];    // [n][cin/8][R][8]    (B, MN-major: N = cin channel)
    float *gacc[WG_MAX_LAYERS];      // default path: fp32 accumulation buffer of the layer [tap][cin/4][cout][4], updated with 16-byte reductions
    float *gab[WG_MAX_LAYERS];       // default path: bias-gradient accumulation buffer [(row - first)/4][cout][4] = sum over positions of dZ (or NULL)
    float *scratch;       // deterministic path (nl == 1, gacc[0] == NULL): partial tiles [combo][split][kx][nb/4][128][4]
    int nl;
    int n, cin, cout;     // cin padded (multiple of 16)
    int R, pitch, pad, ksize, first, ntiles;
    int nb;               // N block (cin % nb == 0), k*nb <= 512
    int n_mblk, n_nblk;   // cout / 128 (rounded up), cin / nb
    int splits;           // split-K CTAs per combo
    int nstages;
    unsigned long long *tl; int tl_id;
    int ngroups;          // CTAs are dealt to `ngroups` layer groups: group g walks layers g, g + ngroups, .. with `splits` CTAs per kernel row
    long long *dbg;       // optional timeline (DG_TIMELINE=1): per CTA {start, mma-loop start, mma-loop end, total wait, end}
};

#define WG_RP_MAX 128     // K rows per stage: a template parameter (96 for the 3x3 geometry: four 50 KB stages instead of three 65 KB ones)

template <int WG_RP>
__global__ void __launch_bounds__(384, 1) wgrad_umma_kernel(const __grid_constant__ WgradParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // accumulator hand-over is per MMA warp (= per group of kernel columns): the drain of warp w's columns overlaps the other
    // warps' MMAs of the next layer
    __shared__ uint64_t s_full[CONV_MAX_STAGES], s_empty[CONV_MAX_STAGES], acc_full[CONV_MMA_WARPS], acc_empty[CONV_MMA_WARPS];
    __shared__ uint32_t tmem_base_s;
    // Bias gradients ride along: sum over positions of dZ = dZ^T x I, so one extra N = 16 MMA per 16 panel rows against a 16x16
    // identity (MN-major, two 8-column chunks) accumulates this CTA's panel of the bias gradient in the spare TMEM columns.  The
    // CTAs of one split (one per kernel row) share the dZ panel and take every ksize-th 16-row step each.
    __shared__ __align__(128) uint16_t ident_s[2][16][8];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Fewer, longer split-K runs per layer: with G layer groups a CTA accumulates G times more units in TMEM before it drains, so
    // the accumulator hand-over (and the L2 reduction traffic) is paid 1/G as often.
    const int grp = blockIdx.x % p.ngroups, bidx = blockIdx.x / p.ngroups;
    const int combo = bidx / p.splits, split = bidx - combo * p.splits;
    // combo -> (ky, mblk, nblk)
    const int nblk = combo % p.n_nblk, mblk = (combo / p.n_nblk) % p.n_mblk, ky = combo / (p.n_nblk * p.n_mblk);
    const int rb = WG_RP + 2 * p.pad;                         // B panel rows incl. the kx halo
    const int rb_pad = (rb + 7) & ~7;
    const uint32_t a_bytes = 16u * WG_RP * 16;                // 16 channel chunks (M = 128) x WG_RP rows
    const uint32_t b_bytes = (uint32_t)(p.nb >> 3) * rb_pad * 16;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    const int panels = p.ntiles * 128 / WG_RP;
    const int units = p.n * panels;
    const int chunks_dz = p.cout >> 3, chunks_x = p.cin >> 3;
    // channel chunks of dZ this M block really has (the rest of the 128 lanes accumulate garbage that is never stored)
    const int a_chunks = min(16, chunks_dz - mblk * 16);
    const bool fused_bias = p.gab[0] != nullptr && nblk == 0;   // splits % panels == 0: every unit of this CTA is the same panel
    const uint32_t bias_col = (uint32_t)p.ksize * p.nb;          // first spare TMEM column (ksize * nb + WG_RP <= 512)

    dg_pdl_launch_dependents();
    DG_TL_IN(p.tl, p.tl_id);
    if (threadIdx.x == 0 && p.dbg) p.dbg[blockIdx.x * 8] = clock64();
    if (threadIdx.x < 256) {
        const int ch = threadIdx.x >> 7, k = (threadIdx.x >> 3) & 15, e = threadIdx.x & 7;
        ident_s[ch][k][e] = k == ch * 8 + e ? 0x3F80u : 0u;      // bf16 1.0
    }
    fence_proxy_async();
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], CONV_MMA_WARPS); }
        for (int w = 0; w < CONV_MMA_WARPS; ++w) { mbar_init(&acc_full[w], 1); mbar_init(&acc_empty[w], 8); }
        fence_barrier_init();
    }
    if (warp == 1) { tmem_alloc(&tmem_base_s, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    dg_pdl_wait();
    DG_TL_GO(p.tl, p.tl_id);

    if (warp == 0) {
        uint32_t s = 0, ph = 0;
        const int pos0 = split / panels, panel0 = split - pos0 * panels;   // one division up front, counters afterwards
        const int dpos = p.splits / panels, dpanel = p.splits - dpos * panels;
        const int brow_off = (ky - p.pad) * p.pitch - p.pad;
        const uint32_t tx = (uint32_t)a_chunks * WG_RP * 16 + (uint32_t)(p.nb >> 3) * rb * 16;
        long long pw = 0, pi = 0;                                  // timeline: cycles waiting for a free stage / issuing copies
        for (int li = grp; li < p.nl; li += p.ngroups) {
            int pos = pos0, panel = panel0;
            for (int u = split; u < units; u += p.splits) {
                const long long t0 = p.dbg ? clock64() : 0;
                mbar_wait(&s_empty[s], ph ^ 1);
                const long long t1 = p.dbg ? clock64() : 0;
                if (elect_one()) {
                    mbar_arrive_expect_tx(&s_full[s], tx);
                    uint8_t *sa = smem + (size_t)s * stage_bytes, *sb = sa + a_bytes;
                    const int row0 = p.first + panel * WG_RP;
                    const uint8_t *asrc = (const uint8_t *)p.dz[li] + (((size_t)pos * chunks_dz + mblk * 16) * p.R + row0) * 16;
                    // (unrolled: each copy needs its operands moved to uniform registers, a dependent ~40-cycle chain per copy otherwise)
                    const uint32_t sa_u = smem_u32(sa), sb_u = smem_u32(sb), bar_u = smem_u32(&s_full[s]);
                    const size_t src_stride = (size_t)p.R * 16;
#pragma unroll 8
                    for (int c = 0; c < a_chunks; ++c)
                        bulk_g2s_u(sa_u + (uint32_t)c * WG_RP * 16, asrc + (size_t)c * src_stride, WG_RP * 16, bar_u);
                    const uint8_t *bsrc = (const uint8_t *)p.x[li] + (((size_t)pos * chunks_x + nblk * (p.nb >> 3)) * p.R + row0 + brow_off) * 16;
#pragma unroll 8
                    for (int c = 0; c < (p.nb >> 3); ++c)
                        bulk_g2s_u(sb_u + (uint32_t)c * rb_pad * 16, bsrc + (size_t)c * src_stride, (uint32_t)rb * 16, bar_u);
                }
                __syncwarp();
                if (p.dbg) { const long long t2 = clock64(); pw += t1 - t0; pi += t2 - t1; }
                if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
                pos += dpos; panel += dpanel;
                if (panel >= panels) { panel -= panels; ++pos; }
            }
        }
        if (lane == 0 && p.dbg) { p.dbg[blockIdx.x * 8 + 5] = pw; p.dbg[blockIdx.x * 8 + 6] = pi; }
    } else if (warp == 1 || warp >= 10) {
        // three MMA warps: warp i issues the MMAs of kernel columns kx = i, i+3, .. (disjoint accumulators), so the shallow
        // tensor-pipe queue always has work while another warp waits on a panel barrier or commits (see conv_umma_kernel)
        const int mma_id = warp == 1 ? 0 : warp - 9;
        const uint32_t idesc = make_idesc_bf16(128, p.nb, 1, 1);
        const uint32_t hi_a = desc_hi(WG_RP * 16), hi_b = desc_hi((uint32_t)rb_pad * 16);
        const uint32_t a_lo_ring = desc_lo(smem_u32(smem), 128), b_lo_ring = desc_lo(smem_u32(smem) + a_bytes, 128);
        const uint32_t stage_units = stage_bytes >> 4;
        const uint32_t idesc_bias = make_idesc_bf16(128, 16, 1, 1);
        const uint64_t ident_desc = desc_join(desc_lo(smem_u32(ident_s), 128), desc_hi(16 * 16));
        uint32_t s = 0, ph = 0;
        uint32_t a_lo_s = a_lo_ring, b_lo_s = b_lo_ring;
        long long wcy = 0;
        if (lane == 0 && p.dbg && mma_id == 0) p.dbg[blockIdx.x * 8 + 1] = clock64();
        for (int li = grp, lc = 0; li < p.nl; li += p.ngroups, ++lc) {
            mbar_wait(&acc_empty[mma_id], (lc & 1) ^ 1);      // the drain warps have read this warp's accumulators of the previous layer
            tc_fence_after();
            uint32_t acc = 0;
            for (int u = split; u < units; u += p.splits) {
                const long long tw0 = p.dbg ? clock64() : 0;
                mbar_wait(&s_full[s], ph);
                if (p.dbg) wcy += clock64() - tw0;
                tc_fence_after();
                if (elect_one()) {
                    for (int kx = mma_id; kx < p.ksize; kx += CONV_MMA_WARPS) {
                        const uint32_t d_tmem = tmem_base + (uint32_t)kx * p.nb;
                        uint32_t a_lo = a_lo_s, b_lo = b_lo_s + kx;
#pragma unroll
                        for (int ks = 0; ks < WG_RP / 16; ++ks, a_lo += 16, b_lo += 16)
                            mma_bf16(d_tmem, desc_join(a_lo, hi_a), desc_join(b_lo, hi_b), idesc, acc | (uint32_t)ks);
                    }
                    if (fused_bias && mma_id == 0)
                        for (int ks = ky; ks < WG_RP / 16; ks += p.ksize)
                            mma_bf16(tmem_base + bias_col + 16u * ks, desc_join(a_lo_s + 16u * ks, hi_a), ident_desc, idesc_bias, acc);
                    mma_commit(&s_empty[s]);
                }
                __syncwarp();
                acc = 1;
                a_lo_s += stage_units; b_lo_s += stage_units;
                if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; a_lo_s = a_lo_ring; b_lo_s = b_lo_ring; }
            }
            if (elect_one()) mma_commit(&acc_full[mma_id]);
            __syncwarp();
        }
        if (lane == 0 && p.dbg && mma_id == 0) { p.dbg[blockIdx.x * 8 + 2] = clock64(); p.dbg[blockIdx.x * 8 + 3] = wcy; }
    } else {
        // ===== drain warps 2..9: two warps per TMEM lane quarter, alternating 32-column blocks =====
        const int q = warp & 3, half = (warp - 2) >> 2;
        const bool any = split < units;                       // a CTA without units holds an unwritten accumulator
        const int co = mblk * 128 + q * 32 + lane;
        const int cin4 = p.cin >> 2;
        for (int li = grp, lc = 0; li < p.nl; li += p.ngroups, ++lc) {
            float4 *gl = reinterpret_cast<float4 *>(p.gacc[li]);
            // scratch tile layout [kx][ci/4][co][4]: lane = co, so one warp-wide 16-byte access covers 512 contiguous bytes
            float4 *sc = reinterpret_cast<float4 *>(p.scratch) + ((size_t)combo * p.splits + split) * p.ksize * 32 * p.nb + (q * 32 + lane);
            if (gl && p.ksize == 3 && p.nb == 128) {
                // Fast path (3x3 layers): the reductions are bound by L2 atomic throughput (all CTAs drain at once), and an MMA warp can
                // only start the next layer when its accumulator has been READ.  So: read two kernel columns into registers and hand
                // them back at once, then issue the reductions -- warps 0 and 1 restart ~100 cycles after they finish, warp 2 after a
                // third of the drain instead of all of it.
                uint32_t ra[2][32], rb2[2][32];
                const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16);
                auto ld_kx = [&](int kx, uint32_t (&r)[2][32]) {
                    tmem_ld32(tq + kx * 128 + half * 32, r[0]);
                    tmem_ld32(tq + kx * 128 + half * 32 + 64, r[1]);
                    tmem_ld_wait();
                };
                auto red_kx = [&](int kx, const uint32_t (&r)[2][32]) {
                    if (!(any && co < p.cout)) return;
#pragma unroll
                    for (int bk = 0; bk < 2; ++bk) {
                        float4 *o = gl + ((size_t)(ky * 3 + kx) * cin4 + nblk * 32 + ((half * 32 + bk * 64) >> 2)) * p.cout + co;
#pragma unroll
                        for (int g = 0; g < 8; ++g)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + (size_t)g * p.cout), "f"(__uint_as_float(r[bk][g * 4])),
                                         "f"(__uint_as_float(r[bk][g * 4 + 1])), "f"(__uint_as_float(r[bk][g * 4 + 2])), "f"(__uint_as_float(r[bk][g * 4 + 3]))
                                         : "memory");
                    }
                };
                auto release = [&](int w) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(&acc_empty[w]); };
                mbar_wait(&acc_full[0], lc & 1);
                tc_fence_after();
                ld_kx(0, ra);
                if (fused_bias) {
                    const int panel = split % panels;
                    float4 *gb4 = reinterpret_cast<float4 *>(p.gab[li]);
                    for (int ks = ky + half * 3; ks < WG_RP / 16; ks += 6) {
                        uint32_t r[32];
                        tmem_ld16(tq + bias_col + 16u * ks, r);
                        tmem_ld_wait();
                        if (any && co < p.cout) {
                            float4 *o = gb4 + (size_t)((panel * WG_RP + 16 * ks) >> 2) * p.cout + co;
#pragma unroll
                            for (int g = 0; g < 4; ++g)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + (size_t)g * p.cout), "f"(__uint_as_float(r[g * 4])),
                                             "f"(__uint_as_float(r[g * 4 + 1])), "f"(__uint_as_float(r[g * 4 + 2])), "f"(__uint_as_float(r[g * 4 + 3]))
                                             : "memory");
                        }
                    }
                }
                release(0);
                mbar_wait(&acc_full[1], lc & 1);
                tc_fence_after();
                ld_kx(1, rb2);
                release(1);
                red_kx(0, ra);
                mbar_wait(&acc_full[2], lc & 1);
                tc_fence_after();
                ld_kx(2, ra);
                release(2);
                red_kx(1, rb2);
                red_kx(2, ra);
            } else
            for (int w = 0; w < CONV_MMA_WARPS; ++w) {
            mbar_wait(&acc_full[w], lc & 1);
            tc_fence_after();
            for (int kx = w; kx < p.ksize; kx += CONV_MMA_WARPS)
                for (int c0 = half * 32; c0 < p.nb; c0 += 64) {
                    uint32_t r[32];
                    const int ncol = min(32, p.nb - c0);          // nb is a multiple of 16
                    const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + kx * p.nb + c0;
                    if (ncol == 32) tmem_ld32(ta, r); else tmem_ld16(ta, r);
                    tmem_ld_wait();
                    if (gl) {
                        // default: add this CTA's partial tile straight into the layer's accumulation buffer with vector reductions
                        // (red.global.add.v4.f32, resolved in L2): no partial-tile round trip through memory, no reduction kernel
                        if (any && co < p.cout) {
                            float4 *o = gl + ((size_t)(ky * p.ksize + kx) * cin4 + nblk * (p.nb >> 2) + (c0 >> 2)) * p.cout + co;
#pragma unroll
                            for (int g = 0; g < 8; ++g)
                                if (g * 4 < ncol)
                                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + (size_t)g * p.cout), "f"(__uint_as_float(r[g * 4])),
                                                 "f"(__uint_as_float(r[g * 4 + 1])), "f"(__uint_as_float(r[g * 4 + 2])), "f"(__uint_as_float(r[g * 4 + 3]))
                                                 : "memory");
                        }
                    } else {
                        float4 *o = sc + ((size_t)kx * (p.nb >> 2) + (c0 >> 2)) * 128;
#pragma unroll
                        for (int g = 0; g < 8; ++g)
                            if (g * 4 < ncol)
                                o[(size_t)g * 128] = any ? make_float4(__uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]), __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]))
                                                         : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            if (w == 0 && fused_bias) {
                // this CTA's share of the bias-gradient panel: 16-row steps ky, ky + ksize, .., alternating between the two warps of a lane quarter
                const int panel = split % panels;
                float4 *gb4 = reinterpret_cast<float4 *>(p.gab[li]);
                for (int ks = ky + half * p.ksize; ks < WG_RP / 16; ks += 2 * p.ksize) {
                    uint32_t r[32];
                    tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + bias_col + 16u * ks, r);
                    tmem_ld_wait();
                    if (any && co < p.cout) {
                        float4 *o = gb4 + (size_t)((panel * WG_RP + 16 * ks) >> 2) * p.cout + co;
#pragma unroll
                        for (int g = 0; g < 4; ++g)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + (size_t)g * p.cout), "f"(__uint_as_float(r[g * 4])),
                                         "f"(__uint_as_float(r[g * 4 + 1])), "f"(__uint_as_float(r[g * 4 + 2])), "f"(__uint_as_float(r[g * 4 + 3]))
                                         : "memory");
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[w]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
    DG_TL_OUT(p.tl, p.tl_id);
    if (threadIdx.x == 0 && p.dbg) p.dbg[blockIdx.x * 8 + 4] = clock64();
}

// gW[(co*cin_true + ci)*KK + ky*k + kx] += sum_split scratch[combo(ky,mblk,nblk)][split][kx][co%128][ci%nb]
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float *__restrict__ scratch, float *__restrict__ gW, int cin_true, int cin,
                                                           int cout, int ksize, int nb, int n_mblk, int n_nblk, int splits) {
    // scratch tile layout per (combo, split): [kx][ci/4][co%128][4].  A block sums one (tap, 64 co, 16 ci) tile over the
    // splits with co-fastest threads (coalesced 16-byte reads), transposes it in shared memory and adds it to the Torch7-order
    // gradient with ci-fastest threads (neighbouring threads 36 bytes apart instead of a whole weight row).
    __shared__ float tile[64][17];
    const int KK = ksize * ksize;
    const int co_tiles = (cout + 63) >> 6, ci_tiles = cin >> 4;
    const int ntiles = KK * co_tiles * ci_tiles;
    const size_t st4 = (size_t)ksize * 32 * nb;                      // float4s between consecutive splits
    const float4 *sc4 = reinterpret_cast<const float4 *>(scratch);
    for (int tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
        const int cit = tl % ci_tiles, cot = (tl / ci_tiles) % co_tiles, t = tl / (ci_tiles * co_tiles);
        const int ky = t / ksize, kx = t - ky * ksize;
        {
            const int co = cot * 64 + (threadIdx.x & 63), c4l = threadIdx.x >> 6, ci = cit * 16 + c4l * 4;
            float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
            if (co < cout) {
                const int combo = (ky * n_mblk + co / 128) * n_nblk + ci / nb;
                const float4 *src = sc4 + (size_t)combo * splits * st4 + ((size_t)kx * (nb >> 2) + ((ci % nb) >> 2)) * 128 + (co & 127);
                int sp = 0;
#pragma unroll 4
                for (; sp + 2 <= splits; sp += 2) {
                    const float4 a = __ldg(src + (size_t)sp * st4), b = __ldg(src + (size_t)(sp + 1) * st4);
                    s0.x += a.x; s0.y += a.y; s0.z += a.z; s0.w += a.w;
                    s1.x += b.x; s1.y += b.y; s1.z += b.z; s1.w += b.w;
                }
                if (sp < splits) {
                    const float4 a = __ldg(src + (size_t)sp * st4);
                    s0.x += a.x; s0.y += a.y; s0.z += a.z; s0.w += a.w;
                }
            }
            float *tr = &tile[threadIdx.x & 63][c4l * 4];
            tr[0] = s0.x + s1.x; tr[1] = s0.y + s1.y; tr[2] = s0.z + s1.z; tr[3] = s0.w + s1.w;
        }
        __syncthreads();
        {
            const int cil = threadIdx.x & 15, ci = cit * 16 + cil;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int col = (threadIdx.x >> 4) + 16 * r, co = cot * 64 + col;
                if (co < cout && ci < cin_true) gW[((size_t)co * cin_true + ci) * KK + t] += tile[col][cil];
            }
        }
        __syncthreads();
    }
}


// gW[(co*cin_true + ci)*KK + t] += gacc[t][ci/4][co][ci%4].  blockIdx.y selects the layer (one launch for the whole net).  A tile is
// (co_t output channels) x (16 input channels) x ALL taps: the accumulation buffer is read with co-fastest 16-byte loads, and
// each output channel's 16*KK gradient elements are contiguous in the Torch7 order, so the read-modify-write of gW is coalesced
// and touches every sector once (a tile per tap wrote 4 of every 36 bytes, nine times over).
#define FIN_SMEM_FLOATS (16 * (16 * 9 + 1))
__global__ void __launch_bounds__(256) wgrad_finalize_kernel(const FinalizeLayer *__restrict__ table, unsigned long long *tl) {
    extern __shared__ float tile[];
    DG_TL_IN(tl, DG_TL_FINALIZE);                  // [co_t][16 * KK + 1]
    const FinalizeLayer fl = table[blockIdx.y];
    const float *__restrict__ gacc = fl.gacc;
    float *__restrict__ gW = fl.gW;
    const int cin_true = fl.cin_true, cin = fl.cin, cout = fl.cout, ksize = fl.ksize;
    const int KK = ksize * ksize, cin4 = cin >> 2;
    const int rowlen = 16 * KK, pitch = rowlen + 1;
    int co_t = 16;
    while (co_t * pitch > FIN_SMEM_FLOATS) co_t >>= 1;                   // 5x5: 4 output channels per tile
    const int co_tiles = (cout + co_t - 1) / co_t, ci_tiles = cin >> 4;
    const float4 *g4 = reinterpret_cast<const float4 *>(gacc);
    for (int tl = blockIdx.x; tl < co_tiles * ci_tiles; tl += gridDim.x) {
        const int cit = tl % ci_tiles, cot = tl / ci_tiles;
        // load: co fastest (coalesced 16-byte reads), then the 4 input-channel quads of the tile, then the taps
        for (int e = threadIdx.x; e < co_t * 4 * KK; e += blockDim.x) {
            const int col = e % co_t, c4l = (e / co_t) & 3, t = e / (co_t * 4);
            const int co = cot * co_t + col;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (co < cout) v = __ldg(g4 + ((size_t)t * cin4 + cit * 4 + c4l) * cout + co);
            float *tr = tile + col * pitch + (c4l * 4) * KK + t;
            tr[0] = v.x; tr[KK] = v.y; tr[2 * KK] = v.z; tr[3 * KK] = v.w;
        }
        __syncthreads();
        // store: for one output channel the tile's 16 * KK elements are contiguous in gW
        const int kmax = min(rowlen, (cin_true - cit * 16) * KK);         // the padded input channels of the first layer do not exist in gW
        for (int e = threadIdx.x; e < co_t * rowlen; e += blockDim.x) {
            const int col = e / rowlen, k = e - col * rowlen;
            const int co = cot * co_t + col;
            if (co < cout && k < kmax) gW[((size_t)co * cin_true + cit * 16) * KK + k] += tile[col * pitch + k];
        }
        __syncthreads();
    }
    // ga[c][p] += gab[(rho(p) - first)/4][c][..] ; gb[c] += sum_p of the same (nn.Add gradBias, conv gradBias)
    if (fl.gab) {
        float *red = tile;
        for (int c = blockIdx.x; c < cout; c += gridDim.x) {
            float s = 0.f;
            for (int pt = threadIdx.x; pt < DG_POINTS; pt += blockDim.x) {
                const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
                const int rr = (x + fl.pad) * fl.pitch + y + fl.pad - fl.first;
                const float v = __ldg(fl.gab + ((size_t)(rr >> 2) * cout + c) * 4 + (rr & 3));
                fl.ga[(size_t)c * DG_POINTS + pt] += v;
                s += v;
            }
#pragma unroll
            for (int sh = 16; sh > 0; sh >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sh);
            if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
            __syncthreads();
            if (threadIdx.x == 0) {
                float t = 0.f;
                for (int w = 0; w < 8; ++w) t += red[w];
                fl.gb[c] += t;
            }
            __syncthreads();
        }
    }
    DG_TL_OUT(tl, DG_TL_FINALIZE);
}

int dg_launch_wgrad_finalize(dg_ctx *c, const FinalizeLayer *table_dev, int nlayers, cudaStream_t st) {
    dg_prof_scope _ps(c, DG_K_WGRAD_REDUCE, st);
    wgrad_finalize_kernel<<<dim3(96, nlayers), 256, FIN_SMEM_FLOATS * 4, st>>>(table_dev, nullptr);
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// per-point / per-channel bias gradients from the blocked bf16 dZ image
// ---------------------------------------------------------------------------------------------
// ga[c][p] += sum_n dz[n][c][rho(p)] ; gb[c] += sum_{n,p} dz  (nn.Add gradBias, conv gradBias)
__global__ void __launch_bounds__(384) bias_grad_bf16_kernel(const bf16 *__restrict__ dz, float *__restrict__ ga, float *__restrict__ gb,
                                                             int n, int cout, int R, int pitch, int pad, int groups) {
    __shared__ float red[12][8];
    const int chunk = blockIdx.x, grp = blockIdx.y, p = threadIdx.x;
    const int chunks = cout >> 3;
    float s[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) s[e] = 0.f;
    if (p < DG_POINTS) {
        const int x = p / DG_BOARD, y = p - x * DG_BOARD;
        const int rho = (x + pad) * pitch + y + pad;
#pragma unroll 4
        for (int i = grp; i < n; i += groups) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(dz + (((size_t)i * chunks + chunk) * R + rho) * 8));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                s[2 * e] += __uint_as_float(w[e] << 16);
                s[2 * e + 1] += __uint_as_float(w[e] & 0xFFFF0000u);
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(&ga[(size_t)(chunk * 8 + e) * DG_POINTS + p], s[e]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        float v = s[e];
#pragma unroll
        for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
        if ((p & 31) == 0) red[p >> 5][e] = v;
    }
    __syncthreads();
    if (p < 8) {
        float t = 0.f;
        for (int w = 0; w < 12; ++w) t += red[w][p];
        atomicAdd(&gb[chunk * 8 + p], t);
    }
}

// ---------------------------------------------------------------------------------------------
// head: last layer (C -> 1, 3x3) + bias + ReLU + log-softmax + NLL + top-1, and its backward
// (experiments.lua:138-151 for the last layer; train.lua:8-10), one CTA per position.
// ---------------------------------------------------------------------------------------------
struct HeadParams {
    const bf16 *in;          // H_{L-1}: [n][C/8][R][8], 3x3 geometry
    const float *W, *b, *a;  // master parameters of the last layer: W[0][ci][3][3], b[0], a[0][19][19]
    const int32_t *targets;  // may be NULL
    float *logp, *loss_n;
    int32_t *argmax;
    bf16 *dz_out;            // dZ_{L-1}: [n][C/8][R_out][8] (NULL: forward only)
    long long *dbg;          // optional phase timestamps (DG_TIMELINE=1)
    unsigned long long *tl;  // optional step timeline (DG_TL=1)
    float *part;             // per-position partial gradients [n][9*C + 361]: weight-gradient partials, then dz of the 1-channel map
                             // (summed over positions by head_grad_reduce_kernel: no same-address atomics from 128 CTAs)
    int n, C, R;
    int R_out, pitch_out, pad_out;
    float scale;
};

__device__ __forceinline__ float block_reduce_max(float v, float *sh) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, s));
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sh[0];
    for (int w = 1; w < 12; ++w) r = fmaxf(r, sh[w]);
    __syncthreads();
    return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float *sh) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    for (int w = 0; w < 12; ++w) r += sh[w];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(384, 1) head_bf16_kernel(const HeadParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ float red[12];
    __shared__ int redi[12];
    __shared__ float dzs[DG_R3 + 32];            // dz of the 1-channel map by image row, zero at pads
    const int pos = blockIdx.x, tid = threadIdx.x;
#define HSTAMP(k) do { if (p.dbg && tid == 0) p.dbg[blockIdx.x * 8 + (k)] = clock64(); } while (0)
    HSTAMP(0);
    const int chunks = p.C >> 3;
    const uint32_t img_bytes = (uint32_t)chunks * p.R * 16;
    uint8_t *img = smem;
    float *ws = reinterpret_cast<float *>(smem + ((img_bytes + 1023) & ~1023u));    // [9][C]
    float *gws = ws + 9 * p.C;                                                         // [9][C] partial weight gradient

    dg_pdl_launch_dependents();
    DG_TL_IN(p.tl, DG_TL_HEAD);
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    __syncthreads();
    dg_pdl_wait();
    DG_TL_GO(p.tl, DG_TL_HEAD);
    if (tid == 0) {
        mbar_arrive_expect_tx(&bar, img_bytes);
        const uint8_t *src = (const uint8_t *)p.in + (size_t)pos * img_bytes;
        for (int c = 0; c < chunks; ++c) bulk_g2s(img + (size_t)c * p.R * 16, src + (size_t)c * p.R * 16, (uint32_t)p.R * 16, &bar);
    }
    for (int i = tid; i < 9 * p.C; i += blockDim.x) {
        const int t = i / p.C, c = i - t * p.C;
        ws[i] = p.W[c * 9 + t];
        gws[i] = 0.f;
    }
    for (int i = tid; i < DG_R3 + 32; i += blockDim.x) dzs[i] = 0.f;
    __syncthreads();
    mbar_wait(&bar, 0);
    HSTAMP(1);

    // ---- forward: one thread per image row 21..404
    const int rho = 21 + tid;
    const int xp = rho / 20, yp = rho - xp * 20;
    const int x = xp - 1, y = yp - 1;
    const bool valid = x >= 0 && x < DG_BOARD && y >= 0 && y < DG_BOARD;
    const int pt = x * DG_BOARD + y;
    float z = 0.f;
    if (valid) {
        float acc = 0.f;
        for (int t = 0; t < 9; ++t) {
            const int row = rho + (t / 3 - 1) * 20 + (t % 3 - 1);
            const float *wt = ws + t * p.C;
            for (int c = 0; c < chunks; ++c) {
                const uint4 v = *reinterpret_cast<const uint4 *>(img + ((size_t)c * p.R + row) * 16);
                // weights as two 16-byte broadcast loads per chunk (the loop is shared-memory-instruction bound)
                const float4 wa = *reinterpret_cast<const float4 *>(wt + c * 8), wb = *reinterpret_cast<const float4 *>(wt + c * 8 + 4);
                acc = fmaf(__uint_as_float(v.x << 16), wa.x, acc); acc = fmaf(__uint_as_float(v.x & 0xFFFF0000u), wa.y, acc);
                acc = fmaf(__uint_as_float(v.y << 16), wa.z, acc); acc = fmaf(__uint_as_float(v.y & 0xFFFF0000u), wa.w, acc);
                acc = fmaf(__uint_as_float(v.z << 16), wb.x, acc); acc = fmaf(__uint_as_float(v.z & 0xFFFF0000u), wb.y, acc);
                acc = fmaf(__uint_as_float(v.w << 16), wb.z, acc); acc = fmaf(__uint_as_float(v.w & 0xFFFF0000u), wb.w, acc);
            }
        }
        z = acc + p.b[0] + p.a[pt];
    }
    HSTAMP(2);
    const float h = valid ? fmaxf(z, 0.f) : -INFINITY;           // ReLU stays before LogSoftMax (experiments.lua:147)
    const float mx = block_reduce_max(h, red);
    const float se = block_reduce_sum(valid ? expf(h - mx) : 0.f, red);
    const float lse = mx + logf(se);
    const float lp = h - lse;
    if (valid) p.logp[(size_t)pos * DG_POINTS + pt] = lp;
    // first index of the maximum
    {
        int cand = (valid && h == mx) ? pt : 0x7fffffff;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, s));
        if ((tid & 31) == 0) redi[tid >> 5] = cand;
        __syncthreads();
        if (tid == 0) {
            int b = redi[0];
            for (int w = 1; w < 12; ++w) b = min(b, redi[w]);
            if (p.argmax) p.argmax[pos] = b;
        }
        __syncthreads();
    }
    const int tgt = p.targets ? p.targets[pos] : -1;
    if (valid && pt == tgt && p.loss_n) p.loss_n[pos] = -lp;
    HSTAMP(3);
    if (!p.dz_out) { DG_TL_OUT(p.tl, DG_TL_HEAD); return; }

    // ---- backward of LogSoftMax + NLL + ReLU on the 1-channel map
    float dzv = 0.f;
    if (valid && z > 0.f) dzv = p.scale * (expf(lp) - (pt == tgt ? 1.f : 0.f));
    dzs[rho] = dzv;
    float *part = p.part + (size_t)pos * (9 * p.C + DG_POINTS);
    if (valid) part[9 * p.C + pt] = dzv;                         // nn.Add gradBias (361) and conv gradBias (1) are sums of these over positions
    __syncthreads();                                             // orders the dzs writes

    // ---- weight gradient of the last layer: gW[ci][t] = sum_rho dz[rho] * in[rho+off(t)][ci]
    // one warp per 8-channel chunk, lanes stride over the rows (conflict-free 16-byte reads), 72 accumulators per lane,
    // warp-shuffle reduction, one writer per element
    {
        const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;
        for (int c = warp; c < chunks; c += nwarps) {
            float acc[9][8];
#pragma unroll
            for (int t = 0; t < 9; ++t)
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[t][e] = 0.f;
            for (int r = 21 + lane; r < 400; r += 32) {
                const float d = dzs[r];
                if (d == 0.f) continue;
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    const int row = r + (t / 3 - 1) * 20 + (t % 3 - 1);
                    const uint4 v = *reinterpret_cast<const uint4 *>(img + ((size_t)c * p.R + row) * 16);
                    const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        acc[t][2 * e] = fmaf(d, __uint_as_float(w4[e] << 16), acc[t][2 * e]);
                        acc[t][2 * e + 1] = fmaf(d, __uint_as_float(w4[e] & 0xFFFF0000u), acc[t][2 * e + 1]);
                    }
                }
            }
#pragma unroll
            for (int t = 0; t < 9; ++t)
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float v = acc[t][e];
#pragma unroll
                    for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
                    if (lane == 0) part[(c * 8 + e) * 9 + t] = v;        // Torch7 order of W[0][ci][ky][kx]
                }
        }
    }

    __syncthreads();
    HSTAMP(4);
    // ---- input gradient + ReLU mask of layer L-1: dZ[rho][c] = (in[rho][c] > 0) * sum_t dz[rho - off(t)] * W[c][t]
    if (valid) {
        float d9[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) d9[t] = dzs[rho - ((t / 3 - 1) * 20 + (t % 3 - 1))];
        const int rho_out = (x + p.pad_out) * p.pitch_out + y + p.pad_out;
        for (int c = 0; c < chunks; ++c) {
            const uint4 hv = *reinterpret_cast<const uint4 *>(img + ((size_t)c * p.R + rho) * 16);
            const uint32_t hw[4] = {hv.x, hv.y, hv.z, hv.w};
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = 0.f;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const float4 wa = *reinterpret_cast<const float4 *>(ws + t * p.C + c * 8), wb = *reinterpret_cast<const float4 *>(ws + t * p.C + c * 8 + 4);
                v[0] = fmaf(d9[t], wa.x, v[0]); v[1] = fmaf(d9[t], wa.y, v[1]); v[2] = fmaf(d9[t], wa.z, v[2]); v[3] = fmaf(d9[t], wa.w, v[3]);
                v[4] = fmaf(d9[t], wb.x, v[4]); v[5] = fmaf(d9[t], wb.y, v[5]); v[6] = fmaf(d9[t], wb.z, v[6]); v[7] = fmaf(d9[t], wb.w, v[7]);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const uint32_t hh = (hw[e >> 1] >> ((e & 1) * 16)) & 0xFFFFu;
                if (!(hh != 0 && (hh & 0x8000u) == 0)) v[e] = 0.f;
            }
            uint4 o;
            __nv_bfloat162 t0 = __floats2bfloat162_rn(v[0], v[1]), t1 = __floats2bfloat162_rn(v[2], v[3]);
            __nv_bfloat162 t2 = __floats2bfloat162_rn(v[4], v[5]), t3 = __floats2bfloat162_rn(v[6], v[7]);
            o.x = *reinterpret_cast<uint32_t *>(&t0); o.y = *reinterpret_cast<uint32_t *>(&t1);
            o.z = *reinterpret_cast<uint32_t *>(&t2); o.w = *reinterpret_cast<uint32_t *>(&t3);
            *reinterpret_cast<uint4 *>(p.dz_out + (((size_t)pos * chunks + c) * p.R_out + rho_out) * 8) = o;
        }
    }
    HSTAMP(5);
    DG_TL_OUT(p.tl, DG_TL_HEAD);
}

// ---------------------------------------------------------------------------------------------
// head on the tensor cores (C <= 128): the same math as head_bf16_kernel, three small GEMMs instead of SIMT loops.
//   logits   G[row][t] = sum_c in[row][c] * W[c][t]            (M = rows, K = C, N = 32), then z[rho] = sum_t G[rho + off(t)][t]
//   gW       D[c][t]   = sum_row in[row][c] * dz[row - off(t)]  (M = C -> 128, K = 432 rows, N = 32)
//   dZ_{L-1} [rho][c]  = sum_t dz[rho - off(t)] * W[c][t]       (M = rows, K = 32, N = C), then the ReLU mask of in
// fp32 factors are split into bf16 hi + lo columns (hi*hi + lo*hi + hi*lo), so the results keep fp32-level accuracy.
// The 32 "tap" columns j are: 0..8 hi(t), 9..17 lo(t), 18..26 hi(t) again (pairs with the lo weights), 27..31 zero.
// ---------------------------------------------------------------------------------------------
#define HU_ROWS DG_R3
__global__ void __launch_bounds__(384, 1) head_umma_kernel(const HeadParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar_img, bar_mma;
    __shared__ uint32_t tmem_base_s;
    __shared__ float red[12];
    __shared__ int redi[12];
    __shared__ float dzs[HU_ROWS + 64];          // dz of the 1-channel map by image row (+32), zero at pads and beyond
    const int pos = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    HSTAMP(0);
    const int chunks = p.C >> 3;
    const uint32_t img_bytes = (uint32_t)chunks * HU_ROWS * 16;
    // shared-memory map (the image region is sized for 16 chunks: the M = 128 weight-gradient MMA reads that far)
    uint8_t *img = smem;                                         // [C/8][432][8] bf16
    uint8_t *bw = smem + 16 * HU_ROWS * 16 + 2048;               // [C/8][32][8] bf16: W hi/lo columns, K-major B of the logits GEMM
    uint8_t *b2 = bw + 16 * 32 * 16;                             // [4][128][8] bf16: W rows by tap column, K-major B of the dZ GEMM
    uint8_t *dzim = b2 + 4 * 128 * 16;                           // [4][432][8] bf16: shifted dz columns (A of dZ, B of gW)
    float *gs = reinterpret_cast<float *>(dzim + 4 * HU_ROWS * 16);   // [432][9] fp32: G

    dg_pdl_launch_dependents();
    DG_TL_IN(p.tl, DG_TL_HEAD);
    if (tid == 0) { mbar_init(&bar_img, 1); mbar_init(&bar_mma, 1); fence_barrier_init(); }
    if (warp == 1) { tmem_alloc(&tmem_base_s, 512); tmem_relinquish(); }
    // weights are parameters (not produced by the previous kernel): pack them before the dependency wait
    for (int i = tid; i < 32 * p.C; i += blockDim.x) {
        const int c = i % p.C, j = i / p.C;                      // tap column j of channel c
        const int t = j < 9 ? j : (j < 18 ? j - 9 : j - 18);
        float hi = 0.f, lo = 0.f;
        if (j < 27) {
            const float w = p.W[c * 9 + t];
            const bf16 h = __float2bfloat16(w);
            hi = __bfloat162float(h);
            lo = w - hi;
        }
        // logits B: columns 0..8 hi, 9..17 lo;  dZ B: rows 0..8 hi (x dz hi), 9..17 hi (x dz lo), 18..26 lo (x dz hi)
        reinterpret_cast<bf16 *>(bw)[((size_t)(c >> 3) * 32 + j) * 8 + (c & 7)] = __float2bfloat16(j < 9 ? hi : (j < 18 ? lo : 0.f));
        reinterpret_cast<bf16 *>(b2)[((size_t)(j >> 3) * 128 + c) * 8 + (j & 7)] = __float2bfloat16(j < 18 ? hi : (j < 27 ? lo : 0.f));
    }
    for (int i = tid; i < HU_ROWS + 64; i += blockDim.x) dzs[i] = 0.f;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    dg_pdl_wait();
    DG_TL_GO(p.tl, DG_TL_HEAD);
    if (tid == 0) {
        mbar_arrive_expect_tx(&bar_img, img_bytes);
        const uint8_t *src = (const uint8_t *)p.in + (size_t)pos * img_bytes;
        for (int c = 0; c < chunks; ++c) bulk_g2s(img + (size_t)c * HU_ROWS * 16, src + (size_t)c * HU_ROWS * 16, (uint32_t)HU_ROWS * 16, &bar_img);
    }
    mbar_wait(&bar_img, 0);
    HSTAMP(1);

    // ---- G = in x W: 4 row tiles (rows 0..511; the last tile runs past the image, those rows are never read back)
    if (warp == 0) {
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, 32, 0, 0);
            const uint32_t a_hi = desc_hi(128), b_hi = desc_hi(128);
            for (int tile = 0; tile < 4; ++tile) {
                uint32_t a_lo = desc_lo(smem_u32(img) + (uint32_t)tile * 128 * 16, (uint32_t)HU_ROWS * 16);
                uint32_t b_lo = desc_lo(smem_u32(bw), 32 * 16);
                for (int ks = 0; ks < (p.C >> 4); ++ks, a_lo += 2 * HU_ROWS, b_lo += 2 * 32)
                    mma_bf16(tmem_base + 384 + tile * 32, desc_join(a_lo, a_hi), desc_join(b_lo, b_hi), idesc, ks != 0);
            }
            mma_commit(&bar_mma);
        }
        __syncwarp();
    }
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    {
        // warp w reads TMEM lanes 32*(w%4)..: tiles w/4 (0..2), and tile 3 again by warps 0..3
        const int q = warp & 3;
        for (int tile = warp >> 2; tile < 4; tile += 3) {
            const int row = tile * 128 + q * 32 + lane;
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 384 + tile * 32, r);
            tmem_ld_wait();
            if (row < HU_ROWS) {
#pragma unroll
                for (int t = 0; t < 9; ++t) gs[row * 9 + t] = __uint_as_float(r[t]) + __uint_as_float(r[9 + t]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    HSTAMP(2);

    // ---- logits, log-softmax, NLL, top-1: one thread per image row 21..404
    const int rho = 21 + tid;
    const int xp = rho / 20, yp = rho - xp * 20;
    const int x = xp - 1, y = yp - 1;
    const bool valid = x >= 0 && x < DG_BOARD && y >= 0 && y < DG_BOARD;
    const int pt = x * DG_BOARD + y;
    float z = 0.f;
    if (valid) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < 9; ++t) acc += gs[(rho + (t / 3 - 1) * 20 + (t % 3 - 1)) * 9 + t];
        z = acc + p.b[0] + p.a[pt];
    }
    const float h = valid ? fmaxf(z, 0.f) : -INFINITY;           // ReLU stays before LogSoftMax (experiments.lua:147)
    const float mx = block_reduce_max(h, red);
    const float se = block_reduce_sum(valid ? expf(h - mx) : 0.f, red);
    const float lse = mx + logf(se);
    const float lp = h - lse;
    if (valid) p.logp[(size_t)pos * DG_POINTS + pt] = lp;
    {
        int cand = (valid && h == mx) ? pt : 0x7fffffff;          // first index of the maximum
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, s));
        if (lane == 0) redi[warp] = cand;
        __syncthreads();
        if (tid == 0) {
            int bst = redi[0];
            for (int w = 1; w < 12; ++w) bst = min(bst, redi[w]);
            if (p.argmax) p.argmax[pos] = bst;
        }
        __syncthreads();
    }
    const int tgt = p.targets ? p.targets[pos] : -1;
    if (valid && pt == tgt && p.loss_n) p.loss_n[pos] = -lp;
    HSTAMP(3);
    if (p.dz_out) {
        // ---- backward of LogSoftMax + NLL + ReLU on the 1-channel map
        float dzv = 0.f;
        if (valid && z > 0.f) dzv = p.scale * (expf(lp) - (pt == tgt ? 1.f : 0.f));
        dzs[32 + rho] = dzv;
        float *part = p.part + (size_t)pos * (9 * p.C + DG_POINTS);
        if (valid) part[9 * p.C + pt] = dzv;                     // nn.Add gradBias (361) and conv gradBias (1) are sums of these over positions
        __syncthreads();
        // shifted dz columns: dzim[row][j] = hi/lo of dz[row - off(t)]
        for (int row = tid; row < HU_ROWS; row += blockDim.x) {
            __align__(16) bf16 col[32];
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const float d = dzs[32 + row - ((t / 3 - 1) * 20 + (t % 3 - 1))];
                const bf16 hh = __float2bfloat16(d);
                col[t] = hh; col[18 + t] = hh;
                col[9 + t] = __float2bfloat16(d - __bfloat162float(hh));
            }
#pragma unroll
            for (int j = 27; j < 32; ++j) col[j] = __float2bfloat16(0.f);
#pragma unroll
            for (int g = 0; g < 4; ++g)
                *reinterpret_cast<uint4 *>(dzim + ((size_t)g * HU_ROWS + row) * 16) = *reinterpret_cast<const uint4 *>(col + 8 * g);
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        if (warp == 0) {
            if (elect_one()) {
                // gW: M = channels (MN-major image, 16 chunks), N = 32 tap columns (MN-major dzim), K = all 432 rows
                const uint32_t idesc_w = make_idesc_bf16(128, 32, 1, 1);
                const uint32_t hi_mn = desc_hi((uint32_t)HU_ROWS * 16);
                uint32_t a_lo = desc_lo(smem_u32(img), 128), b_lo = desc_lo(smem_u32(dzim), 128);
                for (int ks = 0; ks < HU_ROWS / 16; ++ks, a_lo += 16, b_lo += 16)
                    mma_bf16(tmem_base + 384, desc_join(a_lo, hi_mn), desc_join(b_lo, hi_mn), idesc_w, ks != 0);
                // dZ: M = rows 21 + 128*tile, K = 32 tap columns (K-major dzim), N = C channels (K-major b2)
                const uint32_t idesc_d = make_idesc_bf16(128, p.C, 0, 0);
                const uint32_t hi_k = desc_hi(128);
                for (int tile = 0; tile < 3; ++tile) {
                    uint32_t a2 = desc_lo(smem_u32(dzim) + (uint32_t)(21 + tile * 128) * 16, (uint32_t)HU_ROWS * 16);
                    uint32_t bb = desc_lo(smem_u32(b2), 128 * 16);
                    for (int ks = 0; ks < 2; ++ks, a2 += 2 * HU_ROWS, bb += 2 * 128)
                        mma_bf16(tmem_base + tile * p.C, desc_join(a2, hi_k), desc_join(bb, hi_k), idesc_d, ks != 0);
                }
                mma_commit(&bar_mma);
            }
            __syncwarp();
        }
        mbar_wait(&bar_mma, 1);
        tc_fence_after();
        HSTAMP(4);
        const int q = warp & 3;
        if (warp < 4) {
            // weight gradient of the last layer: lane = input channel
            const int ci = q * 32 + lane;
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 384, r);
            tmem_ld_wait();
            if (ci < p.C) {
#pragma unroll
                for (int t = 0; t < 9; ++t) part[ci * 9 + t] = __uint_as_float(r[t]) + __uint_as_float(r[9 + t]);   // Torch7 order of W[0][ci][ky][kx]
            }
        }
        {
            // dZ_{L-1}: warp w drains tile w/4, its lane quarter; ReLU mask from the image in shared memory
            const int tile = warp >> 2;
            const int ro = 21 + tile * 128 + q * 32 + lane;
            const int xq = ro / 20 - 1, yq = ro - (ro / 20) * 20 - 1;
            const bool ok = xq >= 0 && xq < DG_BOARD && yq >= 0 && yq < DG_BOARD;
            const int rho_out = (xq + p.pad_out) * p.pitch_out + yq + p.pad_out;
            const __nv_bfloat162 zero2 = __floats2bfloat162_rn(0.f, 0.f);
            for (int c0 = 0; c0 < p.C; c0 += 32) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + tile * p.C + c0, r);
                tmem_ld_wait();
                if (ok) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int chunk = (c0 >> 3) + g;
                        const uint4 hv = *reinterpret_cast<const uint4 *>(img + ((size_t)chunk * HU_ROWS + ro) * 16);
                        const uint32_t hw[4] = {hv.x, hv.y, hv.z, hv.w};
                        uint32_t w[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(r[g * 8 + 2 * j]), __uint_as_float(r[g * 8 + 2 * j + 1]));
                            w[j] = *reinterpret_cast<const uint32_t *>(&t2) & __hgt2_mask(*reinterpret_cast<const __nv_bfloat162 *>(&hw[j]), zero2);
                        }
                        *reinterpret_cast<uint4 *>(p.dz_out + (((size_t)pos * chunks + chunk) * p.R_out + rho_out) * 8) = make_uint4(w[0], w[1], w[2], w[3]);
                    }
                }
            }
        }
        HSTAMP(5);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
    DG_TL_OUT(p.tl, DG_TL_HEAD);
}

// gW[i] += sum_n part[n][i] (i < 9C) ; ga[pt] += sum_n dz[n][pt] ; gb += sum_{n,pt} dz   -- fixed summation order.
// A block owns 32 columns; its 8 warps take the positions n = w, w+8, .. and the partial sums meet in shared memory.
// The block after the last column block reduces the step's loss (mean NLL, fixed order) and top-1 error count, so reading the loss
// after dg_train_step needs no further kernel.
__global__ void __launch_bounds__(256) head_grad_reduce_kernel(const float *__restrict__ part, int n, int C, float *__restrict__ gW,
                                                               float *__restrict__ gb, float *__restrict__ ga,
                                                               const float *__restrict__ loss_n, const int32_t *__restrict__ argmax,
                                                               const int32_t *__restrict__ targets, float *__restrict__ loss_out,
                                                               int *__restrict__ errors_out) {
    __shared__ float red[8][32];
    const int width = 9 * C + DG_POINTS;
    if (blockIdx.x * 32 >= width) {
        __shared__ float sl[256];
        __shared__ int se[256];
        float s = 0.f;
        int e = 0;
        for (int i = threadIdx.x; i < n; i += 256) {
            s += loss_n[i];
            if (argmax && targets) e += argmax[i] != targets[i];
        }
        sl[threadIdx.x] = s; se[threadIdx.x] = e;
        __syncthreads();
        for (int w = 128; w > 0; w >>= 1) {
            if (threadIdx.x < w) { sl[threadIdx.x] += sl[threadIdx.x + w]; se[threadIdx.x] += se[threadIdx.x + w]; }
            __syncthreads();
        }
        if (threadIdx.x == 0) { *loss_out = sl[0] / (float)n; *errors_out = se[0]; }
        return;
    }
    const int lane = threadIdx.x & 31, grp = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (i < width) {
        int k = grp;
        for (; k + 24 < n; k += 32) {
            s0 += part[(size_t)k * width + i]; s1 += part[(size_t)(k + 8) * width + i];
            s2 += part[(size_t)(k + 16) * width + i]; s3 += part[(size_t)(k + 24) * width + i];
        }
        for (; k < n; k += 8) s0 += part[(size_t)k * width + i];
    }
    red[grp][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (grp != 0) return;
    float s = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) s += red[g][lane];
    float bsum = 0.f;
    if (i < 9 * C) gW[i] += s;
    else if (i < width) { ga[i - 9 * C] += s; bsum = s; }
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) bsum += __shfl_xor_sync(0xffffffffu, bsum, sh);
    if (lane == 0 && bsum != 0.f) atomicAdd(gb, bsum);     // the conv bias of the head: a handful of blocks hold its 361 elements
}

// ---------------------------------------------------------------------------------------------
// staging + parameter packing
// ---------------------------------------------------------------------------------------------

// raw records -> blocked bf16 input image [n][48/8][R][8] (planes 37..47 stay zero); same plane
// definitions as encode.cu (dataloader.lua:50-92).  One block per position:
//   * the 3,249-byte record is pulled into shared memory with 16-byte loads (records are odd-sized, so the block reads the aligned
//     window around its record; the window's last 16 bytes are fetched byte-wise where they would cross the end of the buffer);
//   * one thread per IMAGE ROW from the first to the last board point, pad rows included (they are written as zeros, which they
//     already are): every channel chunk is then ONE contiguous run of 16-byte stores, whole 32-byte sectors throughout, instead of
//     19-row runs with 2-row holes whose edge sectors are partial writes.
__global__ void __launch_bounds__(512) encode_blocked_kernel(const uint8_t *__restrict__ raw, const uint8_t *__restrict__ player,
                                                            const uint8_t *__restrict__ rank, const uint8_t *__restrict__ sym,
                                                            bf16 *__restrict__ out, int R, int pitch, int pad,
                                                            const uint8_t *__restrict__ mx, const uint8_t *__restrict__ my, int32_t *__restrict__ targets,
                                                            size_t raw_bytes) {
    __shared__ __align__(16) uint8_t rec_s[DG_RAW_PLANES * DG_POINTS + 32];
    const int n = blockIdx.x, tid = threadIdx.x;
    dg_pdl_launch_dependents();
    dg_pdl_wait();
    const size_t base = (size_t)n * DG_RAW_PLANES * DG_POINTS, abase = base & ~(size_t)15;
    const int shift = (int)(base - abase);
    const int nvec = (shift + DG_RAW_PLANES * DG_POINTS + 15) >> 4;
    for (int v = tid; v < nvec; v += blockDim.x) {
        const size_t off = abase + (size_t)v * 16;
        if (off + 16 <= raw_bytes) *reinterpret_cast<uint4 *>(rec_s + v * 16) = __ldg(reinterpret_cast<const uint4 *>(raw + off));
        else for (int e = 0; e < 16; ++e) rec_s[v * 16 + e] = off + e < raw_bytes ? raw[off + e] : 0;
    }
    const int sy = sym ? sym[n] : 0;
    if (tid == 511 && targets) {
        // the move label of the position (dataloader.lua:89), transformed like the board: saves the separate labels launch
        int a = mx[n] - 1, b = my[n] - 1;
        if (sy & 1) { int t = a; a = b; b = t; }
        if (sy & 2) a = DG_BOARD - 1 - a;
        if (sy & 4) b = DG_BOARD - 1 - b;
        targets[n] = DG_BOARD * a + b;
    }
    const int pl = player[n], rk = rank[n];
    __syncthreads();
    const int row_first = pad * pitch + pad, row_last = (DG_BOARD - 1 + pad) * pitch + DG_BOARD - 1 + pad;
    for (int rho = row_first + tid; rho <= row_last; rho += blockDim.x) {
        const int xp = rho / pitch, yp = rho - xp * pitch;
        const int x = xp - pad, y = yp - pad;
        unsigned long long m = 0;
        if (y >= 0 && y < DG_BOARD) m = dg_plane_bits(rec_s + shift, dg_src_point_d(x * DG_BOARD + y, sy), pl, rk);      // x is on the board for every row in range
#pragma unroll
        for (int c = 0; c < 6; ++c) __stcs(reinterpret_cast<uint4 *>(out + (((size_t)n * 6 + c) * R + rho) * 8), dg_plane_chunk(m, c));      // written once, read by another kernel: streaming stores
    }
}

// fp32 NCHW planes -> blocked bf16 input image
__global__ void __launch_bounds__(384) planes_to_blocked_kernel(const float *__restrict__ planes, bf16 *__restrict__ out, int cin_true,
                                                                int R, int pitch, int pad) {
    const int n = blockIdx.x, p = threadIdx.x;
    if (p >= DG_POINTS) return;
    const int x = p / DG_BOARD, y = p - x * DG_BOARD;
    const int rho = (x + pad) * pitch + y + pad;
    for (int c = 0; c < 6; ++c) {
        __align__(16) bf16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int ch = c * 8 + e;
            v[e] = __float2bfloat16(ch < cin_true ? planes[((size_t)n * cin_true + ch) * DG_POINTS + p] : 0.f);
        }
        *reinterpret_cast<uint4 *>(out + (((size_t)n * 6 + c) * R + rho) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
}

// master fp32 (Torch7 order) -> bf16 MMA operands of one layer:
//   wf[t][ci/8][co][8]  = W[co][ci][t]            (forward B operand; ci >= cin_true -> 0)
//   wd[t][co/8][ci][8]  = W[co][ci][KK-1-t]       (input-gradient B operand, taps flipped); may be NULL
//   ab[c/8][rho][8]     = bf16(b[c] + a[c][x][y]) (epilogue bias in the OUTPUT geometry)
struct PackLayer {
    const float *W, *b, *a;
    bf16 *wf, *wd;
    bf16 *ab;
    int cin_true, cin, cout, ksize;
};

__global__ void __launch_bounds__(256) pack_layer_kernel(const PackLayer *__restrict__ table, int R_out, int pitch_out, int pad_out) {
    dg_pdl_launch_dependents();
    dg_pdl_wait();
    const PackLayer pl = table[blockIdx.y];
    const float *__restrict__ W = pl.W, *__restrict__ b = pl.b, *__restrict__ a = pl.a;
    bf16 *__restrict__ wf = pl.wf, *__restrict__ wd = pl.wd;
    bf16 *__restrict__ ab = pl.ab;
    const int cin_true = pl.cin_true, cin = pl.cin, cout = pl.cout, ksize = pl.ksize;
    const int KK = ksize * ksize;
    const int nwf = KK * cin * cout;
    const int nab = cout * DG_POINTS;
    const int stride = gridDim.x * blockDim.x;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nwf; idx += stride) {
        const int e = idx & 7, co = (idx >> 3) % cout, c8 = (idx / (8 * cout)) % (cin >> 3), t = idx / (8 * cout * (cin >> 3));
        const int ci = c8 * 8 + e;
        wf[idx] = __float2bfloat16(ci < cin_true ? W[((size_t)co * cin_true + ci) * KK + t] : 0.f);
    }
    if (wd) {
        for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nwf; idx += stride) {
            const int e = idx & 7, ci = (idx >> 3) % cin, c8 = (idx / (8 * cin)) % (cout >> 3), t = idx / (8 * cin * (cout >> 3));
            const int co = c8 * 8 + e;
            wd[idx] = __float2bfloat16(ci < cin_true ? W[((size_t)co * cin_true + ci) * KK + (KK - 1 - t)] : 0.f);
        }
    }
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nab; idx += stride) {
        const int c = idx / DG_POINTS, pt = idx - c * DG_POINTS;
        const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
        const int rho = (x + pad_out) * pitch_out + y + pad_out;
        ab[((size_t)(c >> 3) * R_out + rho) * 8 + (c & 7)] = __float2bfloat16(b[c] + a[idx]);
    }
}

// SGD:step (optimizer.lua:25) fused with the operand repack: block (x = group of 8 output channels, y = conv layer) reads its
// 8 master weight rows and their gradients coalesced, writes p <- p + (-rate)*g back, and emits the bf16 operands from shared
// memory: forward copy as 128-byte runs (8 co x 8 ci), input-gradient copy as contiguous 16-byte (8 co) entries per ci, and
// the fp32 epilogue bias b[c] + a[c][x][y].  The 1-channel last layer (y = gridDim.y - 1) only gets the plain update.
struct SgdPackLayer {
    float *W, *b, *a;            // master parameters of the layer (Torch7 order)
    const float *gW, *gb, *ga;   // gradients
    bf16 *wf, *wd;
    bf16 *ab;
    int cin_true, cin, cout, ksize;
    // fused step: gradients straight from the weight-gradient kernel's accumulation buffers (NULL: read gW / gb / ga)
    const float *gacc;           // [tap][cin/4][cout][4]
    const float *gab;            // [(row - first)/4][cout][4]
    int g_pitch, g_pad, g_first; // geometry of the layer's dZ image
};

// data parallel: after the reduce-scatter over NVLink peer memory the summed gradient of slice r lives on rank r only; the SGD kernel
// pulls what it needs from the owner (world <= 1: everything is local)
struct SgdPeers { const float *base; const float *buf[DG_MAX_RANKS]; size_t n, per4; int world; };
__device__ __forceinline__ const float *sgd_grad_ptr(const SgdPeers &P, const float *local) {
    if (P.world <= 1) return local;
    const size_t off = (size_t)(local - P.base);
    if (off >= P.n) return local;                    // not in the exchanged allocation
    const int owner = min((int)((off >> 2) / P.per4), P.world - 1);
    return P.buf[owner] + off;
}
#ifndef SGD_CO_B
#define SGD_CO_B 4
#define SGD_NT 256
#endif
template <int CO_B, int NT>      // output channels per block, threads per block
__global__ void __launch_bounds__(NT) sgd_pack_kernel(const SgdPackLayer *__restrict__ table, float rate, const float *__restrict__ rate_dev,
                                                       int R_out, int pitch_out, int pad_out, int smem_row, unsigned long long *tl, const __grid_constant__ SgdPeers peers) {
    extern __shared__ float wrow[];                 // [CO_B][smem_row]: the block's updated weight rows
    __shared__ float bnew[CO_B], bsum[CO_B];
    __shared__ float bpart[(CO_B * DG_POINTS + NT - 1) / NT][NT / 32][2];   // per-warp partial sums of the per-point bias gradient
    dg_pdl_launch_dependents();
    DG_TL_IN(tl, DG_TL_SGD);
    dg_pdl_wait();
    DG_TL_GO(tl, DG_TL_SGD);
    const SgdPackLayer L = table[blockIdx.y];
    const float neg_rate = -(rate_dev ? *rate_dev : rate);
    const int KK = L.ksize * L.ksize, row = L.cin_true * KK;
    const int co0 = blockIdx.x * CO_B;
    if (co0 >= L.cout) { DG_TL_OUT(tl, DG_TL_SGD); return; }
    const int nco = min(CO_B, L.cout - co0);
    constexpr int NB = (CO_B * DG_POINTS + NT - 1) / NT;      // per-point bias elements per thread
    const bool from_acc = L.gacc != nullptr;
    if (threadIdx.x < CO_B) bsum[threadIdx.x] = 0.f;
    // the block is a chain of dependent global round trips: the master weights are requested first, together with the gradient
    float *Wp = L.W + (size_t)co0 * row;
    const int total = nco * row;
    constexpr int NPRE = 20;                        // covers CO_B * 9 * 128 / NT = 18 elements per thread
    float wpre[NPRE];
#pragma unroll
    for (int k = 0; k < NPRE; ++k) {
        const int i = threadIdx.x + k * NT;
        wpre[k] = i < total ? Wp[i] : 0.f;
    }
    if (from_acc) {
        // the accumulation buffer [tap][cin/4][cout][4] is read in ITS order (8 output channels x 16 bytes = one 128-byte line) and
        // transposed through shared memory into the rows the sweep below walks
        const int cin4 = L.cin >> 2;
        const float4 *g4 = reinterpret_cast<const float4 *>(L.gacc);
        for (int e = threadIdx.x; e < KK * cin4 * CO_B; e += NT) {
            const int r = e % CO_B, c4 = (e / CO_B) % cin4, t = e / (CO_B * cin4);
            if (r < nco) {
                const float4 g = __ldcg(reinterpret_cast<const float4 *>(sgd_grad_ptr(peers, reinterpret_cast<const float *>(g4 + ((size_t)t * cin4 + c4) * L.cout + co0 + r))));
                const float gj[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c4 * 4 + j < L.cin_true) wrow[r * smem_row + (c4 * 4 + j) * KK + t] = gj[j];
            }
        }
        __syncthreads();
    }
    // ---- weights: update in place, keep the new values in shared memory.  The 8 rows are contiguous in the master vector: a flat,
    // coalesced sweep; the block is latency-bound, so every thread first issues all its loads (weights, gradients, and the
    // per-point bias of the next phase) and only then computes
    const float *Gp = L.gW + (size_t)co0 * row;
    float av[NB], ag[NB];                           // per-point bias: nco * 361 elements
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const int i = threadIdx.x + k * NT;
        av[k] = ag[k] = 0.f;
        if (i < nco * DG_POINTS) {
            const size_t gi = (size_t)co0 * DG_POINTS + i;
            av[k] = L.a[gi];
            if (from_acc) {
                const int r = i / DG_POINTS, pt = i - r * DG_POINTS;
                const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
                const int rr = (x + L.g_pad) * L.g_pitch + y + L.g_pad - L.g_first;
                ag[k] = __ldcg(sgd_grad_ptr(peers, L.gab + ((size_t)(rr >> 2) * L.cout + co0 + r) * 4 + (rr & 3)));
            } else {
                ag[k] = __ldcg(sgd_grad_ptr(peers, L.ga + gi));
            }
        }
        if (from_acc) {
            // conv gradBias = sum over the points of nn.Add's gradBias.  A warp spans at most two channels (361 > 32): two shuffle
            // reductions per warp; the per-warp partial sums are added in a fixed order below (data-parallel replicas must stay
            // bit-identical, so no floating-point atomics here)
            const int r_lo = (int)((threadIdx.x & ~31u) + k * NT) / DG_POINTS;
            const int r_me = i / DG_POINTS;
            float s_lo = (i < nco * DG_POINTS && r_me == r_lo) ? ag[k] : 0.f, s_hi = (i < nco * DG_POINTS && r_me != r_lo) ? ag[k] : 0.f;
#pragma unroll
            for (int sh = 16; sh > 0; sh >>= 1) { s_lo += __shfl_xor_sync(0xffffffffu, s_lo, sh); s_hi += __shfl_xor_sync(0xffffffffu, s_hi, sh); }
            if ((threadIdx.x & 31) == 0) { bpart[k][threadIdx.x >> 5][0] = s_lo; bpart[k][threadIdx.x >> 5][1] = s_hi; }
        }
    }
#pragma unroll
    for (int k = 0; k < NPRE; ++k) {
        const int i = threadIdx.x + k * NT;
        if (i < total) {
            const int r = i / row;
            const float g = from_acc ? wrow[r * smem_row + (i - r * row)] : __ldcg(sgd_grad_ptr(peers, Gp + i));
            const float v = fmaf(neg_rate, g, wpre[k]);
            Wp[i] = v;
            if (L.wf) wrow[r * smem_row + (i - r * row)] = v;
        }
    }
    for (int i = threadIdx.x + NPRE * NT; i < total; i += NT) {      // wider layers: the rest, one element at a time
        const int r = i / row;
        const float g = from_acc ? wrow[r * smem_row + (i - r * row)] : __ldcg(sgd_grad_ptr(peers, Gp + i));
        const float v = fmaf(neg_rate, g, Wp[i]);
        Wp[i] = v;
        if (L.wf) wrow[r * smem_row + (i - r * row)] = v;
    }
    if (from_acc) {
        __syncthreads();
        if (threadIdx.x < nco) {
            float t = 0.f;
            for (int k = 0; k < NB; ++k)
                for (int w = 0; w < NT / 32; ++w) {
                    const int r_lo = (w * 32 + k * NT) / DG_POINTS;
                    if (r_lo == (int)threadIdx.x) t += bpart[k][w][0];
                    else if (r_lo + 1 == (int)threadIdx.x) t += bpart[k][w][1];
                }
            bsum[threadIdx.x] = t;
        }
        __syncthreads();
    }
    if (threadIdx.x < nco) {
        const float v = fmaf(neg_rate, from_acc ? bsum[threadIdx.x] : __ldcg(sgd_grad_ptr(peers, L.gb + co0 + threadIdx.x)), L.b[co0 + threadIdx.x]);
        L.b[co0 + threadIdx.x] = v;
        bnew[threadIdx.x] = v;
    }
    __syncthreads();
    // ---- per-point bias: update, and emit b + a in the blocked epilogue layout (32 contiguous bytes per point)
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const int i = threadIdx.x + k * NT;
        if (i < nco * DG_POINTS) {
            const int r = i / DG_POINTS, pt = i - r * DG_POINTS;
            const float v = fmaf(neg_rate, ag[k], av[k]);
            L.a[(size_t)co0 * DG_POINTS + i] = v;
            if (L.ab) {
                const int x = pt / DG_BOARD, y = pt - x * DG_BOARD;
                const int rho = (x + pad_out) * pitch_out + y + pad_out;
                L.ab[((size_t)(co0 >> 3) * R_out + rho) * 8 + (co0 & 7) + r] = __float2bfloat16(v + bnew[r]);
            }
        }
    }
    if (!L.wf) { DG_TL_OUT(tl, DG_TL_SGD); return; }
    // ---- forward operand wf[t][ci/8][co][8]: one thread per (t, ci8, r) writes 8 ci = 16 bytes; r fastest -> 16 * CO_B byte runs
    const int cin8 = L.cin >> 3;
    for (int i = threadIdx.x; i < KK * cin8 * CO_B; i += NT) {
        const int r = i % CO_B, c8 = (i / CO_B) % cin8, t = i / (CO_B * cin8);
        if (r >= nco) continue;
        __align__(16) bf16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int ci = c8 * 8 + e;
            v[e] = __float2bfloat16(ci < L.cin_true ? wrow[r * smem_row + ci * KK + t] : 0.f);
        }
        *reinterpret_cast<uint4 *>(L.wf + (((size_t)t * cin8 + c8) * L.cout + co0 + r) * 8) = *reinterpret_cast<const uint4 *>(v);
    }
    // ---- input-gradient operand wd[t'][co/8][ci][8] = W[co][ci][KK-1-t']: one thread per (t', ci) writes the CO_B co of this block
    if (L.wd && nco == CO_B) {
        for (int i = threadIdx.x; i < KK * L.cin; i += NT) {
            const int ci = i % L.cin, t = i / L.cin;
            __align__(16) bf16 v[CO_B];
#pragma unroll
            for (int r = 0; r < CO_B; ++r) v[r] = __float2bfloat16(ci < L.cin_true ? wrow[r * smem_row + ci * KK + (KK - 1 - t)] : 0.f);
            bf16 *dst = L.wd + (((size_t)t * (L.cout >> 3) + (co0 >> 3)) * L.cin + ci) * 8 + (co0 & 7);
            if (CO_B == 8) *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(v);
            else if (CO_B == 4) *reinterpret_cast<uint2 *>(dst) = *reinterpret_cast<const uint2 *>(v);
            else *reinterpret_cast<uint32_t *>(dst) = *reinterpret_cast<const uint32_t *>(v);
        }
    }
    DG_TL_OUT(tl, DG_TL_SGD);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
struct LayerOps {
    bf16 *wf = nullptr, *wd = nullptr;
    bf16 *ab = nullptr;
    int cin_pad = 0;
};

struct Bf16Path {
    Geom g_first, g3;
    int cin0_pad = 48;
    bf16 *x0 = nullptr;                    // [B][6][R_first][8]
    std::vector<bf16 *> h;                 // h[l], l = 1..L-1: [B][C/8][R3][8]
    std::vector<bf16 *> dzl;               // dZ of conv layer l (0-based): l = 0 in the first layer's geometry, l >= 1 in the 3x3 geometry
    std::vector<LayerOps> ops;             // per conv layer 0..L-2
    float *head_part = nullptr;            // [B][9*C + 361] per-position partial gradients of the last layer
    PackLayer *pack_table = nullptr;       // device copy, one entry per conv layer
    FinalizeLayer *fin_table = nullptr;    // device copy, one entry per conv layer
    SgdPackLayer *sgd_table = nullptr;     // device copy, one entry per layer (the 1-channel last layer included)
    SgdPackLayer *sgd_table_acc = nullptr; // same, conv layers reading the accumulation buffers (fused step)
    std::vector<SgdPackLayer> sgd_host;
    dg_p2p *p2p = nullptr;                 // data parallel: the peers' accumulation buffers mapped over NVLink (NULL: NCCL)
    bool x0_from_raw = false;              // the staged batch is raw records: the first conv layer encodes them itself (fused encode)
    bool p2p_tried = false;
    bool p2p_pull = false;                 // the exchange is a reduce-scatter, the SGD kernel pulls each slice from its owner
    bool defer_grads = false;              // fused steps hand the conv gradients to the SGD kernel in the accumulation buffers
    bool grads_pending = false;            // the conv layers' gradient of the last fused step is still in gacc / gab, not in m->grads
    int sgd_smem_row = 0, sgd_max_co8 = 0;
    std::vector<FinalizeLayer> fin_host;
    float *gacc = nullptr;                 // weight-gradient accumulation buffers of all conv layers (default path)
    std::vector<size_t> gacc_off;          // per conv layer, in floats
    std::vector<size_t> gab_off;           // bias-gradient accumulation buffers, same allocation
    bool fused_bias = false;               // bias gradients come out of the weight-gradient kernel
    size_t gacc_floats = 0;
    size_t head_n = 0;                     // floats of the head layer's gradient (gW, gb, ga), the first block of gacc
    bool deterministic = false;            // DG_DETERMINISTIC=1: partial tiles + ordered reduction instead of L2 reductions
    float *scratch = nullptr;              // weight-gradient partial tiles, two halves used alternately
    size_t scratch_bytes = 0;              // bytes of one half
    // the small HBM-bound gradient kernels (bias gradients, split-K reduction) run on a side stream, under the
    // tensor-core kernels of the main stream
    cudaStream_t side = nullptr;
    cudaEvent_t ev_head = nullptr, ev_fork = nullptr, ev_memset = nullptr, ev_dz = nullptr, ev_wgrad = nullptr, ev_bias[2] = {nullptr, nullptr}, ev_reduce[2] = {nullptr, nullptr}, ev_join = nullptr;
    int num_sms = 148;
    bool head_bwd_done = false;
};

static size_t img_elems(int C, int R) { return (size_t)C * R; }

static int conv_cfg(const dg_model *m, ConvParams &cp, int cin, int cout) {
    cp.cin = cin; cp.cout = cout;
    cp.nb = cout <= 128 ? cout : (cout % 96 == 0 ? 96 : (cout % 128 == 0 ? 128 : 64));
    DG_REQUIRE(cout % cp.nb == 0 && cp.nb % 16 == 0, "bf16 path: unsupported channel count %d", cout);
    cp.kp = cin <= 128 ? cin : (cin % 64 == 0 ? 64 : cin);   // one ring stage = one tap x kp input channels
    DG_REQUIRE(cin % cp.kp == 0 && cp.kp % 16 == 0 && cin / cp.kp <= CONV_MAX_KPANELS, "bf16 path: unsupported input channel count %d", cin);
    DG_REQUIRE(cp.ntiles * cp.nb <= 512, "bf16 path: accumulator does not fit TMEM");
    return 0;
}

static int launch_conv(dg_model *m, ConvParams cp) {
    dg_ctx *c = m->ctx;
    static long long *dbg = nullptr;
    static const bool timeline = getenv("DG_TIMELINE") != nullptr;
    if (timeline && !dbg) cudaMalloc(&dbg, (148 * 10 + 64) * sizeof(long long));
    cp.dbg = timeline ? dbg : nullptr;
    cp.tl = c->tl; cp.tl_id = cp.mode ? DG_TL_DGRAD : (cp.ksize == 5 || cp.cin < cp.cout ? DG_TL_CONV0 : DG_TL_CONV);
    const size_t img_bytes = ((size_t)(cp.cin / 8) * cp.R_in * 16 + 1023) & ~(size_t)1023;
    // CTA pairs (cta_group::2) when the layer is one N block and one K panel: each CTA keeps half of every weight stage
    // (opt-in, DG_PAIR=1: measured slower at batch 128 -- the two SMs of a TPC then drain their accumulators in lock-step and
    // contend for the TPC's path to L2, which costs more than the MMA phase gains)
    static const bool want_pair = getenv("DG_PAIR") != nullptr;
    const bool pair = want_pair && cp.cout == cp.nb && cp.cin == cp.kp && cp.nb % 32 == 0 && cp.n >= 2;
    const size_t stage = (size_t)cp.kp * (pair ? cp.nb / 2 : cp.nb) * 2;
    const size_t budget = 227 * 1024 - 15 * 1024;   // static smem (barriers, the epilogue's row table) lives in the remaining 8 KB
    DG_REQUIRE(img_bytes + 2 * stage <= budget, "bf16 path: image (%zu B) + weight ring do not fit shared memory", img_bytes);
    int ns = (int)((budget - img_bytes) / stage);
    if (ns > CONV_MAX_STAGES) ns = CONV_MAX_STAGES;
    cp.nstages = ns;
    const size_t smem = img_bytes + ns * stage;
    int grid = cp.n < m->bf->num_sms ? cp.n : m->bf->num_sms;
    if (pair) grid = 2 * std::min((cp.n + 1) / 2, m->bf->num_sms / 2);
    { dg_prof_scope _ps(c, cp.mode ? DG_K_DGRAD : (cp.ksize == 5 || cp.cin < cp.cout ? DG_K_CONV_FIRST : DG_K_CONV));
    if (pair && cp.mode) DG_CUDA(dg_launch_pdl(conv_umma_kernel<true, 1>, dim3(grid), dim3(384), smem, c->stream, cp, 2));
    else if (pair) DG_CUDA(dg_launch_pdl(conv_umma_kernel<true, 0>, dim3(grid), dim3(384), smem, c->stream, cp, 2));
    else if (cp.mode) DG_CUDA(dg_launch_pdl(conv_umma_kernel<false, 1>, dim3(grid), dim3(384), smem, c->stream, cp));
    else DG_CUDA(dg_launch_pdl(conv_umma_kernel<false, 0>, dim3(grid), dim3(384), smem, c->stream, cp));
    }
    c->launches++;
    if (timeline) {
        std::vector<long long> h(148 * 10 + 64);
        cudaStreamSynchronize(c->stream);
        cudaMemcpy(h.data(), dbg, h.size() * 8, cudaMemcpyDeviceToHost);
        double d[7] = {0};
        for (int b = 0; b < grid; ++b) for (int i = 1; i < 7; ++i) d[i] += (double)(h[b * 8 + i] - h[b * 8]) / grid;
        double w = 0;
        for (int b = 0; b < grid; ++b) w += (double)h[b * 8 + 7] / grid;
        long long t0min = h[148 * 8], t0max = h[148 * 8], t1max = h[148 * 8 + 1], t1min = h[148 * 8 + 1], dmax = 0, dmin = 1ll << 60;
        for (int b = 0; b < grid; ++b) {
            const long long a0 = h[148 * 8 + 2 * b], a1 = h[148 * 8 + 2 * b + 1];
            t0min = std::min(t0min, a0); t0max = std::max(t0max, a0); t1max = std::max(t1max, a1); t1min = std::min(t1min, a1);
            dmax = std::max(dmax, a1 - a0); dmin = std::min(dmin, a1 - a0);
        }
        if (cp.nchain > 1) {
            fprintf(stderr, "   chain (CTA 0, cycles): layer: mma-start -> all-issued | epilogue sees acc_full -> drained+handed over\n");
            const long long t0 = h[0];
            for (int l = 0; l < cp.nchain && l < 16; ++l)
                fprintf(stderr, "     l=%d  mma %lld -> %lld (%lld) | epi %lld -> %lld (%lld)\n", l, h[148 * 10 + l * 4] - t0, h[148 * 10 + l * 4 + 1] - t0,
                        h[148 * 10 + l * 4 + 1] - h[148 * 10 + l * 4], h[148 * 10 + l * 4 + 2] - t0, h[148 * 10 + l * 4 + 3] - t0,
                        h[148 * 10 + l * 4 + 3] - h[148 * 10 + l * 4 + 2]);
        }
        fprintf(stderr, "   globaltimer ns: CTA start spread %lld, first start -> last end %lld, first end -> last end %lld, CTA duration min %lld max %lld\n",
                t0max - t0min, t1max - t0min, t1max - t1min, dmin, dmax);
        fprintf(stderr, "conv timeline (cycles since CTA start, mean over %d CTAs) mode=%d k=%d: setup %.0f | img panel0 %.0f panel1 %.0f | mma issued %.0f | acc ready %.0f | epilogue done %.0f | weight-wait %.0f\n",
                grid, cp.mode, cp.ksize, d[1], d[2], d[3], d[4], d[5], d[6], w);
    }
    DG_CUDA(cudaGetLastError());
    return 0;
}

struct WgradPlan { int nb, n_mblk, n_nblk, combos, splits, nstages, rp; size_t smem, scratch_bytes; };

static int wgrad_plan(int num_sms, int cin_pad, int cout, int ksize, const Geom &g, WgradPlan &w) {
    w.nb = cin_pad;
    while (ksize * w.nb > 512) w.nb /= 2;
    DG_REQUIRE(w.nb % 16 == 0 && cin_pad % w.nb == 0 && w.nb <= 256, "bf16 path: unsupported weight-gradient N block for cin %d", cin_pad);
    w.n_nblk = cin_pad / w.nb;
    w.n_mblk = (cout + 127) / 128;
    w.combos = ksize * w.n_mblk * w.n_nblk;
    DG_REQUIRE(w.combos <= num_sms, "bf16 path: too many weight-gradient tiles");
    // a multiple of the panels per position: every CTA then sees one panel index only (its bias-gradient accumulator is one panel)
    // K rows per ring stage.  96 (four 50 KB stages instead of three 65 KB ones, DG_WG_RP96=1) was measured: no gain, the producer
    // then issues a third more copies per byte and the units no longer divide evenly among the CTAs
    static const bool rp96 = getenv("DG_WG_RP96") != nullptr;
    w.rp = (g.ntiles * 128) % 96 == 0 && rp96 ? 96 : 128;
    const int WG_RP = w.rp;
    const int panels = g.ntiles * 128 / WG_RP;
    w.splits = num_sms / w.combos;
    if (w.splits > panels) w.splits -= w.splits % panels;
    const int rb_pad = (WG_RP + 2 * g.pad + 7) & ~7;
    const size_t stage = 16 * WG_RP * 16 + (size_t)(w.nb / 8) * rb_pad * 16;
    int ns = (int)((227 * 1024 - 1024) / stage);
    if (ns > CONV_MAX_STAGES) ns = CONV_MAX_STAGES;
    w.nstages = ns;
    w.smem = ns * stage;
    w.scratch_bytes = (size_t)w.combos * w.splits * ksize * 128 * w.nb * 4;
    return 0;
}

int dg_bf16_create(dg_model *m) {
    DG_REQUIRE(m->C % 32 == 0 && m->C >= 32 && m->C <= 256, "DG_PREC_BF16 needs channelSize %% 32 == 0 and <= 256 (got %d)", m->C);
    DG_REQUIRE(384 % m->C == 0 || m->C > 384, "DG_PREC_BF16 head kernel needs channelSize dividing 384 (32, 64, 96, 128, 192); got %d", m->C);
    Bf16Path *b = new Bf16Path();
    m->bf = b;
    b->g_first = geom_for(m->k_first);
    b->g3 = geom_for(3);
    cudaDeviceProp prop;
    DG_CUDA(cudaGetDeviceProperties(&prop, m->ctx->dev));
    b->num_sms = prop.multiProcessorCount;
    const size_t B = m->max_batch;
    const int L = m->L, C = m->C;
    cudaStream_t st = m->ctx->stream;
    auto alloc0 = [&](void **p, size_t bytes) -> int {
        DG_CUDA(cudaMalloc(p, bytes));
        DG_CUDA(cudaMemsetAsync(*p, 0, bytes, st));
        return 0;
    };
    if (int rc = alloc0((void **)&b->x0, B * img_elems(b->cin0_pad, b->g_first.rows) * 2)) return rc;
    b->h.assign(L, nullptr);
    for (int l = 1; l < L; ++l) if (int rc = alloc0((void **)&b->h[l], B * img_elems(C, DG_R3) * 2)) return rc;
    if (int rc = alloc0((void **)&b->head_part, B * (size_t)(9 * C + DG_POINTS) * 4)) return rc;
    b->dzl.assign(L - 1, nullptr);
    for (int l = 0; l < L - 1; ++l)
        if (int rc = alloc0((void **)&b->dzl[l], B * img_elems(C, l == 0 ? b->g_first.rows : DG_R3) * 2)) return rc;
    b->ops.resize(L - 1);
    size_t scratch = 0;
    for (int l = 0; l < L - 1; ++l) {
        const LayerInfo &li = m->layers[l];
        LayerOps &o = b->ops[l];
        o.cin_pad = l == 0 ? b->cin0_pad : li.cin;
        const size_t nw = (size_t)li.k * li.k * o.cin_pad * li.cout;
        if (int rc = alloc0((void **)&o.wf, nw * 2)) return rc;
        if (l > 0) if (int rc = alloc0((void **)&o.wd, nw * 2)) return rc;
        if (int rc = alloc0((void **)&o.ab, (size_t)li.cout * DG_R3 * 2)) return rc;
        WgradPlan w;
        if (int rc = wgrad_plan(b->num_sms, o.cin_pad, li.cout, li.k, l == 0 ? b->g_first : b->g3, w)) return rc;
        if (w.scratch_bytes > scratch) scratch = w.scratch_bytes;
    }
    {
        std::vector<PackLayer> tab(L - 1);
        for (int l = 0; l < L - 1; ++l) {
            const LayerInfo &li = m->layers[l];
            tab[l].W = m->params + li.offW; tab[l].b = m->params + li.offB; tab[l].a = m->params + li.offA;
            tab[l].wf = b->ops[l].wf; tab[l].wd = b->ops[l].wd; tab[l].ab = b->ops[l].ab;
            tab[l].cin_true = li.cin; tab[l].cin = b->ops[l].cin_pad; tab[l].cout = li.cout; tab[l].ksize = li.k;
        }
        std::vector<FinalizeLayer> ft(L - 1);
        for (int l = 0; l < L - 1; ++l) {
            const LayerInfo &li = m->layers[l];
            ft[l].gacc = nullptr; ft[l].gW = m->grads + li.offW; ft[l].cin_true = li.cin; ft[l].cin = b->ops[l].cin_pad; ft[l].cout = li.cout; ft[l].ksize = li.k;
            const Geom &gl = l == 0 ? b->g_first : b->g3;
            ft[l].gab = nullptr; ft[l].ga = m->grads + li.offA; ft[l].gb = m->grads + li.offB; ft[l].pitch = gl.pitch; ft[l].pad = gl.pad; ft[l].first = gl.first;
        }
        b->fin_host = ft;
        std::vector<SgdPackLayer> sg(L);
        for (int l = 0; l < L; ++l) {
            const LayerInfo &li = m->layers[l];
            sg[l].W = m->params + li.offW; sg[l].b = m->params + li.offB; sg[l].a = m->params + li.offA;
            sg[l].gW = m->grads + li.offW; sg[l].gb = m->grads + li.offB; sg[l].ga = m->grads + li.offA;
            sg[l].wf = l < L - 1 ? b->ops[l].wf : nullptr; sg[l].wd = l < L - 1 ? b->ops[l].wd : nullptr; sg[l].ab = l < L - 1 ? b->ops[l].ab : nullptr;
            sg[l].cin_true = li.cin; sg[l].cin = l < L - 1 ? b->ops[l].cin_pad : li.cin; sg[l].cout = li.cout; sg[l].ksize = li.k;
            sg[l].gacc = nullptr; sg[l].gab = nullptr; sg[l].g_pitch = sg[l].g_pad = sg[l].g_first = 0;
            b->sgd_smem_row = std::max(b->sgd_smem_row, li.cin * li.k * li.k + 1);
            b->sgd_max_co8 = std::max(b->sgd_max_co8, (li.cout + SGD_CO_B - 1) / SGD_CO_B);
        }
        b->sgd_host = sg;
        DG_CUDA(cudaMalloc((void **)&b->sgd_table, sg.size() * sizeof(SgdPackLayer)));
        DG_CUDA(cudaMalloc((void **)&b->sgd_table_acc, sg.size() * sizeof(SgdPackLayer)));
        DG_CUDA(cudaMemcpyAsync(b->sgd_table, sg.data(), sg.size() * sizeof(SgdPackLayer), cudaMemcpyHostToDevice, st));
        DG_CUDA(cudaFuncSetAttribute(sgd_pack_kernel<SGD_CO_B, SGD_NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SGD_CO_B * b->sgd_smem_row * 4));
        DG_CUDA(cudaMalloc((void **)&b->fin_table, ft.size() * sizeof(FinalizeLayer)));
        DG_CUDA(cudaMalloc((void **)&b->pack_table, tab.size() * sizeof(PackLayer)));
        DG_CUDA(cudaMemcpyAsync(b->pack_table, tab.data(), tab.size() * sizeof(PackLayer), cudaMemcpyHostToDevice, st));
        DG_CUDA(cudaStreamSynchronize(st));
    }
    b->deterministic = getenv("DG_DETERMINISTIC") != nullptr;
    // Accumulation buffers, one allocation cleared by one memset: [head layer gradient | layer 0: weights, bias | layer 1: .. ].
    // Each layer's block is contiguous, so a data-parallel bucket is a range of layers; the head block (gW, gb, ga in Torch7
    // order) lets a fused step exchange everything in the buckets' own order.
    b->gacc_off.assign(L - 1, 0);
    b->gab_off.assign(L - 1, 0);
    b->fused_bias = !b->deterministic;
    {
        const LayerInfo &lh = m->layers[L - 1];
        b->head_n = lh.nW + lh.cout + (size_t)lh.cout * DG_POINTS;
        b->gacc_floats = (b->head_n + 3) & ~(size_t)3;
    }
    for (int l = 0; l < L - 1; ++l) {
        const Geom &gl = l == 0 ? b->g_first : b->g3;
        WgradPlan wpl;
        if (int rc = wgrad_plan(b->num_sms, b->ops[l].cin_pad, m->layers[l].cout, m->layers[l].k, gl, wpl)) return rc;
        if (m->layers[l].k * wpl.nb + wpl.rp > 512 || wpl.splits % (gl.ntiles * 128 / wpl.rp) != 0) b->fused_bias = false;
        b->gacc_off[l] = b->gacc_floats; b->gacc_floats += (size_t)m->layers[l].k * m->layers[l].k * b->ops[l].cin_pad * m->layers[l].cout;
        b->gab_off[l] = b->gacc_floats; b->gacc_floats += (size_t)gl.ntiles * 128 * m->layers[l].cout;
    }
    DG_CUDA(cudaMalloc((void **)&b->gacc, b->gacc_floats * 4));
    for (int l = 0; l < L - 1; ++l) { b->fin_host[l].gacc = b->gacc + b->gacc_off[l]; b->fin_host[l].gab = b->fused_bias ? b->gacc + b->gab_off[l] : nullptr; }
    b->defer_grads = b->fused_bias && getenv("DG_NO_DEFER") == nullptr;
    if (b->fused_bias) {
        std::vector<SgdPackLayer> sa = b->sgd_host;
        for (int l = 0; l < L - 1; ++l) {
            const Geom &gl = l == 0 ? b->g_first : b->g3;
            sa[l].gacc = b->gacc + b->gacc_off[l]; sa[l].gab = b->gacc + b->gab_off[l];
            sa[l].g_pitch = gl.pitch; sa[l].g_pad = gl.pad; sa[l].g_first = gl.first;
        }
        {
            const LayerInfo &lh = m->layers[L - 1];
            sa[L - 1].gW = b->gacc; sa[L - 1].gb = b->gacc + lh.nW; sa[L - 1].ga = b->gacc + lh.nW + lh.cout;
        }
        DG_CUDA(cudaMemcpyAsync(b->sgd_table_acc, sa.data(), sa.size() * sizeof(SgdPackLayer), cudaMemcpyHostToDevice, st));
    }
    DG_CUDA(cudaMemcpyAsync(b->fin_table, b->fin_host.data(), b->fin_host.size() * sizeof(FinalizeLayer), cudaMemcpyHostToDevice, st));
    DG_CUDA(cudaStreamSynchronize(st));
    DG_CUDA(cudaMalloc((void **)&b->scratch, 2 * scratch));
    b->scratch_bytes = scratch;
    DG_CUDA(cudaStreamCreateWithFlags(&b->side, cudaStreamNonBlocking));
    cudaEvent_t *evs[] = {&b->ev_head, &b->ev_fork, &b->ev_memset, &b->ev_dz, &b->ev_wgrad, &b->ev_bias[0], &b->ev_bias[1], &b->ev_reduce[0], &b->ev_reduce[1], &b->ev_join};
    for (cudaEvent_t *e : evs) DG_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    DG_CUDA(cudaFuncSetAttribute(conv_umma_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 15 * 1024));
    DG_CUDA(cudaFuncSetAttribute(conv_umma_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 15 * 1024));
    DG_CUDA(cudaFuncSetAttribute(conv_umma_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 15 * 1024));
    DG_CUDA(cudaFuncSetAttribute(conv_umma_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 15 * 1024));
    DG_CUDA(cudaFuncSetAttribute(wgrad_umma_kernel<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
    DG_CUDA(cudaFuncSetAttribute(wgrad_umma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
    DG_CUDA(cudaFuncSetAttribute(head_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 4096));
    DG_CUDA(cudaFuncSetAttribute(head_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 8192));
    return 0;
}

void dg_bf16_destroy(dg_model *m) {
    Bf16Path *b = m->bf;
    if (!b) return;
    cudaFree(b->x0);
    for (bf16 *p : b->h) if (p) cudaFree(p);
    for (bf16 *p : b->dzl) if (p) cudaFree(p);
    for (LayerOps &o : b->ops) { cudaFree(o.wf); if (o.wd) cudaFree(o.wd); cudaFree(o.ab); }
    cudaFree(b->scratch);
    cudaFree(b->gacc);
    cudaFree(b->pack_table);
    cudaFree(b->fin_table);
    cudaFree(b->sgd_table);
    cudaFree(b->sgd_table_acc);
    dg_dp_p2p_destroy(b->p2p);
    cudaFree(b->head_part);
    if (b->side) cudaStreamDestroy(b->side);
    cudaEvent_t evs[] = {b->ev_head, b->ev_fork, b->ev_memset, b->ev_dz, b->ev_wgrad, b->ev_bias[0], b->ev_bias[1], b->ev_reduce[0], b->ev_reduce[1], b->ev_join};
    for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    delete b;
    m->bf = nullptr;
}

int dg_bf16_params_changed(dg_model *m) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    {
        dg_prof_scope _ps(c, DG_K_PACK);
        DG_CUDA(dg_launch_pdl(pack_layer_kernel, dim3(b->num_sms, m->L - 1), dim3(256), 0, c->stream, (const PackLayer *)b->pack_table, (int)DG_R3, b->g3.pitch, b->g3.pad));
    }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// SGD update + operand repack in one launch; rate_dev != NULL: the rate is read from device memory (captured graph)
int dg_bf16_sgd_step(dg_model *m, float rate, const float *rate_dev) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    {
        dg_prof_scope _ps(c, DG_K_SGD);
        SgdPeers peers{};
        if (b->grads_pending && b->p2p && b->p2p_pull) {
            peers.base = b->gacc; peers.n = b->gacc_floats;
            dg_dp_p2p_peers(b->p2p, peers.buf, &peers.world, &peers.per4);
        }
        DG_CUDA(dg_launch_pdl(sgd_pack_kernel<SGD_CO_B, SGD_NT>, dim3(b->sgd_max_co8, m->L), dim3(SGD_NT), (size_t)SGD_CO_B * b->sgd_smem_row * 4, c->stream,
                              (const SgdPackLayer *)(b->grads_pending ? b->sgd_table_acc : b->sgd_table), rate, rate_dev, (int)DG_R3, b->g3.pitch, b->g3.pad, b->sgd_smem_row, c->tl, peers));
    }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

// pull mode: fetch every slice of the summed gradient from its owner into the local buffer (before the gradient vector is built from it)
__global__ void __launch_bounds__(256) p2p_pull_all_kernel(const __grid_constant__ SgdPeers P, float *local) {
    const size_t n4 = P.n >> 2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float *src = sgd_grad_ptr(P, local + 4 * i);
        if (src != local + 4 * i) reinterpret_cast<float4 *>(local)[i] = __ldcg(reinterpret_cast<const float4 *>(src));
    }
}

// data parallel, first fused step (a collective: every rank gets here): map the peers' accumulation buffers.  Falls back to NCCL
// (with the reason on stderr, once) when CUDA IPC or peer access is unavailable; DG_DP_NCCL=1 forces NCCL.
int dg_bf16_dp_prepare(dg_model *m) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    if (c->world <= 1 || b->p2p_tried || !b->defer_grads) return 0;
    b->p2p_tried = true;
    if (getenv("DG_DP_NCCL")) return 0;
    DG_CUDA(cudaStreamSynchronize(c->stream));
    if (dg_dp_p2p_setup(c, b->gacc, b->gacc_floats, &b->p2p) != 0) {
        fprintf(stderr, "deepgo_b200: %s\n", dg_last_error());
        b->p2p = nullptr;
    }
    // pull mode (DG_DP_PULL=1: reduce-scatter only, the SGD kernel fetches each slice from its owner) is implemented and tested but
    // slower: remote reads inside the latency-bound SGD blocks cost 27 us, the write phase it saves 4 us (2 GPUs)
    b->p2p_pull = b->p2p != nullptr && getenv("DG_DP_PULL") != nullptr;
    return 0;
}

// Weight gradients with the bf16 tensor-core kernel for callers outside this path (conv_tf32.cu feeds it hi/lo splits of fp32
// images): `nl` (<= 16) products of one shape in ONE persistent launch -- product j adds dz[j]^T x[j] per tap into gacc[j]
// ([tap][cin_pad/4][cout][4]; several products may name the same block).  dz [n][cout/8][R][8], x [n][cin_pad/8][R][8] in the
// geometry of a k x k layer.  The layer groups are chosen by a small cost model: phases x (units per CTA + a drain).
int dg_bf16_wgrad_multi(dg_ctx *c, int num_sms, int nl, const void *const *dz, const void *const *x, float *const *gacc, int n, int cin_pad, int cout, int ksize) {
    DG_REQUIRE(nl >= 1 && nl <= WG_MAX_LAYERS, "weight gradient: %d products in one launch (1..%d)", nl, WG_MAX_LAYERS);
    const Geom g = geom_for(ksize);
    WgradPlan w;
    if (int rc = wgrad_plan(num_sms, cin_pad, cout, ksize, g, w)) return rc;
    DG_REQUIRE(w.rp == 128, "weight gradient: 128-row stages expected");
    static bool attr_set[64] = {false};
    if (c->dev < 64 && !attr_set[c->dev]) {
        DG_CUDA(cudaFuncSetAttribute(wgrad_umma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 2048));
        attr_set[c->dev] = true;
    }
    const int panels = g.ntiles * 128 / w.rp, units = n * panels, per_row = num_sms / w.combos;
    int best_g = 1, best_st = w.splits;
    double best_t = 1e30;
    for (int G = 1; G <= nl; ++G) {
        if (nl % G) continue;
        const int st = G == 1 ? w.splits : per_row / (G * panels) * (G * panels);
        if (st < G * panels && G > 1) continue;
        if (units < st / G) continue;
        const double t = (double)(nl / G) * ((units * G + st - 1) / st + 5.0);      // microseconds, roughly: 1 per unit, 5 per drain + refill
        if (t < best_t) { best_t = t; best_g = G; best_st = st; }
    }
    WgradParams wp{};
    for (int j = 0; j < nl; ++j) { wp.dz[j] = (const bf16 *)dz[j]; wp.x[j] = (const bf16 *)x[j]; wp.gacc[j] = gacc[j]; wp.gab[j] = nullptr; }
    wp.scratch = nullptr; wp.nl = nl;
    wp.n = n; wp.cin = cin_pad; wp.cout = cout;
    wp.R = g.rows; wp.pitch = g.pitch; wp.pad = g.pad; wp.ksize = ksize; wp.first = g.first; wp.ntiles = g.ntiles;
    wp.ngroups = best_g;
    wp.nb = w.nb; wp.n_mblk = w.n_mblk; wp.n_nblk = w.n_nblk; wp.splits = best_st / best_g; wp.nstages = w.nstages;
    { dg_prof_scope _ps(c, DG_K_WGRAD);
    DG_CUDA(dg_launch_pdl(wgrad_umma_kernel<128>, dim3(w.combos * best_st), dim3(384), w.smem, c->stream, wp)); }
    c->launches++;
    return 0;
}

// one process, several devices: the replicas of a multi-device model map each other's accumulation buffers directly
int dg_bf16_dp_prepare_local(dg_model *const *kids, int n) {
    std::vector<dg_ctx *> ctxs(n);
    std::vector<float *> bufs(n);
    std::vector<dg_p2p *> out(n, nullptr);
    for (int k = 0; k < n; ++k) {
        Bf16Path *b = kids[k]->bf;
        DG_REQUIRE(b && b->defer_grads, "multi-device step: the bf16 path with gradient hand-off is required");
        if (b->p2p) return 0;                                  // already mapped
        ctxs[k] = kids[k]->ctx; bufs[k] = b->gacc;
        DG_CUDA(cudaSetDevice(ctxs[k]->dev));
        DG_CUDA(cudaStreamSynchronize(ctxs[k]->stream));
    }
    if (int rc = dg_dp_p2p_setup_local(ctxs.data(), bufs.data(), kids[0]->bf->gacc_floats, n, out.data())) {
        for (dg_p2p *p : out) dg_dp_p2p_destroy(p);
        return rc;
    }
    for (int k = 0; k < n; ++k) { kids[k]->bf->p2p = out[k]; kids[k]->bf->p2p_tried = true; kids[k]->bf->p2p_pull = false; }
    return 0;
}

// materialise the conv layers' gradient of the last fused step in the Torch7-order vector (no-op when it already is there)
int dg_bf16_sync_grads(dg_model *m) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    if (!b->grads_pending) return 0;
    const LayerInfo &l0 = m->layers[0], &lh = m->layers[m->L - 1];
    if (b->p2p && b->p2p_pull && c->world > 1) {
        // the local buffer holds the summed values of this rank's slice only (the peers keep theirs until every rank has entered its
        // next step, and this call is ordered before ours)
        SgdPeers peers{};
        peers.base = b->gacc; peers.n = b->gacc_floats;
        dg_dp_p2p_peers(b->p2p, peers.buf, &peers.world, &peers.per4);
        p2p_pull_all_kernel<<<2 * b->num_sms, 256, 0, c->stream>>>(peers, b->gacc);
        c->launches++;
    }
    DG_CUDA(cudaMemsetAsync(m->grads + l0.offW, 0, (lh.offW - l0.offW) * 4, c->stream));
    wgrad_finalize_kernel<<<dim3(96, m->L - 1), 256, FIN_SMEM_FLOATS * 4, c->stream>>>(b->fin_table, nullptr);
    c->launches++;
    DG_CUDA(cudaMemcpyAsync(m->grads + lh.offW, b->gacc, b->head_n * 4, cudaMemcpyDeviceToDevice, c->stream));
    DG_CUDA(cudaGetLastError());
    b->grads_pending = false;
    return 0;
}
void dg_bf16_drop_pending_grads(dg_model *m) { m->bf->grads_pending = false; }
// a replayed step graph runs no host code of dg_bf16_backward: restore what the captured step leaves behind
void dg_bf16_after_fused_step(dg_model *m) { m->bf->grads_pending = m->bf->defer_grads; m->bf->x0_from_raw = false; }
bool dg_bf16_defers_grads(const dg_model *m) { return m->bf->defer_grads; }

int dg_bf16_stage_raw(dg_model *m) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    // DG_FUSED_ENCODE=1: the records stay as they are and the first conv kernel of the next forward pass encodes them.  Implemented and
    // tested, but not faster: the separate 7 us encode kernel runs under the launch latency of the step graph, while the in-kernel encode
    // adds 2.3 us to the first conv layer
    static const bool fused = getenv("DG_FUSED_ENCODE") != nullptr;
    if (fused && b->ops[0].cin_pad == 48) { b->x0_from_raw = true; return 0; }
    b->x0_from_raw = false;
    { dg_prof_scope _ps(c, DG_K_ENCODE);
    DG_CUDA(dg_launch_pdl(encode_blocked_kernel, dim3(m->n), dim3(512), 0, c->stream, (const uint8_t *)m->d_raw, (const uint8_t *)m->d_player, (const uint8_t *)m->d_rank, (const uint8_t *)m->d_sym, b->x0, b->g_first.rows, b->g_first.pitch, b->g_first.pad,
                          (const uint8_t *)m->d_x, (const uint8_t *)m->d_y, m->have_targets ? m->d_targets : (int32_t *)nullptr, (size_t)m->max_batch * DG_RAW_PLANES * DG_POINTS));
    }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}
// 1: the next forward pass starts from raw records (fused encode), 0: from the blocked image in HBM -- two different step graphs
int dg_bf16_input_mode(const dg_model *m) { return m->bf->x0_from_raw ? 1 : 0; }
bool dg_bf16_input_mode_fused() { static const bool f = getenv("DG_FUSED_ENCODE") != nullptr; return f; }

int dg_bf16_stage_planes(dg_model *m) {
    Bf16Path *b = m->bf;
    b->x0_from_raw = false;
    dg_ctx *c = m->ctx;
    { dg_prof_scope _ps(c, DG_K_ENCODE);
    planes_to_blocked_kernel<<<m->n, 384, 0, c->stream>>>(m->act[0], b->x0, m->in_planes, b->g_first.rows, b->g_first.pitch, b->g_first.pad);
    }
    c->launches++;
    DG_CUDA(cudaGetLastError());
    return 0;
}

static int launch_head(dg_model *m, bool backward, float scale) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    const LayerInfo &li = m->layers[m->L - 1];
    HeadParams hp;
    hp.in = b->h[m->L - 1];
    hp.W = m->params + li.offW; hp.b = m->params + li.offB; hp.a = m->params + li.offA;
    hp.targets = m->have_targets ? m->d_targets : nullptr;
    hp.logp = m->d_logp; hp.loss_n = m->d_loss_n; hp.argmax = m->d_argmax;
    // dZ_{L-1} lives in the geometry of layer L-1's own kernel: first-layer geometry when L-1 == 1
    const bool into_first = (m->L - 1) == 1;
    const Geom &go = into_first ? b->g_first : b->g3;
    hp.dz_out = backward ? b->dzl[m->L - 2] : nullptr;
    hp.part = b->head_part;
    static long long *hdbg = nullptr;
    static const bool htl = getenv("DG_TIMELINE") != nullptr;
    if (htl && !hdbg) cudaMalloc(&hdbg, 4096 * 8 * sizeof(long long));
    hp.dbg = htl && m->n <= 4096 ? hdbg : nullptr;
    hp.tl = c->tl;
    hp.n = m->n; hp.C = m->C; hp.R = DG_R3;
    hp.R_out = go.rows; hp.pitch_out = go.pitch; hp.pad_out = go.pad;
    hp.scale = scale;
    const size_t smem = (((size_t)(m->C / 8) * DG_R3 * 16 + 1023) & ~(size_t)1023) + (size_t)18 * m->C * 4;
    DG_REQUIRE(smem <= 227 * 1024 - 4096, "bf16 path: head image does not fit shared memory (C=%d)", m->C);
    // C <= 128: the tensor-core head (three small GEMMs); wider nets keep the SIMT kernel (3 x C accumulator columns exceed TMEM)
    static const bool simt_head = getenv("DG_SIMT_HEAD") != nullptr;
    const bool umma_head = !simt_head && m->C <= 128 && m->C % 16 == 0;
    const size_t smem_u = (size_t)16 * DG_R3 * 16 + 2048 + 16 * 32 * 16 + 4 * 128 * 16 + 4 * DG_R3 * 16 + (size_t)DG_R3 * 9 * 4;
    { dg_prof_scope _ps(c, DG_K_HEAD);
    if (umma_head) DG_CUDA(dg_launch_pdl(head_umma_kernel, dim3(m->n), dim3(384), smem_u, c->stream, hp));
    else DG_CUDA(dg_launch_pdl(head_bf16_kernel, dim3(m->n), dim3(384), smem, c->stream, hp));
    }
    c->launches++;
    if (hp.dbg && backward) {
        std::vector<long long> h((size_t)m->n * 8);
        cudaStreamSynchronize(c->stream);
        cudaMemcpy(h.data(), hp.dbg, h.size() * 8, cudaMemcpyDeviceToHost);
        double d[6] = {0};
        for (int i = 0; i < m->n; ++i) for (int k = 1; k < 6; ++k) d[k] += (double)(h[i * 8 + k] - h[i * 8]) / m->n;
        fprintf(stderr, "head timeline (cycles, mean over %d CTAs): image %.0f | logits %.0f | softmax %.0f | weight grad %.0f | input grad %.0f\n", m->n, d[1], d[2], d[3], d[4], d[5]);
    }
    DG_CUDA(cudaGetLastError());
    return 0;
}

static bool chain_ok(const dg_model *m) {
    static const bool off = getenv("DG_NO_CHAIN") != nullptr;
    return !off && m->C <= 128 && m->L >= 4;      // one N block, one K panel, at least two C->C layers
}

int dg_bf16_forward(dg_model *m) {
    Bf16Path *b = m->bf;
    auto fill = [&](ConvParams &cp, int l) {
        const LayerInfo &li = m->layers[l];
        const Geom &gi = l == 0 ? b->g_first : b->g3;
        cp.in = l == 0 ? b->x0 : b->h[l];
        cp.n = m->n;
        cp.R_in = gi.rows; cp.pitch_in = gi.pitch; cp.pad_in = gi.pad; cp.ksize = li.k; cp.first = gi.first; cp.ntiles = gi.ntiles;
        cp.R_out = DG_R3; cp.pitch_out = b->g3.pitch; cp.pad_out = b->g3.pad;
        cp.R_mask = cp.pitch_mask = cp.pad_mask = 0;
        cp.mode = 0;
    };
    int l = 0;
    while (l < m->L - 1) {
        ConvParams cp{};
        fill(cp, l);
        int nchain = 1;
        if (l >= 1 && chain_ok(m)) nchain = std::min(m->L - 1 - l, CONV_MAX_CHAIN);
        for (int j = 0; j < nchain; ++j) {
            cp.w[j] = b->ops[l + j].wf; cp.bias[j] = b->ops[l + j].ab; cp.mask[j] = nullptr; cp.out[j] = b->h[l + j + 1];
        }
        cp.nchain = nchain;
        if (l == 0 && b->x0_from_raw) {
            cp.enc_raw = m->d_raw; cp.enc_player = m->d_player; cp.enc_rank = m->d_rank; cp.enc_sym = m->d_sym; cp.enc_out = b->x0;
        }
        if (int rc = conv_cfg(m, cp, b->ops[l].cin_pad, m->layers[l].cout)) return rc;
        if (int rc = launch_conv(m, cp)) return rc;
        l += nchain;
    }
    b->x0_from_raw = false;               // the image is in HBM now: a second forward pass over the same batch loads it
    b->head_bwd_done = false;
    if (m->fused_step) return 0;          // dg_train_step: the head runs once, in dg_bf16_backward
    return launch_head(m, false, 0.f);
}

int dg_bf16_backward(dg_model *m, float scale) {
    Bf16Path *b = m->bf;
    dg_ctx *c = m->ctx;
    // Inside dg_train_step the SGD kernel reads the accumulation buffers itself: the Torch7-order gradient vector is not
    // materialised (dg_bf16_sync_grads does it on demand, e.g. for dg_model_get_grads)
    const bool defer = m->fused_step && b->defer_grads;
    // ---- the weight-gradient launches of this pass: conv layers L-2 .. 1 have one shape and (default path) go out as ONE launch;
    // layer 0 on its own
    struct WgLaunch { int l, nl, ngroups, grid; WgradPlan w; };
    std::vector<WgLaunch> wl;
    // (Measured and dropped: the top conv layer's weight gradient on the 20 SMs that have no position during the input-gradient chain,
    // on the side stream -- the persistent launch then has 9 layers in 3 phases and is 5 us shorter, but the step as a whole came
    // out 10 us LONGER: the side kernel delays the chain and the join in front of the SGD kernel.)
    for (int l = m->L - 2; l >= 0;) {
        const LayerInfo &li = m->layers[l];
        const Geom &g = l == 0 ? b->g_first : b->g3;
        WgLaunch q; q.l = l;
        if (int rc = wgrad_plan(b->num_sms, b->ops[l].cin_pad, li.cout, li.k, g, q.w)) return rc;
        q.nl = 1;
        if (!b->deterministic && l >= 1) q.nl = std::min(l, WG_MAX_LAYERS);       // layers l, l-1, .., l-nl+1 (all >= 1)
        const int panels_l = g.ntiles * 128 / q.w.rp;
        // G layer groups: group g walks layers g, g + G, ..; the splits of a kernel row are dealt evenly to the groups, every CTA keeps one
        // panel index (splits per group a multiple of the panels).  Default: TWO PHASES (G = layers / 2): a CTA accumulates a layer over
        // many units and drains only twice, and the drain's red.global traffic (the partial tile of every CTA) drops with the number
        // of CTAs per layer: 10 layers in 5 groups of 27 CTAs measured 104 us against 113 us for 2 groups of 72 CTAs, 13 SMs idle
        // notwithstanding.  DG_WG_GROUPS forces a count.
        static const int env_groups = [] { const char *e = getenv("DG_WG_GROUPS"); return e ? atoi(e) : 0; }();
        q.ngroups = 1;
        if (!b->deterministic && q.nl >= 2) {
            // G among the divisors of the layer count by a small cost model: phases x (units per CTA, ~1 us each, + ~5 us per drain and refill)
            const int per_row = b->num_sms / q.w.combos, units = m->n * panels_l;         // CTAs a kernel row can have; split-K units per layer and row
            double best_t = 1e30;
            for (int G = 1; G <= q.nl; ++G) {
                if (q.nl % G || (env_groups > 0 && G != env_groups)) continue;
                const int st = G == 1 ? q.w.splits : per_row / (G * panels_l) * (G * panels_l);   // .. dealt evenly to the groups
                if ((G > 1 && st < G * panels_l) || units < st / G) continue;
                const double t = (double)(q.nl / G) * ((units * G + st - 1) / st + 5.0);
                if (t < best_t) { best_t = t; q.ngroups = G; }
            }
            if (q.ngroups > 1) q.w.splits = per_row / (q.ngroups * panels_l) * (q.ngroups * panels_l);
        }
        q.grid = q.w.combos * q.w.splits;
        wl.push_back(q);
        l -= q.nl;
    }
    // the accumulation buffers are cleared on the side stream while the head and the input-gradient chain run (a memset node between
    // two kernels of the main stream costs its 5 us and the programmatic launch overlap of its neighbours)
    if (!b->deterministic) {
        DG_CUDA(cudaEventRecord(b->ev_fork, c->stream));
        DG_CUDA(cudaStreamWaitEvent(b->side, b->ev_fork, 0));
        // data parallel, pull mode: the peers' SGD kernels of the previous step read this buffer; every rank is past them once all
        // ranks have reached this barrier (it is ordered after this rank's own SGD kernel)
        if (m->fused_step && b->p2p && b->p2p_pull) if (int rc = dg_dp_p2p_barrier(c, b->p2p, b->side)) return rc;
        DG_CUDA(cudaMemsetAsync(b->gacc, 0, b->gacc_floats * 4, b->side));
        DG_CUDA(cudaEventRecord(b->ev_memset, b->side));
    }
    if (int rc = launch_head(m, true, scale)) return rc;      // its per-position gradient partials are summed on the side stream
    DG_CUDA(cudaEventRecord(b->ev_head, c->stream));
    DG_CUDA(cudaStreamWaitEvent(b->side, b->ev_head, 0));
    {
        const LayerInfo &lh = m->layers[m->L - 1];
        cudaStream_t hst = b->side;
        dg_prof_scope _ps(c, DG_K_HEAD, hst);
        // a deferring fused step sums the head layer's gradient into the head block of the accumulation buffers
        float *hW = defer ? b->gacc : m->grads + lh.offW, *hb = defer ? b->gacc + lh.nW : m->grads + lh.offB;
        float *ha = defer ? b->gacc + lh.nW + lh.cout : m->grads + lh.offA;
        head_grad_reduce_kernel<<<(9 * m->C + DG_POINTS + 31) / 32 + 1, 256, 0, hst>>>(b->head_part, m->n, m->C, hW, hb, ha, m->d_loss_n, m->d_argmax,
                                                                                           m->d_targets, m->d_loss, m->d_errors);
        m->loss_reduced = true;
        c->launches++;
    }
    // ---- input gradients first: dZ_{l} -> dZ_{l-1} for conv layers l = L-2 .. 1 (transposed conv with flipped taps, masked by
    // the ReLU of layer l's input h_l); with C <= 128 the whole chain is one kernel with dZ resident in shared memory
    {
        int l = m->L - 2;
        while (l >= 1) {
            ConvParams cp{};
            cp.in = b->dzl[l];
            cp.n = m->n;
            cp.R_in = DG_R3; cp.pitch_in = b->g3.pitch; cp.pad_in = b->g3.pad; cp.ksize = 3; cp.first = b->g3.first; cp.ntiles = b->g3.ntiles;
            cp.R_mask = DG_R3; cp.pitch_mask = b->g3.pitch; cp.pad_mask = b->g3.pad;
            cp.mode = 1;
            int nchain = chain_ok(m) ? std::min(l, CONV_MAX_CHAIN) : 1;
            for (int j = 0; j < nchain; ++j) {
                cp.w[j] = b->ops[l - j].wd; cp.bias[j] = nullptr; cp.mask[j] = b->h[l - j]; cp.out[j] = b->dzl[l - j - 1];
            }
            cp.nchain = nchain;
            const bool into_first = (l - nchain) == 0;                 // the last output of this chain is dZ of conv layer 0
            const Geom &go = into_first ? b->g_first : b->g3;
            cp.R_out = go.rows; cp.pitch_out = go.pitch; cp.pad_out = go.pad;
            if (int rc = conv_cfg(m, cp, m->layers[l].cout, m->layers[l].cin)) return rc;
            if (int rc = launch_conv(m, cp)) return rc;
            l -= nchain;
        }
    }
    // every dZ_l is complete on the main stream: the side stream may read them
    if (!b->fused_bias) {
        DG_CUDA(cudaEventRecord(b->ev_dz, c->stream));
        DG_CUDA(cudaStreamWaitEvent(b->side, b->ev_dz, 0));
    }
    if (!b->deterministic) DG_CUDA(cudaStreamWaitEvent(c->stream, b->ev_memset, 0));     // the weight-gradient kernels add into the cleared buffers
    bool rec_reduce[2] = {false, false};   // events recorded in THIS pass (stream capture forbids waiting on older records)
    static long long *wdbg = nullptr;
    static const bool wtl = getenv("DG_TIMELINE") != nullptr;
    if (wtl && !wdbg) cudaMalloc(&wdbg, 148 * 8 * sizeof(long long));
    // bias gradients of every conv layer: side stream, all dZ are ready
    for (int l = m->L - 2; l >= 0 && !b->fused_bias; --l) {
        const LayerInfo &li = m->layers[l];
        const Geom &g = l == 0 ? b->g_first : b->g3;
        const int groups = m->n >= 64 ? 32 : (m->n >= 8 ? 8 : 1);
        dg_prof_scope _ps(c, DG_K_BIAS_GRAD, b->side);
        bias_grad_bf16_kernel<<<dim3(li.cout / 8, groups), 384, 0, b->side>>>(b->dzl[l], m->grads + li.offA, m->grads + li.offB, m->n, li.cout,
                                                                             g.rows, g.pitch, g.pad, groups);
        c->launches++;
    }
    for (const WgLaunch &q : wl) {
        const int l = q.l, nl = q.nl;
        const WgradPlan &w = q.w;
        const LayerInfo &li = m->layers[l];
        const LayerOps &o = b->ops[l];
        const Geom &g = l == 0 ? b->g_first : b->g3;
        const int par = l & 1;
        float *scratch = b->scratch + (size_t)par * (b->scratch_bytes / 4);
        WgradParams wp{};
        for (int j = 0; j < nl; ++j) {
            wp.dz[j] = b->dzl[l - j];
            wp.x[j] = (l - j) == 0 ? b->x0 : b->h[l - j];
            wp.gacc[j] = b->deterministic ? nullptr : b->gacc + b->gacc_off[l - j];
            wp.gab[j] = b->fused_bias ? b->gacc + b->gab_off[l - j] : nullptr;
        }
        wp.scratch = scratch; wp.nl = nl;
        wp.n = m->n; wp.cin = o.cin_pad; wp.cout = li.cout;
        wp.R = g.rows; wp.pitch = g.pitch; wp.pad = g.pad; wp.ksize = li.k; wp.first = g.first; wp.ntiles = g.ntiles;
        wp.ngroups = q.ngroups;
        wp.nb = w.nb; wp.n_mblk = w.n_mblk; wp.n_nblk = w.n_nblk; wp.splits = w.splits / q.ngroups; wp.nstages = w.nstages;
        wp.dbg = wtl ? wdbg : nullptr;
        wp.tl = c->tl; wp.tl_id = l == 0 ? DG_TL_WGRAD0 : DG_TL_WGRAD;
        // deterministic path: this scratch half was last read by the reduction two layers ago
        if (b->deterministic && rec_reduce[par]) DG_CUDA(cudaStreamWaitEvent(c->stream, b->ev_reduce[par], 0));
        { dg_prof_scope _ps(c, DG_K_WGRAD);
        if (w.rp == 96) DG_CUDA(dg_launch_pdl(wgrad_umma_kernel<96>, dim3(q.grid), dim3(384), w.smem, c->stream, wp));
        else DG_CUDA(dg_launch_pdl(wgrad_umma_kernel<128>, dim3(q.grid), dim3(384), w.smem, c->stream, wp));
        }
        c->launches++;
        if (wp.dbg) {
            std::vector<long long> h(148 * 8);
            cudaStreamSynchronize(c->stream);
            cudaMemcpy(h.data(), wp.dbg, h.size() * 8, cudaMemcpyDeviceToHost);
            double d1 = 0, d2 = 0, d3 = 0, d4 = 0, d5 = 0, d6 = 0; const int gsz = q.grid;
            for (int i = 0; i < gsz; ++i) { d1 += (double)(h[i * 8 + 1] - h[i * 8]) / gsz; d2 += (double)(h[i * 8 + 2] - h[i * 8]) / gsz; d3 += (double)h[i * 8 + 3] / gsz; d4 += (double)(h[i * 8 + 4] - h[i * 8]) / gsz; d5 += (double)h[i * 8 + 5] / gsz; d6 += (double)h[i * 8 + 6] / gsz; }
            fprintf(stderr, "wgrad timeline (cycles, mean over %d CTAs) k=%d layers=%d: mma loop start %.0f end %.0f (waiting for panels %.0f) | CTA end %.0f | producer: waiting for a free stage %.0f, issuing copies %.0f\n", gsz, li.k, nl, d1, d2, d3, d4, d5, d6);
        }
        if (b->deterministic) {
            DG_CUDA(cudaEventRecord(b->ev_wgrad, c->stream));
            DG_CUDA(cudaStreamWaitEvent(b->side, b->ev_wgrad, 0));
            for (int j = 0; j < nl; ++j) {
                const int lj = l - j;
                const LayerInfo &lij = m->layers[lj];
                dg_prof_scope _ps(c, DG_K_WGRAD_REDUCE, b->side);
                wgrad_reduce_kernel<<<b->num_sms * 8, 256, 0, b->side>>>(scratch, m->grads + lij.offW, lij.cin, b->ops[lj].cin_pad, lij.cout, lij.k, w.nb,
                                                                         w.n_mblk, w.n_nblk, w.splits);
                c->launches++;
            }
            DG_CUDA(cudaEventRecord(b->ev_reduce[par], b->side));
            rec_reduce[par] = true;
        }
    }
    b->grads_pending = defer;
    if (!b->deterministic && !defer) {
        // one launch folds every layer's accumulation buffer into the Torch7-order gradient
        DG_CUDA(cudaEventRecord(b->ev_wgrad, c->stream));
        DG_CUDA(cudaStreamWaitEvent(b->side, b->ev_wgrad, 0));
        dg_prof_scope _ps(c, DG_K_WGRAD_REDUCE, b->side);
        wgrad_finalize_kernel<<<dim3(96, m->L - 1), 256, FIN_SMEM_FLOATS * 4, b->side>>>(b->fin_table, c->tl);
        c->launches++;
    }
    // the side stream joins: head-layer gradient, (deterministic path) reductions, gradient vector
    DG_CUDA(cudaEventRecord(b->ev_join, b->side));
    DG_CUDA(cudaStreamWaitEvent(c->stream, b->ev_join, 0));
    // data parallel: ONE allreduce of the whole gradient after the last weight-gradient kernel.  Exchanging layer blocks earlier, under
    // the weight-gradient kernels, was built three ways and measured slower every time: those kernels keep the L2 busy
    // with operand traffic and red.global additions, and whatever shares it with them -- remote loads, fences, flags -- crawls.
    if (m->fused_step && c->world > 1) {
        const LayerInfo &l0 = m->layers[0], &lh = m->layers[m->L - 1];
        const size_t head_n = lh.nW + lh.cout + (size_t)lh.cout * DG_POINTS;
        if (defer) {
            if (b->p2p) { if (int rc = dg_dp_p2p_allreduce(c, b->p2p, c->stream, 0, b->gacc_floats, 0, b->p2p_pull ? 1 : 0)) return rc; }   // one kernel over NVLink peer memory
            else if (int rc = dg_dp_bucket(c, b->gacc, b->gacc_floats)) return rc;
        } else {
            if (int rc = dg_dp_bucket(c, m->grads + l0.offW, lh.offW + head_n - l0.offW)) return rc;
        }
    }
    DG_CUDA(cudaGetLastError());
    return 0;
}
