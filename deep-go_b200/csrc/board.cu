// Board analysis on the GPU: summarize_board of makedata.lua:143-153, i.e. per position
//   liberties of every group (makedata.lua:441-468), kills / liberties-after of both colours at every empty point
//   (:122-141, :304-327) and ladder captures (:393-439, :470-474).
// Not a translation of the reference's table-walking flood fills: the board is held as 361-bit sets.
//   kernel 1 (one CTA of 384 threads per board): connected components by min-label propagation in shared memory, then per
//     group a stone bitset and a neighbourhood bitset; liberties = popcount(neighbourhood & empty).  An imagined move at an
//     empty point e for colour p needs no flood fill at all: captured groups are the adjacent opponent groups whose single
//     liberty is e, and the liberties afterwards are popcount((N(e) | N(own adjacent groups)) & ~merged stones & empty'),
//     empty' = empty - e + captured stones.
//   kernel 2 (one warp per board): the ladder search for every group with exactly two liberties, one group per lane, on
//     bitboards kept in registers/local memory (bit-parallel dilation flood fill), recursion unrolled onto an explicit stack.
// Bit/point index h = 19*x + y with x the first board index, as the reference's hash() (makedata.lua:188-191).
#include "common.cuh"
#include "../../include/deepgo.h"

#define NW 12                       // 361 bits -> 12 words
#define LADDER_MAX_DEPTH 64         // explicit recursion stack (the reference recurses unbounded; the fixture corpus reaches 37)
#define MAX_GROUPS2 180             // groups with exactly two liberties per board (a board has at most 180 groups)

struct BB { uint32_t w[NW]; };

__constant__ uint32_t c_col0[NW], c_col18[NW];     // points with y == 0 / y == 18

__device__ __forceinline__ BB bb_zero() { BB r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = 0; return r; }
__device__ __forceinline__ bool bb_test(const BB &b, int h) { return (b.w[h >> 5] >> (h & 31)) & 1u; }
__device__ __forceinline__ void bb_set(BB &b, int h) { b.w[h >> 5] |= 1u << (h & 31); }
__device__ __forceinline__ int bb_count(const BB &b) { int c = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) c += __popc(b.w[i]); return c; }
__device__ __forceinline__ bool bb_eq(const BB &a, const BB &b) { uint32_t d = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) d |= a.w[i] ^ b.w[i]; return d == 0; }

// all 4-neighbours of the set (bits >= 361 are cleared)
__device__ __forceinline__ BB bb_dilate(const BB &b) {
    BB r;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        uint32_t up = (b.w[i] << 19) | (i > 0 ? b.w[i - 1] >> 13 : 0u);                     // h + 19 (x + 1)
        uint32_t dn = (b.w[i] >> 19) | (i < NW - 1 ? b.w[i + 1] << 13 : 0u);                // h - 19
        const uint32_t nr = b.w[i] & ~c_col18[i], nl = b.w[i] & ~c_col0[i];
        const uint32_t nr_prev = i > 0 ? (b.w[i - 1] & ~c_col18[i - 1]) : 0u;
        const uint32_t nl_next = i < NW - 1 ? (b.w[i + 1] & ~c_col0[i + 1]) : 0u;
        uint32_t rt = (nr << 1) | (nr_prev >> 31);                                          // h + 1 (y + 1)
        uint32_t lf = (nl >> 1) | (nl_next << 31);                                          // h - 1
        r.w[i] = up | dn | rt | lf;
    }
    r.w[NW - 1] &= 0x1FFu;                                                                  // bits 352..360 only
    return r;
}

// connected group of `colour` containing `seed` (bit-parallel flood fill)
__device__ __forceinline__ BB bb_flood(const BB &colour, int seed) {
    BB g = bb_zero();
    bb_set(g, seed);
    for (int it = 0; it < 361; ++it) {
        const BB d = bb_dilate(g);
        BB nx;
#pragma unroll
        for (int i = 0; i < NW; ++i) nx.w[i] = g.w[i] | (d.w[i] & colour.w[i]);
        if (bb_eq(nx, g)) break;
        g = nx;
    }
    return g;
}

__device__ __forceinline__ int group_liberties(const BB &g, const BB &mine, const BB &theirs) {
    const BB d = bb_dilate(g);
    int c = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) c += __popc(d.w[i] & ~(mine.w[i] | theirs.w[i]));
    return c;
}

// play_with_f + apply_f_to_dead_neighbors + apply_f_if_dead (makedata.lua:224-241,388-391): place a stone of the colour whose
// bitboard is `mine`, remove every adjacent opponent group left without liberties (neighbour order -x,+x,-y,+y), then the own
// group if it has none.
__device__ void bb_play(BB &mine, BB &theirs, int h) {
    bb_set(mine, h);
    const int x = h / DG_BOARD, y = h - x * DG_BOARD;
    const int nb[4] = {x > 0 ? h - DG_BOARD : -1, x < DG_BOARD - 1 ? h + DG_BOARD : -1, y > 0 ? h - 1 : -1, y < DG_BOARD - 1 ? h + 1 : -1};
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        const int q = nb[d];
        if (q < 0 || !bb_test(theirs, q)) continue;
        const BB g = bb_flood(theirs, q);
        if (group_liberties(g, mine, theirs) == 0) {
#pragma unroll
            for (int i = 0; i < NW; ++i) theirs.w[i] &= ~g.w[i];
        }
    }
    const BB g = bb_flood(mine, h);
    if (group_liberties(g, mine, theirs) == 0) {
#pragma unroll
        for (int i = 0; i < NW; ++i) mine.w[i] &= ~g.w[i];
    }
}

// liberties of the group at h (0 when h is empty) and, optionally, the liberty set
__device__ int libs_at(const BB &mine, const BB &theirs, int h, BB *libset) {
    const bool m = bb_test(mine, h), t = bb_test(theirs, h);
    if (!m && !t) return 0;
    const BB g = bb_flood(m ? mine : theirs, h);
    const BB d = bb_dilate(g);
    BB l;
    int c = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) { l.w[i] = d.w[i] & ~(mine.w[i] | theirs.w[i]); c += __popc(l.w[i]); }
    if (libset) *libset = l;
    return c;
}

__device__ __forceinline__ void two_bits(const BB &l, int &a, int &b) {
    a = b = -1;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        uint32_t w = l.w[i];
        while (w) {
            const int h = i * 32 + __ffs(w) - 1;
            w &= w - 1;
            if (a < 0) a = h; else b = h;
        }
    }
}

struct LadderFrame { BB def, att; short lib[2]; short i, result; };   // boards at frame entry: defender's and attacker's stones

// ladder_moves (makedata.lua:393-439) for the defender group whose two liberties are lib0, lib1; returns a 2-bit mask
// (bit i: the attacker playing lib[i] captures the group in a ladder).
__device__ int ladder_search(const BB &def0, const BB &att0, int lib0, int lib1, LadderFrame *stack) {
    int top = 0;
    stack[0].def = def0; stack[0].att = att0; stack[0].lib[0] = (short)lib0; stack[0].lib[1] = (short)lib1; stack[0].i = 0; stack[0].result = 0;
    int ret = 0;
    while (true) {
        LadderFrame &f = stack[top];
        if (f.i == 2) {                                   // both orderings tried: return to the caller
            ret = f.result;
            if (top == 0) break;
            --top;
            LadderFrame &par = stack[top];
            if (ret != 0) par.result |= 1 << (par.i - 1);  // the recursive call belonged to the parent's trial i-1
            continue;
        }
        const int i = f.i++;
        const int a = f.lib[i], o = f.lib[1 - i];
        BB def = f.def, att = f.att;
        bb_play(att, def, a);                              // attacker plays a
        if (libs_at(att, def, a, nullptr) > 2) {
            bb_play(def, att, o);                          // defender extends at o
            BB lset;
            const int n = libs_at(def, att, o, &lset);
            if (n == 1) {
                f.result |= 1 << i;
            } else if (n == 2) {
                if (libs_at(att, def, a, nullptr) > 1 && top + 1 < LADDER_MAX_DEPTH) {
                    int l0, l1;
                    two_bits(lset, l0, l1);
                    LadderFrame &nf = stack[top + 1];
                    nf.def = def; nf.att = att; nf.lib[0] = (short)l0; nf.lib[1] = (short)l1; nf.i = 0; nf.result = 0;
                    ++top;
                }
            }
        }
    }
    return ret;
}

struct Group2 { short root, lib0, lib1, size; };    // a group with exactly two liberties; colour = stones[root]

// ---------------------------------------------------------------------------------------------
// kernel 1: groups, liberties, kills, liberties-after
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(384) board_groups_kernel(const uint8_t *__restrict__ stones, uint8_t *__restrict__ out,
                                                           Group2 *__restrict__ g2list, int *__restrict__ g2count) {
    __shared__ uint8_t st[DG_POINTS];
    __shared__ short root[DG_POINTS];
    __shared__ uint32_t gstones[DG_POINTS][NW], gnbr[DG_POINTS][NW];       // indexed by the group's root (min h of the group)
    __shared__ uint32_t empty[NW];
    __shared__ short glibs[DG_POINTS], gsize[DG_POINTS];
    __shared__ int n2;
    const int b = blockIdx.x, h = threadIdx.x;
    const bool on = h < DG_POINTS;
    const int x = on ? h / DG_BOARD : 0, y = on ? h - x * DG_BOARD : 0;
    const int nb[4] = {on && x > 0 ? h - DG_BOARD : -1, on && x < DG_BOARD - 1 ? h + DG_BOARD : -1, on && y > 0 ? h - 1 : -1, on && y < DG_BOARD - 1 ? h + 1 : -1};
    if (on) { st[h] = stones[(size_t)b * DG_POINTS + h]; }
    if (h < NW) empty[h] = 0;
    if (h == 0) n2 = 0;
    for (int i = h; i < DG_POINTS * NW; i += blockDim.x) { (&gstones[0][0])[i] = 0; (&gnbr[0][0])[i] = 0; }
    __syncthreads();
    const int me = on ? st[h] : 0;
    if (on) root[h] = me ? (short)h : (short)-1;
    __syncthreads();
    // min-label propagation with pointer jumping until nothing changes
    for (int it = 0; it < 400; ++it) {
        int r = me ? root[h] : -1;
        bool changed = false;
        if (me) {
#pragma unroll
            for (int d = 0; d < 4; ++d)
                if (nb[d] >= 0 && st[nb[d]] == me) r = min(r, (int)root[nb[d]]);
            r = min(r, (int)root[r]);
            changed = r != root[h];
        }
        __syncthreads();
        if (me) root[h] = (short)r;
        if (!__syncthreads_or(changed)) break;
    }
    if (on) {
        if (me) {
            const int r = root[h];
            atomicOr(&gstones[r][h >> 5], 1u << (h & 31));
#pragma unroll
            for (int d = 0; d < 4; ++d)
                if (nb[d] >= 0) atomicOr(&gnbr[r][nb[d] >> 5], 1u << (nb[d] & 31));
        } else {
            atomicOr(&empty[h >> 5], 1u << (h & 31));
        }
    }
    __syncthreads();
    if (on && me && root[h] == h) {
        int l = 0, s = 0;
        uint32_t lw[NW];
#pragma unroll
        for (int i = 0; i < NW; ++i) { lw[i] = gnbr[h][i] & empty[i]; l += __popc(lw[i]); s += __popc(gstones[h][i]); }
        glibs[h] = (short)l; gsize[h] = (short)s;
        if (l == 2) {
            const int k = atomicAdd(&n2, 1);
            if (k < MAX_GROUPS2) {
                int a = -1, c = -1;
                for (int i = 0; i < NW; ++i) { uint32_t w = lw[i]; while (w) { const int q = i * 32 + __ffs(w) - 1; w &= w - 1; if (a < 0) a = q; else c = q; } }
                Group2 g; g.root = (short)h; g.lib0 = (short)a; g.lib1 = (short)c; g.size = (short)s;
                g2list[(size_t)b * MAX_GROUPS2 + k] = g;
            }
        }
    }
    __syncthreads();
    uint8_t *o = out + (size_t)b * 7 * DG_POINTS;
    if (on) o[h] = me ? (uint8_t)glibs[root[h]] : 0;                       // plane 0: liberties of the group on the point
    if (h == 0) g2count[b] = min(n2, MAX_GROUPS2);
    // imagined moves: 2 colours x 361 points
    for (int t = h; t < 2 * DG_POINTS; t += blockDim.x) {
        const int p = t / DG_POINTS + 1, e = t - (p - 1) * DG_POINTS;
        int kills = 0, la = 0;
        if (st[e] == 0) {
            const int ex = e / DG_BOARD, ey = e - ex * DG_BOARD;
            const int en[4] = {ex > 0 ? e - DG_BOARD : -1, ex < DG_BOARD - 1 ? e + DG_BOARD : -1, ey > 0 ? e - 1 : -1, ey < DG_BOARD - 1 ? e + 1 : -1};
            uint32_t S[NW], Nb[NW], freed[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) { S[i] = 0; Nb[i] = 0; freed[i] = 0; }
            S[e >> 5] |= 1u << (e & 31);
            int seen[4] = {-1, -1, -1, -1};
#pragma unroll
            for (int d = 0; d < 4; ++d) {
                const int q = en[d];
                if (q < 0) continue;
                Nb[q >> 5] |= 1u << (q & 31);
                const int c = st[q];
                if (c == 0) continue;
                const int r = root[q];
                bool dup = false;
#pragma unroll
                for (int k = 0; k < 4; ++k) dup |= seen[k] == r;
                if (dup) continue;
                seen[d] = r;
                if (c == p) {
#pragma unroll
                    for (int i = 0; i < NW; ++i) { S[i] |= gstones[r][i]; Nb[i] |= gnbr[r][i]; }
                } else if (glibs[r] <= 1) {                                // its only liberty is e (or it had none): captured
                    kills += gsize[r];
#pragma unroll
                    for (int i = 0; i < NW; ++i) freed[i] |= gstones[r][i];
                }
            }
#pragma unroll
            for (int i = 0; i < NW; ++i) la += __popc(Nb[i] & ~S[i] & (empty[i] | freed[i]));   // e itself is in S
        }
        o[(size_t)(1 + (p - 1)) * DG_POINTS + e] = (uint8_t)la;            // planes 1-2: liberties_after[p]
        o[(size_t)(3 + (p - 1)) * DG_POINTS + e] = (uint8_t)kills;         // planes 3-4: kills[p]
    }
}

// ---------------------------------------------------------------------------------------------
// kernel 2: ladders
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) board_ladders_kernel(const uint8_t *__restrict__ stones, const Group2 *__restrict__ g2list,
                                                           const int *__restrict__ g2count, uint8_t *__restrict__ out) {
    __shared__ int key[2][DG_POINTS];
    const int b = blockIdx.x, lane = threadIdx.x;
    for (int i = lane; i < 2 * DG_POINTS; i += 32) (&key[0][0])[i] = 0;
    BB black = bb_zero(), white = bb_zero();
    const uint8_t *s = stones + (size_t)b * DG_POINTS;
    for (int h = 0; h < DG_POINTS; ++h) {
        const int c = s[h];
        if (c == 1) bb_set(black, h); else if (c == 2) bb_set(white, h);
    }
    __syncwarp();
    LadderFrame stack[LADDER_MAX_DEPTH];
    const int cnt = g2count[b];
    for (int k = lane; k < cnt; k += 32) {
        const Group2 g = g2list[(size_t)b * MAX_GROUPS2 + k];
        const int owner = s[g.root];
        const int m = owner == 1 ? ladder_search(black, white, g.lib0, g.lib1, stack) : ladder_search(white, black, g.lib0, g.lib1, stack);
        // ladders[3-owner][move] = groupsize, groups visited in scan order of their first stone: the later group wins
        // (makedata.lua:455-474) -> keep the entry with the largest root index
        const int v = ((g.root + 1) << 8) | (g.size & 255);
        if (m & 1) atomicMax(&key[2 - owner][g.lib0], v);
        if (m & 2) atomicMax(&key[2 - owner][g.lib1], v);
    }
    __syncwarp();
    uint8_t *o = out + (size_t)b * 7 * DG_POINTS + 5 * DG_POINTS;         // planes 5-6: ladders[p=1,2]
    for (int i = lane; i < 2 * DG_POINTS; i += 32) o[i] = (uint8_t)((&key[0][0])[i] & 255);
}

// update_board (makedata.lua:329-354): every age in 1..254 grows by one, the stone is placed, captured groups (and the
// own group on suicide) are removed, and every point whose state the move set gets age 1.  One CTA per board.
__global__ void __launch_bounds__(384) update_boards_kernel(uint8_t *__restrict__ stones, uint8_t *__restrict__ age, const uint8_t *__restrict__ mx,
                                                            const uint8_t *__restrict__ my, const uint8_t *__restrict__ mplayer, int *__restrict__ status) {
    __shared__ uint32_t before[2][NW], after[2][NW];
    const int b = blockIdx.x, h = threadIdx.x;
    uint8_t *s = stones + (size_t)b * DG_POINTS, *a = age + (size_t)b * DG_POINTS;
    const int move = (mx[b] - 1) * DG_BOARD + (my[b] - 1), player = mplayer[b];
    if (h < 2 * NW) (&before[0][0])[h] = 0;
    __syncthreads();
    const int c = h < DG_POINTS ? s[h] : 0;
    if (c == 1 || c == 2) atomicOr(&before[c - 1][h >> 5], 1u << (h & 31));
    __syncthreads();
    if (h == 0) {
        if (s[move] != 0) {
            status[b] = 1;                                  // "playing somewhere that's already been played!" (makedata.lua:352)
            for (int i = 0; i < NW; ++i) { after[0][i] = before[0][i]; after[1][i] = before[1][i]; }
        } else {
            status[b] = 0;
            BB mine, theirs;
            for (int i = 0; i < NW; ++i) { mine.w[i] = before[player - 1][i]; theirs.w[i] = before[2 - player][i]; }
            bb_play(mine, theirs, move);
            for (int i = 0; i < NW; ++i) { after[player - 1][i] = mine.w[i]; after[2 - player][i] = theirs.w[i]; }
        }
    }
    __syncthreads();
    if (h < DG_POINTS && status[b] == 0) {
        const int nb = ((after[0][h >> 5] >> (h & 31)) & 1u) ? 1 : (((after[1][h >> 5] >> (h & 31)) & 1u) ? 2 : 0);
        int ag = a[h];
        if (ag > 0 && ag < 255) ag++;
        // set() is called for the placed stone and for every removed stone (also the placed stone itself on suicide)
        if (nb != c || h == move) ag = 1;
        s[h] = (uint8_t)nb;
        a[h] = (uint8_t)ag;
    }
}

static bool g_consts_set[64] = {false};

static int board_consts(dg_ctx *c) {
    if (!g_consts_set[c->dev & 63]) {
        uint32_t c0[NW] = {0}, c18[NW] = {0};
        for (int x = 0; x < DG_BOARD; ++x) {
            const int a = x * DG_BOARD, b = x * DG_BOARD + DG_BOARD - 1;
            c0[a >> 5] |= 1u << (a & 31);
            c18[b >> 5] |= 1u << (b & 31);
        }
        DG_CUDA(cudaMemcpyToSymbol(c_col0, c0, sizeof c0));
        DG_CUDA(cudaMemcpyToSymbol(c_col18, c18, sizeof c18));
        g_consts_set[c->dev & 63] = true;
    }
    return 0;
}

extern "C" int dg_update_boards(dg_ctx *c, uint8_t *stones, uint8_t *age, const uint8_t *x, const uint8_t *y, const uint8_t *player, int n) {
    DG_REQUIRE(c, "dg_update_boards: ctx is NULL");
    DG_REQUIRE(n >= 0, "dg_update_boards: n < 0");
    if (n == 0) return 0;
    DG_REQUIRE(stones && age && x && y && player, "dg_update_boards: NULL buffer");
    for (int i = 0; i < n; ++i) {
        DG_REQUIRE(x[i] >= 1 && x[i] <= 19 && y[i] >= 1 && y[i] <= 19, "move %d: (%d,%d) outside the board", i, x[i], y[i]);
        DG_REQUIRE(player[i] == 1 || player[i] == 2, "move %d: player %d not in {1,2}", i, player[i]);
    }
    DG_CUDA(cudaSetDevice(c->dev));
    if (int rc = board_consts(c)) return rc;
    const size_t bb = (size_t)n * DG_POINTS;
    const size_t off_age = (bb + 255) / 256 * 256, off_m = off_age + (bb + 255) / 256 * 256, off_st = (off_m + 3 * (size_t)n + 255) / 256 * 256;
    char *d = nullptr;
    if (int rc = dg_ctx_scratch(c, off_st + (size_t)n * 4, (void **)&d)) return rc;
    DG_CUDA(cudaMemcpyAsync(d, stones, bb, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + off_age, age, bb, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + off_m, x, n, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + off_m + n, y, n, cudaMemcpyHostToDevice, c->stream));
    DG_CUDA(cudaMemcpyAsync(d + off_m + 2 * (size_t)n, player, n, cudaMemcpyHostToDevice, c->stream));
    update_boards_kernel<<<n, 384, 0, c->stream>>>((uint8_t *)d, (uint8_t *)d + off_age, (const uint8_t *)d + off_m, (const uint8_t *)d + off_m + n,
                                                   (const uint8_t *)d + off_m + 2 * (size_t)n, (int *)(d + off_st));
    c->launches++;
    DG_CUDA(cudaGetLastError());
    std::vector<int> st(n);
    DG_CUDA(cudaMemcpyAsync(stones, d, bb, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaMemcpyAsync(age, d + off_age, bb, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaMemcpyAsync(st.data(), d + off_st, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < n; ++i)
        DG_REQUIRE(st[i] == 0, "dg_update_boards: board %d: playing somewhere that's already been played (makedata.lua:352); that board is left unchanged", i);
    return 0;
}

extern "C" int dg_summarize_boards(dg_ctx *c, const uint8_t *stones, uint8_t *out, int n) {
    DG_REQUIRE(c, "dg_summarize_boards: ctx is NULL");
    DG_REQUIRE(n >= 0, "dg_summarize_boards: n < 0");
    if (n == 0) return 0;
    DG_REQUIRE(stones && out, "dg_summarize_boards: NULL buffer");
    for (size_t i = 0; i < (size_t)n * DG_POINTS; ++i)
        DG_REQUIRE(stones[i] <= 2, "dg_summarize_boards: board %zu has a point value %d outside {0,1,2}", i / DG_POINTS, stones[i]);
    DG_CUDA(cudaSetDevice(c->dev));
    if (int rc = board_consts(c)) return rc;
    const size_t in_b = (size_t)n * DG_POINTS, out_b = (size_t)n * 7 * DG_POINTS;
    const size_t off_out = (in_b + 255) / 256 * 256;
    const size_t off_list = (off_out + out_b + 255) / 256 * 256;
    const size_t off_cnt = off_list + (size_t)n * MAX_GROUPS2 * sizeof(Group2);
    char *d = nullptr;
    if (int rc = dg_ctx_scratch(c, off_cnt + (size_t)n * 4, (void **)&d)) return rc;
    DG_CUDA(cudaMemcpyAsync(d, stones, in_b, cudaMemcpyHostToDevice, c->stream));
    board_groups_kernel<<<n, 384, 0, c->stream>>>((const uint8_t *)d, (uint8_t *)d + off_out, (Group2 *)(d + off_list), (int *)(d + off_cnt));
    board_ladders_kernel<<<n, 32, 0, c->stream>>>((const uint8_t *)d, (const Group2 *)(d + off_list), (const int *)(d + off_cnt), (uint8_t *)d + off_out);
    c->launches += 2;
    DG_CUDA(cudaGetLastError());
    DG_CUDA(cudaMemcpyAsync(out, d + off_out, out_b, cudaMemcpyDeviceToHost, c->stream));
    DG_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}
