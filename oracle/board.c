/*
 * CPU oracle: restatement of deep-go's board analysis (makedata.lua) in plain C.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): checker for dg_summarize_boards /
 * dg_update_boards and the timed CPU baseline for that row.  Never linked into the product.
 *
 * Parity: PINNED.  tests/test_oracle_board.py replays all 4,398 fixture records of the
 * reference (tests/golden/positions.npz) through these functions: planes 2-6,8-9 of every
 * record and stones+age of every successor record must match bit-exactly.
 *
 * Board indexing follows the reference: board[x][y], x first (makedata.lua:63,80); here
 * h = 19*x + y with 0-based x,y (the reference's hash(), makedata.lua:188-191).
 */
#include <stdint.h>
#include <string.h>

#define N 19
#define NN 361

/* neighbour order {-1,0},{1,0},{0,-1},{0,1} (makedata.lua:199) */
static const int DX[4] = {-1, 1, 0, 0};
static const int DY[4] = {0, 0, -1, 1};

typedef struct {
    int count;            /* number of distinct liberties */
    int size;             /* stones in the group */
    int libs[NN];         /* liberty points (h), in discovery order */
    int stones[NN];       /* group points (h) */
} group_t;

/* count_liberties (makedata.lua:245-282): flood fill over same-colour 4-neighbours; liberties are
 * the distinct empty neighbours of the group.  Empty point -> 0. */
static int flood(const uint8_t *b, int h0, group_t *g)
{
    uint8_t seen[NN];
    int stack[NN], sp = 0;
    const uint8_t player = b[h0];
    g->count = 0;
    g->size = 0;
    if (player == 0) return 0;
    memset(seen, 0, sizeof seen);
    stack[sp++] = h0;
    seen[h0] = 1;
    while (sp) {
        const int h = stack[--sp];
        const int x = h / N, y = h % N;
        g->stones[g->size++] = h;
        for (int d = 0; d < 4; ++d) {
            const int a = x + DX[d], c = y + DY[d];
            if (a < 0 || a >= N || c < 0 || c >= N) continue;
            const int nh = a * N + c;
            if (seen[nh]) continue;
            if (b[nh] == player) { seen[nh] = 1; stack[sp++] = nh; }
            else if (b[nh] == 0) { seen[nh] = 2; g->libs[g->count++] = nh; }
        }
    }
    return g->count;
}

typedef struct { int h; uint8_t old; } undo_t;
typedef struct { undo_t e[4 * NN]; int n; } undo_stack_t;

static void set_pt(uint8_t *b, int h, uint8_t p, undo_stack_t *u)
{
    u->e[u->n].h = h;
    u->e[u->n].old = b[h];
    u->n++;
    b[h] = p;
}

static void unwind(uint8_t *b, undo_stack_t *u, int to)
{
    while (u->n > to) { u->n--; b[u->e[u->n].h] = u->e[u->n].old; }
}

/* play_with_f + apply_f_to_dead_neighbors + apply_f_if_dead (makedata.lua:224-241,388-391):
 * place the stone; remove each opposite-colour neighbour group left without liberties; then the
 * own group if it has none (suicide).  Returns opponent stones removed ("kills"). */
static int play(uint8_t *b, int h, uint8_t player, undo_stack_t *u)
{
    group_t g;
    int kills = 0;
    const int x = h / N, y = h % N;
    set_pt(b, h, player, u);
    for (int d = 0; d < 4; ++d) {
        const int a = x + DX[d], c = y + DY[d];
        if (a < 0 || a >= N || c < 0 || c >= N) continue;
        const int nh = a * N + c;
        if (b[nh] == 3 - b[h] && b[nh] != 0) {
            if (flood(b, nh, &g) == 0) {
                for (int i = 0; i < g.size; ++i) { set_pt(b, g.stones[i], 0, u); kills++; }
            }
        }
    }
    if (b[h] != 0 && flood(b, h, &g) == 0)
        for (int i = 0; i < g.size; ++i) set_pt(b, g.stones[i], 0, u);
    return kills;
}

/* ladder_moves (makedata.lua:393-439).  Returns a 2-bit mask: bit i set when liberty i is a
 * ladder move for the attacker (3 - colour of the group at h0). */
static int ladder_moves(uint8_t *b, int h0, const int lib[2])
{
    const uint8_t player = b[h0], opp = (uint8_t)(3 - player);
    int result = 0;
    undo_stack_t u;
    group_t g;
    u.n = 0;
    for (int i = 0; i < 2; ++i) {
        const int a = lib[i], o = lib[1 - i];
        play(b, a, opp, &u);
        if (flood(b, a, &g) > 2) {
            play(b, o, player, &u);
            const int n = flood(b, o, &g);
            if (n == 1) {
                result |= 1 << i;
            } else if (n == 2) {
                const int nl[2] = {g.libs[0], g.libs[1]};
                group_t ga;
                if (flood(b, a, &ga) > 1 && ladder_moves(b, h0, nl) != 0) result |= 1 << i;
            }
        }
        unwind(b, &u, 0);
    }
    return result;
}

/*
 * summarize_board (makedata.lua:143-153) = all_kills_and_liberties_after (:122-141) +
 * all_ladder_moves_and_liberties (:441-479).
 * out planes (each 361 bytes, [x][y]): 0 liberties, 1-2 liberties_after[p=1,2], 3-4 kills[p=1,2],
 * 5-6 ladders[p=1,2] -- record planes 2..6 and 8..9 (dataloader.lua:20-27); values wrap mod 256
 * like torch.ByteTensor(table) (makedata.lua:147-151).
 */
void dg_oracle_summarize(const uint8_t *stones, uint8_t *out)
{
    uint8_t b[NN];
    uint8_t considered[NN];
    group_t g;
    undo_stack_t u;
    memcpy(b, stones, NN);
    memset(out, 0, 7 * NN);
    uint8_t *lib = out, *la = out + NN, *kl = out + 3 * NN, *lad = out + 5 * NN;

    for (int k = 1; k <= 2; ++k)
        for (int h = 0; h < NN; ++h) {
            if (b[h] != 0) continue;
            u.n = 0;
            const int kills = play(b, h, (uint8_t)k, &u);
            const int l = flood(b, h, &g);
            unwind(b, &u, 0);
            kl[(k - 1) * NN + h] = (uint8_t)kills;
            la[(k - 1) * NN + h] = (uint8_t)l;
        }

    memset(considered, 0, sizeof considered);
    for (int h = 0; h < NN; ++h) {
        if (b[h] == 0 || considered[h]) continue;
        const int n = flood(b, h, &g);
        for (int i = 0; i < g.size; ++i) { considered[g.stones[i]] = 1; lib[g.stones[i]] = (uint8_t)n; }
        if (n == 2) {
            const int l2[2] = {g.libs[0], g.libs[1]};
            const int size = g.size;
            const int owner = b[h];
            const int m = ladder_moves(b, h, l2);
            for (int i = 0; i < 2; ++i)
                if (m & (1 << i)) lad[(3 - owner - 1) * NN + l2[i]] = (uint8_t)size;
        }
    }
}

/* update_board (makedata.lua:329-354): ages 1..254 increment; every point set by the move (the
 * stone placed, stones captured) gets age 1.  Returns -1 if the point is occupied. */
int dg_oracle_update_board(uint8_t *stones, uint8_t *age, int x0, int y0, int player)
{
    undo_stack_t u;
    const int h = x0 * N + y0;
    if (stones[h] != 0) return -1;
    for (int i = 0; i < NN; ++i)
        if (age[i] > 0 && age[i] < 255) age[i]++;
    u.n = 0;
    play(stones, h, (uint8_t)player, &u);
    for (int i = 0; i < u.n; ++i) age[u.e[i].h] = 1;
    return 0;
}

/* batch helpers for the timed CPU baseline */
void dg_oracle_summarize_batch(const uint8_t *stones, uint8_t *out, int n)
{
#pragma omp parallel for schedule(dynamic, 8)
    for (int i = 0; i < n; ++i) dg_oracle_summarize(stones + (size_t)i * NN, out + (size_t)i * 7 * NN);
}
