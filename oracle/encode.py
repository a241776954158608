"""CPU oracle: numpy restatement of deep-go's feature-plane encoder (dataloader.lua).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
Parity: UNPINNED for preprocess (the reference has no test or golden output for it and Torch7
cannot run here); the code follows dataloader.lua line by line.  Symmetry ids 1..7 are UNPINNED
by construction: the reference's transform() is the identity (dataloader.lua:41-44); id 0 is the
reference behaviour, ids 1..7 follow the spec in DESIGN.md.
"""
import numpy as np

# channel bases, 0-based (dataloader.lua:6-14 are 1-based)
STONE, LIBERTIES, LIBERTIES_AFTER, KILL, AGE, LADDER, RANK = 0, 3, 7, 14, 21, 26, 27
TOTAL = 37
SIZE = 19
# raw plane indices, 0-based (dataloader.lua:20-27)
IN_STONE, IN_LIB, IN_LIB_AFTER, IN_KILLS, IN_AGE, IN_LADDER, IN_TOTAL = 0, 1, 2, 4, 6, 7, 9


def transform(grid, sym):
    """D4 action on [..., x, y] grids. bits: b0 transpose, b1 flip x, b2 flip y, applied in that order
    (the stone at (x,y) moves to T_s(x,y)).  sym 0 == dataloader.lua:41-44."""
    g = grid
    if sym & 1:
        g = np.swapaxes(g, -1, -2)
    if sym & 2:
        g = g[..., ::-1, :]
    if sym & 4:
        g = g[..., :, ::-1]
    return np.ascontiguousarray(g)


def transform_xy(x, y, sym, size=SIZE):
    """Same action on a 0-based point."""
    if sym & 1:
        x, y = y, x
    if sym & 2:
        x = size - 1 - x
    if sym & 4:
        y = size - 1 - y
    return x, y


def inverse_sym(sym):
    """T_inv: if no transpose the element is an involution; else the two flips swap."""
    if not (sym & 1):
        return sym
    return 1 | ((sym & 2) << 1) | ((sym & 4) >> 1)


def preprocess(raw, player, rank, sym=0, dtype=np.float64):
    """One record -> 37 planes (dataloader.lua:50-92).

    raw u8[9,19,19]; player in {1,2} (data.move.player); rank = data.ranks[player] in 1..9.
    """
    raw = transform(np.asarray(raw, dtype=np.uint8), sym)
    out = np.zeros((TOTAL, SIZE, SIZE), dtype)
    stone = raw[IN_STONE]
    out[STONE] = stone == 0                                   # :57
    out[STONE + 1] = stone == player                          # :58
    out[STONE + 2] = stone == 3 - player                      # :59
    lib = raw[IN_LIB]
    for i in range(1, 4):
        out[LIBERTIES + i - 1] = lib == i                     # :68
    out[LIBERTIES + 3] = lib >= 4                             # :69
    la = raw[IN_LIB_AFTER + player - 1]                       # :72
    out[LIBERTIES_AFTER] = out[STONE] * (la == 0)             # :73
    for i in range(1, 6):
        out[LIBERTIES_AFTER + i] = la == i                    # :74
    out[LIBERTIES_AFTER + 6] = la >= 6                        # :75
    kills = raw[IN_KILLS + player - 1]                        # :77
    for i in range(1, 7):
        out[KILL + i - 1] = kills == i                        # :78
    out[KILL + 6] = kills >= 7                                # :79
    age = raw[IN_AGE]
    for i in range(1, 6):
        out[AGE + i - 1] = age == i                           # :82
    ladder = raw[IN_LADDER + player - 1]                      # :84
    out[LADDER] = ladder >= 1                                 # :85
    if not 1 <= rank <= 9:
        raise IndexError("rank %d out of range (dataloader.lua:87 would index past the tensor)" % rank)
    out[RANK + rank] = 1                                      # :87  (plane RANK itself is never written)
    return out


def label(x, y, sym=0):
    """0-based class index of a 1-based move: reference output = 19*(x-1)+y (dataloader.lua:89)."""
    tx, ty = transform_xy(x - 1, y - 1, sym)
    return SIZE * tx + ty


def load_minibatch(raw, player, rank, x, y, sym=None, dtype=np.float64):
    """Batch form of load_minibatch (dataloader.lua:113-121): -> input[n,37,19,19], output0[n] (0-based)."""
    n = len(raw)
    sym = np.zeros(n, np.uint8) if sym is None else sym
    inp = np.zeros((n, TOTAL, SIZE, SIZE), dtype)
    out = np.zeros(n, np.int32)
    for i in range(n):
        inp[i] = preprocess(raw[i], int(player[i]), int(rank[i]), int(sym[i]), dtype)
        out[i] = label(int(x[i]), int(y[i]), int(sym[i]))
    return inp, out
