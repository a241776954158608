"""Oracle self-checks for the feature-plane encoder restatement (dataloader.lua:50-92) and the D4 spec."""
import numpy as np
import pytest

from oracle import encode


def _sample(golden, k=64, seed=0):
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(golden["raw"]), k)
    pl = golden["player"][idx]
    return idx, golden["raw"][idx], pl, golden["ranks"][idx, pl - 1], golden["x"][idx], golden["y"][idx]


def test_plane_structure(golden):
    idx, raw, pl, rk, x, y = _sample(golden)
    inp, out = encode.load_minibatch(raw, pl, rk, x, y)
    assert inp.shape == (64, 37, 19, 19) and inp.dtype == np.float64      # dataloader.lua:94-111
    assert set(np.unique(inp)) <= {0.0, 1.0}
    assert np.all(inp[:, 0:3].sum(1) == 1)                                 # empty / own / opponent partition
    assert np.all(inp[:, encode.RANK] == 0)                                # plane 28 (1-based) is never written
    for i in range(64):
        assert np.all(inp[i, encode.RANK + rk[i]] == 1)
        assert inp[i, encode.RANK + 1:].sum() == 361
        # the move played is on an empty point and the label addresses it (dataloader.lua:89)
        assert inp[i, 0].reshape(-1)[out[i]] == 1
        assert out[i] == 19 * (int(x[i]) - 1) + int(y[i]) - 1
    # liberties planes only on stones, liberties_after/kills only on empties
    assert np.all(inp[:, 3:7].sum(1) <= 1 - inp[:, 0])
    assert np.all(inp[:, 7:14].sum(1) <= inp[:, 0])


def test_thresholds_on_fuzzed_bytes():
    rng = np.random.default_rng(1)
    raw = rng.integers(0, 256, (9, 19, 19)).astype(np.uint8)
    raw[0] = rng.choice([0, 1, 2], (19, 19), p=[.6, .2, .2])
    for player in (1, 2):
        o = encode.preprocess(raw, player, 5)
        la = raw[2 + player - 1]; kl = raw[4 + player - 1]
        assert np.array_equal(o[6], raw[1] >= 4)
        assert np.array_equal(o[7], (raw[0] == 0) & (la == 0))
        assert np.array_equal(o[13], la >= 6)
        assert np.array_equal(o[20], kl >= 7)
        assert np.array_equal(o[25], raw[6] == 5)
        assert np.array_equal(o[26], raw[7 + player - 1] >= 1)
    with pytest.raises(IndexError):
        encode.preprocess(raw, 1, 10)


def test_symmetry_group_axioms():
    g = np.arange(361, dtype=np.int32).reshape(19, 19)
    images = [encode.transform(g, s) for s in range(8)]
    assert len({im.tobytes() for im in images}) == 8                       # 8 distinct images
    assert np.array_equal(images[0], g)                                    # id 0 == reference (identity)
    for s in range(8):
        inv = encode.inverse_sym(s)
        assert np.array_equal(encode.transform(images[s], inv), g)
        for x, y in ((0, 0), (3, 7), (18, 2)):
            tx, ty = encode.transform_xy(x, y, s)
            assert images[s][tx, ty] == g[x, y]                            # the stone at (x,y) moves to T(x,y)
            assert encode.transform_xy(tx, ty, inv) == (x, y)
    # closure
    table = {im.tobytes(): s for s, im in enumerate(images)}
    for a in range(8):
        for b in range(8):
            assert encode.transform(images[a], b).tobytes() in table


def test_symmetry_commutes_with_preprocess(golden):
    idx, raw, pl, rk, x, y = _sample(golden, 8, seed=2)
    for i in range(8):
        base = encode.preprocess(raw[i], int(pl[i]), int(rk[i]), 0)
        for s in range(8):
            assert np.array_equal(encode.preprocess(raw[i], int(pl[i]), int(rk[i]), s), encode.transform(base, s))
            lab = encode.label(int(x[i]), int(y[i]), s)
            assert encode.transform(base, s)[0].reshape(-1)[lab] == 1
