"""-m gpu: feature-plane encoder through the C ABI vs the oracle; bit-exact (dataloader.lua:50-92)."""
import numpy as np
import pytest

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _unpack_bits(bits):
    n = bits.shape[0]
    b = np.unpackbits(bits.view(np.uint8).reshape(n, 37, 48), axis=-1, bitorder="little")
    return b[:, :, :361].reshape(n, 37, 19, 19)


def test_all_fixtures_sym0_bit_exact(golden):
    from deepgo_b200.host import engine
    from oracle import encode
    pl = golden["player"][:]
    rk = golden["ranks"][np.arange(len(pl)), pl - 1]
    got = U.ctx().encode(golden["raw"], pl, rk)
    bits = U.ctx().encode(golden["raw"], pl, rk, fmt=engine.DG_FMT_BITPLANES)
    assert got.dtype == np.float32 and got.shape == (4398, 37, 19, 19)
    ub = _unpack_bits(bits)
    for i in range(0, 4398, 7):
        ref = encode.preprocess(golden["raw"][i], int(pl[i]), int(rk[i]))
        assert np.array_equal(got[i], ref.astype(np.float32)), i
    assert np.array_equal(ub, got.astype(np.uint8))
    labels = engine.encode_labels(golden["x"], golden["y"])
    assert np.array_equal(labels, 19 * (golden["x"].astype(np.int32) - 1) + golden["y"].astype(np.int32) - 1)


def test_eight_symmetries_bit_exact(golden):
    from deepgo_b200.host import engine
    b = U.sample(golden, 64, 0)
    for s in range(8):
        b["sym"] = np.full(64, s, np.uint8)
        ref, lab = U.oracle_batch(b, np.float32)
        got = U.ctx().encode(b["raw"], b["player"], b["rank"], b["sym"])
        assert np.array_equal(got, ref), s
        assert np.array_equal(engine.encode_labels(b["x"], b["y"], b["sym"]), lab)


def test_fuzzed_bytes_bit_exact():
    from deepgo_b200.host import engine
    b = U.fuzz(257, 1)
    ref, lab = U.oracle_batch(b, np.float32)
    got = U.ctx().encode(b["raw"], b["player"], b["rank"], b["sym"])
    assert np.array_equal(got, ref)
    bits = U.ctx().encode(b["raw"], b["player"], b["rank"], b["sym"], fmt=engine.DG_FMT_BITPLANES)
    assert np.array_equal(_unpack_bits(bits), ref.astype(np.uint8))


def test_edge_cases():
    import deepgo_b200
    b = U.fuzz(3, 2)
    out = U.ctx().encode(b["raw"][:0], b["player"][:0], b["rank"][:0])           # empty batch
    assert out.shape == (0, 37, 19, 19)
    bad = b["rank"].copy(); bad[1] = 10
    with pytest.raises(deepgo_b200.DeepGoError):                                    # dataloader.lua:87 would index out of range
        U.ctx().encode(b["raw"], b["player"], bad)
    badp = b["player"].copy(); badp[0] = 3
    with pytest.raises(deepgo_b200.DeepGoError):
        U.ctx().encode(b["raw"], badp, b["rank"])
