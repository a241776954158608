"""-m gpu, collected last on purpose (a run that stops at the first failure has then already run everything else).

dg_encode_noalloc serves callers that are not allowed to allocate GPU memory: a loader thread working beside a
training step (an allocation would synchronise the device and stall the step), or a process whose HBM already belongs
to another allocator.  It stages through a fixed array inside the library instead of a cudaMalloc'ed buffer, so the
checks here are: it returns exactly what dg_encode returns and what the CPU oracle computes (bit-exact: the planes
are 0/1 integers), on two calls in a row that share that array -- the second smaller than the first, so bytes left
over from the first call would show up if they were ever read -- and it refuses batches beyond its fixed capacity."""
import numpy as np
import pytest

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _unpack_bits(bits):
    """uint32[n,37,12] bit-planes (bit p%32 of word p/32 = plane value at point p) -> u8[n,37,19,19]"""
    n = bits.shape[0]
    b = np.unpackbits(bits.view(np.uint8).reshape(n, 37, 48), axis=-1, bitorder="little")
    return b[:, :, :361].reshape(n, 37, 19, 19)


def test_two_calls_in_a_row_match_dg_encode(golden):
    import deepgo_b200
    from deepgo_b200.host import engine
    ctx = U.ctx()
    first = U.sample(golden, 64, 11)             # fills the whole fixed staging area
    second = U.fuzz(9, 12)                       # smaller, different data, full byte range
    for b in (first, second):
        for fmt in (engine.DG_FMT_F32_NCHW, engine.DG_FMT_BITPLANES):
            ref = ctx.encode(b["raw"], b["player"], b["rank"], b["sym"], fmt=fmt)
            got = ctx.encode_noalloc(b["raw"], b["player"], b["rank"], b["sym"], fmt=fmt)
            assert got.dtype == ref.dtype and got.shape == ref.shape
            assert np.array_equal(got, ref)
            # ... and against the checker itself, not only against the sibling entry point
            want, _ = U.oracle_batch(b, np.float32)
            if fmt == engine.DG_FMT_F32_NCHW:
                assert np.array_equal(got, want)
            else:
                assert np.array_equal(_unpack_bits(got), want.astype(np.uint8))
    # no symmetry array (NULL) after calls that used one: the stale ids in the shared array must not be applied
    ref = ctx.encode(second["raw"], second["player"], second["rank"])
    assert np.array_equal(ctx.encode_noalloc(second["raw"], second["player"], second["rank"]), ref)
    # capacity is fixed: 65 positions are an error, not a silent truncation
    big = U.fuzz(65, 13)
    with pytest.raises(deepgo_b200.DeepGoError, match="at most 64"):
        ctx.encode_noalloc(big["raw"], big["player"], big["rank"], big["sym"])
    assert ctx.encode_noalloc(big["raw"][:0], big["player"][:0], big["rank"][:0]).shape == (0, 37, 19, 19)
