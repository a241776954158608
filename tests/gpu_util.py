"""Shared helpers for the -m gpu parity tests (oracle = checker, C ABI = the thing under test)."""
import os

import numpy as np
import torch

import __graft_entry__ as entry

_ctx = None


def ctx():
    global _ctx
    if _ctx is None:
        entry.build()
        from deepgo_b200.host import engine
        _ctx = engine.Context(0)
    return _ctx


def sample(golden, n, seed, with_sym=True):
    """S1 of SURVEY.md 8(d): seeded draw with replacement from the fixture corpus + random symmetry ids."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(golden["raw"]), n)
    pl = golden["player"][idx]
    sym = rng.integers(0, 8, n).astype(np.uint8) if with_sym else np.zeros(n, np.uint8)
    return dict(raw=golden["raw"][idx], player=pl, rank=golden["ranks"][idx, pl - 1], x=golden["x"][idx], y=golden["y"][idx], sym=sym)


def fuzz(n, seed):
    """S2: iid bytes over the full 0..255 range (threshold / wrap coverage); stones in {0,1,2}."""
    rng = np.random.default_rng(seed)
    raw = rng.integers(0, 256, (n, 9, 19, 19)).astype(np.uint8)
    raw[:, 0] = rng.choice([0, 1, 2], (n, 19, 19), p=[.6, .2, .2])
    return dict(raw=raw, player=rng.integers(1, 3, n).astype(np.uint8), rank=rng.integers(1, 10, n).astype(np.uint8),
                x=rng.integers(1, 20, n).astype(np.uint8), y=rng.integers(1, 20, n).astype(np.uint8),
                sym=rng.integers(0, 8, n).astype(np.uint8))


def oracle_batch(b, dtype=np.float64):
    from oracle import encode
    return encode.load_minibatch(b["raw"], b["player"], b["rank"], b["x"], b["y"], b["sym"], dtype)


def rel(a, b):
    a = np.asarray(a, np.float64).ravel(); b = np.asarray(b, np.float64).ravel()
    # tensors whose reference is (numerically) zero -- e.g. the last conv bias gradient when no logit is clamped,
    # sum(softmax - onehot) = 0 -- are compared absolutely
    nb = np.linalg.norm(b)
    if nb < 1e-9 * np.sqrt(b.size):
        return float(np.linalg.norm(a - b))
    return float(np.linalg.norm(a - b) / nb)
