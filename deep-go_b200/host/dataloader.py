"""Host mirror of dataloader.lua: same constants and function names; the arithmetic runs in libdeepgo_b200.so.

  reference                                   here
  STONE..TOTAL, SIZE, INPUT.* (:6-27)         same names (1-based values kept, as the Lua globals)
  flatten_data(data) (:30-39)                 flatten_data(data)  -> record dict
  transform(grid) (:41-44)                    transform(grid, sym=0)   (identity for sym 0, D4 otherwise)
  preprocess(data) (:50-92)                   preprocess(data)    -> {"input": f32[37,19,19], "output": 19*(x-1)+y}
  empty_minibatch_io(size) (:94-111)          empty_minibatch_io(size)
  load_minibatch(filenames) (:113-121)        load_minibatch(filenames)
  load_and_preprocess(filename) (:123-125)    load_and_preprocess(filename)
`output` keeps the reference's 1-based class index; the C ABI takes 0-based targets (train.py subtracts 1).
"""
import threading

import numpy as np

import deepgo_b200 as _dg
from deepgo_b200 import ffi, check
from . import engine

STONE = 1
LIBERTIES = STONE + 3
LIBERTIES_AFTER = LIBERTIES + 4
KILL = LIBERTIES_AFTER + 7
AGE = KILL + 7
LADDER = AGE + 5
RANK = LADDER + 1
TOTAL = RANK + 9
SIZE = 19
input_dimensions = [TOTAL, SIZE, SIZE]
output_dimensions = []


class _Input:
    stone = 1
    liberties = 2
    liberties_after = 3
    kills = 5
    age = 7
    ladder = 8
    total = 9


INPUT = _Input()
_ctx = None
_ctx_lock = threading.Lock()


def context(device=0):
    """The device context the loader encodes on (created on first use).  A dg_ctx takes ONE caller thread at a time
    (include/deepgo.h): encode calls go through encode_lock()."""
    global _ctx
    with _ctx_lock:
        if _ctx is None:
            _ctx = engine.Context(device)
    return _ctx


_encode_lock = threading.Lock()


def encode_lock():
    return _encode_lock


def flatten_data(data):
    """Pack a board summary into the 9-plane byte record (dataloader.lua:30-39)."""
    result = np.zeros((INPUT.total, SIZE, SIZE), np.uint8)
    result[INPUT.stone - 1] = data["stones"]
    result[INPUT.liberties - 1] = data["liberties"]
    la = data.get("liberties_after", data.get("liberties_after_move"))
    result[INPUT.liberties_after - 1:INPUT.liberties_after + 1] = la
    result[INPUT.kills - 1:INPUT.kills + 1] = data["kills"]
    result[INPUT.age - 1] = data.get("age", data.get("ages"))
    if data.get("ladders") is not None:
        result[INPUT.ladder - 1:INPUT.ladder + 1] = data["ladders"]
    return {"input": result, "move": data["move"], "ranks": data["ranks"], "flat": True}


def transform(grid, sym=0):
    """Symmetry hook (dataloader.lua:41-44 is the identity = sym 0).  bits: 1 transpose, 2 flip x, 4 flip y."""
    g = np.asarray(grid)
    if sym & 1:
        g = np.swapaxes(g, -1, -2)
    if sym & 2:
        g = g[..., ::-1, :]
    if sym & 4:
        g = g[..., :, ::-1]
    return np.ascontiguousarray(g)


def read_record(filename):
    """torch.load(filename) of one record without Torch7 -> record dict (dataloader.lua:124)."""
    raw = np.empty((9, SIZE, SIZE), np.uint8)
    x, y, p = ffi.new("int *"), ffi.new("int *"), ffi.new("int *")
    r = ffi.new("int[2]")
    check(_dg.lib().dg_t7_read_record(str(filename).encode(), ffi.cast("uint8_t *", raw.ctypes.data), x, y, p, r))
    return {"input": raw, "move": {"x": x[0], "y": y[0], "player": p[0]}, "ranks": {1: r[0], 2: r[1]}, "flat": True}


def _rank_of(data):
    ranks, player = data["ranks"], data["move"]["player"]
    return ranks[player] if isinstance(ranks, dict) else ranks[player - 1]


def preprocess(data, sym=0):
    """One record -> 37 planes + label (dataloader.lua:50-92), encoded on the GPU."""
    mv = data["move"]
    with _encode_lock:
        planes = context().encode(data["input"][None], [mv["player"]], [_rank_of(data)], [sym])[0]
    label0 = engine.encode_labels([mv["x"]], [mv["y"]], [sym])[0]
    return {"input": planes, "output": int(label0) + 1}


def empty_minibatch_io(size):
    return np.zeros([size] + input_dimensions, np.float32), np.zeros([size] + output_dimensions, np.float64)


def read_records(filenames):
    """The host-only half of load_minibatch: torch.load of every record (dg_t7_read_record is reentrant, so this is what
    the loader threads of data.lua:14-19 run)."""
    return [read_record(f) for f in filenames]


def encode_records(recs, sym=None):
    """The device half of load_minibatch: ONE encode call for the batch.  The loader's dg_ctx has a single stream and a
    single staging buffer, so concurrent callers are serialised here."""
    inputs, outputs = empty_minibatch_io(len(recs))
    if recs:
        raw = np.stack([r["input"] for r in recs])
        pl = np.array([r["move"]["player"] for r in recs], np.uint8)
        rk = np.array([_rank_of(r) for r in recs], np.uint8)
        xs = np.array([r["move"]["x"] for r in recs], np.uint8)
        ys = np.array([r["move"]["y"] for r in recs], np.uint8)
        with _encode_lock:
            inputs = context().encode(raw, pl, rk, sym)
        outputs = engine.encode_labels(xs, ys, sym).astype(np.float64) + 1
    return {"input": inputs, "output": outputs, "records": recs}


def load_minibatch(filenames, sym=None):
    """dataloader.lua:113-121: read + encode a list of record files -> {"input": [n,37,19,19], "output": [n] (1-based)}."""
    return encode_records(read_records(filenames), sym)


def load_and_preprocess(filename):
    return preprocess(read_record(filename))
