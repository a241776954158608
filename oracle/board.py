"""ctypes wrapper over oracle/board.c (test infrastructure; see oracle/__init__.py)."""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_board.so")
_lib = None


def build(force=False):
    """gcc oracle/board.c -> oracle/_build/liboracle_board.so (idempotent)."""
    src = os.path.join(_HERE, "board.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", _SO, src])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        u8p = ctypes.POINTER(ctypes.c_uint8)
        _lib.dg_oracle_summarize.argtypes = [u8p, u8p]
        _lib.dg_oracle_summarize_batch.argtypes = [u8p, u8p, ctypes.c_int]
        _lib.dg_oracle_update_board.argtypes = [u8p, u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
        _lib.dg_oracle_update_board.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))


def summarize(stones):
    """stones u8[n,19,19] (0/1/2) -> u8[n,7,19,19]: liberties, liberties_after[1,2], kills[1,2], ladders[1,2]
    (makedata.lua:143-153)."""
    stones = np.ascontiguousarray(stones, dtype=np.uint8).reshape(-1, 19, 19)
    out = np.zeros((stones.shape[0], 7, 19, 19), np.uint8)
    lib().dg_oracle_summarize_batch(_p(stones), _p(out), stones.shape[0])
    return out


def update_board(stones, age, x, y, player):
    """In-place update_board (makedata.lua:329-354); x, y 0-based. Returns 0 or -1 (occupied)."""
    assert stones.dtype == np.uint8 and age.dtype == np.uint8 and stones.flags.c_contiguous and age.flags.c_contiguous
    return lib().dg_oracle_update_board(_p(stones), _p(age), int(x), int(y), int(player))
