"""deep-go_b200: B200-native hot path of vipmath/deep-go behind a plain-C ABI (include/deepgo.h).

This module only binds libdeepgo_b200.so (cffi, ABI mode -- the same header text LuaJIT's ffi.cdef loads).
There is no CPU fallback: if the shared library is missing the import of `lib` raises.
"""
import os as _os

import cffi as _cffi

_PKG = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "deep-go_b200") \
    if not _os.path.basename(_os.path.dirname(_os.path.abspath(__file__))) == "deep-go_b200" \
    else _os.path.dirname(_os.path.abspath(__file__))
ROOT = _os.path.dirname(_PKG)
HEADER = _os.path.join(ROOT, "include", "deepgo.h")
LIB_PATH = _os.environ.get("DEEPGO_LIB") or _os.path.join(_PKG, "libdeepgo_b200.so")      # DEEPGO_LIB: as in lua/deepgo.lua

DG_PREC_FP32, DG_PREC_BF16, DG_PREC_TF32 = 0, 1, 2
DG_FMT_F32_NCHW, DG_FMT_BITPLANES = 0, 1


def header_cdef():
    """include/deepgo.h minus preprocessor lines and the extern "C" braces (cffi/LuaJIT cdef take plain C)."""
    out = []
    for line in open(HEADER):
        s = line.strip()
        if s.startswith("#") or s.startswith('extern "C"') or s == "}":
            continue
        out.append(line)
    return "".join(out)


ffi = _cffi.FFI()
ffi.cdef(header_cdef())
_lib = None


class DeepGoError(RuntimeError):
    pass


def lib():
    """dlopen libdeepgo_b200.so (built by __graft_entry__.build / deep-go_b200/build.py)."""
    global _lib
    if _lib is None:
        if not _os.path.exists(LIB_PATH):
            raise DeepGoError("libdeepgo_b200.so is not built (%s); run `python deep-go_b200/build.py`. "
                              "There is no CPU fallback." % LIB_PATH)
        _lib = ffi.dlopen(LIB_PATH)
    return _lib


def check(rc):
    if rc != 0:
        raise DeepGoError(ffi.string(lib().dg_last_error()).decode())


def declared_symbols():
    """Names of every function include/deepgo.h declares."""
    import re
    text = re.sub(r"/\*.*?\*/", "", header_cdef(), flags=re.S)
    return re.findall(r"\b(dg_[a-z0-9_]+)\s*\(", text)
