"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol include/deepgo.h declares,
the header is plain C that a no-preprocessor cdef parser (cffi here, LuaJIT ffi.cdef in the reference) accepts,
and host-only helpers / error paths behave.  No compute call is made without a GPU."""
import ctypes
import os

import numpy as np
import pytest

import __graft_entry__ as entry


@pytest.fixture(scope="module")
def dg():
    entry.build()
    import deepgo_b200
    return deepgo_b200


def test_every_declared_symbol_is_exported(dg):
    names = dg.declared_symbols()
    assert len(names) >= 30 and len(set(names)) == len(names)
    so = ctypes.CDLL(dg.LIB_PATH)
    for n in names:
        assert hasattr(so, n), n
        getattr(dg.lib(), n)


def test_header_is_plain_c(dg):
    text = dg.header_cdef()
    assert "#" not in text.replace("#include", "")
    import cffi
    f = cffi.FFI()
    f.cdef(text)          # would raise on macros / non-C constructs


def test_labels_helper_matches_reference_formula(dg):
    from deepgo_b200.host import engine
    x = np.array([1, 19, 4, 10], np.uint8); y = np.array([1, 19, 17, 3], np.uint8)
    lab = engine.encode_labels(x, y)
    assert list(lab) == [19 * (int(a) - 1) + int(b) - 1 for a, b in zip(x, y)]      # dataloader.lua:89, 0-based
    # symmetry 1 = transpose
    assert list(engine.encode_labels(x, y, np.ones(4, np.uint8))) == [19 * (int(b) - 1) + int(a) - 1 for a, b in zip(x, y)]
    with pytest.raises(dg.DeepGoError):
        engine.encode_labels(np.array([0], np.uint8), np.array([3], np.uint8))
    with pytest.raises(dg.DeepGoError):
        engine.encode_labels(x, y, np.full(4, 8, np.uint8))


def test_no_cpu_fallback(dg):
    from deepgo_b200.host import engine
    if engine.device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(dg.DeepGoError):
        engine.Context(0)


def test_oracle_is_not_imported_by_product():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dp, _, fs in os.walk(os.path.join(root, "deep-go_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".lua", ".h")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, f
