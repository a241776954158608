"""CPU: the two checkpoint readers (no GPU needed).

  * dg_checkpoint_read: the library's own format (written here by hand, byte for byte as csrc/checkpoint.cpp documents it);
  * dg_t7_read_model: a checkpoint as the REFERENCE writes it (experiments.lua:124-127: torch.save of the pickled experiment
    table).  The reference ships no such file and Torch7 cannot run here, so the file is produced by the small writer below,
    which follows the published Torch7 serialisation (torch/File.lua: type tags 0 nil, 1 number, 2 string, 3 table, 4 torch
    object, 5 boolean, 6/7/8 function; objects are memoised by index) -- UNPINNED like the record reader's format would be
    without the 4,398 reference files, which pin tags 1-5 and tensors/storages (tests/test_host_mirror.py)."""
import struct

import numpy as np
import pytest

import __graft_entry__ as entry


@pytest.fixture(scope="module")
def engine():
    entry.build()
    from deepgo_b200.host import engine
    return engine


class T7Writer:
    def __init__(self):
        self.out = bytearray()
        self.next_index = 1

    def i32(self, v): self.out += struct.pack("<i", v)
    def i64(self, v): self.out += struct.pack("<q", v)
    def f64(self, v): self.out += struct.pack("<d", v)
    def string(self, s): b = s.encode(); self.i32(len(b)); self.out += b

    def index(self):
        i = self.next_index
        self.next_index += 1
        return i

    def number(self, v): self.i32(1); self.f64(float(v))
    def str_(self, s): self.i32(2); self.string(s)
    def boolean(self, v): self.i32(5); self.i32(1 if v else 0)

    def table(self, items):
        """items: list of (key, writer-callable); keys are str or int"""
        self.i32(3); self.i32(self.index()); self.i32(len(items))
        for k, w in items:
            if isinstance(k, str):
                self.str_(k)
            else:
                self.number(k)
            w()

    def tensor(self, arr, cls="Float"):
        """a 1-D tensor over its own storage, with a storage offset of `lead` elements to exercise the offset handling"""
        lead = 3
        dt = np.float64 if cls == "Double" else np.float32
        st = np.concatenate([np.full(lead, 7.0, dt), np.asarray(arr, dt)])
        self.i32(4); self.i32(self.index()); self.string("V 1"); self.string("torch.%sTensor" % cls)
        self.i32(1); self.i64(len(arr)); self.i64(1); self.i64(lead + 1)
        self.i32(4); self.i32(self.index()); self.string("V 1"); self.string("torch.%sStorage" % cls)
        self.i64(len(st)); self.out += st.tobytes()

    def function(self, tag=8):
        """tag 8 (and legacy 7): memoised like tables; tag 6, the older form, carries no index"""
        self.i32(tag)
        if tag != 6:
            self.i32(self.index())
        dumped = b"\x1bLJ\x02fake bytecode"
        self.i32(len(dumped)); self.out += dumped
        self.table([])                                 # upvalues

    def module(self, cls, items):
        self.i32(4); self.i32(self.index()); self.string("V 1"); self.string(cls)
        self.table(items)


def write_reference_model(path, params, cls, rate, decay, iterations):
    w = T7Writer()
    w.table([
        ("name", lambda: w.str_("basicGoExperiment")),
        ("new", w.function),                                          # Experiment:new copies the methods into the instance
        ("init", lambda: w.function(6)), ("run", lambda: w.function(7)),
        ("useCuda", lambda: w.boolean(True)),
        ("kernels", lambda: w.table([(1, lambda: w.number(5)), (2, lambda: w.number(3)), (3, lambda: w.number(3))])),
        ("criterion", lambda: w.module("nn.ClassNLLCriterion", [("sizeAverage", lambda: w.boolean(True))])),
        ("model", lambda: w.module("nn.Sequential", [("modules", lambda: w.table([
            (1, lambda: w.module("nn.SpatialZeroPadding", [("pad_l", lambda: w.number(2))])),
            (2, lambda: w.module("nn.ReLU", [("val", lambda: w.number(0))]))]))])),
        ("optimizer", lambda: w.table([("rate", lambda: w.number(rate)), ("rate_decay", lambda: w.number(decay))])),
        ("modelParameters", lambda: w.tensor(params, cls)),
        ("iterations", lambda: w.number(iterations)),
        ("id", lambda: w.number(0.95223742560484)),
    ])
    with open(path, "wb") as f:
        f.write(bytes(w.out))


@pytest.mark.parametrize("cls", ["Float", "Cuda", "Double"])
def test_reference_model_reader(engine, tmp_path, cls):
    rng = np.random.default_rng(0)
    params = rng.normal(0, 1, 1234).astype(np.float32)
    fn = tmp_path / "0.95223742560484.model"
    write_reference_model(fn, params, cls, 0.0099, 1e-7, 4000)
    got, rate, decay, iterations = engine.read_reference_model(fn)
    assert np.array_equal(got, params)
    assert rate == 0.0099 and decay == 1e-7 and iterations == 4000


def test_reference_model_reader_rejects_other_files(engine, tmp_path):
    import deepgo_b200
    bad = tmp_path / "x.model"
    bad.write_bytes(b"not a torch file at all")
    with pytest.raises(deepgo_b200.DeepGoError):
        engine.read_reference_model(bad)
    w = T7Writer()
    w.table([("iterations", lambda: w.number(3))])                       # a table without modelParameters
    bad.write_bytes(bytes(w.out))
    with pytest.raises(deepgo_b200.DeepGoError, match="modelParameters"):
        engine.read_reference_model(bad)
    with pytest.raises(deepgo_b200.DeepGoError):
        engine.read_reference_model(tmp_path / "missing.model")


def test_native_checkpoint_reader(engine, tmp_path):
    import deepgo_b200
    params = np.arange(10, dtype=np.float32) * 0.5
    fn = tmp_path / "ck.model"
    fn.write_bytes(b"DGCKPT02\nnumLayers=3\nchannelSize=64\nkFirst=5\ninPlanes=37\nP=10\nrate=0.01\nname=basicGoExperiment\nend\n" + params.tobytes())
    meta, got = engine.read_checkpoint(fn)
    assert np.array_equal(got, params)
    assert meta["numLayers"] == "3" and meta["rate"] == "0.01" and meta["name"] == "basicGoExperiment" and meta["P"] == "10"
    fn.write_bytes(b"DGCKPT02\nnumLayers=3\nP=10\nend\n" + params.tobytes()[:-4])      # truncated
    with pytest.raises(deepgo_b200.DeepGoError, match="truncated"):
        engine.read_checkpoint(fn)
    fn.write_bytes(b"DGCKPT01\nP=10\nend\n" + params.tobytes())
    with pytest.raises(deepgo_b200.DeepGoError):
        engine.read_checkpoint(fn)
