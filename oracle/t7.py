"""Generic reader for Torch7 binary serialisation (``torch.save``), little-endian.

Test infrastructure (see oracle/__init__.py).  The wire format is upstream Torch7
(torch/File.lua, un-vendored); the layout below was verified by parsing all 4,398
records under the reference's ``data/``.  The records themselves are
what ``makedata.lua:554`` wrote with ``torch.save(target, flatten_data(board))`` and what
``dataloader.lua:123-125`` reads back with ``torch.load``.

Object := int32 tag
  0 nil | 1 number(f64) | 2 string(int32 len, bytes) | 3 table(int32 ref, int32 n, n*(key,value))
  4 torch object(int32 ref, string version "V 1", string class, payload) | 5 boolean(int32)
Tensor payload  := int32 ndim, int64 size[ndim], int64 stride[ndim], int64 storageOffset(1-based), Storage object
Storage payload := int64 n, raw elements
"""
import struct
import numpy as np

_STORAGE_DTYPES = {
    "torch.ByteStorage": np.uint8, "torch.CharStorage": np.int8, "torch.ShortStorage": np.int16,
    "torch.IntStorage": np.int32, "torch.LongStorage": np.int64, "torch.FloatStorage": np.float32,
    "torch.DoubleStorage": np.float64,
}


class _Reader:
    def __init__(self, buf):
        self.b = buf
        self.p = 0
        self.memo = {}

    def i32(self):
        v = struct.unpack_from("<i", self.b, self.p)[0]
        self.p += 4
        return v

    def i64(self):
        v = struct.unpack_from("<q", self.b, self.p)[0]
        self.p += 8
        return v

    def f64(self):
        v = struct.unpack_from("<d", self.b, self.p)[0]
        self.p += 8
        return v

    def string(self):
        n = self.i32()
        s = self.b[self.p:self.p + n].decode("latin-1")
        self.p += n
        return s

    def obj(self):
        tag = self.i32()
        if tag == 0:
            return None
        if tag == 1:
            v = self.f64()
            return int(v) if v == int(v) else v
        if tag == 2:
            return self.string()
        if tag == 5:
            return self.i32() != 0
        if tag == 3:
            ref = self.i32()
            if ref in self.memo:
                return self.memo[ref]
            n = self.i32()
            t = {}
            self.memo[ref] = t
            for _ in range(n):
                k = self.obj()
                t[k] = self.obj()
            return t
        if tag == 4:
            ref = self.i32()
            if ref in self.memo:
                return self.memo[ref]
            version = self.string()
            cls = self.string() if version.startswith("V ") else version
            if cls.endswith("Tensor"):
                nd = self.i32()
                size = [self.i64() for _ in range(nd)]
                stride = [self.i64() for _ in range(nd)]
                off = self.i64() - 1
                storage = self.obj()
                if storage is None or nd == 0:
                    arr = np.zeros(size, dtype=np.float64)
                else:
                    it = storage.dtype.itemsize
                    arr = np.lib.stride_tricks.as_strided(
                        storage[off:], shape=size, strides=[s * it for s in stride]).copy()
                self.memo[ref] = arr
                return arr
            if cls in _STORAGE_DTYPES:
                n = self.i64()
                dt = np.dtype(_STORAGE_DTYPES[cls]).newbyteorder("<")
                arr = np.frombuffer(self.b, dtype=dt, count=n, offset=self.p).copy()
                self.p += n * dt.itemsize
                self.memo[ref] = arr
                return arr
            raise ValueError("unsupported torch class %r" % cls)
        raise ValueError("bad tag %d at %d" % (tag, self.p - 4))


def load(path):
    """Parse one ``torch.save`` file and return plain python / numpy objects."""
    with open(path, "rb") as f:
        return _Reader(f.read()).obj()


def load_record(path):
    """Read one deep-go position record -> (raw u8[9,19,19], x, y, player, (rank1, rank2)).

    Record = flatten_data() output (dataloader.lua:30-39): {input, move={x,y,player}, ranks, flat}.
    x, y, player and ranks keep the reference's 1-based values.
    """
    d = load(path)
    raw = np.ascontiguousarray(d["input"], dtype=np.uint8)
    assert raw.shape == (9, 19, 19), raw.shape
    mv = d["move"]
    rk = d["ranks"]
    return raw, int(mv["x"]), int(mv["y"]), int(mv["player"]), (int(rk[1]), int(rk[2]))
