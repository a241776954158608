"""Host mirror of data.lua: Dataset index (:53-80), random (game, move) sampling (:29-37) and the
queue_on_minibatch / do_queued_tasks pair (:82-96).  The pool threads only READ the record files (the record reader of the C
library is reentrant); the batch is encoded on the device and the callback runs on the caller's thread inside
do_queued_tasks, as `threads` does -- the loader's device context is never entered by two threads at once."""
import os
import random
from concurrent.futures import ThreadPoolExecutor

from . import dataloader

_pool = None
_queued = []


def prepare_data_loaders(number_loaders=32):
    global _pool
    _pool = ThreadPoolExecutor(max_workers=number_loaders) if number_loaders and number_loaders > 1 else None


prepare_data_loaders()


class Dataset:
    root = "/home/ubuntu/ebs_disk"
    directory = "train"

    def __init__(self, **kw):
        self.initialized = False
        self.games = []
        for k, v in kw.items():
            setattr(self, k, v)

    @classmethod
    def new(cls, d=None):
        return cls(**(d or {}))

    def init(self):
        """Read <root>/<directory>_game_counts.txt: `<path>\\t<count>`; games without moves are dropped (:74)."""
        self.games = []
        path = os.path.join(self.root, self.directory)
        file = path + "_game_counts.txt"
        if not os.path.exists(file):
            raise IOError("unable to find counts (probably bad dataset path): " + file)
        with open(file) as f:
            for line in f:
                r = line.rstrip("\n").split("\t")
                size = int(r[1])
                if size > 0:
                    self.games.append({"name": r[0], "size": size})
        self.initialized = True

    def generate_random_filename(self):
        """uniform game, then uniform move (not uniform over positions), data.lua:29-37"""
        if not self.initialized:
            self.init()
        game = random.choice(self.games)
        return "%s/%d" % (game["name"], random.randint(1, game["size"]))


def queue_on_minibatch(f, dataset, n, base=None):
    """Draw n random filenames, load them on a worker, call f(minibatch) on the caller's thread later."""
    if not dataset.initialized:
        dataset.init()
    root = base if base is not None else getattr(dataset, "base", "")
    filenames = [os.path.join(root, dataset.generate_random_filename()) for _ in range(n)]
    job = _pool.submit(dataloader.read_records, filenames) if _pool else None
    _queued.append((job, filenames, f))


def do_queued_tasks():
    while _queued:
        job, filenames, f = _queued.pop(0)
        recs = job.result() if job is not None else dataloader.read_records(filenames)
        f(dataloader.encode_records(recs))


# ---- packed corpus: the same positions, one file, native loader --------------------------------------
class Corpus:
    """One packed file instead of one Torch7 file per position (dg_corpus_* in include/deepgo.h).  `raw_minibatch(n)` returns what
    Model.set_batch_raw takes; positions are drawn uniformly (the reference draws a uniform game, then a uniform move of it,
    data.lua:29-37 -- pass `indices` to reproduce any other sampling)."""

    def __init__(self, path, threads=8):
        from deepgo_b200 import ffi, lib, check
        self._ffi, self._lib, self._check = ffi, lib(), check
        h = ffi.new("dg_corpus **")
        check(self._lib.dg_corpus_open(str(path).encode(), h))
        self.h = ffi.gc(h[0], self._lib.dg_corpus_close)
        n = ffi.new("long long *")
        check(self._lib.dg_corpus_count(self.h, n))
        self.count = int(n[0])
        self.threads = threads

    @staticmethod
    def pack(filenames, out_path, threads=8):
        """Pack Torch7 record files (dataloader.lua:123-125 reads them one by one) into one corpus file."""
        from deepgo_b200 import ffi, lib, check
        keep = [ffi.new("char[]", str(f).encode()) for f in filenames]
        arr = ffi.new("const char *[]", keep)
        check(lib().dg_corpus_pack(arr, len(keep), str(out_path).encode(), threads))
        return out_path

    def read(self, indices, want_moves=True):
        import numpy as np
        idx = np.ascontiguousarray(indices, dtype=np.int64)
        n = len(idx)
        raw = np.empty((n, 9, 19, 19), np.uint8)
        pl, rk = np.empty(n, np.uint8), np.empty(n, np.uint8)
        x, y = (np.empty(n, np.uint8), np.empty(n, np.uint8)) if want_moves else (None, None)
        c = lambda a: self._ffi.cast("uint8_t *", a.ctypes.data) if a is not None else self._ffi.NULL
        self._check(self._lib.dg_corpus_read_batch(self.h, self._ffi.cast("const long long *", idx.ctypes.data), n, c(raw), c(pl), c(rk),
                                                   c(x), c(y), self.threads))
        return {"raw": raw, "player": pl, "rank": rk, "x": x, "y": y}

    def raw_minibatch(self, n, rng=None):
        import numpy as np
        rng = rng or np.random.default_rng()
        return self.read(rng.integers(0, self.count, n))


def write_corpus_from_arrays(path, raw, x, y, player, ranks):
    """The packed format written from arrays (fixtures, synthetic corpora): 64-byte header + 3,264-byte records
    (deep-go_b200/csrc/corpus.cpp)."""
    import numpy as np
    n = len(raw)
    rec = np.zeros((n, 3264), np.uint8)
    rec[:, :3249] = np.asarray(raw, np.uint8).reshape(n, 3249)
    rec[:, 3249] = x; rec[:, 3250] = y; rec[:, 3251] = player
    rec[:, 3252:3254] = np.asarray(ranks, np.uint8).reshape(n, 2)
    hdr = np.zeros(64, np.uint8)
    hdr[:8] = np.frombuffer(b"DGCORP01", np.uint8)
    hdr[8:16] = np.frombuffer(np.uint64(n).tobytes(), np.uint8)
    hdr[16:20] = np.frombuffer(np.uint32(3264).tobytes(), np.uint8)
    with open(path, "wb") as f:
        f.write(hdr.tobytes()); f.write(rec.tobytes())
    return path
