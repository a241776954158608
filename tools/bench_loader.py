"""Loader throughput on the host cores (no GPU): the reference's scheme -- one Torch7 file per position, read with the
reference-format reader -- against the packed corpus (dg_corpus_read_batch).  python tools/bench_loader.py [record_dir]
record_dir defaults to the reference's data/ when it exists (build container), else tests/golden/records."""
import glob
import json
import os
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

entry.build()
from deepgo_b200.host import data, dataloader

rec_dir = sys.argv[1] if len(sys.argv) > 1 else ("/root/reference/data" if os.path.isdir("/root/reference/data") else os.path.join(ROOT, "tests", "golden", "records"))
files = sorted(f for f in glob.glob(os.path.join(rec_dir, "**", "*"), recursive=True) if os.path.isfile(f) and os.path.basename(f).isdigit())
cores = os.cpu_count()
out = {"record_files": len(files), "host_cores": cores, "where": "build container" if rec_dir.startswith("/root/reference") else "repo fixtures"}

# (a) the reference's scheme: one file per position (dataloader.lua:123-125), serial and on a thread pool (data.lua:14-19)
reps = max(1, 4000 // len(files))
t0 = time.perf_counter()
for _ in range(reps):
    for f in files:
        dataloader.read_record(f)
dt = time.perf_counter() - t0
out["per_file_reader_1_thread"] = {"positions_per_s": reps * len(files) / dt}
with ThreadPoolExecutor(max_workers=cores) as ex:
    t0 = time.perf_counter()
    for _ in range(reps):
        list(ex.map(dataloader.read_record, files, chunksize=64))
    dt = time.perf_counter() - t0
out["per_file_reader_thread_pool"] = {"positions_per_s": reps * len(files) / dt, "threads": cores}

# (b) packed corpus
with tempfile.TemporaryDirectory() as td:
    t0 = time.perf_counter()
    path = data.Corpus.pack(files, os.path.join(td, "c.dgc"), threads=cores)
    out["pack"] = {"seconds": time.perf_counter() - t0, "bytes": os.path.getsize(path)}
    rng = np.random.default_rng(0)
    for threads in (1, cores):
        c = data.Corpus(path, threads=threads)
        for n in (128, 2048):
            idx = [rng.integers(0, c.count, n) for _ in range(64)]
            c.read(idx[0])
            t0 = time.perf_counter()
            k = 0
            while time.perf_counter() - t0 < 1.0:
                c.read(idx[k % 64]); k += 1
            dt = time.perf_counter() - t0
            out["corpus_batch%d_threads%d" % (n, threads)] = {"positions_per_s": k * n / dt, "GBps": k * n * 3254 / dt / 1e9}
print(json.dumps(out, indent=1))
