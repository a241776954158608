"""The reference's numGPUs option at scale: ONE process, one model handle over N devices (dg_ctx_create_multi), 128 positions per
device, device-resident records.   python tools/bench_multi.py [N ...]   -> one JSON object per N"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine

ns = [int(a) for a in sys.argv[1:]] or [engine.device_count()]
for n in ns:
    ctx = engine.Context(list(range(n))) if n > 1 else engine.Context(0)
    B = bench.BATCH * n
    m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=B)
    m.set_params(bench.he_params(2))
    b = bench.synthetic_batch(100, B)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    for _ in range(10):
        m.encode_resident(); m.train_step(0.01, want_loss=False)
    ctx.synchronize()
    steps = 200
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(steps):
        m.encode_resident(); m.train_step(0.01, want_loss=False)
    ms = ctx.timer_stop() / steps
    wall = (time.perf_counter() - t0) / steps * 1e3
    pin = {k: np.ascontiguousarray(v) for k, v in b.items()}
    for _ in range(3):
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"]); m.train_step(0.01)
    t0 = time.perf_counter()
    for _ in range(50):
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"]); loss = m.train_step(0.01)
    e2e = (time.perf_counter() - t0) / 50 * 1e3
    print(json.dumps({"devices": n, "global_batch": B, "ms_per_step_device": ms, "ms_per_step_wall": wall, "positions_per_s": B / (ms * 1e-3),
                      "e2e_ms_per_step": e2e, "e2e_positions_per_s": B / (e2e * 1e-3), "loss": loss,
                      "note": "one process, one caller thread; device time = slowest device's event pair around 200 back-to-back steps"}), flush=True)
    del m, ctx
