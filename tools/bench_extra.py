"""Secondary measurements (not the bench.py headline): c5 inference, encode and SGD bandwidth, board analysis.
Prints one JSON object; run on a B200:  python tools/bench_extra.py > gpurun_out/extra.json"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine
from oracle import board as oboard

HBM = bench.peaks()[2]
out = {}
ctx = engine.Context(0)

# ---- c5: inference with on-GPU encode + 8-fold symmetry averaging, batch 4096, 12 x 128 -------------------------
B5 = 4096
m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=B5)
m.set_params(bench.he_params(2))
b = bench.synthetic_batch(5, B5)
m.set_batch_raw(b["raw"], b["player"], b["rank"])
m.predict_sym8(want_probs=False)
ctx.synchronize()
t0 = time.perf_counter()
reps = 3
for _ in range(reps):
    m.set_batch_raw(b["raw"], b["player"], b["rank"])       # host records in, argmax out: end to end
    _, am = m.predict_sym8(want_probs=False)
dt = (time.perf_counter() - t0) / reps
F_FWD = 1.1509e9
out["c5_inference_sym8"] = {"batch": B5, "positions_per_s": B5 / dt, "ms_per_batch": dt * 1e3,
                            "forward_evals_per_s": 8 * B5 / dt, "tflops_algorithmic": 8 * B5 * F_FWD / dt / 1e12,
                            "note": "host records -> H2D -> 8 x (encode, forward) -> average -> argmax -> D2H, wall clock"}

# ---- encode bandwidth (resident records -> blocked bf16 image) -------------------------------------------------
ctx.prof_begin()
for _ in range(10):
    m.encode_resident()
pe = ctx.prof_end()["encode"]
enc_ms = pe[0] / pe[1]
bytes_enc = B5 * (9 * 361 + 3 + 6 * 361 * 16)
out["encode_blocked_bf16"] = {"batch": B5, "ms": enc_ms, "positions_per_s": B5 / (enc_ms * 1e-3), "algorithmic_bytes": bytes_enc,
                              "GBps": bytes_enc / (enc_ms * 1e-3) / 1e9, "frac_of_hbm_peak": bytes_enc / (enc_ms * 1e-3) / 1e9 / HBM}
raw = b["raw"]; pl = b["player"]; rk = b["rank"]
t0 = time.perf_counter(); planes = ctx.encode(raw, pl, rk); dt = time.perf_counter() - t0
out["encode_f32_host_to_host"] = {"batch": B5, "positions_per_s": B5 / dt, "note": "dg_encode: H2D + kernel + D2H of 53 KB/position, wall clock"}

# ---- SGD bandwidth ---------------------------------------------------------------------------------------------
m2 = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=128)
m2.set_params(bench.he_params(2))
ctx.prof_begin()
for _ in range(20):
    m2.sgd_step(0.0)
ps = ctx.prof_end()
sgd_ms = ps["sgd"][0] / ps["sgd"][1]
out["sgd"] = {"params": m2.P, "ms": sgd_ms, "algorithmic_bytes": 12 * m2.P, "GBps": 12 * m2.P / (sgd_ms * 1e-3) / 1e9,
              "frac_of_hbm_peak": 12 * m2.P / (sgd_ms * 1e-3) / 1e9 / HBM, "note": "8.4 MB vectors: L2-resident, launch-latency dominated"}

# ---- board analysis --------------------------------------------------------------------------------------------
g = np.load(os.path.join(bench.ROOT, "tests", "golden", "positions.npz"))
stones = np.ascontiguousarray(np.tile(g["raw"][:, 0], (4, 1, 1)))          # 17,592 boards
ctx.summarize_boards(stones[:64])
t0 = time.perf_counter(); got = ctx.summarize_boards(stones); dt_gpu = time.perf_counter() - t0
t0 = time.perf_counter(); ref = oboard.summarize(stones); dt_cpu = time.perf_counter() - t0
out["board_analysis"] = {"boards": len(stones), "gpu_boards_per_s": len(stones) / dt_gpu, "cpu_oracle_boards_per_s": len(stones) / dt_cpu,
                         "cpu_threads": os.cpu_count(), "bit_exact": bool(np.array_equal(got, ref)),
                         "note": "dg_summarize_boards host to host (wall clock) vs the C oracle with OpenMP over boards"}
print(json.dumps(out, indent=1))
