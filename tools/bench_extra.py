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


# ---- c3: the paper's largest net, 12 x 192, batch 256: bf16 training throughput + bf16-vs-fp32 agreement -----------------
def he_params_lc(seed, L, C):
    rng = np.random.default_rng(seed)
    chans = [37] + [C] * (L - 1) + [1]
    ks = [5] + [3] * (L - 1)
    parts = []
    for i in range(L):
        fan = chans[i] * ks[i] * ks[i]
        parts += [rng.normal(0, np.sqrt(2.0 / fan), chans[i + 1] * fan), rng.normal(0, 0.05, chans[i + 1]), rng.normal(0, 0.05, chans[i + 1] * 361)]
    parts[-2] = parts[-2] + 1.0          # lift the last layer's bias so the final ReLU is not dead at init
    return np.concatenate(parts).astype(np.float32)


B3, L3, C3 = 256, 12, 192
p3 = he_params_lc(3, L3, C3)
b3 = bench.synthetic_batch(33, B3)
m3 = engine.Model(ctx, L3, C3, engine.DG_PREC_BF16, max_batch=B3)
m3.set_params(p3)
m3.set_batch_raw(b3["raw"], b3["player"], b3["rank"], b3["sym"], b3["x"], b3["y"])
lp_bf = m3.forward()
for _ in range(3):
    m3.encode_resident(); m3.train_step(0.0, want_loss=False)
ctx.synchronize()
ctx.timer_start()
for _ in range(10):
    m3.encode_resident(); m3.train_step(0.0, want_loss=False)
ms3 = ctx.timer_stop() / 10
F3 = 3 * (2 * 361 * (25 * 37 * C3 + 10 * 9 * C3 * C3 + 9 * C3)) - 2 * 361 * 25 * 37 * C3
m3f = engine.Model(ctx, L3, C3, engine.DG_PREC_FP32, max_batch=B3)
m3f.set_params(p3)
m3f.set_batch_raw(b3["raw"], b3["player"], b3["rank"], b3["sym"], b3["x"], b3["y"])
lp_f = m3f.forward()
out["c3_12x192_batch256"] = {"batch": B3, "ms_per_step": ms3, "positions_per_s": B3 / (ms3 * 1e-3), "tflops_algorithmic": B3 * F3 / (ms3 * 1e-3) / 1e12,
                             "bf16_vs_fp32_logprob_max_abs": float(np.abs(lp_bf - lp_f).max()),
                             "bf16_vs_fp32_top1_agreement": float((lp_bf.argmax(1) == lp_f.argmax(1)).mean()),
                             "note": "bf16 tensor-core path (per-layer launches: 192 channels need two N blocks) vs the fp32 CUDA-core path of this library, He init, rate 0"}
del m3, m3f

# ---- board analysis --------------------------------------------------------------------------------------------
g = np.load(os.path.join(bench.ROOT, "tests", "golden", "positions.npz"))
stones = np.ascontiguousarray(np.tile(g["raw"][:, 0], (4, 1, 1)))          # 17,592 boards
ctx.summarize_boards(stones[:64])
t0 = time.perf_counter(); got = ctx.summarize_boards(stones); dt_gpu = time.perf_counter() - t0
t0 = time.perf_counter(); ref = oboard.summarize(stones); dt_cpu = time.perf_counter() - t0
out["board_analysis"] = {"boards": len(stones), "gpu_boards_per_s": len(stones) / dt_gpu, "cpu_oracle_boards_per_s": len(stones) / dt_cpu,
                         "cpu_threads": os.cpu_count(), "bit_exact": bool(np.array_equal(got, ref)),
                         "note": "dg_summarize_boards host to host (wall clock) vs the C oracle with OpenMP over boards"}
# ---- SGF -> records (N4): the 22 fixture games re-transcribed from synthesised SGF text, host to host ---------------------
from deepgo_b200.host import makedata
LET = "abcdefghijklmnopqrs"
texts = []
for gi in range(len(g["games"])):
    sel = np.where(g["game"] == gi)[0]
    sel = sel[np.argsort(g["move"][sel])]
    texts.append("(;BR[%dd];WR[%dd]" % tuple(g["ranks"][sel[0]]) + "".join(
        ";%s[%s%s]" % ("BW"[g["player"][i] - 1], LET[g["x"][i] - 1], LET[g["y"][i] - 1]) for i in sel) + ";)")
texts = texts * 8                                                            # 176 games in lockstep
makedata.transcribe(texts[:2], ctx)
t0 = time.perf_counter(); rec = makedata.transcribe(texts, ctx); dt = time.perf_counter() - t0
out["sgf_transcribe"] = {"games": len(texts), "records": int(len(rec["raw"])), "records_per_s": len(rec["raw"]) / dt,
                         "note": "dg_sgf_parse + lockstep dg_update_boards (one call per move number) + one dg_summarize_boards; Python driver, wall clock"}
print(json.dumps(out, indent=1))
