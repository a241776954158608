"""BASELINE config 3: the paper's largest net (12 layers x 192 filters), "tf32 vs bf16 tolerance study".

For the weight sets of SURVEY.md 8(d) -- W0 Torch7-style uniform init, W1 He-scaled, W2 = W1 after 50 SGD steps at rate .01 (run
here with the library's fp32 path on S1 batches of 32, not with the fp64 oracle: 50 oracle steps of this net are minutes of CPU) --
every precision of the library (fp32 CUDA cores, tf32 and bf16 tensor cores) is compared with the un-rounded fp64 oracle on the
same 32 S1 positions: log-prob error, top-1 agreement, loss, per-tensor gradient error.  Then training throughput of the three
precisions at the config's own batch (256) and, for the roofline denominator, the TF32 GEMM rate cuBLAS reaches on this GPU
(torch.matmul 8192^3 with allow_tf32, best of 10 and a 2 s back-to-back loop -- the method of MEASURED_PEAKS.json).

  python tools/precision_study.py > gpurun_out/precision_study.json        (on a B200; the oracle runs on the host cores)"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine
from oracle import encode, net

L, C, NB = 12, 192, 32
PRECS = [("fp32", engine.DG_PREC_FP32), ("tf32", engine.DG_PREC_TF32), ("bf16", engine.DG_PREC_BF16)]
ctx = engine.Context(0)
g = np.load(os.path.join(bench.ROOT, "tests", "golden", "positions.npz"))


def s1(n, seed=0):
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(g["raw"]), n)
    pl = g["player"][idx]
    return dict(raw=g["raw"][idx], player=pl, rank=g["ranks"][idx, pl - 1], sym=rng.integers(0, 8, n).astype(np.uint8), x=g["x"][idx], y=g["y"][idx])


def rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


w0 = net.init_params(L, C, 2, "torch7")
w1 = net.init_params(L, C, 3, "he")
net.split_params(w1, L, C)[-1][1].add_(1.0)            # keeps the final ReLU (experiments.lua:147) from clamping every logit at init
m = engine.Model(ctx, L, C, engine.DG_PREC_FP32, max_batch=NB)
m.set_params(w1.numpy())
rate = 0.01
for step in range(50):
    b = s1(NB, 1000 + step)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    m.train_step(rate)
    rate *= 1 - 1e-7
w2 = torch.from_numpy(m.get_params()).double()
del m

out = {"net": "12 x 192, 5x5 first, 3x3 rest", "positions": NB, "sample": "S1 (seed 0) of the fixture corpus", "sets": {}}
b = s1(NB, 0)
x, t = encode.load_minibatch(b["raw"], b["player"], b["rank"], b["x"], b["y"], b["sym"])
for wname, flat in (("W0 torch7 init", w0), ("W1 He", w1), ("W2 He + 50 SGD steps", w2)):
    t0 = time.perf_counter()
    loss_r, g_r, logp_r = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C)
    oracle_s = time.perf_counter() - t0
    rows = {}
    for pname, prec in PRECS:
        m = engine.Model(ctx, L, C, prec, max_batch=NB)
        m.set_params(flat.numpy())
        m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
        logp = m.forward()
        m.zero_grads()
        loss = m.loss_backward()
        gg = torch.from_numpy(m.get_grads())
        errs = [(rel(W, Wr), rel(a, ar)) for (W, _, a), (Wr, _, ar) in zip(net.split_params(gg, L, C), net.split_params(g_r, L, C))]
        d = np.abs(logp - logp_r.numpy())
        rows[pname] = {"logprob_max_abs_err": float(d.max()), "logprob_mean_abs_err": float(d.mean()),
                       "top1_agreement": float((logp.argmax(1) == logp_r.numpy().argmax(1)).mean()), "loss_abs_err": abs(loss - loss_r),
                       "grad_W_rel_err_max": max(e[0] for e in errs), "grad_W_rel_err_by_layer": [round(e[0], 6) for e in errs],
                       "grad_a_rel_err_max": max(e[1] for e in errs)}
        del m
    out["sets"][wname] = {"oracle_loss": loss_r, "oracle_seconds": oracle_s, "precisions": rows}

# ---- throughput at the config's batch -------------------------------------------------------------------------------------------
B3 = 256
b3 = s1(B3, 33)
F3 = 3 * (2 * 361 * (25 * 37 * C + 10 * 9 * C * C + 9 * C)) - 2 * 361 * 25 * 37 * C
out["throughput_batch256"] = {}
for pname, prec in PRECS:
    m = engine.Model(ctx, L, C, prec, max_batch=B3)
    m.set_params(w1.numpy())
    m.set_batch_raw(b3["raw"], b3["player"], b3["rank"], b3["sym"], b3["x"], b3["y"])
    steps = 3 if pname == "fp32" else 10
    for _ in range(2):
        m.encode_resident(); m.train_step(0.0, want_loss=False)
    ctx.synchronize()
    ctx.timer_start()
    for _ in range(steps):
        m.encode_resident(); m.train_step(0.0, want_loss=False)
    ms = ctx.timer_stop() / steps
    out["throughput_batch256"][pname] = {"ms_per_step": ms, "positions_per_s": B3 / (ms * 1e-3), "tflops_algorithmic": B3 * F3 / (ms * 1e-3) / 1e12}
    del m

# ---- TF32 GEMM rate of this GPU (cuBLAS through torch), burst and sustained ---------------------------------------------------------
torch.backends.cuda.matmul.allow_tf32 = True
N = 8192
a = torch.randn(N, N, device="cuda"); bb = torch.randn(N, N, device="cuda")
for _ in range(3):
    a @ bb
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); a @ bb; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = int(2000 / best) + 1
e0.record()
for _ in range(reps):
    a @ bb
e1.record(); torch.cuda.synchronize()
out["tf32_peak"] = {"tf32_tflops": 2 * N ** 3 / (best * 1e-3) / 1e12, "tf32_tflops_sustained": 2 * N ** 3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12,
                    "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS): best of 10, and back to back for ~2 s"}
print(json.dumps(out, indent=1))
