"""Dev tool: per-kernel-kind device time of a TF32 training step (c3 by default: 12 x 192, batch 256)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from deepgo_b200.host import engine
L, C, B = int(os.environ.get("T_L", 12)), int(os.environ.get("T_C", 192)), int(os.environ.get("T_B", 256))
ctx = engine.Context(0)
rng = np.random.default_rng(0)
m = engine.Model(ctx, L, C, engine.DG_PREC_TF32, max_batch=B)
m.set_params(rng.normal(0, 0.03, m.P).astype(np.float32))
b = bench.synthetic_batch(1, B)
m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
for _ in range(2):
    m.train_step(0.0, want_loss=False)
ctx.synchronize()
ctx.timer_start()
for _ in range(5):
    m.train_step(0.0, want_loss=False)
ms = ctx.timer_stop() / 5
ctx.prof_begin()
for _ in range(3):
    m.train_step(0.0, want_loss=False)
p = ctx.prof_end()
print("tf32 %dx%d batch %d: %.3f ms/step, %.0f positions/s" % (L, C, B, ms, B / ms * 1e3))
for k, (t, n) in sorted(p.items(), key=lambda kv: -kv[1][0]):
    print("  %-14s %8.3f ms/step  %5.1f launches/step" % (k, t / 3, n / 3))
