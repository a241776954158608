"""Few c2 training steps through the C ABI, for ncu captures (no CPU baseline, no timing claims)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
ctx = engine.Context(0)
m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=bench.BATCH)
m.set_params(bench.he_params(2))
b = bench.synthetic_batch(100, bench.BATCH)
m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
for _ in range(steps):
    m.encode_resident()
    loss = m.train_step(0.01)
print("loss", loss, "launches", ctx.launch_count())
