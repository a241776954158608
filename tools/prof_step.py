"""Few c2 training steps through the C ABI, for ncu captures (no CPU baseline, no timing claims)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from deepgo_b200.host import engine

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
ctx = engine.Context(int(os.environ.get("LOCAL_RANK", "0")))
m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=bench.BATCH)
m.set_params(bench.he_params(2))
if world > 1:
    import torch.distributed as dist
    from deepgo_b200.host import dp
    dist.init_process_group("gloo")
    dp.init_data_parallel(ctx, m, dist, rank, world, bench.BATCH * world)
b = bench.synthetic_batch(100, bench.BATCH)
m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
for _ in range(steps):
    m.encode_resident()
    loss = m.train_step(0.01)
print("loss", loss, "launches", ctx.launch_count())

if os.environ.get("DG_TL"):
    # in-graph timeline of the last step (globaltimer ns relative to the first kernel of the step)
    from deepgo_b200 import ffi, lib
    names = ["conv0", "conv chain", "head", "dgrad chain", "wgrad 3x3", "wgrad first", "finalize", "sgd", "head reduce", "peer exchange", "peer exch. 2"]
    buf = ffi.new("unsigned long long[33]")
    lib().dg_prof_timeline(ctx.h, buf, 33)          # drop the stamps of the steps so far
    m.encode_resident()
    m.train_step(0.01)
    lib().dg_prof_timeline(ctx.h, buf, 33)
    v = [int(buf[i]) for i in range(33)]
    t0 = min(x for i, x in enumerate(v) if i % 3 == 0 and x != 2**64 - 1)
    for k, nm in enumerate(names):
        a, g, e = v[3 * k:3 * k + 3]
        if a == 2**64 - 1:
            continue
        print("[rank %d] %-13s in %8.1f us  go %8.1f us  out %8.1f us   (%.1f us)" % (rank, nm, (a - t0) / 1e3, (g - t0) / 1e3, (e - t0) / 1e3, (e - g) / 1e3))
