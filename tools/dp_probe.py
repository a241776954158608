"""Dev tool: time the library's gradient allreduce alone (torchrun, one rank per GPU): 8.4 MB fp32 over NCCL."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from deepgo_b200.host import engine, dp

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
ctx = engine.Context(int(os.environ.get("LOCAL_RANK", 0)))
m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=bench.BATCH)
m.set_params(bench.he_params(2))
dp.init_data_parallel(ctx, m, dist, rank, world, bench.BATCH * world)
for _ in range(5):
    m.allreduce_grads()
ctx.synchronize()
dist.barrier()
ts = []
for _ in range(30):
    ctx.timer_start()
    m.allreduce_grads()
    ts.append(ctx.timer_stop())
ts.sort()
if rank == 0:
    print("allreduce of %.1f MB over %d ranks: median %.1f us, min %.1f us  (NCCL_ALGO=%s NCCL_PROTO=%s)" % (
        m.P * 4 / 1e6, world, ts[len(ts) // 2] * 1e3, ts[0] * 1e3, os.environ.get("NCCL_ALGO"), os.environ.get("NCCL_PROTO")))
dist.destroy_process_group()
