"""Dev tool: where does the host time of an end-to-end step go at N ranks? (torchrun, one rank per GPU)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from deepgo_b200.host import engine, dp
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
ctx = engine.Context(lr)
m = engine.Model(ctx, bench.L, bench.C, engine.DG_PREC_BF16, max_batch=bench.BATCH)
m.set_params(bench.he_params(2))
dp.init_data_parallel(ctx, m, dist, rank, world, bench.BATCH * world)
b = bench.synthetic_batch(100 + rank, bench.BATCH)
pin = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy() for k, v in b.items()}
for mode in ("loss", "noloss_sync", "noloss"):
    for _ in range(5):
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"]); m.train_step(0.01)
    ctx.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ts, tt = [], []
    t_all = time.perf_counter()
    for _ in range(50):
        t0 = time.perf_counter()
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"])
        t1 = time.perf_counter()
        if mode == "loss":
            m.train_step(0.01)
        else:
            m.train_step(0.01, want_loss=False)
            if mode == "noloss_sync":
                ctx.synchronize()
        t2 = time.perf_counter()
        ts.append(t1 - t0); tt.append(t2 - t1)
    ctx.synchronize()
    t_all = time.perf_counter() - t_all
    print("[rank %d] %-12s step %.3f ms | set_batch_raw median %.1f us max %.1f us | train_step median %.1f us max %.1f us" % (
        rank, mode, t_all / 50 * 1e3, np.median(ts) * 1e6, max(ts) * 1e6, np.median(tt) * 1e6, max(tt) * 1e6), flush=True)
print("[rank %d] cpus %s affinity %d" % (rank, os.cpu_count(), len(os.sched_getaffinity(0))))
dist.destroy_process_group()
