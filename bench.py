#!/usr/bin/env python
"""bench.py -- training positions/sec (fwd+bwd+SGD) of the deep-go policy net on N B200s.

Workload (BASELINE.json configs[1]): 12 conv layers x 128 filters (5x5 first, 3x3 rest, 361-way softmax),
batch 128 per GPU, synthetic positions drawn (seeded) from the fixture corpus, He-scaled random weights.
A step = on-device feature-plane encode of the resident records + zero grads + forward + NLL + backward
+ [gradient allreduce] + SGD update + bf16 operand repack.  One JSON line on stdout (rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
N > 1: launched by torch.distributed.run, one process per GPU (weak scaling: 128 positions per GPU).
--impl reference: the reference's CPU path (oracle restatement of the Torch7 nn path; Torch7 itself cannot run
on this image) timed on the host cores, same metric and config.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L, C, BATCH = 12, 128, 128
METRIC = "training positions/sec (fwd+bwd+SGD)"
UNIT = "positions/s"
F_TRAIN = 3.3674e9          # algorithmic FLOPs / position, SURVEY.md 8(d): 3*F_fwd - F_fwd(layer 1)
F_CONV_LAYER = 2 * 361 * 9 * C * C          # one 3x3 C->C layer, per position (106.5 MFLOP)


def config(n_gpus, impl="ours"):
    c = {"workload": "c2: policy net 12 layers x 128 filters (5x5 first, 3x3 rest), batch 128/GPU, 19x19x37 planes",
         "num_layers": L, "channel_size": C, "batch_per_gpu": BATCH, "global_batch": BATCH * n_gpus,
         "parallelism": "dp%d" % n_gpus, "precision": "bf16 operands, fp32 accumulate, fp32 master weights",
         "l2": "no flush: the timed steps run back to back and the data a step touches (stored activations and gradient images of 128 "
               "positions, ~0.3 GB, plus weights) exceeds the 126 MB L2 several times over; one CUDA-event pair on the library stream "
               "around the K steps (the per-step figure with a 256 MB fill between steps is reported under `flushed`)",
         "optimizer": "SGD, momentum 0 (optimizer.lua:24-27)"}
    if impl == "reference":
        c["precision"] = "fp64 (Torch7's CPU tensor type), torch-CPU restatement of the nn path"
        c["l2"] = "n/a (host cores)"
        c["parallelism"] = "host cores, one process (rank 0 only)"
    return c


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", d["bf16_tflops"]), d["bf16_tflops"], d["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    return 1400.0, 1590.0, 6650.0, "fallback (B200_PROFILING.md)"


def synthetic_batch(seed, n):
    g = np.load(os.path.join(ROOT, "tests", "golden", "positions.npz"))
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(g["raw"]), n)
    pl = g["player"][idx]
    return dict(raw=np.ascontiguousarray(g["raw"][idx]), player=pl, rank=g["ranks"][idx, pl - 1],
                sym=rng.integers(0, 8, n).astype(np.uint8), x=g["x"][idx], y=g["y"][idx])


def he_params(seed):
    """He-scaled N(0, 2/fan_in) weights of the c2 architecture in Torch7 flat order (W1 of SURVEY.md 8(d))."""
    rng = np.random.default_rng(seed)
    chans = [37] + [C] * (L - 1) + [1]
    ks = [5] + [3] * (L - 1)
    parts = []
    for i in range(L):
        fan = chans[i] * ks[i] * ks[i]
        parts.append(rng.normal(0, np.sqrt(2.0 / fan), chans[i + 1] * fan))
        parts.append(rng.normal(0, 0.05, chans[i + 1]))
        parts.append(rng.normal(0, 0.05, chans[i + 1] * 361))
    return np.concatenate(parts).astype(np.float32)


class ClockSampler:
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                                       "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class Nvml:
    """Synchronous clock / power / throttle-reason samples through NVML (pynvml): one query takes well under a millisecond,
    so a sample can be taken right after every timed step (the nvidia-smi poller above needs ~20 ms per line and saw nothing
    of a 9 ms region)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap", 0x80: "hw_power_brake"}

    def __init__(self, device):
        self.h = None
        self.sm, self.pw, self.reasons = [], [], set()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[device]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else device
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def sample(self):
        if self.h is None:
            return
        try:
            self.sm.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            self.pw.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, nm in self.REASONS.items():
                if r & bit:
                    self.reasons.add(nm)
        except Exception:
            pass

    def take(self):
        """-> summary of the samples since the last take()"""
        out = {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": getattr(self, "max_sm", None),
               "reasons": sorted(self.reasons), "samples": len(self.sm), "power_w_max": max(self.pw) if self.pw else None,
               "source": "NVML, one sample after every timed step" if self.h is not None else "NVML unavailable"}
        self.sm, self.pw, self.reasons = [], [], set()
        return out


class NvmlThread:
    """NVML samples every 20 ms from a helper thread (the sustained loop has no per-step host work to hang a sample on)."""

    def __init__(self, nvml, period=0.02):
        import threading
        self.nvml, self.stop_flag, self.period = nvml, False, period
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        while not self.stop_flag:
            self.nvml.sample()
            time.sleep(self.period)

    def stop(self):
        self.stop_flag = True
        self.t.join()
        return self.nvml.take()


# ------------------------------------------------------------------------------------------------------------
def cpu_step_fn(batch, threads, dtype="f64"):
    """One fwd+bwd+SGD step of the oracle restatement (fp64 = Torch7's CPU tensor type), one eval per step
    (the reference's duplicate pcall eval, train.lua:106, is not replicated)."""
    import torch
    from oracle import encode, net
    torch.set_num_threads(threads)
    b = synthetic_batch(1, batch)
    x, t = encode.load_minibatch(b["raw"], b["player"], b["rank"], b["x"], b["y"], b["sym"])
    x = torch.from_numpy(x)
    flat = torch.from_numpy(he_params(2)).double()
    if dtype == "f32":
        x, flat = x.float(), flat.float()

    def step():
        _, g, _ = net.loss_and_grads(flat, x, t, L, C)
        net.sgd_step(flat, g, 0.01, 1e-7)
    return step


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # bounded sample: batch sized so that the whole run ends within ~2 minutes
    probe = cpu_step_fn(4, threads)
    probe()
    t0 = time.perf_counter(); probe(); t4 = time.perf_counter() - t0
    budget = 100.0 / max(1, args.steps + args.warmup)
    sb = int(max(2, min(BATCH, 4 * budget / max(t4, 1e-3))))
    step = cpu_step_fn(sb, threads)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    v = sb * args.steps / dt
    sample = "%d steps of %d positions (of the %d-position batch), fp64 torch-CPU restatement of the Torch7 nn path" % (args.steps, sb, BATCH)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config(args.gpus, "reference"),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------------------
def timed_steps(ctx, m, rate, steps, warmup, barrier, nvml=None, flush=True):
    """`steps` device-timed steps (CUDA events on the library stream around encode + dg_train_step, an untimed 256 MB fill between
    them) -> (total ms, launches)."""
    for _ in range(warmup):
        m.encode_resident(); m.train_step(rate, want_loss=False)
    barrier()
    total_ms, launches = 0.0, 0
    for _ in range(steps):
        if flush:
            ctx.flush_l2()
        lc = ctx.launch_count()
        ctx.timer_start()
        m.encode_resident(); m.train_step(rate, want_loss=False)
        total_ms += ctx.timer_stop()
        launches += ctx.launch_count() - lc
        if nvml is not None:
            nvml.sample()
    barrier()
    return total_ms, launches


def timed_region(ctx, m, rate, steps, warmup, barrier):
    """EXACTLY `steps` steps back to back inside one CUDA-event pair on the library stream, a barrier + synchronize on both sides
    -> total ms, launches."""
    for _ in range(warmup):
        m.encode_resident(); m.train_step(rate, want_loss=False)
    barrier()
    # one more untimed step: at N > 1 its gradient exchange is a cross-GPU barrier ON THE DEVICE, so the ranks enter the timed
    # region within microseconds of each other (hosts leave a torch.distributed barrier milliseconds apart now and then, and the
    # first timed step would wait for the last one inside the region)
    m.encode_resident(); m.train_step(rate, want_loss=False)
    lc = ctx.launch_count()
    ctx.timer_start()
    for _ in range(steps):
        m.encode_resident(); m.train_step(rate, want_loss=False)
    total_ms = ctx.timer_stop()
    launches = ctx.launch_count() - lc
    barrier()
    return total_ms, launches


def cpu_baselines(threads):
    """The reference's CPU path on this box's host cores, bounded samples: the conv-net step in fp64 (Torch7's CPU
    tensor type) on all cores, in fp32 (the reference's CUDA type), on one thread; and the encode / board-analysis / SGD pieces."""
    import torch
    from oracle import encode as oenc, board as oboard, net

    def rate_of(step, sb, budget, max_n):
        step()
        t0 = time.perf_counter(); n = 0
        while time.perf_counter() - t0 < budget and n < max_n:
            step(); n += 1
        return sb * n / (time.perf_counter() - t0), n

    sb = 16
    v64, n64 = rate_of(cpu_step_fn(sb, threads), sb, 10.0, 8)
    v32, n32 = rate_of(cpu_step_fn(sb, threads, "f32"), sb, 5.0, 8)
    v1, n1 = rate_of(cpu_step_fn(2, 1), 2, 5.0, 2)
    torch.set_num_threads(threads)
    b = synthetic_batch(3, 256)
    t0 = time.perf_counter(); oenc.load_minibatch(b["raw"], b["player"], b["rank"], b["x"], b["y"], b["sym"], np.float32); enc = 256 / (time.perf_counter() - t0)
    g = np.load(os.path.join(ROOT, "tests", "golden", "positions.npz"))
    stones = np.ascontiguousarray(g["raw"][:2048, 0])
    oboard.summarize(stones[:16])
    t0 = time.perf_counter(); oboard.summarize(stones); brd = len(stones) / (time.perf_counter() - t0)
    flat = torch.from_numpy(he_params(2)).double(); gr = torch.ones_like(flat)
    net.sgd_step(flat, gr, 0.0, 0.0)
    t0 = time.perf_counter()
    for _ in range(20):
        net.sgd_step(flat, gr, 0.0, 0.0)
    sgd = 20 * flat.numel() / (time.perf_counter() - t0)
    return {"value": v64, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d steps of %d positions, fp64 torch-CPU restatement of the Torch7 nn path (Torch7 not runnable here)" % (n64, sb),
            "other": {"fp32_all_cores_positions_per_s": v32, "fp32_sample": "%d steps of %d positions" % (n32, sb),
                      "fp64_1_thread_positions_per_s": v1, "fp64_1_thread_sample": "%d steps of 2 positions" % n1,
                      "encode_positions_per_s": enc, "encode_sample": "256 positions, numpy restatement of dataloader.lua:50-92, 1 thread",
                      "board_analysis_boards_per_s": brd, "board_sample": "2048 boards, C restatement of makedata.lua:122-354, OpenMP on all cores",
                      "sgd_params_per_s": sgd, "sgd_sample": "20 updates of the 3.1 M-parameter vector, fp64, torch-CPU"}}


def run_ours(args, rank, local_rank, world):
    import torch
    import __graft_entry__ as entry
    if not os.path.exists(os.path.join(ROOT, "deep-go_b200", "libdeepgo_b200.so")):
        entry.build()
    from deepgo_b200.host import engine, dp as hdp
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    ctx = engine.Context(local_rank)
    if world > 1:
        obj = [engine.dp_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        ctx.init_dp(rank, world, obj[0])
    m = engine.Model(ctx, L, C, engine.DG_PREC_BF16, max_batch=BATCH)
    m.set_params(he_params(2))
    m.set_global_batch(BATCH * world)
    b = synthetic_batch(100 + rank, BATCH)
    rate = 0.01
    warmup = max(3, args.warmup)

    def barrier():
        ctx.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident throughput ("value") ----------------
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    nvml = Nvml(local_rank)
    nvml.sample()
    sampler = NvmlThread(nvml, 0.002)
    total_ms, launches = timed_region(ctx, m, rate, args.steps, warmup, barrier)
    nvml.sample()                                  # the GPU has just finished the last step
    clocks = sampler.stop()
    clocks["source"] = "NVML, every 2 ms during the timed region and once right after it"
    total_ms = max_over_ranks(total_ms)
    ms_per_step = total_ms / args.steps
    value = BATCH * world / (ms_per_step * 1e-3)
    # ---------------- per-kernel times for the roofline (a few extra steps right after the timed region: the kernel is timed alone, at the same clocks) ----------------
    # every rank runs them (the steps contain the gradient exchange); rank 0 reports
    ctx.prof_begin()
    for _ in range(5):
        m.encode_resident(); m.train_step(rate, want_loss=False)
    prof = ctx.prof_end()
    barrier()

    # ---------------- end to end through the C ABI with host buffers ----------------
    pin = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy() for k, v in b.items()}
    for _ in range(3):
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"]); m.train_step(rate)
    barrier()
    for _ in range(2):      # untimed, after the barrier: the step's exchange aligns the ranks on the device, its loss read is the synchronize
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"]); m.train_step(rate)
    # (a) the reference's own loop shape: stage, step, read the cost, strictly one after the other (train.lua:98-131)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"])
        loss = m.train_step(rate)          # returns the step's loss: D2H read + sync inside
    ctx.synchronize()
    e2e_sync_s = max_over_ranks(time.perf_counter() - t0)
    # (b) the same work per step -- H2D copy of the step's records from pinned memory, the step, D2H read of ITS cost -- with the copy of
    # minibatch k+1 issued while step k runs (dg_train_step_begin / dg_model_set_batch_raw / dg_train_step_result): what a loader
    # that has the next minibatch ready (data.lua queues them ahead) lets the host do
    m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.train_step_begin(rate)
        m.set_batch_raw(pin["raw"], pin["player"], pin["rank"], pin["sym"], pin["x"], pin["y"])      # the NEXT step's records
        loss = m.train_step_result()
    ctx.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = BATCH * world * args.steps / e2e_s
    e2e_sync_value = BATCH * world * args.steps / e2e_sync_s
    h2d = BATCH * (9 * 361 + 5)
    assert np.isfinite(loss)

    # ---------------- c4 (BASELINE config 4): 256 positions per GPU (global 2048 at N=8), and the strong-scaling point ----------------
    c4 = {}
    for name, per_gpu in (("weak_256_per_gpu", 256), ("strong_global_2048", 2048 // world)):
        m4 = engine.Model(ctx, L, C, engine.DG_PREC_BF16, max_batch=per_gpu)
        m4.set_params(he_params(2))
        m4.set_global_batch(per_gpu * world)
        b4 = synthetic_batch(200 + rank, per_gpu)
        m4.set_batch_raw(b4["raw"], b4["player"], b4["rank"], b4["sym"], b4["x"], b4["y"])
        steps4 = 20 if per_gpu <= 512 else 8
        ms4, _ = timed_region(ctx, m4, rate, steps4, 3, barrier)
        ms4 = max_over_ranks(ms4) / steps4
        c4[name] = {"batch_per_gpu": per_gpu, "global_batch": per_gpu * world, "ms_per_step": ms4, "positions_per_s": per_gpu * world / (ms4 * 1e-3)}
        del m4

    # ---------------- the round-1 way of timing, for comparison: every step on its own, an untimed 256 MB fill before it (ranks drift
    # apart during the fill and the step's first cross-GPU barrier then waits for the slowest one inside the timed region).  After the
    # headline measurements: after these fills the end-to-end loop showed millisecond stalls in single steps at N >= 4.
    fl_ms, _ = timed_steps(ctx, m, rate, args.steps, 0, barrier, nvml)
    nvml.take()
    fl_ms = max_over_ranks(fl_ms) / args.steps
    flushed = {"ms_per_step": fl_ms, "positions_per_s": BATCH * world / (fl_ms * 1e-3),
               "note": "per-step CUDA events, untimed 256 MB device fill between the steps"}


    # ---------------- sustained: >= 2 s of back-to-back steps, no flush, no per-step sync (last of the timed parts: it heats the GPU into its power cap) ----------------
    n_sus = int(max(200, 2.4 / (ms_per_step * 1e-3)))
    barrier()
    sampler = NvmlThread(nvml)
    ctx.timer_start()
    for _ in range(n_sus):
        m.encode_resident(); m.train_step(rate, want_loss=False)
    sus_ms = ctx.timer_stop()
    sus_clocks = sampler.stop()
    sus_ms = max_over_ranks(sus_ms)
    sustained_run = {"positions_per_s": BATCH * world * n_sus / (sus_ms * 1e-3), "seconds": sus_ms * 1e-3, "steps": n_sus,
                     "ms_per_step": sus_ms / n_sus, "sm_mhz_median": sus_clocks["sm_mhz"], "power_w_max": sus_clocks["power_w_max"],
                     "reasons": sus_clocks["reasons"], "clock_samples": sus_clocks["samples"],
                     "note": "one CUDA-event pair around the whole loop; the batch stays resident (activations + gradients of a step, 0.4 GB, exceed L2)"}

    # ---------------- data-parallel correctness, on the driver's own multi-GPU run ----------------
    dp_check = None
    if world > 1:
        gb = synthetic_batch(999, BATCH * world)
        mine = hdp.shard_batch(gb, rank, world)
        m.set_params(he_params(2))
        dp_losses = []
        for _ in range(3):
            m.set_batch_raw(mine["raw"], mine["player"], mine["rank"], mine["sym"], mine["x"], mine["y"])
            dp_losses.append(m.train_step(rate))
        p_dp = m.get_params()
        t = torch.tensor(dp_losses, device="cuda", dtype=torch.float64)
        dist.all_reduce(t)                                  # equal shards: the global-batch loss is the mean of the shard losses
        dp_losses = (t / world).tolist()
        words = p_dp.view(np.uint32)
        sums = [None] * world
        dist.all_gather_object(sums, (int(words.astype(np.uint64).sum()), int(np.bitwise_xor.reduce(words))))
        if rank == 0:
            ctx1 = engine.Context(local_rank)               # no communicator: rank 0 alone on the whole global batch
            m1 = engine.Model(ctx1, L, C, engine.DG_PREC_BF16, max_batch=BATCH * world)
            m1.set_params(he_params(2))
            one_losses = []
            for _ in range(3):
                m1.set_batch_raw(gb["raw"], gb["player"], gb["rank"], gb["sym"], gb["x"], gb["y"])
                one_losses.append(m1.train_step(rate))
            p1 = m1.get_params()
            dmax = float(np.abs(p_dp - p1).max())
            dl = float(max(abs(a - c) for a, c in zip(dp_losses, one_losses)))
            dp_check = {"replicas_identical": bool(all(s == sums[0] for s in sums)), "ranks": world,
                        "vs_single": {"steps": 3, "global_batch": BATCH * world, "max_abs_param_diff": dmax, "max_loss_diff": dl,
                                      "tol": 2e-3, "ok": bool(dmax < 2e-3 and dl < 2e-3)}}
            del m1
        barrier()
    if dist is not None and rank != 0:
        dist.destroy_process_group()
        return
    if rank != 0:
        return
    sustained, burst, hbm, peak_src = peaks()
    kern = {k: {"ms_per_step": v[0] / 5, "launches_per_step": v[1] / 5} for k, v in prof.items()}
    conv_ms, conv_n = prof.get("conv", (0.0, 0))
    roofline = None
    if conv_n:
        # the 10 forward 3x3 128->128 layers of a step run as conv_n/5 launches (one launch when the layers are chained
        # with the activation image resident in shared memory)
        launches_per_step = conv_n / 5
        layers_per_launch = (L - 2) / launches_per_step
        avg_ms = conv_ms / conv_n
        flops = F_CONV_LAYER * BATCH * layers_per_launch
        achieved = flops / (avg_ms * 1e-3) / 1e12
        # the kernel is timed alone (event pair around one launch, a handful of isolated steps at full clocks): burst peak
        roofline = {"kernel": "conv_umma_kernel (3x3, 128->128, forward, %g layers per launch)" % layers_per_launch, "bound": "tensor",
                    "achieved": achieved, "peak": burst, "unit": "TFLOP/s", "frac": achieved / burst, "traffic": ncu_traffic(),
                    "traffic_source": "static: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel "
                                      "(profiles/roofline_traffic_r2.json), not re-measured by this run",
                    "peak_source": peak_src + ", burst bf16 (kernel timed alone)", "avg_launch_ms": avg_ms, "flops_per_launch": flops,
                    "frac_of_sustained_peak": achieved / sustained}
    whole = F_TRAIN * BATCH / (ms_per_step * 1e-3) / 1e12
    whole_sus = F_TRAIN * BATCH * world / world / (sustained_run["ms_per_step"] * 1e-3) / 1e12

    # ---------------- CPU baseline: oracle port on the host cores, bounded samples (N = 1 only) ----------------
    cpu = cpu_baselines(os.cpu_count() or 1) if world == 1 else None

    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "config": config(world), "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "how": "dg_train_step_begin, dg_model_set_batch_raw of the next minibatch (H2D under the running step), dg_train_step_result",
                "synchronous_loop_value": e2e_sync_value},
        "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu, "flushed": flushed, "sustained": sustained_run, "c4": c4, "dp_check": dp_check,
        "tensor_frac_whole_step": {"achieved_tflops": whole, "of_burst_peak": whole / burst,
                                   "sustained_run_tflops": whole_sus, "sustained_run_of_sustained_peak": whole_sus / sustained},
        "kernels": kern}))
    if dist is not None:
        dist.destroy_process_group()


def ncu_traffic():
    """DRAM bytes (read + write) of one launch of the roofline kernel, from the committed `ncu --set full` capture
    (profiles/roofline_traffic_r2.json: a measured constant of the kernel, not re-measured by the bench)."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic_r2.json")) as f:
            return json.load(f)["dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        return None


def main():
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL prints its version banner to stdout)
    # is sent to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
