"""world_size-2 gloo tests (CPU) of the data-parallel host logic: sharding as DataParallelTable(1) does
(experiments.lua:160), the communicator-id exchange, and the gradient semantics the library relies on --
per-rank gradients scaled by 1/global_batch and SUMmed equal the global-batch mean gradient."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import __graft_entry__ as entry  # noqa: F401  (puts the repo root on sys.path)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from deepgo_b200.host import dp
        from oracle import encode, net
        # id exchange: every rank ends up with rank 0's 128 bytes
        uid = dp.exchange_unique_id(dist, rank, lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        g = np.load(os.path.join(entry.ROOT, "tests", "golden", "positions.npz"))
        idx = np.arange(0, 70, 10)                       # 7 positions: ragged over 2 ranks (4 + 3)
        pl = g["player"][idx]
        batch = dict(raw=g["raw"][idx], player=pl, rank=g["ranks"][idx, pl - 1], x=g["x"][idx], y=g["y"][idx])
        mine = dp.shard_batch(batch, rank, world)
        L, C = 3, 8
        flat = net.init_params(L, C, 1, "he")
        x, t = encode.load_minibatch(mine["raw"], mine["player"], mine["rank"], mine["x"], mine["y"])
        _, grads, _ = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C, grad_scale=1.0 / len(idx))
        dist.all_reduce(grads, op=dist.ReduceOp.SUM)
        xa, ta = encode.load_minibatch(batch["raw"], batch["player"], batch["rank"], batch["x"], batch["y"])
        _, g_all, _ = net.loss_and_grads(flat, torch.from_numpy(xa), ta, L, C)
        q.put((rank, len(mine["raw"]), float((grads - g_all).abs().max())))
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_semantics():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [4, 3]
    assert all(r[2] < 1e-12 for r in res)


def test_shard_ranges_cover_and_are_disjoint():
    from deepgo_b200.host import dp
    for n in (1, 2, 7, 128, 2048):
        for world in (1, 2, 3, 8):
            spans = [dp.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
