"""CPU: the host mirror's training loop (host/train.py, host/experiments.py, host/optimizer.py, host/data.py) with the device model
replaced by an oracle-backed stand-in.  What runs on the GPU in the product (forward, NLL, backward, SGD) is oracle.net here, so the
test isolates the HOST logic the reference keeps in Lua: minibatch queueing order, the EMA cost (.95/.05, seeded with the first
cost), the print / validation / save cadence, validation before the update of its iteration, the rate decay after use, the fused
path's equivalence with separate calls (train.lua:47-142, optimizer.lua:24-27)."""
import os
import random
import shutil

import numpy as np
import pytest
import torch

import __graft_entry__ as entry
from oracle import encode, net

REC = os.path.join(entry.ROOT, "tests", "golden", "records")


class OracleNet:
    """the interface host/train.py and host/experiments.py use of PolicyNet, computed by the oracle in fp64"""

    class _Engine:
        def __init__(self, owner):
            self.o = owner
            self.P = owner.flat.numel()

        def get_params(self): return self.o.flat.numpy().astype(np.float32)
        def set_params(self, v): self.o.flat.copy_(torch.from_numpy(np.asarray(v, np.float64)))
        def get_grads(self): return self.o.grads.numpy().astype(np.float32)
        def zero_grads(self): self.o.grads.zero_()
        def sgd_step(self, rate): net.sgd_step(self.o.flat, self.o.grads, rate, 0.0)
        def argmax(self): return self.o.logp.argmax(1).numpy().astype(np.int32)

    def __init__(self, L, C):
        self.L, self.C = L, C
        self.flat = net.init_params(L, C, 0, "he")
        self.grads = torch.zeros_like(self.flat)
        self.engine = OracleNet._Engine(self)
        self.calls = []

    def cuda(self): return self
    def getParameters(self):
        from deepgo_b200.host.experiments import FlatVector
        return FlatVector(self, "params"), FlatVector(self, "grads")

    def forward(self, inputs, targets1=None):
        from deepgo_b200.host.experiments import Predictions
        self.x = torch.from_numpy(np.asarray(inputs, np.float64))
        self.t0 = None if targets1 is None else np.asarray(targets1).astype(np.int64) - 1
        self.logp, _ = net.forward(self.flat, self.x, self.L, self.C)
        self.calls.append("forward")
        return Predictions(self, self.logp.numpy())

    def nll(self, targets1):
        t0 = np.asarray(targets1).astype(np.int64) - 1
        return float(-self.logp[torch.arange(len(t0)), torch.from_numpy(t0)].mean())

    def backward(self, inputs, dfdo):
        _, g, _ = net.loss_and_grads(self.flat, self.x, self.t0, self.L, self.C)
        self.grads.add_(g)
        self.calls.append("backward")

    def train_step(self, inputs, targets1, rate):
        x = torch.from_numpy(np.asarray(inputs, np.float64))
        t0 = np.asarray(targets1).astype(np.int64) - 1
        loss, g, _ = net.loss_and_grads(self.flat, x, t0, self.L, self.C)
        net.sgd_step(self.flat, g, rate, 0.0)
        self.calls.append("train_step")
        return loss


@pytest.fixture()
def host(monkeypatch, tmp_path):
    entry.build()
    from deepgo_b200.host import dataloader, data, experiments, train
    root = tmp_path / "root"
    shutil.copytree(os.path.join(REC, "data"), root / "data")
    for split in ("train", "validation"):
        shutil.copyfile(root / "data" / "test_game_counts.txt", root / "data" / (split + "_game_counts.txt"))

    def encode_records(recs, sym=None):            # the device half of load_minibatch, by the oracle
        raw = np.stack([r["input"] for r in recs])
        pl = np.array([r["move"]["player"] for r in recs]); rk = np.array([r["ranks"][p] for r, p in zip(recs, pl)])
        xs = np.array([r["move"]["x"] for r in recs]); ys = np.array([r["move"]["y"] for r in recs])
        x, t = encode.load_minibatch(raw, pl, rk, xs, ys, None, np.float32)
        return {"input": x, "output": (t + 1).astype(np.float64), "records": recs}

    monkeypatch.setattr(dataloader, "encode_records", encode_records)
    monkeypatch.setattr(experiments, "getBasicModel", lambda numLayers, kernels, channels, **kw: OracleNet(numLayers, channels[1]))
    monkeypatch.chdir(tmp_path)
    return dict(experiments=experiments, train=train, data=data, root=root)


@pytest.mark.parametrize("separate", [False, True])
def test_training_loop_bookkeeping(host, capsys, separate):
    ex = host["experiments"]
    e = ex.basicGoExperiment.new(dict(useCuda=False, data_root=str(host["root"] / "data"), validation_interval=20, validationSize=10,
                                      batchSize=4, rate=0.01))
    e.save = lambda opt_filename=None: saved.append(e.iterations)            # checkpoint writing needs the device library's model handle
    saved = []
    e.init()
    flat0 = e.model.flat.clone()
    random.seed(21)
    costs, vcosts = e.run({"iters": 30, "base": str(host["root"]), "separate_calls": separate})
    out = capsys.readouterr().out
    # replay with the oracle alone
    random.seed(21)
    from deepgo_b200.host import dataloader
    names = lambda ds, n: [os.path.join(str(host["root"]), ds.generate_random_filename()) for _ in range(n)]
    vb = dataloader.encode_records(dataloader.read_records(names(e.datasets["validation"], 10)))
    tb = [dataloader.encode_records(dataloader.read_records(names(e.datasets["train"], 4))) for _ in range(30)]
    p, rate, avg, want, vwant = flat0.clone(), 0.01, None, [], []
    for it in range(1, 31):
        x, t0 = torch.from_numpy(tb[it - 1]["input"]).double(), (tb[it - 1]["output"] - 1).astype(np.int64)
        cost, g, _ = net.loss_and_grads(p, x, t0, e.numLayers, e.channelSize)
        avg = cost if avg is None else avg
        avg = .95 * avg + .05 * cost
        if it % 10 == 0:
            if it % 20 == 0:
                vwant.append(net.eval_validation(p, torch.from_numpy(vb["input"]).double(), (vb["output"] - 1).astype(np.int64), 4, e.numLayers, e.channelSize)[0])
            want.append(avg)
        rate = net.sgd_step(p, g, rate, e.rateDecay)
    assert np.allclose(costs, want, atol=1e-9) and np.allclose(vcosts, vwant, atol=1e-9)
    assert abs(e.optimizer.rate - rate) < 1e-18 and saved == [20]
    assert torch.allclose(e.model.flat, p, atol=1e-12)
    assert out.count("training ") == 2 and out.count("validation at iteration 20") == 1 and "total samples per second" in out
    calls = e.model.calls
    if separate:
        assert "train_step" not in calls
    else:
        # every iteration is one fused call except the one that validates between eval() and the update
        assert calls.count("train_step") == 29 and calls.count("backward") == 1
