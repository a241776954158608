"""Host mirror of experiments.lua: the Experiment prototype and its option surface (useCuda, numGPUs, directories,
data_root, validationSize, validation_interval, batchSize, name, kernels, strides, channels, numLayers,
channelSize, rate, rateDecay, criterion), getBasicModel and makeDataParallel.  The model is a handle on a
dg_model: forward/backward/getParameters keep the nn.Module call shape used by train.lua."""
import copy
import os
import random
import time

import numpy as np

from . import engine, data as _data
from .optimizer import SGD
from .train import train


class ClassNLLCriterion:
    """nn.ClassNLLCriterion() (experiments.lua:45): sizeAverage mean NLL; forward/backward are fused into the
    device step, these methods only expose the numbers in the reference's call order (train.lua:8-9,37)."""

    fused = True        # mean NLL with sizeAverage is what dg_train_step computes: train() may fuse eval() with the update

    def cuda(self):
        return self

    def forward(self, preds, targets):
        return preds.model.nll(targets)

    def backward(self, preds, targets):
        return ("dfdo", preds.model)


class FlatVector:
    """One of the two vectors returned by model:getParameters() (experiments.lua:107), Torch7 flat order."""

    def __init__(self, model, kind):
        self.model, self.kind = model, kind

    def zero(self):
        assert self.kind == "grads"
        self.model.engine.zero_grads()

    def numpy(self):
        return self.model.engine.get_grads() if self.kind == "grads" else self.model.engine.get_params()

    def copy(self, values):
        assert self.kind == "params"
        self.model.engine.set_params(values)


class Predictions:
    def __init__(self, model, logp):
        self.model, self.logp = model, logp

    def max(self, dim=2):
        """preds:max(2) (train.lua:29): (max log-prob, 1-based argmax) per row"""
        am = self.model.engine.argmax()
        return self.logp[np.arange(len(am)), am], am + 1


class PolicyNet:
    """getBasicModel(numLayers, kernels, channels) (experiments.lua:133-153) as a device model."""

    def __init__(self, numLayers, kernels, channels, precision=engine.DG_PREC_BF16, max_batch=2048, device=0):
        assert len(kernels) == numLayers and len(channels) == numLayers + 1
        assert all(k == 3 for k in kernels[1:]) and all(c == channels[1] for c in channels[1:-1]) and channels[-1] == 1
        self.spec = dict(numLayers=numLayers, kernels=list(kernels), channels=list(channels), precision=precision, max_batch=max_batch)
        self.ctx = engine.Context(device)
        self.engine = engine.Model(self.ctx, numLayers, channels[1], precision, max_batch, channels[0], kernels[0])
        self._targets = None

    def cuda(self):
        return self

    def getParameters(self):
        return FlatVector(self, "params"), FlatVector(self, "grads")

    def stage(self, inputs, targets1=None):
        t0 = None if targets1 is None else np.asarray(targets1).astype(np.int32) - 1
        self._targets = t0
        self.engine.set_batch_planes(inputs, t0)

    def forward(self, inputs, targets1=None):
        self.stage(inputs, targets1 if targets1 is not None else None)
        return Predictions(self, self.engine.forward())

    def nll(self, targets1):
        t0 = np.asarray(targets1).astype(np.int64) - 1
        lp = self.engine.forward() if self._targets is None else None
        if lp is None:
            _, cost = self.engine.eval_batch()
            return cost
        return float(-lp[np.arange(len(t0)), t0].mean())

    def backward(self, inputs, dfdo):
        return self.engine.loss_backward()

    def train_step(self, inputs, targets1, rate):
        """eval() + optimizer:step of train.lua:111,131 as ONE device call (dg_train_step: the captured step graph, and with
        several GPUs the in-graph gradient exchange); returns the cost eval() would have returned."""
        self.stage(inputs, targets1)
        return self.engine.train_step(rate)


default_device = 0        # cutorch.setDevice (default-experiment.lua:6, experiments/repeated.lua:12), 0-based here


def getBasicModel(numLayers, kernels, channels, **kw):
    kw.setdefault("device", default_device)
    return PolicyNet(numLayers, kernels, channels, **kw)


def makeDataParallel(model, nGPU):
    """experiments.lua:155-168: for nGPU > 1 the model becomes nGPU replicas, one per device of THIS process, behind the same
    interface (nn.DataParallelTable(1): batches split on the batch dimension, gradients summed); the parameters are carried
    over.  nGPU == 1 returns the model unchanged, as the reference does.  (One process per GPU is the other way to the same
    result: host/dp.py.)"""
    if nGPU > 1:
        assert nGPU <= engine.device_count(), "number of GPUs less than nGPU specified"
        params = model.engine.get_params()
        spec = model.spec
        model = PolicyNet(spec["numLayers"], spec["kernels"], spec["channels"], precision=spec["precision"],
                          max_batch=spec["max_batch"], device=list(range(nGPU)))
        model.engine.set_params(params)
    return model


class Experiment:
    defaults = dict(useCuda=False, numGPUs=1, directories={"train": "train", "validation": "validation", "test": "test"},
                    data_root="/home/ubuntu/ebs_disk", validationSize=2000, validation_interval=2000, batchSize=32)

    def __init__(self, proto=None, **kw):
        src = dict(Experiment.defaults) if proto is None else dict(proto.__dict__)
        for k, v in src.items():
            setattr(self, k, copy.copy(v))
        for k, v in kw.items():
            setattr(self, k, v)
        self.datasets = {}
        self.initialized = False

    def new(self, d=None):
        """Prototype copy (experiments.lua:19-31)."""
        return Experiment(self, **(d or {}))

    def init_data(self):
        for group, directory in self.directories.items():
            ds = _data.Dataset(root=self.data_root, directory=directory)
            ds.init()
            self.datasets[group] = ds

    def init(self):
        """experiments.lua:75-108"""
        self.id = random.random()
        self.iterations = 0
        self.initialized = True
        self.optimizer = SGD.new(self.rate, self.rateDecay)
        self.validation_costs = []
        self.init_data()
        kernels, strides, channels = list(self.kernels), list(self.strides), list(self.channels)
        for _ in range(2, self.numLayers + 1):
            kernels.append(3); strides.append(1); channels.append(self.channelSize)
        channels.append(1)
        self.kernels, self.strides, self.channels = kernels, strides, channels
        self.model = getBasicModel(self.numLayers, kernels, channels, precision=getattr(self, "precision", engine.DG_PREC_BF16),
                                   max_batch=max(self.batchSize, 256))
        if self.useCuda:
            self.model = makeDataParallel(self.model.cuda(), self.numGPUs)
        self.modelParameters, self.grads = self.model.getParameters()

    def run(self, params):
        assert params["iters"] > 0
        if not self.initialized:
            self.init()
        start = time.time()
        train_costs, validation_costs = train(self, params)
        self.runningTime = time.time() - start
        return train_costs, validation_costs

    def pickle(self):
        """experiments.lua:55-61: everything but the data sets; the model goes as its flat parameter vector"""
        skip = {"datasets", "model", "modelParameters", "grads", "criterion"}
        d = {k: v for k, v in self.__dict__.items() if k not in skip}
        d["parameters"] = self.modelParameters.numpy()
        return d

    _SCALARS = ("name", "useCuda", "numGPUs", "data_root", "validationSize", "validation_interval", "batchSize", "rate", "rateDecay",
                "id", "iterations", "precision")

    def save(self, opt_filename=None):
        """experiment:save() (experiments.lua:124-127).  One checkpoint format for every host of the library
        (dg_checkpoint_save): the flat parameters plus what a resumed run needs -- options, optimizer rate, iterations, costs."""
        filename = opt_filename or "%s.model" % self.id
        meta = {k: getattr(self, k) for k in self._SCALARS if hasattr(self, k)}
        meta["kernels0"], meta["channels0"] = self.kernels[0], self.channels[0]
        meta["optimizer_rate"] = repr(self.optimizer.rate)
        meta["optimizer_rate_decay"] = repr(self.optimizer.rate_decay)
        meta["validation_costs"] = ",".join(repr(float(v)) for v in self.validation_costs)
        meta["directories"] = ",".join("%s:%s" % kv for kv in sorted(self.directories.items()))
        self.model.engine.save_checkpoint(filename, meta)
        return filename

    @staticmethod
    def load(filename, **overrides):
        """Experiment:load (experiments.lua:129-131 + unpickle :63-72): rebuild the experiment and its model from a checkpoint of
        this library, or -- for a file the reference wrote with torch.save -- take the flat parameter vector, rate and
        iteration count from it and the options from basicGoExperiment."""
        with open(filename, "rb") as f:
            native = f.read(9) == b"DGCKPT02\n"
        if not native:
            return Experiment.load_reference(filename, **overrides)
        meta, params = engine.read_checkpoint(filename)
        conv = {"useCuda": lambda v: v == "True", "numGPUs": int, "validationSize": int, "validation_interval": int, "batchSize": int,
                "rate": float, "rateDecay": float, "id": lambda v: v, "iterations": int, "precision": int, "name": str, "data_root": str}
        kw = {k: conv[k](meta[k]) for k in conv if k in meta}
        for k in ("id",):
            try:
                kw[k] = float(kw[k])
            except (KeyError, ValueError):
                pass
        if meta.get("directories"):
            kw["directories"] = dict(kv.split(":", 1) for kv in meta["directories"].split(","))
        kw.update(overrides)                      # e.g. data_root on another machine
        iterations = kw.pop("iterations", 0)
        ident = kw.pop("id", None)
        e = Experiment(criterion=ClassNLLCriterion(), kernels=[int(meta.get("kernels0", meta["kFirst"]))], strides=[1],
                       channels=[int(meta.get("channels0", meta["inPlanes"]))], numLayers=int(meta["numLayers"]),
                       channelSize=int(meta["channelSize"]), **kw)
        e.init()
        e.iterations = iterations
        if ident is not None:
            e.id = ident
        e.validation_costs = [float(v) for v in meta.get("validation_costs", "").split(",") if v]
        e.optimizer = SGD.new(float(meta.get("optimizer_rate", e.rate)), float(meta.get("optimizer_rate_decay", e.rateDecay)))
        e.modelParameters.copy(params)
        return e

    @staticmethod
    def load_reference(filename, proto=None, **kw):
        """A `.model` file of the reference itself: the architecture follows from the length of its flat vector
        (numLayers x channelSize of `proto`, default basicGoExperiment, must match)."""
        params, rate, decay, iterations = engine.read_reference_model(filename)
        e = (proto or basicGoExperiment).new(kw)
        e.init()
        assert e.model.engine.P == len(params), "the checkpoint holds %d parameters, the experiment's model has %d" % (len(params), e.model.engine.P)
        e.modelParameters.copy(params)
        if rate is not None:
            e.optimizer = SGD.new(rate, decay if decay is not None else e.rateDecay)
        if iterations is not None:
            e.iterations = iterations
        return e


basicGoExperiment = Experiment(name="basicGoExperiment", kernels=[5], strides=[1], channels=[37], numLayers=3, channelSize=64,
                               rate=.01, rateDecay=1e-7, criterion=ClassNLLCriterion())
