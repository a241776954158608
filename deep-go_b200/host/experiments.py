"""Host mirror of experiments.lua: the Experiment prototype and its option surface (useCuda, numGPUs, directories,
data_root, validationSize, validation_interval, batchSize, name, kernels, strides, channels, numLayers,
channelSize, rate, rateDecay, criterion), getBasicModel and makeDataParallel.  The model is a handle on a
dg_model: forward/backward/getParameters keep the nn.Module call shape used by train.lua."""
import copy
import os
import pickle
import random
import time

import numpy as np

from . import engine, data as _data
from .optimizer import SGD
from .train import train


class ClassNLLCriterion:
    """nn.ClassNLLCriterion() (experiments.lua:45): sizeAverage mean NLL; forward/backward are fused into the
    device step, these methods only expose the numbers in the reference's call order (train.lua:8-9,37)."""

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


def getBasicModel(numLayers, kernels, channels, **kw):
    return PolicyNet(numLayers, kernels, channels, **kw)


def makeDataParallel(model, nGPU):
    """experiments.lua:155-168.  Here data parallelism is one process per GPU (host/dp.py); inside one process the
    model is returned unchanged, as the reference does for nGPU == 1."""
    if nGPU > 1:
        assert nGPU <= engine.device_count(), "number of GPUs less than nGPU specified"
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
        skip = {"datasets", "model", "modelParameters", "grads", "criterion"}
        d = {k: v for k, v in self.__dict__.items() if k not in skip}
        d["parameters"] = self.modelParameters.numpy()
        return d

    def save(self, opt_filename=None):
        """experiment:save() (experiments.lua:124-127): flat parameters + optimizer rate + iterations + costs"""
        filename = opt_filename or "%s.model" % self.id
        with open(filename, "wb") as f:
            pickle.dump(self.pickle(), f)
        return filename

    @staticmethod
    def load(filename):
        with open(filename, "rb") as f:
            d = pickle.load(f)
        params = d.pop("parameters")
        e = Experiment(**d)
        e.initialized = False
        keep = (e.iterations, e.validation_costs, e.optimizer, e.id)
        e.kernels, e.strides, e.channels = [e.kernels[0]], [1], [e.channels[0]]
        e.init()
        e.iterations, e.validation_costs, e.optimizer, e.id = keep
        e.modelParameters.copy(params)
        return e


basicGoExperiment = Experiment(name="basicGoExperiment", kernels=[5], strides=[1], channels=[37], numLayers=3, channelSize=64,
                               rate=.01, rateDecay=1e-7, criterion=ClassNLLCriterion())
