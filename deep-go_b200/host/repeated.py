"""Host mirror of experiments/repeated.lua: continue training from the best previous checkpoint.

  th experiments/repeated.lua -gpu 1 -num 3 -iters 100          (reference)
  python -m deepgo_b200.host.repeated -gpu 1 -num 3 -iters 100 [-model 0.95223742560484.model]

As the reference does (repeated.lua:14-22): load the checkpoint (one written by this library, or the reference's own torch.save
file), rename the run "repeatedRun<num>", give it a FRESH SGD.new(rate, rateDecay) -- the decayed rate of the loaded run is
dropped on purpose -- and run `iters` more iterations."""
import argparse

from . import experiments
from .optimizer import SGD

BEST_PREVIOUS = "0.95223742560484.model"        # repeated.lua:14


def repeated(model_file=BEST_PREVIOUS, gpu=1, num=1, iters=100, base="", **overrides):
    """overrides: options that replace the checkpoint's (data_root on another machine, validationSize for a small test ..)"""
    experiments.default_device = gpu - 1
    experiment = experiments.Experiment.load(model_file, **overrides)
    experiment.id = "repeatedRun%s" % num
    experiment.optimizer = SGD.new(experiment.rate, experiment.rateDecay)
    print("experiment rate", experiment.rate)
    costs = experiment.run({"iters": iters, "base": base})
    return experiment, costs


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-gpu", type=int, default=1, help="gpu to run on (1-based, as cutorch.setDevice)")
    ap.add_argument("-num", type=int, default=1, help="experiment number")
    ap.add_argument("-iters", type=int, default=100, help="iterations")
    ap.add_argument("-model", default=BEST_PREVIOUS, help="checkpoint to continue from")
    opt = ap.parse_args(argv)
    repeated(opt.model, opt.gpu, opt.num, opt.iters)


if __name__ == "__main__":
    main()
