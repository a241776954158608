"""Host mirror of data.lua: Dataset index (:53-80), random (game, move) sampling (:29-37) and the
queue_on_minibatch / do_queued_tasks pair (:82-96).  Jobs run on a thread pool (the record reader of the C
library is reentrant); callbacks run on the caller's thread inside do_queued_tasks, as `threads` does."""
import os
import random
from concurrent.futures import ThreadPoolExecutor

from . import dataloader

_pool = None
_queued = []


def prepare_data_loaders(number_loaders=32):
    global _pool
    _pool = ThreadPoolExecutor(max_workers=number_loaders) if number_loaders and number_loaders > 1 else None


prepare_data_loaders()


class Dataset:
    root = "/home/ubuntu/ebs_disk"
    directory = "train"

    def __init__(self, **kw):
        self.initialized = False
        self.games = []
        for k, v in kw.items():
            setattr(self, k, v)

    @classmethod
    def new(cls, d=None):
        return cls(**(d or {}))

    def init(self):
        """Read <root>/<directory>_game_counts.txt: `<path>\\t<count>`; games without moves are dropped (:74)."""
        self.games = []
        path = os.path.join(self.root, self.directory)
        file = path + "_game_counts.txt"
        if not os.path.exists(file):
            raise IOError("unable to find counts (probably bad dataset path): " + file)
        with open(file) as f:
            for line in f:
                r = line.rstrip("\n").split("\t")
                size = int(r[1])
                if size > 0:
                    self.games.append({"name": r[0], "size": size})
        self.initialized = True

    def generate_random_filename(self):
        """uniform game, then uniform move (not uniform over positions), data.lua:29-37"""
        if not self.initialized:
            self.init()
        game = random.choice(self.games)
        return "%s/%d" % (game["name"], random.randint(1, game["size"]))


def queue_on_minibatch(f, dataset, n, base=None):
    """Draw n random filenames, load them on a worker, call f(minibatch) on the caller's thread later."""
    if not dataset.initialized:
        dataset.init()
    root = base if base is not None else getattr(dataset, "base", "")
    filenames = [os.path.join(root, dataset.generate_random_filename()) for _ in range(n)]
    job = _pool.submit(dataloader.load_minibatch, filenames) if _pool else None
    _queued.append((job, filenames, f))


def do_queued_tasks():
    while _queued:
        job, filenames, f = _queued.pop(0)
        f(job.result() if job is not None else dataloader.load_minibatch(filenames))
