"""Host mirror of the reference's Lua modules (deep-go_b200/host): CPU-side checks here, the localtest.lua replay
and loader parity under -m gpu."""
import os
import shutil

import numpy as np
import pytest

import __graft_entry__ as entry

REC = os.path.join(entry.ROOT, "tests", "golden", "records")
GAME = "data/test/1993/2000-03-24b.sgf"


@pytest.fixture(scope="module")
def host():
    entry.build()
    from deepgo_b200.host import dataloader, data, optimizer, experiments, train, engine
    return dict(dataloader=dataloader, data=data, optimizer=optimizer, experiments=experiments, train=train, engine=engine)


def test_channel_constants_match_reference(host):
    d = host["dataloader"]        # dataloader.lua:6-27
    assert (d.STONE, d.LIBERTIES, d.LIBERTIES_AFTER, d.KILL, d.AGE, d.LADDER, d.RANK, d.TOTAL, d.SIZE) == (1, 4, 8, 15, 22, 27, 28, 37, 19)
    assert (d.INPUT.stone, d.INPUT.liberties, d.INPUT.liberties_after, d.INPUT.kills, d.INPUT.age, d.INPUT.ladder, d.INPUT.total) == (1, 2, 3, 5, 7, 8, 9)
    assert d.input_dimensions == [37, 19, 19]


def test_record_reader_matches_packed_corpus(host, golden):
    d = host["dataloader"]
    sel = np.where(golden["split"] == 2)[0]
    for mv in (1, 2, 3, 60, 125):
        i = sel[golden["move"][sel] == mv][0]
        r = d.read_record(os.path.join(REC, GAME, str(mv)))
        assert np.array_equal(r["input"], golden["raw"][i])
        assert (r["move"]["x"], r["move"]["y"], r["move"]["player"]) == (golden["x"][i], golden["y"][i], golden["player"][i])
        assert (r["ranks"][1], r["ranks"][2]) == tuple(golden["ranks"][i])
    import deepgo_b200
    with pytest.raises(deepgo_b200.DeepGoError):
        d.read_record(os.path.join(REC, "nope"))
    with pytest.raises(deepgo_b200.DeepGoError):          # not a Torch7 file
        d.read_record(os.path.join(REC, "data", "test_game_counts.txt"))


def test_dataset_index(host):
    ds = host["data"].Dataset(root=os.path.join(REC, "data"), directory="test")
    ds.init()
    assert ds.games == [{"name": GAME, "size": 3}]        # the zero-move game is dropped (data.lua:74)
    for _ in range(20):
        name = ds.generate_random_filename()
        assert name.startswith(GAME + "/") and 1 <= int(name.rsplit("/", 1)[1]) <= 3
    with pytest.raises(IOError):
        host["data"].Dataset(root=REC, directory="missing").init()


def test_flatten_data_layout(host):
    d = host["dataloader"]
    z = np.zeros((19, 19), np.uint8)
    rec = d.flatten_data(dict(stones=z + 1, liberties=z + 2, liberties_after=np.stack([z + 3, z + 4]), kills=np.stack([z + 5, z + 6]),
                              age=z + 7, ladders=np.stack([z + 8, z + 9]), move={"x": 1, "y": 2, "player": 1}, ranks={1: 3, 2: 4}))
    assert [int(rec["input"][i, 0, 0]) for i in range(9)] == [1, 2, 3, 4, 5, 6, 7, 8, 9] and rec["flat"]


def _make_tree(tmp_path):
    """a data_root with train/validation/test index files over the 3 consecutive fixture records"""
    root = tmp_path / "root"
    shutil.copytree(os.path.join(REC, "data"), root / "data")
    for split in ("train", "validation"):
        shutil.copyfile(root / "data" / "test_game_counts.txt", root / "data" / (split + "_game_counts.txt"))
    return root


@pytest.mark.gpu
def test_load_minibatch_matches_oracle(host, golden):
    from oracle import encode
    files = [os.path.join(REC, GAME, str(m)) for m in (1, 2, 3, 60, 125)]
    mb = host["dataloader"].load_minibatch(files)
    assert mb["input"].shape == (5, 37, 19, 19) and mb["output"].shape == (5,)
    for k, r in enumerate(mb["records"]):
        pl = r["move"]["player"]
        ref = encode.preprocess(r["input"], pl, r["ranks"][pl])
        assert np.array_equal(mb["input"][k], ref.astype(np.float32))
        assert mb["output"][k] == 19 * (r["move"]["x"] - 1) + r["move"]["y"]          # 1-based, dataloader.lua:89
    one = host["dataloader"].load_and_preprocess(files[3])
    assert np.array_equal(one["input"], mb["input"][3]) and one["output"] == mb["output"][3]


@pytest.mark.gpu
def test_queued_minibatches_load_concurrently_and_match_oracle(host, tmp_path):
    """data.lua:82-96: many minibatches queued at once on the loader pool (train() queues all its iterations up front).  Every
    callback must see the planes of ITS files -- different batch sizes, so a shared staging buffer that is re-grown or
    overwritten by another job would show."""
    from oracle import encode
    d, dl = host["data"], host["dataloader"]
    root = _make_tree(tmp_path)
    ds = d.Dataset(root=str(root / "data"), directory="test")
    got = []
    import random
    random.seed(3)
    sizes = [1, 7, 2, 33, 5, 64, 3, 12] * 3
    for n in sizes:
        d.queue_on_minibatch(got.append, ds, n, str(root))
    d.do_queued_tasks()
    assert [len(mb["records"]) for mb in got] == sizes
    for mb in got:
        assert mb["input"].shape == (len(mb["records"]), 37, 19, 19)
        for k, r in enumerate(mb["records"]):
            pl = r["move"]["player"]
            assert np.array_equal(mb["input"][k], encode.preprocess(r["input"], pl, r["ranks"][pl]).astype(np.float32))
            assert mb["output"][k] == 19 * (r["move"]["x"] - 1) + r["move"]["y"]


@pytest.mark.gpu
@pytest.mark.parametrize("precision", [0, 1])
def test_localtest_replay(host, tmp_path, capsys, precision):
    """localtest.lua: basicGoExperiment (3 layers x 64), batchSize 2, validation_interval 20, iters 20."""
    ex = host["experiments"]
    root = _make_tree(tmp_path)
    e = ex.basicGoExperiment.new(dict(useCuda=False, data_root=str(root / "data"), validation_interval=20, validationSize=4, batchSize=2,
                                      precision=precision, rate=0.05))
    os.chdir(tmp_path)
    import random
    random.seed(0)                       # data.py samples (game, move) with Python's RNG, as data.lua uses math.random
    e.init()
    rng = np.random.default_rng(0)
    e.modelParameters.copy(rng.normal(0, 0.05, e.model.engine.P).astype(np.float32))
    costs, vcosts = e.run({"iters": 20, "base": str(root)})
    out = capsys.readouterr().out
    assert "validation at iteration 20" in out and "training" in out and "total samples per second" in out
    assert e.iterations == 20 and len(costs) == 2 and len(vcosts) == 1
    assert np.isfinite(costs).all() and max(costs) < 8.0                  # EMA cost (train.lua:114-115): finite, not diverging
    assert np.isfinite(vcosts).all()
    assert abs(e.optimizer.rate - 0.05 * (1 - 1e-7) ** 20) < 1e-12       # optimizer.lua:26
    # checkpoint / resume in the flat parameter layout (experiments.lua:124-131)
    fn = e.save(str(tmp_path / "ck.model"))
    e2 = ex.Experiment.load(fn)
    assert e2.iterations == 20 and np.array_equal(e2.modelParameters.numpy(), e.modelParameters.numpy())


def _experiment(host, root, tmp_path, **kw):
    ex = host["experiments"]
    opts = dict(useCuda=False, data_root=str(root / "data"), validation_interval=20, validationSize=10, batchSize=4, precision=0, rate=0.01)
    opts.update(kw)
    e = ex.basicGoExperiment.new(opts)
    os.chdir(tmp_path)
    e.init()
    rng = np.random.default_rng(5)
    flat = rng.normal(0, 0.05, e.model.engine.P).astype(np.float32)
    e.modelParameters.copy(flat)
    return e, flat


def _oracle_batch_of(host, names, root):
    from oracle import encode
    recs = [host["dataloader"].read_record(os.path.join(str(root), n)) for n in names]
    raw = np.stack([r["input"] for r in recs])
    pl = np.array([r["move"]["player"] for r in recs]); rk = np.array([r["ranks"][p] for r, p in zip(recs, pl)])
    xs = np.array([r["move"]["x"] for r in recs]); ys = np.array([r["move"]["y"] for r in recs])
    return encode.load_minibatch(raw, pl, rk, xs, ys)


@pytest.mark.gpu
def test_eval_validation_equals_oracle(host, tmp_path):
    """train.lua:14-45 on the device against oracle.net.eval_validation, aggregate quirks included: V = 10, batch 4 -> two batches
    (the tail of 2 is dropped, :24), weights batchSize - 1 (:39), divisor V (:44)."""
    import torch
    from oracle import net
    root = _make_tree(tmp_path)
    e, flat = _experiment(host, root, tmp_path)
    import random
    random.seed(11)
    ds = e.datasets["validation"]
    names = [ds.generate_random_filename() for _ in range(10)]
    x, t = _oracle_batch_of(host, names, root)
    cost, acc = host["train"].eval_validation(e, x.astype(np.float32), (t + 1).astype(np.float64))
    cost_o, acc_o = net.eval_validation(torch.from_numpy(flat).double(), torch.from_numpy(x), t, 4, e.numLayers, e.channelSize)
    assert abs(cost - cost_o) < 1e-4 and acc == acc_o
    assert abs(cost_o - (3 * 2 * 5.88) / 10) < 0.5          # the quirk is really in: 2 batches x weight 3 over V = 10, costs near ln 361


@pytest.mark.gpu
@pytest.mark.parametrize("separate", [False, True])
def test_train_bookkeeping_equals_oracle_replay(host, tmp_path, capsys, separate):
    """30 iterations of train() (train.lua:47-142) against a pure-oracle replay of the same minibatches: the EMA cost list (one entry
    per 10 iterations, .95/.05, seeded with the first cost), the validation at iteration 20 BEFORE that iteration's update, the
    decayed rate.  Both ways through the library: eval() and the update fused into dg_train_step, and as separate calls."""
    import random
    import torch
    from oracle import net
    root = _make_tree(tmp_path)
    e, flat = _experiment(host, root, tmp_path)
    iters, B, V = 30, e.batchSize, e.validationSize
    random.seed(21)
    costs, vcosts = e.run({"iters": iters, "base": str(root), "separate_calls": separate})
    out = capsys.readouterr().out
    # ---- replay
    random.seed(21)
    vnames = [e.datasets["validation"].generate_random_filename() for _ in range(V)]
    tnames = [[e.datasets["train"].generate_random_filename() for _ in range(B)] for _ in range(iters)]
    vx, vt = _oracle_batch_of(host, vnames, root)
    p = torch.from_numpy(flat).double()
    rate, avg, want, vwant = 0.01, None, [], []
    for it in range(1, iters + 1):
        x, t = _oracle_batch_of(host, tnames[it - 1], root)
        cost, g, _ = net.loss_and_grads(p, torch.from_numpy(x), t, e.numLayers, e.channelSize)
        avg = cost if avg is None else avg
        avg = .95 * avg + .05 * cost
        if it % 10 == 0:
            if it % e.validation_interval == 0:
                vwant.append(net.eval_validation(p, torch.from_numpy(vx), vt, B, e.numLayers, e.channelSize)[0])
            want.append(avg)
        rate = net.sgd_step(p, g, rate, e.rateDecay)
    assert len(costs) == 3 and len(vcosts) == 1
    assert np.allclose(costs, want, atol=2e-4), (costs, want)
    assert abs(vcosts[0] - vwant[0]) < 2e-4
    assert abs(e.optimizer.rate - rate) < 1e-15
    assert np.abs(e.modelParameters.numpy() - p.numpy()).max() < 1e-4
    assert out.count("training ") == 2 and out.count("validation at iteration 20") == 1        # print cadence, train.lua:117-127


@pytest.mark.gpu
def test_checkpoint_resume_and_repeated_run(host, tmp_path, capsys):
    """experiment:save / Experiment:load (experiments.lua:124-131) and experiments/repeated.lua: reload the best model, fresh
    SGD.new(rate, rateDecay), run more iterations."""
    import random
    from deepgo_b200.host import repeated
    root = _make_tree(tmp_path)
    e, flat = _experiment(host, root, tmp_path, precision=1)
    random.seed(3)
    e.run({"iters": 20, "base": str(root)})
    assert os.path.exists("%s.model" % e.id)                  # saved at the validation of iteration 20 (train.lua:123)
    fn = e.save(str(tmp_path / "best.model"))
    e2 = host["experiments"].Experiment.load(fn)
    assert e2.iterations == 20 and e2.numLayers == 3 and e2.channelSize == 64 and e2.batchSize == 4 and e2.validation_costs == e.validation_costs
    assert np.array_equal(e2.modelParameters.numpy(), e.modelParameters.numpy())
    assert e2.optimizer.rate == e.optimizer.rate and e2.optimizer.rate < 0.01
    r, _ = repeated.repeated(fn, gpu=1, num=7, iters=10, base=str(root))
    assert r.id == "repeatedRun7" and r.iterations == 30
    assert abs(r.optimizer.rate - 0.01 * (1 - 1e-7) ** 10) < 1e-15          # the loaded run's decayed rate was dropped (repeated.lua:17)
    assert "experiment rate 0.01" in capsys.readouterr().out
    assert not np.array_equal(r.modelParameters.numpy(), e.modelParameters.numpy())
    # a checkpoint in the reference's own format goes through the same door
    from tests.test_checkpoint import write_reference_model
    ref_fn = tmp_path / "ref.model"
    write_reference_model(ref_fn, e.modelParameters.numpy(), "Cuda", 0.004, 1e-7, 1234)
    e3 = host["experiments"].Experiment.load(str(ref_fn), data_root=str(root / "data"), precision=1)
    assert e3.iterations == 1234 and e3.optimizer.rate == 0.004 and np.array_equal(e3.modelParameters.numpy(), e.modelParameters.numpy())
