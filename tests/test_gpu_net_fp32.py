"""-m gpu: fp32 path (reference CUDA arithmetic type, train.lua:99) through the C ABI vs the fp64 oracle.
Tolerances: forward log-probs 1e-4 abs; gradients 1e-4 relative per tensor; SGD update 1e-6 abs."""
import numpy as np
import pytest
import torch

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _run(golden, L, C, n, kind, seed):
    from deepgo_b200.host import engine
    from oracle import net
    b = U.sample(golden, n, seed)
    x, t = U.oracle_batch(b)
    flat = net.init_params(L, C, seed + 2, kind)
    loss_ref, g_ref, logp_ref = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C)
    m = engine.Model(U.ctx(), L, C, engine.DG_PREC_FP32, max_batch=n)
    m.set_params(flat.numpy())
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    logp = m.forward()
    assert np.abs(logp - logp_ref.numpy()).max() < 1e-4
    assert np.allclose(np.exp(logp.astype(np.float64)).sum(1), 1.0, atol=1e-5)
    assert np.array_equal(m.argmax(), logp.argmax(1))
    m.zero_grads()
    loss = m.loss_backward()
    assert abs(loss - loss_ref) < 1e-5
    g = m.get_grads()
    for (W, bb, a), (Wr, br, ar) in zip(net.split_params(torch.from_numpy(g), L, C), net.split_params(g_ref, L, C)):
        assert U.rel(W, Wr) < 1e-4 and U.rel(a, ar) < 1e-4 and U.rel(bb, br) < 1e-4
    # SGD:step (optimizer.lua:25)
    m.sgd_step(0.05)
    p_ref = flat.clone()
    net.sgd_step(p_ref, g_ref, 0.05, 0.0)
    assert np.abs(m.get_params() - p_ref.numpy()).max() < 1e-6
    return m, b, flat, t


def test_localtest_shape_c1(golden):            # localtest.lua: 3 layers x 64, batch 2
    _run(golden, 3, 64, 2, "torch7", 0)


def test_he_weights_6x64(golden):               # default-experiment.lua shape, signal kept alive
    _run(golden, 6, 64, 5, "he", 1)


def test_planes_staging_matches_raw_staging(golden):
    from deepgo_b200.host import engine
    m, b, flat, t = _run(golden, 3, 16, 4, "he", 2)
    lp_raw = m.forward()
    x, t = U.oracle_batch(b, np.float32)
    m.set_batch_planes(x, t)
    assert np.array_equal(m.forward(), lp_raw)
    errs, cost = m.eval_batch()
    assert errs == int((lp_raw.argmax(1) != t).sum())
    assert abs(cost - float(-lp_raw[np.arange(4), t].mean())) < 1e-5


def test_train_steps_follow_oracle(golden):
    """5 fused train steps (eval + SGD:step, train.lua:111,131) track the oracle's fp64 trajectory."""
    from deepgo_b200.host import engine
    from oracle import net
    L, C, n = 4, 32, 8
    flat = net.init_params(L, C, 9, "he")
    m = engine.Model(U.ctx(), L, C, engine.DG_PREC_FP32, max_batch=n)
    m.set_params(flat.numpy())
    rate = 0.05
    for step in range(5):
        b = U.sample(golden, n, 100 + step)
        x, t = U.oracle_batch(b)
        loss_ref, g_ref, _ = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C)
        m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
        loss = m.train_step(rate)
        assert abs(loss - loss_ref) < 1e-4
        rate = net.sgd_step(flat, g_ref, rate, 1e-7)
    assert np.abs(m.get_params() - flat.numpy()).max() < 1e-4


def test_api_errors(golden):
    import deepgo_b200
    from deepgo_b200.host import engine
    m = engine.Model(U.ctx(), 3, 8, engine.DG_PREC_FP32, max_batch=2)
    with pytest.raises(deepgo_b200.DeepGoError):
        m.set_params(np.zeros(5, np.float32))
    b = U.sample(golden, 3, 0)
    with pytest.raises(deepgo_b200.DeepGoError):             # batch larger than max_batch
        m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    with pytest.raises(deepgo_b200.DeepGoError):             # backward before forward
        m.set_batch_raw(b["raw"][:2], b["player"][:2], b["rank"][:2], None, b["x"][:2], b["y"][:2])
        m.loss_backward()
