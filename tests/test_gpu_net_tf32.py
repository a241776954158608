"""-m gpu: DG_PREC_TF32 (tcgen05 kind::tf32, fp32 storage and accumulation) through the C ABI -- the tensor-core type closest to
the reference's CUDA arithmetic (fp32, train.lua:99-100).

Two comparisons, as for the bf16 path:
  (1) against the un-rounded fp64 oracle: forward log-probs 2e-3 abs (TF32 rounds both factors of every product to 10
      mantissa bits, 2^-11 relative each; sums are fp32), loss 1e-3, top-1 equal on >= 95 % of the positions;
  (2) gradients against the oracle with the SAME rounding points (weight operands, stored activations and stored dZ rounded to
      TF32): 5e-3 relative per tensor for nets up to 6 layers (measured 4e-5 .. 1e-3), 1e-1 for the 12-layer net (measured 4-5e-2: ten
      masked layers, and the masks of two forward passes that differ in summation order still differ in a few places).  Against the un-rounded oracle a
      hidden ReLU whose input sits within the rounding noise of zero passes on one side only, and a fraction f of flipped masks
      costs sqrt(f) in every gradient below (measured 2-8e-2; cf. test_gpu_net_bf16.py).
Three SGD steps stay within 1e-3 of the oracle's parameters."""
import numpy as np
import pytest
import torch

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _model(L, C, n, flat):
    from deepgo_b200.host import engine
    m = engine.Model(U.ctx(), L, C, engine.DG_PREC_TF32, max_batch=n)
    m.set_params(flat.numpy())
    return m


def _check(golden, L, C, n, kind, seed, gtol):
    from oracle import net
    b = U.sample(golden, n, seed)
    x, t = U.oracle_batch(b)
    flat = net.init_params(L, C, seed + 2, kind)
    if kind == "he":
        net.split_params(flat, L, C)[-1][1].add_(1.0)          # logits kept off the final ReLU's clamp, as in the bf16 tests
    loss_r, _, logp_r = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C)
    _, g_r, logp_q = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C, quant="tf32")
    m = _model(L, C, n, flat)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    logp = m.forward()
    assert np.isfinite(logp).all()
    print("  vs the oracle with TF32 rounding points: max|dlogp| %.3e" % np.abs(logp - logp_q.numpy()).max())
    assert np.abs(logp - logp_q.numpy()).max() < max(5e-4, 2e-4 * L)      # a stored activation on a rounding boundary goes the other way
    print("tf32 L=%d C=%d n=%d: max|dlogp| %.3e" % (L, C, n, np.abs(logp - logp_r.numpy()).max()))
    assert np.abs(logp - logp_r.numpy()).max() < 2e-3
    assert np.array_equal(m.argmax(), logp.argmax(1))
    assert (logp.argmax(1) == logp_r.numpy().argmax(1)).mean() >= 0.95
    m.zero_grads()
    loss = m.loss_backward()
    assert abs(loss - loss_r) < 1e-3
    g = torch.from_numpy(m.get_grads())
    errs = []
    for li, ((W, bb, a), (Wr, br, ar)) in enumerate(zip(net.split_params(g, L, C), net.split_params(g_r, L, C))):
        errs.append((li, U.rel(W, Wr), U.rel(a, ar), U.rel(bb, br)))
    print("  per-layer relative gradient error (W, a, b):", errs)
    assert max(max(e[1:]) for e in errs) < gtol, errs
    return m, b, flat


def test_localtest_shape(golden):                    # 3 x 64, batch 2 (localtest.lua)
    _check(golden, 3, 64, 2, "he", 0, 5e-3)


def test_c2_width_shallow(golden):                   # 128 filters: image panels of 32 channels, double-buffered
    _check(golden, 4, 128, 6, "he", 1, 5e-3)


def test_c3_width_shallow(golden):                   # 192 filters: two N blocks of 96, two M blocks in the weight gradient
    _check(golden, 4, 192, 3, "he", 2, 5e-3)


def test_torch7_init_6x64(golden):
    _check(golden, 6, 64, 5, "torch7", 3, 5e-3)


def test_c2_depth(golden):                           # 12 x 128
    _check(golden, 12, 128, 6, "he", 1, 1e-1)


def test_more_positions_than_sms_and_train_steps(golden):
    from oracle import net
    L, C, n = 3, 64, 301
    b = U.sample(golden, n, 5)
    x, t = U.oracle_batch(b)
    flat = net.init_params(L, C, 11, "he")
    logp_r, _ = net.forward(flat, torch.from_numpy(x), L, C)
    m = _model(L, C, n, flat)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    assert np.abs(m.forward() - logp_r.numpy()).max() < 2e-3
    L, C, n = 4, 64, 8
    flat = net.init_params(L, C, 9, "he")
    m = _model(L, C, n, flat)
    rate = 0.05
    for step in range(3):
        b = U.sample(golden, n, 200 + step)
        x, t = U.oracle_batch(b)
        loss_r, g_r, _ = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C)
        if step % 2:
            m.set_batch_planes(x.astype(np.float32), t)
        else:
            m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
        loss = m.train_step(rate)
        assert abs(loss - loss_r) < 1e-3, step
        rate = net.sgd_step(flat, g_r, rate, 1e-7)
    assert np.abs(m.get_params() - flat.numpy()).max() < 1e-3
