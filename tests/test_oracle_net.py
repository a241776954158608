"""Oracle self-checks for the net restatement: shapes pinned by the reference notebook, explicit backward vs
autograd, finite differences (fp64), SGD formula, eval_validation quirks."""
import numpy as np
import torch

from oracle import encode, net


def _batch(golden, n, seed=0):
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(golden["raw"]), n)
    pl = golden["player"][idx]
    x, t = encode.load_minibatch(golden["raw"][idx], pl, golden["ranks"][idx, pl - 1], golden["x"][idx], golden["y"][idx])
    return torch.from_numpy(x), t


def test_shapes_pinned_by_notebook():
    # "Run Experiment.ipynb":95-96,131,175-176,263-264,289 : 3-layer/64 model
    specs = net.layer_specs(3, 64)
    assert specs == [(37, 64, 5), (64, 64, 3), (64, 1, 3)]
    flat = torch.zeros(net.param_count(3, 64), dtype=torch.float64)
    (W1, b1, a1), (W2, b2, a2), (W3, b3, a3) = net.split_params(flat, 3, 64)
    assert W1.reshape(64, -1).shape == (64, 925) and a1.numel() == 23104 and a3.numel() == 361
    assert W2.reshape(64, -1).shape == (64, 576) and W3.reshape(1, -1).shape == (1, 576)
    assert net.param_count(3, 64) == 143338 and net.param_count(12, 128) == 2104170 and net.param_count(12, 192) == 4261994


def test_backward_matches_autograd(golden):
    x, t = _batch(golden, 5)
    L, C = 4, 8
    flat = net.init_params(L, C, 3, "he")
    loss, grads, logp = net.loss_and_grads(flat, x, t, L, C)
    f2 = flat.clone().requires_grad_(True)
    lp, _ = net.forward(f2, x, L, C)
    l2 = torch.nn.functional.nll_loss(lp, torch.as_tensor(t, dtype=torch.long))
    l2.backward()
    assert abs(loss - l2.item()) < 1e-14
    assert torch.allclose(grads, f2.grad, atol=1e-14, rtol=1e-12)
    assert torch.allclose(torch.exp(logp).sum(1), torch.ones(5, dtype=torch.float64), atol=1e-12)


def test_finite_differences(golden):
    x, t = _batch(golden, 3, seed=1)
    L, C = 3, 4
    flat = net.init_params(L, C, 5, "he")
    _, grads, _ = net.loss_and_grads(flat, x, t, L, C)
    rng = np.random.default_rng(0)
    for j in rng.integers(0, flat.numel(), 25):
        e = 1e-6
        fp, fm = flat.clone(), flat.clone()
        fp[j] += e; fm[j] -= e
        lp = net.loss_and_grads(fp, x, t, L, C)[0]
        lm = net.loss_and_grads(fm, x, t, L, C)[0]
        assert abs((lp - lm) / (2 * e) - grads[j].item()) < 1e-6


def test_sgd_formula():
    p = torch.tensor([1.0, 2.0], dtype=torch.float64); g = torch.tensor([0.5, -1.0], dtype=torch.float64)
    r = net.sgd_step(p, g, 0.1, 1e-2)                   # optimizer.lua:25-26
    assert torch.allclose(p, torch.tensor([0.95, 2.1], dtype=torch.float64))
    assert abs(r - 0.1 * 0.99) < 1e-15


def test_eval_validation_quirks(golden):
    x, t = _batch(golden, 10, seed=4)
    L, C = 3, 4
    flat = net.init_params(L, C, 7, "he")
    cost, acc = net.eval_validation(flat, x, t, 4, L, C)     # 10/4 -> 2 batches, tail of 2 dropped
    lp, _ = net.forward(flat, x[:8], L, C)
    tt = torch.as_tensor(t[:8], dtype=torch.long)
    c1 = -lp[:4][torch.arange(4), tt[:4]].mean().item(); c2 = -lp[4:][torch.arange(4), tt[4:]].mean().item()
    assert abs(cost - (c1 * 3 + c2 * 3) / 10) < 1e-12       # size = batchSize-1, divisor = V (train.lua:39,44)
    errs = int((lp.argmax(1) != tt).sum())
    assert abs(acc - (1 - errs / 10)) < 1e-12


def test_bf16_quant_mode_is_close(golden):
    x, t = _batch(golden, 4, seed=2)
    L, C = 4, 16
    flat = net.init_params(L, C, 3, "he")
    l0, g0, _ = net.loss_and_grads(flat, x, t, L, C)
    l1, g1, _ = net.loss_and_grads(flat, x, t, L, C, quant="bf16")
    # bf16 rounding flips a few ReLU masks on the (clamped) logits, which dominates the gradient difference
    assert abs(l0 - l1) < 0.05
    assert (g0 - g1).norm() / g0.norm() < 0.25
