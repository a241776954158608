"""-m gpu: tcgen05 bf16 path through the C ABI.

Two comparisons, tolerances stated here:
  (1) against the oracle with the SAME rounding points (weights, stored activations and stored dZ rounded to
      bf16, fp32/fp64 accumulation): forward log-probs 1e-3 * numLayers abs (an activation that sits on a bf16
      rounding boundary rounds the other way under a different accumulation order, and each layer adds its
      share), per-tensor gradient error 2e-2 relative; for the 12-layer net 2e-1 relative and cosine > 0.96
      (see test_c2_shape_small_batch: hidden ReLU masks that differ between two bf16 forward passes);
  (2) against the un-rounded fp64 reference arithmetic: log-probs 5e-2 abs, top-1 agreement >= 90% on
      He-scaled weights (north_star: "within a stated fp32 (tf32/bf16) tolerance, top-1 move agreement").
"""
import numpy as np
import pytest
import torch

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _model(L, C, n, flat):
    from deepgo_b200.host import engine
    m = engine.Model(U.ctx(), L, C, engine.DG_PREC_BF16, max_batch=n)
    m.set_params(flat.numpy())
    return m


def _check(golden, L, C, n, kind, seed, gtol=2e-2):
    from oracle import net
    b = U.sample(golden, n, seed)
    x, t = U.oracle_batch(b)
    flat = net.init_params(L, C, seed + 2, kind)
    if kind == "he":
        # The reference keeps a ReLU on the logits (experiments.lua:147).  With logits straddling 0 a bf16-level
        # difference flips that mask and dominates every gradient below it (ill-conditioned by construction, not a
        # kernel property), so the He-scaled parity nets lift the last layer's bias to keep the logits positive.
        net.split_params(flat, L, C)[-1][1].add_(1.0)
    xq = torch.from_numpy(x)
    loss_q, g_q, logp_q = net.loss_and_grads(flat, xq, t, L, C, quant="bf16")
    loss_r, g_r, logp_r = net.loss_and_grads(flat, xq, t, L, C)
    m = _model(L, C, n, flat)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    logp = m.forward()
    assert np.isfinite(logp).all()
    assert np.abs(logp - logp_q.numpy()).max() < 1e-3 * L
    assert np.abs(logp - logp_r.numpy()).max() < 5e-2
    assert np.allclose(np.exp(logp.astype(np.float64)).sum(1), 1.0, atol=1e-4)
    assert np.array_equal(m.argmax(), logp.argmax(1))
    if kind == "he":
        assert (logp.argmax(1) == logp_r.numpy().argmax(1)).mean() >= 0.9
    m.zero_grads()
    loss = m.loss_backward()
    assert abs(loss - loss_q) < 1e-3 * L
    g = torch.from_numpy(m.get_grads())
    errs, cos = [], []
    for li, ((W, bb, a), (Wq, bq, aq)) in enumerate(zip(net.split_params(g, L, C), net.split_params(g_q, L, C))):
        errs.append((li, U.rel(W, Wq), U.rel(a, aq), U.rel(bb, bq)))
        cos.append(float((W.double().flatten() @ Wq.flatten()) / (W.double().norm() * Wq.norm())))
    print("per-layer relative gradient error (W, a, b) L=%d C=%d n=%d:" % (L, C, n), errs, "cos(W):", cos)
    worst = max(max(e[1:]) for e in errs)
    assert worst < gtol, errs
    assert min(cos) > 1 - gtol * gtol, cos
    return m, b, flat


def test_localtest_shape_c1(golden):                 # 3 x 64, batch 2 (localtest.lua)
    _check(golden, 3, 64, 2, "he", 0)


def test_c2_width_shallow(golden):                   # 128 filters, 4 layers: every GEMM shape of c2, few mask flips
    _check(golden, 4, 128, 6, "he", 1)


def test_c2_shape_small_batch(golden):               # 12 x 128 (BASELINE config 2 architecture), 6 positions
    # Deep net: GPU and oracle forward passes differ by bf16 rounding decisions, so a fraction f of the hidden
    # ReLU masks differs and every dZ below carries an L2 error of about sqrt(f) (f ~ 1% at depth 10 -> ~10%),
    # unbiased and independent of the kernels (the 4-layer net of the same width is held to 2e-2 above).
    _check(golden, 12, 128, 6, "he", 1, gtol=2e-1)


def test_192_filters(golden):                        # c3 architecture width, shallow
    _check(golden, 4, 192, 3, "he", 2)


def test_torch7_init_6x64(golden):
    _check(golden, 6, 64, 5, "torch7", 3)


def test_multi_position_per_cta_and_ragged_batch(golden):
    """more positions than SMs: every CTA loops over several positions; batch not a multiple of anything"""
    from oracle import net
    L, C, n = 3, 64, 301
    b = U.sample(golden, n, 5)
    x, t = U.oracle_batch(b)
    flat = net.init_params(L, C, 11, "he")
    logp_q, _ = net.forward(flat, torch.from_numpy(x), L, C, quant="bf16")
    m = _model(L, C, n, flat)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    logp = m.forward()
    assert np.abs(logp - logp_q.numpy()).max() < 2e-3


def test_planes_staging_and_train_steps(golden):
    from oracle import net
    L, C, n = 4, 64, 8
    flat = net.init_params(L, C, 9, "he")
    m = _model(L, C, n, flat)
    rate = 0.05
    for step in range(4):
        b = U.sample(golden, n, 200 + step)
        x, t = U.oracle_batch(b)
        loss_q, g_q, _ = net.loss_and_grads(flat, torch.from_numpy(x), t, L, C, quant="bf16")
        if step % 2:
            m.set_batch_planes(x.astype(np.float32), t)
        else:
            m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
        loss = m.train_step(rate)
        assert abs(loss - loss_q) < 5e-3, step
        rate = net.sgd_step(flat, g_q, rate, 1e-7)
    assert U.rel(m.get_params(), flat.numpy()) < 1e-3


def test_full_size_properties(golden):
    """c2 at its full batch (128): properties that need no oracle run -- probabilities sum to 1, the loss of a
    repeated step decreases, gradients of a duplicated batch equal those of the single batch."""
    from oracle import net
    L, C, n = 12, 128, 128
    flat = net.init_params(L, C, 4, "he")
    m = _model(L, C, n, flat)
    b = U.sample(golden, n, 7)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    logp = m.forward()
    assert np.isfinite(logp).all()
    assert np.allclose(np.exp(logp.astype(np.float64)).sum(1), 1.0, atol=1e-4)
    m.zero_grads(); m.loss_backward()
    g_full = m.get_grads()
    # mean gradient of [batch, batch] equals the mean gradient of batch (linearity of the batch reduction)
    half = {k: np.concatenate([v[:64], v[:64]]) for k, v in b.items()}
    m.set_batch_raw(half["raw"], half["player"], half["rank"], half["sym"], half["x"], half["y"])
    m.forward(False); m.zero_grads(); m.loss_backward()
    g_dup = m.get_grads()
    one = {k: v[:64] for k, v in b.items()}
    m.set_batch_raw(one["raw"], one["player"], one["rank"], one["sym"], one["x"], one["y"])
    m.forward(False); m.zero_grads(); m.loss_backward()
    g_one = m.get_grads()
    assert U.rel(g_dup, g_one) < 1e-3
    assert np.linalg.norm(g_full) > 0
    l0 = m.train_step(0.02); l1 = m.train_step(0.02); l2 = m.train_step(0.02)
    assert l2 < l0


def test_gradient_vector_after_fused_step(golden):
    """dg_train_step leaves the conv gradients in the accumulation buffers (the SGD kernel reads them there); the Torch7-order
    vector is built on demand.  With rate 0 the parameters do not move, so the vector read after a fused step must equal the one
    of the separate calls zero_grads / forward / loss_backward on the same batch -- twice in a row (graph replay) and after
    switching between the two call styles."""
    from oracle import net
    L, C, n = 5, 128, 16
    flat = net.init_params(L, C, 21, "he")
    m = _model(L, C, n, flat)
    b = U.sample(golden, n, 31)
    m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    m.forward(False); m.zero_grads(); loss_sep = m.loss_backward()
    g_sep = m.get_grads()
    assert np.linalg.norm(g_sep) > 0
    for _ in range(2):
        m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
        loss_fused = m.train_step(0.0)
        g_fused = m.get_grads()
        assert abs(loss_fused - loss_sep) < 1e-5
        assert U.rel(g_fused, g_sep) < 1e-4                      # same kernels, only the summation order of the L2 reductions differs
    assert np.array_equal(m.get_params(), flat.numpy().astype(np.float32))
    m.forward(False); m.zero_grads(); m.loss_backward()
    assert U.rel(m.get_grads(), g_sep) < 1e-4


@pytest.mark.parametrize("prec", [0, 1, 2])
def test_begin_result_pipeline_equals_synchronous_steps(golden, prec):
    """dg_train_step_begin / dg_model_set_batch_raw (next minibatch, copied on the second stream into the other record buffer while
    the step runs) / dg_train_step_result: the same losses and the same parameters as staging and stepping one after the other,
    over six DIFFERENT minibatches of two sizes."""
    from deepgo_b200.host import engine
    from oracle import net
    L, C = 4, 64
    flat = net.init_params(L, C, 17, "he").numpy().astype(np.float32)
    batches = [U.sample(golden, 9 if i % 2 else 12, 300 + i) for i in range(6)]
    stage = lambda m, b: m.set_batch_raw(b["raw"], b["player"], b["rank"], b["sym"], b["x"], b["y"])
    ms = engine.Model(U.ctx(), L, C, prec, max_batch=12); ms.set_params(flat)
    want = []
    for b in batches:
        stage(ms, b)
        want.append(ms.train_step(0.01))
    mp = engine.Model(U.ctx(), L, C, prec, max_batch=12); mp.set_params(flat)
    got = []
    stage(mp, batches[0])
    for i in range(len(batches)):
        mp.train_step_begin(0.01)
        if i + 1 < len(batches):
            stage(mp, batches[i + 1])            # lands in the other buffer while step i runs
        got.append(mp.train_step_result())
    tol = 1e-5 if prec == 0 else 2e-3            # the gradients are summed by atomics: the order differs run to run, and six steps amplify it
    assert np.allclose(got, want, atol=tol), (got, want)
    assert np.abs(mp.get_params() - ms.get_params()).max() < (1e-5 if prec == 0 else 5e-3)
    import deepgo_b200
    mp.train_step_begin(0.0)
    with pytest.raises(deepgo_b200.DeepGoError):         # one step may be outstanding
        mp.train_step_begin(0.0)
    mp.train_step_result()
    with pytest.raises(deepgo_b200.DeepGoError):
        mp.train_step_result()
