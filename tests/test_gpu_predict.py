"""-m gpu: c5 inference path -- 8-fold symmetry-averaged move probabilities (dg_predict_sym8) vs the oracle
(encode with each symmetry id, forward, map every point back, average).  Symmetry ids 1..7 are this repo's spec
(the reference's transform() is the identity), so this pins kernel vs oracle, not vs the reference."""
import numpy as np
import pytest
import torch

from tests import gpu_util as U

pytestmark = pytest.mark.gpu


def _oracle_sym8(flat, b, L, C, quant):
    from oracle import encode, net
    n = len(b["raw"])
    acc = np.zeros((n, 361))
    for s in range(8):
        x = np.stack([encode.preprocess(b["raw"][i], int(b["player"][i]), int(b["rank"][i]), s) for i in range(n)])
        logp, _ = net.forward(flat, torch.from_numpy(x), L, C, quant=quant)
        p = np.exp(logp.numpy())
        for q in range(361):
            tx, ty = encode.transform_xy(q // 19, q % 19, s)
            acc[:, q] += p[:, 19 * tx + ty] / 8
    return acc


@pytest.mark.parametrize("prec,quant,tol", [(0, None, 1e-6), (1, "bf16", 2e-4)])
def test_sym8_average_matches_oracle(golden, prec, quant, tol):
    from deepgo_b200.host import engine
    from oracle import net
    L, C, n = 4, 64, 7
    b = U.sample(golden, n, 21)
    flat = net.init_params(L, C, 5, "he")
    ref = _oracle_sym8(flat, b, L, C, quant)
    m = engine.Model(U.ctx(), L, C, prec, max_batch=n)
    m.set_params(flat.numpy())
    m.set_batch_raw(b["raw"], b["player"], b["rank"])
    probs, am = m.predict_sym8()
    assert np.abs(probs - ref).max() < tol
    assert np.allclose(probs.sum(1), 1.0, atol=1e-4)
    assert np.array_equal(am, probs.argmax(1))
    # a symmetric net input gives a symmetric answer: averaging is invariant to pre-transforming the record
    from oracle import encode
    raw_t = np.stack([encode.transform(r, 5) for r in b["raw"]])
    m.set_batch_raw(raw_t, b["player"], b["rank"])
    probs_t, _ = m.predict_sym8()
    back = np.stack([encode.transform(probs_t[i].reshape(19, 19), encode.inverse_sym(5)).reshape(-1) for i in range(n)])
    assert np.abs(back - probs).max() < (1e-5 if prec == 0 else 5e-3)
