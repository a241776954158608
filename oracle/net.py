"""CPU oracle: PyTorch-CPU restatement of deep-go's policy net, loss, backward and SGD.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
Parity: UNPINNED.  The arithmetic lives in un-vendored, un-pinned Torch7 rocks (nn, torch; no
rockspec in the reference) and the reference has no test for it; this file restates the published
module semantics at the reference's call sites:
  experiments.lua:88-94,133-153  model construction      train.lua:4-12   eval (fwd, NLL, bwd)
  experiments.lua:45             ClassNLLCriterion       train.lua:14-45  eval_validation
  optimizer.lua:24-27            SGD step                experiments.lua:105-107  flat parameter vector
It is self-checked by autograd + finite differences in tests/test_oracle_net.py.

`quant="bf16"` rounds the same tensors the B200 bf16 path rounds (weights, stored activations h_l,
stored dZ_l) so the GPU can be compared tightly; `quant="tf32"` does the same for the TF32 path (weight operands, stored
activations and stored dZ rounded to 10 mantissa bits; bias and sums in full precision); `quant=None` is the reference
fp64 arithmetic.
"""
import math
import numpy as np
import torch
import torch.nn.functional as F

BOARD = 19
POINTS = 361


def layer_specs(num_layers, channel_size, in_planes=37, k_first=5):
    """(cin, cout, k) per layer: kernels={5,3,..}, channels={37,C,..,C,1} (experiments.lua:88-93)."""
    chans = [in_planes] + [channel_size] * (num_layers - 1) + [1]
    ks = [k_first] + [3] * (num_layers - 1)
    return [(chans[i], chans[i + 1], ks[i]) for i in range(num_layers)]


def param_count(num_layers, channel_size, in_planes=37, k_first=5):
    return sum(co * ci * k * k + co + co * POINTS for ci, co, k in layer_specs(num_layers, channel_size, in_planes, k_first))


def split_params(flat, num_layers, channel_size, in_planes=37, k_first=5):
    """Views (W[co,ci,k,k], b[co], a[co,19,19]) into the Torch7 flat vector: per layer W, b (conv), then
    the Add bias (module order of experiments.lua:137-145; getParameters flattens in module order)."""
    out, o = [], 0
    for ci, co, k in layer_specs(num_layers, channel_size, in_planes, k_first):
        nw = co * ci * k * k
        W = flat[o:o + nw].view(co, ci, k, k); o += nw
        b = flat[o:o + co]; o += co
        a = flat[o:o + co * POINTS].view(co, BOARD, BOARD); o += co * POINTS
        out.append((W, b, a))
    assert o == flat.numel()
    return out


def init_params(num_layers, channel_size, seed, kind="torch7", dtype=torch.float64):
    """kind 'torch7': U(+-1/sqrt(cin*k*k)) for W,b and U(+-1/sqrt(cout*361)) for the Add bias (nn reset()s);
    kind 'he': N(0, 2/fan_in) weights, small biases (keeps the signal alive over 12 layers)."""
    rng = np.random.default_rng(seed)
    parts = []
    for ci, co, k in layer_specs(num_layers, channel_size):
        fan = ci * k * k
        if kind == "torch7":
            s = 1.0 / math.sqrt(fan)
            parts += [rng.uniform(-s, s, co * fan), rng.uniform(-s, s, co)]
            s2 = 1.0 / math.sqrt(co * POINTS)
            parts.append(rng.uniform(-s2, s2, co * POINTS))
        else:
            parts += [rng.normal(0, math.sqrt(2.0 / fan), co * fan), rng.normal(0, 0.05, co)]
            parts.append(rng.normal(0, 0.05, co * POINTS))
    return torch.from_numpy(np.concatenate(parts)).to(dtype)


def _q(t, quant):
    if quant == "bf16":
        return t.to(torch.bfloat16).to(t.dtype)
    if quant == "tf32":
        # round to nearest, ties away from zero, to 10 mantissa bits (cvt.rna.tf32.f32): add half an ulp of the kept part to the
        # magnitude and clear the 13 dropped bits
        i = t.to(torch.float32).contiguous().view(torch.int32)
        i = (i + 0x1000) & ~0x1FFF
        return i.view(torch.float32).to(t.dtype)
    return t


def forward(flat, x, num_layers, channel_size, quant=None):
    """x[n,37,19,19] -> (logp[n,361], cache). Every layer incl. the last: pad, conv, +b[c], +a[c,x,y], ReLU
    (experiments.lua:135-147); then Reshape(361)+LogSoftMax (:150-151)."""
    layers = split_params(flat, num_layers, channel_size)
    h = x.to(flat.dtype)
    hs, zs = [h], []
    for i, (W, b, a) in enumerate(layers):
        k = W.shape[-1]
        Wq = _q(W, quant) if i < len(layers) - 1 else W         # the 1-channel head keeps fp32 weights on the GPU
        bias = b.view(-1, 1, 1) + a                               # per-channel + per-point bias (experiments.lua:138,143-145)
        if i < len(layers) - 1 and quant == "bf16":
            bias = _q(bias, quant)                                # the GPU's bf16 epilogue operand; the head keeps fp32
        z = F.conv2d(h, Wq, None, padding=(k - 1) // 2)
        if i < len(layers) - 1 and quant == "bf16":
            z = _q(z, quant)                                      # the GPU rounds the fp32 accumulator to bf16, then adds the
        z = z + bias.unsqueeze(0)                                 # bf16 bias with one more rounding (packed fma + ReLU)
        zs.append(z)
        h = torch.relu(z)
        if i < len(layers) - 1:
            h = _q(h, quant)
        hs.append(h)
    logits = h.reshape(h.shape[0], POINTS)
    return F.log_softmax(logits, dim=1), (hs, zs)


def loss_and_grads(flat, x, targets0, num_layers, channel_size, quant=None, grad_scale=None):
    """train.lua:4-12 eval(): mean NLL (ClassNLLCriterion sizeAverage) + explicit module-by-module backward.
    targets0 are 0-based.  Returns (loss, grads_flat, logp).  grad_scale overrides 1/n (data-parallel shards
    use 1/global_batch)."""
    n = x.shape[0]
    logp, (hs, zs) = forward(flat, x, num_layers, channel_size, quant)
    idx = torch.arange(n)
    t = torch.as_tensor(np.asarray(targets0), dtype=torch.long)
    loss = -logp[idx, t].mean()
    scale = (1.0 / n) if grad_scale is None else grad_scale
    # LogSoftMax backward of dlogp = -scale at target: dlogit = scale * (softmax - onehot)
    dlogit = torch.exp(logp) * scale
    dlogit[idx, t] -= scale
    dh = dlogit.view(n, 1, BOARD, BOARD)
    layers = split_params(flat, num_layers, channel_size)
    grads = torch.zeros_like(flat)
    gl = split_params(grads, num_layers, channel_size)
    for i in range(num_layers - 1, -1, -1):
        W, b, a = layers[i]
        k = W.shape[-1]
        dz = dh * (zs[i] > 0).to(dh.dtype)                       # ReLU = Threshold(0,0), strict >
        if i < num_layers - 1:
            dz = _q(dz, quant)
        gW, gb, ga = gl[i]
        ga.copy_(dz.sum(0))                                       # Add: gradBias = sum over batch
        gb.copy_(dz.sum((0, 2, 3)))                               # conv bias: sum over batch and space
        gW.copy_(torch.nn.grad.conv2d_weight(hs[i], W.shape, dz, padding=(k - 1) // 2))
        if i > 0:
            Wq = _q(W, quant) if i < num_layers - 1 else W
            dh = torch.nn.grad.conv2d_input(hs[i].shape, Wq, dz, padding=(k - 1) // 2)
    return float(loss), grads, logp


def sgd_step(flat, grads, rate, rate_decay):
    """optimizer.lua:24-27: p += -rate*g ; rate *= (1-decay)  (decay applied after use). Returns new rate."""
    flat.add_(grads, alpha=-rate)
    return rate * (1.0 - rate_decay)


def eval_validation(flat, x, targets0, batch_size, num_layers, channel_size, quant=None):
    """train.lua:14-45 including its aggregate quirks: `for i=1,V/batchSize` (float bound -> tail dropped),
    size = range[2]-range[1] = batchSize-1 (:39), divisor = full V (:44)."""
    V = x.shape[0]
    sum_costs, sum_errors = 0.0, 0
    t = torch.as_tensor(np.asarray(targets0), dtype=torch.long)
    for i in range(1, int(V // batch_size) + 1):
        lo, hi = (i - 1) * batch_size, min(i * batch_size, V)
        logp, _ = forward(flat, x[lo:hi], num_layers, channel_size, quant)
        top = logp.argmax(1)
        errors = int((top != t[lo:hi]).sum())
        cost = float(-logp[torch.arange(hi - lo), t[lo:hi]].mean())
        size = (hi) - (lo + 1)
        sum_costs += cost * size
        sum_errors += errors
    return sum_costs / V, 1 - sum_errors / V
