"""Thin numpy-facing wrapper over the C ABI (include/deepgo.h).  Every call goes through
libdeepgo_b200.so; nothing here computes."""
import numpy as np

import deepgo_b200 as _dg
from deepgo_b200 import ffi, check, DG_PREC_FP32, DG_PREC_BF16, DG_PREC_TF32, DG_FMT_F32_NCHW, DG_FMT_BITPLANES  # noqa: F401


KERNEL_KINDS = ["encode", "conv_first", "conv", "dgrad", "wgrad", "wgrad_reduce", "bias_grad", "head", "sgd", "pack",
                "allreduce", "other"]


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, ffi.cast("const uint8_t *", a.ctypes.data)


def _opt_u8(a):
    if a is None:
        return None, ffi.NULL
    return _u8(a)


def device_count():
    n = ffi.new("int *")
    check(_dg.lib().dg_device_count(n))
    return n[0]


def encode_labels(x, y, sym=None):
    """0-based labels of 1-based moves (dataloader.lua:89) under symmetry `sym`."""
    xa, xp = _u8(x); ya, yp = _u8(y); sa, sp = _opt_u8(sym)
    out = np.empty(len(xa), np.int32)
    check(_dg.lib().dg_encode_labels(xp, yp, sp, len(xa), ffi.cast("int32_t *", out.ctypes.data)))
    return out


class Context:
    def __init__(self, device=0):
        """device: an int, or a list of ints for several GPUs inside this process (dg_ctx_create_multi: the reference's
        numGPUs option, experiments.lua:10,155-168)."""
        self.L = _dg.lib()
        p = ffi.new("dg_ctx **")
        if isinstance(device, (list, tuple)):
            devs = ffi.new("int[]", [int(d) for d in device])
            check(self.L.dg_ctx_create_multi(len(device), devs, p))
            self.devices = list(device)
            device = self.devices[0]
        else:
            check(self.L.dg_ctx_create(device, p))
            self.devices = [device]
        self.h = ffi.gc(p[0], self.L.dg_ctx_destroy)
        self.device = device

    def synchronize(self):
        check(self.L.dg_ctx_synchronize(self.h))

    def launch_count(self):
        n = ffi.new("long long *")
        check(self.L.dg_ctx_launch_count(self.h, n))
        return int(n[0])

    def timer_start(self):
        check(self.L.dg_timer_start(self.h))

    def timer_stop(self):
        ms = ffi.new("float *")
        check(self.L.dg_timer_stop(self.h, ms))
        return float(ms[0])

    def prof_begin(self):
        check(self.L.dg_prof_begin(self.h))

    def prof_end(self):
        """-> {kind: (total_ms, launches)} for the launches since prof_begin"""
        check(self.L.dg_prof_end(self.h))
        out = {}
        for k, name in enumerate(KERNEL_KINDS):
            ms = ffi.new("float *"); n = ffi.new("long long *")
            check(self.L.dg_prof_get(self.h, k, ms, n))
            if n[0]:
                out[name] = (float(ms[0]), int(n[0]))
        return out

    def flush_l2(self):
        check(self.L.dg_ctx_flush_l2(self.h))

    def init_dp(self, rank, world, uid):
        check(self.L.dg_ctx_init_dp(self.h, rank, world, ffi.from_buffer(uid)))

    def encode(self, raw, player, rank, sym=None, fmt=DG_FMT_F32_NCHW):
        """preprocess() over a batch (dataloader.lua:50-92): -> float32[n,37,19,19] or uint32[n,37,12] bit-planes."""
        ra, rp = _u8(raw); pa, pp = _u8(player); ka, kp = _u8(rank); sa, sp = _opt_u8(sym)
        n = len(pa)
        out = np.empty((n, 37, 19, 19), np.float32) if fmt == DG_FMT_F32_NCHW else np.empty((n, 37, 12), np.uint32)
        check(self.L.dg_encode(self.h, rp, pp, kp, sp, n, ffi.cast("void *", out.ctypes.data), fmt))
        return out

    def encode_noalloc(self, raw, player, rank, sym=None, fmt=DG_FMT_F32_NCHW):
        """dg_encode_noalloc: same result as encode() for up to 64 positions, staged through the library's fixed
        device array instead of an allocated buffer (for callers that must not cudaMalloc)."""
        ra, rp = _u8(raw); pa, pp = _u8(player); ka, kp = _u8(rank); sa, sp = _opt_u8(sym)
        n = len(pa)
        out = np.empty((n, 37, 19, 19), np.float32) if fmt == DG_FMT_F32_NCHW else np.empty((n, 37, 12), np.uint32)
        check(self.L.dg_encode_noalloc(self.h, rp, pp, kp, sp, n, ffi.cast("void *", out.ctypes.data), fmt))
        return out

    def summarize_boards(self, stones):
        sa, sp = _u8(stones)
        n = sa.reshape(-1, 19, 19).shape[0]
        out = np.empty((n, 7, 19, 19), np.uint8)
        check(self.L.dg_summarize_boards(self.h, sp, ffi.cast("uint8_t *", out.ctypes.data), n))
        return out


def _update_boards(self, stones, age, x, y, player):
    """dg_update_boards: in-place update_board (makedata.lua:329-354) of n boards; x, y 1-based."""
    assert stones.dtype == np.uint8 and age.dtype == np.uint8 and stones.flags.c_contiguous and age.flags.c_contiguous
    xa, xp = _u8(x); ya, yp = _u8(y); pa, pp = _u8(player)
    check(self.L.dg_update_boards(self.h, ffi.cast("uint8_t *", stones.ctypes.data), ffi.cast("uint8_t *", age.ctypes.data), xp, yp, pp, len(xa)))


Context.update_boards = _update_boards


def dp_unique_id():
    buf = bytearray(128)
    check(_dg.lib().dg_dp_unique_id(ffi.from_buffer(buf)))
    return bytes(buf)


class Model:
    """getBasicModel (experiments.lua:133-153) + its flat parameter / gradient vectors (:105-107)."""

    def __init__(self, ctx, num_layers, channel_size, precision=DG_PREC_BF16, max_batch=128, in_planes=37, k_first=5):
        self.ctx, self.L = ctx, ctx.L
        p = ffi.new("dg_model **")
        check(self.L.dg_model_create(ctx.h, num_layers, channel_size, in_planes, k_first, precision, max_batch, p))
        self.h = ffi.gc(p[0], self.L.dg_model_destroy)
        n = ffi.new("size_t *")
        check(self.L.dg_model_param_count(self.h, n))
        self.P = int(n[0])
        self.n = 0

    def set_params(self, flat):
        a = np.ascontiguousarray(flat, dtype=np.float32)
        check(self.L.dg_model_set_params(self.h, ffi.cast("const float *", a.ctypes.data), a.size))

    def _get(self, fn):
        out = np.empty(self.P, np.float32)
        check(fn(self.h, ffi.cast("float *", out.ctypes.data), out.size))
        return out

    def get_params(self):
        return self._get(self.L.dg_model_get_params)

    def get_grads(self):
        return self._get(self.L.dg_model_get_grads)

    def zero_grads(self):
        check(self.L.dg_zero_grads(self.h))

    def set_global_batch(self, gb):
        check(self.L.dg_model_set_global_batch(self.h, gb))

    def set_batch_raw(self, raw, player, rank, sym=None, x=None, y=None):
        ra, rp = _u8(raw); pa, pp = _u8(player); ka, kp = _u8(rank)
        sa, sp = _opt_u8(sym); xa, xp = _opt_u8(x); ya, yp = _opt_u8(y)
        self.n = len(pa)
        check(self.L.dg_model_set_batch_raw(self.h, rp, pp, kp, sp, xp, yp, self.n))

    def encode_resident(self):
        check(self.L.dg_model_encode_resident(self.h))

    def set_batch_planes(self, planes, targets0=None):
        a = np.ascontiguousarray(planes, dtype=np.float32)
        self.n = a.shape[0]
        if targets0 is None:
            tp = ffi.NULL
        else:
            t = np.ascontiguousarray(targets0, dtype=np.int32)
            tp = ffi.cast("const int32_t *", t.ctypes.data)
        check(self.L.dg_model_set_batch_planes(self.h, ffi.cast("const float *", a.ctypes.data), tp, self.n))

    def forward(self, want_logp=True):
        if not want_logp:
            check(self.L.dg_forward(self.h, ffi.NULL))
            return None
        out = np.empty((self.n, 361), np.float32)
        check(self.L.dg_forward(self.h, ffi.cast("float *", out.ctypes.data)))
        return out

    def loss_backward(self):
        loss = ffi.new("float *")
        check(self.L.dg_loss_backward(self.h, loss))
        return float(loss[0])

    def allreduce_grads(self):
        check(self.L.dg_dp_allreduce_grads(self.h))

    def sgd_step(self, rate):
        check(self.L.dg_sgd_step(self.h, rate))

    def train_step(self, rate, want_loss=True):
        if not want_loss:
            check(self.L.dg_train_step(self.h, rate, ffi.NULL))
            return None
        loss = ffi.new("float *")
        check(self.L.dg_train_step(self.h, rate, loss))
        return float(loss[0])

    def train_step_begin(self, rate):
        """queue eval() + optimizer:step; stage the next minibatch while it runs; collect the cost with train_step_result()"""
        check(self.L.dg_train_step_begin(self.h, rate))

    def train_step_result(self):
        loss = ffi.new("float *")
        check(self.L.dg_train_step_result(self.h, loss))
        return float(loss[0])

    def argmax(self):
        out = np.empty(self.n, np.int32)
        check(self.L.dg_argmax(self.h, ffi.cast("int32_t *", out.ctypes.data)))
        return out

    def eval_batch(self):
        e = ffi.new("int *"); c = ffi.new("float *")
        check(self.L.dg_eval_batch(self.h, e, c))
        return int(e[0]), float(c[0])

    def save_checkpoint(self, path, meta=None):
        """dg_checkpoint_save: flat parameters + key=value header lines (meta: dict of scalars / strings)."""
        text = "".join("%s=%s\n" % (k, v) for k, v in (meta or {}).items())
        check(self.L.dg_checkpoint_save(self.h, str(path).encode(), text.encode()))

    def predict_sym8(self, want_probs=True):
        probs = np.empty((self.n, 361), np.float32) if want_probs else None
        am = np.empty(self.n, np.int32)
        check(self.L.dg_predict_sym8(self.h, ffi.cast("float *", probs.ctypes.data) if want_probs else ffi.NULL,
                                     ffi.cast("int32_t *", am.ctypes.data)))
        return probs, am


def read_checkpoint(path):
    """dg_checkpoint_read -> (meta dict of strings, float32 parameters)."""
    L = _dg.lib()
    n = ffi.new("size_t *")
    meta = ffi.new("char[]", 1 << 16)
    check(L.dg_checkpoint_read(str(path).encode(), meta, 1 << 16, ffi.NULL, 0, n))
    params = np.empty(int(n[0]), np.float32)
    check(L.dg_checkpoint_read(str(path).encode(), ffi.NULL, 0, ffi.cast("float *", params.ctypes.data), params.size, n))
    d = {}
    for line in ffi.string(meta).decode().splitlines():
        if "=" in line:
            k, v = line.split("=", 1)
            d[k] = v
    return d, params


def read_reference_model(path):
    """dg_t7_read_model: a checkpoint the reference wrote with torch.save -> (float32 parameters, rate, rate_decay, iterations);
    scalars the file does not hold come back as None."""
    L = _dg.lib()
    n = ffi.new("size_t *")
    rate, decay, it = ffi.new("double *", float("nan")), ffi.new("double *", float("nan")), ffi.new("long long *", -1)
    check(L.dg_t7_read_model(str(path).encode(), ffi.NULL, 0, n, ffi.NULL, ffi.NULL, ffi.NULL))
    params = np.empty(int(n[0]), np.float32)
    check(L.dg_t7_read_model(str(path).encode(), ffi.cast("float *", params.ctypes.data), params.size, n, rate, decay, it))
    nn = lambda v: None if v != v else v
    return params, nn(rate[0]), nn(decay[0]), (None if it[0] < 0 else int(it[0]))
