"""Keras-like object protocol on torch device memory.

The reference's only seam is ``Layer.__init__(hparams)`` / ``build(input_shape)`` creating named
weights via ``add_weight`` / ``call(inputs)`` invoked as ``obj(x)`` (e.g.
DL/layer/interaction.py:87-106).  This base keeps that protocol; tensors are ``torch.Tensor`` on
CUDA and every ``call`` goes to libctrhot through ctypes.  There is no CPU compute path: calling a
layer without a CUDA device raises.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch

from .. import _lib

_GEN = {"seed": 1234}


def set_seed(seed):
    """Seed for the (otherwise Keras-default) weight initialisers."""
    _GEN["seed"] = int(seed)


def _next_gen():
    g = torch.Generator(device="cpu")
    g.manual_seed(_GEN["seed"])
    _GEN["seed"] += 1
    return g


def default_device():
    return torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")


def require_cuda(t, what):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise _lib.CtrhotError("%s: a CUDA device is required -- ctrhot has no CPU fallback" % what)


def as_device_f32(x, device=None):
    """numpy / list / torch (any float dtype, any device) -> contiguous fp32 tensor on the device."""
    device = device or default_device()
    if isinstance(x, torch.Tensor):
        t = x
    else:
        a = np.asarray(x)
        if a.dtype == object:
            a = a.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype != torch.float32:
        t = t.to(torch.float32)
    t = t.to(device, non_blocking=True)
    return t if _rows_contiguous(t) else t.contiguous()


def _rows_contiguous(t):
    """True when every dim but the first is densely packed (kernels take an explicit leading dimension)."""
    if t.dim() == 0:
        return True
    expect = 1
    for d in range(t.dim() - 1, 0, -1):
        if t.shape[d] != 1 and t.stride(d) != expect:
            return False
        expect *= t.shape[d]
    return t.dim() == 1 and t.stride(0) == 1 or t.dim() > 1 and t.stride(0) >= expect


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return 0 if t is None else t.data_ptr()


def ids_arg(ids):
    """-> (tensor, kind, ld) for an ids matrix given as int32 or as the reference's float columns."""
    if ids.dtype == torch.int32:
        kind = _lib.IDS_I32
    elif ids.dtype == torch.float32:
        kind = _lib.IDS_F32
    else:
        ids = ids.to(torch.int32)
        kind = _lib.IDS_I32
    if ids.dim() == 1:
        ids = ids.unsqueeze(1)
    if ids.stride(-1) != 1:
        ids = ids.contiguous()
    return ids, kind, ids.stride(0) if ids.dim() == 2 else ids.shape[-1]


# ---------------------------------------------------------------- initialisers (Keras defaults)
def init_zeros(shape, gen):
    return torch.zeros(shape, dtype=torch.float32)


def init_ones(shape, gen):
    return torch.ones(shape, dtype=torch.float32)


def init_random_normal(shape, gen, stddev=0.05):
    """tf.random_normal_initializer() defaults: mean 0, stddev 0.05."""
    return torch.randn(shape, generator=gen, dtype=torch.float32) * stddev


def init_uniform(shape, gen, lim=0.05):
    """keras Embedding default 'uniform': U(-0.05, 0.05)."""
    return (torch.rand(shape, generator=gen, dtype=torch.float32) * 2 - 1) * lim


def init_glorot_uniform(shape, gen):
    lim = math.sqrt(6.0 / (shape[0] + shape[-1]))
    return (torch.rand(shape, generator=gen, dtype=torch.float32) * 2 - 1) * lim


class Layer:
    """build()/call()/__call__ with named weights, ``get_weights``/``set_weights``/``load_npz``."""

    def __init__(self, device=None):
        self.built = False
        self.device = torch.device(device) if device is not None else default_device()
        self._weights = OrderedDict()
        self._sublayers = []
        self._err_flag = None
        self._version = 0          # bumped whenever weights change (invalidates packed copies)

    # -- weights
    def add_weight(self, name, shape, initializer=init_zeros, **_ignored):
        t = initializer(tuple(shape), _next_gen()).to(self.device)
        self._weights[name] = t
        return t

    def track(self, layer):
        self._sublayers.append(layer)
        return layer

    def named_weights(self, prefix=""):
        out = OrderedDict()
        for k, v in self._weights.items():
            out[prefix + k] = v
        for i, sub in enumerate(self._sublayers):
            out.update(sub.named_weights("%s%s_%d/" % (prefix, type(sub).__name__, i)))
        return out

    def get_weights(self):
        return [w.detach().cpu().numpy() for w in self.named_weights().values()]

    def set_weights(self, weights):
        cur = self.named_weights()
        if isinstance(weights, dict):
            items = [(k, weights[k]) for k in weights]
        else:
            weights = list(weights)
            if len(weights) != len(cur):
                raise ValueError("expected %d weight arrays, got %d" % (len(cur), len(weights)))
            items = list(zip(cur.keys(), weights))
        for k, w in items:
            if k not in cur:
                raise KeyError(k)
            src = torch.as_tensor(np.asarray(w, dtype=np.float32))
            if tuple(src.shape) != tuple(cur[k].shape):
                raise ValueError("weight %s: expected shape %s, got %s" % (k, tuple(cur[k].shape), tuple(src.shape)))
            cur[k].copy_(src)
        self._bump()

    def load_npz(self, path):
        """Flat .npz keyed by the names used in add_weight (checkpoint import)."""
        with np.load(path) as z:
            self.set_weights({k: z[k] for k in z.files})

    def _bump(self):
        self._version += 1
        for s in self._sublayers:
            s._bump()

    # -- protocol
    def __init_subclass__(cls, **kw):
        """Like Keras, calling build() directly marks the layer built (so __call__ does not re-create weights)."""
        super().__init_subclass__(**kw)
        if "build" in cls.__dict__:
            raw = cls.__dict__["build"]

            def build(self, input_shape=None, _raw=raw):
                _raw(self, input_shape)
                self.built = True
            build.__doc__ = raw.__doc__
            cls.build = build

    def build(self, input_shape=None):
        self.built = True

    def call(self, inputs, *args, **kwargs):
        raise NotImplementedError

    def __call__(self, inputs, *args, **kwargs):
        if not self.built:
            self.build(_shape_of(inputs))
            self.built = True
        return self.call(inputs, *args, **kwargs)

    # -- out-of-range id reporting (CPU-TF-like IndexError; GPU-TF-like zeros when check=False)
    def err_flag(self):
        if self._err_flag is None:
            self._err_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        return self._err_flag

    def check_ids(self):
        rc = _lib.load().ctrhot_check_flag(ptr(self.err_flag()), stream_ptr())
        _lib.check(rc, type(self).__name__)


def _shape_of(x):
    if isinstance(x, (list, tuple)):
        return [_shape_of(t) for t in x]
    if isinstance(x, dict):
        return {k: _shape_of(v) for k, v in x.items()}
    if hasattr(x, "shape"):
        return tuple(x.shape)
    return tuple(np.asarray(x).shape)


class Model(Layer):
    """keras.Model stand-in: ``call(inputs, training=None, mask=None)``; forward only."""

    def predict(self, inputs):
        return self(inputs)
