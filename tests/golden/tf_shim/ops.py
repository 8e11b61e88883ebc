"""``tf.*`` functions the reference's hot path calls, on numpy arrays -- TEST INFRASTRUCTURE ONLY.

Each function states the TensorFlow behaviour it encodes.  Arrays keep whatever float dtype the caller's weights have
(float64 weights give the float64 golden run; float32 weights give the reference's own precision), except where the
reference itself names a dtype (``tf.cast``, ``tf.one_hot``'s float32 default).
"""
from __future__ import annotations

import numpy as np

__all__ = ["matmul", "pow", "reduce_sum", "reduce_mean", "reduce_max", "tensordot", "one_hot", "gather", "transpose",
           "concat", "cast", "expand_dims", "tile", "squeeze", "where", "equal", "not_equal", "ones_like",
           "zeros_like", "reshape", "stack", "convert_to_tensor", "multiply", "sigmoid", "zeros_initializer",
           "random_normal_initializer", "float32", "float64", "int32", "int64", "string", "constant", "zeros",
           "divide", "as_string"]

float32, float64, int32, int64 = np.float32, np.float64, np.int32, np.int64
string = np.str_


class Tensor(np.ndarray):
    """ndarray with the two tf.Tensor methods the reference calls."""

    def get_shape(self):
        return tuple(self.shape)

    def numpy(self):
        return np.asarray(self)


def _t(x):
    return np.asarray(x).view(Tensor)


def _dtype(d):
    return np.dtype({"float32": np.float32, "float64": np.float64, "int32": np.int32, "int64": np.int64}.get(d, d))


def convert_to_tensor(value, dtype=None):
    """A python list of equal-shape tensors is packed along a new axis 0 (tf.convert_to_tensor / ops.pack)."""
    return _t(np.asarray(value, dtype=dtype))


constant = convert_to_tensor


def zeros(shape, dtype=np.float32):
    return _t(np.zeros(shape, dtype=_dtype(dtype)))


def matmul(a, b):
    """Batched matrix product over the last two axes, leading axes broadcast (TF >= 1.14)."""
    return _t(np.matmul(np.asarray(a), np.asarray(b)))


def pow(x, y):  # noqa: A001 - the reference calls tf.pow
    return _t(np.power(np.asarray(x), y))


def multiply(x, y):
    return _t(np.asarray(x) * np.asarray(y))


def divide(x, y):
    return _t(np.asarray(x) / np.asarray(y))


def reduce_sum(x, axis=None, keepdims=False, name=None):
    return _t(np.sum(np.asarray(x), axis=axis, keepdims=keepdims))


def reduce_mean(x, axis=None, keepdims=False, name=None):
    return _t(np.mean(np.asarray(x), axis=axis, keepdims=keepdims))


def reduce_max(x, axis=None, keepdims=False, name=None):
    return _t(np.max(np.asarray(x), axis=axis, keepdims=keepdims))


def tensordot(a, b, axes):
    """``axes=1``: contract a's last axis with b's first; ``axes=(-1, 0)`` the same, spelled out."""
    if isinstance(axes, (tuple, list)):
        axes = ([axes[0]] if np.isscalar(axes[0]) else list(axes[0]), [axes[1]] if np.isscalar(axes[1]) else list(axes[1]))
    return _t(np.tensordot(np.asarray(a), np.asarray(b), axes=axes))


def one_hot(indices, depth, dtype=np.float32):
    """float32 by default; an index outside [0, depth) gives an all-zero row (no error)."""
    idx = np.asarray(indices).astype(np.int64)
    out = np.zeros(idx.shape + (int(depth),), dtype=_dtype(dtype))
    ok = (idx >= 0) & (idx < depth)
    out[ok, idx[ok]] = 1
    return _t(out)


def gather(params, indices, axis=0):
    """Slices along ``axis`` (default 0).  CPU TensorFlow raises on an out-of-range index."""
    params, idx = np.asarray(params), np.asarray(indices)
    if idx.size and (idx.min() < 0 or idx.max() >= params.shape[axis]):
        raise IndexError("tf.gather: index out of range")
    return _t(np.take(params, idx, axis=axis))


def transpose(a, perm=None):
    return _t(np.transpose(np.asarray(a), perm))


def concat(values, axis):
    """Like TensorFlow, every input must have the same rank (numpy agrees) -- and the same dtype: TF refuses mixed
    float32/float64; here the narrower operand is promoted, which is exact."""
    return _t(np.concatenate([np.asarray(v) for v in values], axis=axis))


def stack(values, axis=0):
    return _t(np.stack([np.asarray(v) for v in values], axis=axis)) if isinstance(values, (list, tuple)) else _t(values)


def cast(x, dtype):
    """float -> int truncates toward zero (the C cast TF uses)."""
    return _t(np.asarray(x).astype(_dtype(dtype)))


def expand_dims(x, axis):
    return _t(np.expand_dims(np.asarray(x), axis))


def tile(x, multiples):
    return _t(np.tile(np.asarray(x), tuple(int(m) for m in multiples)))


def squeeze(x, axis=None):
    return _t(np.squeeze(np.asarray(x), axis=axis))


def where(condition, x, y):
    return _t(np.where(np.asarray(condition), np.asarray(x), np.asarray(y)))


def equal(x, y):
    return _t(np.asarray(x) == np.asarray(y))


def not_equal(x, y):
    return _t(np.asarray(x) != np.asarray(y))


def ones_like(x):
    return _t(np.ones_like(np.asarray(x)))


def zeros_like(x):
    return _t(np.zeros_like(np.asarray(x)))


def reshape(x, shape):
    return _t(np.reshape(np.asarray(x), tuple(int(s) for s in shape)))


def sigmoid(x):
    """1 / (1 + exp(-x)), evaluated without overflow on either side."""
    x = np.asarray(x)
    e = np.exp(-np.abs(x))
    return _t(np.where(x >= 0, 1.0 / (1.0 + e), e / (1.0 + e)))


def relu(x):
    return _t(np.maximum(np.asarray(x), 0))


def softmax(logits, axis=-1, name=None):
    """``tf.nn.softmax``: over the LAST axis unless told otherwise; max-subtracted like TF's kernel."""
    x = np.asarray(logits)
    e = np.exp(x - np.max(x, axis=axis, keepdims=True))
    return _t(e / np.sum(e, axis=axis, keepdims=True))


def bias_add(value, bias, data_format=None):
    return _t(np.asarray(value) + np.asarray(bias))


def repeat_elements(x, rep, axis):
    """keras.backend.repeat_elements == np.repeat."""
    return _t(np.repeat(np.asarray(x), int(rep), axis=axis))


def as_string(x):
    """tf.as_string on integers: decimal text."""
    return np.asarray(x).astype(np.int64).astype(str)


def to_hash_bucket_fast(x, num_buckets, name=None):
    raise NotImplementedError("tf_shim: FarmHash Fingerprint64 is not restated here")


# ------------------------------------------------------------------ initialisers (weights are injected afterwards)
_rng = np.random.default_rng(20220806)


class _Init:
    def __init__(self, *a, **k):
        pass


class Zeros(_Init):
    def __call__(self, shape, dtype=np.float64):
        return np.zeros(shape, dtype=dtype)


zeros_initializer = Zeros


class random_normal_initializer(_Init):
    """mean 0, stddev 0.05 (TF's default)."""

    def __call__(self, shape, dtype=np.float64):
        return _rng.normal(0.0, 0.05, size=shape).astype(dtype)


class glorot_uniform(_Init):
    def __call__(self, shape, dtype=np.float64):
        lim = np.sqrt(6.0 / (shape[0] + shape[-1])) if len(shape) >= 2 else 0.05
        return _rng.uniform(-lim, lim, size=shape).astype(dtype)


class glorot_normal(_Init):
    def __call__(self, shape, dtype=np.float64):
        sd = np.sqrt(2.0 / (shape[0] + shape[-1])) if len(shape) >= 2 else 0.05
        return _rng.normal(0.0, sd, size=shape).astype(dtype)
