"""``keras.layers`` / ``keras.Model`` stand-ins on numpy -- TEST INFRASTRUCTURE ONLY.

Keras behaviour encoded here (each is the documented default the reference leans on):

* ``Layer.__call__``: ``build(input_shape)`` once, on the first call, then ``call(inputs, *args, **kwargs)``.
  ``add_weight(name, shape, initializer, ...)`` returns the variable (here: an ndarray the generator overwrites).
* ``Dense(units, activation)``: ``activation(inputs . kernel + bias)`` over the LAST axis; kernel ``[in, units]``;
  ``activation`` may be a name, None or any callable -- including a Layer instance (``Dense(h, activation=PReLU())``).
* ``Embedding(input_dim, output_dim)``: a non-integer input is cast to int32 (truncation), then a row lookup;
  an id outside ``[0, input_dim)`` raises, as TensorFlow's CPU kernel does.
* ``PReLU()``: ``alpha`` has the input's shape minus the batch axis (``shared_axes=None``), initial value 0;
  ``relu(x) - alpha * relu(-x)``.
* ``BatchNormalization(axis=-1, epsilon=1e-3, center=True, scale=True)`` outside ``fit`` (``training`` None/False):
  ``(x - moving_mean) * gamma / sqrt(moving_variance + epsilon) + beta`` with gamma/beta only when scale/center.
* ``Dropout``: identity outside training.
"""
from __future__ import annotations

import numpy as np

from . import ops as O

FLOATX = [np.float64]      # the dtype add_weight draws in; the generator sets float32 for the fp32 run


def l2(*a, **k):
    return None


def get_activation(a):
    if a is None or a == "linear":
        return lambda x: x
    if callable(a):
        return a
    return {"relu": O.relu, "sigmoid": O.sigmoid, "tanh": lambda x: O._t(np.tanh(np.asarray(x))),
            "softmax": O.softmax,
            "hard_sigmoid": lambda x: O._t(np.clip(0.2 * np.asarray(x) + 0.5, 0.0, 1.0))}[a]      # keras.backend.hard_sigmoid


class Layer:
    def __init__(self, trainable=True, name=None, dtype=None, **kwargs):
        self.built = False
        self.name = name or type(self).__name__
        self._weights = []

    def add_weight(self, name=None, shape=None, dtype=None, initializer=None, regularizer=None, trainable=True,
                   **kwargs):
        init = initializer if initializer is not None else O.glorot_uniform()
        if isinstance(init, type):
            init = init()
        w = np.asarray(init(tuple(int(s) for s in shape), FLOATX[0]), dtype=FLOATX[0])
        self._weights.append((name, w))
        return w

    def build(self, input_shape):
        self.built = True

    def call(self, inputs, *args, **kwargs):
        return inputs

    @staticmethod
    def _shape_of(x):
        if isinstance(x, dict):
            return {k: Layer._shape_of(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return [Layer._shape_of(v) for v in x]
        return tuple(np.shape(x))

    def __call__(self, inputs, *args, **kwargs):
        if not self.built:
            self.build(self._shape_of(inputs))
            self.built = True
        return self.call(inputs, *args, **kwargs)

    def get_config(self):
        return {}


class Model(Layer):
    pass


class Dense(Layer):
    def __init__(self, units, activation=None, use_bias=True, **kwargs):
        super().__init__(**kwargs)
        self.units = int(units)
        self.activation = get_activation(activation)
        self.use_bias = use_bias

    def build(self, input_shape):
        self.kernel = self.add_weight("kernel", (input_shape[-1], self.units), initializer=O.glorot_uniform())
        self.bias = self.add_weight("bias", (self.units,), initializer=O.Zeros()) if self.use_bias else None

    def call(self, inputs, *args, **kwargs):
        y = np.tensordot(np.asarray(inputs), self.kernel, axes=([-1], [0]))
        if self.bias is not None:
            y = y + self.bias
        return self.activation(O._t(y))


class Embedding(Layer):
    def __init__(self, input_dim, output_dim, **kwargs):
        super().__init__(**kwargs)
        self.input_dim, self.output_dim = int(input_dim), int(output_dim)

    def build(self, input_shape):
        lim = 0.05     # keras 'uniform' initialiser
        self.embeddings = self.add_weight(
            "embeddings", (self.input_dim, self.output_dim),
            initializer=lambda s, d: O._rng.uniform(-lim, lim, size=s).astype(d))

    def call(self, inputs, *args, **kwargs):
        idx = np.asarray(inputs)
        if idx.dtype not in (np.int32, np.int64):
            idx = idx.astype(np.int32)
        if idx.size and (idx.min() < 0 or idx.max() >= self.input_dim):
            raise IndexError("Embedding: indices out of range [0, %d)" % self.input_dim)
        return O._t(self.embeddings[idx])


class PReLU(Layer):
    def __init__(self, shared_axes=None, **kwargs):
        super().__init__(**kwargs)
        assert shared_axes is None

    def build(self, input_shape):
        self.alpha = self.add_weight("alpha", tuple(input_shape[1:]), initializer=O.Zeros())

    def call(self, inputs, *args, **kwargs):
        x = np.asarray(inputs)
        return O._t(np.maximum(x, 0) - self.alpha * np.maximum(-x, 0))


class BatchNormalization(Layer):
    def __init__(self, axis=-1, momentum=0.99, epsilon=1e-3, center=True, scale=True, **kwargs):
        super().__init__(**kwargs)
        assert axis == -1
        self.epsilon, self.center, self.scale = epsilon, center, scale

    def build(self, input_shape):
        n = (input_shape[-1],)
        self.gamma = self.add_weight("gamma", n, initializer=lambda s, d: np.ones(s, d)) if self.scale else None
        self.beta = self.add_weight("beta", n, initializer=O.Zeros()) if self.center else None
        self.moving_mean = self.add_weight("moving_mean", n, initializer=O.Zeros())
        self.moving_variance = self.add_weight("moving_variance", n, initializer=lambda s, d: np.ones(s, d))

    def call(self, inputs, training=None, *args, **kwargs):
        if training:
            raise NotImplementedError("tf_shim: BatchNormalization is inference-only")
        x = np.asarray(inputs)
        inv = 1.0 / np.sqrt(self.moving_variance + self.epsilon)
        if self.gamma is not None:
            inv = inv * self.gamma
        y = (x - self.moving_mean) * inv
        if self.beta is not None:
            y = y + self.beta
        return O._t(y)


class Dropout(Layer):
    def __init__(self, rate=0.0, seed=None, **kwargs):
        super().__init__(**kwargs)

    def call(self, inputs, training=None, *args, **kwargs):
        if training:
            raise NotImplementedError("tf_shim: Dropout is inference-only")
        return inputs


class Flatten(Layer):
    def call(self, inputs, *args, **kwargs):
        x = np.asarray(inputs)
        return O._t(x.reshape(x.shape[0], -1))


class Concatenate(Layer):
    def __init__(self, axis=-1, **kwargs):
        super().__init__(**kwargs)
        self.axis = axis

    def call(self, inputs, *args, **kwargs):
        return O.concat(inputs, self.axis)


class Add(Layer):
    def call(self, inputs, *args, **kwargs):
        return O._t(sum(np.asarray(v) for v in inputs))


class Activation(Layer):
    def __init__(self, activation, **kwargs):
        super().__init__(**kwargs)
        self.activation = get_activation(activation)

    def call(self, inputs, *args, **kwargs):
        return self.activation(inputs)


class Lambda(Layer):
    def __init__(self, function, **kwargs):
        super().__init__(**kwargs)
        self.function = function

    def call(self, inputs, *args, **kwargs):
        return self.function(inputs)


class GRUCell(Layer):
    """Only the constructor behaviour the reference's AUGRUCell (DL/layer/activation.py:91-100) runs into: under TF 2,
    ``tensorflow.python.keras.layers.GRUCell`` is recurrent_v2.GRUCell, whose __init__(units, activation='tanh',
    recurrent_activation='sigmoid', ...) assigns ``self.activation`` / ``self.recurrent_activation`` from ITS OWN arguments.
    AUGRUCell sets both attributes first and then calls ``super().__init__(units, **kwargs)`` without forwarding them, so the
    base class overwrites them with tanh / sigmoid whatever AUGRUCell was given."""

    def __init__(self, units, activation="tanh", recurrent_activation="sigmoid", use_bias=True, **kwargs):
        super().__init__(**kwargs)
        self.units = units
        self.activation = get_activation(activation)
        self.recurrent_activation = get_activation(recurrent_activation)
        self.use_bias = use_bias
