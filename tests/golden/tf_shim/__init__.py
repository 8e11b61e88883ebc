"""NumPy stand-ins for the TensorFlow / Keras symbols the reference imports -- TEST INFRASTRUCTURE ONLY.

The reference (Hcyand/recommender_system) is pure Python; only its ``import tensorflow`` / ``import keras`` lines stop
its ``call()`` bodies from running in this container (no TensorFlow wheel, no network).  ``install()`` registers
module objects under those names so that the reference's UNMODIFIED files

    algorithm/deep_learning/layer/interaction.py, layer/core.py, layer/utils.py, layer/activation.py
    algorithm/deep_learning/model/{fm,ffm,deepFM,wideDeep,dcn,pnn,din,nfm,afm}.py

import and execute eagerly on ``numpy.ndarray``.  ``tests/golden/make_ref_golden.py`` uses this to emit golden
outputs of the reference's own code (op order, axes, pair order, mask fill, weight shapes) for the oracle tests.

What this pins and what it does not: every line of the reference's Python runs as written; the semantics of the
~40 TF/Keras symbols below are restated from their documented behaviour (``ops.py``, ``layers.py`` say which
default each one encodes).  Anything the reference never reaches on the hot path is a stub that raises on use.
Nothing under ``recommender_system_b200/`` may import this package.
"""
from __future__ import annotations

import sys
import types

from . import layers as L
from . import ops as O


class _Stub(types.ModuleType):
    """A module whose unknown attributes are placeholder classes: importing a name always works (the reference's
    module headers import many symbols its hot path never touches), using one raises."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        full = "%s.%s" % (self.__name__, name)

        def _refuse(*a, **k):
            raise NotImplementedError("tf_shim: %s is not implemented (not on the pinned path)" % full)

        cls = type(name, (), {"__init__": _refuse, "__call__": _refuse})
        setattr(self, name, cls)
        return cls


def _mod(name, **attrs):
    m = _Stub(name)
    m.__dict__.update(attrs)
    m.__path__ = []          # so that "import a.b.c" treats it as a package
    sys.modules[name] = m
    parent, _, leaf = name.rpartition(".")
    if parent:
        setattr(sys.modules[parent], leaf, m)
    return m


_LAYER_NAMES = ("Layer", "Dense", "Dropout", "Embedding", "BatchNormalization", "PReLU", "Flatten", "Concatenate",
                "Add", "Activation", "Lambda", "GRUCell")


def install():
    """Register the stand-in modules.  Idempotent.  Refuses to shadow a real TensorFlow."""
    if "tensorflow" in sys.modules and not isinstance(sys.modules["tensorflow"], _Stub):
        raise RuntimeError("a real tensorflow is already imported; the shim is for boxes without one")
    if isinstance(sys.modules.get("tensorflow"), _Stub):
        return sys.modules["tensorflow"]
    layer_attrs = {n: getattr(L, n) for n in _LAYER_NAMES}
    tf = _mod("tensorflow", **{n: getattr(O, n) for n in O.__all__})
    tf.__version__ = "2.9.0"          # the reference branches on it (layer/sequence.py:268, layer/utils.py); a TF 2 answer
    _mod("tensorflow.nn", softmax=O.softmax, sigmoid=O.sigmoid, relu=O.relu, bias_add=O.bias_add)
    _mod("tensorflow.strings", to_hash_bucket_fast=O.to_hash_bucket_fast, as_string=O.as_string)
    _mod("tensorflow.python")
    _mod("tensorflow.python.ops")
    _mod("tensorflow.python.ops.init_ops", Zeros=O.Zeros, RandomNormal=O.random_normal_initializer,
         glorot_normal_initializer=O.glorot_normal, glorot_uniform_initializer=O.glorot_uniform)
    _mod("tensorflow.python.ops.init_ops_v2", Zeros=O.Zeros, glorot_normal=O.glorot_normal,
         glorot_uniform=O.glorot_uniform)
    for n in ("lookup_ops", "math_ops"):
        _mod("tensorflow.python.ops." + n)
    import numpy as _np
    _mod("tensorflow.python.ops.array_ops", transpose=O.transpose, stack=O.stack, expand_dims=O.expand_dims, tile=O.tile,
         unstack=lambda x, axis=0: [O._t(v) for v in _np.moveaxis(_np.asarray(x), axis, 0)],
         shape=lambda x: _np.asarray(_np.shape(x)), where_v2=O.where)
    _mod("tensorflow.python.util")
    # the three nest functions rnn_augru (DL/layer/activation.py:145-263) uses, for flat (non-nested) structures and lists
    _seq = (list, tuple)
    _mod("tensorflow.python.util.nest", is_nested=lambda x: isinstance(x, _seq),
         flatten=lambda x: [v for y in x for v in ([y] if not isinstance(y, _seq) else y)] if isinstance(x, _seq) else [x],
         map_structure=lambda f, x: type(x)(f(v) for v in x) if isinstance(x, _seq) else f(x),
         pack_sequence_as=lambda ref, flat: type(ref)(flat) if isinstance(ref, _seq) else flat[0])
    _mod("tensorflow.python.framework")
    _mod("tensorflow.python.framework.dtypes")
    _mod("tensorflow.python.framework.ops")
    for root in ("tensorflow.python.keras", "keras"):
        _mod(root, Model=L.Model)
        _mod(root + ".layers", **layer_attrs)
        _mod(root + ".models", Model=L.Model)
        _mod(root + ".regularizers", l2=L.l2)
        _mod(root + ".backend", repeat_elements=O.repeat_elements, zeros_like=O.zeros_like,
             expand_dims=O.expand_dims, dot=O.matmul)
        _mod(root + ".activations", get=L.get_activation)
        _mod(root + ".optimizers")
        _mod(root + ".losses")
        _mod(root + ".engine")
        _mod(root + ".engine.base_layer", Layer=L.Layer)
        _mod(root + ".utils")
        _mod(root + ".utils.generic_utils")
        _mod(root + ".preprocessing")
        _mod(root + ".preprocessing.sequence")
    tf.keras = sys.modules["tensorflow.python.keras"]
    return tf


def uninstall():
    for name in [n for n, m in sys.modules.items() if isinstance(m, _Stub)]:
        del sys.modules[name]
