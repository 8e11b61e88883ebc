"""Drop-in for the reference's ``model/deepCrossing.py`` (DL/model/deepCrossing.py:15-32): EmbedLayer -> [dense, embed] ->
``res_layer_num`` residual units -> ``Dense(1, sigmoid)``."""
from __future__ import annotations

import torch

from .. import _lib
from ..layer.base import Model, ptr, require_cuda, stream_ptr
from ..layer.core import DenseStack, EmbedLayer, dense_params
from ..layer.interaction import ResLayer
from .criteo import split_inputs


class DeepCrossing(Model):
    def __init__(self, feature_columns, k, hidden_units, res_layer_num, embed_dim=8, precision="tf32", check_ids=True,
                 device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        # like the reference, ``k`` is accepted and unused: its EmbedLayer keeps its own default width
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=check_ids, device=self.device))
        self.res_layer = [self.track(ResLayer(hidden_units, precision=precision, device=self.device)) for _ in range(res_layer_num)]
        self.stack = DenseStack(precision)

    def build(self, input_shape):
        n = len(self.dense_feature_columns) + self.embed_layer.F * self.embed_layer.k
        for r in self.res_layer:
            if not r.built:
                r.build((None, n))
                r.built = True
        self.output_layer = dense_params(self, "output_layer", n, 1)

    def call(self, inputs, training=None, mask=None):
        D = len(self.dense_feature_columns)
        dense, ids, _ = split_inputs(inputs, self.device, D)
        require_cuda(dense, "DeepCrossing")
        x = self.embed_layer.gather(ids, dense)                    # [dense | embed], deepCrossing.py:26-28
        for layer in self.res_layer:
            x = layer(x)
        self.stack.set_layers([{"kernel": self.output_layer[0], "bias": self.output_layer[1], "act": _lib.ACT_NONE}], self._version)
        return self.stack.run(x, final_sigmoid=True)
