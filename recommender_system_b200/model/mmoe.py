"""Drop-in for the reference's ``model/mmoe.py`` (DL/model/mmoe.py:10-32): ``mmoe_layer`` then one ``tower_layer`` per task;
returns the list of ``num_tasks`` tensors ``[B, output_dim]``."""
from __future__ import annotations

from ..layer.base import Model
from ..layer.interaction import mmoe_layer, tower_layer


class MMOE(Model):
    def __init__(self, mmoe_hidden_units, num_experts, num_tasks, tower_hidden_units, output_dim, activation="relu",
                 use_expert_bias=True, use_gate_bias=True, precision="tf32", device=None, **kwargs):
        super().__init__(device)
        self.mmoe_hidden_units = mmoe_hidden_units
        self.mmoe_layer = self.track(mmoe_layer(mmoe_hidden_units, num_experts, num_tasks, use_expert_bias, use_gate_bias,
                                                precision=precision, device=self.device))
        self.tower_layer = [self.track(tower_layer(tower_hidden_units, output_dim, activation, precision=precision,
                                                   device=self.device)) for _ in range(num_tasks)]

    def build(self, input_shape):
        if not self.mmoe_layer.built:
            self.mmoe_layer.build(input_shape)
            self.mmoe_layer.built = True
        for t in self.tower_layer:
            if not t.built:
                t.build((None, self.mmoe_hidden_units))
                t.built = True

    def call(self, inputs, training=None, mask=None):
        outs = self.mmoe_layer(inputs)
        return [layer(o) for layer, o in zip(self.tower_layer, outs)]
