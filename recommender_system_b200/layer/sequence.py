"""Drop-in for the attention pooling of the reference's ``layer/sequence.py`` (DL/layer/sequence.py:180-290):

``AttentionSequencePoolingLayer(att_hidden_units=(80, 40), att_activation='sigmoid', weight_normalization=False,
return_score=False, supports_masking=False)``; ``call([query [B,1,E], keys [B,T,E], keys_length [B,1]])`` -> ``[B,1,E]``
(``[B,1,T]`` with ``return_score``); with ``supports_masking=True``: ``call([query, keys], mask=[None, key_mask [B,T]])``.

The LocalActivationUnit scores run on the tensor-pipe tower (layer/core.py), the mask / softmax / weighted sum is one launch
(ctrhot_seq_attn_pool_f32).  The reference's ``keys_length`` branch raises as written (``as_list[1]`` subscripts a method,
sequence.py:248); it is implemented here as DeepCTR means it: ``tf.sequence_mask(keys_length, T)``.
"""
from __future__ import annotations

import torch

from .. import _lib
from .base import Layer, as_device_f32, init_glorot_uniform, ptr, require_cuda, stream_ptr
from .core import LocalActivationUnit


class AttentionSequencePoolingLayer(Layer):
    def __init__(self, att_hidden_units=(80, 40), att_activation="sigmoid", weight_normalization=False, return_score=False,
                 supports_masking=False, precision="tf32", device=None, **kwargs):
        super().__init__(device)
        self.att_hidden_units = list(att_hidden_units)
        self.att_activation = att_activation
        self.weight_normalization = bool(weight_normalization)
        self.return_score = bool(return_score)
        self.supports_masking = bool(supports_masking)
        # LocalActivationUnit(hidden, activation, l2_reg=0, dropout_rate=0, use_bn=False, seed=1024) -- sequence.py:238-239
        self.local_att = self.track(LocalActivationUnit(self.att_hidden_units, att_activation, 0, 0, False, seed=1024,
                                                        precision=precision, device=self.device))

    def build(self, input_shape):
        if not self.supports_masking:
            if not isinstance(input_shape, (list, tuple)) or len(input_shape) != 3:
                raise ValueError("A `AttentionSequencePoolingLayer` layer should be called on a list of 3 inputs")
        if not self.local_att.built:
            self.local_att.build(list(input_shape[:2]))
            self.local_att.built = True

    def call(self, inputs, mask=None, training=None, **kwargs):
        if self.supports_masking:
            if mask is None:
                raise ValueError("When supports_masking=True,input must support masking")
            queries, keys = inputs
            key_mask, lengths = mask[-1], None
        else:
            queries, keys, lengths = inputs
            key_mask = None
        keys = as_device_f32(keys, self.device)
        require_cuda(keys, "AttentionSequencePoolingLayer")
        if keys.stride(2) != 1 or keys.stride(1) != keys.shape[2]:
            keys = keys.contiguous()
        B, T, E = keys.shape
        score = self.local_att([queries, keys], training=training).reshape(B, T)
        if key_mask is not None:
            m = torch.as_tensor(key_mask).to(self.device)
            m = (m != 0).to(torch.uint8).reshape(B, T).contiguous()
            mp, mld, lp = ptr(m), T, None
        else:
            ln = torch.as_tensor(lengths).to(self.device).reshape(B).to(torch.int32).contiguous()
            mp, mld, lp = None, 0, ptr(ln)
        out = torch.empty((B, 1, T if self.return_score else E), dtype=torch.float32, device=self.device)
        _lib.call("ctrhot_seq_attn_pool_f32", ptr(score), score.stride(0), ptr(keys), keys.stride(0), mp, mld, lp, B, T, E,
                  int(self.weight_normalization), int(self.return_score), ptr(out), out.stride(0), stream_ptr())
        return out


class AUGRU(Layer):
    """``AUGRU(units, gru_type='AUGRU', return_sequence=False, activation='tanh', recurrent_activation='hard_sigmoid')`` --
    DL/layer/sequence.py:293-395 around ``AUGRUCell`` (DL/layer/activation.py:91-142); ``call(inputs [B,T,E], att_score [B,T,1])``
    -> last output ``[B, units]`` or, with ``return_sequence``, ``[B, T, units]``.  Weights ``kernel [E, 3 units]``,
    ``recurrent_kernel [units, 3 units]`` (z | r | h blocks, no biases).  One launch (ctrhot_augru_f32), zero initial state.

    Under TF 2 the reference's cell runs with tanh / SIGMOID whatever it is given (keras' GRUCell constructor overwrites the
    attributes AUGRUCell sets first, activation.py:97-100; pinned by the committed reference golden, case 'augru'); this class does
    the same by default.  ``as_written=False`` uses the constructor's activations instead.  The masked branch of rnn_augru
    (activation.py:214-236) is not implemented: DIEN passes no mask."""

    def __init__(self, units, gru_type="AUGRU", return_sequence=False, activation="tanh", recurrent_activation="hard_sigmoid",
                 as_written=True, device=None, **kwargs):
        super().__init__(device)
        if activation not in ("tanh", "linear", None) or recurrent_activation not in ("hard_sigmoid", "sigmoid"):
            raise ValueError("AUGRU: activation tanh / linear, recurrent_activation hard_sigmoid / sigmoid")
        self.units = int(units)
        self.gru_type = gru_type
        self.return_sequences = bool(return_sequence)
        self.activation, self.recurrent_activation = ("tanh", "sigmoid") if as_written else (activation, recurrent_activation)

    def build(self, input_shape):
        E = int(input_shape[-1])
        self.kernel = self.add_weight("kernel", (E, 3 * self.units), init_glorot_uniform)
        self.recurrent_kernel = self.add_weight("recurrent_kernel", (self.units, 3 * self.units), init_glorot_uniform)

    def call(self, inputs, att_score=None, mask=None, training=None, initial_state=None, **kwargs):
        if mask is not None:
            raise NotImplementedError("AUGRU: the masked branch of rnn_augru is not implemented")
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "AUGRU")
        if x.stride(2) != 1 or x.stride(1) != x.shape[2]:
            x = x.contiguous()
        B, T, E = x.shape
        att = as_device_f32(att_score, self.device).reshape(B, T).contiguous()
        h0 = None if initial_state is None else as_device_f32(initial_state, self.device).reshape(B, self.units).contiguous()
        H = self.units
        seq = torch.empty((B, T, H), dtype=torch.float32, device=self.device) if self.return_sequences else None
        last = None if self.return_sequences else torch.empty((B, H), dtype=torch.float32, device=self.device)
        _lib.call("ctrhot_augru_f32", ptr(x), x.stride(0), ptr(att), att.stride(0), ptr(self.kernel), ptr(self.recurrent_kernel),
                  ptr(h0), B, T, E, H, int(self.recurrent_activation == "hard_sigmoid"), int(self.activation != "tanh"),
                  ptr(seq), ptr(last), stream_ptr())
        return seq if self.return_sequences else last
