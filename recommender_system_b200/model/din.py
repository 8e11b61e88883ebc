"""Drop-in for the reference's ``model/din.py`` (DL/model/din.py:15-95).

``DIN(feature_columns, behavior_feature_list, att_hidden_units=(80,40), dnn_hidden_units=(256,128,64),
att_attention='prelu', dnn_activation='prelu', dnn_dropout=0.0)``; ``call(inputs)`` takes the
reference's dict ``name -> [B,1] | [B,T]`` with the candidate under the literal key 'movie_id'
(din.py:74).  History and candidate share the behaviour tables (din.py:76-79); the mask comes from
the first behaviour feature (din.py:80); BatchNormalization runs in inference mode (eps 1e-3) and is
folded into the first tower layer; Dropout is the identity.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from ..layer.base import (Model, as_device_f32, init_ones, init_uniform, init_zeros, ptr, require_cuda, stream_ptr)
from ..layer.core import DenseStack, _ids_to_device, dense_params
from ..layer.interaction import Attention


class DIN(Model):
    def __init__(self, feature_columns, behavior_feature_list, att_hidden_units=(80, 40),
                 dnn_hidden_units=(256, 128, 64), att_attention="prelu", dnn_activation="prelu",
                 dnn_dropout=0.0, precision="tf32", check_ids=True, candidate_key="movie_id", device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.behavior_feature_list = list(behavior_feature_list)
        self.other_sparse = [f for f in self.sparse_feature_columns if f["feat"] not in self.behavior_feature_list]
        self.behavior = [f for f in self.sparse_feature_columns if f["feat"] in self.behavior_feature_list]
        self.other_sparse_num = len(self.other_sparse)
        self.dense_num = len(self.dense_feature_columns)
        self.behavior_num = len(self.behavior)
        self.candidate_key = candidate_key
        # keras Embedding(feat_onehot_dim, embed_dim), default 'uniform' init (din.py:39-45)
        self.embed_sparse_layers = [self.add_weight("embed_sparse_%d/embeddings" % i,
                                                    (f["feat_onehot_dim"], f["embed_dim"]), init_uniform)
                                    for i, f in enumerate(self.other_sparse)]
        self.embed_seq_layers = [self.add_weight("embed_seq_%d/embeddings" % i,
                                                 (f["feat_onehot_dim"], f["embed_dim"]), init_uniform)
                                 for i, f in enumerate(self.behavior)]
        self.att_layer = self.track(Attention(att_hidden_units, att_attention, device=self.device, precision=precision))
        self.dnn_hidden_units = list(dnn_hidden_units)
        self.dnn_activation = dnn_activation
        self.dnn_dropout = dnn_dropout
        self.check = check_ids
        self.stack = DenseStack(precision)
        self._folded = None

    # ---------------------------------------------------------------- weights
    def build(self, input_shape):
        K = sum(f["embed_dim"] for f in self.behavior)
        dcat = 2 * K + sum(f["embed_dim"] for f in self.other_sparse) + self.dense_num
        self.K, self.dcat = K, dcat
        self.bn_gamma = self.add_weight("bn/gamma", (dcat,), init_ones)
        self.bn_beta = self.add_weight("bn/beta", (dcat,), init_zeros)
        self.bn_mean = self.add_weight("bn/moving_mean", (dcat,), init_zeros)
        self.bn_var = self.add_weight("bn/moving_variance", (dcat,), init_ones)
        dims = [dcat] + self.dnn_hidden_units
        self.dense_layer = []
        for i in range(len(self.dnn_hidden_units)):
            k, b = dense_params(self, "dense_%d" % i, dims[i], dims[i + 1])
            if self.dnn_activation == "prelu":
                a = self.add_weight("dense_%d/prelu_alpha" % i, (dims[i + 1],), init_zeros)
                self.dense_layer.append((k, b, a))
            else:   # Dense(unit, activation=Dice()) -- din.py:51
                a = self.add_weight("dense_%d/dice_alpha" % i, (dims[i + 1],), init_zeros)
                m = self.add_weight("dense_%d/dice_bn/moving_mean" % i, (dims[i + 1],), init_zeros)
                v = self.add_weight("dense_%d/dice_bn/moving_variance" % i, (dims[i + 1],), init_ones)
                self.dense_layer.append((k, b, (a, m, v)))
        self.out_kernel, self.out_bias = dense_params(self, "out", dims[-1], 1)

    def _tower_layers(self):
        """BN (inference, eps 1e-3) folded into the first Dense: x_bn = x*s + t  =>  W' = s[:,None]*W, b' = b + t@W."""
        if self._folded is None or self._folded[0] != self._version:
            s = self.bn_gamma / torch.sqrt(self.bn_var + 1e-3)
            t = self.bn_beta - self.bn_mean * s
            layers = []
            for i, (k, b, a) in enumerate(self.dense_layer):
                if i == 0:
                    b = b + (t.double() @ k.double()).float()
                    k = (s[:, None] * k).contiguous()
                if self.dnn_activation == "prelu":
                    layers.append({"kernel": k, "bias": b, "act": _lib.ACT_PRELU, "act_param": a, "alpha_rows": 1})
                else:
                    layers.append({"kernel": k, "bias": b, "act": _lib.ACT_DICE, "act_param": torch.stack(a).contiguous(),
                                   "dice_eps": 1e-9})
            # final Dense(1, sigmoid) (din.py:54): the sigmoid rides in the tower kernel's post stage
            layers.append({"kernel": self.out_kernel, "bias": self.out_bias, "act": _lib.ACT_NONE})
            self._folded = (self._version, layers)
        return self._folded[1]

    # ---------------------------------------------------------------- forward
    def _col(self, inputs, name):
        v = inputs[name]
        if isinstance(v, torch.Tensor):
            return v
        a = np.asarray(v)
        return a

    def _gather1(self, key, tab, ids, out_view, out_ld, flag):
        """One table's rows for one id per row of ``ids`` [n, 1] into ``out_view`` (row stride ``out_ld``): ctrhot_gather_f32."""
        from ..layer.base import ids_arg
        if not hasattr(self, "_row_offs"):
            self._row_offs = {}
        cached = self._row_offs.get(key)               # built once: torch.tensor(..., device=) is a blocking copy
        if cached is None or cached[0] != tab.shape[0]:
            cached = self._row_offs[key] = (tab.shape[0], torch.tensor([0, tab.shape[0]], dtype=torch.int64, device=self.device))
        ids2, kind, ld = ids_arg(ids)
        _lib.call("ctrhot_gather_f32", ptr(tab), ptr(cached[1]), 1, tab.shape[1], ptr(ids2), kind, ld, ids2.shape[0], 0, 0, 0,
                  ptr(out_view), out_ld, ptr(flag), stream_ptr())

    def call(self, inputs, training=None, mask=None):
        dev = self.device
        if not self.att_layer.built:
            T = int(np.asarray(inputs[self.behavior[0]["feat"]]).shape[1]) if not isinstance(
                inputs[self.behavior[0]["feat"]], torch.Tensor) else int(inputs[self.behavior[0]["feat"]].shape[1])
            self.att_layer.build_for(T, self.K)
        hist = [_ids_to_device(self._col(inputs, f["feat"]), dev) for f in self.behavior]          # each [B,T]
        cand = _ids_to_device(self._col(inputs, self.candidate_key), dev)                          # [B,nb] (din.py:74)
        if cand.dim() == 1:
            cand = cand.unsqueeze(1)
        require_cuda(cand, "DIN")
        B = cand.shape[0]
        flag = self.err_flag()
        ld = (self.dcat + 3) // 4 * 4                           # 16-byte aligned rows for the tower kernel
        emb = torch.empty((B, ld), dtype=torch.float32, device=dev)[:, :self.dcat]
        K = self.K
        T = hist[0].shape[1]
        others = [_ids_to_device(self._col(inputs, f["feat"]), dev).reshape(B, -1)[:, :1] for f in self.other_sparse]
        dense = None
        if self.dense_num:
            cols = [as_device_f32(self._col(inputs, f["feat"]), dev).reshape(B, -1) for f in self.dense_feature_columns]
            dense = cols[0] if len(cols) == 1 else torch.cat(cols, dim=-1)
        if self.behavior_num == 1 and len(others) <= 8 and self.att_layer.front_fits(T, K):
            # ONE launch: candidate / history / other lookups, the attention unit and the dense columns write the tower's
            # input matrix [att | item | other | dense] (din.py:84-90's concat) -- csrc/din_fwd.cu
            extras, col = [], 2 * K
            for tab, ids in zip(self.embed_sparse_layers, others):
                extras.append((tab, ids, col))
                col += tab.shape[1]
            self.att_layer.front(self.embed_seq_layers[0], hist[0], cand[:, 0:1], emb, extras, dense, col, flag)
        else:
            if self.behavior_num == 1 and K % 4 == 0:
                # the attention kernel writes [att | item] straight into the tower's input matrix
                self.att_layer.call_gather(self.embed_seq_layers[0], hist[0], cand[:, 0:1], flag, out=(emb[:, :K], emb[:, K:2 * K]))
            elif self.behavior_num == 1:
                att, item = self.att_layer.call_gather(self.embed_seq_layers[0], hist[0], cand[:, 0:1], flag)
                emb[:, :K] = att
                emb[:, K:2 * K] = item
            else:
                # several behaviour features: key = concat of their rows (din.py:76-79).  Each table's rows go through the
                # gather kernel into its columns of key [B*T, K] and of the item block of emb; bad ids raise the flag
                key = torch.empty((B, T, K), dtype=torch.float32, device=dev)
                off = 0
                for i, tab in enumerate(self.embed_seq_layers):
                    e = tab.shape[1]
                    self._gather1(i + 1000, tab, hist[i].reshape(B * T, 1), key.view(B * T, K)[:, off:off + e], K, flag)
                    self._gather1(i + 1000, tab, cand[:, i:i + 1], emb[:, K + off:K + off + e], emb.stride(0), flag)
                    off += e
                m = (hist[0] != 0).to(torch.float32)
                att = self.att_layer([emb[:, K:2 * K], key, key, m])
                emb[:, :K] = att
            col = 2 * K
            for i, (tab, ids) in enumerate(zip(self.embed_sparse_layers, others)):
                e = tab.shape[1]
                self._gather1(i, tab, ids, emb[:, col:col + e], emb.stride(0), flag)
                col += e
            if dense is not None:
                emb[:, col:col + self.dense_num] = dense
        self.stack.set_layers(self._tower_layers(), self._version)
        out = self.stack.run(emb, final_sigmoid=True)
        if self.check:
            self.check_ids()
        return out
