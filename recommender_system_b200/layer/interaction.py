"""Drop-ins for the hot-path classes of the reference's ``layer/interaction.py``:
WideLayer :11, DNNLayer :30, CrossLayer :49, FMLayer :86, FFMLayer :117, InnerProductLayer :166,
OuterProductLayer :186, Attention :355, Dice :410 -- same constructor signatures, weight names and
shapes, ``call()`` input/output layouts; arithmetic runs in libctrhot (sm_100a CUDA).
Regularisers are accepted and ignored (forward only).
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .base import (Layer, as_device_f32, ids_arg, init_glorot_uniform, init_ones, init_random_normal, init_zeros,
                   ptr, require_cuda, stream_ptr)
from .core import DenseStack, _ids_to_device, dense_params


def _out(B, device, width=1):
    return torch.empty((B, width), dtype=torch.float32, device=device)


class WideLayer(Layer):
    """``w0 + X @ w`` -- DL/layer/interaction.py:11-26."""

    def __init__(self, device=None):
        super().__init__(device)

    def build(self, input_shape):
        self.w0 = self.add_weight("w0", (1,), init_zeros)
        self.w = self.add_weight("w", (input_shape[-1], 1), init_random_normal)

    def call(self, inputs, *args, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "WideLayer")
        out = _out(x.shape[0], self.device)
        _lib.call("ctrhot_wide_dense_f32", ptr(x), x.stride(0), x.shape[1], x.shape[0], ptr(self.w0), ptr(self.w),
                  ptr(out), stream_ptr())
        return out

    def call_onehot(self, dense, dense_col, ids, col_off, w=None):
        """Same result without materialising the one-hot block: w0 + sum dense_j w[dense_col_j] + sum_f w[col_off_f+id_f].
        ``w``: a weight vector of the same layout to use instead of ``self.w`` (WideDeep folds its duplicated dense block)."""
        ids, kind, ids_ld = ids_arg(ids)
        B = ids.shape[0]
        out = _out(B, self.device)
        _lib.call("ctrhot_wide_onehot_f32", ptr(dense), dense.stride(0), dense.shape[1], ptr(dense_col), ptr(ids), kind,
                  ids_ld, ids.shape[1], ptr(col_off), B, ptr(self.w0), ptr(self.w if w is None else w), ptr(out), stream_ptr())
        return out


class DNNLayer(Layer):
    """``DNNLayer(hidden_units, output_dim, activation='relu', dropout=0.2)`` -- interaction.py:30-46.
    Dropout is the identity at inference.  ``precision='tf32'`` runs the tcgen05 tower, 'fp32' the
    exact CUDA-core one."""

    def __init__(self, hidden_units, output_dim, activation="relu", dropout=0.2, precision="tf32", device=None):
        super().__init__(device)
        self.hidden_units = list(hidden_units)
        self.output_dim = output_dim
        if activation not in _lib.ACT_BY_NAME or activation in ("prelu", "dice"):
            raise ValueError("unsupported activation %r" % (activation,))
        self.activation = activation
        self.dropout = dropout
        self.stack = DenseStack(precision)

    def build(self, input_shape):
        dims = [int(input_shape[-1])] + self.hidden_units
        self.hidden = [dense_params(self, "dense_%d" % i, dims[i], dims[i + 1]) for i in range(len(self.hidden_units))]
        self.output = dense_params(self, "output", dims[-1], self.output_dim)

    def layer_dicts(self):
        act = _lib.ACT_BY_NAME[self.activation]
        ls = [{"kernel": k, "bias": b, "act": act} for k, b in self.hidden]
        ls.append({"kernel": self.output[0], "bias": self.output[1], "act": _lib.ACT_NONE})
        return ls

    def call(self, inputs, *args, x_ld=None, **kwargs):
        x = as_device_f32(inputs, self.device)
        self.stack.set_layers(self.layer_dicts(), self._version)
        return self.stack.run(x, x_ld)


class CrossLayer(Layer):
    """``x_{l+1} = x0 * (x_l . w_l) + b_l + x_l`` -- interaction.py:49-83; all layers in one launch."""

    def __init__(self, layer_num, reg_w=1e-4, reg_b=1e-4, device=None):
        super().__init__(device)
        self.layer_num = layer_num
        self.reg_w, self.reg_b = reg_w, reg_b
        self._stacked = None

    def build(self, input_shape):
        d = int(input_shape[1])
        self.cross_weight = [self.add_weight("w%d" % i, (d, 1), init_random_normal) for i in range(self.layer_num)]
        self.cross_bias = [self.add_weight("b%d" % i, (d, 1), init_random_normal) for i in range(self.layer_num)]

    def stacked(self):
        if self._stacked is None or self._stacked[0] != self._version:
            d = self.cross_weight[0].shape[0] if self.layer_num else 1
            w = torch.stack([t[:, 0] for t in self.cross_weight]) if self.layer_num else torch.zeros((1, d), device=self.device)
            b = torch.stack([t[:, 0] for t in self.cross_bias]) if self.layer_num else torch.zeros((1, d), device=self.device)
            self._stacked = (self._version, w.contiguous(), b.contiguous())
        return self._stacked[1], self._stacked[2]

    def call(self, inputs, *args, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "CrossLayer")
        w, b = self.stacked()
        d = x.shape[1]
        # rows padded to whole 16-byte pieces (a [B, d] view of them is returned): every store of the kernel is a full float4
        out = torch.empty((x.shape[0], (d + 3) // 4 * 4), dtype=torch.float32, device=x.device)[:, :d]
        _lib.call("ctrhot_cross_f32", ptr(x), x.stride(0), x.shape[1], x.shape[0], self.layer_num, ptr(w), ptr(b),
                  ptr(out), out.stride(0), 0, 0, stream_ptr())
        return out


class FMLayer(Layer):
    """``FMLayer(k, reg_w, reg_b)`` -- interaction.py:86-114; weights w0[1], w1[n,1], v[n,k]."""

    def __init__(self, k, reg_w=1e-4, reg_b=1e-4, device=None):
        super().__init__(device)
        self.k = k
        self.reg_w, self.reg_b = reg_w, reg_b
        self._planes = None

    def build(self, input_shape):
        n = int(input_shape[-1])
        self.w0 = self.add_weight("w0", (1,), init_zeros)
        self.w1 = self.add_weight("w1", (n, 1), init_random_normal)
        self.v = self.add_weight("v", (n, self.k), init_random_normal)

    def call(self, inputs, *args, apply_sigmoid=False, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "FMLayer")
        out = _out(x.shape[0], self.device)
        need = _lib.query("ctrhot_fm_dense_ws_bytes", x.shape[1], x.shape[0], self.k) if x.shape[0] > 0 else 0
        if need > 0:
            # few rows of many columns (the reference's FM on its dense one-hot matrix): columns split across CTAs
            ws = getattr(self, "_split_ws", None)
            if ws is None or ws.numel() * 4 < need or ws.device != x.device:
                ws = self._split_ws = torch.empty(need // 4, dtype=torch.float32, device=x.device)
            _lib.call("ctrhot_fm_dense_ws_f32", ptr(x), x.stride(0), x.shape[1], x.shape[0], ptr(self.w0), ptr(self.w1),
                      ptr(self.v), self.k, ptr(out), int(apply_sigmoid), ptr(ws), ws.numel() * 4, stream_ptr())
            return out
        _lib.call("ctrhot_fm_dense_f32", ptr(x), x.stride(0), x.shape[1], x.shape[0], ptr(self.w0), ptr(self.w1),
                  ptr(self.v), self.k, ptr(out), int(apply_sigmoid), stream_ptr())
        return out

    def call_onehot(self, dense, dense_col, ids, col_off, apply_sigmoid=False):
        """north_star's (sum v)^2 - sum v^2 over gathered rows: FM on [dense, one-hot] without the one-hot."""
        ids, kind, ids_ld = ids_arg(ids)
        B = ids.shape[0]
        out = _out(B, self.device)
        _lib.call("ctrhot_fm_onehot_f32", ptr(dense), dense.stride(0), dense.shape[1], ptr(dense_col), ptr(ids), kind,
                  ids_ld, ids.shape[1], ptr(col_off), B, ptr(self.w0), ptr(self.w1), ptr(self.v), self.k, ptr(out),
                  int(apply_sigmoid), stream_ptr())
        return out

    def planes(self, D):
        """[k+2][n_pad] planes in the fused front end's internal column order (cached)."""
        if self._planes is None or self._planes[0] != (self._version, D):
            n = self.w1.shape[0]
            nbytes = _lib.query("ctrhot_fm_planes_bytes", n, self.k)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=self.device)
            _lib.call("ctrhot_fm_pack_planes", ptr(self.w1), ptr(self.v), n, self.k, D, ptr(buf), stream_ptr())
            self._planes = ((self._version, D), buf)
        return self._planes[1]


class FFMLayer(Layer):
    """``FFMLayer(feature_columns, k, w_reg, v_reg)`` -- interaction.py:117-163 (the reference's own
    formula).  Input: the ``[B, 13+26]`` matrix with ids stored as floats."""

    def __init__(self, feature_columns, k, w_reg=1e-4, v_reg=1e-4, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.k = k
        self.w_reg, self.v_reg = w_reg, v_reg
        self.feature_num = sum(f["feat_onehot_dim"] for f in self.sparse_feature_columns) + len(self.dense_feature_columns)
        self.field_num = len(self.dense_feature_columns) + len(self.sparse_feature_columns)
        D = len(self.dense_feature_columns)
        off = D + np.concatenate([[0], np.cumsum([f["feat_onehot_dim"] for f in self.sparse_feature_columns])])
        self.col_off = torch.from_numpy(off.astype(np.int64)).to(self.device)

    def build(self, input_shape):
        self.w0 = self.add_weight("w0", (1,), init_zeros)
        self.w = self.add_weight("w", (self.feature_num, 1), init_random_normal)
        self.v = self.add_weight("v", (self.feature_num, self.field_num, self.k), init_random_normal)

    def call(self, inputs, *args, apply_sigmoid=False, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "FFMLayer")
        D = len(self.dense_feature_columns)
        F = len(self.sparse_feature_columns)
        B = x.shape[0]
        out = _out(B, self.device)
        ids = x[:, D:]
        _lib.call("ctrhot_ffm_onehot_f32", ptr(x), x.stride(0), D, ptr(ids), _lib.IDS_F32, x.stride(0), F,
                  ptr(self.col_off), B, ptr(self.w0), ptr(self.w), ptr(self.v), self.field_num, self.k, ptr(out),
                  int(apply_sigmoid), stream_ptr())
        return out


class InnerProductLayer(Layer):
    """``[B,F,k] -> [B, F(F-1)/2]`` pair dots, row-major i<j -- interaction.py:166-183."""

    def __init__(self, device=None):
        super().__init__(device)

    def call(self, inputs, *args, out=None, **kwargs):
        e = as_device_f32(inputs, self.device)
        require_cuda(e, "InnerProductLayer")
        B, F, k = e.shape
        P = F * (F - 1) // 2
        if out is None:
            out = _out(B, self.device, P)
        _lib.call("ctrhot_pnn_inner_f32", ptr(e), e.stride(0), F, k, B, ptr(out), out.stride(0), stream_ptr())
        return out


class OuterProductLayer(Layer):
    """``out[b,p] = q_p^T W_p p_p`` with ``W:[k,P,k]`` -- interaction.py:186-215."""

    def __init__(self, device=None):
        super().__init__(device)
        self.tensor_pipe = True          # k = 8: tcgen05 path (False: the CUDA-core kernels, any k)

    def build(self, input_shape):
        self.field_num = int(input_shape[1])
        self.k = int(input_shape[2])
        self.pair_num = self.field_num * (self.field_num - 1) // 2
        self.w = self.add_weight("W", (self.k, self.pair_num, self.k), init_random_normal)

    def call(self, inputs, *args, out=None, **kwargs):
        e = as_device_f32(inputs, self.device)
        require_cuda(e, "OuterProductLayer")
        B, F, k = e.shape
        if out is None:
            out = _out(B, self.device, self.pair_num)
        if (self.tensor_pipe and self.front_fits() and B >= 128 and e.stride(2) == 1 and e.stride(1) == k
                and e.stride(0) % 4 == 0 and e.data_ptr() % 16 == 0 and out.stride(1) == 1):
            # k = 8: the contraction on tcgen05 as a three-term TF32 split (fp32-accurate), csrc/pnn_fwd.cu
            _lib.call("ctrhot_pnn_products_f32", ptr(e), e.stride(0), F, k, B, ptr(self._pack()), 0, ptr(out), out.stride(0),
                      stream_ptr())
            return out
        _lib.call("ctrhot_pnn_outer_f32", ptr(e), e.stride(0), F, k, B, ptr(self.w), ptr(out), out.stride(0),
                  stream_ptr())
        return out

    def _pack(self):
        pk = getattr(self, "_front_pack", None)
        if pk is None or pk[0] != self._version:
            n = _lib.query("ctrhot_pnn_pack_bytes", self.field_num, self.k)
            buf = torch.empty(n, dtype=torch.uint8, device=self.device)
            _lib.call("ctrhot_pnn_pack", ptr(self.w), self.field_num, self.k, ptr(buf), stream_ptr())
            pk = self._front_pack = (self._version, buf)
        return pk[1]

    # ---- the one-launch front of PNN.call (gather -> flat | inner | outer), csrc/pnn_fwd.cu
    def front_fits(self):
        return self.built and self.k == 8 and 2 <= self.field_num <= 28

    def front(self, embed_layer, ids, do_inner, out, out_ld, signal_flags=None, epoch=0):
        """``out[b] = [rows | <e_i, e_j> | e_j^T W_p e_i | 0]`` straight from the tables (DL/model/pnn.py:28-49).
        ``signal_flags`` / ``epoch``: publish ``out`` per 128-row tile for a tower launched behind (chained launches)."""
        pk = (None, self._pack())
        ids, kind, ids_ld = ids_arg(ids)
        E = embed_layer
        _lib.call("ctrhot_pnn_front_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ids.shape[0],
                  ptr(pk[1]), 1 if do_inner else 0, ptr(out), out_ld, ptr(E.err_flag()), ptr(signal_flags), int(epoch),
                  _lib.LAUNCH_OVERLAP if signal_flags is not None else 0, stream_ptr())
        if E.check:
            E.check_ids()
        return out


class Dice(Layer):
    """``Dice(axis=-1, epsilon=1e-9)`` -- interaction.py:410-425; inference BN (moving stats, no gamma/beta)."""

    def __init__(self, axis=-1, epsilon=1e-9, device=None):
        super().__init__(device)
        self.axis = axis
        self.epsilon = epsilon

    def build(self, input_shape):
        h = int(input_shape[-1])
        self.alphas = self.add_weight("dice_alpha", (h,), init_zeros)
        self.moving_mean = self.add_weight("bn/moving_mean", (h,), init_zeros)
        self.moving_variance = self.add_weight("bn/moving_variance", (h,), init_ones)

    def params(self):
        return torch.stack([self.alphas, self.moving_mean, self.moving_variance]).contiguous()

    def call(self, inputs, training=None, *args, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "Dice")
        h = x.shape[-1]
        out = torch.empty_like(x)
        _lib.call("ctrhot_dice_f32", ptr(x), x.numel() // h, h, ptr(self.alphas), ptr(self.moving_mean),
                  ptr(self.moving_variance), float(self.epsilon), ptr(out), stream_ptr())
        return out


class Attention(Layer):
    """DIN local-activation unit, ``Attention(hidden_units, activation='prelu')`` -- interaction.py:355-406.

    ``call([query [B,K], key [B,T,K], value (== key), mask [B,T]]) -> [B,K]``.  'prelu': Dense(h,
    PReLU) layers whose alpha is [T,h] (Keras builds PReLU over input_shape[1:]); 'dice': Dice
    layers only, no Dense projection, then Dense(1)."""

    debug_checks = False        # True: also compare value with key element-wise when they are distinct tensors (device sync)

    def __init__(self, hidden_units, activation="prelu", device=None, precision="tf32"):
        super().__init__(device)
        self.hidden_units = list(hidden_units)
        if activation not in ("prelu", "dice"):
            raise ValueError("activation must be 'prelu' or 'dice'")
        self.activation = activation
        self.precision = precision          # 'tf32': tcgen05 unit (csrc/din_tc.cu) when the shape fits; 'fp32': CUDA cores
        self._descs = None
        self._tc = None
        self._fwd = None

    def build(self, input_shape):
        q_shape, k_shape = input_shape[0], input_shape[1]
        self.build_for(int(k_shape[1]), int(q_shape[-1]))

    def build_for(self, T, K):
        self.T, self.K = T, K
        w = 4 * K
        self.layers = []
        if self.activation == "prelu":
            for i, h in enumerate(self.hidden_units):
                k, b = dense_params(self, "dense_%d" % i, w, h)
                a = self.add_weight("dense_%d/prelu_alpha" % i, (T, h), init_zeros)
                self.layers.append((k, b, a))
                w = h
        else:
            self.dice_layers = [self.track(Dice(device=self.device)) for _ in self.hidden_units]
            for d in self.dice_layers:
                d.build((None, T, w))
                d.built = True
        self.out_kernel, self.out_bias = dense_params(self, "out", w, 1)
        self.built = True

    def descs(self):
        if self._descs is not None and self._descs[0] == self._version:
            return self._descs[1], self._descs[2]
        keep = []
        if self.activation == "prelu":
            arr = (_lib.DenseDesc * len(self.layers))()
            for i, (k, b, a) in enumerate(self.layers):
                d = arr[i]
                d.weight, d.bias, d.act_param = ptr(k), ptr(b), ptr(a)
                d.in_dim, d.out_dim = int(k.shape[0]), int(k.shape[1])
                d.act, d.alpha_rows = _lib.ACT_PRELU, self.T
        else:
            arr = (_lib.DenseDesc * len(self.dice_layers))()
            for i, dl in enumerate(self.dice_layers):
                p = dl.params()
                keep.append(p)
                d = arr[i]
                d.weight, d.bias, d.act_param = 0, 0, ptr(p)
                d.in_dim = d.out_dim = 4 * self.K
                d.act, d.alpha_rows, d.dice_eps = _lib.ACT_DICE, 1, float(dl.epsilon)
        self._descs = (self._version, arr, keep)
        return arr, keep

    def call(self, inputs, *args, **kwargs):
        query, key, value, mask = inputs
        query = as_device_f32(query, self.device)
        key = as_device_f32(key, self.device)
        mask = as_device_f32(mask, self.device)
        require_cuda(key, "Attention")
        if value is not key:
            value = as_device_f32(value, self.device)
            # DIN passes seq_embed for both (DL/model/din.py:81).  Comparing two distinct tensors element-wise would synchronise
            # the device on every call: shapes are checked always, contents only under Attention.debug_checks
            if value.data_ptr() != key.data_ptr() and (tuple(value.shape) != tuple(key.shape) or
                                                        (Attention.debug_checks and not torch.equal(value, key))):
                raise ValueError("Attention: value must equal key (DIN passes seq_embed for both, DL/model/din.py:81)")
        B, T, K = key.shape
        arr, _keep = self.descs()
        out = _out(B, self.device, K)
        _lib.call("ctrhot_din_attention_f32", ptr(query), query.stride(0), ptr(key), key.stride(0), ptr(mask), B, T, K,
                  arr, len(arr), ptr(self.out_kernel), ptr(self.out_bias), ptr(out), out.stride(0), stream_ptr())
        return out

    def call_gather(self, table, hist_ids, cand_ids, check_flag=None, out=None):
        """Fused history/candidate gather + attention: keys never reach HBM.  -> (att [B,K], item [B,K]);
        ``out=(att, item)``: write into these (row-strided) views instead, e.g. columns of the tower's input matrix."""
        hist_ids, kind, hist_ld = ids_arg(hist_ids)
        cand_ids, kind2, cand_ld = ids_arg(cand_ids)
        if kind2 != kind:
            cand_ids = cand_ids.to(hist_ids.dtype)
        B, T = hist_ids.shape
        V, K = table.shape
        arr, _keep = self.descs()
        att, item = out if out is not None else (_out(B, self.device, K), _out(B, self.device, K))
        if self.precision == "tf32" and self.activation == "prelu" and len(self.layers) == 2:
            h1, h2 = self.hidden_units
            nbytes = _lib.load().ctrhot_din_pack_bytes(T, K, h1, h2)
            if nbytes > 0:
                if self._tc is None or self._tc[0] != self._version:
                    blob = torch.empty(nbytes // 4, dtype=torch.float32, device=self.device)
                    _lib.call("ctrhot_din_pack", arr, len(arr), ptr(self.out_kernel), ptr(self.out_bias), T, K, ptr(blob),
                              stream_ptr())
                    self._tc = (self._version, blob)
                ws = torch.empty(_lib.query("ctrhot_din_tc_workspace_bytes", B, h1) // 4 + 4, dtype=torch.float32,
                                 device=self.device)
                _lib.call("ctrhot_din_attention_gather_tf32", ptr(table), V, K, ptr(hist_ids), kind, hist_ld,
                          ptr(cand_ids), cand_ld, B, T, ptr(self._tc[1]), h1, h2, ptr(att), att.stride(0), ptr(item),
                          item.stride(0), ptr(ws), ws.numel() * 4, ptr(check_flag), stream_ptr())
                _lib.LAUNCHES[0] += 1                       # prep + main kernel
                return att, item
        _lib.call("ctrhot_din_attention_gather_f32", ptr(table), V, K, ptr(hist_ids), kind, hist_ld, ptr(cand_ids),
                  cand_ld, B, T, arr, len(arr), ptr(self.out_kernel), ptr(self.out_bias), ptr(att), att.stride(0),
                  ptr(item), item.stride(0), ptr(check_flag), stream_ptr())
        return att, item

    def front_fits(self, T, K):
        """Does ctrhot_din_fwd_tf32 (the one-launch DIN front end) take this unit at history length T, key width K?"""
        if not (self.precision == "tf32" and self.activation == "prelu" and len(self.layers) == 2):
            return False
        h1, h2 = self.hidden_units
        return _lib.load().ctrhot_din_fwd_pack_bytes(T, K, h1, h2) > 0

    def front(self, table, hist_ids, cand_ids, emb, extras=(), dense=None, dense_col=0, check_flag=None):
        """DIN.call's front end in ONE launch: ``emb[:, :K]`` = attention output, ``emb[:, K:2K]`` = candidate rows,
        ``extras`` = [(table [rows, dim], ids [B,1], first column)] other sparse lookups, ``dense`` [B, n] copied to
        ``emb[:, dense_col:]`` (DL/model/din.py:56-90; csrc/din_fwd.cu).  ``emb`` is a [B, ld] matrix with ld % 4 == 0."""
        hist_ids, kind, hist_ld = ids_arg(hist_ids)
        cand_ids, kind2, cand_ld = ids_arg(cand_ids)
        if kind2 != kind:
            cand_ids = cand_ids.to(hist_ids.dtype)
        B, T = hist_ids.shape
        V, K = table.shape
        h1, h2 = self.hidden_units
        arr, _keep = self.descs()
        if self._fwd is None or self._fwd[0] != (self._version, T, K):
            nbytes = _lib.load().ctrhot_din_fwd_pack_bytes(T, K, h1, h2)
            if nbytes <= 0:
                raise ValueError("Attention.front: this shape needs call_gather (see front_fits)")
            blob = torch.empty(nbytes // 4, dtype=torch.float32, device=self.device)
            _lib.call("ctrhot_din_fwd_pack", arr, len(arr), ptr(self.out_kernel), ptr(self.out_bias), T, K, ptr(blob),
                      stream_ptr())
            self._fwd = ((self._version, T, K), blob)
        ex = np.zeros((max(len(extras), 1), 6), dtype=np.int64)
        keep = []
        for i, (tab, ids, col) in enumerate(extras):
            ids, k3, ld3 = ids_arg(ids)
            if k3 != kind:
                ids = ids.to(hist_ids.dtype)
            keep.append(ids)
            ex[i] = (tab.data_ptr(), tab.shape[0], tab.shape[1], ids.data_ptr(), ids.stride(0), col)
        ws = torch.empty(_lib.query("ctrhot_din_fwd_workspace_bytes", B, h1) // 4 + 4, dtype=torch.float32, device=self.device)
        n_dense = 0 if dense is None else dense.shape[1]
        _lib.call("ctrhot_din_fwd_tf32", ptr(table), V, K, ptr(hist_ids), kind, hist_ld, ptr(cand_ids), cand_ld, B, T,
                  ptr(self._fwd[1]), h1, h2, ptr(emb), emb.stride(0), ex.ctypes.data, len(extras),
                  ptr(dense) if n_dense else None, dense.stride(0) if n_dense else 0, n_dense, dense_col,
                  ptr(ws), ws.numel() * 4, ptr(check_flag), stream_ptr())
        return emb


# ---------------------------------------------------------------- the remaining towers
class ResLayer(Layer):
    """``ResLayer(hidden_units)`` -- interaction.py:261-278: ``relu(x + Dense_in(relu-Dense chain(x)))``.  The chain runs on
    the tensor pipe (DenseStack), the residual add + ReLU is one more pass (ctrhot_residual_relu_f32)."""

    def __init__(self, hidden_units, precision="tf32", device=None):
        super().__init__(device)
        self.hidden_units = list(hidden_units)
        self.stack = DenseStack(precision)

    def build(self, input_shape):
        n = int(input_shape[-1])
        dims = [n] + self.hidden_units
        self.dense_layer = [dense_params(self, "dense_%d" % i, dims[i], dims[i + 1]) for i in range(len(self.hidden_units))]
        self.output_layer = dense_params(self, "output", dims[-1], n)                # Dense(input width), interaction.py:269

    def call(self, inputs, *args, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "ResLayer")
        ls = [{"kernel": k, "bias": b, "act": _lib.ACT_RELU} for k, b in self.dense_layer]
        ls.append({"kernel": self.output_layer[0], "bias": self.output_layer[1], "act": _lib.ACT_NONE})
        self.stack.set_layers(ls, self._version)
        y = self.stack.run(x)
        out = torch.empty((x.shape[0], x.shape[1]), dtype=torch.float32, device=x.device)
        _lib.call("ctrhot_residual_relu_f32", ptr(x), x.stride(0), ptr(y), y.stride(0), x.shape[0], x.shape[1], ptr(out),
                  out.stride(0), stream_ptr())
        return out


class mmoe_layer(Layer):
    """``mmoe_layer(hidden_units, num_experts, num_tasks, use_expert_bias=True, use_gate_bias=True)`` -- interaction.py:429-509.
    The reference multiplies by ``expert_matrix[:, :, i]`` expert by expert and by every gate separately; here the experts are
    ONE Dense layer (the [in, hidden, E] weight is an [in, hidden*E] matrix in memory) and the gates another, both on the
    tensor pipe, followed by one kernel that applies each task's softmax over the experts and sums (ctrhot_mmoe_combine_f32).
    Returns the list of ``num_tasks`` tensors ``[B, hidden_units]``."""

    def __init__(self, hidden_units, num_experts, num_tasks, use_expert_bias=True, use_gate_bias=True, precision="tf32",
                 device=None, **kwargs):
        super().__init__(device)
        self.hidden_units, self.num_experts, self.num_tasks = int(hidden_units), int(num_experts), int(num_tasks)
        self.use_expert_bias, self.use_gate_bias = use_expert_bias, use_gate_bias
        self.experts, self.gates = DenseStack(precision), DenseStack(precision)
        self._cat = None

    def build(self, input_shape):
        n, H, E = int(input_shape[-1]), self.hidden_units, self.num_experts
        self.expert_matrix = self.add_weight("expert_matrix", (n, H, E), init_random_normal)
        self.expert_bias = self.add_weight("expert_bias", (H, E), init_random_normal) if self.use_expert_bias else None
        self.gate_matrix = [self.add_weight("gate_matrix%d" % i, (n, E), init_random_normal) for i in range(self.num_tasks)]
        self.gate_bias = ([self.add_weight("gate_bias%d" % i, (E,), init_random_normal) for i in range(self.num_tasks)]
                          if self.use_gate_bias else None)

    def call(self, inputs, *args, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "mmoe_layer")
        B, H, E, T = x.shape[0], self.hidden_units, self.num_experts, self.num_tasks
        if self._cat is None or self._cat[0] != self._version:
            n = self.expert_matrix.shape[0]
            ek = self.expert_matrix.reshape(n, H * E).contiguous()
            eb = (self.expert_bias.reshape(-1) if self.expert_bias is not None
                  else torch.zeros(H * E, device=self.device)).contiguous()
            gk = torch.cat(self.gate_matrix, dim=1).contiguous()
            gb = (torch.cat(self.gate_bias) if self.gate_bias is not None else torch.zeros(T * E, device=self.device)).contiguous()
            self._cat = (self._version, ek, eb, gk, gb)
        _, ek, eb, gk, gb = self._cat
        self.experts.set_layers([{"kernel": ek, "bias": eb, "act": _lib.ACT_RELU}], self._version)
        self.gates.set_layers([{"kernel": gk, "bias": gb, "act": _lib.ACT_NONE}], self._version)
        ex = self.experts.run(x)
        g = self.gates.run(x)
        out = torch.empty((T, B, H), dtype=torch.float32, device=x.device)
        _lib.call("ctrhot_mmoe_combine_f32", ptr(ex), ex.stride(0), ptr(g), g.stride(0), B, H, E, T, ptr(out), out.stride(0),
                  out.stride(1), stream_ptr())
        return [out[t] for t in range(T)]


class tower_layer(DNNLayer):
    """``tower_layer(hidden_units, output_dim, activation='relu')`` -- interaction.py:512-523: DNNLayer without the dropout."""

    def __init__(self, hidden_units, output_dim, activation="relu", precision="tf32", device=None):
        super().__init__(hidden_units, output_dim, activation, 0.0, precision=precision, device=device)
