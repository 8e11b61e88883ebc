"""Criteo-shaped models of the reference zoo on the fused B200 path:
FM (DL/model/fm.py:14-23), FFM (ffm.py:14-23), DeepFM (deepFM.py:15-31), WideDeep
(wideDeep.py:14-34), DCN (dcn.py:15-34), PNN (pnn.py:14-53).  Constructor signatures, weight names
and ``call()`` layouts follow the reference; additive kwargs: ``embed_dim`` (the reference never
passes EmbedLayer's k), ``precision`` ('tf32' tcgen05 tower | 'fp32'), ``device``.

Inputs: the reference's single ``[B, 39]`` matrix (cols 0-12 dense floats, 13-38 label-encoded ids
stored as floats) as numpy or torch; or the fast path ``(dense_f32 [B,13], ids_i32 [B,26])``.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib
from ..layer.base import Model, as_device_f32, ids_arg, ptr, require_cuda, stream_ptr
from ..layer.core import EmbedLayer, _ids_to_device
from ..layer.interaction import (CrossLayer, DNNLayer, FFMLayer, FMLayer, InnerProductLayer, OuterProductLayer,
                                 WideLayer)

from ..utils.transport import PackedRows  # noqa: E402

N_DENSE = 13   # hard-coded in every reference call(): deepFM.py:24, dcn.py:25, wideDeep.py:23, pnn.py:29


def split_inputs(inputs, device, n_dense=N_DENSE, n_sparse=None):
    """-> (dense [B,D] fp32 view, ids [B,F] view (float or int32), rest or None), all on device."""
    if isinstance(inputs, PackedRows):                     # transport records: one unpack launch (utils/transport.py)
        dense, ids = inputs.layout.unpack(inputs.records)
        return dense, ids, None
    if isinstance(inputs, (tuple, list)) and len(inputs) == 2:
        dense = as_device_f32(inputs[0], device)
        ids = _ids_to_device(inputs[1], device)
        return dense, ids, None
    x = as_device_f32(inputs, device)
    end = x.shape[1] if n_sparse is None else n_dense + n_sparse
    rest = x[:, end:] if end < x.shape[1] else None
    return x[:, :n_dense], x[:, n_dense:end], rest


def _front_supported(k, kfm, F, D):
    return k % 4 == 0 and F * (k // 4) <= 26 * 32 and D <= 32 and (kfm is None or kfm in (4, 8, 10, 16))


def _pad32(n):
    return (n + 31) // 32 * 32


def _permute_rows(W, D):
    """Reference column order [dense D | emb] -> internal [emb | dense D] (rows of a [n, *] weight)."""
    return torch.cat([W[D:], W[:D]], dim=0).contiguous()


class FM(Model):
    """``sigmoid(FMLayer(k)(x))`` on the dense one-hot matrix -- DL/model/fm.py:14-23."""

    def __init__(self, k, w_reg=1e-4, v_reg=1e-4, device=None):
        super().__init__(device)
        self.fm = self.track(FMLayer(k, w_reg, v_reg, device=self.device))

    def build(self, input_shape):
        if not self.fm.built:
            self.fm.build(input_shape)
            self.fm.built = True

    def call(self, inputs, training=None, mask=None):
        return self.fm(inputs, apply_sigmoid=True)

    def call_sparse(self, dense, ids, onehot_offsets):
        """Same result from (dense [B,13], ids [B,26]) + the one-hot column offset of each field."""
        dense = as_device_f32(dense, self.device)
        ids = _ids_to_device(ids, self.device)
        dense_col = torch.arange(dense.shape[1], dtype=torch.int32, device=self.device)
        col_off = torch.as_tensor(np.asarray(onehot_offsets, dtype=np.int64)).to(self.device)
        return self.fm.call_onehot(dense, dense_col, ids, col_off, apply_sigmoid=True)


class FFM(Model):
    """``sigmoid(FFMLayer(feature_columns, k)(x))`` -- DL/model/ffm.py:14-23."""

    def __init__(self, feature_columns, k, w_reg=1e-4, v_reg=1e-4, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.ffm = self.track(FFMLayer(feature_columns, k, w_reg, v_reg, device=self.device))

    def build(self, input_shape):
        if not self.ffm.built:
            self.ffm.build(input_shape)
            self.ffm.built = True

    def call(self, inputs, **kwargs):
        return self.ffm(inputs, apply_sigmoid=True)


class _EmbedTowerModel(Model):
    """Shared plumbing: embedding arena + tower whose first kernel is permuted to the internal order."""

    def _tower_layers(self, dnn, D):
        key = (dnn._version, D)
        if getattr(self, "_tl", None) is None or self._tl[0] != key:
            ls = dnn.layer_dicts()
            ls[0] = dict(ls[0], kernel=_permute_rows(ls[0]["kernel"], D))
            self._tl = (key, ls)
        return self._tl[1]

    def _ensure_built(self, dnn, n):
        if not dnn.built:
            dnn.build((None, n))
            dnn.built = True


class DeepFM(_EmbedTowerModel):
    """``sigmoid(0.5 * (FM(x) + DNN(x)))``, x = [dense, Embed(ids)] -- DL/model/deepFM.py:15-31.
    The FM term is a dense FM over x with its own v:[d,k]."""

    def __init__(self, feature_columns, k, w_reg, v_reg, hidden_units, output_dim, activation,
                 embed_dim=8, precision="tf32", check_ids=True, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=False,
                                                 device=self.device))
        self.fm = self.track(FMLayer(k, w_reg, v_reg, device=self.device))
        self.dnn = self.track(DNNLayer(hidden_units, output_dim, activation, precision=precision, device=self.device))
        self.check = check_ids
        self.fused = True
        self.whole = True           # one persistent kernel for the whole forward (ctrhot_ctr_fused_tf32) when it fits
        self.cluster = 0            # CTAs sharing each weight chunk by TMA multicast (0 = auto)
        # True: consecutive forward calls on one stream may overlap at their edges (CTRHOT_LAUNCH_OVERLAP, include/ctrhot.h).
        # Only for independent batches: the inputs of a call must not be produced by the previous call.
        self.overlap_launches = False
        self.events = None          # bench hook: list receiving (front_start, front_end, tower_end) CUDA events
        self._side = None
        self.embed_layer._err_flag = self.err_flag()

    def _fm_side(self, D):
        """FM weights as the whole-model kernel wants them: side [n,16] = [v | w1 | 0], sq_w[j] = sum_c v[j,c]^2
        (internal column order), combine = {w0}  (FMLayer.call, DL/layer/interaction.py:106-114)."""
        key = (self.fm._version, D)
        if self._side is None or self._side[0] != key:
            n, k = self.fm.v.shape
            side = torch.zeros((n, 16), dtype=torch.float32, device=self.device)
            side[:, :k] = self.fm.v
            side[:, k] = self.fm.w1[:, 0]
            n_pad = _pad32(n)
            sq_w = self.fm.planes(D)[(k + 1) * n_pad:(k + 2) * n_pad]
            self._side = (key, side, sq_w, [float(self.fm.w0.item())])
        return self._side[1:]

    def build(self, input_shape):
        D = len(self.dense_feature_columns)
        n = D + self.embed_layer.F * self.embed_layer.k
        if not self.fm.built:
            self.fm.build((None, n))
            self.fm.built = True
        self._ensure_built(self.dnn, n)

    def call(self, inputs, training=None, mask=None):
        E = self.embed_layer
        D = len(self.dense_feature_columns)
        records, fits = None, False
        if self.device.type != "cuda":                      # before anything touches torch.cuda: fail as the library, loudly
            raise _lib.CtrhotError("DeepFM: inputs/weights must live on a CUDA device -- there is no CPU fallback")
        if self.fused and self.whole and self.dnn.output_dim == 1 and self.fm.k < 16:
            stack = self.dnn.stack
            stack.set_layers(self._tower_layers(self.dnn, D), (self.dnn._version, D))
            fits = stack.ctr_fused_ok(E.F, E.k)
        if isinstance(inputs, PackedRows) and fits and inputs.layout.D == D and inputs.layout.F == E.F:
            # transport records go straight into the whole-model kernel: its gather warps decode the ids
            records = inputs
            require_cuda(records.records, "DeepFM")
            dense, ids, B = records.records, None, records.records.shape[0]
        else:
            dense, ids, _ = split_inputs(inputs, self.device, D)
            require_cuda(dense, "DeepFM")
            B = dense.shape[0]
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        if fits:
            if True:        # (kept as a block: the body below predates the `fits` hoist)
                # K-fused: gather -> FM (split-TF32 side columns + x^2 term) -> tower -> sigmoid(0.5*(fm+dnn)), ONE launch
                side, sq_w, cb = self._fm_side(D)
                if records is None:
                    ids, kind, ids_ld = ids_arg(ids)
                else:
                    kind, ids_ld = _lib.IDS_PACKED, 0
                ev = None
                if self.events is not None:
                    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                    ev[0].record()
                stack.run_ctr_fused(E, ids, kind, ids_ld, dense, B, self.dnn.hidden[0][0], side, sq_w, mode=1,
                                    n_side=self.fm.k, combine=cb, sigmoid=True, err_flag=self.err_flag(), out=out,
                                    cluster=self.cluster | (_lib.LAUNCH_OVERLAP if self.overlap_launches else 0),
                                    side_key=("fm", self.fm._version), records=records, n_dense=D)
                if ev:
                    ev[1].record()
                    self.events.append(tuple(ev))
                if self.check:
                    self.check_ids()
                return out
        if self.fused and self.dnn.output_dim == 1 and _front_supported(E.k, self.fm.k, E.F, D):
            ids, kind, ids_ld = ids_arg(ids)
            n = D + E.F * E.k
            fm_out = torch.empty((B,), dtype=torch.float32, device=self.device)
            planes = self.fm.planes(D)
            stack = self.dnn.stack
            stack.set_layers(self._tower_layers(self.dnn, D), (self.dnn._version, D))
            one_kernel_tower = stack.fused_ok() and E.F <= 32
            x, x_ld = None, 0
            if not one_kernel_tower:
                x_ld = _pad32(n)
                x = torch.empty((B, x_ld), dtype=torch.float32, device=self.device)
            ev = None
            if self.events is not None:
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                ev[0].record()
            # K2: gather -> FM in registers (x is written only if the tower cannot gather for itself)
            _lib.call("ctrhot_deepfm_front_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld,
                      ptr(dense), dense.stride(0), D, B, ptr(self.fm.w0), ptr(planes), self.fm.k,
                      ptr(x), x_ld, ptr(fm_out), ptr(self.err_flag()), stream_ptr())
            if ev:
                ev[1].record()
            if one_kernel_tower:
                # K7f: gather -> 3 Dense layers -> Dense(1) -> sigmoid(0.5*(fm+dnn)), one persistent tcgen05 kernel
                stack.run_fused_gather(E, ids, kind, ids_ld, dense, B,
                                       {"add_vec": fm_out, "scale": 0.5, "sigmoid": True}, self.err_flag(), out)
                dnn_out = None
            else:
                dnn_out = stack.run(x, x_ld)
            if ev:
                ev[2].record()
                self.events.append(tuple(ev))
            if dnn_out is None:
                if self.check:
                    self.check_ids()
                return out
        else:
            x = E.gather(ids, dense)
            fm_out = self.fm(x)
            dnn_out = self.dnn(x)
        _lib.call("ctrhot_combine_sigmoid_f32", ptr(fm_out), ptr(dnn_out), 0, 0, 0.5, B, ptr(out), stream_ptr())
        if self.check:
            self.check_ids()
        return out


class WideDeep(_EmbedTowerModel):
    """``sigmoid(0.5 * (Wide([dense, onehot]) + DNN(Embed(ids))))`` -- DL/model/wideDeep.py:14-34.
    The deep side sees embeddings only.  Matrix input: ``[B, 39 + n_onehot]``; fast path
    ``(dense, ids)`` uses ``onehot_width`` (default: feat_onehot_dim) and never builds the one-hot."""

    def __init__(self, feature_columns, hidden_units, output_dim, activation, embed_dim=8, precision="tf32",
                 check_ids=True, onehot_width=None, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=False,
                                                 device=self.device))
        self.wide = self.track(WideLayer(device=self.device))
        self.deep = self.track(DNNLayer(hidden_units, output_dim, activation, precision=precision, device=self.device))
        self.check = check_ids
        widths = onehot_width if onehot_width is not None else [f["feat_onehot_dim"] for f in self.sparse_feature_columns]
        self.set_onehot_layout(widths)

    def set_onehot_layout(self, widths):
        D = len(self.dense_feature_columns)
        self.onehot_width = [int(w) for w in widths]
        # wide input = [dense(D), get_dummies block = dense(D) again + one-hot fields] (dataset.py:53-58)
        off = 2 * D + np.concatenate([[0], np.cumsum(self.onehot_width)])
        self.col_off = torch.from_numpy(off.astype(np.int64)).to(self.device)
        self.dense_col = torch.arange(2 * D, dtype=torch.int32, device=self.device)
        self.n_wide = int(off[-1])

    def build(self, input_shape):
        if not self.wide.built:
            self.wide.build((None, self.n_wide))
            self.wide.built = True
        self._ensure_built(self.deep, self.embed_layer.F * self.embed_layer.k)

    def call(self, inputs, training=None, mask=None):
        E = self.embed_layer
        D = len(self.dense_feature_columns)
        dense, ids, onehot = split_inputs(inputs, self.device, D, E.F)
        require_cuda(dense, "WideDeep")
        B = dense.shape[0]
        stack = self.deep.stack
        stack.set_layers(self.deep.layer_dicts(), self.deep._version)
        if (onehot is None and getattr(self, "one_launch", True) and self.deep.output_dim == 1 and D <= 16 and E.k >= 8
                and (E.k & (E.k - 1)) == 0 and stack.ctr_fused_ok(E.F, E.k)):
            # ONE launch: the wide logit is formed inside the persistent kernel (ctrhot_widedeep_fused_tf32)
            wkey = self.wide._version
            if getattr(self, "_wide_host", None) is None or self._wide_host[0] != wkey:
                w = self.wide.w.reshape(-1)
                dw = (w[:D] + w[D:2 * D]).cpu().tolist()       # the dense block appears twice in the wide input
                self._wide_host = (wkey, (ctypes.c_float * max(D, 1))(*dw), float(self.wide.w0.item()), w.contiguous())
            _, dw, w0, wflat = self._wide_host
            out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
            ids2, kind, ids_ld = ids_arg(ids)
            # overlap_launches: as for DeepFM -- consecutive calls on independent batches may overlap at their edges
            stack.run_widedeep(E, ids2, kind, ids_ld, B, self.deep.hidden[0][0], wflat, self.col_off, dense, D, dw, w0, 0.5,
                               E.err_flag(), out,
                               cluster=_lib.LAUNCH_OVERLAP if getattr(self, "overlap_launches", False) else 0)
            if self.check:
                E.check_ids()
            return out
        if onehot is not None:
            wide_in = torch.cat([dense, onehot], dim=1)
            if wide_in.shape[1] != self.wide.w.shape[0]:
                raise ValueError("wide input has %d columns, weights expect %d" % (wide_in.shape[1], self.wide.w.shape[0]))
            wide_out = self.wide(wide_in)
        else:
            # the wide input carries the dense block twice (dataset.py:53-58): x_j w_j + x_j w_{D+j} = x_j (w_j + w_{D+j}), so
            # the two weight blocks are summed once per weight version instead of concatenating [dense, dense] per call
            wkey = self.wide._version
            if getattr(self, "_wide_folded", None) is None or self._wide_folded[0] != wkey:
                w2 = self.wide.w.clone()
                w2[:D] += w2[D:2 * D]
                w2[D:2 * D] = 0
                self._wide_folded = (wkey, w2)
            wide_out = self.wide.call_onehot(dense, self.dense_col[:D], ids, self.col_off, w=self._wide_folded[1])
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        stack = self.deep.stack
        stack.set_layers(self.deep.layer_dicts(), self.deep._version)
        if self.deep.output_dim == 1 and stack.ctr_fused_ok(E.F, E.k):
            # deep tower fed from the arena by the persistent kernel, wide logit added in its epilogue: two launches in all
            ids2, kind, ids_ld = ids_arg(ids)
            stack.run_ctr_tower(E, ids2, kind, ids_ld, None, B, self.deep.hidden[0][0], add_vec=wide_out, scale=0.5,
                                sigmoid=True, err_flag=E.err_flag(), out=out)
            if self.check:
                E.check_ids()
            return out
        emb = E.gather(ids)
        deep_out = self.deep(emb)
        _lib.call("ctrhot_combine_sigmoid_f32", ptr(wide_out), ptr(deep_out), 0, 0, 0.5, B, ptr(out), stream_ptr())
        if self.check:
            E.check_ids()
        return out


class DCN(_EmbedTowerModel):
    """``sigmoid(Dense_1([Cross(x), DNN(x)]))`` -- DL/model/dcn.py:15-34; all cross layers fused with the gather."""

    def __init__(self, feature_columns, hidden_units, output_dim, activation, layer_num, reg_w=1e-4, reg_b=1e-4,
                 embed_dim=8, precision="tf32", check_ids=True, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=False,
                                                 device=self.device))
        self.dense_layer = self.track(DNNLayer(hidden_units, output_dim, activation, precision=precision,
                                               device=self.device))
        self.cross_layer = self.track(CrossLayer(layer_num, reg_w, reg_b, device=self.device))
        self.check = check_ids
        self.fused = True
        self.whole = True
        self.cluster = 0
        self._packed = None
        self._side = None
        self.embed_layer._err_flag = self.err_flag()

    def build(self, input_shape):
        from ..layer.core import dense_params
        D = len(self.dense_feature_columns)
        n = D + self.embed_layer.F * self.embed_layer.k
        if not self.cross_layer.built:
            self.cross_layer.build((None, n))
            self.cross_layer.built = True
        self._ensure_built(self.dense_layer, n)
        self.out_kernel, self.out_bias = dense_params(self, "output_layer", n + self.dense_layer.output_dim, 1)

    def _cross_side(self, n):
        """Cross layers as scalars for the whole-model kernel (include/ctrhot.h, mode 2): x_l = a_l x_0 + c_l, so
        side = [w_0 .. w_{L-1} | final_w], combine = {c_l . w_l, c_L . final_w, wy, bias}, c_l = sum_{i<l} b_i."""
        key = (self._version, self.cross_layer._version)
        if self._side is None or self._side[0] != key:
            L = self.cross_layer.layer_num
            w, b = self.cross_layer.stacked()
            fw = self.out_kernel[:n, 0]
            side = torch.zeros((n, 16), dtype=torch.float32, device=self.device)
            cb = []
            c = torch.zeros(n, dtype=torch.float64, device=self.device)
            for l in range(L):
                side[:, l] = w[l]
                cb.append(float((c * w[l].double()).sum().item()))
                c = c + b[l].double()
            side[:, L] = fw
            cb.append(float((c * fw.double()).sum().item()))
            cb += [float(self.out_kernel[n, 0].item()), float(self.out_bias[0].item())]
            self._side = (key, side, cb)
        return self._side[1:]

    def _cc_side(self, side, D, n):
        """[round32(n), 4] copy of the first four side columns in the kernels' internal column order [embeddings | dense | 0]."""
        key = (self._version, self.cross_layer._version)
        if getattr(self, "_cc", None) is None or self._cc[0] != key:
            cc = torch.zeros((_pad32(n), 4), dtype=torch.float32, device=self.device)
            cc[:n - D] = side[D:n, :4]
            cc[n - D:n] = side[:D, :4]
            self._cc = (key, cc)
        return self._cc[1]

    def _packed_cross(self, D, n):
        key = (self._version, self.cross_layer._version)
        if self._packed is None or self._packed[0] != key:
            n_pad = _pad32(n)
            w, b = self.cross_layer.stacked()
            L = max(self.cross_layer.layer_num, 1)
            cw = torch.empty((L, n_pad), dtype=torch.float32, device=self.device)
            cb = torch.empty((L, n_pad), dtype=torch.float32, device=self.device)
            fw = torch.empty((1, n_pad), dtype=torch.float32, device=self.device)
            _lib.call("ctrhot_permute_cols", ptr(w), L, n, D, ptr(cw), stream_ptr())
            _lib.call("ctrhot_permute_cols", ptr(b), L, n, D, ptr(cb), stream_ptr())
            fsrc = self.out_kernel[:n, 0].contiguous()
            _lib.call("ctrhot_permute_cols", ptr(fsrc), 1, n, D, ptr(fw), stream_ptr())
            self._packed = (key, cw, cb, fw)
        return self._packed[1:]

    def call(self, inputs, training=None, mask=None):
        E = self.embed_layer
        D = len(self.dense_feature_columns)
        dense, ids, _ = split_inputs(inputs, self.device, D)
        require_cuda(dense, "DCN")
        B = dense.shape[0]
        n = D + E.F * E.k
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        if self.fused and self.whole and self.dense_layer.output_dim == 1 and self.cross_layer.layer_num < 16:
            stack = self.dense_layer.stack
            stack.set_layers(self._tower_layers(self.dense_layer, D), (self.dense_layer._version, D))
            if stack.ctr_fused_ok(E.F, E.k):
                side, cb = self._cross_side(n)
                ids, kind, ids_ld = ids_arg(ids)
                stack.run_ctr_fused(E, ids, kind, ids_ld, dense, B, self.dense_layer.hidden[0][0], side, None, mode=2,
                                    n_side=self.cross_layer.layer_num, combine=cb, sigmoid=True,
                                    err_flag=self.err_flag(), out=out,
                                    cluster=self.cluster | (_lib.LAUNCH_OVERLAP if getattr(self, "overlap_launches", False) else 0),
                                    side_key=("dcn", self._version, self.cross_layer._version))
                if self.check:
                    self.check_ids()
                return out
        if (self.fused and self.whole and self.dense_layer.output_dim == 1 and self.cross_layer.layer_num <= 3
                and len(self.dense_layer.hidden) >= 2 and E.k % 4 == 0 and E.F <= 63):
            stack = self.dense_layer.stack
            stack.set_layers(self._tower_layers(self.dense_layer, D), (self.dense_layer._version, D))
            if stack.wide_rows_ok():
                # tower too wide to chain through TMEM (cfg 3): gather + cross scalars + first layer in one launch, then one
                # launch per remaining hidden layer; h1, h2 and the per-row cross term are the only intermediates in HBM
                side, cb = self._cross_side(n)
                cc_w = self._cc_side(side, D, n)
                ids, kind, ids_ld = ids_arg(ids)
                stack.run_dcn_wide(E, ids, kind, ids_ld, dense, B, self.dense_layer.hidden[0][0], cc_w,
                                   self.cross_layer.layer_num, cb, cb[-2], cb[-1], self.err_flag(), out, cluster=self.cluster)
                if self.check:
                    self.check_ids()
                return out
        if self.fused and self.dense_layer.output_dim == 1 and _front_supported(E.k, None, E.F, D):
            ids, kind, ids_ld = ids_arg(ids)
            cw, cb, fw = self._packed_cross(D, n)
            stack = self.dense_layer.stack
            stack.set_layers(self._tower_layers(self.dense_layer, D), (self.dense_layer._version, D))
            one_kernel_tower = stack.fused_ok() and E.F <= 32
            x, x_ld = None, 0
            if not one_kernel_tower:
                x_ld = _pad32(n)
                x = torch.empty((B, x_ld), dtype=torch.float32, device=self.device)
            cdot = torch.empty((B,), dtype=torch.float32, device=self.device)
            # K5: gather -> all L cross layers in registers -> dot with the cross half of the final Dense
            _lib.call("ctrhot_dcn_front_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld,
                      ptr(dense), dense.stride(0), D, B, self.cross_layer.layer_num, ptr(cw), ptr(cb), ptr(fw),
                      ptr(x), x_ld, 0, 0, ptr(cdot), ptr(self.err_flag()), stream_ptr())
            if one_kernel_tower:
                stack.run_fused_gather(E, ids, kind, ids_ld, dense, B,
                                       {"y_scale": self.out_kernel[n:], "add_vec": cdot, "add_bias": self.out_bias,
                                        "scale": 1.0, "sigmoid": True}, self.err_flag(), out)
            else:
                dnn_out = stack.run(x, x_ld)
                _lib.call("ctrhot_combine_sigmoid_f32", ptr(cdot), ptr(dnn_out), ptr(self.out_kernel[n:]),
                          ptr(self.out_bias), 1.0, B, ptr(out), stream_ptr())
        else:
            x = E.gather(ids, dense)
            cross = self.cross_layer(x)
            dnn_out = self.dense_layer(x)
            cat = torch.cat([cross, dnn_out], dim=1)
            logit = torch.empty((B, 1), dtype=torch.float32, device=self.device)
            _lib.call("ctrhot_wide_dense_f32", ptr(cat), cat.stride(0), cat.shape[1], B, ptr(self.out_bias),
                      ptr(self.out_kernel), ptr(logit), stream_ptr())
            _lib.call("ctrhot_combine_sigmoid_f32", ptr(logit), 0, 0, 0, 1.0, B, ptr(out), stream_ptr())
        if self.check:
            self.check_ids()
        return out


class NFM(_EmbedTowerModel):
    """``sigmoid(Dense_1(DNN(BN([dense, BiInteraction(Embed(ids))]))))`` -- DL/model/nfm.py:15-34.
    Bi-Interaction = 0.5 * ((sum_f e_f)^2 - sum_f e_f^2) over the [B, F, k] view of the embeddings (the view the reference's
    own comment names; its EmbedLayer flattens).  Two launches: gather + pooling in registers
    (ctrhot_nfm_bi_f32, the F rows never reach HBM), then the tower with the inference BatchNormalization folded into its
    first layer and the final Dense(1) + sigmoid in its epilogue."""

    def __init__(self, feature_columns, hidden_units, output_dim, activation="relu", dropout=0, embed_dim=8,
                 precision="tf32", check_ids=True, device=None):
        super().__init__(device)
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.emb_layers = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=False,
                                                device=self.device))
        self.dnn_layers = self.track(DNNLayer(hidden_units, output_dim, activation, dropout, precision=precision,
                                              device=self.device))
        self.check = check_ids
        self._folded = None
        self.emb_layers._err_flag = self.err_flag()

    def build(self, input_shape):
        from ..layer.base import init_ones, init_zeros
        from ..layer.core import dense_params
        n = len(self.dense_feature_columns) + self.emb_layers.k
        self.bn_gamma = self.add_weight("bn/gamma", (n,), init_ones)
        self.bn_beta = self.add_weight("bn/beta", (n,), init_zeros)
        self.bn_mean = self.add_weight("bn/moving_mean", (n,), init_zeros)
        self.bn_var = self.add_weight("bn/moving_variance", (n,), init_ones)
        self._ensure_built(self.dnn_layers, n)
        self.out_kernel, self.out_bias = dense_params(self, "output_layer", self.dnn_layers.output_dim, 1)

    def _tower(self):
        """BN (inference, eps 1e-3) folded into the first Dense; DNN's activation-free output layer and the model's final
        Dense(1) are two linear maps in a row and are multiplied out: one [h_last, 1] layer."""
        key = (self._version, self.dnn_layers._version)
        if self._folded is None or self._folded[0] != key:
            s = (self.bn_gamma / torch.sqrt(self.bn_var + 1e-3)).double()
            t = self.bn_beta.double() - self.bn_mean.double() * s
            ls = [dict(l) for l in self.dnn_layers.layer_dicts()]
            k0 = ls[0]["kernel"].double()
            ls[0]["bias"] = (ls[0]["bias"].double() + t @ k0).float().contiguous()
            ls[0]["kernel"] = (s[:, None] * k0).float().contiguous()
            wo, bo = ls[-1]["kernel"].double(), ls[-1]["bias"].double()
            fk, fb = self.out_kernel.double(), self.out_bias.double()
            ls[-1] = dict(ls[-1], kernel=(wo @ fk).float().contiguous(), bias=(bo @ fk + fb).float().contiguous())
            self._folded = (key, ls)
        return self._folded[1]

    def call(self, inputs, training=None, mask=None):
        E = self.emb_layers
        D = len(self.dense_feature_columns)
        dense, ids, _ = split_inputs(inputs, self.device, D)
        require_cuda(dense, "NFM")
        B = dense.shape[0]
        n = D + E.k
        x_ld = (n + 3) // 4 * 4
        x = torch.empty((B, x_ld), dtype=torch.float32, device=self.device)
        ids, kind, ids_ld = ids_arg(ids)
        _lib.call("ctrhot_nfm_bi_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, B, ptr(dense),
                  dense.stride(0), D, ptr(x), x_ld, ptr(self.err_flag()), stream_ptr())
        stack = self.dnn_layers.stack
        stack.set_layers(self._tower(), (self._version, self.dnn_layers._version, "nfm"))
        out = stack.run(x[:, :n], x_ld, final_sigmoid=True)
        if self.check:
            self.check_ids()
        return out


class AFMLayer(Model):
    """``AFMLayer(feature_columns, mode)`` -- DL/layer/interaction.py:322-351 with its InteractionLayer / AttentionLayer
    (:281-319): sigmoid(Dense_1(pool over all e_i * e_j)), mode in 'avg' | 'max' | 'att'.  One launch (ctrhot_afm_pool_f32):
    the [B, P, k] pair tensor is never built.  In 'att' mode the reference's softmax runs over a size-1 axis, so the attention
    weights are all 1 and its two Dense layers (kept here as weights, for checkpoints) do not reach the output."""

    def __init__(self, feature_columns, mode, n_sigmoid=1, check_ids=True, device=None):
        super().__init__(device)
        if mode not in ("avg", "max", "att"):
            raise ValueError("mode must be 'avg', 'max' or 'att'")
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.mode = mode
        ks = {int(f.get("embed_dim", 8)) for f in self.sparse_feature_columns}
        if len(ks) != 1:
            raise ValueError("AFMLayer: all sparse features must share one embed_dim")
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=ks.pop(), check_ids=False, device=self.device))
        self.n_sigmoid = n_sigmoid
        self.check = check_ids
        self.embed_layer._err_flag = self.err_flag()

    def build(self, input_shape):
        from ..layer.core import dense_params
        F, k = self.embed_layer.F, self.embed_layer.k
        P = F * (F - 1) // 2
        if self.mode == "att":
            self.att_w = dense_params(self, "attention_w", k, P)
            self.att_h = dense_params(self, "attention_h", P, 1)
        self.out_kernel, self.out_bias = dense_params(self, "output_layer", k, 1)

    def call(self, inputs, training=None, mask=None):
        E = self.embed_layer
        D = len(self.dense_feature_columns)
        _, ids, _ = split_inputs(inputs, self.device, D)
        require_cuda(ids, "AFMLayer")
        B = ids.shape[0]
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        ids, kind, ids_ld = ids_arg(ids)
        _lib.call("ctrhot_afm_pool_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, B,
                  {"avg": 0, "att": 1, "max": 2}[self.mode], ptr(self.out_kernel), ptr(self.out_bias), int(self.n_sigmoid),
                  ptr(out), ptr(self.err_flag()), stream_ptr())
        if self.check:
            self.check_ids()
        return out


class AFM(AFMLayer):
    """``AFM(feature_columns, mode)`` -- DL/model/afm.py: sigmoid(AFMLayer(x)), i.e. the sigmoid is applied twice."""

    def __init__(self, feature_columns, mode, check_ids=True, device=None):
        super().__init__(feature_columns, mode, n_sigmoid=2, check_ids=check_ids, device=device)


class PNN(Model):
    """``DNN([flat(E), inner(E) and/or outer(E)])``, no sigmoid -- DL/model/pnn.py:14-53.  The product
    layers get the [B,F,k] view of the gathered buffer; the FGCNN branch is out of scope."""

    def __init__(self, feature_columns, mode, hidden_units, output_dim, activation="relu", dropout=0.2,
                 use_fgcnn=False, embed_dim=8, precision="tf32", check_ids=True, device=None):
        super().__init__(device)
        if mode not in ("inner", "outer", "both"):
            raise ValueError("Please choice mode's value in 'inner', 'outer', 'both'.")
        if use_fgcnn:
            raise NotImplementedError("FGCNN branch (pnn.py:33-35) is outside the hot path (SURVEY 2.1 C8)")
        self.mode = mode
        self.fused_front = True          # False: gather / inner / outer as separate launches (the only path for k != 8)
        self.chain_launches = True       # front -> tower under per-tile flags with overlapping launches (large batches)
        self._chain = None
        self.dense_feature_columns, self.sparse_feature_columns = feature_columns
        self.dnn_layer = self.track(DNNLayer(hidden_units, output_dim, activation, dropout, precision=precision,
                                             device=self.device))
        self.inner_product_layer = self.track(InnerProductLayer(device=self.device))
        self.outer_product_layer = self.track(OuterProductLayer(device=self.device))
        self.embed_layer = self.track(EmbedLayer(self.sparse_feature_columns, k=embed_dim, check_ids=check_ids,
                                                 device=self.device))

    def build(self, input_shape):
        E = self.embed_layer
        P = E.F * (E.F - 1) // 2
        if not self.outer_product_layer.built:
            self.outer_product_layer.build((None, E.F, E.k))
            self.outer_product_layer.built = True
        width = E.F * E.k + P * {"inner": 1, "outer": 1, "both": 2}[self.mode]
        if not self.dnn_layer.built:
            self.dnn_layer.build((None, width))
            self.dnn_layer.built = True
        self.width = width

    def call(self, inputs, training=None, mask=None):
        E = self.embed_layer
        D = len(self.dense_feature_columns)
        _dense, ids, _ = split_inputs(inputs, self.device, D)
        B = ids.shape[0]
        fk = E.F * E.k
        P = E.F * (E.F - 1) // 2
        ld = _pad32(self.width)
        # every column below `width` is written by the gather / product kernels; only the K padding of the tower's first layer
        # has to be zero (garbage times a zero weight row could be NaN), not the whole [B, ld] buffer
        if self.fused_front and self.mode != "inner" and self.outer_product_layer.front_fits():
            # gather + both product layers in one launch, the rows never reach HBM as an intermediate (csrc/pnn_fwd.cu)
            st = self.dnn_layer.stack
            arr = st.descs()
            n_tiles = (B + 127) // 128
            chain = (self.chain_launches and n_tiles >= 148 and st.precision == "tf32" and ld % 4 == 0
                     and st.ctr_fused_ok(1, 4, arr) and arr[-1].out_dim == 1)
            if chain:
                # the tower starts on the SMs the front frees and waits per tile: two alternating product buffers (the next
                # call's front must not write what this call's tower still reads) and a new epoch per call
                c = self._chain
                if c is None or c[0].shape[0] < 2 * B or c[0].shape[1] != ld:      # capacity: two buffers of B rows each
                    c = self._chain = [torch.empty((2 * B, ld), dtype=torch.float32, device=self.device),
                                       torch.zeros(n_tiles, dtype=torch.int32, device=self.device), 0]
                c[2] += 1
                half = c[0].shape[0] // 2
                buf = c[0][(c[2] & 1) * half:(c[2] & 1) * half + B]
                self.outer_product_layer.front(E, ids, self.mode == "both", buf, ld, c[1], c[2])
                out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
                return st.run_rows(buf[:, :self.width], ld, False, out, c[1], c[2])
            buf = torch.empty((B, ld), dtype=torch.float32, device=self.device)
            self.outer_product_layer.front(E, ids, self.mode == "both", buf, ld)
            return self.dnn_layer(buf[:, :self.width], x_ld=ld)
        buf = torch.empty((B, ld), dtype=torch.float32, device=self.device)
        if ld > self.width:
            buf[:, self.width:].zero_()
        E.gather(ids, out=buf, out_ld=ld)
        z = buf.as_strided((B, E.F, E.k), (ld, E.k, 1))
        col = fk
        if self.mode in ("inner", "both"):
            self.inner_product_layer(z, out=buf[:, col:col + P])
            col += P
        if self.mode in ("outer", "both"):
            self.outer_product_layer(z, out=buf[:, col:col + P])
        return self.dnn_layer(buf[:, :self.width], x_ld=ld)
