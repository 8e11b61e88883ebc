"""Drop-in for the reference's ``layer/core.py`` hot-path class: ``EmbedLayer``
(DL/layer/core.py:267-280), plus ``DenseStack``, the host-side driver of the Dense-tower kernels
that DNNLayer / DIN / Attention share.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib
from .base import (Layer, as_device_f32, ids_arg, init_glorot_uniform, init_uniform, init_zeros, ptr,
                   require_cuda, stream_ptr)


class EmbedLayer(Layer):
    """``EmbedLayer(sparse_feature_columns, k=8)`` -- DL/layer/core.py:268.

    The reference builds one keras ``Embedding(feat_onehot_dim, k)`` per field and ignores
    ``feat['embed_dim']``; so does this.  All tables live in one arena
    ``[sum(V_f), k]`` so a single kernel serves the 26 lookups; ``emb_layers[i]`` is a view of
    table i.  ``call`` returns the flattened ``[B, F*k]`` of core.py:279 (``flatten=False`` gives
    the ``[B, F, k]`` view the PNN product layers are specified on).
    Out-of-range ids raise IndexError (CPU-TF behaviour); ``check_ids=False`` gives GPU-TF's
    silent zero rows.
    """

    def __init__(self, sparse_feature_columns, k=8, flatten=True, check_ids=True, device=None, init="uniform"):
        super().__init__(device)
        self.k = int(k)
        self.flatten = flatten
        self.check = check_ids
        self.vocab = [int(f["feat_onehot_dim"]) for f in sparse_feature_columns]
        self.F = len(self.vocab)
        off = np.concatenate([[0], np.cumsum(self.vocab)]).astype(np.int64)
        self.row_off_host = off
        self.row_off = torch.from_numpy(off).to(self.device)
        total = int(off[-1])
        if init == "uniform" and total * self.k > (1 << 24) and self.device.type == "cuda":
            # large arenas: same U(-0.05, 0.05) law, generated on the device (no multi-GB host staging)
            g = torch.Generator(device=self.device)
            from .base import _GEN
            g.manual_seed(_GEN["seed"] * 1000003 + total)          # follows set_seed like every other initialiser
            _GEN["seed"] += 1
            self.arena = torch.empty((total, self.k), dtype=torch.float32, device=self.device)
            self.arena.uniform_(-0.05, 0.05, generator=g)
            self._weights["embeddings"] = self.arena
        elif init == "uniform":
            self.arena = self.add_weight("embeddings", (total, self.k), init_uniform)
        else:                                   # caller fills the arena (e.g. hash-initialised shards)
            self.arena = torch.empty((total, self.k), dtype=torch.float32, device=self.device)
            self._weights["embeddings"] = self.arena
        self.emb_layers = [self.arena[off[i]:off[i + 1]] for i in range(self.F)]
        self.built = True

    def set_tables(self, tables):
        for dst, src in zip(self.emb_layers, tables):
            dst.copy_(torch.as_tensor(np.asarray(src, dtype=np.float32)))
        self._bump()

    def gather(self, ids, dense=None, out=None, out_ld=None):
        """x = [dense, embed(ids)] in the reference column order (dense may be None)."""
        ids = ids if isinstance(ids, torch.Tensor) and ids.is_cuda else _ids_to_device(ids, self.device)
        require_cuda(ids, "EmbedLayer")
        ids, kind, ids_ld = ids_arg(ids)
        B = ids.shape[0]
        D = 0 if dense is None else dense.shape[1]
        width = D + self.F * self.k
        if out is None:
            out_ld = width
            out = torch.empty((B, width), dtype=torch.float32, device=self.device)
        _lib.call("ctrhot_gather_f32", ptr(self.arena), ptr(self.row_off), self.F, self.k, ptr(ids), kind, ids_ld, B,
                  ptr(dense), 0 if dense is None else dense.stride(0), D, ptr(out), out_ld, ptr(self.err_flag()),
                  stream_ptr())
        if self.check:
            self.check_ids()
        return out

    def call(self, inputs, *args, **kwargs):
        emb = self.gather(inputs)
        if not self.flatten:
            emb = emb.view(emb.shape[0], self.F, self.k)
        return emb


def _ids_to_device(ids, device):
    if isinstance(ids, torch.Tensor):
        t = ids
    else:
        a = np.asarray(ids)
        if a.dtype == object:
            a = a.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype in (torch.int64, torch.int16, torch.uint8, torch.int8):
        t = t.to(torch.int32)
    elif t.dtype in (torch.float64, torch.float16, torch.bfloat16):
        t = t.to(torch.float32)
    return t.to(device, non_blocking=True)


class _Plan:
    """Owner of one ctrhot_fused_plan handle (host memory inside the library)."""

    def __init__(self, handle):
        self.handle = handle

    def __del__(self):
        try:
            _lib.load().ctrhot_fused_plan_destroy(self.handle)
        except Exception:
            pass


class DenseStack:
    """Host driver for a chain of Dense layers: descriptor array, packed weights, workspace.

    ``layers``: list of dicts {kernel, bias, act, act_param, alpha_rows, dice_eps}.  ``precision`` is
    'tf32' (tcgen05/TMEM path) or 'fp32' (CUDA-core path).
    """

    def __init__(self, precision="tf32"):
        if precision not in ("tf32", "fp32"):
            raise ValueError("precision must be 'tf32' or 'fp32'")
        self.precision = precision
        self.layers = []
        self._packed = {}
        self._plans = {}
        self._ws = None
        self._flags, self._epoch = None, 0
        self.chain_launches = True      # wide towers: launch l + 1 may start under the tail of launch l (per-tile ready flags)
        self._key = None
        self.allow_fused = True

    def set_layers(self, layers, version):
        if self._key != version:
            self.layers = layers
            self._packed = {}
            self._plans = {}               # plans hold pointers into the packed weights of the old version
            self._key = version
            self._descs_tmpl = None

    def descs(self):
        """Fresh descriptor array (callers overwrite ``packed`` of layer 0); built once per weight version and copied after
        that -- filling the ctypes structs was a third of the host time of a forward call."""
        tmpl = getattr(self, "_descs_tmpl", None)
        if tmpl is not None and tmpl[0] is self.layers:
            return type(tmpl[1]).from_buffer_copy(tmpl[1])
        arr = self._build_descs()
        self._descs_tmpl = (self.layers, arr)
        return type(arr).from_buffer_copy(arr)

    def _build_descs(self):
        n = len(self.layers)
        arr = (_lib.DenseDesc * n)()
        for i, L in enumerate(self.layers):
            k = L["kernel"]
            d = arr[i]
            d.weight = ptr(k)
            d.bias = ptr(L.get("bias"))
            d.act_param = ptr(L.get("act_param"))
            d.in_dim, d.out_dim = int(k.shape[0]), int(k.shape[1])
            d.act = int(L.get("act", _lib.ACT_NONE))
            d.alpha_rows = int(L.get("alpha_rows", 1))
            d.dice_eps = float(L.get("dice_eps", 1e-9))
            d.packed = 0
            if self.precision == "tf32" and d.out_dim >= 8:
                if i not in self._packed:
                    nbytes = _lib.query("ctrhot_mlp_pack_bytes", d.in_dim, d.out_dim)
                    buf = torch.empty(nbytes // 4, dtype=torch.float32, device=k.device)
                    _lib.call("ctrhot_mlp_pack", ptr(k), d.in_dim, d.out_dim, ptr(buf), stream_ptr())
                    self._packed[i] = buf
                d.packed = ptr(self._packed[i])
        return arr

    def fused_ok(self, arr=None):
        """Shape fits the single persistent tcgen05 tower kernel (ctrhot_tower_fused_tf32 / _gather_tf32)."""
        arr = arr if arr is not None else self.descs()
        n = len(arr)
        if self.precision != "tf32" or not (2 <= n <= 4) or not self.allow_fused:
            return False
        if arr[n - 1].out_dim != 1 or arr[n - 1].act != _lib.ACT_NONE:
            return False
        return all(8 <= arr[i].out_dim <= 256 for i in range(n - 1))

    @staticmethod
    def _post(post):
        post = post or {}
        return (ptr(post.get("y_scale")), ptr(post.get("add_vec")), ptr(post.get("add_bias")),
                float(post.get("scale", 1.0)), int(bool(post.get("sigmoid", False))))

    def run_fused(self, x, x_ld=None, post=None, out=None):
        """Whole tower (+ optional post) in one kernel; x [M, >=K] fp32, 16-byte aligned rows."""
        M = x.shape[0]
        x_ld = x.stride(0) if x_ld is None else x_ld
        arr = self.descs()
        if out is None:
            out = torch.empty((M, 1), dtype=torch.float32, device=x.device)
        _lib.call("ctrhot_tower_fused_tf32", ptr(x), x_ld, M, arr, len(arr), *self._post(post), ptr(out), stream_ptr())
        return out

    def run_rows(self, x, x_ld, sigmoid, out, wait_flags=None, epoch=0):
        """Tower over a matrix in HBM through the persistent whole-model kernel (no ids: row m of x feeds row m).
        ``wait_flags`` / ``epoch``: x is being written by the launch in front (chained launches, include/ctrhot.h): this launch
        may start before that one has finished and waits per 128-row tile."""
        arr = self.descs()
        K, n1 = arr[0].in_dim, arr[0].out_dim
        r = K % 4
        key = ("rows", r)
        if key not in self._packed:
            W1 = self.layers[0]["kernel"]
            W1r = torch.cat([W1[K - r:], W1[:K - r]], dim=0).contiguous() if r else W1.contiguous()
            nbytes = _lib.query("ctrhot_fused_pack_bytes", K, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=W1.device)
            _lib.call("ctrhot_fused_pack", ptr(W1r), None, K, n1, r, ptr(buf), stream_ptr())
            self._packed[key] = buf
        arr[0].packed = ptr(self._packed[key])
        if wait_flags is not None:
            _lib.call("ctrhot_tower_rows_post_tf32", ptr(x), x_ld, K, x.shape[0], arr, len(arr), None, 1.0, 0.0, 0.0,
                      int(bool(sigmoid)), _lib.LAUNCH_OVERLAP, ptr(out), ptr(wait_flags), int(epoch), stream_ptr())
            return out
        _lib.call("ctrhot_tower_rows_tf32", ptr(x), x_ld, K, x.shape[0], arr, len(arr), int(bool(sigmoid)), 0, ptr(out),
                  stream_ptr())
        return out

    def _rows_pack(self, l, K, n1):
        """Layer ``l`` packed for the persistent kernel in rows mode (last K % 4 rows of W first)."""
        r = K % 4
        key = ("rows", l, r)
        if key not in self._packed:
            W = self.layers[l]["kernel"]
            Wr = torch.cat([W[K - r:], W[:K - r]], dim=0).contiguous() if r else W.contiguous()
            nbytes = _lib.query("ctrhot_fused_pack_bytes", K, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=W.device)
            _lib.call("ctrhot_fused_pack", ptr(Wr), None, K, n1, r, ptr(buf), stream_ptr())
            self._packed[key] = buf
        return self._packed[key]

    def wide_rows_ok(self, arr=None):
        """Tower too wide to chain through TMEM but every hidden layer fits one persistent launch (<= 416 units):
        ctrhot_dense_rows_tf32 per hidden layer, the last one + the final Dense(h -> 1) by ctrhot_tower_rows_tf32."""
        arr = arr if arr is not None else self.descs()
        n = len(arr)
        if self.precision != "tf32" or n < 2 or not self.allow_fused:
            return False
        if arr[n - 1].out_dim != 1 or arr[n - 1].act != _lib.ACT_NONE:
            return False
        for i in range(n - 1):
            d = arr[i]
            if not (8 <= d.out_dim <= 416) or d.in_dim < 4:
                return False
            if d.act == _lib.ACT_PRELU:
                if d.alpha_rows > 1:
                    return False
            elif d.act not in (_lib.ACT_NONE, _lib.ACT_RELU):
                return False
        return True

    def run_rows_wide(self, x, x_ld, sigmoid, out):
        arr = self.descs()
        n = len(arr)
        M = x.shape[0]
        mx = max((arr[i].out_dim + 31) // 32 * 32 for i in range(n - 1))
        if n > 2 and (self._ws is None or self._ws.numel() < 2 * M * mx or self._ws.device != x.device):
            self._ws = torch.empty(2 * M * mx, dtype=torch.float32, device=x.device)
        cur, cur_ld = x, x_ld
        for l in range(n - 1):
            K, n1 = arr[l].in_dim, arr[l].out_dim
            arr[l].packed = ptr(self._rows_pack(l, K, n1))
            if l < n - 2:
                h_ld = (n1 + 31) // 32 * 32
                h = self._ws[(l & 1) * M * mx:(l & 1) * M * mx + M * h_ld].view(M, h_ld)
                _lib.call("ctrhot_dense_rows_tf32", ptr(cur), cur_ld, K, M, ctypes.pointer(arr[l]), ptr(h), h_ld, 0,
                          None, None, 0, stream_ptr())
                cur, cur_ld = h, h_ld
            else:
                _lib.call("ctrhot_tower_rows_tf32", ptr(cur), cur_ld, K, M, ctypes.pointer(arr[l]), 2, int(bool(sigmoid)), 0,
                          ptr(out), stream_ptr())
        return out

    def run_dcn_wide(self, E, ids, kind, ids_ld, dense, B, first_kernel_ref, cc_w, n_side, combine, wy, bias, err_flag, out,
                     cluster=0):
        """DCN with a wide tower (hidden layers of up to 416 units): ctrhot_dcn_first_layer_tf32 (gather + cross scalars +
        first layer), ctrhot_dense_rows_tf32 per middle layer, ctrhot_tower_rows_post_tf32 (last hidden + final Dense +
        sigmoid(side + wy * y + bias)).  Neither x nor the cross output exist in HBM."""
        arr = self.descs()
        n = len(arr)
        D = dense.shape[1]
        n_in, n1 = int(first_kernel_ref.shape[0]), int(first_kernel_ref.shape[1])
        key = ("fused", None)
        if key not in self._packed:
            nbytes = _lib.query("ctrhot_fused_pack_bytes", n_in, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=first_kernel_ref.device)
            _lib.call("ctrhot_fused_pack", ptr(first_kernel_ref.contiguous()), None, n_in, n1, D, ptr(buf), stream_ptr())
            self._packed[key] = buf
        mx = max((arr[i].out_dim + 31) // 32 * 32 for i in range(n - 1))
        # Chained launches: with per-tile "ready" flags launch l + 1 starts on the SMs launch l frees while it finishes its last
        # tiles (include/ctrhot.h).  Needs every launch to fill the GPU (then "all CTAs of a launch have started" implies that
        # every earlier launch is complete, so only neighbours overlap) and buffers no neighbour writes: two workspaces used in
        # turn, because the first launch of a call overlaps the last launch of the previous call.
        n_tiles = (B + 127) // 128
        chain = self.chain_launches and n_tiles >= 148 and n >= 3
        per = (2 * B * mx + B + 31) // 32 * 32              # (the second workspace starts 16-byte aligned for any B)
        if self._ws is None or self._ws.numel() < 2 * per or self._ws.device != dense.device:
            if self._ws is not None:
                torch.cuda.current_stream().synchronize()   # the previous call may still read the buffers being replaced
            self._ws = torch.empty(2 * per, dtype=torch.float32, device=dense.device)
            self._flags, self._epoch = None, 0
        # the two workspaces sit at FIXED offsets of the allocation, whatever this call's batch size: a call's first launch may
        # start while the previous call's last one still reads its buffers (overlapping launches), so the regions of consecutive
        # calls must not move against each other when B changes (a ragged last batch followed by a full one)
        per = self._ws.numel() // 2 // 32 * 32
        if chain and (self._flags is None or self._flags.shape[2] < n_tiles):
            self._flags = torch.zeros((2, n, n_tiles), dtype=torch.int32, device=dense.device)
            self._epoch = 0
        self._epoch += 1
        par = self._epoch & 1
        ws = self._ws[par * per:(par + 1) * per]
        fl = (lambda l: ptr(self._flags[par, l])) if chain else (lambda l: None)
        ep = self._epoch if chain else 0
        cl = int(cluster) | (_lib.LAUNCH_OVERLAP if chain else 0)
        side = ws[2 * B * mx:2 * B * mx + B]
        cb = (ctypes.c_float * len(combine))(*[float(c) for c in combine])
        h_ld = (n1 + 31) // 32 * 32
        h = ws[:B * h_ld].view(B, h_ld)
        arr[0].packed = ptr(self._packed[key])
        _lib.call("ctrhot_dcn_first_layer_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, B, ctypes.pointer(arr[0]), ptr(cc_w), int(n_side), cb, len(combine), ptr(h), h_ld,
                  ptr(side), cl, ptr(err_flag), fl(0), ep, stream_ptr())
        cur, cur_ld = h, h_ld
        for l in range(1, n - 1):
            K, nl = arr[l].in_dim, arr[l].out_dim
            arr[l].packed = ptr(self._rows_pack(l, K, nl))
            if l < n - 2:
                h_ld = (nl + 31) // 32 * 32
                h = ws[(l & 1) * B * mx:(l & 1) * B * mx + B * h_ld].view(B, h_ld)
                _lib.call("ctrhot_dense_rows_tf32", ptr(cur), cur_ld, K, B, ctypes.pointer(arr[l]), ptr(h), h_ld, cl,
                          fl(l - 1), fl(l), ep, stream_ptr())
                cur, cur_ld = h, h_ld
            else:
                _lib.call("ctrhot_tower_rows_post_tf32", ptr(cur), cur_ld, K, B, ctypes.pointer(arr[l]), 2, ptr(side),
                          float(wy), 1.0, float(bias), 1, cl, ptr(out), fl(l - 1), ep, stream_ptr())
        return out

    def run_ctr_tower(self, E, ids, kind, ids_ld, dense, B, first_kernel_ref, add_vec=None, scale=1.0, sigmoid=False,
                      err_flag=None, out=None, cluster=0):
        """Tower fed from EmbedLayer ``E`` by the whole-model kernel; out = (sigmoid of) scale * (add_vec + y)."""
        arr = self.descs()
        n, n1 = int(first_kernel_ref.shape[0]), int(first_kernel_ref.shape[1])
        D = 0 if dense is None else dense.shape[1]
        key = ("fused", None)
        if key not in self._packed:
            nbytes = _lib.query("ctrhot_fused_pack_bytes", n, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=first_kernel_ref.device)
            _lib.call("ctrhot_fused_pack", ptr(first_kernel_ref.contiguous()), None, n, n1, D, ptr(buf), stream_ptr())
            self._packed[key] = buf
        arr[0].packed = ptr(self._packed[key])
        if out is None:
            out = torch.empty((B, 1), dtype=torch.float32, device=E.arena.device)
        _lib.call("ctrhot_ctr_tower_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  0 if dense is None else dense.stride(0), D, B, arr, len(arr), ptr(add_vec), float(scale),
                  int(bool(sigmoid)), int(cluster), ptr(out), ptr(err_flag), stream_ptr())
        return out

    def run_widedeep(self, E, ids, kind, ids_ld, B, first_kernel_ref, wide_w, wide_off, dense, D, h_dw, w0, scale, err_flag,
                     out, cluster=0):
        """Wide&Deep in one launch: deep tower from the arena + the wide logit formed in the same kernel."""
        arr = self.descs()
        n, n1 = int(first_kernel_ref.shape[0]), int(first_kernel_ref.shape[1])
        key = ("fused", None)
        if key not in self._packed:
            nbytes = _lib.query("ctrhot_fused_pack_bytes", n, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=first_kernel_ref.device)
            _lib.call("ctrhot_fused_pack", ptr(first_kernel_ref.contiguous()), None, n, n1, 0, ptr(buf), stream_ptr())
            self._packed[key] = buf
        arr[0].packed = ptr(self._packed[key])
        _lib.call("ctrhot_widedeep_fused_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, B, arr, len(arr),
                  ptr(wide_w), ptr(wide_off), ptr(dense), dense.stride(0), int(D), h_dw, float(w0), float(scale), 1,
                  int(cluster), ptr(out), ptr(err_flag), stream_ptr())
        return out

    def run_fused_gather(self, E, ids, kind, ids_ld, dense, B, post=None, err_flag=None, out=None):
        """Tower fed straight from the embedding arena of EmbedLayer ``E`` (x never exists in HBM)."""
        arr = self.descs()
        D = 0 if dense is None else dense.shape[1]
        if out is None:
            out = torch.empty((B, 1), dtype=torch.float32, device=E.arena.device)
        _lib.call("ctrhot_tower_gather_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld,
                  ptr(dense), 0 if dense is None else dense.stride(0), D, B, arr, len(arr), *self._post(post),
                  ptr(out), ptr(err_flag), stream_ptr())
        return out

    # -- whole-model kernel (ctrhot_ctr_fused_tf32): gather -> side products -> tower -> combine, one launch
    def ctr_fused_ok(self, F, k, arr=None):
        arr = arr if arr is not None else self.descs()
        n = len(arr)
        if self.precision != "tf32" or not (2 <= n <= 4) or not self.allow_fused or k % 4 != 0 or F > 63:
            return False
        if arr[n - 1].out_dim != 1 or arr[n - 1].act != _lib.ACT_NONE:
            return False
        cols = 32                                            # side accumulator
        for i in range(n - 1):
            d = arr[i]
            if not (8 <= d.out_dim <= 256):
                return False
            if d.act == _lib.ACT_PRELU:
                if d.alpha_rows > 1:
                    return False
            elif d.act not in (_lib.ACT_NONE, _lib.ACT_RELU):
                return False
            cols += (d.out_dim + 31) // 32 * 32
        return cols <= 512

    def run_ctr_fused(self, E, ids, kind, ids_ld, dense, B, first_kernel_ref, side=None, sq_w=None, mode=0, n_side=0,
                      combine=(), sigmoid=False, err_flag=None, out=None, cluster=0, side_key=None, records=None, n_dense=None):
        """``first_kernel_ref``: layer-0 kernel [n, n1] in the REFERENCE column order [dense | embeddings];
        ``side`` [n, 16] (reference order) / ``sq_w`` / ``combine`` as described in include/ctrhot.h."""
        arr = self.descs()
        n, n1 = int(first_kernel_ref.shape[0]), int(first_kernel_ref.shape[1])
        D = (0 if dense is None else dense.shape[1]) if records is None else int(n_dense)
        key = ("fused", side_key)
        if key not in self._packed:
            nbytes = _lib.query("ctrhot_fused_pack_bytes", n, n1)
            buf = torch.empty(nbytes // 4, dtype=torch.float32, device=first_kernel_ref.device)
            _lib.call("ctrhot_fused_pack", ptr(first_kernel_ref.contiguous()), ptr(side), n, n1, D, ptr(buf), stream_ptr())
            self._packed[key] = buf
        arr[0].packed = ptr(self._packed[key])
        if out is None:
            out = torch.empty((B, 1), dtype=torch.float32, device=E.arena.device)
        if getattr(E, "peers", None) is None:
            # unsharded: a prepared launch (ctrhot_fused_plan_*), created once per (weights, input layout, launch options)
            if records is not None:
                batch, dl, lay = records.records, 0, records.layout
                layout_key = (_lib.IDS_PACKED, int(batch.stride(0)), lay.dense_off, id(lay))
            else:
                batch, dl, lay = ids, 0 if dense is None else int(dense.stride(0)), None
                layout_key = (int(kind), int(ids_ld), dl, None)
            pkey = (key, layout_key, D, int(mode), int(n_side), tuple(float(c) for c in combine), bool(sigmoid), int(cluster),
                    ptr(E.arena), ptr(err_flag), ptr(sq_w))
            plan = self._plans.get(pkey)
            if plan is None:
                cb = (ctypes.c_float * max(len(combine), 1))(*[float(c) for c in combine])
                h = ctypes.c_void_p()
                _lib.check(_lib.load().ctrhot_fused_plan_create(
                    ptr(E.arena), ptr(E.row_off), E.F, E.k, layout_key[0], layout_key[1],
                    ptr(lay.desc(batch.device)) if lay is not None else None, lay.dense_off if lay is not None else 0, dl, D,
                    arr, len(arr), ptr(sq_w), int(mode), int(n_side), cb, len(combine), int(bool(sigmoid)), int(cluster),
                    ptr(err_flag), ctypes.byref(h)), "ctrhot_fused_plan_create")
                plan = self._plans[pkey] = _Plan(h)
            _lib.call("ctrhot_fused_plan_run", plan.handle, ptr(batch), ptr(dense) if records is None else None, B, ptr(out),
                      stream_ptr())
            return out
        cb = (ctypes.c_float * max(len(combine), 1))(*[float(c) for c in combine])
        if records is not None:
            # compact transport records (utils/transport.py): ids decoded by the gather warps, dense read in place
            lay, rec = records.layout, records.records
            _lib.call("ctrhot_ctr_fused_packed_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(rec), rec.stride(0),
                      ptr(lay.desc(rec.device)), lay.dense_off, D, B, arr, len(arr), ptr(sq_w), int(mode), int(n_side), cb,
                      len(combine), int(bool(sigmoid)), int(cluster), ptr(out), ptr(err_flag), stream_ptr())
            return out
        tail = (ptr(ids), kind, ids_ld, ptr(dense), 0 if dense is None else dense.stride(0), D, B, arr, len(arr), ptr(sq_w),
                int(mode), int(n_side), cb, len(combine), int(bool(sigmoid)), int(cluster), ptr(out), ptr(err_flag),
                stream_ptr())
        if getattr(E, "peers", None) is not None:
            # row-sharded tables (model/sharded.py): the gather warps read the owners' arenas over NVLink
            _lib.call("ctrhot_ctr_fused_sharded_tf32", E.peers, ptr(E.row_off_all), E.G, E.rank, E.F, E.k, *tail)
        else:
            _lib.call("ctrhot_ctr_fused_tf32", ptr(E.arena), ptr(E.row_off), E.F, E.k, *tail)
        return out

    def run(self, x, x_ld=None, out=None, final_sigmoid=False):
        """``final_sigmoid``: sigmoid on the (activation-free) last layer's output -- inside the single-kernel tower when
        the shape fits it, otherwise as one more elementwise pass."""
        require_cuda(x, "DenseStack")
        M = x.shape[0]
        x_ld = x.stride(0) if x_ld is None else x_ld
        arr = self.descs()
        n = len(arr)
        n_out = arr[n - 1].out_dim
        if out is None:
            out = torch.empty((M, n_out), dtype=torch.float32, device=x.device)
        aligned = x_ld % 4 == 0 and x.data_ptr() % 16 == 0 and out.stride(0) == 1
        K = arr[0].in_dim
        if aligned and M >= 128 and K >= 4 and self.ctr_fused_ok(1, 4, arr):
            # the whole-model kernel with x as its one 'table' (ctrhot_tower_rows_tf32): warp-specialised, TMEM-chained
            return self.run_rows(x, x_ld, final_sigmoid, out)
        if self.fused_ok(arr) and aligned:
            return self.run_fused(x, x_ld, {"sigmoid": True} if final_sigmoid else None, out)
        if aligned and K >= 4 and self.wide_rows_ok(arr):
            return self.run_rows_wide(x, x_ld, final_sigmoid, out)
        need = _lib.query("ctrhot_mlp_workspace_bytes", M, arr, n)
        if self._ws is None or self._ws.numel() * 4 < need or self._ws.device != x.device:
            self._ws = torch.empty((need + 3) // 4, dtype=torch.float32, device=x.device)
        fn = "ctrhot_mlp_tf32" if self.precision == "tf32" else "ctrhot_mlp_f32"
        _lib.call(fn, ptr(x), x_ld, M, arr, n, ptr(out), out.stride(0), ptr(self._ws), self._ws.numel() * 4, stream_ptr())
        fused_dot = (self.precision == "tf32" and n >= 2 and n_out == 1 and arr[n - 1].act == _lib.ACT_NONE
                     and out.stride(0) == 1 and arr[n - 2].out_dim >= 8)
        _lib.LAUNCHES[0] += (n - 1 if fused_dot else n) - 1      # one kernel per layer; a trailing Dense(h->1) is fused
        if final_sigmoid:
            torch.sigmoid_(out)
        return out


def dense_params(layer, name, in_dim, out_dim):
    """keras Dense defaults: glorot_uniform kernel [in,out], zero bias."""
    k = layer.add_weight(name + "/kernel", (in_dim, out_dim), init_glorot_uniform)
    b = layer.add_weight(name + "/bias", (out_dim,), init_zeros)
    return k, b


# ---------------------------------------------------------------- the DeepCTR-lineage layers of layer/core.py
from .base import init_ones, init_zeros as _zeros  # noqa: E402


class DNN(Layer):
    """``DNN(hidden_units, activation='relu', l2_reg=0, dropout_rate=0, use_bn=False, output_activation=None, seed=1024)`` --
    DL/layer/core.py:123-220, the DeepCTR multi-layer perceptron on the LAST axis of an nD input.  Per layer: Dense, optional
    BatchNormalization BEFORE the activation (inference form, eps 1e-3: folded into the layer's weights), activation
    ('relu' / 'sigmoid' / 'tanh' / 'linear' / 'dice'); the last layer takes ``output_activation`` when given.  Dropout is the
    identity.  Weight names follow the reference: kernel0.., bias0.. (core.py:153-161)."""

    def __init__(self, hidden_units, activation="relu", l2_reg=0, dropout_rate=0, use_bn=False, output_activation=None,
                 seed=1024, precision="tf32", device=None, **kwargs):
        super().__init__(device)
        self.hidden_units = list(hidden_units)
        self.activation, self.output_activation = activation, output_activation
        self.use_bn = use_bn
        self.stack = DenseStack(precision)
        self._layers = None

    def _act_of(self, i):
        a = self.output_activation if (self.output_activation and i == len(self.hidden_units) - 1) else self.activation
        a = "dice" if a in ("dice", "Dice") else a
        if a not in _lib.ACT_BY_NAME or a == "prelu":
            raise ValueError("unsupported activation %r" % (a,))
        return a

    def build(self, input_shape):
        dims = [int(input_shape[-1])] + self.hidden_units
        n = len(self.hidden_units)
        self.kernels = [self.add_weight("kernel%d" % i, (dims[i], dims[i + 1]), init_glorot_uniform) for i in range(n)]
        self.bias = [self.add_weight("bias%d" % i, (dims[i + 1],), _zeros) for i in range(n)]
        self.bn = ([tuple(self.add_weight("bn_%d/%s" % (i, nm), (dims[i + 1],), init)
                          for nm, init in (("moving_mean", _zeros), ("moving_variance", init_ones), ("gamma", init_ones),
                                           ("beta", _zeros))) for i in range(n)] if self.use_bn else None)
        self.dice = {i: tuple(self.add_weight("dice_%d/%s" % (i, nm), (dims[i + 1],), init)
                              for nm, init in (("alpha", _zeros), ("moving_mean", _zeros), ("moving_variance", init_ones)))
                     for i in range(n) if self._act_of(i) == "dice"}

    def layer_dicts(self):
        if self._layers is None or self._layers[0] != self._version:
            ls = []
            for i, (k, b) in enumerate(zip(self.kernels, self.bias)):
                if self.bn is not None:
                    mean, var, gamma, beta = self.bn[i]
                    s = gamma / torch.sqrt(var + 1e-3)
                    k, b = (k * s[None, :]).contiguous(), ((b - mean) * s + beta).contiguous()
                a = self._act_of(i)
                d = {"kernel": k, "bias": b, "act": _lib.ACT_BY_NAME[a]}
                if a == "dice":
                    d["act_param"] = torch.stack(self.dice[i]).contiguous()
                    d["dice_eps"] = 1e-9
                ls.append(d)
            self._layers = (self._version, ls)
        return self._layers[1]

    def call(self, inputs, training=None, **kwargs):
        x = as_device_f32(inputs, self.device)
        require_cuda(x, "DNN")
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        if not self.hidden_units:
            return x
        self.stack.set_layers(self.layer_dicts(), self._version)
        y = self.stack.run(x2)
        return y.reshape(*lead, y.shape[-1])


class LocalActivationUnit(Layer):
    """``LocalActivationUnit(hidden_units=(64, 32), activation='sigmoid', l2_reg=0, dropout_rate=0, use_bn=False, seed=1024)``
    -- DL/layer/core.py:28-120: ``[q, k, q-k, q*k]`` (query repeated over T) -> DNN -> Dense(1) -> ``[B, T, 1]`` scores (no
    mask, no softmax).  The concat is built by ctrhot_lau_concat_f32; DNN and the final ``kernel [h,1]`` / ``bias`` run as one
    Dense stack."""

    def __init__(self, hidden_units=(64, 32), activation="sigmoid", l2_reg=0, dropout_rate=0, use_bn=False, seed=1024,
                 precision="tf32", device=None, **kwargs):
        super().__init__(device)
        self.hidden_units = list(hidden_units)
        self.dnn = self.track(DNN(hidden_units, activation, l2_reg, dropout_rate, use_bn, seed=seed, precision=precision,
                                  device=self.device))
        self.stack = DenseStack(precision)

    def build(self, input_shape):
        if not isinstance(input_shape, (list, tuple)) or len(input_shape) != 2:
            raise ValueError("A `LocalActivationUnit` layer should be called on a list of 2 inputs")
        K = int(input_shape[0][-1])
        size = 4 * K if len(self.hidden_units) == 0 else self.hidden_units[-1]
        self.kernel = self.add_weight("kernel", (size, 1), init_glorot_uniform)
        self.bias = self.add_weight("bias", (1,), _zeros)
        if not self.dnn.built:
            self.dnn.build((None, 4 * K))
            self.dnn.built = True

    def call(self, inputs, training=None, **kwargs):
        query, keys = inputs
        q = as_device_f32(query, self.device)
        k = as_device_f32(keys, self.device)
        require_cuda(k, "LocalActivationUnit")
        if q.dim() != 3 or k.dim() != 3 or q.shape[1] != 1 or q.shape[2] != k.shape[2]:
            raise ValueError("LocalActivationUnit needs inputs of shape (None,1,embedding_size) and (None,T,embedding_size)")
        B, T, K = k.shape
        q2, k2 = q.reshape(B, K), k.reshape(B * T, K)
        a = torch.empty((B * T, 4 * K), dtype=torch.float32, device=k.device)
        _lib.call("ctrhot_lau_concat_f32", ptr(q2), q2.stride(0), ptr(k2), k2.stride(0), B, T, K, ptr(a), a.stride(0), stream_ptr())
        ls = list(self.dnn.layer_dicts()) if self.hidden_units else []
        ls.append({"kernel": self.kernel, "bias": self.bias, "act": _lib.ACT_NONE})
        self.stack.set_layers(ls, (self._version, self.dnn._version))
        return self.stack.run(a).reshape(B, T, 1)
