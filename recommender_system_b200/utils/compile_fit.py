"""Drop-in for the reference's ``utils/compile_fit.py`` (DL/utils/compile_fit.py:9-15):

    compile_fit(model, X, y, batch_size=32, epochs=10, sgd=0.01) -> model

``tf.data`` batches of ``batch_size`` in order (no shuffling: keras does not shuffle a Dataset), ``SGD(sgd)``, loss
``binary_crossentropy``, metric accuracy, ``epochs`` passes.  Each batch is ONE call of ``ctrhot_deepfm_train_step_f32``
(csrc/train.cu): forward, clipped BCE + FMLayer's l2 regularisers, exact fp32 gradients, in-place SGD on the embedding
rows that were looked up, the FM weights and the tower.  SURVEY 8(f) rank 3; built for DeepFM (the reference's flagship
``__main__``, DL/model/deepFM.py:34-57) and for FM on its dense one-hot matrix (DL/model/fm.py:26-45, BASELINE config 1;
``ctrhot_fm_train_step_f32``) and for DCN (DL/model/dcn.py:37-49: cross layers + tower + the final Dense over their concat;
``ctrhot_dcn_train_step_f32``) and for PNN (the loop of DL/model/pnn.py:74-85, which compile_fit reproduces for it: BCE on
the raw output, no regularisers; ``ctrhot_pnn_train_step_f32``) and for WideDeep (DL/model/wideDeep.py:37-49; ``ctrhot_widedeep_train_step_f32``) and DeepCrossing
(DL/model/deepCrossing.py:35-47; ``ctrhot_deepcrossing_train_step_f32``) and NFM (DL/model/nfm.py:37-50, BatchNormalization in training
mode; ``ctrhot_nfm_train_step_f32``) and AFM (``ctrhot_afm_train_step_f32``), FFM (``ctrhot_ffm_train_step_f32``) and DIN on its dict
input (DL/model/din.py:98-108; ``ctrhot_din_train_step_f32``) -- other models raise NotImplementedError.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from ..layer.base import as_device_f32, ids_arg, ptr, stream_ptr


class _DeepFMTrainer:
    """Owns the layer descriptors and the workspace of one (model, batch size)."""

    def __init__(self, model, dropout=None):
        from ..model.criteo import DeepFM
        if not isinstance(model, DeepFM):
            raise NotImplementedError("compile_fit: the training step exists for DeepFM, DCN, WideDeep, DeepCrossing, NFM, AFM, PNN, FM, FFM and DIN only (SURVEY 8(f) rank 3)")
        self.m = model
        if not model.dnn.built:
            model.build(None)
        if model.dnn.activation != "relu" or model.dnn.output_dim != 1:
            raise NotImplementedError("compile_fit: DNNLayer(activation='relu', output_dim=1) only")
        self.dropout = float(model.dnn.dropout if dropout is None else dropout)
        layers = [(k, b, _lib.ACT_RELU) for k, b in model.dnn.hidden] + [(model.dnn.output[0], model.dnn.output[1], _lib.ACT_NONE)]
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.err_flag()
        self.step = 0

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        """dense [B, D] float32, ids [B, F] int32 / float, y [B] float32, all on the device -> device tensor
        [mean BCE before the update, number of correct predictions] (read it when you need it: no sync here)."""
        m, E = self.m, self.m.embed_layer
        ids, kind, ids_ld = ids_arg(ids)
        B, D = ids.shape[0], dense.shape[1]
        n = D + E.F * E.k
        need = _lib.query("ctrhot_deepfm_train_workspace_bytes", B, n, m.fm.k, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_deepfm_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, ptr(y), B, ptr(m.fm.w0), ptr(m.fm.w1), ptr(m.fm.v), m.fm.k, float(m.fm.reg_w),
                  float(m.fm.reg_b), self.desc, len(self.desc), float(lr), self.dropout, int(seed) & (2 ** 64 - 1),
                  ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss

    # ---- one captured CUDA graph per (data set, batch size), replayed once per batch
    def capture(self, X, y, batch_size, lr, seed=0):
        """``X`` [N, D + F] float32 on the device (the reference's matrix: ids as float columns), ``y`` [N].  Captures ONE graph
        of ctrhot_deepfm_train_step_ctl_f32 whose batch is ``rows [row0, row0 + batch_size)`` with row0 in a device control
        block that the step itself advances: `replay()` then costs one graph launch per batch instead of ~25 kernel launches."""
        m, E = self.m, self.m.embed_layer
        D = X.shape[1] - E.F
        n = D + E.F * E.k
        need = _lib.query("ctrhot_deepfm_train_workspace_bytes", batch_size, n, m.fm.k, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        self.ctl = torch.zeros(4, dtype=torch.int64, device=m.device)          # {step, row0, (acc_loss, acc_correct), pad}
        self._graph_keep = (X, y)
        ids = X[:, D:]
        args = (ptr(E.arena), ptr(E.row_off), E.F, E.k, ids.data_ptr(), _lib.IDS_F32, X.stride(0), X.data_ptr(), X.stride(0), D,
                ptr(y), batch_size, ptr(m.fm.w0), ptr(m.fm.w1), ptr(m.fm.v), m.fm.k, float(m.fm.reg_w), float(m.fm.reg_b),
                self.desc, len(self.desc), float(lr), self.dropout, int(seed) & (2 ** 64 - 1), ptr(self.loss), ptr(self.ws),
                self.ws.numel() * 4, ptr(self.flag), ptr(self.ctl))
        torch.cuda.synchronize(m.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            _lib.call("ctrhot_deepfm_train_step_ctl_f32", *args, stream_ptr())
        return self.graph

    def epoch_begin(self):
        self.ctl[1:3].zero_()                                                  # row0 and the two accumulators; step keeps counting

    def replay(self, n_batches):
        for _ in range(n_batches):
            self.graph.replay()
        self.step += n_batches

    def epoch_totals(self):
        """-> (sum of the per-sample BCE, number of correct predictions) of the replayed batches (synchronises)."""
        acc = self.ctl[2:3].view(torch.float32).tolist()
        return float(acc[0]), float(acc[1])

    def done(self):
        """The weights moved in place: invalidate every packed copy the forward path keeps."""
        self.m._bump()


class _FMTrainer:
    """FM (DL/model/fm.py:14-23) on its dense one-hot matrix: ctrhot_fm_train_step_f32 per batch."""

    def __init__(self, model):
        self.m, self.ws, self.step = model, None, 0
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)

    def train_on_batch(self, x, y, lr):
        fm = self.m.fm
        if not fm.built:
            self.m.build((None, x.shape[1]))
        B, n = x.shape
        need = _lib.query("ctrhot_fm_train_workspace_bytes", B, n, fm.k)
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=self.m.device)
        _lib.call("ctrhot_fm_train_step_f32", ptr(x), x.stride(0), n, ptr(y), B, ptr(fm.w0), ptr(fm.w1), ptr(fm.v), fm.k,
                  float(fm.reg_w), float(fm.reg_b), float(lr), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, stream_ptr())
        self.step += 1
        return self.loss


class _FFMTrainer:
    """FFM (DL/model/ffm.py:14-23) on the reference's [B, 13 + 26] matrix (ids stored as floats): ctrhot_ffm_train_step_f32 per
    batch.  The one-hot expansion FFMLayer.call builds (interaction.py:145-151) is implied by the ids."""

    def __init__(self, model):
        self.m, self.ws, self.step = model, None, 0
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)

    def train_on_batch(self, x, y, lr):
        L = self.m.ffm
        if not L.built:
            self.m.build((None, x.shape[1]))
        D, F = len(L.dense_feature_columns), len(L.sparse_feature_columns)
        B = x.shape[0]
        need = _lib.query("ctrhot_ffm_train_workspace_bytes", B, L.field_num, L.k)
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=self.m.device)
        ids = x[:, D:]
        _lib.call("ctrhot_ffm_train_step_f32", ptr(x), x.stride(0), D, ptr(ids), _lib.IDS_F32, x.stride(0), F, ptr(L.col_off),
                  ptr(y), B, ptr(L.w0), ptr(L.w), ptr(L.v), L.feature_num, L.field_num, L.k, float(L.w_reg), float(L.v_reg),
                  float(lr), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, stream_ptr())
        self.step += 1
        return self.loss


def _accumulate(tot, out, rows, cache={}):
    """tot (float64 [2]) += (out[0] * rows, out[1]) in two launches (a product and a promoting add) instead of the seven of
    ``tot[0] += out[0].double() * rows; tot[1] += out[1].double()`` -- at the reference's batch of 32 a whole step is ~25."""
    key = (out.device, rows)
    w = cache.get(key)
    if w is None:                                                   # (_StagedStep makes this call once before its capture)
        w = cache[key] = torch.tensor([float(rows), 1.0], dtype=torch.float32, device=out.device)
    if tot is not None:
        tot.add_(out * w)


class _StagedStep:
    """One training step captured as a CUDA graph over static staging buffers: per batch the inputs are copied into the buffers
    (one small copy each) and the graph is replayed -- 3-6 launches instead of the 25-60 of a call-by-call step, which is what
    bounds the reference's batch of 32.  Only for steps without dropout (the seed is a launch parameter); the epoch's running
    (sum of per-sample losses, correct predictions) accumulates inside the graph."""

    def __init__(self, dev, buffers, step, batch_size):
        self.buffers = buffers
        self.acc = torch.zeros(2, dtype=torch.float64, device=dev)
        _accumulate(None, self.acc, batch_size)                      # the weight vector exists before the capture
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            out = step(*buffers)
            _accumulate(self.acc, out, batch_size)

    def run(self, *sources):
        for b, src in zip(self.buffers, sources):
            b.copy_(src)
        self.graph.replay()


def _no_dropout(tr):
    # `staged = False` on a trainer: measured slower as a graph (DeepCrossing's ~70 small dependent kernels at B = 32 are
    # execution-bound, not launch-bound: 1 890 steps/s call by call, 1 470 replayed; profiles/time_fit.py)
    return float(getattr(tr, "dropout", 0.0) or 0.0) == 0.0 and getattr(tr, "staged", True)


def _epochs(model, tr, N, batch_size, epochs, verbose, use_graph, dev, make_buffers, slices, step):
    """The batch loop shared by every trainer but DeepFM's: batches in order, the full ones through one replayed graph when the
    step has no dropout, the ragged last one (and everything when ``use_graph`` is off) call by call."""
    n_full = N // batch_size if (use_graph and _no_dropout(tr) and N >= 2 * batch_size) else 0
    staged = _StagedStep(dev, make_buffers(), step, batch_size) if n_full else None
    model.history = []
    for ep in range(epochs):
        tot = torch.zeros(2, dtype=torch.float64, device=dev)
        if staged is not None:
            staged.acc.zero_()
        for bi, s in enumerate(range(0, N, batch_size)):
            e = min(N, s + batch_size)
            if bi < n_full:
                staged.run(*slices(s, e))
                continue
            out = step(*slices(s, e))
            _accumulate(tot, out, e - s)
        if staged is not None:
            tot += staged.acc
        loss, acc = (tot / N).tolist()
        if hasattr(tr, "flag"):
            _lib.check(_lib.load().ctrhot_check_flag(ptr(tr.flag), stream_ptr()), "compile_fit")
        model.history.append((loss, acc))
        if verbose:
            print("Epoch %d/%d - loss: %.4f - accuracy: %.4f" % (ep + 1, epochs, loss, acc))
    model._bump()
    return model


def _fit_fm(model, X, y, batch_size, epochs, sgd, verbose, trainer=None, use_graph=True):
    tr = trainer or _FMTrainer(model)
    dev = model.device
    Xd = as_device_f32(X, dev).contiguous()
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = Xd.shape[0]
    layer = getattr(model, "fm", None) or getattr(model, "ffm")
    if not layer.built:                                             # (before any capture: build() allocates and fills weights)
        model.build((None, Xd.shape[1]))
    return _epochs(model, tr, N, batch_size, epochs, verbose, use_graph, dev,
                   lambda: [torch.zeros((batch_size, Xd.shape[1]), dtype=torch.float32, device=dev),
                            torch.zeros(batch_size, dtype=torch.float32, device=dev)],
                   lambda a, b: (Xd[a:b], yd[a:b]), lambda x, t: tr.train_on_batch(x, t, sgd))


class _DCNTrainer:
    """DCN (DL/model/dcn.py:15-34): ctrhot_dcn_train_step_f32 per batch -- gather, L cross layers, ReLU tower with Dropout, the
    final Dense over [cross | dnn]; backward through all of it, sparse SGD on the embedding rows."""

    def __init__(self, model, dropout=None):
        self.m = model
        if not model.dense_layer.built:
            model.build(None)
        dl = model.dense_layer
        if dl.activation != "relu" or dl.output_dim != 1:
            raise NotImplementedError("compile_fit: DNNLayer(activation='relu', output_dim=1) only")
        self.dropout = float(dl.dropout if dropout is None else dropout)
        layers = [(k, b, _lib.ACT_RELU) for k, b in dl.hidden] + [(dl.output[0], dl.output[1], _lib.ACT_NONE)]
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        cl = model.cross_layer
        self.cw = np.array([t.data_ptr() for t in cl.cross_weight], dtype=np.int64)
        self.cb = np.array([t.data_ptr() for t in cl.cross_bias], dtype=np.int64)
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.err_flag()
        self.step = 0

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        m, E, cl = self.m, self.m.embed_layer, self.m.cross_layer
        ids, kind, ids_ld = ids_arg(ids)
        B, D = ids.shape[0], dense.shape[1]
        n = D + E.F * E.k
        need = _lib.query("ctrhot_dcn_train_workspace_bytes", B, n, cl.layer_num, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_dcn_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, ptr(y), B, self.cw.ctypes.data, self.cb.ctypes.data, cl.layer_num, float(cl.reg_w),
                  float(cl.reg_b), self.desc, len(self.desc), ptr(m.out_kernel), ptr(m.out_bias), float(lr), self.dropout,
                  int(seed) & (2 ** 64 - 1), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


def _fit_dcn(model, X, y, batch_size, epochs, sgd, dropout, seed, verbose):
    tr = _DCNTrainer(model, dropout)
    dev = model.device
    D = len(model.dense_feature_columns)
    Xd = as_device_f32(X, dev).contiguous()
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = Xd.shape[0]
    model.history = []
    for ep in range(epochs):
        tot = torch.zeros(2, dtype=torch.float64, device=dev)
        for s in range(0, N, batch_size):
            e = min(N, s + batch_size)
            out = tr.train_on_batch(Xd[s:e, :D], Xd[s:e, D:], yd[s:e], sgd, seed=seed * 1000003 + tr.step)
            _accumulate(tot, out, e - s)
        loss, acc = (tot / N).tolist()
        _lib.check(_lib.load().ctrhot_check_flag(ptr(tr.flag), stream_ptr()), "compile_fit")
        model.history.append((loss, acc))
        if verbose:
            print("Epoch %d/%d - loss: %.4f - accuracy: %.4f" % (ep + 1, epochs, loss, acc))
    model._bump()
    model.cross_layer._bump()
    return model


class _PNNTrainer:
    """PNN's own loop (DL/model/pnn.py:74-85) as ctrhot_pnn_train_step_f32 per batch: BCE on the model's raw output (PNN.call has
    no sigmoid; keras clips), no regularisers, sparse SGD on rows / outer-product weights / tower."""

    _MODE = {"inner": 1, "outer": 2, "both": 3}

    def __init__(self, model, dropout=None):
        self.m = model
        if not model.dnn_layer.built:
            model.build(None)
        dl = model.dnn_layer
        if dl.activation != "relu" or dl.output_dim != 1:
            raise NotImplementedError("compile_fit: DNNLayer(activation='relu', output_dim=1) only")
        self.dropout = float(dl.dropout if dropout is None else dropout)
        layers = [(k, b, _lib.ACT_RELU) for k, b in dl.hidden] + [(dl.output[0], dl.output[1], _lib.ACT_NONE)]
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.embed_layer.err_flag()
        self.step = 0

    def train_on_batch(self, ids, y, lr, seed=0):
        m, E = self.m, self.m.embed_layer
        ids, kind, ids_ld = ids_arg(ids)
        B = ids.shape[0]
        mode = self._MODE[m.mode]
        need = _lib.query("ctrhot_pnn_train_workspace_bytes", B, E.F, E.k, mode, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_pnn_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(y), B,
                  ptr(m.outer_product_layer.w), mode, self.desc, len(self.desc), float(lr), self.dropout,
                  int(seed) & (2 ** 64 - 1), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


def _fit_pnn(model, X, y, batch_size, epochs, sgd, dropout, seed, verbose):
    tr = _PNNTrainer(model, dropout)
    dev = model.device
    D = len(model.dense_feature_columns)
    Xd = as_device_f32(X, dev).contiguous()
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = Xd.shape[0]
    model.history = []
    for ep in range(epochs):
        tot = torch.zeros(2, dtype=torch.float64, device=dev)
        for s in range(0, N, batch_size):
            e = min(N, s + batch_size)
            out = tr.train_on_batch(Xd[s:e, D:], yd[s:e], sgd, seed=seed * 1000003 + tr.step)
            _accumulate(tot, out, e - s)
        loss, acc = (tot / N).tolist()
        _lib.check(_lib.load().ctrhot_check_flag(ptr(tr.flag), stream_ptr()), "compile_fit")
        model.history.append((loss, acc))
        if verbose:
            print("Epoch %d/%d - loss: %.4f - accuracy: %.4f" % (ep + 1, epochs, loss, acc))
    model._bump()
    return model


class _WideDeepTrainer:
    """WideDeep (DL/model/wideDeep.py:14-34): ctrhot_widedeep_train_step_f32 per batch -- the wide logit and its gradient on the
    ids (the one-hot matrix is never built), the deep tower on the embeddings, WideLayer's l2(1e-4) on w."""

    REG_W = 1e-4                                                  # WideLayer.build, DL/layer/interaction.py:19-22

    def __init__(self, model, dropout=None):
        self.m = model
        if not model.deep.built:
            model.build(None)
        dl = model.deep
        if dl.activation != "relu" or dl.output_dim != 1:
            raise NotImplementedError("compile_fit: DNNLayer(activation='relu', output_dim=1) only")
        self.dropout = float(dl.dropout if dropout is None else dropout)
        layers = [(k, b, _lib.ACT_RELU) for k, b in dl.hidden] + [(dl.output[0], dl.output[1], _lib.ACT_NONE)]
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.embed_layer.err_flag()
        self.step = 0

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        m, E = self.m, self.m.embed_layer
        ids, kind, ids_ld = ids_arg(ids)
        B, D = ids.shape[0], dense.shape[1]
        n = E.F * E.k
        need = _lib.query("ctrhot_widedeep_train_workspace_bytes", B, n, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_widedeep_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, ptr(y), B, ptr(m.wide.w0), ptr(m.wide.w), ptr(m.col_off), m.n_wide, self.REG_W, self.desc,
                  len(self.desc), float(lr), self.dropout, int(seed) & (2 ** 64 - 1), ptr(self.loss), ptr(self.ws),
                  self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


def _fit_rows(tr, model, X, y, batch_size, epochs, sgd, seed, verbose, use_graph=True):
    """The trainers that take (dense, ids, y): see _epochs."""
    dev = model.device
    D = len(model.dense_feature_columns)
    F = model.embed_layer.F
    Xd = as_device_f32(X, dev).contiguous()
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = Xd.shape[0]
    return _epochs(model, tr, N, batch_size, epochs, verbose, use_graph, dev,
                   lambda: [torch.zeros((batch_size, D), dtype=torch.float32, device=dev),
                            torch.zeros((batch_size, F), dtype=torch.float32, device=dev),
                            torch.zeros(batch_size, dtype=torch.float32, device=dev)],
                   lambda a, b: (Xd[a:b, :D], Xd[a:b, D:D + F], yd[a:b]),
                   lambda d_, i_, t: tr.train_on_batch(d_, i_, t, sgd, seed=seed * 1000003 + tr.step))


class _DeepCrossingTrainer:
    """DeepCrossing (DL/model/deepCrossing.py:14-32): ctrhot_deepcrossing_train_step_f32 per batch -- the residual units'
    Dense chains, their relu(x + h), the final Dense(1, sigmoid); no dropout, no regularisers."""

    def __init__(self, model, dropout=None):
        self.m = model
        if not hasattr(model, "output_layer"):
            model.build(None)
        layers = []
        for r in model.res_layer:
            layers += [(k, b, _lib.ACT_RELU) for k, b in r.dense_layer] + [(r.output_layer[0], r.output_layer[1], _lib.ACT_NONE)]
        self.per_res = len(model.res_layer[0].dense_layer) + 1
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.embed_layer.err_flag()
        self.step = 0
        self.staged = False

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        m, E = self.m, self.m.embed_layer
        ids, kind, ids_ld = ids_arg(ids)
        B, D = ids.shape[0], dense.shape[1]
        n = D + E.F * E.k
        n_res = len(m.res_layer)
        need = _lib.query("ctrhot_deepcrossing_train_workspace_bytes", B, n, self.desc, n_res, self.per_res)
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_deepcrossing_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, ptr(y), B, self.desc, n_res, self.per_res, ptr(m.output_layer[0]), ptr(m.output_layer[1]),
                  float(lr), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


class _NFMTrainer:
    """NFM (DL/model/nfm.py:15-34) in training mode: ctrhot_nfm_train_step_f32 per batch -- Bi-Interaction, BatchNormalization on
    the batch's own statistics (moving statistics updated with keras' momentum 0.99), DNNLayer, the final Dense(1)."""

    def __init__(self, model, dropout=None):
        self.m = model
        if not model.dnn_layers.built:
            model.build(None)
        dl = model.dnn_layers
        if dl.activation != "relu":
            raise NotImplementedError("compile_fit: DNNLayer(activation='relu') only")
        self.dropout = float(dl.dropout if dropout is None else dropout)
        layers = [(k, b, _lib.ACT_RELU) for k, b in dl.hidden] + [(dl.output[0], dl.output[1], _lib.ACT_NONE)]
        self.desc = (_lib.DenseDesc * len(layers))()
        for d, (k, b, act) in zip(self.desc, layers):
            d.weight, d.bias, d.in_dim, d.out_dim, d.act = ptr(k), ptr(b), int(k.shape[0]), int(k.shape[1]), act
        self._keep = layers
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.err_flag()
        self.step = 0

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        m, E = self.m, self.m.emb_layers
        ids, kind, ids_ld = ids_arg(ids)
        B, D = ids.shape[0], dense.shape[1]
        need = _lib.query("ctrhot_nfm_train_workspace_bytes", B, D, E.F, E.k, self.desc, len(self.desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_nfm_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(dense),
                  dense.stride(0), D, ptr(y), B, ptr(m.bn_gamma), ptr(m.bn_beta), ptr(m.bn_mean), ptr(m.bn_var), 0.99, 1e-3,
                  self.desc, len(self.desc), ptr(m.out_kernel), ptr(m.out_bias), float(lr), self.dropout,
                  int(seed) & (2 ** 64 - 1), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


class _AFMTrainer:
    """AFM / AFMLayer (DL/model/afm.py, DL/layer/interaction.py:322-351): ctrhot_afm_train_step_f32 per batch.  In 'att' mode
    AttentionLayer's two Dense layers receive zero gradient (softmax over a size-1 axis) and stay as they are."""

    def __init__(self, model, dropout=None):
        self.m = model
        if not hasattr(model, "out_kernel"):
            model.build(None)
        self.ws = None
        self.loss = torch.zeros(2, dtype=torch.float32, device=model.device)
        self.flag = model.err_flag()
        self.step = 0

    def train_on_batch(self, dense, ids, y, lr, seed=0):
        m, E = self.m, self.m.embed_layer
        ids, kind, ids_ld = ids_arg(ids)
        B = ids.shape[0]
        need = _lib.query("ctrhot_afm_train_workspace_bytes", B, E.F, E.k)
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=m.device)
        _lib.call("ctrhot_afm_train_step_f32", ptr(E.arena), ptr(E.row_off), E.F, E.k, ptr(ids), kind, ids_ld, ptr(y), B,
                  {"avg": 0, "att": 1, "max": 2}[m.mode], int(m.n_sigmoid), ptr(m.out_kernel), ptr(m.out_bias), float(lr),
                  ptr(self.loss), ptr(self.ws), self.ws.numel() * 4, ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


class _DINTrainer:
    """DIN (DL/model/din.py:56-95 and its ``__main__``, :98-108): ctrhot_din_train_step_f32 per batch.  One behaviour feature,
    PReLU attention unit and tower, dnn_dropout 0 -- the reference's defaults; BatchNormalization runs on the batch's statistics
    and its moving statistics are the ones the forward path serves afterwards."""

    def __init__(self, model):
        import ctypes
        m = self.m = model
        if m.behavior_num != 1 or m.att_layer.activation != "prelu" or m.dnn_activation != "prelu" or m.dnn_dropout:
            raise NotImplementedError("compile_fit(DIN): one behaviour feature, 'prelu' attention and tower, dnn_dropout = 0 only")
        self.ws, self.step, self.ct = None, 0, ctypes
        self.loss = torch.zeros(2, dtype=torch.float32, device=m.device)
        self.flag = m.err_flag()
        self.desc = None

    def _descs(self):
        m, ct = self.m, self.ct
        att, tw = m.att_layer.layers, m.dense_layer
        self.att_desc = (_lib.DenseDesc * len(att))()
        self.tw_desc = (_lib.DenseDesc * len(tw))()
        for arr, layers, rows in ((self.att_desc, att, m.att_layer.T), (self.tw_desc, tw, 1)):
            for d, (k, b, a) in zip(arr, layers):
                d.weight, d.bias, d.act_param = ptr(k), ptr(b), ptr(a)
                d.in_dim, d.out_dim, d.act, d.alpha_rows = int(k.shape[0]), int(k.shape[1]), _lib.ACT_PRELU, rows
        tabs = m.embed_sparse_layers
        self.n_other = len(tabs)
        self.o_ptr = (ct.c_void_p * max(1, len(tabs)))(*[ptr(t) for t in tabs])
        self.o_rows = (ct.c_int64 * max(1, len(tabs)))(*[int(t.shape[0]) for t in tabs])
        self.o_dims = (ct.c_int32 * max(1, len(tabs)))(*[int(t.shape[1]) for t in tabs])
        self.desc = m._version

    def train_on_batch(self, inputs, y, lr):
        from ..layer.core import _ids_to_device
        m, dev = self.m, self.m.device
        hist = _ids_to_device(m._col(inputs, m.behavior[0]["feat"]), dev).to(torch.int32).contiguous()
        B, T = hist.shape
        if not m.att_layer.built:
            if not hasattr(m, "bn_gamma"):
                m.build(None)
            m.att_layer.build_for(T, m.K)
        elif not hasattr(m, "bn_gamma"):
            m.build(None)
        if T != m.att_layer.T:
            raise ValueError("DIN: history length %d, the attention unit was built for %d" % (T, m.att_layer.T))
        if self.desc is None:
            self._descs()
        cand = _ids_to_device(m._col(inputs, m.candidate_key), dev).reshape(B, -1)[:, 0].to(torch.int32).contiguous()
        if self.n_other:
            oid = torch.stack([_ids_to_device(m._col(inputs, f["feat"]), dev).reshape(B, -1)[:, 0].to(torch.int32)
                               for f in m.other_sparse], dim=1).contiguous()
        else:
            oid = None
        dense = None
        if m.dense_num:
            cols = [as_device_f32(m._col(inputs, f["feat"]), dev).reshape(B, -1) for f in m.dense_feature_columns]
            dense = (cols[0] if len(cols) == 1 else torch.cat(cols, dim=-1)).contiguous()
        yd = as_device_f32(y, dev).reshape(-1)
        need = _lib.query("ctrhot_din_train_workspace_bytes", B, T, m.K, m.dcat, self.att_desc, len(self.att_desc), self.tw_desc,
                          len(self.tw_desc))
        if self.ws is None or self.ws.numel() * 4 < need:
            self.ws = torch.empty(need // 4 + 4, dtype=torch.float32, device=dev)
        seq = m.embed_seq_layers[0]
        A = m.att_layer
        _lib.call("ctrhot_din_train_step_f32", ptr(seq), int(seq.shape[0]), m.K, ptr(hist), hist.stride(0), T, ptr(cand),
                  self.o_ptr if self.n_other else None, self.o_rows if self.n_other else None,
                  self.o_dims if self.n_other else None, self.n_other, ptr(oid) if oid is not None else None,
                  self.n_other, ptr(dense) if dense is not None else None, dense.stride(0) if dense is not None else 0,
                  m.dense_num, ptr(yd), B, self.att_desc, len(self.att_desc), ptr(A.out_kernel), ptr(A.out_bias),
                  ptr(m.bn_gamma), ptr(m.bn_beta), ptr(m.bn_mean), ptr(m.bn_var), 0.99, 1e-3, self.tw_desc, len(self.tw_desc),
                  ptr(m.out_kernel), ptr(m.out_bias), float(lr), ptr(self.loss), ptr(self.ws), self.ws.numel() * 4,
                  ptr(self.flag), stream_ptr())
        self.step += 1
        return self.loss


def _fit_din(model, X, y, batch_size, epochs, sgd, verbose, use_graph=True):
    """X: the reference's dict of per-feature lists / arrays (DL/model/din.py:101-105); batches in order."""
    from ..layer.core import _ids_to_device
    tr = _DINTrainer(model)
    dev = model.device
    names = [f["feat"] for f in model.dense_feature_columns] + [f["feat"] for f in model.sparse_feature_columns]
    if model.candidate_key not in names:
        names.append(model.candidate_key)
    dense_names = set(f["feat"] for f in model.dense_feature_columns)
    Xd = {n: (as_device_f32(np.asarray(X[n], dtype=np.float32) if not isinstance(X[n], torch.Tensor) else X[n], dev)
              if n in dense_names else _ids_to_device(X[n], dev)) for n in names}
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = yd.shape[0]
    keys = list(Xd)
    if not hasattr(model, "bn_gamma"):                              # (before any capture: build() allocates and fills weights)
        model.build(None)
    if not model.att_layer.built:
        model.att_layer.build_for(int(Xd[model.behavior[0]["feat"]].shape[1]), model.K)
    tr._descs()
    return _epochs(model, tr, N, batch_size, epochs, verbose, use_graph, dev,
                   lambda: [torch.zeros((batch_size,) + tuple(Xd[n].shape[1:]), dtype=Xd[n].dtype, device=dev) for n in keys] +
                           [torch.zeros(batch_size, dtype=torch.float32, device=dev)],
                   lambda a, b: tuple(Xd[n][a:b] for n in keys) + (yd[a:b],),
                   lambda *args: tr.train_on_batch(dict(zip(keys, args[:-1])), args[-1], sgd))


def compile_fit(model, X, y, batch_size=32, epochs=10, sgd=0.01, dropout=None, seed=0, verbose=True, use_graph=True):
    """``X`` the reference's [N, 39] matrix (13 dense floats, 26 label-encoded ids stored as floats), numpy or torch; ``y`` [N].
    ``dropout=None`` uses the model's DNNLayer rate (0.2 in the reference).  Returns the model, trained in place; the per-epoch
    (loss, accuracy) pairs are left in ``model.history``.  ``use_graph``: the full batches of an epoch replay one captured CUDA
    graph (the step is launch-bound at the reference's batch size); the ragged last batch is an ordinary call."""
    from ..model.criteo import AFMLayer, DCN, FFM, FM, NFM, PNN, WideDeep
    from ..model.deepCrossing import DeepCrossing
    from ..model.din import DIN
    if isinstance(model, DIN):
        return _fit_din(model, X, y, batch_size, epochs, sgd, verbose, use_graph)
    if isinstance(model, FFM):
        return _fit_fm(model, X, y, batch_size, epochs, sgd, verbose, trainer=_FFMTrainer(model), use_graph=use_graph)
    if isinstance(model, AFMLayer):                               # AFM is a subclass
        return _fit_rows(_AFMTrainer(model, dropout), model, X, y, batch_size, epochs, sgd, seed, verbose, use_graph)
    if isinstance(model, NFM):
        model.embed_layer = model.emb_layers                      # (_fit_rows reads the field count from .embed_layer)
        return _fit_rows(_NFMTrainer(model, dropout), model, X, y, batch_size, epochs, sgd, seed, verbose, use_graph)
    if isinstance(model, DeepCrossing):
        return _fit_rows(_DeepCrossingTrainer(model, dropout), model, X, y, batch_size, epochs, sgd, seed, verbose, use_graph)
    if isinstance(model, WideDeep):
        # the one-hot columns of the reference's [B, 39 + n_onehot] matrix are implied by the ids (model.onehot_width)
        return _fit_rows(_WideDeepTrainer(model, dropout), model, X, y, batch_size, epochs, sgd, seed, verbose, use_graph)
    if isinstance(model, PNN):
        return _fit_pnn(model, X, y, batch_size, epochs, sgd, dropout, seed, verbose)
    if isinstance(model, FM):
        return _fit_fm(model, X, y, batch_size, epochs, sgd, verbose, use_graph=use_graph)
    if isinstance(model, DCN):
        return _fit_dcn(model, X, y, batch_size, epochs, sgd, dropout, seed, verbose)
    tr = _DeepFMTrainer(model, dropout)
    dev = model.device
    D = len(model.dense_feature_columns)
    Xd = as_device_f32(X, dev).contiguous()
    yd = as_device_f32(np.asarray(y, dtype=np.float32).reshape(-1) if not isinstance(y, torch.Tensor) else y.reshape(-1), dev)
    N = Xd.shape[0]
    n_full = N // batch_size if use_graph else 0
    if n_full:
        tr.capture(Xd, yd, batch_size, sgd, seed=seed * 1000003)
    model.history = []
    for ep in range(epochs):
        tot = torch.zeros(2, dtype=torch.float64, device=dev)
        if n_full:
            tr.epoch_begin()
            tr.replay(n_full)
        for s in range(n_full * batch_size, N, batch_size):
            e = min(N, s + batch_size)
            out = tr.train_on_batch(Xd[s:e, :D], Xd[s:e, D:], yd[s:e], sgd, seed=seed * 1000003 + tr.step)
            if n_full:
                tr.ctl[0] += 1                                 # the graph's step counter numbers EVERY batch (dropout seeds)
            _accumulate(tot, out, e - s)
        loss_sum, correct = tot.tolist()
        if n_full:
            a, b = tr.epoch_totals()
            loss_sum, correct = loss_sum + a, correct + b
        loss, acc = loss_sum / N, correct / N
        _lib.check(_lib.load().ctrhot_check_flag(ptr(tr.flag), stream_ptr()), "compile_fit")
        model.history.append((loss, acc))
        if verbose:
            print("Epoch %d/%d - loss: %.4f - accuracy: %.4f" % (ep + 1, epochs, loss, acc))
    tr.done()
    return model
