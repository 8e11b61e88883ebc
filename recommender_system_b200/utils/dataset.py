"""Drop-in for the Criteo half of the reference's ``utils/dataset.py`` (DL/utils/dataset.py:19-75): the input step on the left
of the hot path, on the GPU.

    create_criteo_dataset(t, file_path, test_size=0.3)      same signature, same return value as the reference
    features_dict(file_path, embed_dim=8)                   same
    CriteoEncoder                                           the pieces: parse (CUDA) -> fit (torch, once) -> transform (CUDA)

The reference reads the TSV with pandas, fills missing values (0 / '-1'), MinMax-scales the 13 counts with sklearn and
label-encodes the 26 categorical columns with sklearn, one Python call per column: 17 s for the bundled 10 000 rows in
'fm' mode, 1 s in 'DeepFM' mode.  Here the file's bytes go to the device once; a kernel parses every line
(csrc/encode.cu), the fit (min / max per count, sorted distinct tokens per column) uses torch reductions / ``torch.unique``
on the device, and a second kernel produces ``dense`` float32 and ``ids`` int32 -- what every model's fast path takes.
The split itself is sklearn's ``train_test_split`` on row numbers (host, a permutation): same unseeded-by-default behaviour
as the reference (dataset.py:63).  There is no CPU path: without a CUDA device the encoder raises.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from ..layer.base import default_device, ptr, stream_ptr

criteo_dense_features = ["I" + str(i) for i in range(1, 14)]
criteo_sparse_features = ["C" + str(i) for i in range(1, 27)]
LABEL_ENCODED = ("ffm", "DeepCrossing", "pnn", "dcn", "DeepFM", "nfm")        # dataset.py:49
ONE_HOT = ("fm", "fnn")                                                       # dataset.py:47


def sparseFeature(feat, feat_onehot_dim, embed_dim):
    return {"feat": feat, "feat_onehot_dim": feat_onehot_dim, "embed_dim": embed_dim}


def denseFeature(feat):
    return {"feat": feat}


class CriteoEncoder:
    def __init__(self, device=None):
        self.device = torch.device(device) if device is not None else default_device()
        if self.device.type != "cuda":
            raise _lib.CtrhotError("CriteoEncoder: a CUDA device is required -- ctrhot has no CPU fallback")
        self.err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.fitted = False

    # ---------------------------------------------------------------- parse
    def parse(self, text, has_label=True):
        """text: bytes / uint8 array / uint8 tensor holding the TSV file.  -> (label int8 [N] or None, dense_raw float64 [N, 13]
        with NaN for empty fields, tokens int64 [N, 26] with -1 for empty fields), all on the device."""
        if isinstance(text, (bytes, bytearray, memoryview)):
            text = torch.frombuffer(bytearray(text), dtype=torch.uint8)
        elif isinstance(text, np.ndarray):
            text = torch.from_numpy(np.ascontiguousarray(text, dtype=np.uint8))
        buf = text.to(self.device).contiguous()
        n_bytes = buf.numel()
        nl = torch.nonzero(buf == 10).flatten()
        ends = nl + 1
        if n_bytes and (nl.numel() == 0 or int(nl[-1]) != n_bytes - 1):
            ends = torch.cat([ends, torch.tensor([n_bytes], dtype=torch.int64, device=self.device)])
        line_start = torch.cat([torch.zeros(1, dtype=torch.int64, device=self.device), ends]).contiguous()
        N = line_start.numel() - 1
        label = torch.empty(N, dtype=torch.int8, device=self.device) if has_label else None
        dense_raw = torch.empty((N, 13), dtype=torch.float64, device=self.device)
        tokens = torch.empty((N, 26), dtype=torch.int64, device=self.device)
        _lib.call("ctrhot_criteo_parse_tsv", ptr(buf), ptr(line_start), N, int(has_label), ptr(label), ptr(dense_raw), ptr(tokens),
                  ptr(self.err), stream_ptr())
        self._check("malformed field in the TSV text")
        return label, dense_raw, tokens

    def _check(self, what):
        if int(self.err.item()) != 0:
            self.err.zero_()
            raise ValueError("CriteoEncoder: " + what)

    # ---------------------------------------------------------------- fit
    def fit(self, dense_raw, tokens):
        """MinMaxScaler.fit on the filled counts (dataset.py:42,46) and LabelEncoder.fit per column (dataset.py:43,51-52)."""
        filled = torch.nan_to_num(dense_raw, nan=0.0)
        lo, hi = filled.amin(0).cpu().numpy(), filled.amax(0).cpu().numpy()
        rng = hi - lo
        rng[rng == 0.0] = 1.0                                  # sklearn: a constant column keeps scale 1
        scale = 1.0 / rng
        self.scale = torch.from_numpy(scale).to(self.device)
        self.min_ = torch.from_numpy(0.0 - lo * scale).to(self.device)
        vocabs = [torch.unique(tokens[:, f], sorted=True) for f in range(26)]
        self.classes = vocabs                                   # LabelEncoder.classes_ (as token values; -1 = '-1')
        off = np.concatenate([[0], np.cumsum([v.numel() for v in vocabs])]).astype(np.int64)
        self.vocab = torch.cat(vocabs).contiguous()
        self.vocab_off = torch.from_numpy(off).to(self.device)
        self.n_classes = [int(v.numel()) for v in vocabs]
        # features_dict (dataset.py:69-75) counts distinct NON-missing tokens of the raw file, plus one
        self.feat_onehot_dim = [int(v.numel()) - int((v[0] == -1).item()) + 1 if v.numel() else 1 for v in vocabs]
        self.fitted = True
        return self

    # ---------------------------------------------------------------- transform
    def transform(self, dense_raw, tokens, dense=None, ids=None):
        """-> dense float32 [N, 13] (MinMax-scaled), ids int32 [N, 26] (label-encoded), on the device."""
        if not self.fitted:
            raise RuntimeError("CriteoEncoder.transform before fit")
        N = dense_raw.shape[0]
        if dense is None:
            dense = torch.empty((N, 13), dtype=torch.float32, device=self.device)
        if ids is None:
            ids = torch.empty((N, 26), dtype=torch.int32, device=self.device)
        _lib.call("ctrhot_criteo_encode", ptr(dense_raw), ptr(tokens), N, ptr(self.scale), ptr(self.min_), ptr(self.vocab),
                  ptr(self.vocab_off), ptr(dense), dense.stride(0), ptr(ids), ids.stride(0), ptr(self.err), stream_ptr())
        self._check("a token that the encoder was not fitted on (LabelEncoder raises on unseen labels)")
        return dense, ids

    def fit_transform_text(self, text, has_label=True):
        label, dense_raw, tokens = self.parse(text, has_label)
        self.fit(dense_raw, tokens)
        dense, ids = self.transform(dense_raw, tokens)
        return label, dense, ids


def _read(file_path):
    with open(file_path, "rb") as f:
        return f.read()


def features_dict(file_path, embed_dim=8, device=None):
    """DL/utils/dataset.py:69-75: ``feat_onehot_dim = nunique + 1`` over the raw (unfilled) column."""
    enc = CriteoEncoder(device)
    _, dense_raw, tokens = enc.parse(_read(file_path))
    enc.fit(dense_raw, tokens)
    return [[denseFeature(f) for f in criteo_dense_features],
            [sparseFeature(f, v, embed_dim) for f, v in zip(criteo_sparse_features, enc.feat_onehot_dim)]]


def create_criteo_dataset(t, file_path, test_size=0.3, device=None):
    """DL/utils/dataset.py:36-66.  Returns ``(X_train, y_train), (X_test, y_test)`` as the reference does: numpy, X float64
    ``[n, 39]`` = 13 scaled counts then 26 label-encoded ids for the label-encoding modes; for 'fm' / 'fnn' the dense one-hot
    matrix of ``pd.get_dummies`` (13 counts, then one 0/1 column per distinct value of every categorical column, in sorted
    order); for 'WideDeep' the two side by side (dataset.py:53-58).  The scaled counts are the float32 the models consume,
    widened to float64 (the reference keeps sklearn's float64; the difference is below 6e-8 and disappears at the models' own
    float32 cast)."""
    from sklearn.model_selection import train_test_split
    enc = CriteoEncoder(device)
    label, dense, ids = enc.fit_transform_text(_read(file_path))
    y = label.cpu().numpy().astype(np.int64)
    dense_h, ids_h = dense.cpu().numpy().astype(np.float64), ids.cpu().numpy()
    if t in LABEL_ENCODED:
        X = np.concatenate([dense_h, ids_h.astype(np.float64)], axis=1)
    elif t in ONE_HOT or t == "WideDeep":
        off = 13 + np.concatenate([[0], np.cumsum(enc.n_classes)])
        onehot = torch.zeros((ids.shape[0], int(off[-1])), dtype=torch.float32, device=enc.device)
        onehot[:, :13] = dense
        cols = ids.long() + torch.from_numpy(off[:-1]).to(enc.device)[None, :]
        onehot.scatter_(1, cols, 1.0)
        oh = onehot.cpu().numpy().astype(np.float64)
        X = oh if t in ONE_HOT else np.concatenate([dense_h, ids_h.astype(np.float64), oh], axis=1)
    else:
        raise ValueError("unknown dataset mode %r" % (t,))
    X_train, X_test, y_train, y_test = train_test_split(X, y, test_size=test_size)
    return (X_train, y_train), (X_test, y_test)
