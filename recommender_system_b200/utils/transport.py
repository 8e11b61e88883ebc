"""Compact transport records for Criteo-shaped batches (csrc/transport.cu): the host-to-device copy on the left of the hot
path.  The reference feeds ``model(X)`` one float64 ``[B, 39]`` matrix (DL/utils/dataset.py:61,
DL/model/deepFM.py:24); a record keeps the 13 dense float32 bit-exact and stores every id at the bit width its
``feat_onehot_dim`` needs -- 96 B per sample instead of 156 (int32 ids) or 312 (the reference's matrix) on the
Criteo-Kaggle cardinalities (104 B with whole-byte fields, ``bits=False``).

    layout = RowLayout.from_feature_columns(feature_columns)         # widths from feat_onehot_dim
    rec = layout.pack(dense, ids)                                    # host: uint8 [B, rec_bytes] (C, multi-threaded)
    p = model(PackedRows(layout, rec_on_device))                     # DeepFM: decoded inside the whole-model kernel
    dense_d, ids_d = layout.unpack(rec_on_device)                    # any other model: one extra launch
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib


class RowLayout:
    def __init__(self, vocab, n_dense=13, bits=True):
        """``bits=True``: every id at the BIT width its vocabulary needs (96 B per sample on the Criteo-Kaggle cardinalities);
        ``bits=False``: whole bytes (104 B).  ``widths``: per field, in bits / in bytes."""
        self.vocab = [int(v) for v in vocab]
        self.F, self.D = len(self.vocab), int(n_dense)
        self.bits = bool(bits)
        v = (ctypes.c_int64 * self.F)(*self.vocab)
        d = (ctypes.c_int32 * self.F)()
        self.rec_bytes = _lib.query("ctrhot_rows_layout_bits" if self.bits else "ctrhot_rows_layout", v, self.F, self.D, d)
        self.desc_host = np.frombuffer(d, dtype=np.int32).copy()
        self.widths = [(int(x) >> 24) & 0x7F if self.bits else int(x) >> 16 for x in self.desc_host]
        self.dense_off = 0
        self._dev = {}

    @classmethod
    def from_feature_columns(cls, feature_columns):
        dense_cols, sparse_cols = feature_columns
        return cls([f["feat_onehot_dim"] for f in sparse_cols], len(dense_cols))

    def desc(self, device):
        """The per-field descriptor on ``device`` (uploaded once per device)."""
        key = str(device)
        if key not in self._dev:
            self._dev[key] = torch.from_numpy(self.desc_host).to(device)
        return self._dev[key]

    def pack(self, dense, ids, out=None, threads=0):
        """Host: dense float32 [B, D], ids int32 [B, F] -> uint8 [B, rec_bytes] (``out``: e.g. a pinned torch tensor).
        IndexError if an id does not fit its field's width (i.e. is >= 2^(8 width) >= feat_onehot_dim)."""
        dense = np.ascontiguousarray(dense, dtype=np.float32)
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        B = ids.shape[0]
        if out is None:
            out = np.empty((B, self.rec_bytes), dtype=np.uint8)
        buf = out.numpy() if isinstance(out, torch.Tensor) else out
        assert buf.shape == (B, self.rec_bytes) and buf.dtype == np.uint8 and buf.flags["C_CONTIGUOUS"]
        rc = _lib.load().ctrhot_rows_pack_host(dense.ctypes.data, dense.shape[1] if self.D else 0, ids.ctypes.data, ids.shape[1],
                                               B, self.desc_host.ctypes.data, self.F, self.D, self.rec_bytes, buf.ctypes.data,
                                               int(threads))
        _lib.check(rc, "ctrhot_rows_pack_host")
        return out

    def unpack(self, records, dense=None, ids=None):
        """Device: uint8 [M, rec_bytes] -> (dense float32 [M, D], ids int32 [M, F]) by one kernel launch."""
        from ..layer.base import ptr, stream_ptr
        if not records.is_cuda:
            raise RuntimeError("RowLayout.unpack needs device records (there is no CPU path)")
        M = records.shape[0]
        dev = records.device
        if dense is None:
            dense = torch.empty((M, max(self.D, 1)), dtype=torch.float32, device=dev)[:, :self.D]
        if ids is None:
            ids = torch.empty((M, self.F), dtype=torch.int32, device=dev)
        _lib.call("ctrhot_rows_unpack", ptr(records), records.stride(0), ptr(self.desc(dev)), self.F, self.dense_off, self.D, M,
                  ptr(dense) if self.D else 0, dense.stride(0) if self.D else 0, ptr(ids), ids.stride(0), stream_ptr())
        return dense, ids


class PackedRows:
    """A batch in transport-record form: ``records`` uint8 [B, layout.rec_bytes] on the device."""

    def __init__(self, layout, records):
        if records.dtype != torch.uint8 or records.dim() != 2 or records.shape[1] != layout.rec_bytes or records.stride(1) != 1:
            raise ValueError("records must be uint8 [B, %d]" % layout.rec_bytes)
        self.layout, self.records = layout, records

    @property
    def shape(self):
        return (self.records.shape[0], self.layout.D + self.layout.F)
