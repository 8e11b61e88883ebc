"""Drop-in for the reference's ``Hash`` layer (DL/layer/utils.py:21-56) on libctrhot.

``Hash(num_buckets, mask_zero=False)(x)``: bucket of FarmHash Fingerprint64 of each value's string form
(``tf.strings.to_hash_bucket_fast``; integers as their decimal string, ``tf.as_string``), modulo ``num_buckets`` -- or modulo
``num_buckets - 1``, shifted up by one, with the value ``0`` / ``"0"`` mapped to 0, when ``mask_zero``.  Returns int64 like TensorFlow.
The ``vocabulary_path`` branch (a ``StaticHashTable`` read from a text file, utils.py:27-29, 42-44) is a file-backed dictionary,
not the hash: it raises NotImplementedError.  Float inputs raise too: ``tf.as_string`` of a float is a printf format this layer
does not restate (the reference feeds int32 ids, DL/utils/inputs.py:60-70)."""
from __future__ import annotations

import numpy as np
import torch

from .. import _lib
from .base import Layer, ptr, stream_ptr


class Hash(Layer):
    def __init__(self, num_buckets, mask_zero=False, vocabulary_path=None, default_value=0, device=None, **kwargs):
        super().__init__(device)
        if vocabulary_path:
            raise NotImplementedError("Hash(vocabulary_path=...): the file-backed StaticHashTable branch is outside the hot path")
        self.num_buckets = int(num_buckets)
        self.mask_zero = bool(mask_zero)
        self.vocabulary_path = vocabulary_path
        self.default_value = default_value
        self.built = True
        self._err = None

    def compute_output_shape(self, input_shape):
        return input_shape

    def get_config(self):
        return {"num_buckets": self.num_buckets, "mask_zero": self.mask_zero, "vocabulary_apth": self.vocabulary_path,
                "default_value": self.default_value}

    def call(self, x, mask=None, **kwargs):
        if self.device.type != "cuda":
            raise _lib.CtrhotError("Hash: a CUDA device is required -- ctrhot has no CPU fallback")
        if isinstance(x, torch.Tensor):
            t = x
        else:
            a = np.asarray(x)
            if a.dtype.kind in ("S", "U", "O"):
                return self._call_strings(a)
            t = torch.from_numpy(np.ascontiguousarray(a))
        if t.dtype in (torch.int8, torch.int16, torch.uint8):
            t = t.to(torch.int32)
        if t.dtype not in (torch.int32, torch.int64):
            raise NotImplementedError("Hash: integer or string input (tf.as_string of %s is not restated)" % t.dtype)
        t = t.to(self.device, non_blocking=True).contiguous()
        out = torch.empty(t.shape, dtype=torch.int64, device=self.device)
        _lib.call("ctrhot_hash_bucket_int", ptr(t), 1 if t.dtype == torch.int64 else 0, t.numel(), self.num_buckets,
                  1 if self.mask_zero else 0, ptr(out), stream_ptr())
        return out

    def _call_strings(self, a):
        flat = [v if isinstance(v, bytes) else str(v).encode() for v in a.reshape(-1)]
        n = len(flat)
        width = max([len(v) for v in flat] + [1])
        if width > 64:
            raise ValueError("Hash: strings over 64 bytes are not supported (got %d)" % width)
        buf = np.zeros((n, width), np.uint8)
        lens = np.zeros(n, np.int32)
        for i, v in enumerate(flat):
            buf[i, :len(v)] = np.frombuffer(v, np.uint8)
            lens[i] = len(v)
        d_buf, d_len = torch.from_numpy(buf).to(self.device), torch.from_numpy(lens).to(self.device)
        out = torch.empty(n, dtype=torch.int64, device=self.device)
        if self._err is None:
            self._err = torch.zeros(1, dtype=torch.int32, device=self.device)
        _lib.call("ctrhot_hash_bucket_bytes", ptr(d_buf), width, ptr(d_len), n, self.num_buckets, 1 if self.mask_zero else 0,
                  ptr(out), ptr(self._err), stream_ptr())
        return out.view(a.shape)
