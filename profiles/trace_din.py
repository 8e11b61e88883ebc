"""Pipeline timeline of the tcgen05 DIN attention kernel (CTA 0), needs a DT_TRACE tuning build:

    python profiles/build_variants.py src=din_tc.cu TRACE=1
    CTRHOT_LIB=recommender_system_b200/build/variants/libctrhot_din_tc_TRACE1.so python profiles/trace_din.py
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import _lib  # noqa: E402
from recommender_system_b200.layer import Attention, set_seed  # noqa: E402

dev = torch.device("cuda", 0)
B, T, K, V = 16384, 100, 64, 63002
set_seed(1)
L = Attention((80, 40), "prelu", device=dev)
L.build_for(T, K)
rng = np.random.default_rng(0)
table = torch.from_numpy((rng.random((V, K), dtype=np.float32) - 0.5) * 0.1).to(dev)
hist = torch.from_numpy(rng.integers(1, V, size=(B, T)).astype(np.int32)).to(dev)
cand = torch.from_numpy(rng.integers(1, V, size=(B, 1)).astype(np.int32)).to(dev)
for _ in range(3):
    L.call_gather(table, hist, cand, None)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (8 * 128))()
assert _lib.load().ctrhot_debug_din_trace(buf) == 0
t = np.frombuffer(buf, dtype=np.int64).reshape(8, 128).astype(np.float64)
names = ["gather issued", "stager sees rows", "A staged", "ep1 sees D1", "H1 ready", "ep2 sees D2", "sample done"]
n = 100
t0 = t[0, 0]
print("per-sample period (cycles), samples 10..%d: %.0f" % (n, (t[6, n] - t[6, 10]) / (n - 10)))
print("event times relative to 'gather issued' of the same sample, mean over samples 10..%d:" % n)
for e in range(1, 7):
    print("  %-18s +%7.0f cycles" % (names[e], np.mean(t[e, 10:n] - t[0, 10:n])))
print("stage spacing: mean cycles between consecutive events")
for e in range(1, 7):
    print("  %-18s -> %-18s %7.0f" % (names[e - 1], names[e], np.mean(t[e, 10:n] - t[e - 1, 10:n])))
# whole-launch time against (samples per CTA) x (period of CTA 0): how far the other CTAs / the start-up are from CTA 0's steady state
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    L.call_gather(table, hist, cand, None)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
per = (t[6, n] - t[6, 10]) / (n - 10)
n_my = (B + 147) // 148
print("call_gather (prep + attention kernels): %.1f us; %d samples per CTA x %.0f cycles = %.1f us at 1.965 GHz; CTA 0 first gather -> last sample done: %.0f cycles"
      % (ms * 1e3, n_my, per, n_my * per / 1965.0, t[6, min(n_my, 128) - 1] - t[0, 0]))
