"""Pipeline timeline of the one-launch DIN front end (CTA 0), needs a DF_TRACE tuning build:

    python profiles/build_variants.py src=din_fwd.cu D:DF_TRACE
    CTRHOT_LIB=recommender_system_b200/build/variants/libctrhot_din_fwd_DF_TRACE.so python profiles/trace_din_fwd.py
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
B, T, K, V = 16384, 100, 64, 10_000_000
FULL = "full" in sys.argv[1:]                                # full histories instead of len ~ U{0..T}
if "small" in sys.argv[1:]:
    V = 63002                                                  # a 16 MB table: L2-resident, few pages          # full histories instead of len ~ U{0..T}
set_seed(1)
L = Attention((80, 40), "prelu", device=dev)
L.build_for(T, K)
rng = np.random.default_rng(0)
table = (torch.rand((V, K), device=dev) - 0.5) * 0.1
hist = rng.integers(1, V, size=(B, T)).astype(np.int32)
if not FULL:
    lens = rng.integers(0, T + 1, size=B)
    hist[np.arange(T)[None, :] >= lens[:, None]] = 0
hist = torch.from_numpy(hist).to(dev)
cand = torch.from_numpy(rng.integers(1, V, size=(B, 1)).astype(np.int32)).to(dev)
emb = torch.empty((B, 196), dtype=torch.float32, device=dev)
f = lambda: L.front(table, hist, cand, emb, (), None, 2 * K, None)
for _ in range(3):
    f()
torch.cuda.synchronize()
lib = _lib.load()
lib.ctrhot_debug_din_fwd_trace.restype = ctypes.c_int
lib.ctrhot_debug_din_fwd_trace.argtypes = [ctypes.c_void_p]
buf = (ctypes.c_longlong * (14 * 128))()
assert lib.ctrhot_debug_din_fwd_trace(buf) == 0
t = np.frombuffer(buf, dtype=np.int64).reshape(14, 128).astype(np.float64)
names = ["gather issued", "builder sees q", "Ws built", "ep1 sees D1", "H1 ready", "ep2 sees D2", "P ready", "att written"]
t7 = t[7, :].copy(); t[7, :] = t[9, :]                # event 7 of the table below = mark 9 (att written by the gather warps)
n = 100
print("history: %s" % ("full" if FULL else "len ~ U{0..T}"))
print("per-sample period (cycles), samples 10..%d: %.0f" % (n, (t[7, n] - t[7, 10]) / (n - 10)))
print("event times relative to 'gather issued' of the same sample, mean over samples 10..%d:" % n)
for e in range(1, 8):
    print("  %-18s +%7.0f cycles" % (names[e], np.mean(t[e, 10:n] - t[0, 10:n])))
print("stage spacing: mean cycles between consecutive events")
for e in range(1, 8):
    print("  %-18s -> %-18s %7.0f" % (names[e - 1], names[e], np.mean(t[e, 10:n] - t[e - 1, 10:n])))
# the gather warps' iteration i: [wait P(i-4)] wsum(i-4) -> att(i-4) written -> ids(i) published -> rows of sample i issued
g = slice(14, 100)
gi = np.arange(14, 100)
print("gather warps, per iteration: rows issued(i-1) -> P(i-4) seen %.0f | weighted sum %.0f | reduce + store %.0f | ids published %.0f | row copies issued %.0f"
      % (np.mean(t[11, gi - 4] - t[0, gi - 1]), np.mean(t[12, gi - 4] - t[11, gi - 4]), np.mean(t[9, gi - 4] - t[12, gi - 4]),
         np.mean(t[10, gi] - t[9, gi - 4]), np.mean(t[0, gi] - t[10, gi])))
print("kernel start -> prologue done %.0f cycles; prologue done -> first gather issued %.0f; last sample done -> CTA exit %.0f"
      % (t[8, 0] - t7[0], t[0, 0] - t[8, 0], t[13, 0] - t[7, min((B + 147) // 148, 128) - 1]))
print("first gather issued -> first sample done: %.0f cycles;  prologue (kernel start is not marked): see whole-launch time"
      % (t[7, 0] - t[0, 0]))
lib.ctrhot_debug_din_fwd_cta.restype = ctypes.c_int
lib.ctrhot_debug_din_fwd_cta.argtypes = [ctypes.c_void_p]
cb = (ctypes.c_ulonglong * (148 * 4))()
assert lib.ctrhot_debug_din_fwd_cta(cb) == 0
c = np.frombuffer(cb, dtype=np.uint64).reshape(148, 4).astype(np.float64)
t00 = c[:, 0].min()
print("per CTA (us from the first CTA's start): start min/max %.1f/%.1f, prologue end min/med/max %.1f/%.1f/%.1f, exit min/med/max %.1f/%.1f/%.1f"
      % ((c[:, 0].min() - t00) / 1e3, (c[:, 0].max() - t00) / 1e3, (c[:, 1].min() - t00) / 1e3, (np.median(c[:, 1]) - t00) / 1e3,
         (c[:, 1].max() - t00) / 1e3, (c[:, 3].min() - t00) / 1e3, (np.median(c[:, 3]) - t00) / 1e3, (c[:, 3].max() - t00) / 1e3))
print("  exit time by CTA (us): " + " ".join("%.0f" % ((c[k, 3] - t00) / 1e3) for k in range(148)))
print("  CTA 0: prologue end %.1f, exit %.1f; slowest CTAs: %s" % ((c[0, 1] - t00) / 1e3, (c[0, 3] - t00) / 1e3,
      ", ".join("%d:%.0f" % (k, (c[k, 3] - t00) / 1e3) for k in np.argsort(-c[:, 3])[:8])))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    f()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
per = (t[7, n] - t[7, 10]) / (n - 10)
n_my = (B + 147) // 148
print("launch: %.1f us; %d samples per CTA x %.0f cycles = %.1f us at 1.965 GHz; CTA 0 first gather -> last sample done: %.0f cycles = %.1f us"
      % (ms * 1e3, n_my, per, n_my * per / 1965.0, t[7, min(n_my, 128) - 1] - t[0, 0], (t[7, min(n_my, 128) - 1] - t[0, 0]) / 1965.0))
