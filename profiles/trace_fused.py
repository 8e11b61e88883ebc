"""Where does the MMA thread of the whole-model kernel wait?  Needs a TRACE=1 tuning build:

    python profiles/build_variants.py TRACE=1
    CTRHOT_LIB=recommender_system_b200/build/variants/libctrhot_TRACE1.so python profiles/trace_fused.py [cluster]
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from recommender_system_b200 import _lib  # noqa: E402
from recommender_system_b200.utils import synth  # noqa: E402

dev = torch.device("cuda", 0)
model, vocab = bench.build_deepfm(dev, "kaggle", 16, "tf32")
model.cluster = int(sys.argv[1]) if len(sys.argv) > 1 else 1
B = 65536
ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, vocab, seed=5 + i)) for i in range(3)]
for i in range(3):
    model(ins[i])
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_ulonglong * (256 * 8))()
assert lib.ctrhot_debug_fused_trace(buf) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.float64)
t = t[t[:, 7] > 0]
names = ["W(L1)", "raw", "lo", "h_ready", "last_free", "W(L2+)", "tiles", "total"]
print("CTAs", len(t))
for lab, sel in (("all CTAs mean", t.mean(0)), ("CTAs with most tiles, mean", t[t[:, 6] == t[:, 6].max()].mean(0))):
    print(lab)
    for n, v in zip(names, sel):
        print("   %-10s %10.0f %s" % (n, v, "" if n == "tiles" else "cyc (%.1f%% of total)" % (100 * v / sel[7])))

buf2 = (ctypes.c_ulonglong * (256 * 8))()
if hasattr(lib, "ctrhot_debug_fused_trace2") and lib.ctrhot_debug_fused_trace2(buf2) == 0:
    t2 = np.frombuffer(buf2, dtype=np.uint64).reshape(256, 8).astype(np.float64)
    t2 = t2[t2[:, 3] > 0]
    n2 = ["stager: wait full_raw", "stager: wait a_free", "stager: wait last_free", "stager total", "gather: wait empty_raw",
          "gather total", "epilogue: wait d_full", "epilogue total"]
    m = t2.mean(0)
    for i, n in enumerate(n2):
        tot = m[3] if i < 4 else m[5] if i < 6 else m[7]
        print("   %-24s %10.0f cyc (%.1f%% of its role)" % (n, m[i], 100 * m[i] / tot))
buf3 = (ctypes.c_ulonglong * (256 * 8))()
if hasattr(lib, "ctrhot_debug_fused_trace3") and lib.ctrhot_debug_fused_trace3(buf3) == 0:
    t3 = np.frombuffer(buf3, dtype=np.uint64).reshape(256, 8).astype(np.float64)
    t3 = t3[t3[:, 2] > 0]
    print("   stager per chunk: load + math %.0f cyc, TMEM store + hand-off %.0f cyc (%.0f chunks per CTA)" % (
        (t3[:, 0] / t3[:, 2]).mean(), (t3[:, 1] / t3[:, 2]).mean(), t3[:, 2].mean()))
    print("   gather per chunk: id fetch %.0f cyc, copy issue (after the stage is free) %.0f cyc" % (
        (t3[:, 3] / t3[:, 2]).mean(), (t3[:, 4] / t3[:, 2]).mean()))
tb = (ctypes.c_ulonglong * (256 * 4))()
if lib.ctrhot_debug_fused_times(tb) == 0:
    tt = np.frombuffer(tb, dtype=np.uint64).reshape(256, 4).astype(np.float64)
    tt = tt[tt[:, 3] > 0]
    print("CTA entry -> MMA loop begin %.1f us, MMA loop %.1f us, loop end -> exit %.1f us; first entry -> last exit %.1f us" % (
        (tt[:, 1] - tt[:, 0]).mean() / 1e3, (tt[:, 2] - tt[:, 1]).mean() / 1e3, (tt[:, 3] - tt[:, 2]).mean() / 1e3,
        (tt[:, 3].max() - tt[:, 0].min()) / 1e3))
if lib.ctrhot_debug_fused_times(tb) == 0:
    t0 = tt[:, 0].min()
    ent, lb, le, ex = (tt[:, i] - t0 for i in range(4))
    q = lambda a: "min %.1f  p10 %.1f  median %.1f  p90 %.1f  max %.1f us" % tuple(np.percentile(a, [0, 10, 50, 90, 100]) / 1e3)
    print("per-CTA timeline (globaltimer, relative to the first CTA's entry):")
    print("   entry        ", q(ent))
    print("   MMA loop end ", q(le))
    print("   exit         ", q(ex))
    print("   loop length  ", q(le - lb))
