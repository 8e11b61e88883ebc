"""MMA-thread wait causes of the persistent kernel on a wide single layer (DCN cfg 3's 400-wide tower, rows mode).
    python profiles/build_variants.py TRACE=1
    CTRHOT_LIB=recommender_system_b200/build/variants/libctrhot_TRACE1.so python profiles/trace_wide.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.layer.interaction import DNNLayer

dev = torch.device("cuda", 0)
M, K = 65536, 429
hidden = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [400, 400, 400]
set_seed(1)
L = DNNLayer(hidden, 1, "relu", precision="tf32", device=dev)
L.build((None, K))
xs = [torch.rand((M, 448), device=dev)[:, :K] for _ in range(3)]
for x in xs:
    L(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(20):
    L(xs[i % 3])
e1.record(); torch.cuda.synchronize()
print("K=%d hidden=%s: %.1f us per forward" % (K, hidden, e0.elapsed_time(e1) / 20 * 1e3))
lib = ctypes.CDLL(_lib.LIB_PATH)
buf = (ctypes.c_ulonglong * (256 * 8))()
assert lib.ctrhot_debug_fused_trace(buf) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.float64)
t = t[t[:, 7] > 0]
names = ["W(L1)", "A staged", "-", "h_ready", "last_free", "W(L2+)", "tiles", "total"]
print("CTAs", len(t))
for lab, sel in (("all CTAs mean", t.mean(0)), ("CTAs with most tiles, mean", t[t[:, 6] == t[:, 6].max()].mean(0))):
    print(lab)
    for n, v in zip(names, sel):
        print("   %-10s %10.0f %s" % (n, v, "" if n == "tiles" else "cyc (%.1f%% of total)" % (100 * v / sel[7])))

tb = (ctypes.c_ulonglong * (256 * 4))()
if lib.ctrhot_debug_fused_times(tb) == 0:
    tt = np.frombuffer(tb, dtype=np.uint64).reshape(256, 4).astype(np.float64)[:len(t)]
    t0 = tt[:, 0].min()
    print("globaltimer (us after the first CTA's entry): entry %.1f..%.1f | MMA loop begin %.1f..%.1f | MMA loop end %.1f..%.1f | exit %.1f..%.1f"
          % tuple(v / 1e3 for c in range(4) for v in ((tt[:, c] - t0).min(), (tt[:, c] - t0).max())))
    end = (tt[:, 2] - t0) / 1e3
    print("MMA loop end per cluster of 4 (us):", " ".join("%.0f" % end[i:i + 4].max() for i in range(0, len(end), 4)))
    print("tiles per CTA:", " ".join("%d" % v for v in t[:, 6]))
    print("wait A staged per CTA (kcyc):", " ".join("%.0f" % (v / 1e3) for v in t[:, 1]))
    print("wait W per CTA (kcyc):", " ".join("%.0f" % (v / 1e3) for v in t[:, 0]))
    print("wait last_free per CTA (kcyc):", " ".join("%.0f" % (v / 1e3) for v in t[:, 4]))

# launch-overhead-free timing: the same forward captured in a CUDA graph
try:
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        L(xs[0])
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(g):
        y = L(xs[0])
    g.replay(); torch.cuda.synchronize()
    e0.record()
    for i in range(20):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    print("CUDA graph replay: %.1f us per forward" % (e0.elapsed_time(e1) / 20 * 1e3))
except Exception as ex:  # noqa: BLE001
    print("graph capture failed:", ex)
