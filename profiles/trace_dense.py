"""Where the MMA thread of the per-layer tower kernel (mlp_tc.cu) spends its time: python profiles/trace_dense.py
Layer [65536, 448] x [448, 400] (DCN cfg 3), counters from ctrhot_debug_tc_trace."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.layer.interaction import DNNLayer

dev = torch.device("cuda", 0)
M, K = 65536, int(sys.argv[1]) if len(sys.argv) > 1 else 429
hidden = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [400]
set_seed(1)
L = DNNLayer(hidden, 1, "relu", precision="tf32", device=dev)
L.build((None, K))
L.stack.allow_fused = False
xs = [torch.rand((M, (K + 31) // 32 * 32), device=dev)[:, :K] for _ in range(3)]
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
buf = torch.zeros(148 * 4, dtype=torch.int64, device=dev)
lib.ctrhot_debug_tc_trace(ctypes.c_void_p(buf.data_ptr()))
L(xs[0]); torch.cuda.synchronize()
lib.ctrhot_debug_tc_trace(ctypes.c_void_p(0))
t = buf.cpu().numpy().reshape(148, 4).astype(np.float64)
t = t[t[:, 3] > 0]
print("last layer launch: MMA thread cycles total %.0f (max %.0f) | waiting x/W stage %.0f (%.0f%%) | waiting free accumulator %.0f (%.0f%%) | steps %.1f"
      % (t[:, 2].mean(), t[:, 2].max(), t[:, 0].mean(), 100 * t[:, 0].sum() / t[:, 2].sum(), t[:, 1].mean(),
         100 * t[:, 1].sum() / t[:, 2].sum(), t[:, 3].mean()))
