"""Attention-output error of din_fwd_kernel against the float64 oracle at the BASELINE config-4 shape, for the library in
CTRHOT_LIB (default build: H1 rounded to nearest TF32; -DDF_ROUND_H1=0: left to the MMA's truncation)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from test_gpu_din_fwd import front
from test_gpu_din_tc import make_case

for scale in (1.0, 3.0):
    errs = []
    for seed in range(4):
        T, K, hidden, B, V = 100, 64, (80, 40), 2048, 63002
        table, hist, cand, ws, want, q = make_case(B, T, K, hidden, V, 100 + seed, scale)
        got, flag, L = front(table, hist, cand, ws, T, K, hidden)
        e = np.abs(got[:, :K] - want)
        errs.append((e.max(), e.mean(), np.abs(want).max()))
    print(os.environ.get("CTRHOT_LIB", "default"), "scale", scale, "max abs err %.3e  mean abs err %.3e  (max |att| %.2f)" %
          (max(x[0] for x in errs), np.mean([x[1] for x in errs]), max(x[2] for x in errs)))
