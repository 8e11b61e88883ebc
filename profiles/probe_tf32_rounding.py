"""Probe: does tcgen05.mma kind::tf32 truncate or round the low 13 mantissa bits of an fp32 operand it reads from
shared memory?  (The split-TF32 side products of fused.cu mask the operand themselves, so they do not depend on the
answer; this records what the hardware does.)  Run on the GPU box: python profiles/probe_tf32_rounding.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recommender_system_b200.layer import DNNLayer  # noqa: E402

K = 32
x = np.zeros((128, K), np.float32)
x[:, 0] = 1.0 + 2.0 ** -11 + 2.0 ** -12            # needs rounding to fit 10 explicit mantissa bits
W1 = np.zeros((K, 8), np.float32); W1[0, 0] = 1.0
Wo = np.zeros((8, 1), np.float32); Wo[0, 0] = 1.0
L = DNNLayer([8], 1, "relu", precision="tf32")
L.build((None, K))
L.set_weights([W1, np.zeros(8, np.float32), Wo, np.zeros(1, np.float32)])
L.stack.allow_fused = True                          # persistent kernel: A operand = raw fp32 bits via cp.async
y = float(L(x).cpu().numpy()[0, 0])
print("operand 1+2^-11+2^-12 ->", repr(y), "=> hardware", "TRUNCATES" if y == 1.0 else
      ("ROUNDS-TO-NEAREST" if y == 1.0 + 2.0 ** -10 else "other"))
