"""DIN cfg 4 forward (B = 16384, T = 100, K = 64, attention 80-40, tower 256-128-64, item vocabulary 10 M): ms per step and the
split between the front-end launch and the tower, CUDA events.   python profiles/time_din.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200 import _lib  # noqa: E402
from recommender_system_b200.layer import set_seed  # noqa: E402
from recommender_system_b200.model.din import DIN  # noqa: E402
from recommender_system_b200.utils import synth  # noqa: E402

dev = torch.device("cuda", 0)
set_seed(1)
Bd, T, v_item = 16384, 100, 10_000_000
m = DIN(synth.din_feature_columns(v_item=v_item), ["movies_seq"], (80, 40), (256, 128, 64), "prelu", "prelu", 0.0,
        precision="tf32", check_ids=False, device=dev)
ins = []
for i in range(3):
    d = synth.din_batch(Bd, T, v_item=v_item, seed=11 + i)
    ins.append({k: torch.from_numpy(v).to(dev) for k, v in d.items()})
m(ins[0])


def timeit(fn, steps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


it = [0]


def run():
    m(ins[it[0] % 3]); it[0] += 1


l0 = _lib.LAUNCHES[0]
ms = timeit(run)
print("DIN cfg4 whole forward: %.1f us/step = %.1f M samples/s, %.2f C-ABI launches per step"
      % (ms * 1e3, Bd / ms / 1e3, (_lib.LAUNCHES[0] - l0) / 23.0))
# the front end alone
A = m.att_layer
K = m.K
emb = torch.empty((Bd, (m.dcat + 3) // 4 * 4), dtype=torch.float32, device=dev)
x = ins[0]
if A.front_fits(T, K):
    ex = [(m.embed_sparse_layers[0], x["user_id"], 2 * K)]
    f = lambda: A.front(m.embed_seq_layers[0], x["movies_seq"], x["movie_id"], emb, ex, x["age"], 2 * K + 64, None)
    print("front end (ctrhot_din_fwd_tf32): %.1f us" % (timeit(f) * 1e3))
f2 = lambda: A.call_gather(m.embed_seq_layers[0], x["movies_seq"], x["movie_id"], None)
print("round-1 unit (din_prep + din_tc): %.1f us" % (timeit(f2) * 1e3))
m.stack.set_layers(m._tower_layers(), m._version)
f3 = lambda: m.stack.run(emb[:, :m.dcat], final_sigmoid=True)
print("tower: %.1f us" % (timeit(f3) * 1e3))
