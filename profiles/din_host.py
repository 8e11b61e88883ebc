"""Host time vs device time of one DIN forward (cfg 4): python profiles/din_host.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.din import DIN
from recommender_system_b200.utils import synth
dev = torch.device("cuda", 0)
Bd, T, v_item = 16384, 100, 10_000_000
set_seed(1)
m = DIN(synth.din_feature_columns(v_item=v_item), ["movies_seq"], (80, 40), (256, 128, 64), "prelu", "prelu", 0.0,
        precision="tf32", check_ids=False, device=dev)
ins = [{k: torch.from_numpy(v).to(dev) for k, v in synth.din_batch(Bd, T, v_item=v_item, seed=11 + i).items()} for i in range(3)]
for i in range(20):
    m(ins[i % 3])
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(100):
    m(ins[i % 3])
t_host = (time.perf_counter() - t0) / 100
torch.cuda.synchronize()
t_all = (time.perf_counter() - t0) / 100
print("DIN forward: host issue time %.0f us per call, wall incl. device %.0f us per call" % (t_host * 1e6, t_all * 1e6))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for i in range(50):
    m(ins[i % 3])
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
