"""DeepFM cfg 2a (B = 65 536, kaggle vocabulary, k = 16): step time per cluster size with and without overlapped launches.
python profiles/cluster_overlap_sweep.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from recommender_system_b200.utils import synth  # noqa: E402

dev = torch.device("cuda", 0)
m, vocab = bench.build_deepfm(dev, "kaggle", 16, "tf32")
B = 65536
ins = []
for i in range(4):
    d, ids = synth.criteo_batch(B, vocab, seed=100 + i)
    ins.append((torch.from_numpy(d).to(dev), torch.from_numpy(ids).to(dev)))
m.check = False
for cl in (0, 1, 2, 4):
    for ov in (False, True):
        m.cluster, m.overlap_launches = cl, ov
        for i in range(5):
            m(ins[i % 4])
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(50):
                m(ins[i % 4])
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 50)
        print("cluster %d overlap %d: %.2f us/step = %.0f M samples/s" % (cl, ov, best * 1e3, B / best / 1e3))
