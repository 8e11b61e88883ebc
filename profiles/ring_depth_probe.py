"""Does a deeper row ring help once the weight ring does not have to shrink for it?  DeepFM with a 128-wide first layer (weight
stage 20 KB instead of 36 KB, so NR = 8 rows stages fit beside NB = 4 weight stages) timed on tuning builds:
    python profiles/build_variants.py NR=4,NB=4 NR=8,NB=4 NR=12,NB=4
    for v in NR4_NB4 NR8_NB4 NR12_NB4; do CTRHOT_LIB=recommender_system_b200/build/variants/libctrhot_$v.so python profiles/ring_depth_probe.py; done
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recommender_system_b200.layer import set_seed  # noqa: E402
from recommender_system_b200.model.criteo import DeepFM  # noqa: E402
from recommender_system_b200.utils import synth  # noqa: E402

dev = torch.device("cuda", 0)
B, k = 65536, 16
vocab = synth.KAGGLE_VOCAB
for hidden in ([128, 128, 64], [256, 128, 64]):
    set_seed(1)
    m = DeepFM(synth.criteo_feature_columns(vocab, k), 10, 1e-4, 1e-4, hidden, 1, "relu", embed_dim=k, check_ids=False, device=dev)
    m.build(None)
    ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, vocab, seed=5 + i)) for i in range(4)]
    for i in range(5):
        m(ins[i % 4])
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(50):
            m(ins[i % 4])
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 50)
    print("%-40s hidden %-16s %.2f us/step" % (os.path.basename(os.environ.get("CTRHOT_LIB", "libctrhot.so")), hidden, best * 1e3))
    del m
