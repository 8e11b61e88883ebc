"""Stress of the chained-launch paths (per-tile flags + overlapping launches): many back-to-back forwards over rotating batches,
every result compared bit for bit with plain stream order.  PNN (front -> tower) and DCN cfg 3 (three launches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import DCN, PNN
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
vocab = [5000] * 26
bad = 0
for name, make, ks in (("PNN", lambda: PNN(synth.criteo_feature_columns(vocab, 8), "both", [256, 128, 64], 1, embed_dim=8, check_ids=False, device=dev), 8),
                       ("DCN", lambda: DCN(synth.criteo_feature_columns(vocab, 16), [400, 400, 400], 1, "relu", 3, embed_dim=16, check_ids=False, device=dev), 16)):
    set_seed(1)
    m = make()
    m.build(None)
    for B in (65536, 148 * 128 + 1, 40000):
        ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, vocab, seed=50 + i)) for i in range(4)]
        if name == "PNN":
            m.chain_launches = False
        else:
            m.dense_layer.stack.chain_launches = False
        want = [m(x).clone() for x in ins]
        torch.cuda.synchronize()
        if name == "PNN":
            m.chain_launches = True
        else:
            m.dense_layer.stack.chain_launches = True
        outs = []
        for i in range(200):
            outs.append(m(ins[i % 4]))
            if i % 25 == 24:                      # keep only a window alive, compare in bulk
                torch.cuda.synchronize()
                for j, o in enumerate(outs):
                    if not torch.equal(o, want[(i - len(outs) + 1 + j) % 4]):
                        bad += 1
                outs = []
        print(name, "B", B, "200 chained forwards, mismatches so far:", bad, flush=True)
print("STRESS", "OK" if bad == 0 else "FAILED")
