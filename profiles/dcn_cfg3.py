"""DCN cfg 3 (L=3, 400-400-400 tower, B=65536) alone: CUDA-event time, for launch lists under ncu.
    python profiles/dcn_cfg3.py [reps]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import DCN
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
B, k = 65536, 16
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
vocab = synth.KAGGLE_VOCAB
fc = synth.criteo_feature_columns(vocab, k)
ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, vocab, seed=5 + i)) for i in range(3)]
set_seed(1)
m = DCN(fc, [400, 400, 400], 1, "relu", 3, embed_dim=k, check_ids=False, device=dev)
for i in range(3):
    m(ins[i % 3])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
l0 = _lib.LAUNCHES[0]
e0.record()
for i in range(reps):
    m(ins[i % 3])
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(json.dumps({"model": "DCN cfg3", "B": B, "ms": round(ms, 4), "samples_per_s": B / ms * 1e3,
                  "launches_per_step": (_lib.LAUNCHES[0] - l0) / reps,
                  "TFLOPs": (2 * (429 * 400 + 400 * 400 * 2 + 400) + 6 * 429) * B / ms / 1e9}))
