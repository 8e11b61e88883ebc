"""CUDA-event times of the secondary models at BASELINE's shapes (device-resident inputs):
WideDeep cfg 2b, DCN cfg 3 (400-wide tower), PNN, FFM, FM one-hot -- python profiles/time_models.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import DCN, NFM, PNN, WideDeep
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
B, k = 65536, 16
vocab = synth.KAGGLE_VOCAB
fc = synth.criteo_feature_columns(vocab, k)
batches = [synth.criteo_batch(B, vocab, seed=5 + i) for i in range(3)]
ins = [(torch.from_numpy(d).to(dev), torch.from_numpy(i).to(dev)) for d, i in batches]


def t(fn, reps=20, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = _lib.LAUNCHES[0]
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, (_lib.LAUNCHES[0] - l0) / reps


def line(name, ms, launches, flops):
    print(json.dumps({"model": name, "B": B, "ms": round(ms, 4), "samples_per_s": B / ms * 1e3, "launches_per_step": launches,
                      "TFLOPs": flops * B / ms / 1e9}), flush=True)


set_seed(1)
m = WideDeep(fc, [256, 128, 64], 1, "relu", embed_dim=k, check_ids=False, device=dev)
ms, l = t(lambda i: m(ins[i % 3]))
line("WideDeep cfg2b", ms, l, 2 * (416 * 256 + 256 * 128 + 128 * 64 + 64))
del m
set_seed(1)
m = DCN(fc, [400, 400, 400], 1, "relu", 3, embed_dim=k, check_ids=False, device=dev)
ms, l = t(lambda i: m(ins[i % 3]))
line("DCN cfg3 (L=3, tower 400x3)", ms, l, 2 * (429 * 400 + 400 * 400 * 2 + 400) + 6 * 429)
del m
set_seed(1)
m = DCN(fc, [256, 128, 64], 1, "relu", 3, embed_dim=k, check_ids=False, device=dev)
ms, l = t(lambda i: m(ins[i % 3]))
line("DCN (L=3, tower 256-128-64: whole-model kernel)", ms, l, 2 * (429 * 256 + 256 * 128 + 128 * 64 + 64) + 6 * 429)
del m
del m
set_seed(1)
fc8 = synth.criteo_feature_columns(vocab, 8)
m = NFM(fc8, [256, 128, 64], 1, "relu", embed_dim=8, check_ids=False, device=dev)
m.build(None)
ms, l = t(lambda i: m(ins[i % 3]))
line("NFM (k=8, Bi-Interaction + tower 256-128-64; SURVEY 8(f) rank 1)", ms, l, 2 * (21 * 256 + 256 * 128 + 128 * 64 + 64) + 26 * 8 * 3)
