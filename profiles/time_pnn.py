"""CUDA-event times of PNN (DL/model/pnn.py, the reference's embed_dim = 8) at the Criteo shape, B = 65 536, kaggle vocabulary:
the one-launch front (csrc/pnn_fwd.cu) vs gather + InnerProductLayer + OuterProductLayer as separate launches, and the whole
model either way -- python profiles/time_pnn.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import PNN
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
B, k, F = 65536, 8, 26
vocab = synth.KAGGLE_VOCAB
fc = synth.criteo_feature_columns(vocab, k)
ins = [torch.from_numpy(synth.criteo_batch(B, vocab, seed=5 + i)[1]).to(dev) for i in range(3)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


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


P = F * (F - 1) // 2
for mode in ("both", "outer"):
    set_seed(1)
    m = PNN(fc, mode, [256, 128, 64], 1, embed_dim=k, check_ids=False, device=dev)
    m.build(None)
    E, OP, IP = m.embed_layer, m.outer_product_layer, m.inner_product_layer
    width = F * k + P * (2 if mode == "both" else 1)
    ld = (width + 31) // 32 * 32
    buf = torch.empty((B, ld), dtype=torch.float32, device=dev)

    def fused(i):
        OP.front(E, ins[i % 3], mode == "both", buf, ld)

    def separate(i):
        E.gather(ins[i % 3], out=buf, out_ld=ld)
        z = buf.as_strided((B, F, k), (ld, k, 1))
        col = F * k
        if mode == "both":
            IP(z, out=buf[:, col:col + P]); col += P
        OP(z, out=buf[:, col:col + P])

    x = [torch.cat([torch.zeros((B, 13), device=dev), ins[i].float()], dim=1) for i in range(3)]
    for name, fn in (("front, one launch", fused), ("front, separate launches", separate)):
        ms, l = t(fn)
        algo = B * (F * 4 + F * k * 4 + width * 4)            # ids + rows read, product vector written
        print(json.dumps({"what": name, "mode": mode, "B": B, "ms": round(ms, 4), "launches": l, "samples_per_s": B / ms * 1e3,
                          "algo_GBs": algo / ms / 1e6, "outer_TFLOPs_fp32_equiv": 2 * P * k * k * B / ms / 1e9}), flush=True)
    for ff in (True, False):
        m.fused_front = ff
        ms, l = t(lambda i: m(x[i % 3]))
        print(json.dumps({"what": "PNN model, fused_front=%s" % ff, "mode": mode, "B": B, "ms": round(ms, 4), "launches": l,
                          "samples_per_s": B / ms * 1e3}), flush=True)
    del m
