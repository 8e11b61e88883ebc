"""Per-kernel timing of the interaction layers at B = 65 536 (device-resident inputs, CUDA events, 200 iterations after 50
warm-ups): algorithmic GB/s (one pass over the layer's input + output, gathered rows for the one-hot layers) against the
measured HBM peak.   python profiles/interaction_bench.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from recommender_system_b200.layer import set_seed
from recommender_system_b200.layer.interaction import (CrossLayer, Dice, FFMLayer, FMLayer, InnerProductLayer,
                                                      OuterProductLayer)
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
PK = bench.peaks()
B, F, k = 65536, 26, 16


def t(fn, reps=200, warm=50):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def line(name, ms, bytes_per_sample, note=""):
    gbs = bytes_per_sample * B / ms / 1e6
    print(json.dumps({"kernel": name, "B": B, "ms": round(ms, 4), "algo_bytes_per_sample": bytes_per_sample,
                      "algo_GBs": round(gbs, 1), "hbm_frac": round(gbs / PK["hbm_gbs"], 3), "note": note}), flush=True)


set_seed(1)
g = torch.Generator(device=dev); g.manual_seed(0)
es = [torch.rand((B, F, k), device=dev, generator=g) - 0.5 for _ in range(3)]
it = [0]


def rot(xs):
    it[0] += 1
    return xs[it[0] % len(xs)]


ip = InnerProductLayer(device=dev)
out_p = torch.empty((B, 325), dtype=torch.float32, device=dev)
line("pnn_inner_f32 [B,26,16] -> [B,325]", t(lambda: ip(rot(es), out=out_p)), F * k * 4 + 325 * 4)
op = OuterProductLayer(device=dev)
op.build(es[0].shape)
line("pnn_outer_f32 [B,26,16] -> [B,325]", t(lambda: op(rot(es), out=out_p)), F * k * 4 + 325 * 4, "W [16,325,16] = 333 KB in L2")
# the reference's own PNN shape (embed_dim 8): OuterProductLayer on the tensor pipe (csrc/pnn_fwd.cu) vs the CUDA-core kernel
e8 = [torch.rand((B, F, 8), device=dev, generator=g) - 0.5 for _ in range(3)]
op8 = OuterProductLayer(device=dev)
op8.build(e8[0].shape)
line("pnn_outer_f32 [B,26,8] -> [B,325], tcgen05 3xTF32", t(lambda: op8(rot(e8), out=out_p)), F * 8 * 4 + 325 * 4, "W [8,325,8] = 83 KB")
op8.tensor_pipe = False
line("pnn_outer_f32 [B,26,8] -> [B,325], CUDA cores", t(lambda: op8(rot(e8), out=out_p)), F * 8 * 4 + 325 * 4, "")
xs = [torch.rand((B, 448), device=dev, generator=g)[:, :429] for _ in range(3)]
cl = CrossLayer(3, device=dev)
cl.build((None, 429))
line("cross_f32 L=3 [B,429] -> [B,429]", t(lambda: cl(rot(xs))), 2 * 429 * 4)
fm = FMLayer(10, device=dev)
fm.build((None, 429))
line("fm_dense_f32 [B,429] -> [B,1] (k_fm=10)", t(lambda: fm(rot(xs))), 429 * 4 + 4)
hs = [torch.rand((B, 256), device=dev, generator=g) - 0.5 for _ in range(3)]
di = Dice(device=dev)
di.build((None, 256))
line("dice_f32 [B,256]", t(lambda: di(rot(hs))), 2 * 256 * 4)
# one-hot layers on the bundled sample's vocabulary (43 605 rows): FFM v = [n_feat, 39, 8]
vocab = synth.SAMPLE_VOCAB
fc = synth.criteo_feature_columns(vocab, 8)
ms_in = [torch.from_numpy(synth.criteo_matrix(*synth.criteo_batch(B, vocab, seed=3 + i)).astype(np.float32)).to(dev) for i in range(3)]
ffm = FFMLayer(fc, 8, device=dev)
ffm.build(None)
line("ffm_onehot_f32 (39 fields, k=8, sample vocabulary)", t(lambda: ffm(rot(ms_in)), reps=50, warm=10),
     39 * 4 + 26 * 39 * 8 * 4 + 13 * 39 * 8 * 4 + 4, "26 gathered + 13 dense rows of 1 248 B; v = 54 MB (L2-resident)")
