"""Per-kernel timing of the other BASELINE configs on one B200 (device-resident inputs, CUDA events).

    python profiles/kernels_bench.py [gather] [dcn] [din] [widedeep]

Prints one JSON line per measurement: samples/s, ms, algorithmic GB/s and fraction of the measured HBM peak.
bench.py stays the contract line (DeepFM cfg 2a); this is the map of everything else.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from recommender_system_b200 import _lib  # noqa: E402
from recommender_system_b200.layer import set_seed  # noqa: E402
from recommender_system_b200.layer.base import ptr, stream_ptr  # noqa: E402
from recommender_system_b200.utils import synth  # noqa: E402

dev = torch.device("cuda", 0)
PK = bench.peaks()


def timeit(fn, steps=30, warm=5):
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


def report(name, B, ms, algo_bytes=None, flops=None, **kw):
    d = {"kernel": name, "B": B, "ms": round(ms, 5), "samples_per_s": B / ms * 1e3}
    if algo_bytes:
        d["algo_GBs"] = algo_bytes * B / ms / 1e6
        d["hbm_frac"] = d["algo_GBs"] / PK["hbm_gbs"]
    if flops:
        d["TFLOPs"] = flops * B / ms / 1e9
    d.update(kw)
    print(json.dumps(d), flush=True)


def gather(k, vocab, B, tag):
    from recommender_system_b200.layer.core import EmbedLayer
    set_seed(1)
    E = EmbedLayer(synth.criteo_feature_columns(vocab, k)[1], k=k, check_ids=False, device=dev)
    ins = [torch.from_numpy(synth.criteo_batch(B, vocab, seed=7 + i)[1]).to(dev) for i in range(4)]
    it = [0]

    def run():
        E(ins[it[0] % 4])
        it[0] += 1
    ms = timeit(run)
    # ids + rows read + rows written (the standalone EmbedLayer must write its output)
    report("gather_f32 " + tag, B, ms, algo_bytes=26 * 4 + 2 * 26 * 4 * k, k=k, note="read+write of the rows counted")


def main(which):
    B = 65536
    if "gather" in which:
        gather(16, synth.KAGGLE_VOCAB, B, "k=16 kaggle (64 B rows)")
        gather(128, [1000000] * 26, B, "k=128 26 x 1M rows (512 B rows)")
    if "dcn" in which:
        from recommender_system_b200.model.criteo import DCN
        set_seed(1)
        m = DCN(synth.criteo_feature_columns(synth.KAGGLE_VOCAB, 16), [400, 400, 400], 1, "relu", 3, embed_dim=16,
                precision="tf32", check_ids=False, device=dev)
        m.build(None)
        ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, synth.KAGGLE_VOCAB, seed=3 + i)) for i in range(4)]
        it = [0]

        def run():
            m(ins[it[0] % 4]); it[0] += 1
        l0 = _lib.LAUNCHES[0]
        ms = timeit(run)
        report("DCN cfg3 (L=3, tower 400x3)", B, ms, algo_bytes=1824, flops=991000, launches_per_step=(_lib.LAUNCHES[0] - l0) / 35)
    if "widedeep" in which:
        from recommender_system_b200.model.criteo import WideDeep
        set_seed(1)
        m = WideDeep(synth.criteo_feature_columns(synth.KAGGLE_VOCAB, 16), [256, 128, 64], 1, "relu", embed_dim=16,
                     precision="tf32", check_ids=False, device=dev)
        m.build(None)
        ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, synth.KAGGLE_VOCAB, seed=3 + i)) for i in range(4)]
        it = [0]

        def run():
            m(ins[it[0] % 4]); it[0] += 1
        ms = timeit(run)
        report("WideDeep cfg2b", B, ms, algo_bytes=1928, flops=295000)
    if "din" in which:
        from recommender_system_b200.model.din import DIN
        Bd, T = 16384, 100
        for v_item, tag in ((63002, "item vocab 63k (L2-resident)"), (10_000_000, "item vocab 10M (HBM)")):
            set_seed(1)
            m = DIN(synth.din_feature_columns(v_item=v_item), ["movies_seq"], (80, 40), (256, 128, 64), "prelu", "prelu",
                    0.0, precision="tf32", check_ids=False, device=dev)
            ins = []
            for i in range(3):
                d = synth.din_batch(Bd, T, v_item=v_item, seed=11 + i)
                ins.append({k: torch.from_numpy(v).to(dev) for k, v in d.items()})
            m(ins[0])
            it = [0]

            def run():
                m(ins[it[0] % 3]); it[0] += 1
            ms = timeit(run, steps=10, warm=3)
            report("DIN cfg4 " + tag, Bd, ms, algo_bytes=26528, flops=4950000)


if __name__ == "__main__":
    main(sys.argv[1:] or ["gather", "dcn", "din", "widedeep"])
