"""DeepFM SGD step at B = 32 / 4096 / 32768 (bench.py's deepfm_train leg at more sizes): ms per step, call by call."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import DeepFM
from recommender_system_b200.utils import synth
from recommender_system_b200.utils.compile_fit import _DeepFMTrainer
dev = torch.device("cuda:0")


def timeit(fn, steps, warm):
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


vocab = [1000] * 26
out = {}
for B in [int(a) for a in sys.argv[1:]] or (32, 4096, 32768):
    set_seed(5)
    m = DeepFM(synth.criteo_feature_columns(vocab, 8), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=8, precision="fp32", check_ids=False, device=dev)
    m.build(None)
    dense, ids = synth.criteo_batch(B, vocab, seed=3)
    y = (np.random.default_rng(4).random(B) < 0.3).astype(np.float32)
    d, i, yy = (torch.from_numpy(a).to(dev) for a in (dense, ids, y))
    tr = _DeepFMTrainer(m, dropout=0.2)
    ms = timeit(lambda: tr.train_on_batch(d, i, yy, 0.01), 30, 5)
    out["B%d" % B] = {"ms": ms, "samples_per_s": B / ms * 1e3}
print(json.dumps(out))
