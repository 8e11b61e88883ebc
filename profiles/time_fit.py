"""compile_fit at the reference's batch of 32, staged graph replay against call by call: steps/s over whole epochs (host loop,
copies into the staging buffers and epoch bookkeeping included), for the trainers whose step has no dropout."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import AFM, FFM
from recommender_system_b200.model.deepCrossing import DeepCrossing
from recommender_system_b200.model.din import DIN
from recommender_system_b200.utils import synth
from recommender_system_b200.utils.compile_fit import compile_fit
vocab = [1000] * 26
fc = synth.criteo_feature_columns(vocab, 8)
N = 32 * 300
dense, ids = synth.criteo_batch(N, vocab, seed=3)
X = np.concatenate([dense, ids.astype(np.float32)], axis=1)
y = (np.random.default_rng(4).random(N) < 0.3).astype(np.float32)
dd = synth.din_batch(N, 10, seed=6)
out = {}


def run(name, make, data):
    res = {}
    for g in (False, True):
        set_seed(5)
        m = make()
        compile_fit(m, data, y, batch_size=32, epochs=1, sgd=0.01, verbose=False, use_graph=g)      # warm: build, capture cost
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        compile_fit(m, data, y, batch_size=32, epochs=3, sgd=0.01, verbose=False, use_graph=g)
        torch.cuda.synchronize()
        res["graph" if g else "calls"] = 3 * (N // 32) / (time.perf_counter() - t0)
    out[name] = res


run("deepcrossing_3res", lambda: DeepCrossing(fc, 32, [256, 128], 3, embed_dim=8, precision="fp32"), X)
run("afm_att", lambda: AFM(fc, "att"), X)
run("ffm_k8", lambda: FFM(fc, 8, 1e-4, 1e-4), X)
run("din_T10", lambda: DIN(synth.din_feature_columns(embed_dim=8), ["movies_seq"], (80, 40), (256, 128, 64), "prelu", "prelu", 0.0,
                           precision="fp32"), dd)
print(json.dumps(out))
