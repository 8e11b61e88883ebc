"""Upper bound for dropping the FM side MMAs: the whole-model kernel on the cfg-2 shape with and without side products.
    python profiles/time_tower_mode.py     (with a TRACE=1 tuning build in CTRHOT_LIB it also prints the MMA thread's waits)"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from recommender_system_b200 import _lib
from recommender_system_b200.layer import set_seed
from recommender_system_b200.layer.base import ids_arg
from recommender_system_b200.model.criteo import WideDeep
from recommender_system_b200.utils import synth

dev = torch.device("cuda", 0)
B, k = 65536, 16
model, vocab = bench.build_deepfm(dev, "kaggle", k, "tf32")
ins = [tuple(torch.from_numpy(a).to(dev) for a in synth.criteo_batch(B, vocab, seed=5 + i)) for i in range(4)]
set_seed(1)
wd = WideDeep(synth.criteo_feature_columns(vocab, k), [256, 128, 64], 1, "relu", embed_dim=k, check_ids=False, device=dev)
wd.build(None)
wd.one_launch = False          # this script times the bare tower launch (and must run on A/B builds without the fused entry)
wd(ins[0])
E = wd.embed_layer
stack = wd.deep.stack
out = torch.empty((B, 1), dtype=torch.float32, device=dev)
zeros = torch.zeros((B,), dtype=torch.float32, device=dev)


def tower(i):
    ids2, kind, ids_ld = ids_arg(ins[i % 4][1])
    stack.run_ctr_tower(E, ids2, kind, ids_ld, None, B, wd.deep.hidden[0][0], add_vec=zeros, scale=0.5, sigmoid=True,
                        err_flag=E.err_flag(), out=out)


def t(fn, reps=300, warm=150):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def dump(tag):
    try:
        lib = ctypes.CDLL(_lib.LIB_PATH)
        buf = (ctypes.c_ulonglong * (256 * 8))()
        if lib.ctrhot_debug_fused_trace(buf) != 0:
            return
    except AttributeError:
        return
    a = np.frombuffer(buf, dtype=np.uint64).reshape(256, 8).astype(np.float64)
    a = a[a[:, 7] > 0]
    a = a[a[:, 6] == a[:, 6].max()].mean(0)
    names = ["W(L1)", "A staged", "-", "h_ready", "last_free", "W(L2+)", "tiles", "total"]
    print(tag, " ".join("%s=%.0f" % (n, v) for n, v in zip(names, a)))


print("DeepFM whole-model kernel (side MMAs): %.1f us" % t(lambda i: model(ins[i % 4])))
dump("  trace:")
print("tower mode, 416 -> 256 -> 128 -> 64 -> 1 from the arena (no side MMAs): %.1f us" % t(tower))
dump("  trace:")
