"""Times the dense side of config 5 (FMLayer and DNNLayer on x [B, 3341]) separately.  python profiles/time_cfg5_dense.py [B]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from recommender_system_b200.layer import set_seed
from recommender_system_b200.layer.interaction import DNNLayer, FMLayer

B = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
dev = torch.device("cuda", 0)
n, ld = 3341, 3344
set_seed(1)
x = (torch.rand((B, ld), device=dev) - 0.5) * 0.1
xv = x[:, :n]
fm = FMLayer(10, device=dev); fm.build((None, n)); fm.built = True
dn = DNNLayer([256, 128, 64], 1, "relu", precision="tf32", device=dev); dn.build((None, n)); dn.built = True


def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print("B=%d  x = %.2f GB" % (B, B * ld * 4 / 1e9))
print("FMLayer   %.3f ms" % t(lambda: fm(xv)))
print("DNNLayer  %.3f ms" % t(lambda: dn(xv, x_ld=ld)))
