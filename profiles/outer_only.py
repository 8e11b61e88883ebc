import os, sys
sys.path.insert(0, "/root/repo")
import torch
from recommender_system_b200.layer.interaction import OuterProductLayer
dev = torch.device("cuda", 0)
e = torch.rand((65536, 26, 16), device=dev) - 0.5
op = OuterProductLayer(device=dev); op.build(e.shape)
out = torch.empty((65536, 325), device=dev)
for _ in range(3): op(e, out=out)
torch.cuda.synchronize()
