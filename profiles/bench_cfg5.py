"""Only the row-sharded config-5 measurement of bench.py (all three exchanges), 1 process or under torchrun:

    python profiles/bench_cfg5.py [rows_per_table [modes, e.g. fused,peer]]      (CTRHOT_CFG5_CLUSTER=1|2|4 for the fused kernel)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 profiles/bench_cfg5.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import bench

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
modes = tuple(sys.argv[2].split(",")) if len(sys.argv) > 2 else ("fused", "peer", "a2a")
r = bench.sharded_cfg5(dev, rank, world, rows_per_table=rows, modes=modes)
if rank == 0:
    print(json.dumps(r), flush=True)
if world > 1:
    dist.destroy_process_group()
