"""Multi-GPU check of the row-sharded DeepFM (cfg 5), one process per GPU over NCCL:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29533 \
        tests/run_sharded_nccl.py

Every rank runs its B/G samples through ShardedDeepFM (tables hash-initialised, row r on rank r mod G), regenerates the rows
it needed from the hash on the host and compares with the float64 oracle; then all ranks agree that (a) the assembled x is a
bit-exact copy of the global rows and (b) the probabilities match the oracle.  Prints one "SHARDED OK" line from rank 0.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ctr_oracle as O  # noqa: E402
from oracle import synth as S  # noqa: E402
from recommender_system_b200.layer import set_seed  # noqa: E402
from recommender_system_b200.model.sharded import ShardedDeepFM  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    vocab = [100_003, 17, 5_000, 1, 2_000_000, 333] + [1000 + 7 * i for i in range(20)]
    k, B, seed = 128, 4096, 11
    rng = np.random.default_rng(1000 + rank)
    Bl = B // world
    dense = rng.random((Bl, 13), dtype=np.float32)
    ids = np.stack([rng.integers(0, v, Bl) for v in vocab], 1).astype(np.int32)
    rows = np.concatenate([S.hash_table_value(seed, f, ids[:, f][:, None], np.arange(k)[None, :]) for f in range(len(vocab))], 1)
    bad_x = bad_p = 0
    worst = 0.0
    for mode in ("a2a", "peer", "fused"):
        set_seed(1)                                           # same dense-side weights on every rank
        m = ShardedDeepFM(vocab, k, fm_k=10, hidden_units=(256, 128, 64), device=dev, rank=rank, world=world, exchange=mode)
        m.build()
        m.embed.init_hash(seed)
        got = m((dense, ids)).cpu().numpy().astype(np.float64)
        m.embed.check_ids()
        # x itself: rerun the lookup into a fresh buffer and compare with rows regenerated from the hash
        x = torch.zeros((Bl, m.x_ld), dtype=torch.float32, device=dev)
        m.embed.lookup_into(torch.from_numpy(ids), x, m.D)
        exact = bool(np.array_equal(x.cpu().numpy()[:, m.D:m.D + len(vocab) * k], rows))
        w = {n: t.detach().cpu().numpy().astype(np.float64) for n, t in m.named_weights().items()}
        names = sorted(w)
        fm = next(n for n in names if n.endswith("/w0")).rsplit("/", 1)[0]
        dn = next(n for n in names if n.endswith("dense_0/kernel")).rsplit("/", 2)[0]
        xfull = np.concatenate([dense.astype(np.float64), rows.astype(np.float64)], 1)
        y = O.fm_layer(xfull, w[fm + "/w0"], w[fm + "/w1"], w[fm + "/v"]) + O.dnn_layer(
            xfull, [(w["%s/dense_%d/kernel" % (dn, i)], w["%s/dense_%d/bias" % (dn, i)]) for i in range(3)],
            (w[dn + "/output/kernel"], w[dn + "/output/bias"]))
        want = O.sigmoid(0.5 * y)
        err = float(np.max(np.abs(got - want)))
        worst = max(worst, err)
        bad_x += 0 if exact else 1
        bad_p += 0 if err <= 5e-4 else 1
        dist.barrier()
        m.embed.close()
        del m, x
    flags = torch.tensor([bad_x, bad_p], dtype=torch.int32, device=dev)
    dist.all_reduce(flags)
    errs = torch.tensor([worst], dtype=torch.float64, device=dev)
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    if rank == 0:
        ok = int(flags[0].item()) == 0 and int(flags[1].item()) == 0
        print("SHARDED %s world=%d exchanges=a2a,peer,fused x_bit_exact=%s max|p - oracle|=%.3g" % ("OK" if ok else "FAIL", world,
                                                                          int(flags[0].item()) == 0, float(errs.item())), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if (int(flags[0].item()) == 0 and int(flags[1].item()) == 0) else 1)


if __name__ == "__main__":
    main()
