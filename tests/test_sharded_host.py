"""Host-side logic of the row-sharded lookup (cfg 5) on CPU: world_size-2 (and 3) process groups over gloo.

``exchange_rows`` (recommender_system_b200/model/sharded.py) is the product's all-to-all plumbing: split sizes from the
routed counts, request order, reply order.  Here it runs with CPU stand-ins for the three CUDA steps (route / gather / unpack
-- the stand-ins restate csrc/api.cu's kernels in torch and live in this test only), and every rank's assembled x must equal
a direct lookup in the full tables (bit-exact: rows are copies).  The GPU twin is tests/test_gpu_sharded.py.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from recommender_system_b200.model.sharded import exchange_rows


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _tables(vocab, k, seed=0):
    rng = np.random.default_rng(seed)
    return [rng.standard_normal((v, k)).astype(np.float32) for v in vocab]


def _worker(rank, world, port, vocab, k, B, D, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        F = len(vocab)
        tables = _tables(vocab, k)
        local = [torch.from_numpy(t[rank::world].copy()) for t in tables]      # row r on rank r % G at r // G
        rng = np.random.default_rng(100 + rank)
        Bl = B // world + (1 if rank == 0 else 0)                               # ragged: rank 0 has one more sample
        ids = torch.from_numpy(np.stack([rng.integers(0, v, Bl) for v in vocab], 1).astype(np.int32))
        x = torch.zeros((Bl, D + F * k), dtype=torch.float32)

        def route(ids_):                                                        # ctrhot_shard_route_ids, restated
            flat = ids_.reshape(-1).to(torch.int64)
            owner = flat % world
            order = torch.argsort(owner, stable=True)
            counts = torch.bincount(owner, minlength=world).to(torch.int32)
            return ((flat // world)[order].to(torch.int32), (torch.arange(flat.numel()) % F)[order].to(torch.int32),
                    order.to(torch.int32), counts)

        def gather(local_row, field):                                           # ctrhot_shard_gather_pack, restated
            out = torch.empty((local_row.numel(), k), dtype=torch.float32)
            for i in range(local_row.numel()):
                out[i] = local[int(field[i])][int(local_row[i])]
            return out

        def unpack(rows, slot):                                                 # ctrhot_shard_unpack, restated
            s = slot.to(torch.int64)
            b, f = s // F, s % F
            for i in range(s.numel()):
                x[b[i], D + f[i] * k: D + (f[i] + 1) * k] = rows[i]

        served = exchange_rows(ids, F, k, world, route, gather, unpack)
        want = np.concatenate([np.zeros((Bl, D), np.float32)] + [tables[f][ids[:, f].numpy()] for f in range(F)], 1)
        ok = np.array_equal(x.numpy(), want)
        tot = torch.tensor([served], dtype=torch.int64)
        dist.all_reduce(tot)
        q.put((rank, bool(ok), int(tot.item()), Bl * F))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_rows_over_gloo(world):
    vocab, k, B, D = [7, 1, 40, 13, 2], 4, 30, 3                                # a 1-row and a 2-row table: empty shards
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, vocab, k, B, D, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _, _ in res), res
    # every request was served exactly once, somewhere
    assert res[0][2] == sum(n for _, _, _, n in res)


def test_hash_table_value_is_deterministic_and_in_range():
    from oracle import synth as S
    v = S.hash_table_value(1234, np.arange(3)[:, None, None], np.arange(1000)[None, :, None], np.arange(8)[None, None, :])
    assert v.dtype == np.float32 and v.shape == (3, 1000, 8)
    assert float(v.min()) >= -0.05 and float(v.max()) < 0.05
    assert abs(float(v.mean())) < 2e-3 and np.unique(v).size > 0.99 * v.size
    np.testing.assert_array_equal(v, S.hash_table_value(1234, np.arange(3)[:, None, None], np.arange(1000)[None, :, None],
                                                        np.arange(8)[None, None, :]))
