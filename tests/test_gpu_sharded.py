"""GPU parity for the row-sharded lookup kernels (K8: ctrhot_shard_route_ids / _gather_pack / _unpack, csrc/api.cu) and the
ShardedDeepFM model (cfg 5).

* G = 1 through the product path (no collective) against the float64 DeepFM oracle with tables regenerated from the hash;
* G = 2 and 4 simulated in ONE process: one ShardedEmbedding per virtual rank on the same device, the all-to-all replaced by
  slicing -- the assembled x of every virtual rank must be a bit-exact copy of the rows a direct lookup gives
  (the real NCCL run is tests/run_sharded_nccl.py under torchrun; the gloo twin of the plumbing is test_sharded_host.py).
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from oracle import synth as S
from recommender_system_b200.model.sharded import ShardedDeepFM, ShardedEmbedding

pytestmark = pytest.mark.gpu


def hash_tables(vocab, k, seed):
    return [S.hash_table_value(seed, f, np.arange(v)[:, None], np.arange(k)[None, :]) for f, v in enumerate(vocab)]


def test_hash_init_matches_numpy_twin(lib):
    dev = torch.device("cuda", 0)
    vocab, k = [5, 1000, 33, 2], 8
    for G in (1, 3):
        for rank in range(G):
            E = ShardedEmbedding(vocab, k, dev, rank, G)
            E.init_hash(99)
            got = E.arena.cpu().numpy()
            want = np.concatenate([t[rank::G] for t in hash_tables(vocab, k, 99)], 0)
            np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("G", [2, 4])
def test_route_gather_unpack_virtual_ranks(lib, G):
    dev = torch.device("cuda", 0)
    vocab, k, D, Bl = [7, 1, 4000, 130, 2, 999], 16, 13, 257
    F = len(vocab)
    tables = hash_tables(vocab, k, 7)
    embeds = [ShardedEmbedding(vocab, k, dev, r, G) for r in range(G)]
    for E in embeds:
        E.init_hash(7)
    rng = np.random.default_rng(3)
    ids = [np.stack([rng.integers(0, v, Bl) for v in vocab], 1).astype(np.int32) for _ in range(G)]
    routed = [embeds[r]._route(torch.from_numpy(ids[r]).to(dev)) for r in range(G)]
    counts = [[int(c) for c in routed[r][3].tolist()] for r in range(G)]
    for r in range(G):
        assert sum(counts[r]) == Bl * F
    xs = []
    for r in range(G):                                            # requester r
        local_row, field, slot, _ = routed[r]
        off = np.concatenate([[0], np.cumsum(counts[r])])
        rows = torch.empty((Bl * F, k), dtype=torch.float32, device=dev)
        for o in range(G):                                        # owner o serves its slice of r's requests
            a, b = int(off[o]), int(off[o + 1])
            rows[a:b] = embeds[o]._gather(local_row[a:b].contiguous(), field[a:b].contiguous())
        x = torch.zeros((Bl, D + F * k + 3), dtype=torch.float32, device=dev)
        from recommender_system_b200 import _lib
        from recommender_system_b200.layer.base import ptr, stream_ptr
        _lib.call("ctrhot_shard_unpack", ptr(rows), ptr(slot), Bl * F, F, k, D, ptr(x), x.stride(0), stream_ptr())
        xs.append(x.cpu().numpy())
    for r in range(G):
        want = np.concatenate([tables[f][ids[r][:, f]] for f in range(F)], 1)
        np.testing.assert_array_equal(xs[r][:, D:D + F * k], want)
        assert not xs[r][:, :D].any() and not xs[r][:, D + F * k:].any()
    for E in embeds:
        E.check_ids()


def test_sharded_deepfm_single_rank_matches_oracle(lib):
    dev = torch.device("cuda", 0)
    vocab, k, B = [50, 3, 2000, 77, 1000, 9] + [40] * 20, 32, 777
    m = ShardedDeepFM(vocab, k, fm_k=10, hidden_units=(64, 32), device=dev, rank=0, world=1)
    m.build()
    m.embed.init_hash(5)
    rng = np.random.default_rng(1)
    dense = rng.random((B, 13), dtype=np.float32)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], 1).astype(np.int32)
    got = m((dense, ids)).cpu().numpy().astype(np.float64)
    w = {n: t.detach().cpu().numpy().astype(np.float64) for n, t in m.named_weights().items()}
    tables = [t.astype(np.float64) for t in hash_tables(vocab, k, 5)]
    names = sorted(w)
    fm = next(n for n in names if n.endswith("/w0")).rsplit("/", 1)[0]
    dn = next(n for n in names if n.endswith("dense_0/kernel")).rsplit("/", 2)[0]
    hidden = [(w["%s/dense_%d/kernel" % (dn, i)], w["%s/dense_%d/bias" % (dn, i)]) for i in range(2)]
    out = (w[dn + "/output/kernel"], w[dn + "/output/bias"])
    x = np.concatenate([dense.astype(np.float64), ids.astype(np.float64)], 1)
    want = O.deepfm_model(x, tables, (w[fm + "/w0"], w[fm + "/w1"], w[fm + "/v"]), hidden, out, n_dense=13)
    assert np.max(np.abs(got - want)) <= 5e-4, np.max(np.abs(got - want))
    m.embed.check_ids()


@pytest.mark.parametrize("G,k,D,pad", [(1, 16, 13, 3), (2, 128, 13, 3), (4, 16, 13, 0), (3, 8, 0, 0), (4, 12, 2, 5), (2, 256, 1, 1)])
def test_peer_lookup_virtual_ranks(lib, G, k, D, pad):
    """ctrhot_shard_lookup_peer_f32 with the G arenas on one device: every virtual rank's x must be a bit-exact copy of the
    global rows, for aligned and misaligned destinations (D = 13 -> rows start 4 bytes past a 16-byte boundary)."""
    dev = torch.device("cuda", 0)
    vocab, Bl = [7, 1, 4000, 130, 2, 999], 301
    F = len(vocab)
    tables = hash_tables(vocab, k, 21)
    embeds = [ShardedEmbedding(vocab, k, dev, r, G, peer=("ipc" if r % 2 == 0 else False)) for r in range(G)]
    for E in embeds:
        E.init_hash(21)
        E.connect_local(embeds)
    rng = np.random.default_rng(5)
    for r in range(G):
        ids = np.stack([rng.integers(0, v, Bl) for v in vocab], 1).astype(np.int32)
        dense = rng.random((Bl, max(D, 1)), dtype=np.float32)[:, :D]
        x = torch.full((Bl, D + F * k + pad), 7.0, dtype=torch.float32, device=dev)
        dt = torch.from_numpy(np.ascontiguousarray(dense)).to(dev) if D else None
        embeds[r].lookup_into(torch.from_numpy(ids), x, D, dense=dt)
        got = x.cpu().numpy()
        want = np.concatenate([tables[f][ids[:, f]] for f in range(F)], 1)
        np.testing.assert_array_equal(got[:, D:D + F * k], want)
        if D:
            np.testing.assert_array_equal(got[:, :D], dense)
            assert not got[:, D + F * k:].any()
        embeds[r].check_ids()
    for E in embeds:
        E.close()


def test_peer_lookup_flags_bad_ids(lib):
    dev = torch.device("cuda", 0)
    vocab, k = [10, 20], 8
    E = ShardedEmbedding(vocab, k, dev, 0, 1, peer="ipc")
    E.init_hash(1)
    E.connect_local([E])
    x = torch.empty((2, 16), dtype=torch.float32, device=dev)
    E.lookup_into(torch.tensor([[3, 20], [-1, 0]], dtype=torch.int32), x, 0)
    with pytest.raises(IndexError):
        E.check_ids()
    got = x.cpu().numpy()
    assert not got[0, 8:].any() and not got[1, :8].any()
    E.close()


def test_sharded_deepfm_peer_single_rank_matches_a2a(lib):
    dev = torch.device("cuda", 0)
    vocab, k, B = [50, 3, 2000, 77, 1000, 9] + [40] * 20, 32, 500
    from recommender_system_b200.layer import set_seed
    outs = []
    rng = np.random.default_rng(1)
    dense = rng.random((B, 13), dtype=np.float32)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], 1).astype(np.int32)
    for mode in ("a2a", "peer"):
        set_seed(3)
        m = ShardedDeepFM(vocab, k, fm_k=10, hidden_units=(64, 32), device=dev, rank=0, world=1, exchange=mode)
        m.build()
        m.embed.init_hash(5)
        outs.append(m((dense, ids)).cpu().numpy())
        m.embed.check_ids()
        m.embed.close()
    np.testing.assert_array_equal(outs[0], outs[1])


@pytest.mark.parametrize("G,k,B", [(1, 32, 777), (2, 32, 300), (3, 16, 129), (4, 128, 260)])
def test_sharded_deepfm_fused_virtual_ranks(lib, G, k, B):
    """exchange='fused' (ctrhot_ctr_fused_sharded_tf32): G virtual ranks on one device, each with its own samples; the
    probabilities must match the float64 DeepFM oracle over the hash tables (2e-3 on logits -> 5e-4 on probabilities)."""
    from recommender_system_b200.layer import set_seed
    dev = torch.device("cuda", 0)
    vocab = [50, 3, 2000, 77, 1000, 9] + [40 + i for i in range(20)]
    models = []
    for r in range(G):
        set_seed(4)
        m = ShardedDeepFM(vocab, k, fm_k=10, hidden_units=(64, 32), device=dev, rank=r, world=G, exchange="fused", arena="torch")
        m.build()
        m.embed.init_hash(5)
        models.append(m)
    for m in models:
        m.embed.connect_local([q.embed for q in models])
    tables = [t.astype(np.float64) for t in hash_tables(vocab, k, 5)]
    for r, m in enumerate(models):
        rng = np.random.default_rng(10 + r)
        dense = rng.random((B, 13), dtype=np.float32)
        ids = np.stack([rng.integers(0, v, B) for v in vocab], 1).astype(np.int32)
        got = m((dense, ids)).cpu().numpy().astype(np.float64)
        m.embed.check_ids()
        w = {n: t.detach().cpu().numpy().astype(np.float64) for n, t in m.named_weights().items()}
        names = sorted(w)
        fm = next(n for n in names if n.endswith("/w0")).rsplit("/", 1)[0]
        dn = next(n for n in names if n.endswith("dense_0/kernel")).rsplit("/", 2)[0]
        hidden = [(w["%s/dense_%d/kernel" % (dn, i)], w["%s/dense_%d/bias" % (dn, i)]) for i in range(2)]
        out = (w[dn + "/output/kernel"], w[dn + "/output/bias"])
        x = np.concatenate([dense.astype(np.float64), ids.astype(np.float64)], 1)
        want = O.deepfm_model(x, tables, (w[fm + "/w0"], w[fm + "/w1"], w[fm + "/v"]), hidden, out, n_dense=13)
        assert np.max(np.abs(got - want)) <= 5e-4, (r, np.max(np.abs(got - want)))
    for m in models:
        m.embed.close()


def test_sharded_fused_flags_bad_ids(lib):
    from recommender_system_b200.layer import set_seed
    dev = torch.device("cuda", 0)
    vocab, k = [10, 20, 30, 40], 16
    set_seed(1)
    ms = [ShardedDeepFM(vocab, k, fm_k=4, hidden_units=(32, 16), n_dense=2, device=dev, rank=r, world=2, exchange="fused", arena="torch")
          for r in range(2)]
    for m in ms:
        m.build(); m.embed.init_hash(1)
    for m in ms:
        m.embed.connect_local([q.embed for q in ms])
    ids = np.array([[1, 2, 3, 4], [9, 19, 30, 0]], dtype=np.int32)          # 30 is one past the end of table 2
    ms[0]((np.zeros((2, 2), np.float32), ids))
    with pytest.raises(IndexError):
        ms[0].embed.check_ids()
    for m in ms:
        m.embed.close()
