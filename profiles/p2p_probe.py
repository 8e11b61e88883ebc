"""One process, two GPUs: how fast can a kernel on cuda:0 read rows that live on cuda:1 over NVLink?
(a) torch copy cuda:1 -> cuda:0 (copy engine), (b) ctrhot_shard_lookup_peer_f32 with random / sequential ids, several row widths.
python profiles/p2p_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from recommender_system_b200.model.sharded import ShardedEmbedding

d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
print("can access peer:", torch.cuda.can_device_access_peer(0, 1))


def t(fn, reps=5):
    fn(); torch.cuda.synchronize(d0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize(d0)
    return e0.elapsed_time(e1) / reps


with torch.cuda.device(d0):
    from recommender_system_b200 import _lib
    _lib.check(_lib.load().ctrhot_peer_enable(1), "peer_enable")
    a = torch.empty(1 << 28, dtype=torch.float32, device=d1)      # 1 GiB on GPU 1
    b = torch.empty(1 << 28, dtype=torch.float32, device=d0)
    ms = t(lambda: b.copy_(a))
    print("copy engine 1->0: %.1f GB/s" % (a.numel() * 4 / ms / 1e6))
    del a, b
    for k in (128, 32):
        V, F, Bl = 4_000_000, 26, 65536
        vocab = [V] * F
        E = [ShardedEmbedding(vocab, k, d0, 0, 2), None]
        with torch.cuda.device(d1):
            E[1] = ShardedEmbedding(vocab, k, d1, 1, 2)
            E[1].init_hash(3)
            torch.cuda.synchronize(d1)
        E[0].init_hash(3)
        E[0].connect_local(E)
        x = torch.empty((Bl, F * k), dtype=torch.float32, device=d0)
        rng = np.random.default_rng(0)
        for name, ids in (("random, all remote", rng.integers(0, V // 2, (Bl, F)) * 2 + 1),
                          ("random, all local", rng.integers(0, V // 2, (Bl, F)) * 2),
                          ("random, half remote", rng.integers(0, V, (Bl, F))),
                          ("sequential, all remote", (np.arange(Bl * F).reshape(Bl, F) % (V // 2)) * 2 + 1)):
            idt = torch.from_numpy(ids.astype(np.int32)).to(d0)
            ms = t(lambda: E[0].lookup_into(idt, x, 0))
            print("k=%3d %-24s %.3f ms  rows read %.1f GB/s" % (k, name, ms, Bl * F * k * 4 / ms / 1e6))
        E[0].check_ids()
        del E, x
        torch.cuda.empty_cache()
