"""Two processes (torchrun), one GPU each: bandwidth of (a) NCCL all_to_all_single, (b) the peer lookup kernel through a
CUDA-IPC mapping (ctrhot_peer_open) for small and large tables, (c) the same kernel through torch symmetric memory
(cuMem VMM mapping).  NCCL_DEBUG=INFO shows which transport NCCL picked."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from recommender_system_b200 import _lib
from recommender_system_b200.layer.base import ptr, stream_ptr
from recommender_system_b200.model.sharded import ShardedEmbedding

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def say(*a):
    if rank == 0:
        print(*a, flush=True)


n = 1 << 27
a = torch.empty(n, dtype=torch.float32, device=dev); b = torch.empty_like(a)
ms = t(lambda: dist.all_to_all_single(b, a))
say("NCCL all_to_all_single %d MB: %.3f ms, %.1f GB/s sent to the peer" % (n * 4 >> 20, ms, n * 2 / ms / 1e6))
del a, b
k, F, Bl = 128, 26, 65536
for V in (100_000, 4_000_000):
    vocab = [V] * F
    E = ShardedEmbedding(vocab, k, dev, rank, world, peer=True)
    E.init_hash(3)
    E.connect()
    x = torch.empty((Bl, F * k), dtype=torch.float32, device=dev)
    rng = np.random.default_rng(rank)
    other = 1 - rank
    for name, ids in (("random, all remote", rng.integers(0, V // 2, (Bl, F)) * 2 + other),
                      ("random, all local", rng.integers(0, V // 2, (Bl, F)) * 2 + rank),
                      ("sequential, all remote", (np.arange(Bl * F).reshape(Bl, F) % (V // 2)) * 2 + other)):
        idt = torch.from_numpy(ids.astype(np.int32)).to(dev)
        ms = t(lambda: E.lookup_into(idt, x, 0))
        say("IPC  V=%8d %-24s %.3f ms  rows read %.1f GB/s" % (V, name, ms, Bl * F * k * 4 / ms / 1e6))
    E.check_ids()
    E.close()
    del E, x
    torch.cuda.empty_cache()
# (c) symmetric memory
try:
    import torch.distributed._symmetric_memory as symm
    V = 4_000_000
    vocab = [V] * F
    E = ShardedEmbedding(vocab, k, dev, rank, world)              # torch arena, replaced below
    rows = int(E.row_off_all[:, -1].max().item())
    buf = symm.empty((rows, k), dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(buf, dist.group.WORLD)
    E.arena = buf[:E.total_rows]
    E.init_hash(3)
    torch.cuda.synchronize(); dist.barrier()
    ptrs = [int(p) for p in hdl.buffer_ptrs]
    say("symm buffer_ptrs", [hex(p) for p in ptrs])
    E.peers = (ctypes.c_void_p * world)(*ptrs)
    x = torch.empty((Bl, F * k), dtype=torch.float32, device=dev)
    rng = np.random.default_rng(rank)
    other = 1 - rank
    for name, ids in (("random, all remote", rng.integers(0, V // 2, (Bl, F)) * 2 + other),
                      ("random, all local", rng.integers(0, V // 2, (Bl, F)) * 2 + rank),
                      ("sequential, all remote", (np.arange(Bl * F).reshape(Bl, F) % (V // 2)) * 2 + other)):
        idt = torch.from_numpy(ids.astype(np.int32)).to(dev)
        ms = t(lambda: E.lookup_into(idt, x, 0))
        say("SYMM V=%8d %-24s %.3f ms  rows read %.1f GB/s" % (V, name, ms, Bl * F * k * 4 / ms / 1e6))
    E.check_ids()
    # bit-exact check of remote rows against the hash
    from oracle import synth as S
    got = x.cpu().numpy()[:64]
    want = np.concatenate([S.hash_table_value(3, f, ids[:64, f][:, None], np.arange(k)[None, :]) for f in range(F)], 1)
    say("SYMM rows bit-exact:", bool(np.array_equal(got, want)))
except Exception as e:                                            # diagnostics only
    say("symmetric memory path failed:", repr(e))
dist.barrier()
dist.destroy_process_group()
