"""Row-sharded embedding tables + batch-data-parallel DeepFM for the large-vocabulary config (cfg 5).

Row ``r`` of every table lives on rank ``r mod G`` at local row ``r div G`` (round-robin beats range partitioning under
skew).  Rank ``g`` owns samples ``[g*B/G, (g+1)*B/G)`` for the dense part (FM + tower), so the only data-path collectives
are the two all-to-alls of one lookup:

    1. route   : bucket this rank's B_l x F ids by owner                      (ctrhot_shard_route_ids)
    2. a2a ids : counts, then (local_row, field) pairs, owner-major           (torch.distributed, NCCL over NVLink)
    3. gather  : the owner reads its rows                                     (ctrhot_shard_gather_pack)
    4. a2a rows: back to the requester                                        (torch.distributed)
    5. unpack  : straight into x[b, D + f*k : ...]                            (ctrhot_shard_unpack)

then ``sigmoid(0.5 * (FM(x) + DNN(x)))`` exactly as DeepFM.call (DL/model/deepFM.py:23-31) on the local samples.

``exchange="peer"`` replaces steps 1-5 by ONE kernel (ctrhot_shard_lookup_peer_f32, csrc/peer.cu): every rank maps its
peers' arenas once at load time (CUDA IPC handles exchanged over the process group) and its gather kernel then reads the
rows it needs straight out of the owner's HBM over NVLink/NVSwitch -- no staging buffers, no host round trip for counts.
``exchange="fused"`` goes one step further: the whole-model kernel of the unsharded DeepFM (csrc/fused.cu) gathers from
the mapped arenas itself (ctrhot_ctr_fused_sharded_tf32), so x never exists in HBM and the NVLink reads of one tile run
under the tensor work of the previous ones: one launch per forward, no collective on the data path.
``exchange_rows`` is written against three callables so that the host-side plumbing (split sizes, ordering) is testable
with the gloo backend on CPU (tests/test_sharded_host.py); the product passes the CUDA entry points below -- there is no
CPU compute path in the package.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch
import torch.distributed as dist

from .. import _lib
from ..layer.base import Model, ptr, require_cuda, stream_ptr
from ..layer.interaction import DNNLayer, FMLayer
from .criteo import _pad32, _permute_rows


def exchange_rows(ids, F, k, G, route, gather, unpack, group=None):
    """The two all-to-alls of one sharded lookup.  ``route(ids) -> (local_row, field, slot, counts)`` (owner-major, counts
    [G]); ``gather(local_row, field) -> rows [n, k]``; ``unpack(rows, slot)`` consumes the rows in request order.
    Returns the number of rows this rank served (for accounting)."""
    local_row, field, slot, counts = route(ids)
    send_counts = [int(c) for c in counts.tolist()]
    if G == 1:
        rows = gather(local_row, field)
        unpack(rows, slot)
        return int(local_row.shape[0])
    cnt_in = torch.tensor(send_counts, dtype=torch.int64, device=counts.device)
    cnt_out = torch.empty_like(cnt_in)
    dist.all_to_all_single(cnt_out, cnt_in, group=group)                       # how many requests each peer sends me
    recv_counts = [int(c) for c in cnt_out.tolist()]
    n_recv = sum(recv_counts)
    # one message per peer: [local_row | field] pairs
    req_out = torch.stack([local_row, field], dim=1).contiguous()              # [n_send, 2] int32
    req_in = torch.empty((n_recv, 2), dtype=req_out.dtype, device=req_out.device)
    dist.all_to_all_single(req_in, req_out, recv_counts, send_counts, group=group)
    rows_out = gather(req_in[:, 0].contiguous(), req_in[:, 1].contiguous())    # [n_recv, k] rows I own
    rows_in = torch.empty((int(local_row.shape[0]), k), dtype=rows_out.dtype, device=rows_out.device)
    dist.all_to_all_single(rows_in, rows_out, send_counts, recv_counts, group=group)
    unpack(rows_in, slot)
    return n_recv


class _PeerArena:
    """A ctrhot_peer_alloc'd (cudaMalloc) block seen by torch through ``__cuda_array_interface__``."""

    def __init__(self, rows, k):
        p = ctypes.c_void_p()
        _lib.check(_lib.load().ctrhot_peer_alloc(int(rows) * int(k) * 4, ctypes.byref(p)), "ctrhot_peer_alloc")
        self.ptr = int(p.value)
        self.__cuda_array_interface__ = {"shape": (int(rows), int(k)), "typestr": "<f4", "data": (self.ptr, False),
                                         "version": 2, "strides": None}

    def free(self):
        if self.ptr:
            _lib.load().ctrhot_peer_free(ctypes.c_void_p(self.ptr))
            self.ptr = 0


class ShardedEmbedding:
    """Local shard of F tables of ``vocab[f]`` global rows each, width k, on ``device``; rank/world from the process group.
    ``peer=True`` allocates the arena so that it can be mapped by the other ranks (``connect`` / ``connect_local``):
    torch symmetric memory when world > 1 (``peer="ipc"``: a cudaMalloc block shared through CUDA IPC handles instead)."""

    def __init__(self, vocab, k, device, rank=0, world=1, group=None, peer=False):
        self.vocab, self.k, self.F = [int(v) for v in vocab], int(k), len(vocab)
        self.device, self.rank, self.G, self.group = device, rank, world, group
        every = [[(v - r + world - 1) // world for v in self.vocab] for r in range(world)]   # rows r with r % G == rank
        local = every[rank]
        self.local_rows = local
        offs = np.stack([np.concatenate([[0], np.cumsum(e)]) for e in every]).astype(np.int64)
        self.row_off_all = torch.tensor(offs, dtype=torch.int64, device=device)              # [G, F+1]
        self.row_off = self.row_off_all[rank].contiguous()
        self.total_rows = int(sum(local))
        self._block = None
        self._symm = None
        if peer and world > 1 and peer != "ipc":
            # torch symmetric memory = cuMem (VMM) allocations mapped into every rank with 2 MB pages.  Measured on 2 B200s
            # (profiles/p2p_probe_mp.py): random 512-B rows out of a 26 GB peer shard read at 634 GB/s through this mapping
            # but at 100 GB/s through a cudaIpcOpenMemHandle mapping (small pages: TLB-bound), so IPC is opt-in only.
            import torch.distributed._symmetric_memory as symm
            rows = int(offs[:, -1].max())                                      # symmetric: same size on every rank
            self._symm = symm.empty((max(rows, 1), k), dtype=torch.float32, device=device)
            self.arena = self._symm[:self.total_rows]
        elif peer == "ipc" and self.total_rows > 0:
            with torch.cuda.device(device):
                self._block = _PeerArena(self.total_rows, k)
            self.arena = torch.as_tensor(self._block, device=device)
        else:
            self.arena = torch.empty((self.total_rows, k), dtype=torch.float32, device=device)
        self.err = torch.zeros(1, dtype=torch.int32, device=device)
        self._ws = None
        self.peers = None                                                      # ctypes array of G arena pointers
        self._opened = []

    # -- peer mapping (load time; not on the data path)
    def connect(self):
        """Exchange CUDA IPC handles over the process group and map every peer's arena into this process."""
        lib = _lib.load()
        if self.G == 1:
            return self.connect_local([self])
        if self._symm is not None:
            import torch.distributed._symmetric_memory as symm
            torch.cuda.synchronize(self.device)
            hdl = symm.rendezvous(self._symm, self.group if self.group is not None else dist.group.WORLD)
            self._hdl = hdl
            self.peers = (ctypes.c_void_p * self.G)(*[int(p) for p in hdl.buffer_ptrs])
            dist.barrier(group=self.group)
            return
        if self._block is None:
            raise _lib.CtrhotError("ShardedEmbedding.connect: build with peer=True")
        nb = lib.ctrhot_peer_handle_bytes()
        buf = (ctypes.c_ubyte * nb)()
        with torch.cuda.device(self.device):
            _lib.check(lib.ctrhot_peer_export(ctypes.c_void_p(self._block.ptr), buf), "ctrhot_peer_export")
        mine = torch.tensor(list(bytes(buf)), dtype=torch.uint8, device=self.device)
        every = torch.empty((self.G, nb), dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(every, mine, group=self.group)
        every = every.cpu().numpy()
        ptrs = []
        with torch.cuda.device(self.device):
            for g in range(self.G):
                if g == self.rank:
                    ptrs.append(self._block.ptr)
                    continue
                h = (ctypes.c_ubyte * nb)(*every[g].tolist())
                p = ctypes.c_void_p()
                _lib.check(lib.ctrhot_peer_open(h, ctypes.byref(p)), "ctrhot_peer_open(rank %d)" % g)
                self._opened.append(int(p.value))
                ptrs.append(int(p.value))
        self.peers = (ctypes.c_void_p * self.G)(*ptrs)
        dist.barrier(group=self.group)

    def connect_local(self, embeds):
        """All G shards live in THIS process (tests, G = 1): take their arena pointers directly."""
        assert len(embeds) == self.G
        self.peers = (ctypes.c_void_p * self.G)(*[int(e.arena.data_ptr()) or 1 for e in embeds])
        self._keep = embeds

    def close(self):
        """Unmap the peers, then (after everyone has unmapped) free the arena."""
        lib = _lib.load()
        had_peers = bool(self._opened)
        for p in self._opened:
            lib.ctrhot_peer_close(ctypes.c_void_p(p))
        self._opened = []
        self.peers = None
        if (had_peers or self._symm is not None) and dist.is_initialized():
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
        if self._symm is not None:
            self.arena = None
            self._hdl = None
            self._symm = None
        if self._block is not None:
            torch.cuda.synchronize(self.device)
            self._block.free()
            self._block = None
        self.arena = None                                                      # an ordinary torch arena is released too

    def init_hash(self, seed=1234):
        """value(f, global row, col) from the counter-based hash: any rank can regenerate any row (oracle.synth.hash_table_value)."""
        _lib.call("ctrhot_shard_init_hash_f32", ptr(self.arena), ptr(self.row_off), self.total_rows, self.F, self.k,
                  self.G, self.rank, int(seed), stream_ptr())
        self._fence_peers()

    def set_global_tables(self, tables):
        """Tests: take this rank's rows out of full tables (list of [V_f, k] arrays)."""
        parts = [torch.as_tensor(np.asarray(t)[self.rank::self.G], dtype=torch.float32) for t in tables]
        self.arena.copy_(torch.cat(parts, dim=0).to(self.device))
        self._fence_peers()

    def _fence_peers(self):
        """The arena is peer-visible once connected: a fill must have finished on EVERY rank before any rank's kernels may read
        rows over NVLink again (filling before connect() needs nothing: connect() itself synchronises and rendezvouses)."""
        if self.peers is not None:
            torch.cuda.synchronize(self.device)
            import torch.distributed as dist
            if self.G > 1 and dist.is_available() and dist.is_initialized():
                dist.barrier()

    # -- the three CUDA steps
    def _route(self, ids):
        B, F = ids.shape
        n = B * F
        dev = ids.device
        local_row = torch.empty(n, dtype=torch.int32, device=dev)
        field = torch.empty(n, dtype=torch.int32, device=dev)
        slot = torch.empty(n, dtype=torch.int32, device=dev)
        counts = torch.empty(self.G, dtype=torch.int32, device=dev)
        need = _lib.query("ctrhot_shard_route_workspace_bytes", B, F, self.G)
        if self._ws is None or self._ws.numel() * 4 < need:
            self._ws = torch.empty((need + 3) // 4, dtype=torch.int32, device=dev)
        _lib.call("ctrhot_shard_route_ids", ptr(ids), B, F, self.G, ptr(local_row), ptr(field), ptr(slot), ptr(counts),
                  ptr(self._ws), self._ws.numel() * 4, stream_ptr())
        return local_row, field, slot, counts

    def _gather(self, local_row, field):
        n = int(local_row.shape[0])
        rows = torch.empty((n, self.k), dtype=torch.float32, device=self.device)
        _lib.call("ctrhot_shard_gather_pack", ptr(self.arena), ptr(self.row_off), self.F, self.k, ptr(local_row),
                  ptr(field), n, ptr(rows), ptr(self.err), stream_ptr())
        return rows

    def lookup_into(self, ids, x, D, dense=None):
        """x[b, D + f*k : D + (f+1)*k] = table_f[ids[b, f]] for this rank's samples; x is [B_l, >= D + F*k] fp32.
        With mapped peers it is one kernel (which also copies ``dense`` into x[:, :D] when given)."""
        require_cuda(x, "ShardedEmbedding")
        ids = ids.to(device=self.device, dtype=torch.int32).contiguous()
        if self.peers is not None:
            _lib.call("ctrhot_shard_lookup_peer_f32", self.peers, ptr(self.row_off_all), self.F, self.k, self.G, ptr(ids),
                      int(ids.shape[0]), D, ptr(dense) if dense is not None else None,
                      dense.stride(0) if dense is not None else 0, ptr(x), x.stride(0), ptr(self.err), stream_ptr())
            return int(ids.numel())
        if dense is not None:
            x[:, :D] = dense

        def unpack(rows, slot):
            _lib.call("ctrhot_shard_unpack", ptr(rows), ptr(slot), int(slot.shape[0]), self.F, self.k, D, ptr(x),
                      x.stride(0), stream_ptr())
        return exchange_rows(ids, self.F, self.k, self.G, self._route, self._gather, unpack, self.group)

    def check_ids(self):
        rc = _lib.load().ctrhot_check_flag(ptr(self.err), stream_ptr())
        _lib.check(rc, "ShardedEmbedding")


class ShardedDeepFM(Model):
    """DeepFM over row-sharded tables: ``call((dense [B_l,13], ids [B_l,26]))`` on every rank with ITS samples returns
    their probabilities [B_l,1].  Dense-side weights (FM, tower) are replicated: build them with the same seed on all ranks
    (or broadcast once at load time -- not on the data path).  ``arena``: how a peer-readable shard is allocated -- "auto"
    (symmetric memory when world > 1), "ipc" (cudaMalloc + CUDA IPC) or "torch" (ordinary allocation: all shards in one
    process, wired with ``embed.connect_local``)."""

    def __init__(self, vocab, k, fm_k=10, hidden_units=(256, 128, 64), activation="relu", n_dense=13, precision="tf32",
                 device=None, rank=None, world=None, group=None, exchange="a2a", arena="auto"):
        super().__init__(device)
        if exchange not in ("a2a", "peer", "fused"):
            raise ValueError("exchange must be 'a2a' (NCCL all-to-all pair), 'peer' (one gather kernel with P2P loads over "
                             "NVLink) or 'fused' (whole-model kernel gathering from the peers)")
        self.exchange = exchange
        self.cluster = 0                                        # CTAs sharing a weight chunk in the fused kernel (0 = auto)
        self.overlap_launches = False                           # exchange='fused': consecutive calls on independent batches may overlap (CTRHOT_LAUNCH_OVERLAP)
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        if world is None:
            world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.D, self.k, self.F = n_dense, k, len(vocab)
        self.n = n_dense + self.F * k
        self.embed = ShardedEmbedding(vocab, k, self.device, rank, world, group, peer=(False if exchange == "a2a" else {"auto": True, "torch": False, "ipc": "ipc"}[arena]))
        self.fm = self.track(FMLayer(fm_k, device=self.device))
        self.dnn = self.track(DNNLayer(list(hidden_units), 1, activation, precision=precision, device=self.device))
        self.x_ld = (self.n + 3) // 4 * 4
        self.built_dense = False

    def build(self, input_shape=None):
        if not self.built_dense:
            self.fm.build((None, self.n)); self.fm.built = True
            self.dnn.build((None, self.n)); self.dnn.built = True
            self.built_dense = True

    def call(self, inputs, training=None, mask=None):
        dense, ids = inputs
        dense = torch.as_tensor(dense).to(device=self.device, dtype=torch.float32)
        ids = torch.as_tensor(ids)
        self.build()
        B = dense.shape[0]
        if self.exchange == "fused":
            return self._call_fused(dense, ids, B)
        x = torch.empty((B, self.x_ld), dtype=torch.float32, device=self.device)
        if self.exchange == "peer":
            if self.embed.peers is None:
                self.embed.connect()
            self.embed.lookup_into(ids, x, self.D, dense=dense.contiguous())   # dense copy + pad fused into the kernel
        else:
            x[:, :self.D] = dense                               # reference column order [dense | embeddings] (deepFM.py:26)
            if self.x_ld > self.n:
                x[:, self.n:] = 0
            self.embed.lookup_into(ids, x, self.D)
        xv = x[:, :self.n]
        fm = self.fm(xv)                                        # [B,1] fp32 (FMLayer, interaction.py:106-114)
        dn = self.dnn(xv, x_ld=self.x_ld)                       # [B,1] tower
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        _lib.call("ctrhot_combine_sigmoid_f32", ptr(fm), ptr(dn), 0, 0, 0.5, B, ptr(out), stream_ptr())
        return out

    def _call_fused(self, dense, ids, B):
        """gather (local + peer arenas) -> split-TF32 FM -> TMEM-chained tower -> sigmoid(0.5 (fm + dnn)), ONE launch."""
        E, D = self.embed, self.D
        if E.peers is None:
            E.connect()
        stack = self.dnn.stack
        key = (self.dnn._version, self.fm._version)
        if getattr(self, "_fz", None) is None or self._fz[0] != key:
            ls = self.dnn.layer_dicts()
            ls[0] = dict(ls[0], kernel=_permute_rows(ls[0]["kernel"], D))      # internal column order [embeddings | dense]
            n, kf = self.fm.v.shape
            side = torch.zeros((n, 16), dtype=torch.float32, device=self.device)
            side[:, :kf] = self.fm.v
            side[:, kf] = self.fm.w1[:, 0]
            n_pad = _pad32(n)
            sq_w = self.fm.planes(D)[(kf + 1) * n_pad:(kf + 2) * n_pad]
            self._fz = (key, ls, side, sq_w, [float(self.fm.w0.item())])
        _, ls, side, sq_w, cb = self._fz
        stack.set_layers(ls, (self.dnn._version, D))
        if self.fm.k >= 16 or not stack.ctr_fused_ok(E.F, E.k):
            raise _lib.CtrhotError("ShardedDeepFM(exchange='fused'): this tower does not fit the whole-model kernel; "
                                   "use exchange='peer'")
        ids = ids.to(device=self.device, dtype=torch.int32).contiguous()
        out = torch.empty((B, 1), dtype=torch.float32, device=self.device)
        stack.run_ctr_fused(E, ids, _lib.IDS_I32, ids.stride(0), dense.contiguous(), B, self.dnn.hidden[0][0], side, sq_w,
                            mode=1, n_side=self.fm.k, combine=cb, sigmoid=True, err_flag=E.err, out=out,
                            cluster=self.cluster | (_lib.LAUNCH_OVERLAP if self.overlap_launches else 0),
                            side_key=("fm", self.fm._version))
        return out
