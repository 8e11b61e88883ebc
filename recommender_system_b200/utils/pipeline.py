"""Host-side feeder for the Criteo models: the call a serving user makes when the batches live in host memory.

``HostPipeline(model, batch).submit(dense, ids)`` copies one batch from (pinned) host memory, runs ``model((dense, ids))``
and copies the probabilities back, on three CUDA streams with ``depth`` slots in flight, so that the PCIe transfers of
batch i+1 overlap the kernel of batch i.  Every batch still pays its own H2D and D2H copy; nothing is cached across
batches.  ``result(ticket)`` blocks until that batch's probabilities are in host memory and returns them.  Use ``check_ids=False`` models
here: the per-call id check synchronises the device and would serialise the pipeline (ids can be checked once per epoch with
``model.check_ids()``).

(The reference has no counterpart: its ``model.predict(X)`` hands the whole numpy matrix to TensorFlow.  This is the
input step on the left of the hot path, SURVEY 8(f) rank 2, reduced to what the bench's end-to-end number needs.)
"""
from __future__ import annotations

import torch


class HostPipeline:
    def __init__(self, model, batch, n_dense=13, n_sparse=26, depth=3, device=None, layout=None):
        self.model = model
        self.dev = device if device is not None else model.device
        if self.dev.type != "cuda":
            raise RuntimeError("HostPipeline needs a CUDA device (there is no CPU path)")
        self.depth = depth
        self.s_in, self.s_run, self.s_out = (torch.cuda.Stream(self.dev) for _ in range(3))
        self.slots = []
        for _ in range(depth):
            self.slots.append({
                "dense": torch.empty((batch, n_dense), dtype=torch.float32, device=self.dev),
                "ids": torch.empty((batch, n_sparse), dtype=torch.int32, device=self.dev),
                "host": torch.empty((batch, 1), dtype=torch.float32).pin_memory(),
                "e_in": torch.cuda.Event(), "e_run": torch.cuda.Event(), "e_out": torch.cuda.Event(),
                "busy": False, "out": None})
        self.n = 0
        self.h2d_bytes = batch * (n_dense + n_sparse) * 4
        self.d2h_bytes = batch * 4
        # compact transport records (utils/transport.py): ONE copy of rec_bytes per sample instead of two of 156 B in all
        self.layout = layout
        if layout is not None:
            from .transport import PackedRows
            for s in self.slots:
                s["rec"] = torch.empty((batch, layout.rec_bytes), dtype=torch.uint8, device=self.dev)
                s["packed"] = PackedRows(layout, s["rec"])
            self.h2d_bytes_packed = batch * layout.rec_bytes

    def submit(self, dense, ids):
        """dense: host float32 [B, n_dense]; ids: host int32 [B, n_sparse] (pinned for a truly asynchronous copy)."""
        k = self.n % self.depth
        s = self.slots[k]
        if s["busy"]:
            s["e_out"].synchronize()                  # the slot's previous batch must have left the device
        if not isinstance(dense, torch.Tensor):
            dense = torch.from_numpy(dense)
        if not isinstance(ids, torch.Tensor):
            ids = torch.from_numpy(ids)
        B = dense.shape[0]
        with torch.cuda.stream(self.s_in):
            s["dense"][:B].copy_(dense, non_blocking=True)
            s["ids"][:B].copy_(ids, non_blocking=True)
            s["e_in"].record(self.s_in)
        self.s_run.wait_event(s["e_in"])
        with torch.cuda.stream(self.s_run):
            s["out"] = self.model((s["dense"][:B], s["ids"][:B]))
            s["e_run"].record(self.s_run)
        self.s_out.wait_event(s["e_run"])
        with torch.cuda.stream(self.s_out):
            s["host"][:B].copy_(s["out"], non_blocking=True)
            s["out"].record_stream(self.s_out)        # allocated on s_run, read on s_out: keep it alive until the copy is done
            s["e_out"].record(self.s_out)
        s["busy"], s["B"], s["seq"] = True, B, self.n
        self.n += 1
        return (k, s["seq"])

    def submit_packed(self, records):
        """records: host uint8 [B, layout.rec_bytes] (``RowLayout.pack`` output, pinned for an asynchronous copy)."""
        from .transport import PackedRows
        k = self.n % self.depth
        s = self.slots[k]
        if s["busy"]:
            s["e_out"].synchronize()
        if not isinstance(records, torch.Tensor):
            records = torch.from_numpy(records)
        B = records.shape[0]
        with torch.cuda.stream(self.s_in):
            s["rec"][:B].copy_(records, non_blocking=True)
            s["e_in"].record(self.s_in)
        self.s_run.wait_event(s["e_in"])
        with torch.cuda.stream(self.s_run):
            batch = s["packed"] if B == s["rec"].shape[0] else PackedRows(self.layout, s["rec"][:B])
            s["out"] = self.model(batch)
            s["e_run"].record(self.s_run)
        self.s_out.wait_event(s["e_run"])
        with torch.cuda.stream(self.s_out):
            s["host"][:B].copy_(s["out"], non_blocking=True)
            s["out"].record_stream(self.s_out)        # allocated on s_run, read on s_out: keep it alive until the copy is done
            s["e_out"].record(self.s_out)
        s["busy"], s["B"], s["seq"] = True, B, self.n
        self.n += 1
        return (k, s["seq"])

    def result(self, ticket):
        """Blocks until that batch's probabilities are in host memory.  A ticket is (slot, sequence number): asking for a
        batch whose slot has been reused since (more than ``depth`` submits ago) raises instead of returning another batch."""
        k, seq = ticket
        s = self.slots[k]
        if s.get("seq") != seq:
            raise RuntimeError("HostPipeline: the slot of ticket %r now holds batch %r (collect results within %d submits)"
                               % (ticket, s.get("seq"), self.depth))
        s["e_out"].synchronize()
        s["busy"] = False
        return s["host"][:s["B"]].numpy()

    def drain(self):
        """Results of the batches still in flight, oldest first."""
        live = sorted((s["seq"], k) for k, s in enumerate(self.slots) if s["busy"])
        return [self.result((k, seq)) for seq, k in live]
