"""The retrieval hot path as one call: ids in, top-k evidence out.

``RetrievalPipeline.run(batch)`` is the public entry point ``bench.py`` times end to end. Per batch
of questions it runs, on the current CUDA stream:

  1. edge segmentation / validation               (graph_utils.py:49-104)
  2. multi-source BFS hop distance from the seeds (build_retrieval_pipeline.py:610-631), depth-capped
  3. DDE structure features + projections + triple scoring (retriever.py:185-289)
  4. ragged top-k and the writer's gather         (retriever_topk_edge_writer.py:201-328)

Inputs may be host (pinned) or device tensors with the reference's batch attribute names; outputs
are fixed-shape ``[B, k]`` arrays (what the reference pickles per graph, as arrays).
"""
from __future__ import annotations

import types
from typing import Any, Dict, Optional

import numpy as np
import torch

from . import graph as G
from .retriever import Retriever
from .topk import segmented_topk, writer_payload

_BATCH_FIELDS = ("edge_index", "edge_attr", "ptr", "node_embedding_ids", "node_global_ids", "topic_one_hot",
                 "question_emb", "q_local_indices", "q_ptr")


def host_batch(qb, pin: bool = True) -> types.SimpleNamespace:
    """A synthetic ``QuestionBatch`` as (pinned) host tensors under the reference's attribute names."""
    b = types.SimpleNamespace()
    for name in _BATCH_FIELDS:
        t = torch.from_numpy(np.ascontiguousarray(getattr(qb, name)))
        if pin and torch.cuda.is_available():
            t = t.pin_memory()
        setattr(b, name, t)
    b.num_graphs = qb.num_graphs
    b.q_ptr_host = np.asarray(qb.q_ptr)
    b.ptr_host = np.asarray(qb.ptr)
    b.edge_ptr_host = np.asarray(qb.edge_ptr)  # what a PyG Batch carries as _slice_dict["edge_index"]
    return b


def device_batch(hb, device) -> types.SimpleNamespace:
    """Async H2D of every field of a host batch (non_blocking from pinned memory)."""
    b = types.SimpleNamespace()
    for name in _BATCH_FIELDS:
        setattr(b, name, getattr(hb, name).to(device, non_blocking=True))
    b.num_graphs, b.q_ptr_host, b.ptr_host = hb.num_graphs, hb.q_ptr_host, hb.ptr_host
    return b


def h2d_bytes(hb) -> int:
    return int(sum(getattr(hb, n).numel() * getattr(hb, n).element_size() for n in _BATCH_FIELDS))


class RetrievalPipeline:
    def __init__(self, model: Retriever, top_k: int = 100, bfs_max_depth: int = 2, writer_scores: bool = True,
                 validate: bool = True) -> None:
        self.model = model
        self.top_k = int(top_k)
        self.bfs_max_depth = int(bfs_max_depth)
        self.writer_scores = bool(writer_scores)
        self.validate = bool(validate)
        self._out_pinned: Optional[Dict[str, torch.Tensor]] = None
        self._copy_stream = None
        self._d2h_stream = None
        self._slots = None
        self._slot_free = None
        self._out_slots = None
        # optional {"bfs" | "topk": (torch.cuda.Event, torch.cuda.Event)} recorded around those stages (bench.py)
        self.stage_events: Optional[Dict[str, Any]] = None

    @torch.no_grad()
    def run_device(self, b: Any) -> Dict[str, torch.Tensor]:
        """Inputs already on the device. Returns device tensors."""
        dev = b.edge_index.device
        with torch.cuda.device(dev):  # the C ABI launches on the current device
            return self._run_device(b, dev)

    def _run_device(self, b: Any, dev) -> Dict[str, torch.Tensor]:
        gb = G.GraphBatch.from_edge_index(b.edge_index, b.ptr_host, device=dev, validate=False, ptr_dev=b.ptr)
        # jobs = one per question; seeds stay batch-global, the kernel subtracts the question's node offset
        dist = self._bfs(gb, b, b.q_local_indices)
        b._evi_graph = gb
        was = self.model.validate
        self.model.validate = False
        try:
            out = self.model(b)
        finally:
            self.model.validate = was
        self._mark("topk", 0)
        tk = segmented_topk(out.logits, gb.edge_ptr, self.top_k, sigmoid=self.writer_scores)
        res = writer_payload(tk, gb.edge_ptr, edge_index=b.edge_index, node_global_ids=b.node_global_ids,
                             edge_attr=b.edge_attr, labels=None, logits_fwd=out.logits_fwd,
                             logits_bwd=out.logits_bwd)
        self._mark("topk", 1)
        res["bfs_dist"] = dist
        # one validation word per batch: the segmentation's (cross-graph / out-of-range / unsorted edges) OR the scorer's;
        # the scorer's accumulator is cleared for the next batch, all in stream order (no host synchronisation)
        res["flags"] = torch.bitwise_or(self.model._flags, gb.flags) if gb.flags is not None else self.model._flags.clone()
        self.model._flags.zero_()
        return res

    def _mark(self, stage: str, which: int) -> None:
        ev = self.stage_events.get(stage) if self.stage_events else None
        if ev is not None:
            ev[which].record()

    def _bfs(self, gb, b, seeds_global):
        """Multi-source hop distance from each question's seeds, undirected, capped at ``bfs_max_depth``:
        edge-list sweeps for small caps (no CSR), the CSR frontier kernel when uncapped."""
        from . import lib as _lib

        L = _lib.load()
        dev = gb.head32.device
        B = b.num_graphs
        # every temporary passed by address stays referenced until the call returns: a tensor freed
        # between two argument evaluations can be handed back by the caching allocator for the next one
        ptr32 = b.ptr.to(torch.int32)
        node_end32 = ptr32[1:].contiguous()
        src_ptr32 = b.q_ptr.to(torch.int32).contiguous()
        seeds32 = seeds_global.to(torch.int32).contiguous()
        if seeds32.numel() == 0:
            seeds32 = torch.zeros(1, dtype=torch.int32, device=dev)
        dist_off = b.ptr.to(torch.long).contiguous()
        n_total = int(b.ptr_host[-1])
        dist = torch.empty(max(1, n_total), dtype=torch.int32, device=dev)
        if 0 <= self.bfs_max_depth <= G.SWEEP_MAX_DEPTH:
            edge_end = gb.edge_ptr[1:].contiguous()
            self._mark("bfs", 0)
            st = L.evi_bfs_edge_sweep(gb.head32.data_ptr(), gb.tail32.data_ptr(), B, gb.edge_ptr.data_ptr(),
                                      edge_end.data_ptr(), ptr32.data_ptr(), node_end32.data_ptr(), dist_off.data_ptr(),
                                      src_ptr32.data_ptr(), seeds32.data_ptr(), 1, 0, self.bfs_max_depth,
                                      dist.data_ptr(), _lib.stream_handle())
            _lib.check(st, "evi_bfs_edge_sweep")
            self._mark("bfs", 1)
            return dist[:n_total]
        row_ptr, adj = G.build_csr(gb.head32, gb.tail32, gb.num_nodes, G.CSR_UNDIRECTED)
        queue = torch.empty(max(1, 2 * n_total), dtype=torch.int32, device=dev)
        st = L.evi_bfs_multi_source(row_ptr.data_ptr(), adj.data_ptr(), B, ptr32.data_ptr(), node_end32.data_ptr(),
                                    dist_off.data_ptr(), src_ptr32.data_ptr(), seeds32.data_ptr(), 1,
                                    self.bfs_max_depth, dist.data_ptr(), queue.data_ptr(), _lib.stream_handle())
        _lib.check(st, "evi_bfs_multi_source")
        return dist[:n_total]

    @torch.no_grad()
    def run(self, hb: Any, device=None, post_device=None) -> Dict[str, torch.Tensor]:
        """Host batch in, host (pinned) results out: H2D of the ids, the hot path, D2H of the
        selected triples and BFS distances. Synchronises before returning."""
        dev = torch.device(device) if device is not None else self.model._device()
        b = device_batch(hb, dev)
        res = self.run_device(b)
        if post_device is not None:
            post_device(res)  # e.g. the NCCL gather of the selected triples
        keys = ("score", "index", "count", "head_id", "tail_id", "relation_id", "logit_fwd", "logit_bwd",
                "bfs_dist", "flags")
        out = {k: res[k].to("cpu", non_blocking=True) for k in keys}
        torch.cuda.current_stream(dev).synchronize()
        if self.validate:
            word = int(out["flags"].item())
            if word != 0:
                self.model.raise_for_flags(word)
        return out

    @torch.no_grad()
    def run_stream(self, host_batches, device=None, post_device=None):
        """Generator over ``run`` results for a sequence of host batches, software-pipelined: the H2D copy
        of batch i+1 (own stream, two reusable device slots) overlaps the kernels of batch i, and the host
        waits for the D2H of batch i only after batch i+1 has been queued. Same copies, same kernels and
        same outputs as calling ``run`` per batch."""
        dev = torch.device(device) if device is not None else self.model._device()
        compute = torch.cuda.current_stream(dev)
        if self._copy_stream is None or self._copy_stream.device != dev:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._d2h_stream = torch.cuda.Stream(device=dev)
            self._slots = [dict(), dict()]
            self._slot_free = [None, None]
            self._out_slots = [dict(), dict(), dict()]
        keys = ("score", "index", "count", "head_id", "tail_id", "relation_id", "logit_fwd", "logit_bwd",
                "bfs_dist", "flags")
        pending = None
        for i, hb in enumerate(host_batches):
            slot = self._slots[i & 1]
            with torch.cuda.stream(self._copy_stream):
                if self._slot_free[i & 1] is not None:  # the kernels that last read this slot have finished
                    self._copy_stream.wait_event(self._slot_free[i & 1])
                b = types.SimpleNamespace()
                for name in _BATCH_FIELDS:
                    src = getattr(hb, name)
                    buf = slot.get(name)
                    if buf is None or buf.numel() < src.numel() or buf.dtype != src.dtype:
                        buf = torch.empty(max(1, int(src.numel() * 1.25)), dtype=src.dtype, device=dev)
                        slot[name] = buf
                    view = buf[: src.numel()].view(src.shape)
                    view.copy_(src, non_blocking=True)
                    setattr(b, name, view)
                copied = torch.cuda.Event()
                copied.record(self._copy_stream)
            b.num_graphs, b.q_ptr_host, b.ptr_host = hb.num_graphs, hb.q_ptr_host, hb.ptr_host
            compute.wait_event(copied)
            res = self.run_device(b)
            if post_device is not None:
                post_device(res)
            # results land in one of three rotating sets of pinned host buffers (allocating pinned memory per
            # batch would call cudaHostAlloc, which synchronises the device); a yielded dict stays valid until
            # two more batches have been yielded
            oslot = self._out_slots[i % 3]
            out = {}
            kernels_done = torch.cuda.Event()
            kernels_done.record(compute)
            self._slot_free[i & 1] = kernels_done  # the input slot is free once this batch's kernels have run
            # the ten small device->host copies go to their own stream: the compute stream starts the next batch's
            # kernels instead of waiting behind the copy engine
            with torch.cuda.stream(self._d2h_stream):
                self._d2h_stream.wait_event(kernels_done)
                for k in keys:
                    src = res[k]
                    buf = oslot.get(k)
                    if buf is None or buf.numel() < src.numel() or buf.dtype != src.dtype:
                        buf = torch.empty(max(1, int(src.numel() * 1.25)), dtype=src.dtype, pin_memory=True)
                        oslot[k] = buf
                    view = buf[: src.numel()].view(src.shape)
                    src.record_stream(self._d2h_stream)  # the allocator must not hand this block out before the copy ran
                    view.copy_(src, non_blocking=True)
                    out[k] = view
                done = torch.cuda.Event()
                done.record(self._d2h_stream)
            if pending is not None:
                yield self._finish(pending)
            pending = (done, out)
        if pending is not None:
            yield self._finish(pending)

    def _finish(self, pending):
        """Wait for a batch's device->host copies; with ``validate`` raise for its validation word like ``run``."""
        done, out = pending
        done.synchronize()
        if self.validate:
            word = int(out["flags"].item())
            if word != 0:
                self.model.raise_for_flags(word)
        return out

    @staticmethod
    def d2h_bytes(res: Dict[str, torch.Tensor]) -> int:
        return int(sum(t.numel() * t.element_size() for t in res.values()))
