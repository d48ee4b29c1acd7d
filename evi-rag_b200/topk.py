"""Ragged per-question top-k evidence selection on the device.

One call replaces the reference's per-graph Python loops around ``torch.topk``
(``src/metrics/retriever_metrics.py:125-157,234-258``, ``src/metrics/reachability.py:129-179``,
``src/callbacks/retriever_topk_edge_writer.py:190-328``). Order inside a question: descending
score, ascending edge index on ties (the stable-argsort rule of
``src/data/components/g_agent_builder.py:640-652``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import torch

from . import lib as _lib

MAX_K = 1024


@dataclass
class TopK:
    values: torch.Tensor  # [B, k] f32, padded with -inf
    indices: torch.Tensor  # [B, k] i32 question-local edge index, padded with -1
    counts: torch.Tensor  # [B] i32 = min(E_q, k)


@_lib.device_guard
def segmented_topk(scores: torch.Tensor, edge_ptr: torch.Tensor, k: int, *, sigmoid: bool = False) -> TopK:
    """``segmented_topk(scores[E], edge_ptr[B+1], k)``. ``sigmoid=True`` ranks
    ``sigmoid(scores)`` like the writer (retriever_topk_edge_writer.py:203) and returns those."""
    if k <= 0 or k > MAX_K:
        raise ValueError(f"k must be in [1, {MAX_K}], got {k}")
    if scores.device.type != "cuda":
        raise RuntimeError("segmented_topk runs on CUDA tensors only (no CPU fallback).")
    L = _lib.load()
    dev = scores.device
    scores = scores.detach().to(torch.float32).contiguous().view(-1)
    edge_ptr = edge_ptr.to(device=dev, dtype=torch.long).contiguous()
    B = int(edge_ptr.numel() - 1)
    values = torch.empty((B, k), dtype=torch.float32, device=dev)
    indices = torch.empty((B, k), dtype=torch.int32, device=dev)
    counts = torch.empty(B, dtype=torch.int32, device=dev)
    st = L.evi_segmented_topk(scores.data_ptr(), edge_ptr.data_ptr(), B, int(k),
                              _lib.TOPK_SIGMOID if sigmoid else _lib.TOPK_RAW, values.data_ptr(),
                              indices.data_ptr(), counts.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_segmented_topk")
    return TopK(values, indices, counts)


@_lib.device_guard
def writer_payload(topk: TopK, edge_ptr: torch.Tensor, *, edge_index: torch.Tensor, node_global_ids: torch.Tensor,
                   edge_attr: torch.Tensor, labels: Optional[torch.Tensor] = None,
                   logits_fwd: Optional[torch.Tensor] = None, logits_bwd: Optional[torch.Tensor] = None
                   ) -> Dict[str, torch.Tensor]:
    """Fixed-shape ``[B, k]`` arrays of what ``_select_topk_edges`` gathers per graph
    (retriever_topk_edge_writer.py:303-310): head / relation / tail ids, label, fwd/bwd logits.
    Position j of question b has ``rank = j + 1`` (:334-349)."""
    L = _lib.load()
    dev = topk.indices.device
    B, k = topk.indices.shape
    ei = edge_index.to(device=dev, dtype=torch.long).contiguous()
    E = int(ei.size(1))
    ngid = node_global_ids.to(device=dev, dtype=torch.long).contiguous()
    ea = edge_attr.to(device=dev, dtype=torch.long).contiguous()
    f32 = lambda t: None if t is None else t.to(device=dev, dtype=torch.float32).contiguous().view(-1)
    labels, logits_fwd, logits_bwd = f32(labels), f32(logits_fwd), f32(logits_bwd)
    edge_ptr = edge_ptr.to(device=dev, dtype=torch.long).contiguous()
    out = {
        "head_id": torch.empty((B, k), dtype=torch.long, device=dev),
        "tail_id": torch.empty((B, k), dtype=torch.long, device=dev),
        "relation_id": torch.empty((B, k), dtype=torch.long, device=dev),
        "label": torch.empty((B, k), dtype=torch.float32, device=dev),
        "logit_fwd": torch.empty((B, k), dtype=torch.float32, device=dev),
        "logit_bwd": torch.empty((B, k), dtype=torch.float32, device=dev),
    }
    st = L.evi_topk_gather_payload(
        topk.indices.data_ptr(), topk.counts.data_ptr(), edge_ptr.data_ptr(), B, k, ei.data_ptr(), E,
        ngid.data_ptr(), ea.data_ptr(), _lib.ptr(labels), _lib.ptr(logits_fwd), _lib.ptr(logits_bwd),
        out["head_id"].data_ptr(), out["tail_id"].data_ptr(), out["relation_id"].data_ptr(),
        out["label"].data_ptr(), out["logit_fwd"].data_ptr(), out["logit_bwd"].data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_topk_gather_payload")
    out.update(score=topk.values, index=topk.indices, count=topk.counts)
    return out


@_lib.device_guard
def recall_at_k(topk: TopK, labels: torch.Tensor, edge_ptr: torch.Tensor, k_values: Sequence[int]) -> Dict[int, float]:
    """``EdgeRecallAtK`` (retriever_metrics.py:117-170) from one top-k at ``max(k_values)``:
    hits@k are prefix sums of the gathered labels; torch ops on ``[B, k]`` arrays only."""
    dev = topk.indices.device
    B, k = topk.indices.shape
    edge_ptr = edge_ptr.to(dev)
    labels = labels.to(device=dev, dtype=torch.float32).view(-1)
    gidx = (topk.indices.long() + edge_ptr[:-1].unsqueeze(1)).clamp(min=0, max=max(0, labels.numel() - 1))
    valid = torch.arange(k, device=dev).unsqueeze(0) < topk.counts.unsqueeze(1)
    top_labels = torch.where(valid, labels[gidx] if labels.numel() else torch.zeros_like(gidx, dtype=torch.float32),
                             torch.zeros((), device=dev))
    cum = torch.cumsum(top_labels, dim=1)
    pos = torch.zeros(B, dtype=torch.float32, device=dev)
    seg = torch.repeat_interleave(torch.arange(B, device=dev), (edge_ptr[1:] - edge_ptr[:-1]))
    pos.index_add_(0, seg, labels)
    has_edges = (edge_ptr[1:] - edge_ptr[:-1]) > 0
    n = int(has_edges.sum())
    out: Dict[int, float] = {}
    for kv in k_values:
        ke = torch.minimum(torch.full_like(topk.counts, int(kv)), topk.counts).long()
        hits = torch.where(ke > 0, cum.gather(1, (ke - 1).clamp(min=0).unsqueeze(1)).squeeze(1), torch.zeros((), device=dev))
        rec = hits / pos.clamp(min=1e-6)
        out[int(kv)] = float(rec[has_edges].sum() / n) if n else 0.0
    return out
