"""Retriever metrics from one ragged top-k per batch, on the device.

The reference computes these inside Python loops over graphs — one ``torch.topk`` plus several host
syncs per graph and per metric, and a CPU union-find per graph for reachability
(``src/metrics/retriever_metrics.py:117-170, 173-268, 331-395``, ``src/metrics/reachability.py:129-179,
329-381``). Here one ``segmented_topk`` and one ``evi_topk_metrics`` launch (a CTA per question) produce the
per-question numerators; the final means are a handful of torch reductions on ``[B]`` / ``[B, num_k]`` arrays.
Accumulator semantics follow the torchmetrics states of the reference (sums + counts, ``dist_reduce_fx="sum"``),
so multi-GPU totals are ``sharding.allreduce_sums`` of the returned ``sums`` / ``count``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import torch

from . import lib as _lib
from .topk import MAX_K, TopK, segmented_topk

MAX_K_VALUES = 16


def normalize_k_values(raw_values) -> list:
    """``normalize_k_values`` (src/utils/metrics.py:25-40): positive ints, deduplicated, ascending."""
    out, seen = [], set()
    for item in raw_values or []:
        try:
            k = int(item)
        except (TypeError, ValueError):
            continue
        if k <= 0 or k in seen:
            continue
        out.append(k)
        seen.add(k)
    return sorted(out)


@dataclass
class TopKMetricArrays:
    k_values: list
    hits: Optional[torch.Tensor] = None  # [B, num_k] f32
    pos: Optional[torch.Tensor] = None  # [B] f32
    reach: Optional[torch.Tensor] = None  # [B, num_k] u8
    reach_valid: Optional[torch.Tensor] = None  # [B] u8
    margin: Optional[torch.Tensor] = None  # [B] f32
    margin_valid: Optional[torch.Tensor] = None  # [B] u8


@_lib.device_guard
def topk_metric_arrays(topk: TopK, edge_ptr: torch.Tensor, k_values: Sequence[int], *,
                       labels: Optional[torch.Tensor] = None, scores: Optional[torch.Tensor] = None,
                       head32: Optional[torch.Tensor] = None, tail32: Optional[torch.Tensor] = None,
                       node_ptr: Optional[torch.Tensor] = None, q_local_indices: Optional[torch.Tensor] = None,
                       q_ptr: Optional[torch.Tensor] = None, a_local_indices: Optional[torch.Tensor] = None,
                       a_ptr: Optional[torch.Tensor] = None, num_nodes: Optional[int] = None) -> TopKMetricArrays:
    """Per-question numerators of the retriever metrics (one kernel launch). ``labels`` enables hits / pos,
    ``labels`` + ``scores`` the score margin, the graph arguments reachability."""
    ks = normalize_k_values(k_values)
    if not ks or len(ks) > MAX_K_VALUES:
        raise ValueError(f"k_values must hold 1..{MAX_K_VALUES} positive ints, got {list(k_values)!r}")
    L = _lib.load()
    dev = topk.indices.device
    B, k = topk.indices.shape
    nk = len(ks)
    edge_ptr = edge_ptr.to(device=dev, dtype=torch.long).contiguous()
    f32 = lambda t: None if t is None else t.to(device=dev, dtype=torch.float32).contiguous().view(-1)
    i64 = lambda t: None if t is None else torch.as_tensor(t).to(device=dev, dtype=torch.long).contiguous().view(-1)
    labels, scores = f32(labels), f32(scores)
    out = TopKMetricArrays(k_values=ks)
    if labels is not None:
        out.hits = torch.empty((B, nk), dtype=torch.float32, device=dev)
        out.pos = torch.empty(B, dtype=torch.float32, device=dev)
        if scores is not None:
            out.margin = torch.empty(B, dtype=torch.float32, device=dev)
            out.margin_valid = torch.empty(B, dtype=torch.uint8, device=dev)
    want_reach = head32 is not None
    node_ptr, seeds, seed_ptr, answers, answer_ptr = i64(node_ptr), i64(q_local_indices), i64(q_ptr), \
        i64(a_local_indices), i64(a_ptr)
    scratch = None
    if want_reach:
        if any(t is None for t in (tail32, node_ptr, seeds, seed_ptr, answers, answer_ptr)):
            raise ValueError("reachability needs tail32, node_ptr, q_local_indices/q_ptr and a_local_indices/a_ptr")
        out.reach = torch.empty((B, nk), dtype=torch.uint8, device=dev)
        out.reach_valid = torch.empty(B, dtype=torch.uint8, device=dev)
        if seeds.numel() == 0:
            seeds = torch.zeros(1, dtype=torch.long, device=dev)
        if answers.numel() == 0:
            answers = torch.zeros(1, dtype=torch.long, device=dev)
        n_total = int(num_nodes) if num_nodes is not None else int(node_ptr[-1].item())
        scratch = torch.empty(max(1, n_total), dtype=torch.int32, device=dev)
    ks_arr = (C.c_int * nk)(*ks)
    st = L.evi_topk_metrics(
        topk.indices.data_ptr(), topk.counts.data_ptr(), edge_ptr.data_ptr(), B, k, ks_arr, nk,
        _lib.ptr(labels), _lib.ptr(scores if out.margin is not None else None),
        _lib.ptr(head32 if want_reach else None), _lib.ptr(tail32 if want_reach else None),
        _lib.ptr(node_ptr if want_reach else None), _lib.ptr(seeds if want_reach else None),
        _lib.ptr(seed_ptr if want_reach else None), _lib.ptr(answers if want_reach else None),
        _lib.ptr(answer_ptr if want_reach else None), _lib.ptr(scratch), _lib.ptr(out.hits), _lib.ptr(out.pos),
        _lib.ptr(out.reach), _lib.ptr(out.reach_valid), _lib.ptr(out.margin), _lib.ptr(out.margin_valid),
        _lib.stream_handle())
    _lib.check(st, "evi_topk_metrics")
    return out


def _has_edges(edge_ptr: torch.Tensor) -> torch.Tensor:
    return (edge_ptr[1:] - edge_ptr[:-1]) > 0


def edge_recall_sums(arr: TopKMetricArrays, edge_ptr: torch.Tensor, *, require_positive: bool = False):
    """Accumulators of ``EdgeRecallAtK`` (``require_positive=False``, retriever_metrics.py:134-157: every graph
    with an edge counts, recall = hits / max(pos, 1)) or ``BridgeEdgeRecallAtK`` (``True``, :234-258: graphs
    without a positive are skipped). Returns ``(sums [num_k], count)`` as device tensors."""
    edge_ptr = edge_ptr.to(arr.hits.device)
    keep = _has_edges(edge_ptr)
    if require_positive:
        keep = keep & (arr.pos > 0)
    rec = arr.hits / arr.pos.clamp(min=1.0).unsqueeze(1)
    sums = (rec * keep.unsqueeze(1)).sum(dim=0)
    return sums, keep.sum().to(torch.float32)


@_lib.device_guard
def retriever_metrics(logits: torch.Tensor, labels: torch.Tensor, edge_ptr: torch.Tensor, k_values: Sequence[int], *,
                      graph=None, node_ptr=None, q_local_indices=None, q_ptr=None, a_local_indices=None, a_ptr=None,
                      bridge: bool = True) -> Dict[str, float]:
    """The eval metrics of ``RetrieverModule._update_metrics`` (retriever_module.py:157-176) for one batch:
    ``edge/recall@k``, ``answer/reachability@k``, ``edge/score_margin`` and (``bridge``) ``bridge/recall@k``,
    from raw logits like the reference. ``graph`` is a ``graph.GraphBatch`` of the same batch."""
    ks = normalize_k_values(k_values)
    dev = logits.device
    edge_ptr = edge_ptr.to(device=dev, dtype=torch.long)
    if max(ks) > MAX_K:
        raise ValueError(f"k up to {MAX_K} is supported by the selection kernel (one CTA sorts a question's candidates "
                         f"in shared memory); got max k = {max(ks)}. The reference's configs stop at k = 500.")
    tk = segmented_topk(logits, edge_ptr, max(ks))
    gargs = {}
    if graph is not None:
        gargs = dict(head32=graph.head32, tail32=graph.tail32, node_ptr=node_ptr, q_local_indices=q_local_indices,
                     q_ptr=q_ptr, a_local_indices=a_local_indices, a_ptr=a_ptr, num_nodes=int(graph.ptr_host[-1]))
    arr = topk_metric_arrays(tk, edge_ptr, ks, labels=labels, scores=logits, **gargs)
    out: Dict[str, float] = {}
    sums, cnt = edge_recall_sums(arr, edge_ptr)
    for i, k in enumerate(ks):
        out[f"edge/recall@{k}"] = float(sums[i] / cnt.clamp(min=1.0))
    mv = arr.margin_valid.bool()
    out["edge/score_margin"] = float((arr.margin * mv).sum() / mv.sum().clamp(min=1).to(torch.float32))
    if arr.reach is not None:
        valid = arr.reach_valid.bool()
        denom = valid.sum().clamp(min=1).to(torch.float32)
        hits = (arr.reach.bool() & valid.unsqueeze(1)).sum(dim=0).to(torch.float32)
        for i, k in enumerate(ks):
            out[f"answer/reachability@{k}"] = float(hits[i] / denom)
    if bridge and graph is not None:
        # bridge edges touch neither a seed nor an answer node (compute_qa_edge_mask, graph_utils.py:107-153);
        # the reference drops the others before the per-graph top-k (retriever_metrics.py:197-204)
        n_total = int(graph.ptr_host[-1])
        node_mask = torch.zeros(max(1, n_total), dtype=torch.bool, device=dev)
        qa = torch.cat([torch.as_tensor(q_local_indices).view(-1), torch.as_tensor(a_local_indices).view(-1)]).to(dev)
        if qa.numel():
            node_mask[qa.long()] = True
        bmask = ~(node_mask[graph.head32.long()] | node_mask[graph.tail32.long()])
        b_counts = torch.zeros(edge_ptr.numel() - 1, dtype=torch.long, device=dev)
        b_counts.index_add_(0, graph.edge_batch[bmask], torch.ones(int(bmask.sum()), dtype=torch.long, device=dev))
        b_ptr = torch.zeros_like(edge_ptr)
        b_ptr[1:] = torch.cumsum(b_counts, 0)
        b_scores = logits[bmask].contiguous()
        b_labels = labels.to(dev)[bmask].contiguous()
        if b_scores.numel():
            btk = segmented_topk(b_scores, b_ptr, min(max(ks), 1024))
            barr = topk_metric_arrays(btk, b_ptr, ks, labels=b_labels)
            bsums, bcnt = edge_recall_sums(barr, b_ptr, require_positive=True)
        else:
            bsums, bcnt = torch.zeros(len(ks), device=dev), torch.zeros((), device=dev)
        for i, k in enumerate(ks):
            out[f"bridge/recall@{k}"] = float(bsums[i] / bcnt.clamp(min=1.0))
    return out


__all__ = ["normalize_k_values", "topk_metric_arrays", "edge_recall_sums", "retriever_metrics", "TopKMetricArrays"]
