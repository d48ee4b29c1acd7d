"""Multi-GPU: questions are independent units, so the batch shards across ranks with no data-path
collective (no edge crosses graphs, src/utils/graph_utils.py:78-88). The only exchange is the
gather of the selected triples and the metric sums — the reference does it with
``all_gather_object`` pickles (src/callbacks/retriever_topk_edge_writer.py:449-462) and
torchmetrics ``dist_reduce_fx="sum"`` (src/metrics/retriever_metrics.py:90-91); here it is one
all-gather of fixed-shape ``[Q, k]`` arrays and one all-reduce of a small vector.

Works with any ``torch.distributed`` backend (NCCL over NVLink on the GPU box, gloo in CPU tests).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist


def partition_by_edges(edge_counts: Sequence[int], world_size: int) -> List[Tuple[int, int]]:
    """Contiguous split of the question list so every rank gets ~sum(E_q)/world edges
    (prefix sum over E_q). Returns ``[(q_begin, q_end)] * world_size``."""
    counts = np.asarray(edge_counts, dtype=np.int64)
    n = int(counts.shape[0])
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    prefix = np.concatenate([[0], np.cumsum(counts)])
    total = int(prefix[-1])
    bounds = [0]
    for r in range(1, world_size):
        target = total * r / world_size
        q = int(np.searchsorted(prefix, target, side="left"))
        # pick the closer of the two neighbouring cut points, keep cuts monotone
        if q > 0 and abs(prefix[q - 1] - target) <= abs(prefix[min(q, n)] - target):
            q -= 1
        bounds.append(min(max(q, bounds[-1]), n))
    bounds.append(n)
    return [(bounds[r], bounds[r + 1]) for r in range(world_size)]


def gather_topk(local: Dict[str, torch.Tensor], num_local_questions: int, group=None,
                counts: Optional[Sequence[int]] = None) -> Dict[str, torch.Tensor]:
    """All-gather per-rank ``[Q_r, k]`` (or ``[Q_r]``) arrays into global question order with ONE collective:
    the arrays are packed row-wise into a single ``[Q_max, row_bytes]`` byte matrix, exchanged with
    ``all_gather_into_tensor`` and unpacked. Ranks may hold different Q_r; when the caller already knows every
    rank's question count (``counts``, e.g. from ``partition_by_edges``) no size exchange and no host
    synchronisation happen, otherwise one small all-gather of the counts precedes the exchange."""
    world = dist.get_world_size(group)
    dev = next(iter(local.values())).device
    if counts is None:
        q_local = torch.tensor([num_local_questions], dtype=torch.long, device=dev)
        q_all_t = [torch.zeros_like(q_local) for _ in range(world)]
        dist.all_gather(q_all_t, q_local, group=group)
        q_all = [int(t.item()) for t in q_all_t]
    else:
        q_all = [int(c) for c in counts]
        if len(q_all) != world or q_all[dist.get_rank(group)] != num_local_questions:
            raise ValueError(f"counts {q_all} do not match world size {world} / local count {num_local_questions}")
    q_max = max(q_all)
    names = list(local)
    rows, widths = [], []
    for name in names:
        t = local[name].contiguous()
        if t.shape[0] != num_local_questions:
            raise ValueError(f"{name}: leading dimension {t.shape[0]} != {num_local_questions} questions")
        width = int(np.prod(t.shape[1:])) if t.dim() > 1 else 1  # explicit: reshape(0, -1) is ambiguous for 0 rows
        b = t.reshape(num_local_questions, width).view(torch.uint8)
        rows.append(b)
        widths.append(int(b.shape[1]))
    packed = torch.zeros((q_max, sum(widths)), dtype=torch.uint8, device=dev)
    if num_local_questions:
        torch.cat(rows, dim=1, out=packed[:num_local_questions])
    gathered = torch.empty((world * q_max, sum(widths)), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(gathered, packed, group=group)
    if any(q != q_max for q in q_all):
        gathered = torch.cat([gathered[r * q_max: r * q_max + q] for r, q in enumerate(q_all)], dim=0)
    out: Dict[str, torch.Tensor] = {}
    col = 0
    total = sum(q_all)
    for name, wdt in zip(names, widths):
        t = local[name]
        out[name] = gathered[:, col: col + wdt].contiguous().view(t.dtype).reshape((total,) + tuple(t.shape[1:]))
        col += wdt
    return out


def allreduce_sums(values: torch.Tensor, group=None) -> torch.Tensor:
    """Sum a small vector of metric accumulators over ranks (``dist_reduce_fx='sum'``)."""
    out = values.clone()
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out
