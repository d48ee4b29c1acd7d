"""Graph side of the hot path: edge segmentation, DDE features, BFS hop distances and
shortest-path edge labels, all computed by the sm_100a kernels in ``csrc/graph_kernels.cu``.

Two levels:
  * batched device API (``GraphBatch`` ...): what the pipeline should call — one launch covers
    every question of a batch;
  * single-graph functions with the reference's names and argument meaning
    (``scripts/build_retrieval_pipeline.py``: ``_bfs_dist`` :610-631,
    ``_shortest_path_union_mask_by_pair`` :691-752, ``..._directed`` :755-815,
    ``has_connectivity`` :946-979; ``src/utils/graph_utils.py``: ``compute_edge_batch`` :49-104)
    so call sites in ``build_graph`` can switch one import.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import lib as _lib

_PATH_MODES = ("undirected", "qa_directed")
CSR_UNDIRECTED, CSR_FORWARD, CSR_REVERSE = 0, 1, 2
SWEEP_MAX_DEPTH = 8  # hop caps up to this use the edge-list sweep kernel (O(depth * E)); beyond it, CSR frontiers


def _dev(device=None) -> torch.device:
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise RuntimeError("evi_rag_b200 graph kernels need a CUDA device (no CPU fallback).")
    return torch.device("cuda", torch.cuda.current_device())


# ------------------------------------------------------------------------------------------
# a1: compute_edge_batch
# ------------------------------------------------------------------------------------------
def _raise_edge_flags(word: int) -> None:
    if word & _lib.FLAG_OUT_OF_RANGE:
        raise ValueError("edge_batch contains out-of-range indices; edge_index outside [0, ptr[-1]).")
    if word & _lib.FLAG_CROSS_GRAPH:
        raise ValueError("edge_index crosses graph boundaries; head/tail graph assignments differ. ")
    if word & _lib.FLAG_NOT_MONOTONE:
        raise ValueError(
            "edge_batch is not non-decreasing along the flattened edge list, which breaks per-graph slicing; "
            "Ensure edges are concatenated per-graph (PyG Batch)."
        )


@_lib.device_guard
def compute_edge_batch(
    edge_index: torch.Tensor,
    *,
    node_ptr: torch.Tensor,
    num_graphs: int,
    device: Optional[torch.device] = None,
    debug_batch: object | None = None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same contract as ``src/utils/graph_utils.py:49-104``: returns ``(edge_batch, edge_ptr)``
    int64 on ``device`` and raises ``ValueError`` for out-of-range / cross-graph / unsorted edges."""
    if edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError(f"edge_index must have shape [2, E], got {tuple(edge_index.shape)}")
    if node_ptr.numel() != num_graphs + 1:
        raise ValueError(f"node_ptr length mismatch: got {node_ptr.numel()} expected {num_graphs + 1}")
    dev = _dev(device if device is not None and torch.device(device).type == "cuda" else None)
    L = _lib.load()
    ei = edge_index.to(device=dev, dtype=torch.long).contiguous()
    ptr = node_ptr.to(device=dev, dtype=torch.long).contiguous()
    E = int(ei.size(1))
    eb = torch.empty(E, dtype=torch.long, device=dev)
    ep = torch.empty(num_graphs + 1, dtype=torch.long, device=dev)
    flags = torch.zeros(1, dtype=torch.int32, device=dev)
    st = L.evi_edge_batch(ei.data_ptr(), E, ptr.data_ptr(), int(num_graphs), eb.data_ptr(), None, None, None,
                          ep.data_ptr(), flags.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_edge_batch")
    _raise_edge_flags(int(flags.item()))
    return eb, ep


# ------------------------------------------------------------------------------------------
# batched device API
# ------------------------------------------------------------------------------------------
@dataclass
class GraphBatch:
    """A ragged batch of question subgraphs on the device (batch-global int32 node ids)."""

    head32: torch.Tensor
    tail32: torch.Tensor
    qid32: torch.Tensor
    edge_batch: torch.Tensor  # int64 [E]
    edge_ptr: torch.Tensor  # int64 [B+1] device
    ptr_host: np.ndarray  # int64 [B+1]
    edge_ptr_host: Optional[np.ndarray] = None
    flags: Optional[torch.Tensor] = None  # int32 [1]: the validation word of evi_edge_batch (EVI_FLAG_* bits)

    @property
    def num_nodes(self) -> int:
        return int(self.ptr_host[-1])

    @property
    def num_edges(self) -> int:
        return int(self.head32.numel())

    @property
    def num_graphs(self) -> int:
        return int(self.ptr_host.shape[0] - 1)

    @classmethod
    def from_edge_index(cls, edge_index, ptr, device=None, validate: bool = True, ptr_dev=None) -> "GraphBatch":
        """``ptr`` is the host copy of the node offsets; ``ptr_dev`` an optional device copy (saves a
        synchronous pageable H2D when the caller already has one)."""
        dev = _dev(device)
        with torch.cuda.device(dev):  # the C ABI launches on the current device
            return cls._from_edge_index(edge_index, ptr, dev, validate, ptr_dev)

    @classmethod
    def _from_edge_index(cls, edge_index, ptr, dev, validate, ptr_dev) -> "GraphBatch":
        L = _lib.load()
        ei = torch.as_tensor(edge_index).to(device=dev, dtype=torch.long).contiguous()
        ptr_host = np.asarray(torch.as_tensor(ptr).cpu().numpy(), dtype=np.int64)
        if ptr_dev is None:
            ptr_dev = torch.from_numpy(ptr_host).to(dev)
        ptr_dev = ptr_dev.to(device=dev, dtype=torch.long).contiguous()
        E, B = int(ei.size(1)), int(ptr_host.shape[0] - 1)
        head32 = torch.empty(E, dtype=torch.int32, device=dev)
        tail32 = torch.empty(E, dtype=torch.int32, device=dev)
        qid32 = torch.empty(E, dtype=torch.int32, device=dev)
        eb = torch.empty(E, dtype=torch.long, device=dev)
        ep = torch.empty(B + 1, dtype=torch.long, device=dev)
        flags = torch.zeros(1, dtype=torch.int32, device=dev)
        st = L.evi_edge_batch(ei.data_ptr(), E, ptr_dev.data_ptr(), B, eb.data_ptr(), qid32.data_ptr(),
                              head32.data_ptr(), tail32.data_ptr(), ep.data_ptr(), flags.data_ptr(),
                              _lib.stream_handle())
        _lib.check(st, "evi_edge_batch")
        if validate:
            _raise_edge_flags(int(flags.item()))
        return cls(head32, tail32, qid32, eb, ep, ptr_host, flags=flags)

    def edge_ptr_np(self) -> np.ndarray:
        if self.edge_ptr_host is None:
            self.edge_ptr_host = self.edge_ptr.cpu().numpy()
        return self.edge_ptr_host


@_lib.device_guard
def dde_node_features(gb: GraphBatch, topic_one_hot: torch.Tensor, num_rounds: int = 2,
                      num_reverse_rounds: int = 2) -> torch.Tensor:
    """``Retriever._build_node_structure_features`` (retriever.py:519-553) on the device."""
    L = _lib.load()
    dev = gb.head32.device
    topic = topic_one_hot.to(device=dev, dtype=torch.float32).contiguous()
    N = gb.num_nodes
    ns = torch.empty((N, 2 * (1 + num_rounds + num_reverse_rounds)), dtype=torch.float32, device=dev)
    ws = torch.empty(int(L.evi_dde_workspace_bytes(N)), dtype=torch.uint8, device=dev)
    st = L.evi_dde_forward(gb.head32.data_ptr(), gb.tail32.data_ptr(), gb.num_edges, N, topic.data_ptr(),
                           topic.size(1), num_rounds, num_reverse_rounds, ns.data_ptr(), ws.data_ptr(), ws.numel(),
                           _lib.stream_handle())
    _lib.check(st, "evi_dde_forward")
    return ns


@_lib.device_guard
def build_csr(head32: torch.Tensor, tail32: torch.Tensor, num_nodes: int, mode: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """CSR adjacency over batch-global ids (build_retrieval_pipeline.py:570-603)."""
    L = _lib.load()
    dev = head32.device
    E = int(head32.numel())
    row_ptr = torch.empty(num_nodes + 1, dtype=torch.int32, device=dev)
    adj = torch.empty(max(1, (2 if mode == CSR_UNDIRECTED else 1) * E), dtype=torch.int32, device=dev)
    scratch = torch.empty(max(1, num_nodes), dtype=torch.int32, device=dev)
    st = L.evi_build_csr(head32.data_ptr(), tail32.data_ptr(), E, num_nodes, mode, row_ptr.data_ptr(),
                         adj.data_ptr(), scratch.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_build_csr")
    return row_ptr, adj


@_lib.device_guard
def bfs_jobs(row_ptr: torch.Tensor, adj: torch.Tensor, job_node_begin: np.ndarray, job_node_end: np.ndarray,
             job_src_ptr: np.ndarray, sources_local: np.ndarray, max_depth: int = -1,
             dist: Optional[torch.Tensor] = None, dist_off: Optional[np.ndarray] = None):
    """Run one BFS per job (CTA per job). Returns ``(dist_flat int32 device, job_dist_off int64 host)``.
    ``dist`` / ``dist_off`` let several launches (different CSR flavours) fill one buffer."""
    L = _lib.load()
    dev = row_ptr.device
    nb = np.asarray(job_node_begin, dtype=np.int32)
    ne = np.asarray(job_node_end, dtype=np.int32)
    if dist_off is None:
        sizes = (ne - nb).astype(np.int64)
        dist_off = np.zeros(len(nb) + 1, dtype=np.int64)
        np.cumsum(sizes, out=dist_off[1:])
        total = int(dist_off[-1])
        dist = torch.empty(max(1, total), dtype=torch.int32, device=dev)
        job_off = dist_off[:-1]
    else:
        job_off = np.asarray(dist_off, dtype=np.int64)
        total = int(dist.numel())
    queue = torch.empty(max(1, 2 * total), dtype=torch.int32, device=dev)
    up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    d_nb, d_ne, d_off = up(nb, np.int32), up(ne, np.int32), up(job_off, np.int64)
    d_sp, d_src = up(job_src_ptr, np.int32), up(sources_local if len(sources_local) else np.zeros(1), np.int32)
    st = L.evi_bfs_multi_source(row_ptr.data_ptr(), adj.data_ptr(), len(nb), d_nb.data_ptr(), d_ne.data_ptr(),
                                d_off.data_ptr(), d_sp.data_ptr(), d_src.data_ptr(), 0, int(max_depth), dist.data_ptr(),
                                queue.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_bfs_multi_source")
    return dist[:total], dist_off


@_lib.device_guard
def batched_bfs(gb: GraphBatch, sources: np.ndarray, src_ptr: np.ndarray, *, directed: bool = False,
                max_depth: int = -1) -> torch.Tensor:
    """Multi-source hop distance for every question of the batch: ``sources`` are batch-global
    node ids grouped by ``src_ptr [B+1]``. Returns ``dist [N] int32`` (-1 unreachable)."""
    ptr = gb.ptr_host
    B = gb.num_graphs
    if 0 <= max_depth <= SWEEP_MAX_DEPTH:
        return _sweep_bfs(gb, sources, src_ptr, directed=directed, max_depth=max_depth)
    row_ptr, adj = build_csr(gb.head32, gb.tail32, gb.num_nodes, CSR_FORWARD if directed else CSR_UNDIRECTED)
    q_of_src = np.repeat(np.arange(B), np.diff(src_ptr))
    local = np.asarray(sources, dtype=np.int64) - ptr[:-1][q_of_src]
    dist, _ = bfs_jobs(row_ptr, adj, ptr[:-1], ptr[1:], src_ptr, local, max_depth)
    return dist


@_lib.device_guard
def _sweep_bfs(gb: GraphBatch, sources: np.ndarray, src_ptr: np.ndarray, *, directed: bool, max_depth: int) -> torch.Tensor:
    """Depth-capped multi-source BFS straight from the edge list (``evi_bfs_edge_sweep``)."""
    L = _lib.load()
    dev = gb.head32.device
    up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    ptr = gb.ptr_host
    n_total = int(ptr[-1])
    d_nb, d_ne, d_off = up(ptr[:-1], np.int32), up(ptr[1:], np.int32), up(ptr[:-1], np.int64)
    d_sp = up(src_ptr, np.int32)
    d_src = up(sources if len(sources) else np.zeros(1), np.int32)
    e_begin = gb.edge_ptr[:-1].contiguous()
    e_end = gb.edge_ptr[1:].contiguous()
    dist = torch.empty(max(1, n_total), dtype=torch.int32, device=dev)
    st = L.evi_bfs_edge_sweep(gb.head32.data_ptr(), gb.tail32.data_ptr(), gb.num_graphs, e_begin.data_ptr(),
                              e_end.data_ptr(), d_nb.data_ptr(), d_ne.data_ptr(), d_off.data_ptr(), d_sp.data_ptr(),
                              d_src.data_ptr(), 1, int(directed), int(max_depth), dist.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_bfs_edge_sweep")
    return dist[:n_total]


@_lib.device_guard
def batched_shortest_path_labels(gb: GraphBatch, seeds: np.ndarray, seed_ptr: np.ndarray, answers: np.ndarray,
                                 answer_ptr: np.ndarray, *, path_mode: str = "undirected",
                                 return_pairs: bool = False):
    """Per question, the union over (seed, answer) pairs of the edges on a shortest path
    (build_retrieval_pipeline.py:691-830). Returns the positive mask ``[E] uint8`` (device) and,
    with ``return_pairs``, the per-pair arrays needed to rebuild the reference's pair lists."""
    if path_mode not in _PATH_MODES:
        raise ValueError(f"Unsupported path_mode: {path_mode}. Expected one of {_PATH_MODES}.")
    directed = path_mode == "qa_directed"
    L = _lib.load()
    dev = gb.head32.device
    ptr, eptr = gb.ptr_host, gb.edge_ptr_np()
    B = gb.num_graphs
    j_nb, j_ne, j_src, j_rev = [], [], [], []
    pairs: List[Tuple[int, int, int, int, int]] = []  # (graph, job_s, job_a, s_local, a_local)
    for g in range(B):
        n = int(ptr[g + 1] - ptr[g])
        s_loc = sorted({int(s) - int(ptr[g]) for s in seeds[seed_ptr[g]:seed_ptr[g + 1]] if 0 <= int(s) - int(ptr[g]) < n})
        a_loc = sorted({int(a) - int(ptr[g]) for a in answers[answer_ptr[g]:answer_ptr[g + 1]] if 0 <= int(a) - int(ptr[g]) < n})
        if not s_loc or not a_loc or eptr[g + 1] == eptr[g]:
            continue
        s_jobs, a_jobs = [], []
        for s in s_loc:
            s_jobs.append(len(j_nb)); j_nb.append(ptr[g]); j_ne.append(ptr[g + 1]); j_src.append(s); j_rev.append(False)
        for a in a_loc:
            a_jobs.append(len(j_nb)); j_nb.append(ptr[g]); j_ne.append(ptr[g + 1]); j_src.append(a); j_rev.append(True)
        for si, s in enumerate(s_loc):
            for ai, a in enumerate(a_loc):
                pairs.append((g, s_jobs[si], a_jobs[ai], s, a))
    mask = torch.zeros(max(1, gb.num_edges), dtype=torch.uint8, device=dev)
    if not pairs:
        return (mask[: gb.num_edges], None) if return_pairs else mask[: gb.num_edges]
    j_nb, j_ne = np.asarray(j_nb, np.int64), np.asarray(j_ne, np.int64)
    j_src, j_rev = np.asarray(j_src, np.int64), np.asarray(j_rev, bool)
    sizes = j_ne - j_nb
    off = np.zeros(len(j_nb) + 1, np.int64)
    np.cumsum(sizes, out=off[1:])
    dist = torch.empty(int(off[-1]), dtype=torch.int32, device=dev)
    # one launch per CSR flavour; undirected uses the same adjacency for both job kinds
    groups = [(CSR_UNDIRECTED, np.ones(len(j_nb), bool))] if not directed else \
        [(CSR_FORWARD, ~j_rev), (CSR_REVERSE, j_rev)]
    for mode, sel in groups:
        idx = np.nonzero(sel)[0]
        if idx.size == 0:
            continue
        row_ptr, adj = build_csr(gb.head32, gb.tail32, gb.num_nodes, mode)
        bfs_jobs(row_ptr, adj, j_nb[idx], j_ne[idx], np.arange(idx.size + 1), j_src[idx], -1,
                 dist=dist, dist_off=off[:-1][idx])
    P = len(pairs)
    pg = np.asarray([p[0] for p in pairs])
    up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    e_cnt = (eptr[pg + 1] - eptr[pg]).astype(np.int64)
    out_off = np.zeros(P + 1, np.int64)
    np.cumsum(e_cnt, out=out_off[1:])
    member = torch.empty(max(1, int(out_off[-1])), dtype=torch.uint8, device=dev)
    p_cnt = torch.empty(P, dtype=torch.int32, device=dev)
    p_len = torch.empty(P, dtype=torch.int32, device=dev)
    args = [up(eptr[pg], np.int32), up(eptr[pg + 1], np.int32), up(ptr[pg], np.int32),
            up(ptr[pg + 1] - ptr[pg], np.int32), up(off[[p[1] for p in pairs]], np.int64),
            up(off[[p[2] for p in pairs]], np.int64), up([p[4] for p in pairs], np.int32), up(out_off[:-1], np.int64)]
    st = L.evi_shortest_path_pairs(gb.head32.data_ptr(), gb.tail32.data_ptr(), P, *[a.data_ptr() for a in args],
                                   dist.data_ptr(), int(directed), mask.data_ptr(), member.data_ptr(),
                                   p_cnt.data_ptr(), p_len.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_shortest_path_pairs")
    mask = mask[: gb.num_edges]
    if not return_pairs:
        return mask
    info = {"pairs": pairs, "member": member, "out_off": out_off, "count": p_cnt.cpu().numpy(),
            "length": p_len.cpu().numpy()}
    return mask, info


# ------------------------------------------------------------------------------------------
# single-graph functions with the reference's signatures
# ------------------------------------------------------------------------------------------
def _single_graph_batch(num_nodes: int, edge_src: Sequence[int], edge_dst: Sequence[int]):
    """Device arrays for one graph. Out-of-range endpoints are kept (the kernels drop them like
    build_retrieval_pipeline.py:579-580), so edge ids stay aligned with the caller's lists."""
    dev = _dev()
    src = torch.as_tensor(np.asarray(edge_src, dtype=np.int64).clip(-1, 2**31 - 2), dtype=torch.int32, device=dev)
    dst = torch.as_tensor(np.asarray(edge_dst, dtype=np.int64).clip(-1, 2**31 - 2), dtype=torch.int32, device=dev)
    return src, dst


def bfs_dist_edges(num_nodes: int, edge_src: Sequence[int], edge_dst: Sequence[int], sources: Sequence[int], *,
                   directed: bool = False, max_depth: int = -1) -> List[int]:
    """``_bfs_dist(num_nodes, _build_(un)directed_adjacency(num_nodes, src, dst), sources)``."""
    if num_nodes <= 0:
        return []
    src, dst = _single_graph_batch(num_nodes, edge_src, edge_dst)
    row_ptr, adj = build_csr(src, dst, num_nodes, CSR_FORWARD if directed else CSR_UNDIRECTED)
    srcs = np.asarray([int(s) for s in sources], dtype=np.int64).clip(-1, 2**31 - 2)
    dist, _ = bfs_jobs(row_ptr, adj, [0], [num_nodes], [0, len(srcs)], srcs, max_depth)
    return dist.cpu().tolist()


def _bfs_dist(num_nodes: int, adjacency: Sequence[Sequence[int]], sources: Sequence[int]) -> List[int]:
    """Exact reference signature (adjacency lists in, python list out; :610-631). The lists are
    flattened to CSR on the host; prefer ``bfs_dist_edges`` / ``batched_bfs`` in new code."""
    if num_nodes <= 0:
        return [-1] * max(0, num_nodes)
    dev = _dev()
    deg = np.fromiter((len(a) for a in adjacency), dtype=np.int64, count=num_nodes)
    row = np.zeros(num_nodes + 1, dtype=np.int32)
    np.cumsum(deg, out=row[1:])
    flat = np.fromiter((v for a in adjacency for v in a), dtype=np.int32, count=int(row[-1]))
    row_ptr = torch.from_numpy(row).to(dev)
    adj = torch.from_numpy(flat if flat.size else np.zeros(1, np.int32)).to(dev)
    srcs = np.asarray([int(s) for s in sources], dtype=np.int64).clip(-1, 2**31 - 2)
    dist, _ = bfs_jobs(row_ptr, adj, [0], [num_nodes], [0, len(srcs)], srcs, -1)
    return dist.cpu().tolist()


def _shortest_path_union_mask_by_pair_mode(num_nodes: int, edge_src: Sequence[int], edge_dst: Sequence[int],
                                           sources: Sequence[int], targets: Sequence[int], *, path_mode: str):
    """Reference dispatcher (:818-830): returns ``(mask, pair_start_nodes, pair_answer_nodes,
    pair_edge_local_ids, pair_edge_counts, pair_shortest_lengths)`` as python lists."""
    if path_mode not in _PATH_MODES:
        raise ValueError(f"Unsupported path_mode: {path_mode}. Expected one of {_PATH_MODES}.")
    num_edges = len(edge_src)
    empty = ([False] * num_edges, [], [], [], [], [])
    if num_nodes <= 0 or num_edges == 0 or len(sources) == 0 or len(targets) == 0:
        return empty
    es_np, ed_np = np.asarray(edge_src, dtype=np.int64), np.asarray(edge_dst, dtype=np.int64)
    if not bool(((es_np >= 0) & (ed_np >= 0) & (es_np < num_nodes) & (ed_np < num_nodes)).any()):
        return empty  # no valid edge (:704-706)
    src, dst = _single_graph_batch(num_nodes, edge_src, edge_dst)
    ptr = np.asarray([0, num_nodes], dtype=np.int64)
    gb = GraphBatch(src, dst, torch.zeros_like(src), torch.zeros(num_edges, dtype=torch.long, device=src.device),
                    torch.tensor([0, num_edges], dtype=torch.long, device=src.device), ptr,
                    np.asarray([0, num_edges], dtype=np.int64))
    seeds = np.asarray([int(s) for s in sources], dtype=np.int64)
    answers = np.asarray([int(t) for t in targets], dtype=np.int64)
    mask, info = batched_shortest_path_labels(gb, seeds, np.asarray([0, len(seeds)]), answers,
                                              np.asarray([0, len(answers)]), path_mode=path_mode, return_pairs=True)
    if info is None:
        return empty
    member = info["member"].cpu().numpy().astype(bool)
    p_s, p_a, p_ids, p_cnt, p_len = [], [], [], [], []
    for i, (_, _, _, s, a) in enumerate(info["pairs"]):
        if info["count"][i] < 0:
            continue
        ids = np.nonzero(member[info["out_off"][i]:info["out_off"][i + 1]])[0]
        p_s.append(s), p_a.append(a), p_len.append(int(info["length"][i])), p_cnt.append(int(ids.size))
        p_ids.extend(ids.tolist())
    return mask.cpu().numpy().astype(bool).tolist(), p_s, p_a, p_ids, p_cnt, p_len


def _shortest_path_union_mask_by_pair(num_nodes, edge_src, edge_dst, sources, targets):
    return _shortest_path_union_mask_by_pair_mode(num_nodes, edge_src, edge_dst, sources, targets,
                                                  path_mode="undirected")


def _shortest_path_union_mask_by_pair_directed(num_nodes, edge_src, edge_dst, sources, targets):
    return _shortest_path_union_mask_by_pair_mode(num_nodes, edge_src, edge_dst, sources, targets,
                                                  path_mode="qa_directed")


def has_connectivity(graph: Sequence[Tuple[str, str, str]], seeds: Sequence[str], answers: Sequence[str], *,
                     path_mode: str = "undirected") -> bool:
    """build_retrieval_pipeline.py:946-979: first-seen local ids, multi-source BFS from the seeds."""
    if not graph or not seeds or not answers:
        return False
    node_index: Dict[str, int] = {}
    es, ed = [], []
    for h, _, t in graph:
        es.append(node_index.setdefault(h, len(node_index)))
        ed.append(node_index.setdefault(t, len(node_index)))
    s_ids = [node_index[s] for s in seeds if s in node_index]
    a_ids = [node_index[a] for a in answers if a in node_index]
    if not s_ids or not a_ids:
        return False
    if path_mode not in _PATH_MODES:
        raise ValueError(f"Unsupported path_mode: {path_mode}. Expected one of {_PATH_MODES}.")
    dist = bfs_dist_edges(len(node_index), es, ed, s_ids, directed=(path_mode == "qa_directed"))
    return any(dist[a] >= 0 for a in a_ids)


# ------------------------------------------------------------------------------------------
# a16: deterministic single shortest path
# ------------------------------------------------------------------------------------------
def _next_pow2(x: np.ndarray) -> np.ndarray:
    x = np.maximum(np.asarray(x, dtype=np.int64), 1)
    return (1 << np.ceil(np.log2(x)).astype(np.int64)).astype(np.int64)


@_lib.device_guard
def shortest_path_single_jobs(head32: torch.Tensor, tail32: torch.Tensor, edge_begin, edge_end, node_begin, node_end,
                              sources, src_ptr, targets, tgt_ptr, *, ids_are_global: bool = True,
                              max_len: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """``evi_shortest_path_single`` for a list of jobs (one per question graph). Returns host arrays
    ``path_len [J]`` (-1: no path), ``path_edges [J, max_len]`` (job-local edge ids) and ``path_nodes [J, max_len+1]``
    (job-local node ids, source first)."""
    L = _lib.load()
    dev = head32.device
    node_begin = np.asarray(node_begin, dtype=np.int64)
    node_end = np.asarray(node_end, dtype=np.int64)
    J = int(node_begin.shape[0])
    nn = node_end - node_begin
    if max_len is None:
        max_len = int(max(1, nn.max(initial=1)))
    pow2 = _next_pow2(nn)
    sort_off = np.zeros(J, dtype=np.int64)
    np.cumsum(pow2[:-1], out=sort_off[1:])
    n_total = int(max(1, node_end.max(initial=1)))
    up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dtype=dt)).to(dev)
    nz = lambda a: a if len(a) else np.zeros(1)
    d_eb, d_ee = up(edge_begin, np.int64), up(edge_end, np.int64)
    d_nb, d_ne = up(node_begin, np.int32), up(node_end, np.int32)
    d_sp, d_src = up(src_ptr, np.int32), up(nz(np.asarray(sources, dtype=np.int64).clip(-1, 2**31 - 2)), np.int32)
    d_tp, d_tgt = up(tgt_ptr, np.int32), up(nz(np.asarray(targets, dtype=np.int64).clip(-1, 2**31 - 2)), np.int32)
    d_so = up(sort_off, np.int64)
    dist = torch.empty(n_total, dtype=torch.int32, device=dev)
    order = torch.empty(n_total, dtype=torch.int32, device=dev)
    key = torch.empty(n_total, dtype=torch.int64, device=dev)
    sortbuf = torch.empty(int(max(1, pow2.sum())), dtype=torch.int64, device=dev)
    plen = torch.empty(J, dtype=torch.int32, device=dev)
    pedges = torch.zeros((J, max(1, max_len)), dtype=torch.int32, device=dev)
    pnodes = torch.zeros((J, max_len + 1), dtype=torch.int32, device=dev)
    st = L.evi_shortest_path_single(
        head32.data_ptr(), tail32.data_ptr(), J, d_eb.data_ptr(), d_ee.data_ptr(), d_nb.data_ptr(), d_ne.data_ptr(),
        d_sp.data_ptr(), d_src.data_ptr(), d_tp.data_ptr(), d_tgt.data_ptr(), int(ids_are_global), d_so.data_ptr(),
        dist.data_ptr(), order.data_ptr(), key.data_ptr(), sortbuf.data_ptr(), int(max_len), plen.data_ptr(),
        pedges.data_ptr(), pnodes.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_shortest_path_single")
    plen_h = plen.cpu().numpy()
    if (plen_h == -2).any():
        raise RuntimeError("shortest path longer than max_len; raise max_len")
    return plen_h, pedges.cpu().numpy(), pnodes.cpu().numpy()


def batched_shortest_path_single(gb: GraphBatch, seeds: np.ndarray, seed_ptr: np.ndarray, answers: np.ndarray,
                                 answer_ptr: np.ndarray, *, max_len: Optional[int] = None):
    """One deterministic shortest path per question of a batch (seeds / answers are batch-global node ids)."""
    ptr = gb.ptr_host
    ep = gb.edge_ptr_np()
    return shortest_path_single_jobs(gb.head32, gb.tail32, ep[:-1], ep[1:], ptr[:-1], ptr[1:], seeds, seed_ptr, answers,
                                     answer_ptr, ids_are_global=True, max_len=max_len)


def _shortest_path_single(num_nodes: int, edge_src: Sequence[int], edge_dst: Sequence[int], sources: Sequence[int],
                          targets: Sequence[int]) -> Tuple[List[int], List[int]]:
    """Reference signature and return convention (build_retrieval_pipeline.py:453-530): ``(edges, nodes)``,
    ``([], [])`` when there is no path, ``([], [t])`` when the best target is itself a source."""
    if not len(sources) or not len(targets) or num_nodes <= 0:
        return [], []
    src, dst = _single_graph_batch(num_nodes, edge_src, edge_dst)
    E = int(src.numel())
    if E == 0:
        src = torch.zeros(1, dtype=torch.int32, device=src.device)
        dst = torch.zeros(1, dtype=torch.int32, device=src.device)
    plen, pe, pn = shortest_path_single_jobs(src, dst, [0], [E], [0], [num_nodes], [int(s) for s in sources],
                                             [0, len(sources)], [int(t) for t in targets], [0, len(targets)],
                                             ids_are_global=False)
    n = int(plen[0])
    if n < 0:
        return [], []
    return pe[0, :n].tolist(), pn[0, :n + 1].tolist()


def shortest_path_edge_indices_undirected(num_nodes, edge_src, edge_dst, seeds, answers):
    """build_retrieval_pipeline.py:935-943."""
    return _shortest_path_single(num_nodes, edge_src, edge_dst, seeds, answers)
