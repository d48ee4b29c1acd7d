"""g_agent materialisation on the device.

Drop-in for ``src/data/components/g_agent_builder.py``: the same ``GAgentSettings`` / ``GAgentBuilder`` names,
``process_batch(batch, model_output)``, ``save(path)`` and the same record schema
(``_sample_to_record``, g_agent_builder.py:543-569), but the per-question CPU loops — node-softmax scores
(:594-626), stable top-k (:640-652), per-start-node top edges (:654-724), the Python-dict (h, r, t) dedup
(:338-354) and the dict re-indexing (:364-385) — run as three kernels over the whole batch
(``evi_node_softmax_logit``, ``evi_gagent_start_edges``, ``evi_gagent_build``) plus ``evi_segmented_topk``.
One device->host copy of the fixed-shape result arrays replaces ~10 ``.cpu()`` calls per question.

Per-sample metadata (question text / embedding, seed and answer entity ids) still comes from the caller's
``embedding_store.load_sample(sample_id)`` exactly as in the reference (:252-281); any object with that method
works. ``select_batch`` is the array-level entry point for callers that already hold the ids.
"""
from __future__ import annotations

import statistics
from dataclasses import asdict, dataclass
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence

import torch

from . import lib as _lib
from .topk import segmented_topk

SCORE_MODE_LOGITS = "logits"
SCORE_MODE_NODE_SOFTMAX = "node_softmax"


@dataclass
class GAgentSettings:
    """Same fields, defaults and validation as ``GAgentSettings`` (g_agent_builder.py:31-91)."""

    enabled: bool = True
    allow_empty_answer: bool = False
    edge_top_k: int = 50
    max_hops: int = 2
    score_temperature: float = 1.0
    score_bias: float = 0.0
    start_keep_ratio: float = 0.25
    start_min_edges: int = 1
    start_max_edges: Optional[int] = None
    score_mode: str = SCORE_MODE_NODE_SOFTMAX
    output_path: Path = Path("g_agent/g_agent_samples.pt")

    def __post_init__(self) -> None:
        self.enabled = bool(self.enabled)
        self.allow_empty_answer = bool(self.allow_empty_answer)
        self.edge_top_k = int(self.edge_top_k)
        self.max_hops = int(self.max_hops)
        self.score_temperature = float(self.score_temperature)
        self.score_bias = float(self.score_bias)
        if self.max_hops < 0:
            raise ValueError(f"max_hops must be >= 0, got {self.max_hops}")
        if self.edge_top_k <= 0:
            raise ValueError(f"edge_top_k must be > 0, got {self.edge_top_k}")
        if not (self.score_temperature > 0.0):
            raise ValueError(f"score_temperature must be positive, got {self.score_temperature}")
        self.start_keep_ratio = float(self.start_keep_ratio)
        if not (0.0 <= self.start_keep_ratio <= 1.0):
            raise ValueError(f"start_keep_ratio must be in [0, 1], got {self.start_keep_ratio}")
        self.start_min_edges = int(self.start_min_edges)
        if self.start_min_edges < 0:
            raise ValueError(f"start_min_edges must be >= 0, got {self.start_min_edges}")
        self.start_max_edges = int(self.edge_top_k) if self.start_max_edges is None else int(self.start_max_edges)
        if self.start_max_edges < 0:
            raise ValueError(f"start_max_edges must be >= 0, got {self.start_max_edges}")
        if self.start_max_edges != 0 and self.start_min_edges > self.start_max_edges:
            raise ValueError(
                f"start_min_edges must be <= start_max_edges, got {self.start_min_edges} > {self.start_max_edges}")
        if self.score_mode not in {SCORE_MODE_LOGITS, SCORE_MODE_NODE_SOFTMAX}:
            raise ValueError(
                f"score_mode must be one of {SCORE_MODE_LOGITS}/{SCORE_MODE_NODE_SOFTMAX}, got {self.score_mode}")
        self.output_path = Path(self.output_path).expanduser()

    def to_metadata(self) -> Dict[str, Any]:
        payload = asdict(self)
        payload["output_path"] = str(payload.get("output_path"))
        return payload


@dataclass
class GAgentSample:
    """Field-for-field ``GAgentSample`` (src/data/g_agent_dataset.py:20-52)."""

    sample_id: str
    question: str
    question_emb: torch.Tensor
    node_entity_ids: torch.Tensor
    node_embedding_ids: torch.Tensor
    edge_head_locals: torch.Tensor
    edge_tail_locals: torch.Tensor
    edge_relations: torch.Tensor
    edge_scores: torch.Tensor
    edge_labels: torch.Tensor
    start_entity_ids: torch.Tensor
    answer_entity_ids: torch.Tensor
    gt_path_edge_local_ids: torch.Tensor
    start_node_locals: torch.Tensor
    answer_node_locals: torch.Tensor
    pair_start_node_locals: torch.Tensor
    pair_answer_node_locals: torch.Tensor
    pair_edge_local_ids: torch.Tensor
    pair_edge_counts: torch.Tensor
    pair_shortest_lengths: torch.Tensor
    gt_path_exists: bool = False
    is_answer_reachable: bool = False
    is_dummy_agent: bool = False


@_lib.device_guard
def node_softmax_logit(scores: torch.Tensor, head: torch.Tensor, tail: torch.Tensor, num_nodes: int, *,
                       temperature: float = 1.0, bias: float = 0.0, score_mode: str = SCORE_MODE_NODE_SOFTMAX):
    """``_node_softmax_logit`` (g_agent_builder.py:594-626) over any edge list with node ids in
    ``[0, num_nodes)``; returns ``(calibrated, normalised)`` fp32 device tensors."""
    if scores.device.type != "cuda":
        raise RuntimeError("node_softmax_logit runs on CUDA tensors only (no CPU fallback).")
    L = _lib.load()
    dev = scores.device
    s = scores.detach().to(torch.float32).contiguous().view(-1)
    E = int(s.numel())
    h32 = head.to(device=dev, dtype=torch.int32).contiguous().view(-1)
    t32 = tail.to(device=dev, dtype=torch.int32).contiguous().view(-1)
    cal = torch.empty_like(s)
    out = torch.empty_like(s)
    nbytes = int(L.evi_node_softmax_workspace_bytes(int(num_nodes)))
    ws = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=dev)
    mode = _lib.GAGENT_LOGITS if score_mode == SCORE_MODE_LOGITS else _lib.GAGENT_NODE_SOFTMAX
    st = L.evi_node_softmax_logit(s.data_ptr(), h32.data_ptr(), t32.data_ptr(), E, int(num_nodes), mode,
                                  float(temperature), float(bias), ws.data_ptr(), nbytes, cal.data_ptr(),
                                  out.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_node_softmax_logit")
    return cal, out


def _ragged(lists: Sequence[Sequence[int]], dev) -> tuple:
    ptr = [0]
    flat: List[int] = []
    for x in lists:
        flat.extend(int(v) for v in x)
        ptr.append(len(flat))
    return (torch.tensor(flat, dtype=torch.long, device=dev).contiguous() if flat
            else torch.zeros(1, dtype=torch.long, device=dev)), torch.tensor(ptr, dtype=torch.long, device=dev)


@_lib.device_guard
def select_batch(*, logits: torch.Tensor, edge_index: torch.Tensor, edge_ptr: torch.Tensor, ptr: torch.Tensor,
                 edge_attr: torch.Tensor, labels: Optional[torch.Tensor], node_global_ids: torch.Tensor,
                 node_embedding_ids: torch.Tensor, start_entity_ids: Sequence[Sequence[int]],
                 answer_entity_ids: Sequence[Sequence[int]], settings: GAgentSettings,
                 cap: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """Steps A, C, D of ``_build_and_add_sample`` (g_agent_builder.py:298-428) for every question of a batch.
    Returns host tensors: per-question counts / status and the padded ``[B, cap]`` / ``[B, 2 cap]`` arrays
    (see ``evi_gagent_args``). Edges must be grouped by question (``edge_ptr``), as the dataset collates them."""
    if logits.device.type != "cuda":
        raise RuntimeError("g_agent selection runs on CUDA tensors only (no CPU fallback).")
    L = _lib.load()
    dev = logits.device
    cfg = settings
    B = int(ptr.numel() - 1)
    ei = edge_index.to(device=dev, dtype=torch.long)
    h32 = ei[0].to(torch.int32).contiguous()
    t32 = ei[1].to(torch.int32).contiguous()
    E = int(h32.numel())
    N = int(node_global_ids.numel())
    edge_ptr = edge_ptr.to(device=dev, dtype=torch.long).contiguous()
    node_ptr = ptr.to(device=dev, dtype=torch.long).contiguous()
    ea = edge_attr.to(device=dev, dtype=torch.long).contiguous()
    ngid = node_global_ids.to(device=dev, dtype=torch.long).contiguous()
    nemb = node_embedding_ids.to(device=dev, dtype=torch.long).contiguous()
    lab = None if labels is None else labels.to(device=dev, dtype=torch.float32).contiguous().view(-1)
    cal, sel = node_softmax_logit(logits, h32, t32, N, temperature=cfg.score_temperature, bias=cfg.score_bias,
                                  score_mode=cfg.score_mode)
    k = int(cfg.edge_top_k)
    top = segmented_topk(sel, edge_ptr, k)
    start_flat, start_ptr = _ragged(start_entity_ids, dev)
    ans_flat, ans_ptr = _ragged(answer_entity_ids, dev)
    mark = torch.empty(max(E, 1), dtype=torch.uint8, device=dev)
    status = torch.zeros(B, dtype=torch.int32, device=dev)
    st = L.evi_gagent_start_edges(sel.data_ptr(), h32.data_ptr(), t32.data_ptr(), edge_ptr.data_ptr(),
                                  node_ptr.data_ptr(), B, ngid.data_ptr(), start_flat.data_ptr(), start_ptr.data_ptr(),
                                  float(cfg.start_keep_ratio), int(cfg.start_min_edges),
                                  -1 if cfg.start_max_edges is None else int(cfg.start_max_edges), mark.data_ptr(),
                                  status.data_ptr(), _lib.stream_handle())
    _lib.check(st, "evi_gagent_start_edges")
    max_starts = max(1, max((len(x) for x in start_entity_ids), default=1))
    max_answers = max(1, max((len(x) for x in answer_entity_ids), default=1))
    if cap is None:  # top-k plus at most start_max_edges per start node
        per_start = k if cfg.start_max_edges is None else int(cfg.start_max_edges)
        cap = min(_lib.GAGENT_MAX_CAP, k + per_start * min(max_starts, 64))
    i32 = lambda *shape: torch.zeros(shape, dtype=torch.int32, device=dev)
    i64 = lambda *shape: torch.zeros(shape, dtype=torch.long, device=dev)
    f32 = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)
    out = {
        "status": status, "num_triples": i32(B), "num_nodes": i32(B), "num_start_locals": i32(B),
        "num_answers": i32(B), "num_answer_locals": i32(B), "env_edge_ids": i32(B, cap),
        "edge_heads_global": i64(B, cap), "edge_tails_global": i64(B, cap), "edge_relations": i64(B, cap),
        "edge_scores": f32(B, cap), "edge_labels": f32(B, cap), "edge_head_locals": i32(B, cap),
        "edge_tail_locals": i32(B, cap), "node_entity_ids": i64(B, 2 * cap), "node_embedding_ids": i64(B, 2 * cap),
        "node_batch_index": i32(B, 2 * cap), "start_node_locals": i32(B, max_starts),
        "answer_entity_ids": i64(B, max_answers), "answer_node_locals": i32(B, max_answers),
    }
    a = _lib.GAgentArgs()
    a.num_graphs, a.cap, a.topk_stride, a.max_starts, a.max_answers = B, int(cap), k, max_starts, max_answers
    a.score_mode = _lib.GAGENT_LOGITS if cfg.score_mode == SCORE_MODE_LOGITS else _lib.GAGENT_NODE_SOFTMAX
    a.raw_scores, a.labels = cal.data_ptr(), _lib.ptr(lab)
    a.head32, a.tail32, a.edge_attr = h32.data_ptr(), t32.data_ptr(), ea.data_ptr()
    a.edge_ptr, a.node_ptr = edge_ptr.data_ptr(), node_ptr.data_ptr()
    a.node_global_ids, a.node_embedding_ids = ngid.data_ptr(), nemb.data_ptr()
    a.start_entity_ids, a.start_ptr = start_flat.data_ptr(), start_ptr.data_ptr()
    a.answer_entity_ids, a.answer_ptr = ans_flat.data_ptr(), ans_ptr.data_ptr()
    a.topk_indices, a.topk_counts = top.indices.data_ptr(), top.counts.data_ptr()
    a.mark, a.status = mark.data_ptr(), status.data_ptr()
    a.num_triples, a.num_nodes_out = out["num_triples"].data_ptr(), out["num_nodes"].data_ptr()
    a.num_start_locals, a.num_answers = out["num_start_locals"].data_ptr(), out["num_answers"].data_ptr()
    a.num_answer_locals, a.env_edge_ids = out["num_answer_locals"].data_ptr(), out["env_edge_ids"].data_ptr()
    a.edge_heads_global, a.edge_tails_global = out["edge_heads_global"].data_ptr(), out["edge_tails_global"].data_ptr()
    a.edge_relations, a.edge_scores = out["edge_relations"].data_ptr(), out["edge_scores"].data_ptr()
    a.edge_labels, a.edge_head_locals = out["edge_labels"].data_ptr(), out["edge_head_locals"].data_ptr()
    a.edge_tail_locals, a.node_entity_ids = out["edge_tail_locals"].data_ptr(), out["node_entity_ids"].data_ptr()
    a.node_embedding_ids_out, a.node_batch_index = out["node_embedding_ids"].data_ptr(), out["node_batch_index"].data_ptr()
    a.start_node_locals, a.answer_ids_unique = out["start_node_locals"].data_ptr(), out["answer_entity_ids"].data_ptr()
    a.answer_node_locals = out["answer_node_locals"].data_ptr()
    import ctypes as C

    st = L.evi_gagent_build(C.byref(a), _lib.stream_handle())
    _lib.check(st, "evi_gagent_build")
    host = {key: v.cpu() for key, v in out.items()}  # the one device->host copy of the selection
    host["selection_scores"] = sel
    host["calibrated"] = cal
    return host


class GAgentBuilder:
    """``GAgentBuilder`` (g_agent_builder.py:116-571) with the selection on the device."""

    def __init__(self, settings: GAgentSettings, embedding_store: Any = None, aux_embedding_store: Any = None) -> None:
        self.cfg = settings
        self.embedding_store = embedding_store
        self.aux_embedding_store = aux_embedding_store
        self.samples: List[GAgentSample] = []
        self.stats: Dict[str, Any] = {"num_samples": 0, "path_exists": 0, "retrieval_failed": 0, "edge_counts": [],
                                      "path_lengths": []}

    def reset(self) -> None:
        self.samples = []
        for key in ("num_samples", "path_exists", "retrieval_failed"):
            self.stats[key] = 0
        self.stats["edge_counts"] = []
        self.stats["path_lengths"] = []

    @staticmethod
    def _select_aux_field(aux_data, core_data, field: str, default):
        if aux_data is not None and field in aux_data:
            return aux_data[field]
        return core_data.get(field, default)

    def process_batch(self, batch: Any, model_output: Any) -> None:
        """Same contract and error messages as g_agent_builder.py:148-236."""
        if not hasattr(batch, "ptr"):
            raise ValueError("Batch must have 'ptr' for slicing.")
        logits = getattr(model_output, "logits", None)
        if logits is None:
            raise ValueError("Retriever output missing logits; g_agent requires logit scores.")
        scores = logits.detach().view(-1)
        if scores.numel() == 0:
            return
        query_ids = getattr(model_output, "query_ids", None)
        if query_ids is None:
            raise ValueError("Retriever output missing query_ids; g_agent requires per-edge graph mapping.")
        if not hasattr(batch, "node_embedding_ids"):
            raise AttributeError("Batch missing node_embedding_ids; retriever dataset must provide embedding ids per node.")
        if self.embedding_store is None:
            raise ValueError("EmbeddingStore must be provided; builder cannot proceed without per-sample metadata.")
        dev = scores.device
        query_ids = torch.as_tensor(query_ids, device=dev).to(torch.long).view(-1)
        if query_ids.numel() != scores.numel():
            raise ValueError(f"query_ids/logits shape mismatch: {query_ids.shape} vs {scores.shape}")
        B = int(batch.ptr.numel() - 1)
        counts = torch.bincount(query_ids, minlength=B)
        if counts.numel() != B:
            raise ValueError("Invalid query_ids: edge assignments exceed batch_size.")
        if bool((query_ids[1:] < query_ids[:-1]).any()):
            raise ValueError("g_agent selection expects edges grouped by question (PyG collate order).")
        edge_ptr = torch.zeros(B + 1, dtype=torch.long, device=dev)
        edge_ptr[1:] = torch.cumsum(counts, 0)
        sample_ids = list(getattr(batch, "sample_id", []))
        metas: List[Optional[Dict[str, Any]]] = []
        starts: List[Sequence[int]] = []
        answers: List[Sequence[int]] = []
        counts_host = counts.tolist()
        for i in range(B):
            sid = str(sample_ids[i]) if i < len(sample_ids) else f"unknown_{i}"
            meta = self._load_meta(sid) if counts_host[i] > 0 else None
            metas.append(meta)
            starts.append(meta["start_entity_ids"].tolist() if meta else [])
            answers.append(meta["answer_entity_ids"].tolist() if meta else [])
        res = select_batch(logits=scores, edge_index=batch.edge_index, edge_ptr=edge_ptr, ptr=batch.ptr,
                           edge_attr=batch.edge_attr, labels=batch.labels, node_global_ids=batch.node_global_ids,
                           node_embedding_ids=batch.node_embedding_ids, start_entity_ids=starts,
                           answer_entity_ids=answers, settings=self.cfg)
        for i in range(B):
            if metas[i] is not None:
                self._add_sample(i, metas[i], res)

    def _load_meta(self, sample_id: str) -> Optional[Dict[str, Any]]:
        """Metadata fetch and validation of g_agent_builder.py:252-281."""
        try:
            raw = self.embedding_store.load_sample(sample_id)
        except KeyError:
            return None
        aux = self.aux_embedding_store.load_sample(sample_id) if self.aux_embedding_store is not None else None
        q = torch.as_tensor(raw.get("question_emb", []), dtype=torch.float32).detach().clone()
        if q.dim() == 1:
            q = q.unsqueeze(0)
        elif q.dim() != 2:
            raise ValueError(f"question_emb must be 1D or 2D, got shape {tuple(q.shape)} for {sample_id}")
        text = self._select_aux_field(aux, raw, "question", "")
        if not isinstance(text, str):
            raise TypeError(f"question must be string for {sample_id}, got {type(text).__name__}")
        start = torch.as_tensor(self._select_aux_field(aux, raw, "seed_entity_ids", []), dtype=torch.long).clone()
        if start.numel() == 0:
            raise ValueError(f"Sample {sample_id} missing seed_entity_ids (start_entity_ids). "
                             "If using split LMDBs, pass aux_lmdb_path to GAgentMaterializationCallback.")
        ans_raw = raw.get("answer_entity_ids")
        if ans_raw is None:
            ans_raw = self._select_aux_field(aux, raw, "answer_entity_ids", [])
        ans = torch.as_tensor(ans_raw, dtype=torch.long).clone()
        if ans.numel() == 0:
            raise ValueError(f"Sample {sample_id} missing answer_entity_ids.")
        return {"sample_id": sample_id, "question": text, "question_emb": q, "start_entity_ids": start.view(-1),
                "answer_entity_ids": ans.view(-1)}

    def _add_sample(self, i: int, meta: Dict[str, Any], res: Dict[str, torch.Tensor]) -> None:
        st = int(res["status"][i])
        sid = meta["sample_id"]
        if st & _lib.GAGENT_EMPTY:
            return
        if st & _lib.GAGENT_NO_START_NODE:
            raise ValueError(f"Start entities missing from retrieval graph (sample_id={sid}).")
        if st & _lib.GAGENT_OVERFLOW:
            raise RuntimeError(f"g_agent selection overflow for {sid}: raise `cap` (more selected edges than the "
                               "fixed-shape output holds) or reduce the number of start entities (<= 64).")
        if st & _lib.GAGENT_NO_START_LOCAL:
            self.stats["retrieval_failed"] += 1
            return
        m, nu = int(res["num_triples"][i]), int(res["num_nodes"][i])
        nal = int(res["num_answer_locals"][i])
        dummy = nal == 0
        if dummy and not self.cfg.allow_empty_answer:
            self.stats["retrieval_failed"] += 1
            return
        empty = torch.empty(0, dtype=torch.long)
        edge_labels = res["edge_labels"][i, :m].clone()
        if dummy:
            edge_labels = torch.zeros(m, dtype=torch.float32)
        sample = GAgentSample(
            sample_id=sid, question=meta["question"], question_emb=meta["question_emb"],
            node_entity_ids=res["node_entity_ids"][i, :nu].clone(),
            node_embedding_ids=res["node_embedding_ids"][i, :nu].clone(),
            edge_head_locals=res["edge_head_locals"][i, :m].long(), edge_tail_locals=res["edge_tail_locals"][i, :m].long(),
            edge_relations=res["edge_relations"][i, :m].clone(), edge_scores=res["edge_scores"][i, :m].clone(),
            edge_labels=edge_labels, start_entity_ids=meta["start_entity_ids"],
            answer_entity_ids=res["answer_entity_ids"][i, :int(res["num_answers"][i])].clone(),
            gt_path_edge_local_ids=empty, start_node_locals=res["start_node_locals"][i, :int(res["num_start_locals"][i])].long(),
            answer_node_locals=res["answer_node_locals"][i, :nal].long() if nal else empty,
            pair_start_node_locals=empty, pair_answer_node_locals=empty, pair_edge_local_ids=empty,
            pair_edge_counts=empty, pair_shortest_lengths=empty, gt_path_exists=False,
            is_answer_reachable=not dummy, is_dummy_agent=dummy)
        self.samples.append(sample)
        self.stats["num_samples"] += 1
        self.stats["edge_counts"].append(m)

    def _sample_to_record(self, s: GAgentSample) -> Dict[str, Any]:
        """Deterministic record schema of g_agent_builder.py:543-569."""
        keys = ["sample_id", "question", "question_emb", "edge_relations", "edge_scores", "edge_labels",
                "edge_head_locals", "edge_tail_locals", "node_entity_ids", "node_embedding_ids", "start_entity_ids",
                "start_node_locals", "answer_entity_ids", "answer_node_locals", "pair_start_node_locals",
                "pair_answer_node_locals", "pair_edge_local_ids", "pair_edge_counts", "pair_shortest_lengths",
                "gt_path_edge_local_ids"]
        rec = {key: getattr(s, key) for key in keys}
        rec.update(gt_path_exists=bool(s.gt_path_exists), is_answer_reachable=bool(s.is_answer_reachable),
                   is_dummy_agent=bool(s.is_dummy_agent))
        return rec

    def save(self, output_path: Path) -> Optional[Dict[str, Any]]:
        """``{"settings", "stats", "samples"}`` payload of g_agent_builder.py:512-541."""
        if not self.samples:
            return None
        output_path = Path(output_path)
        output_path.parent.mkdir(parents=True, exist_ok=True)
        attempted = int(self.stats["num_samples"]) + int(self.stats["retrieval_failed"])
        final_stats = {
            "num_samples": self.stats["num_samples"],
            "path_exists_ratio": self.stats["path_exists"] / max(1, self.stats["num_samples"]),
            "retrieval_failed_ratio": self.stats["retrieval_failed"] / max(1, attempted),
            "avg_edges": statistics.mean(self.stats["edge_counts"]) if self.stats["edge_counts"] else 0,
            "avg_gt_len": statistics.mean(self.stats["path_lengths"]) if self.stats["path_lengths"] else 0,
        }
        payload = {"settings": self.cfg.to_metadata(), "stats": final_stats,
                   "samples": [self._sample_to_record(s) for s in self.samples]}
        torch.save(payload, output_path)
        return final_stats


__all__ = ["GAgentBuilder", "GAgentSettings", "GAgentSample", "node_softmax_logit", "select_batch"]
