"""Top-k evidence writer: the ``eval_retriever/<split>.pt`` artifact.

Drop-in for ``src/callbacks/retriever_topk_edge_writer.py``: same constructor keywords, hook names
(``on_predict_start`` / ``write_on_batch_end`` / ``on_predict_end`` and the ``on_test_*`` trio), record schema
(:277-349) and files — ``<split>.pt`` = ``{"settings": {split, topk_values}, "samples": [...]}`` and
``<split>.manifest.json`` = ``{artifact, schema_version, file, created_at, producer}`` (:155-170), which
``ReasonerOracleDataModule`` consumes (src/data/reasoner_oracle_datamodule.py:85-171).

What changes is where the work happens: instead of one ``torch.topk`` and seven ``.cpu().tolist()`` calls per
question (:295-319), a batch costs one ragged top-k launch, one payload-gather launch
(``evi_segmented_topk`` + ``evi_topk_gather_payload``) and one device->host copy of six ``[B, max_k]`` arrays.
The class has no Lightning dependency (``trainer`` only needs ``global_rank``); with Lightning installed it can
be listed under ``callbacks`` as is, because Lightning calls these hooks by name.
"""
from __future__ import annotations

import json
from datetime import datetime, timezone
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from .topk import MAX_K, segmented_topk, writer_payload

try:  # plain-C record builder (csrc/records.c, built by build.sh); the Python loop below is the same thing, slower
    from . import _evi_records
except ImportError:  # pragma: no cover - depends on the build
    _evi_records = None

_DEFAULT_TOPK_VALUES = (1, 5, 10)


def normalize_k_values(raw_values: Any, default: Optional[Sequence[int]] = None) -> List[int]:
    """``src/utils/metrics.py:25-40``: positive unique ints, ascending; ``default`` when nothing is left."""
    if raw_values is None:
        items: Sequence[Any] = []
    elif isinstance(raw_values, (list, tuple, set, range)):
        items = list(raw_values)
    elif isinstance(raw_values, str):
        items = [raw_values]
    else:
        try:
            items = list(raw_values)
        except TypeError:
            items = [raw_values]
    out: List[int] = []
    for item in items:
        try:
            k = int(item)
        except (TypeError, ValueError):
            continue
        if k > 0 and k not in out:
            out.append(k)
    if not out and default is not None:
        return normalize_k_values(default, default=None)
    return sorted(out)


def _sample_ids(batch: Any, num_graphs: int) -> List[str]:
    raw = getattr(batch, "sample_id", None)  # extract_sample_ids, src/utils/metrics.py:46-55
    if raw is None:
        return [str(i) for i in range(int(getattr(batch, "num_graphs", 0)))]
    if isinstance(raw, (list, tuple)):
        return [str(s) for s in raw]
    if torch.is_tensor(raw):
        return [str(s.item()) for s in raw]
    return [str(raw)]


def _questions(batch: Any, num_graphs: int) -> List[str]:
    raw = getattr(batch, "question", None)  # _extract_questions, retriever_topk_edge_writer.py:371-381
    if raw is None:
        return ["" for _ in range(num_graphs)]
    if isinstance(raw, (list, tuple)):
        return [str(q) for q in raw]
    if torch.is_tensor(raw):
        if raw.numel() == num_graphs:
            return [str(v.item()) for v in raw]
        return [str(raw.cpu().tolist()) for _ in range(num_graphs)]
    return [str(raw) for _ in range(num_graphs)]


_EDGE_KEYS = ("head_entity_id", "relation_id", "tail_entity_id", "head_text", "relation_text", "tail_text", "score",
              "label", "rank")
_HOST_KEYS = ("score", "count", "head_id", "tail_id", "relation_id", "label", "logit_fwd", "logit_bwd")


class _PendingBatch:
    """Selected triples of one batch on their way to the host: pinned arrays, the event that marks the end of the
    device->host copies, and the per-question metadata the records need."""

    __slots__ = ("done", "host", "sample_ids", "questions", "answers", "answer_ptr", "has_fwd", "has_bwd", "num_graphs")


class RetrieverTopKEdgeWriter:
    """Persist retriever top-k edges during predict/test (retriever_topk_edge_writer.py:68-170).

    Per batch the hook enqueues one ragged top-k, one payload gather and the device->host copies into rotating pinned
    buffers and returns without synchronising. ``materialise="batch"`` builds the records of batch i-1 while the device
    works on batch i (``pipeline_depth`` batches in flight); ``materialise="end"`` keeps the selected triples columnar
    ([B, max_k] host arrays) until ``on_predict_end`` / ``on_test_end`` turns them into records in one go. The default
    ``"auto"`` picks "batch" when the plain-C record builder (csrc/records.c: 5.8 ms per 128 questions x 110 dicts,
    less than the device needs for the batch) has been built, else "end" (the Python loop costs 16.5 ms per batch and
    would stall the next forward)."""

    def __init__(self, *, output_dir, split: str = "test", enabled: bool = True, artifact_name: str = "eval_retriever",
                 schema_version: int = 1, topk_values: Optional[Sequence[int]] = None, textualize: bool = False,
                 entity_vocab_path: Optional[str] = None, relation_vocab_path: Optional[str] = None,
                 overwrite: bool = True, pipeline_depth: int = 1, materialise: str = "auto") -> None:
        name = str(artifact_name).strip()
        version = int(schema_version)
        ks = normalize_k_values(topk_values, default=_DEFAULT_TOPK_VALUES)
        if not name:
            raise ValueError("artifact_name must be a non-empty string.")
        if version <= 0:
            raise ValueError("schema_version must be a positive integer.")
        if not ks:
            raise ValueError("topk_values must be a non-empty list of positive integers.")
        if ks[-1] > MAX_K:
            raise ValueError(f"topk_values up to {MAX_K} are supported by the selection kernel (one CTA sorts a "
                             f"question's candidates in shared memory); got max k = {ks[-1]}.")
        self.enabled, self.overwrite, self.textualize = bool(enabled), bool(overwrite), bool(textualize)
        self.output_dir, self.split = Path(output_dir), str(split)
        self.artifact_name, self.schema_version = name, version
        self.topk_values, self._max_topk = ks, ks[-1]
        self.entity_vocab_path, self.relation_vocab_path = entity_vocab_path, relation_vocab_path
        self.pipeline_depth = max(0, int(pipeline_depth))
        if materialise not in ("auto", "end", "batch"):
            raise ValueError(f"materialise must be 'auto', 'end' or 'batch', got {materialise!r}")
        self.materialise = ("batch" if _evi_records is not None else "end") if materialise == "auto" else materialise
        self._columnar: List[_PendingBatch] = []
        self.use_native_records = True  # csrc/records.c when built; False forces the Python loop (tests compare both)
        self._records: List[Dict[str, Any]] = []
        self._pending: List[_PendingBatch] = []
        self._pinned: List[Dict[str, torch.Tensor]] = []
        self._batches_seen = 0
        self._entity_map: Optional[Dict[int, str]] = None
        self._relation_map: Optional[Dict[int, str]] = None
        self._output_path: Optional[Path] = None
        self._manifest_path: Optional[Path] = None

    # ------------------------------------------------------------------ hooks
    def on_predict_start(self, trainer=None, pl_module=None) -> None:
        if not self.enabled:
            return
        self._records, self._pending, self._columnar, self._batches_seen = [], [], [], 0
        self.output_dir.mkdir(parents=True, exist_ok=True)
        target = self.output_dir / f"{self.split}.pt"
        if target.exists() and not self.overwrite:
            raise FileExistsError(f"Output path already exists: {target}")
        self._output_path, self._manifest_path = target, target.with_name(f"{self.split}.manifest.json")
        if self.textualize:
            self._entity_map, self._relation_map = self._resolve_vocab_maps()

    def write_on_batch_end(self, trainer, pl_module, prediction, batch_indices, batch, batch_idx, dataloader_idx=0) -> None:
        self._collect_prediction(prediction=prediction, batch=batch)

    def on_predict_end(self, trainer=None, pl_module=None) -> None:
        if not self.enabled:
            return
        self._drain(0)
        for item in self._columnar:  # materialise="end": every batch's records now, in batch order
            self._records.extend(self._materialise(item))
        self._columnar = []
        if not self._records:
            return
        records = self._gather_records(self._records)
        is_writer_rank = trainer is None or int(getattr(trainer, "global_rank", 0)) == 0
        if not records or not is_writer_rank:
            return
        if self._output_path is None or self._manifest_path is None:
            raise RuntimeError("RetrieverTopKEdgeWriter missing output path; on_predict_start was not called.")
        torch.save({"settings": {"split": self.split, "topk_values": list(map(int, self.topk_values))},
                    "samples": records}, self._output_path)
        stamp = datetime.now(timezone.utc).replace(tzinfo=None).isoformat(timespec="seconds") + "Z"
        self._manifest_path.write_text(json.dumps({
            "artifact": self.artifact_name, "schema_version": self.schema_version, "file": self._output_path.name,
            "created_at": stamp, "producer": "retriever_topk_edge_writer"}, indent=2), encoding="utf-8")

    def on_test_start(self, trainer=None, pl_module=None) -> None:
        self.on_predict_start(trainer, pl_module)

    def on_test_batch_end(self, trainer, pl_module, outputs, batch, batch_idx, dataloader_idx: int = 0) -> None:
        self._collect_prediction(prediction=outputs, batch=batch)

    def on_test_end(self, trainer=None, pl_module=None) -> None:
        self.on_predict_end(trainer, pl_module)

    # ------------------------------------------------------------------ records
    def _collect_prediction(self, *, prediction: Any, batch: Any) -> None:
        if not self.enabled or prediction is None:
            return
        if isinstance(prediction, dict):
            if "logits" not in prediction or "query_ids" not in prediction:
                return
            import types

            prediction = types.SimpleNamespace(
                logits=prediction["logits"], query_ids=prediction["query_ids"],
                logits_fwd=prediction.get("logits_fwd"), logits_bwd=prediction.get("logits_bwd"))
        elif not hasattr(prediction, "logits") or not hasattr(prediction, "query_ids"):
            return
        self._pending.append(self._enqueue(batch=batch, output=prediction))
        self._drain(self.pipeline_depth)

    def _build_records(self, *, batch: Any, output: Any) -> List[Dict[str, Any]]:
        """Records of one batch, synchronously (the per-batch path goes through ``_enqueue`` / ``_drain``)."""
        return self._materialise(self._enqueue(batch=batch, output=output))

    def _drain(self, keep: int) -> None:
        """Turn all but the newest ``keep`` in-flight batches into records."""
        while len(self._pending) > keep:
            item = self._pending.pop(0)
            if self.materialise == "batch":
                self._records.extend(self._materialise(item))
            else:  # keep the columns, free the pinned buffers for the next batches
                item.done.synchronize()
                item.host = {k: v.clone() for k, v in item.host.items()}
                self._columnar.append(item)

    def _host_buffers(self, pay: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Pinned destination of one batch's payload: a rotating set per in-flight batch (allocating pinned memory per
        batch would call cudaHostAlloc, which synchronises the device)."""
        n_sets = self.pipeline_depth + 2
        while len(self._pinned) < n_sets:
            self._pinned.append({})
        slot = self._pinned[self._batches_seen % n_sets]
        self._batches_seen += 1
        out = {}
        for k in _HOST_KEYS:
            src = pay[k]
            buf = slot.get(k)
            if buf is None or buf.numel() < src.numel() or buf.dtype != src.dtype:
                buf = torch.empty(max(1, int(src.numel() * 1.25)), dtype=src.dtype, pin_memory=True)
                slot[k] = buf
            view = buf[: src.numel()].view(src.shape)
            view.copy_(src, non_blocking=True)
            out[k] = view
        return out

    def _edge_ptr(self, batch: Any, query_ids: torch.Tensor, num_edges: int, B: int, dev) -> torch.Tensor:
        """Edge offsets per question: the collate's slice dict (:389-395), a precomputed ``batch.edge_ptr``
        (loader.py:81-99), else the counts of the (sorted) query ids."""
        slice_dict = getattr(batch, "_slice_dict", None)
        for cand in (slice_dict.get("edge_index") if isinstance(slice_dict, dict) else None,
                     getattr(batch, "edge_ptr", None)):
            if cand is not None:
                cand = torch.as_tensor(cand, dtype=torch.long).view(-1)
                if cand.numel() == B + 1:
                    # non_blocking: a blocking copy from pageable memory synchronises the stream, i.e. waits for the
                    # forward that was just queued, and the record building below would then run with the device idle
                    return cand.to(dev, non_blocking=True)
        if query_ids.numel() != num_edges:
            raise ValueError(f"query_ids/scores mismatch: {tuple(query_ids.shape)} vs {num_edges}")
        if bool((query_ids[1:] < query_ids[:-1]).any()):
            raise ValueError("writer expects edges grouped by question (PyG collate order).")
        edge_ptr = torch.zeros(B + 1, dtype=torch.long, device=dev)
        edge_ptr[1:] = torch.cumsum(torch.bincount(query_ids, minlength=B)[:B], 0)
        return edge_ptr

    def _enqueue(self, *, batch: Any, output: Any) -> _PendingBatch:
        logits = output.logits.detach().view(-1)
        if logits.device.type != "cuda":
            raise RuntimeError("RetrieverTopKEdgeWriter selects on the device: logits must be a CUDA tensor.")
        dev = logits.device
        labels = batch.labels.detach().view(-1).to(device=dev, dtype=torch.float32)
        if logits.numel() != labels.numel():
            raise ValueError(f"scores/labels shape mismatch: {logits.shape} vs {labels.shape}")
        for name in ("edge_index", "edge_attr", "node_global_ids", "answer_entity_ids"):
            if getattr(batch, name, None) is None:
                raise ValueError(f"Batch missing {name} required for persistence.")
        query_ids = output.query_ids.detach().view(-1).to(dev)
        B = self._num_graphs(batch, query_ids)
        edge_ptr = self._edge_ptr(batch, query_ids, logits.numel(), B, dev)
        slice_dict = getattr(batch, "_slice_dict", None)
        answer_ptr_raw = getattr(batch, "answer_entity_ids_ptr", None)
        if answer_ptr_raw is None and isinstance(slice_dict, dict):
            answer_ptr_raw = slice_dict.get("answer_entity_ids")
        if answer_ptr_raw is None:
            raise ValueError("Batch missing answer_entity_ids_ptr required for persistence.")
        item = _PendingBatch()
        item.answer_ptr = torch.as_tensor(answer_ptr_raw, dtype=torch.long).view(-1).tolist()
        if len(item.answer_ptr) != B + 1:
            raise ValueError(f"answer_entity_ids_ptr length mismatch: {len(item.answer_ptr)} vs expected {B + 1}")
        item.answers = batch.answer_entity_ids.detach().cpu().view(-1).tolist()
        item.has_fwd = getattr(output, "logits_fwd", None) is not None
        item.has_bwd = getattr(output, "logits_bwd", None) is not None
        item.sample_ids, item.questions, item.num_graphs = _sample_ids(batch, B), _questions(batch, B), B
        with torch.cuda.device(dev):
            top = segmented_topk(logits, edge_ptr, self._max_topk, sigmoid=True)  # scores = sigmoid(logits) (:203)
            pay = writer_payload(top, edge_ptr, edge_index=batch.edge_index, node_global_ids=batch.node_global_ids,
                                 edge_attr=batch.edge_attr, labels=labels,
                                 logits_fwd=getattr(output, "logits_fwd", None),
                                 logits_bwd=getattr(output, "logits_bwd", None))
            item.host = self._host_buffers(pay)
            item.done = torch.cuda.Event()
            item.done.record(torch.cuda.current_stream(dev))
        return item

    def _materialise(self, item: _PendingBatch) -> List[Dict[str, Any]]:
        """The reference's per-question record (:321-349) from the [B, max_k] host arrays. Columns become Python lists
        once per batch; an edge dict is one zip step (the values are Python ints / floats already)."""
        item.done.synchronize()
        ent, rel = self._entity_map, self._relation_map
        if _evi_records is not None and self.use_native_records:
            h = item.host
            arr = lambda k: h[k].numpy()
            by_k = _evi_records.build(arr("count"), arr("head_id"), arr("relation_id"), arr("tail_id"), arr("score"),
                                      arr("label"), arr("logit_fwd") if item.has_fwd else None,
                                      arr("logit_bwd") if item.has_bwd else None, int(h["score"].shape[1]),
                                      tuple(int(k) for k in self.topk_values), ent, rel)
            return [{"sample_id": item.sample_ids[gid] if gid < len(item.sample_ids) else str(gid),
                     "question": item.questions[gid] if gid < len(item.questions) else "",
                     "triplets_by_k": tbk,
                     "answer_entity_ids": item.answers[item.answer_ptr[gid]:item.answer_ptr[gid + 1]]}
                    for gid, tbk in enumerate(by_k) if tbk is not None]
        cols = {k: v.tolist() for k, v in item.host.items()}
        keys = _EDGE_KEYS + (("logit_fwd",) if item.has_fwd else ()) + (("logit_bwd",) if item.has_bwd else ())
        ks = [int(k) for k in self.topk_values]
        records: List[Dict[str, Any]] = []
        for gid in range(item.num_graphs):
            kt = cols["count"][gid]
            if kt <= 0:
                continue
            h, r, t = cols["head_id"][gid][:kt], cols["relation_id"][gid][:kt], cols["tail_id"][gid][:kt]
            none = [None] * kt
            fields = [h, r, t,
                      none if ent is None else [ent.get(x) for x in h],
                      none if rel is None else [rel.get(x) for x in r],
                      none if ent is None else [ent.get(x) for x in t],
                      cols["score"][gid][:kt], cols["label"][gid][:kt], range(1, kt + 1)]
            if item.has_fwd:
                fields.append(cols["logit_fwd"][gid][:kt])
            if item.has_bwd:
                fields.append(cols["logit_bwd"][gid][:kt])
            rows = list(zip(*fields))
            # every k gets its own dict objects, like the reference's per-k list comprehension (:324-327)
            triplets_by_k = {k: [dict(zip(keys, row)) for row in rows[: min(k, kt)]] for k in ks}
            records.append({
                "sample_id": item.sample_ids[gid] if gid < len(item.sample_ids) else str(gid),
                "question": item.questions[gid] if gid < len(item.questions) else "",
                "triplets_by_k": triplets_by_k,
                "answer_entity_ids": item.answers[item.answer_ptr[gid]:item.answer_ptr[gid + 1]],
            })
        return records

    @staticmethod
    def _num_graphs(batch: Any, query_ids: torch.Tensor) -> int:
        for attr in ("num_graphs", "batch_size"):  # infer_batch_size, src/utils/logging_utils.py:90-101
            v = getattr(batch, attr, None)
            if v is not None and int(v) > 0:
                return int(v)
        ptr = getattr(batch, "ptr", None)
        if ptr is not None:
            return int(ptr.numel() - 1)
        return int(query_ids.max().item()) + 1 if query_ids.numel() else 0

    def _resolve_vocab_maps(self):
        """entity / relation parquet vocabularies ``(id, label)`` (:409-441)."""
        if not self.entity_vocab_path or not self.relation_vocab_path:
            raise ValueError("textualize=true requires both entity_vocab_path and relation_vocab_path.")
        ent_path, rel_path = Path(self.entity_vocab_path), Path(self.relation_vocab_path)
        if not ent_path.exists():
            raise FileNotFoundError(f"entity_vocab_path not found: {ent_path}")
        if not rel_path.exists():
            raise FileNotFoundError(f"relation_vocab_path not found: {rel_path}")
        import pyarrow.parquet as pq

        def load(path, id_col):
            tab = pq.read_table(path, columns=[id_col, "label"])
            return {int(i): str(l) for i, l in zip(tab.column(id_col).to_pylist(), tab.column("label").to_pylist())
                    if i is not None and l is not None}

        return load(ent_path, "entity_id"), load(rel_path, "relation_id")

    @staticmethod
    def _gather_records(records: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        """Rank 0 receives every rank's records in rank order (:443-456)."""
        if dist.is_available() and dist.is_initialized():
            world = dist.get_world_size()
            gathered: List[Optional[List[Dict[str, Any]]]] = [None for _ in range(world)]
            dist.all_gather_object(gathered, records)
            if dist.get_rank() != 0:
                return []
            return [r for part in gathered if part for r in part]
        return records


def load_artifact(output_dir, split: str = "test") -> Dict[str, Any]:
    """Read ``<split>.pt`` through its manifest the way the consumer does
    (src/data/reasoner_oracle_datamodule.py:85-171): manifest ``file`` relative to the directory."""
    output_dir = Path(output_dir)
    manifest = json.loads((output_dir / f"{split}.manifest.json").read_text(encoding="utf-8"))
    payload = torch.load(output_dir / manifest["file"], map_location="cpu", weights_only=False)
    return {"manifest": manifest, **payload}


__all__ = ["RetrieverTopKEdgeWriter", "normalize_k_values", "load_artifact"]
