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

from .topk import segmented_topk, writer_payload

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


class RetrieverTopKEdgeWriter:
    """Persist retriever top-k edges during predict/test (retriever_topk_edge_writer.py:68-170)."""

    def __init__(self, *, output_dir, split: str = "test", enabled: bool = True, artifact_name: str = "eval_retriever",
                 schema_version: int = 1, topk_values: Optional[Sequence[int]] = None, textualize: bool = False,
                 entity_vocab_path: Optional[str] = None, relation_vocab_path: Optional[str] = None,
                 overwrite: bool = True) -> None:
        self.enabled = bool(enabled)
        self.output_dir = Path(output_dir)
        self.split = str(split)
        self.artifact_name = str(artifact_name).strip()
        if not self.artifact_name:
            raise ValueError("artifact_name must be a non-empty string.")
        self.schema_version = int(schema_version)
        if self.schema_version <= 0:
            raise ValueError("schema_version must be a positive integer.")
        self.topk_values = normalize_k_values(topk_values, default=_DEFAULT_TOPK_VALUES)
        if not self.topk_values:
            raise ValueError("topk_values must be a non-empty list of positive integers.")
        self._max_topk = max(int(k) for k in self.topk_values)
        self.textualize = bool(textualize)
        self.entity_vocab_path = entity_vocab_path
        self.relation_vocab_path = relation_vocab_path
        self.overwrite = bool(overwrite)
        self._records: List[Dict[str, Any]] = []
        self._output_path: Optional[Path] = None
        self._manifest_path: Optional[Path] = None
        self._entity_map: Optional[Dict[int, str]] = None
        self._relation_map: Optional[Dict[int, str]] = None

    # ------------------------------------------------------------------ hooks
    def on_predict_start(self, trainer=None, pl_module=None) -> None:
        if not self.enabled:
            return
        self._records = []
        self.output_dir.mkdir(parents=True, exist_ok=True)
        self._output_path = self.output_dir / f"{self.split}.pt"
        self._manifest_path = self.output_dir / f"{self.split}.manifest.json"
        if self._output_path.exists() and not self.overwrite:
            raise FileExistsError(f"Output path already exists: {self._output_path}")
        if self.textualize:
            self._entity_map, self._relation_map = self._resolve_vocab_maps()

    def write_on_batch_end(self, trainer, pl_module, prediction, batch_indices, batch, batch_idx, dataloader_idx=0) -> None:
        self._collect_prediction(prediction=prediction, batch=batch)

    def on_predict_end(self, trainer=None, pl_module=None) -> None:
        if not self.enabled or not self._records:
            return
        records = self._gather_records(self._records)
        if not records:
            return
        if trainer is not None and int(getattr(trainer, "global_rank", 0)) != 0:
            return
        if self._output_path is None or self._manifest_path is None:
            raise RuntimeError("RetrieverTopKEdgeWriter missing output path; on_predict_start was not called.")
        payload = {"settings": {"split": self.split, "topk_values": [int(k) for k in self.topk_values]},
                   "samples": records}
        torch.save(payload, self._output_path)
        manifest = {
            "artifact": self.artifact_name,
            "schema_version": self.schema_version,
            "file": self._output_path.name,
            "created_at": datetime.now(timezone.utc).replace(tzinfo=None).isoformat(timespec="seconds") + "Z",
            "producer": "retriever_topk_edge_writer",
        }
        self._manifest_path.write_text(json.dumps(manifest, indent=2), encoding="utf-8")

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
        self._records.extend(self._build_records(batch=batch, output=prediction))

    def _build_records(self, *, batch: Any, output: Any) -> List[Dict[str, Any]]:
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
        slice_dict = getattr(batch, "_slice_dict", None)
        edge_ptr = None
        if isinstance(slice_dict, dict) and "edge_index" in slice_dict:
            cand = torch.as_tensor(slice_dict["edge_index"], dtype=torch.long).view(-1)
            if cand.numel() == B + 1:
                edge_ptr = cand.to(dev)
        if edge_ptr is None:
            if query_ids.numel() != logits.numel():
                raise ValueError(f"query_ids/scores mismatch: {query_ids.shape} vs {logits.shape}")
            if bool((query_ids[1:] < query_ids[:-1]).any()):
                raise ValueError("writer expects edges grouped by question (PyG collate order).")
            edge_ptr = torch.zeros(B + 1, dtype=torch.long, device=dev)
            edge_ptr[1:] = torch.cumsum(torch.bincount(query_ids, minlength=B)[:B], 0)
        answer_ptr_raw = getattr(batch, "answer_entity_ids_ptr", None)
        if answer_ptr_raw is None and isinstance(slice_dict, dict):
            answer_ptr_raw = slice_dict.get("answer_entity_ids")
        if answer_ptr_raw is None:
            raise ValueError("Batch missing answer_entity_ids_ptr required for persistence.")
        answer_ptr = torch.as_tensor(answer_ptr_raw, dtype=torch.long).view(-1).tolist()
        if len(answer_ptr) != B + 1:
            raise ValueError(f"answer_entity_ids_ptr length mismatch: {len(answer_ptr)} vs expected {B + 1}")
        answers = batch.answer_entity_ids.detach().cpu().view(-1).tolist()

        top = segmented_topk(logits, edge_ptr, self._max_topk, sigmoid=True)  # scores = sigmoid(logits) (:203)
        pay = writer_payload(top, edge_ptr, edge_index=batch.edge_index, node_global_ids=batch.node_global_ids,
                             edge_attr=batch.edge_attr, labels=labels,
                             logits_fwd=getattr(output, "logits_fwd", None), logits_bwd=getattr(output, "logits_bwd", None))
        host = {k: pay[k].cpu().tolist() for k in ("score", "count", "head_id", "tail_id", "relation_id", "label",
                                                   "logit_fwd", "logit_bwd")}
        has_fwd = getattr(output, "logits_fwd", None) is not None
        has_bwd = getattr(output, "logits_bwd", None) is not None
        sample_ids = _sample_ids(batch, B)
        questions = _questions(batch, B)
        records: List[Dict[str, Any]] = []
        for gid in range(B):
            kt = int(host["count"][gid])
            if kt <= 0:
                continue
            edges = [self._edge_record(host, gid, j, has_fwd, has_bwd) for j in range(kt)]
            triplets_by_k = {int(k): [dict(e) for e in edges[: min(int(k), kt)]] for k in self.topk_values}
            records.append({
                "sample_id": sample_ids[gid] if gid < len(sample_ids) else str(gid),
                "question": questions[gid] if gid < len(questions) else "",
                "triplets_by_k": triplets_by_k,
                "answer_entity_ids": [int(x) for x in answers[answer_ptr[gid]:answer_ptr[gid + 1]]],
            })
        return records

    def _edge_record(self, host, gid: int, j: int, has_fwd: bool, has_bwd: bool) -> Dict[str, Any]:
        h, r, t = int(host["head_id"][gid][j]), int(host["relation_id"][gid][j]), int(host["tail_id"][gid][j])
        rec = {
            "head_entity_id": h, "relation_id": r, "tail_entity_id": t,
            "head_text": self._lookup_text(self._entity_map, h),
            "relation_text": self._lookup_text(self._relation_map, r),
            "tail_text": self._lookup_text(self._entity_map, t),
            "score": float(host["score"][gid][j]), "label": float(host["label"][gid][j]), "rank": j + 1,
        }
        if has_fwd:
            rec["logit_fwd"] = float(host["logit_fwd"][gid][j])
        if has_bwd:
            rec["logit_bwd"] = float(host["logit_bwd"][gid][j])
        return rec

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

    @staticmethod
    def _lookup_text(mapping: Optional[Dict[int, str]], key: Optional[int]) -> Optional[str]:
        if mapping is None or key is None:
            return None
        return mapping.get(key)

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
