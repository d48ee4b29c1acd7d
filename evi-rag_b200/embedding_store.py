"""Device-resident embedding feed ((f) rank 1): what replaces ``GlobalEmbeddingStore`` + the embedding half of
``UnifiedDataLoader.__iter__`` when the retriever runs on this package's kernels.

Reference (``src/data/components/embedding_store.py:12-150``, ``src/data/components/loader.py:166-185``): the two
tables ``entity_embeddings.pt [T+1, 1024]`` (row 0 = non-text) and ``relation_embeddings.pt [R, 1024]`` are loaded
(memory-mapped) on the host; every batch does a CPU ``index_select`` of ``node_embedding_ids`` / ``edge_attr`` into
pinned buffers and ships ``[N, 1024] + [E, 1024]`` fp32 rows over PCIe (655 MB for a 32-question batch).

Here the tables are converted once, slab by slab, into the ``[hi | lo]`` 16-bit rows the projection kernels read
(``Retriever.attach_embedding_tables``), a batch carries ids only, and the row gather is ``evi_gather_rows2`` on the
device. The host-side getters keep the reference's signatures for callers that want raw rows (they are plain data
movement: ``index_select`` on the memory-mapped tensors)."""
from __future__ import annotations

from pathlib import Path
from typing import Any, Iterable, Iterator, Optional, Union

import torch

ENTITY_FILE, RELATION_FILE = "entity_embeddings.pt", "relation_embeddings.pt"


def _load_table(path: Path, mmap: bool) -> torch.Tensor:
    if not path.exists():
        raise FileNotFoundError(f"Embedding file missing: {path}")
    if mmap:
        try:  # zipfile-format checkpoints map lazily; legacy pickles fall back to a full read
            t = torch.load(path, map_location="cpu", mmap=True, weights_only=True)
        except (TypeError, RuntimeError, ValueError):
            t = torch.load(path, map_location="cpu", weights_only=True)
    else:
        t = torch.load(path, map_location="cpu", weights_only=True)
    if not torch.is_tensor(t) or t.dim() != 2:
        raise ValueError(f"{path} must hold one 2-D tensor, got {type(t).__name__}")
    return t


class DeviceEmbeddingStore:
    """Holds the two tables (host, possibly memory-mapped) and puts them into a retriever's HBM."""

    def __init__(self, entity_embeddings: torch.Tensor, relation_embeddings: torch.Tensor,
                 embeddings_dir: Optional[Union[str, Path]] = None) -> None:
        if entity_embeddings.dim() != 2 or relation_embeddings.dim() != 2:
            raise ValueError("embedding tables must be 2-D")
        if entity_embeddings.size(1) != relation_embeddings.size(1):
            raise ValueError(f"entity dim {entity_embeddings.size(1)} != relation dim {relation_embeddings.size(1)}")
        self.entity_embeddings = entity_embeddings
        self.relation_embeddings = relation_embeddings
        self.embeddings_dir = Path(embeddings_dir) if embeddings_dir is not None else None

    @classmethod
    def from_dir(cls, embeddings_dir: Union[str, Path], *, mmap: bool = True) -> "DeviceEmbeddingStore":
        """``embeddings_dir/{entity,relation}_embeddings.pt`` as written by scripts/text_encode_utils.py:125-134."""
        d = Path(embeddings_dir)
        return cls(_load_table(d / ENTITY_FILE, mmap), _load_table(d / RELATION_FILE, mmap), d)

    # ---- the device side -------------------------------------------------------------------------
    def attach(self, model: Any, *, split: bool = True, slab_rows: int = 1 << 20) -> Any:
        """Convert and upload both tables into ``model`` (an ``evi_rag_b200.Retriever`` already on its CUDA device),
        ``slab_rows`` rows at a time so a memory-mapped 100 GB table never needs a host copy. Returns ``model``."""
        model.attach_embedding_tables(self.entity_embeddings, self.relation_embeddings, split=split, slab_rows=slab_rows)
        return model

    def device_bytes(self, *, split: bool = True) -> int:
        rows = int(self.entity_embeddings.size(0)) + int(self.relation_embeddings.size(0))
        return rows * int(self.entity_dim) * (4 if split else 2)

    def ids_only(self, batches: Iterable[Any]) -> Iterator[Any]:
        """Stand-in for ``UnifiedDataLoader.__iter__`` (loader.py:166-185): yields the loader's batches WITHOUT
        attaching ``node_embeddings`` / ``edge_embeddings`` (and drops them if a collate worker attached them), so
        only ids cross PCIe. ``node_embedding_ids`` / ``edge_attr`` are normalised to int64 tensors like the
        reference does before its lookup."""
        for batch in batches:
            for name in ("node_embeddings", "edge_embeddings"):
                if hasattr(batch, name):
                    delattr(batch, name)
            batch.node_embedding_ids = torch.as_tensor(batch.node_embedding_ids, dtype=torch.long)
            batch.edge_attr = torch.as_tensor(batch.edge_attr, dtype=torch.long)
            yield batch

    # ---- reference-shaped host getters (embedding_store.py:101-150) -------------------------------
    @staticmethod
    def _rows(table: torch.Tensor, ids: torch.Tensor, device: Optional[torch.device]) -> torch.Tensor:
        target = device if device is not None else ids.device
        cpu_ids = ids.detach().to("cpu", dtype=torch.long)
        if cpu_ids.numel() == 0:
            return torch.empty((0, int(table.size(1))), dtype=table.dtype, device=target)
        return table.index_select(0, cpu_ids).to(target, non_blocking=True)

    def get_entity_embeddings(self, entity_ids: torch.Tensor, *, device: Optional[torch.device] = None) -> torch.Tensor:
        return self._rows(self.entity_embeddings, entity_ids, device)

    def get_relation_embeddings(self, relation_ids: torch.Tensor, *, device: Optional[torch.device] = None) -> torch.Tensor:
        return self._rows(self.relation_embeddings, relation_ids, device)

    def clear_device_cache(self) -> None:  # nothing transient is kept
        return None

    @property
    def entity_dim(self) -> int:
        return int(self.entity_embeddings.size(-1))

    @property
    def relation_dim(self) -> int:
        return int(self.relation_embeddings.size(-1))


__all__ = ["DeviceEmbeddingStore"]
