"""Import the UNMODIFIED reference from /root/reference (build container only).

Test infrastructure: used by ``oracle/make_golden.py`` and by the container-only test
``tests/test_oracle_vs_reference.py`` (skipped when /root/reference is absent, i.e. on the GPU
box). Two stand-ins are installed before import:

* ``torch_geometric.nn.MessagePassing`` — only ``aggr="mean"``, ``node_dim=0``,
  ``flow="source_to_target"`` as used by ``src/models/components/graph.py:13-23``;
  the real package is tried first.
* an empty ``lmdb`` module — ``scripts/build_retrieval_pipeline.py:36`` imports it at top level
  but the BFS functions never touch it.
"""
from __future__ import annotations

import os
import sys
import types

REFERENCE_ROOT = os.environ.get("EVI_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "models", "components"))


def _install_pyg_shim() -> str:
    try:
        import torch_geometric  # noqa: F401

        return "torch_geometric"
    except Exception:
        pass
    import torch

    class MessagePassing(torch.nn.Module):
        def __init__(self, aggr: str = "mean", node_dim: int = 0, **_):
            super().__init__()
            if aggr != "mean" or node_dim != 0:
                raise NotImplementedError("shim covers aggr='mean', node_dim=0 only")

        def propagate(self, edge_index, x):
            src, dst = edge_index[0], edge_index[1]
            msg = self.message(x[src])
            out = torch.zeros_like(x)
            out.index_add_(0, dst, msg)
            deg = torch.zeros(x.size(0), dtype=x.dtype, device=x.device)
            deg.index_add_(0, dst, torch.ones_like(dst, dtype=x.dtype))
            return out / deg.clamp(min=1).unsqueeze(-1)

        def message(self, x_j):
            return x_j

    tg = types.ModuleType("torch_geometric")
    tgnn = types.ModuleType("torch_geometric.nn")
    tgnn.MessagePassing = MessagePassing
    tg.nn = tgnn
    sys.modules["torch_geometric"] = tg
    sys.modules["torch_geometric.nn"] = tgnn
    return "shim"


def load_reference():
    """Returns ``(Retriever, compute_edge_batch, brp_module, pyg_kind)`` from the reference."""
    if not reference_available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    kind = _install_pyg_shim()
    if "lmdb" not in sys.modules:
        try:
            import lmdb  # noqa: F401
        except Exception:
            sys.modules["lmdb"] = types.ModuleType("lmdb")
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from src.models.components.retriever import Retriever
    from src.utils.graph_utils import compute_edge_batch
    import scripts.build_retrieval_pipeline as brp

    return Retriever, compute_edge_batch, brp, kind
