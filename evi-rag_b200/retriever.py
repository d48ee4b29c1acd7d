"""Drop-in ``Retriever`` for EVI-RAG's retrieval hot path, computed by sm_100a kernels.

Mirrors ``src/models/components/retriever.py`` of the reference: same constructor keywords and
validation (:105-183, :555-570), same ``state_dict`` keys and shapes (so reference checkpoints load
with ``strict=True`` and ``GraphEmbedder`` can still read them by suffix,
``src/models/components/gflownet_embedder.py:489-533``), same duck-typed batch attributes
(:201-204, :419-433, :526-545, :580-584), same ``RetrieverOutput`` fields (:80-99) and the same
``ValueError`` conditions. A Hydra config switches over by changing one ``_target_``
(``configs/model/retriever_module.yaml:8``).

What differs (B200 design):
  * all arithmetic runs in ``libevi_rag_b200.so``; there is no eager-PyTorch / CPU fallback;
  * dense contractions use 16-bit tensor-core operands with fp32 accumulation
    (``operand_dtype="fp16"`` by default, ``"bf16"`` selectable). The per-node / per-relation /
    per-question contractions run over ``[hi | lo]`` operand pairs (three tensor passes, fp32-grade);
    the per-edge ones over single 16-bit rows with the structure context centred on its batch mean
    (``precision="fast"``, scores within ~3e-4 of the fp32 reference) or over pairs as well
    (``precision="strict"``, ~2e-5, about twice the time);
  * entity rows are projected once per row of the attached entity table and checkpoint
    (``project_entity_table``), or once per distinct ``node_embedding_id`` of the batch when that
    table would not fit;
  * with ``attach_embedding_tables`` the 1024-d text-embedding tables live in HBM and a batch only
    needs ids (``node_embedding_ids``, ``edge_attr``) — the reference's per-edge CPU gather + H2D
    of ``edge_embeddings`` (``src/data/components/embedding_store.py:101-150``) disappears;
  * ``edge_embeddings`` ([E, H] fp32 features) are produced only when asked
    (``return_edge_embeddings=True`` or ``extract_edge_tokens``): ``RetrieverModule`` drops them
    in eval (``src/models/retriever_module.py:447-450``).
"""
from __future__ import annotations

import os

import ctypes as C
from dataclasses import dataclass
from typing import Any, Dict, Optional

import torch
from torch import nn

from . import lib as _lib

_NUM_TOPICS_BINARY = 2
_MAX_DDE_ROUNDS = 4
_FP16_SAFE_MAX = 6.0e4


@dataclass
class RetrieverOutput:
    """Same container as the reference (retriever.py:80-99)."""

    logits: torch.Tensor
    query_ids: torch.Tensor
    relation_ids: Optional[torch.Tensor] = None
    logits_fwd: Optional[torch.Tensor] = None
    logits_bwd: Optional[torch.Tensor] = None
    edge_embeddings: Optional[torch.Tensor] = None

    def detach(self) -> "RetrieverOutput":
        d = lambda t: t.detach() if t is not None else None
        return RetrieverOutput(
            logits=self.logits.detach(),
            query_ids=self.query_ids.detach(),
            relation_ids=d(self.relation_ids),
            logits_fwd=d(self.logits_fwd),
            logits_bwd=d(self.logits_bwd),
            edge_embeddings=d(self.edge_embeddings),
        )


class EmbeddingProjector(nn.Module):
    """Linear + Tanh holder with the reference's parameter names (projections.py:9-40)."""

    def __init__(self, output_dim: int, *, input_dim: int) -> None:
        super().__init__()
        self.network = nn.Sequential(nn.Linear(input_dim, output_dim), nn.Tanh())


class _DDEConfig(nn.Module):
    """Round counts of the reference DDE (graph.py:26-39); the propagation itself is a kernel."""

    def __init__(self, num_rounds: int = 2, num_reverse_rounds: int = 2) -> None:
        super().__init__()
        self.num_rounds = int(max(0, num_rounds))
        self.num_reverse_rounds = int(max(0, num_reverse_rounds))
        if self.num_rounds > _MAX_DDE_ROUNDS or self.num_reverse_rounds > _MAX_DDE_ROUNDS:
            raise ValueError(
                f"DDE supports at most {_MAX_DDE_ROUNDS} rounds per direction; "
                f"got num_rounds={self.num_rounds}, num_reverse_rounds={self.num_reverse_rounds}."
            )


class Retriever(nn.Module):
    """Retriever with the reference interface. ``eval()`` / ``no_grad`` runs the fused inference kernels; in training
    mode with autograd enabled ``forward`` builds the reference's graph with tensor-core contractions (training.py)."""

    def __init__(
        self,
        emb_dim: int,
        hidden_dim: int,
        topic_pe: bool = True,
        num_topics: int = 2,
        dde_cfg: Optional[Dict[str, int]] = None,
        dropout_p: float = 0.1,
        core_mode: str = "geometry",
        direction_mode: str = "bidirectional",
        hide_seek_cfg: Optional[Dict[str, Any]] = None,
        *,
        operand_dtype: str = "fp16",
        precision: str = "fast",
        unique_entity_rows: bool = True,
        project_entity_table: Optional[bool] = None,
        projected_table_max_bytes: int = 16 << 30,
        center_structure_context: bool = True,
        return_edge_embeddings: bool = False,
        max_workspace_bytes: int = 12 << 30,
        pipeline_chunks: int = 4,
        overlap_streams: bool = False,
        validate: bool = True,
        **_: Any,
    ) -> None:
        super().__init__()
        self.emb_dim = int(emb_dim)
        self.hidden_dim = int(hidden_dim)
        self.use_topic_pe = bool(topic_pe)
        if not self.use_topic_pe:
            raise ValueError("topic_pe must be enabled; retriever requires topic_one_hot + DDE.")
        self.num_topics = int(num_topics)
        if self.num_topics != _NUM_TOPICS_BINARY:
            raise ValueError(f"num_topics must be {_NUM_TOPICS_BINARY} (seed vs non-seed), got {self.num_topics}")
        _ = self._normalize_core_mode(core_mode)
        self.direction_mode = self._normalize_direction_mode(direction_mode)
        if operand_dtype not in ("fp16", "bf16"):
            raise ValueError(f"operand_dtype must be 'fp16' or 'bf16', got {operand_dtype!r}")
        self.operand_dtype = operand_dtype
        if precision not in ("fast", "strict"):
            raise ValueError(f"precision must be 'fast' or 'strict', got {precision!r}")
        self.precision = precision
        self.unique_entity_rows = bool(unique_entity_rows)
        # entity rows projected once per TABLE row and checkpoint instead of once per batch (None: when the projected
        # table fits `projected_table_max_bytes` of HBM) -- see _projected_entity_table
        self.project_entity_table = None if project_entity_table is None else bool(project_entity_table)
        self.projected_table_max_bytes = int(projected_table_max_bytes)
        self.center_structure_context = bool(center_structure_context)
        self.return_edge_embeddings = bool(return_edge_embeddings)
        self.max_workspace_bytes = int(max_workspace_bytes)
        self.pipeline_chunks = max(1, int(pipeline_chunks))
        # off by default: measured on B200 the step is power-capped, overlapping the tensor-bound and HBM-bound stages
        # lowers the clocks and leaves the step time unchanged (profiles/r01i_overlap_ab.md)
        self.overlap_streams = bool(overlap_streams) or os.environ.get("EVI_OVERLAP", "0") == "1"
        self._aux_stream = None
        self.validate = bool(validate)

        d, h = self.emb_dim, self.hidden_dim
        self.entity_proj = EmbeddingProjector(d, input_dim=d)
        self.relation_proj = EmbeddingProjector(d, input_dim=d)
        self.query_proj = EmbeddingProjector(d, input_dim=d)
        self.non_text_entity_emb = nn.Embedding(1, d)
        self.dde = _DDEConfig(**(dde_cfg or {}))
        num_rounds, num_rev = self.dde.num_rounds, self.dde.num_reverse_rounds
        self._topic_struct_dim = self.num_topics * (1 + num_rounds + num_rev)
        struct_raw_dim = 2 * self._topic_struct_dim
        self.register_buffer(
            "parity_meta",
            torch.tensor([int(self.use_topic_pe), self.num_topics, num_rounds, num_rev], dtype=torch.long),
        )
        self.q_gate = nn.Sequential(nn.Linear(d, d), nn.Sigmoid())
        self.q_bias = nn.Sequential(nn.Linear(d, d), nn.Tanh())
        self.struct_proj = nn.Sequential(nn.Linear(struct_raw_dim, d), nn.LayerNorm(d), nn.GELU())
        self.struct_gate_net = nn.Sequential(nn.Linear(d, 1), nn.Sigmoid())
        self.state_net = nn.Sequential(
            nn.Linear(3 * d + 1, h), nn.LayerNorm(h), nn.GELU(), nn.Dropout(dropout_p), nn.Linear(h, h)
        )
        self.score_head = nn.Linear(h, 1)
        self._init_hide_seek(hide_seek_cfg)

        self._prepared: Optional[Dict[str, Any]] = None
        self._prepared_key = None
        self._tables: Optional[Dict[str, torch.Tensor]] = None
        self._slot_map: Optional[torch.Tensor] = None
        self._ptable: Optional[Dict[str, Any]] = None
        self._ws: Optional[torch.Tensor] = None
        self._ws_score: Optional[torch.Tensor] = None
        self.gemm_events = None
        self.tail_events = None
        self.dde_events = None
        self.last_counts = None
        self.last_chunk_edges = 0
        self.last_num_chunks = 0
        self._flags: Optional[torch.Tensor] = None
        self._pending_flags = False

    # ------------------------------------------------------------------ reference API
    def forward(self, batch: Any) -> RetrieverOutput:
        output, _ = self._forward_impl(batch, return_features=self.return_edge_embeddings)
        return output

    def extract_edge_tokens(self, batch: Any) -> torch.Tensor:
        _, features = self._forward_impl(batch, return_features=True)
        if features is None:
            raise RuntimeError("extract_edge_tokens expected non-empty features but got None.")
        return features

    # ------------------------------------------------------------------ B200 additions
    def attach_embedding_tables(self, entity_embeddings: torch.Tensor, relation_embeddings: torch.Tensor,
                                *, split: bool = True, slab_rows: int = 1 << 20) -> None:
        """Keep the text-embedding tables in HBM. ``entity_embeddings`` is ``[max_id+1, D]`` with row 0
        reserved for non-text entities and ``relation_embeddings`` ``[R, D]``
        (scripts/text_encode_utils.py:60-67, embedding_store.py:60-99). Batches then only need
        ``node_embedding_ids`` and ``edge_attr``.

        ``split=True`` (default) stores every row as a ``[hi | lo]`` pair of 16-bit halves (4 KB per row;
        the projections then see the fp32 table values to ~2^-22); ``split=False`` stores 2 KB per row
        (for tables that would not fit otherwise). Tables may be CPU (also memory-mapped) tensors: they
        are converted and copied ``slab_rows`` at a time."""
        dev = self._device()
        if dev.type != "cuda":
            raise RuntimeError("evi_rag_b200.Retriever runs on CUDA only (no CPU fallback); call .cuda() first.")
        dt = self._op_dtype()
        if entity_embeddings.dim() != 2 or entity_embeddings.size(1) != self.emb_dim:
            raise ValueError(f"entity_embeddings must be [T, {self.emb_dim}], got {tuple(entity_embeddings.shape)}")
        if relation_embeddings.dim() != 2 or relation_embeddings.size(1) != self.emb_dim:
            raise ValueError(f"relation_embeddings must be [R, {self.emb_dim}], got {tuple(relation_embeddings.shape)}")
        d = self.emb_dim

        def upload(name: str, t: torch.Tensor) -> torch.Tensor:
            out = torch.empty((t.size(0), 2 * d if split else d), dtype=dt, device=dev)
            for r0 in range(0, t.size(0), slab_rows):
                x = t[r0:r0 + slab_rows].to(device=dev, dtype=torch.float32)
                if dt == torch.float16 and x.numel() and float(x.abs().max()) > _FP16_SAFE_MAX:
                    raise ValueError(f"{name} embeddings exceed the fp16 range; use operand_dtype='bf16'.")
                hi = x.to(dt)
                out[r0:r0 + slab_rows, :d] = hi
                if split:
                    out[r0:r0 + slab_rows, d:] = (x - hi.float()).to(dt)
            return out

        self._tables = {"entity": upload("entity", entity_embeddings), "relation": upload("relation", relation_embeddings),
                        "split": bool(split)}
        # evi_unique_rows: one int32 per table row, all -1 between calls
        self._slot_map = torch.full((int(entity_embeddings.size(0)),), -1, dtype=torch.int32, device=dev)
        self._ptable = None  # projected rows of the previous table

    def uses_projected_table(self) -> bool:
        """True when the attached entity table is projected once per table row (``project_entity_table``; by default
        when the projected rows, 8 KB per table row, fit ``projected_table_max_bytes``)."""
        if self._tables is None:
            return False
        if self.project_entity_table is not None:
            return self.project_entity_table
        return int(self._tables["entity"].size(0)) * (4 * self.emb_dim + 4 * self.hidden_dim) <= self.projected_table_max_bytes

    def invalidate_projected_table(self) -> None:
        """Forget the projected entity rows (their buffers are kept): the next forward projects the table again.
        bench.py calls this at the start of every timed region so that the projection is paid inside it."""
        if self._ptable is not None:
            self._ptable["key"] = None

    def _projected_entity_table(self, p: Dict[str, Any], dev: torch.device, slab_rows: int = 1 << 18):
        """``node_repr = tanh(W_e x + b_e)`` as [hi | lo] rows and ``P = W_err node_repr`` for EVERY row of the attached
        entity table (row 0 = the learned non-text embedding, retriever.py:497-507), computed by the same kernels and
        therefore bit-identical to the per-batch rows. Both depend on the table and the checkpoint only, so an eval run
        projects each entity once instead of once per batch it appears in (the reference re-projects every node of
        every batch, retriever.py:441). Rebuilt when a parameter or the table changes; 8 KB of HBM per table row."""
        ent = self._tables["entity"]
        key = (self._prepared_key, ent.data_ptr(), int(ent.size(0)))
        pt = self._ptable
        if pt is not None and pt["key"] == key:
            return pt["repr"], pt["p"]
        L = _lib.load()
        d, dt, fmt = self.emb_dim, self._op_dtype(), self._fmt()
        T1 = int(ent.size(0))
        tsplit = bool(self._tables["split"])
        if pt is None or pt["repr"].size(0) != T1 or pt["repr"].device != dev or pt["repr"].dtype != dt:
            pt = {"repr": torch.empty((T1, 2 * d), dtype=dt, device=dev),
                  "p": torch.empty((T1, self.hidden_dim), dtype=torch.float32, device=dev)}
        stream = _lib.stream_handle(dev)
        tanh4, none4 = [_lib.ACT_TANH] * 4, [_lib.ACT_NONE] * 4
        x = torch.empty((min(slab_rows, T1), 2 * d if tsplit else d), dtype=dt, device=dev)
        for r0 in range(0, T1, slab_rows):
            n = min(slab_rows, T1 - r0)
            ids = torch.arange(r0, r0 + n, dtype=torch.long, device=dev)
            _lib.check(L.evi_gather_rows2(ent.data_ptr(), 0, T1, int(ent.stride(0)), ids.data_ptr(), ids.data_ptr(),
                                          p["non_text"].data_ptr(), n, None, d, x.data_ptr(), int(x.stride(0)),
                                          int(tsplit), fmt, self._flags.data_ptr(), stream), "evi_gather_rows2(table)")
            rows = pt["repr"][r0:r0 + n]
            self._project(x[:n], p["we"], p["be"], tanh4, out16=rows, out_lo=d, p=p, x_lo=d if tsplit else -1, w_lo=d)
            self._project(rows, p["w1"], None, none4, out32=pt["p"][r0:r0 + n], p=p, col0=2 * d, in_dim=d, x_lo=d,
                          w_lo=3 * d)
        pt["key"] = key
        self._ptable = pt
        return pt["repr"], pt["p"]

    def check_flags(self) -> None:
        """Read back the device-side validation word (one host sync) and raise like the
        reference (graph_utils.py:66-99, retriever.py:594-604)."""
        if not self._pending_flags or self._flags is None:
            return
        self._pending_flags = False
        word = int(self._flags.item())
        if word == 0:
            return
        self._flags.zero_()
        self.raise_for_flags(word)

    @staticmethod
    def raise_for_flags(word: int) -> None:
        """The reference's ``ValueError`` for a non-zero validation word (``EVI_FLAG_*`` bits)."""
        if word & _lib.FLAG_OUT_OF_RANGE:
            raise ValueError("Invalid edge_index range; edge_index / embedding ids out of range for this batch.")
        if word & _lib.FLAG_CROSS_GRAPH:
            raise ValueError("edge_index crosses graph boundaries; head/tail graph assignments differ. ")
        if word & _lib.FLAG_NOT_MONOTONE:
            raise ValueError(
                "edge_batch is not non-decreasing along the flattened edge list, which breaks per-graph slicing; "
                "Ensure edges are concatenated per-graph (PyG Batch)."
            )

    # ------------------------------------------------------------------ internals
    def _device(self) -> torch.device:
        return next(self.parameters()).device

    def _fmt(self) -> int:
        return _lib.FMT_F16 if self.operand_dtype == "fp16" else _lib.FMT_BF16

    def _op_dtype(self) -> torch.dtype:
        return torch.float16 if self.operand_dtype == "fp16" else torch.bfloat16

    def _prepare_weights(self) -> Dict[str, Any]:
        """16-bit / fp32 kernel-side copies of the parameters; rebuilt when any parameter changes."""
        params = list(self.parameters())
        key = (self.operand_dtype,) + tuple((p.data_ptr(), p._version) for p in params)
        if self._prepared is not None and self._prepared_key == key:
            return self._prepared
        dev = self._device()
        if dev.type != "cuda":
            raise RuntimeError("evi_rag_b200.Retriever runs on CUDA only (no CPU fallback); call .cuda() first.")
        if self.emb_dim != 1024 or self.hidden_dim != 1024:
            raise ValueError("the sm_100a kernels are built for emb_dim == hidden_dim == 1024")
        dt = self._op_dtype()
        d = self.emb_dim

        def w16(t: torch.Tensor) -> torch.Tensor:
            t = t.detach().to(device=dev, dtype=torch.float32)
            if dt == torch.float16 and float(t.abs().max()) > _FP16_SAFE_MAX:
                raise ValueError("weights exceed the fp16 range; construct with operand_dtype='bf16'.")
            return t.to(dt).contiguous()

        def w16_split(t: torch.Tensor) -> torch.Tensor:
            """[out, 2 in] = [hi | lo], lo = W - hi: the second half of a three-pass contraction (gemm::SplitK)."""
            t = t.detach().to(device=dev, dtype=torch.float32)
            hi = w16(t)
            return torch.cat([hi, (t - hi.float()).to(dt)], dim=1).contiguous()

        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        sn0 = self.state_net[0]
        sn4 = self.state_net[4]
        wo = self.score_head.weight.detach().to(dev).double()[0]
        p: Dict[str, Any] = {
            "we": w16_split(self.entity_proj.network[0].weight), "be": f32(self.entity_proj.network[0].bias),
            "wr": w16_split(self.relation_proj.network[0].weight), "br": f32(self.relation_proj.network[0].bias),
            "wq": w16_split(self.query_proj.network[0].weight), "bq": f32(self.query_proj.network[0].bias),
            "wgb": w16_split(torch.cat([self.q_gate[0].weight, self.q_bias[0].weight], dim=0)),
            "bgb": f32(torch.cat([self.q_gate[0].bias, self.q_bias[0].bias], dim=0)),
            "non_text": f32(self.non_text_entity_emb.weight[0]),
            "ws_t": f32(self.struct_proj[0].weight.t()), "bs": f32(self.struct_proj[0].bias),
            "ls_g": f32(self.struct_proj[1].weight), "ls_b": f32(self.struct_proj[1].bias),
            "wn": f32(self.struct_gate_net[0].weight[0]), "bn": float(self.struct_gate_net[0].bias.detach()[0]),
            # [H, 6D] = [iv | sc | err | iv_lo | sc_lo | err_lo]
            "w1": w16_split(sn0.weight[:, : 3 * d]), "w1d": f32(sn0.weight[:, 3 * d]), "b1": f32(sn0.bias),
            "w_sc32": f32(sn0.weight[:, d: 2 * d]),
            "l1_g": f32(self.state_net[1].weight), "l1_b": f32(self.state_net[1].bias),
            "w2": w16_split(sn4.weight), "b2": f32(sn4.bias),
            "v": (sn4.weight.detach().to(dev).double().t() @ wo).float().contiguous(),
            "c0": float(wo @ sn4.bias.detach().to(dev).double() + self.score_head.bias.detach().to(dev).double()[0]),
            "wo": f32(self.score_head.weight[0]), "bo": float(self.score_head.bias.detach()[0]),
            "num_sms": torch.cuda.get_device_properties(dev).multi_processor_count,
        }
        self._prepared, self._prepared_key = p, key
        return p

    def _workspace(self, nbytes: int, dev: torch.device) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != dev:
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
        return self._ws

    def _score_workspace(self, nbytes: int, dev: torch.device) -> torch.Tensor:
        if self._ws_score is None or self._ws_score.numel() < nbytes or self._ws_score.device != dev:
            self._ws_score = None
            self._ws_score = torch.empty(int(nbytes), dtype=torch.uint8, device=dev)
        return self._ws_score

    def _project(self, x16, w16, bias, acts, *, p, out16=None, out32=None, rows_dev=None, col0=0, in_dim=None,
                 x_lo=-1, w_lo=-1, out_lo=-1):
        """out = act(x16[:, :in_dim] @ w16[:, col0:col0+in_dim]^T + bias). ``x_lo`` / ``w_lo``: column offset of the
        residual half of x16 / w16 rows (three-pass contraction, -1 = none; ``w_lo`` is relative to ``col0``);
        ``out_lo``: write out16 as [hi | lo] rows with lo at this column; ``rows_dev``: device int bounding the rows."""
        L = _lib.load()
        act_arr = (C.c_int * 8)(*(list(acts) + [0] * (8 - len(acts))))
        if in_dim is None:
            in_dim = int(w16.size(1)) if w_lo < 0 else int(w_lo)
        st = L.evi_project_split(
            x16.data_ptr(), int(x16.stride(0)), int(x_lo), x16.size(0), _lib.ptr(rows_dev),
            w16.data_ptr() + 2 * int(col0), int(w16.stride(0)), int(w_lo), w16.size(0), int(in_dim), _lib.ptr(bias),
            act_arr, _lib.ptr(out16), int(out16.stride(0)) if out16 is not None else 0, int(out_lo), _lib.ptr(out32),
            self._fmt(), p["num_sms"], _lib.stream_handle(x16.device),
        )
        _lib.check(st, "evi_project_split")

    def _forward_impl(self, batch: Any, *, return_features: bool):
        if self.training:
            # training_step (retriever_module.py:290-334): the reference's per-edge formulation under autograd, every
            # contraction (forward and backward) on the tcgen05 GEMM over [hi | lo] pairs — evi_rag_b200/training.py
            from .training import forward_train

            return forward_train(self, batch, return_features)
        with torch.no_grad():
            return self._forward_eval(batch, return_features=return_features)

    def _forward_eval(self, batch: Any, *, return_features: bool):
        edge_index = getattr(batch, "edge_index", None)
        if edge_index is None:
            raise ValueError("Batch missing edge_index required for scoring.")
        dev = self._device()
        if dev.type != "cuda":
            raise RuntimeError("evi_rag_b200.Retriever runs on CUDA only (no CPU fallback); call .cuda() first.")
        # the C ABI launches on the calling thread's current device: make it the device that owns the parameters
        # (the reference module is plain torch and works from any current device)
        with torch.cuda.device(dev):
            return self._forward_on_device(batch, edge_index, dev, return_features)

    def _forward_on_device(self, batch: Any, edge_index: torch.Tensor, dev: torch.device, return_features: bool):
        edge_index = edge_index.to(device=dev)
        if edge_index.dim() != 2 or edge_index.size(0) != 2:
            raise ValueError(f"edge_index must have shape [2, E], got {tuple(edge_index.shape)}")
        if edge_index.size(1) == 0:
            return self._empty_output(head_idx=edge_index[0], return_features=return_features)
        edge_index = edge_index.to(torch.long).contiguous()
        E = int(edge_index.size(1))

        question_emb = getattr(batch, "question_emb", None)
        node_embedding_ids = getattr(batch, "node_embedding_ids", None)
        edge_attr = getattr(batch, "edge_attr", None)
        if question_emb is None or node_embedding_ids is None or edge_attr is None:
            raise ValueError("Batch must provide question_emb, node_embedding_ids, and edge_attr.")
        node_embeddings = getattr(batch, "node_embeddings", None)
        edge_embeddings = getattr(batch, "edge_embeddings", None)
        use_tables = node_embeddings is None or edge_embeddings is None
        if use_tables and self._tables is None:
            raise ValueError("Batch must provide node_embeddings and edge_embeddings.")

        L = _lib.load()
        p = self._prepare_weights()
        stream = _lib.stream_handle(dev)
        fmt, dt = self._fmt(), self._op_dtype()
        d = self.emb_dim
        strict = self.precision == "strict"
        if self._flags is None or self._flags.device != dev:
            self._flags = torch.zeros(1, dtype=torch.int32, device=dev)
        flags = self._flags
        i32 = lambda n: torch.empty(n, dtype=torch.int32, device=dev)

        question_emb = question_emb.to(device=dev, dtype=torch.float32).contiguous()
        if not torch.is_tensor(node_embedding_ids):
            node_embedding_ids = torch.as_tensor(node_embedding_ids, dtype=torch.long)
        node_embedding_ids = node_embedding_ids.to(device=dev, dtype=torch.long).contiguous()
        edge_attr_dev = edge_attr.to(device=dev, dtype=torch.long).contiguous() if torch.is_tensor(edge_attr) else \
            torch.as_tensor(edge_attr, dtype=torch.long, device=dev)
        B = int(question_emb.size(0))
        N = int(node_embedding_ids.numel())

        # ---- a1/a2: question id per edge (+ int32 copies for the kernels) ------------------
        graph = getattr(batch, "_evi_graph", None)  # a precomputed evi_rag_b200.graph.GraphBatch
        edge_batch = getattr(batch, "edge_batch", None)
        sorted_by_kernel = False  # edges known to be grouped by question (validated on the device)
        if graph is not None and graph.num_edges == E:
            head32, tail32, qid32, query_ids = graph.head32, graph.tail32, graph.qid32, graph.edge_batch
            sorted_by_kernel = True
            if edge_batch is None:
                batch.edge_batch = query_ids
        elif edge_batch is None:
            head32, tail32, qid32 = i32(E), i32(E), i32(E)
            node_ptr = getattr(batch, "ptr", None)
            if node_ptr is None:
                raise ValueError("Batch missing ptr required for edge_batch computation.")
            node_ptr = node_ptr.to(device=dev, dtype=torch.long).contiguous()
            num_graphs = int(node_ptr.numel() - 1)
            if num_graphs <= 0:
                raise ValueError(f"num_graphs must be positive, got {num_graphs}")
            query_ids = torch.empty(E, dtype=torch.long, device=dev)
            st = L.evi_edge_batch(edge_index.data_ptr(), E, node_ptr.data_ptr(), num_graphs, query_ids.data_ptr(),
                                  qid32.data_ptr(), head32.data_ptr(), tail32.data_ptr(), None, flags.data_ptr(),
                                  stream)
            _lib.check(st, "evi_edge_batch")
            self._pending_flags = True
            sorted_by_kernel = True
            batch.edge_batch = query_ids  # retriever.py:612
        else:
            head32, tail32, qid32 = i32(E), i32(E), i32(E)
            if not torch.is_tensor(edge_batch):
                edge_batch = torch.as_tensor(edge_batch, dtype=torch.long)
            query_ids = edge_batch.to(device=dev, dtype=torch.long).view(-1)
            if query_ids.numel() != E:
                raise ValueError(f"edge_batch length mismatch: {query_ids.numel()} vs edges {E}")
            head32.copy_(edge_index[0])
            tail32.copy_(edge_index[1])
            qid32.copy_(query_ids)

        # ---- a4: projections (tanh(xW^T+b)) at their true granularity ---------------------
        # entity rows once per distinct embedding id of the batch (device-side count: buffers are sized by N)
        rows_dev = None
        head_rep32 = tail_rep32 = None
        rel_slot32 = i32(E)
        tanh4 = [_lib.ACT_TANH] * 4
        use_ptable = use_tables and self.uses_projected_table()
        rep_rows = N
        if use_ptable:
            # rows of the projected TABLE: an edge endpoint's row is its embedding id (ids outside the table become
            # row T1, which the operand kernel reports as out of range)
            node_repr, node_p = self._projected_entity_table(p, dev)
            rep_rows = int(node_repr.size(0))
            in_table = (node_embedding_ids >= 0) & (node_embedding_ids < rep_rows)
            node_slot = torch.where(in_table, node_embedding_ids, rep_rows).to(torch.int32)
            # an id outside the table is an error even on a node no edge touches (the reference's row lookup fails)
            flags.bitwise_or_((~in_table).any().to(torch.int32) * _lib.FLAG_OUT_OF_RANGE)
            head_rep32, tail_rep32 = i32(E), i32(E)
            _lib.check(L.evi_edge_rep_index(head32.data_ptr(), tail32.data_ptr(), E, node_slot.data_ptr(), N,
                                            edge_attr_dev.data_ptr(), head_rep32.data_ptr(), tail_rep32.data_ptr(),
                                            rel_slot32.data_ptr(), stream), "evi_edge_rep_index")
        elif use_tables:
            ent = self._tables["entity"]
            tsplit = bool(self._tables["split"])
            row_ids = node_embedding_ids
            node_slot = None
            if self.unique_entity_rows:
                if self._slot_map is None or self._slot_map.device != dev or self._slot_map.numel() != ent.size(0):
                    self._slot_map = torch.full((int(ent.size(0)),), -1, dtype=torch.int32, device=dev)
                row_ids = torch.empty(N, dtype=torch.long, device=dev)
                node_slot, rows_dev = i32(N), i32(1)
                _lib.check(L.evi_unique_rows(node_embedding_ids.data_ptr(), N, self._slot_map.data_ptr(), int(ent.size(0)),
                                             row_ids.data_ptr(), node_slot.data_ptr(), rows_dev.data_ptr(),
                                             flags.data_ptr(), stream), "evi_unique_rows")
                head_rep32, tail_rep32 = i32(E), i32(E)
            _lib.check(L.evi_edge_rep_index(head32.data_ptr(), tail32.data_ptr(), E, _lib.ptr(node_slot), N,
                                            edge_attr_dev.data_ptr(), _lib.ptr(head_rep32), _lib.ptr(tail_rep32),
                                            rel_slot32.data_ptr(), stream), "evi_edge_rep_index")
            x_nodes = torch.empty((N, 2 * d if tsplit else d), dtype=dt, device=dev)
            st = L.evi_gather_rows2(ent.data_ptr(), 0, ent.size(0), int(ent.stride(0)), row_ids.data_ptr(),
                                    row_ids.data_ptr(), p["non_text"].data_ptr(), N, _lib.ptr(rows_dev), d,
                                    x_nodes.data_ptr(), int(x_nodes.stride(0)), int(tsplit), fmt, flags.data_ptr(), stream)
            x_lo = d if tsplit else -1
        else:
            node_embeddings = node_embeddings.to(device=dev, dtype=torch.float32, non_blocking=True).contiguous()
            if node_embeddings.size(0) != N:
                raise ValueError(f"node_embeddings rows {node_embeddings.size(0)} != node_embedding_ids {N}")
            x_nodes = torch.empty((N, 2 * d), dtype=dt, device=dev)
            st = L.evi_gather_rows2(node_embeddings.data_ptr(), 1, N, d, None, node_embedding_ids.data_ptr(),
                                    p["non_text"].data_ptr(), N, None, d, x_nodes.data_ptr(), 2 * d, 1, fmt,
                                    flags.data_ptr(), stream)
            x_lo = d
        self._pending_flags = True
        if not use_ptable:
            _lib.check(st, "evi_gather_rows2(nodes)")
            node_repr = torch.empty((N, 2 * d), dtype=dt, device=dev)  # [hi | lo] rows, one per distinct embedding id
            self._project(x_nodes, p["we"], p["be"], tanh4, out16=node_repr, out_lo=d, p=p, rows_dev=rows_dev, x_lo=x_lo,
                          w_lo=d)
            del x_nodes

        if use_tables:
            rel_in = self._tables["relation"]
            rel_x_lo = d if self._tables["split"] else -1
        else:
            # per-edge relation rows, as the reference projects them (retriever.py:441)
            edge_embeddings = edge_embeddings.to(device=dev, dtype=torch.float32, non_blocking=True).contiguous()
            if edge_embeddings.size(0) != E:
                raise ValueError(f"edge_embeddings rows {edge_embeddings.size(0)} != edges {E}")
            rel_in = torch.empty((E, 2 * d), dtype=dt, device=dev)
            _lib.check(L.evi_gather_rows2(edge_embeddings.data_ptr(), 1, E, d, None, None, None, E, None, d,
                                          rel_in.data_ptr(), 2 * d, 1, fmt, flags.data_ptr(), stream),
                       "evi_gather_rows2(edges)")
            rel_x_lo = d
            rel_slot32 = torch.arange(E, dtype=torch.int32, device=dev)
        R = int(rel_in.size(0))
        rel_repr = torch.empty((R, 2 * d if strict else d), dtype=dt, device=dev)
        self._project(rel_in, p["wr"], p["br"], tanh4, out16=rel_repr, out_lo=d if strict else -1, p=p, x_lo=rel_x_lo,
                      w_lo=d)

        q16 = torch.empty((B, 2 * d), dtype=dt, device=dev)
        _lib.check(L.evi_gather_rows2(question_emb.data_ptr(), 1, B, d, None, None, None, B, None, d, q16.data_ptr(),
                                      2 * d, 1, fmt, flags.data_ptr(), stream), "evi_gather_rows2(questions)")
        qp16 = torch.empty((B, 2 * d), dtype=dt, device=dev)
        self._project(q16, p["wq"], p["bq"], tanh4, out16=qp16, out_lo=d, p=p, x_lo=d, w_lo=d)
        gate_bias = torch.empty((B, 2 * d), dtype=torch.float32, device=dev)
        self._project(qp16, p["wgb"], p["bgb"], [_lib.ACT_SIGMOID] * 4 + [_lib.ACT_TANH] * 4, out32=gate_bias, p=p,
                      x_lo=d, w_lo=d)

        # ---- err block of state_net.0, linear in err = h + r_ctx - t (retriever.py:477-481): once per distinct
        # embedding id and once per distinct (question, relation) pair instead of once per edge and direction
        none4 = [_lib.ACT_NONE] * 4
        if not use_ptable:
            node_p = torch.empty((N, self.hidden_dim), dtype=torch.float32, device=dev)
            self._project(node_repr, p["w1"], None, none4, out32=node_p, p=p, rows_dev=rows_dev, col0=2 * d, in_dim=d,
                          x_lo=d, w_lo=3 * d)
        pair_slot32 = torch.zeros(E, dtype=torch.int32, device=dev)
        if use_tables and graph is not None and graph.num_edges == E:
            edge_ptr_dev = graph.edge_ptr
        elif use_tables:
            if not sorted_by_kernel:
                # a caller-supplied edge_batch: the pair slots below need edges grouped by question; the reference
                # indexes query_repr[query_ids] in any order, so an unsorted list is reported, not mis-scored
                bad = (query_ids[1:] < query_ids[:-1]).any().to(torch.int32) * _lib.FLAG_NOT_MONOTONE
                flags.bitwise_or_(bad)
                self._pending_flags = True
            edge_ptr_dev = torch.searchsorted(query_ids, torch.arange(B + 1, device=dev, dtype=torch.long))
        if use_tables:
            cap = E
            pair_q32, pair_rel32, num_pairs = i32(cap), i32(cap), i32(1)
            st = L.evi_relation_pairs(rel_slot32.data_ptr(), edge_ptr_dev.data_ptr(), B, R, cap,
                                      pair_slot32.data_ptr(), pair_q32.data_ptr(), pair_rel32.data_ptr(),
                                      num_pairs.data_ptr(), flags.data_ptr(), stream)
            _lib.check(st, "evi_relation_pairs")
        else:
            cap = E  # per-edge relation rows: every edge is its own pair
            pair_slot32.copy_(rel_slot32)
            pair_q32, pair_rel32 = qid32, rel_slot32
            num_pairs = torch.full((1,), E, dtype=torch.int32, device=dev)
        rc16 = torch.empty((cap, 2 * d if strict else d), dtype=dt, device=dev)
        st = L.evi_pair_context2(pair_q32.data_ptr(), pair_rel32.data_ptr(), num_pairs.data_ptr(), cap,
                                 rel_repr.data_ptr(), int(rel_repr.stride(0)), R, gate_bias.data_ptr(), B,
                                 rc16.data_ptr(), int(rc16.stride(0)), int(strict), fmt, flags.data_ptr(), stream)
        _lib.check(st, "evi_pair_context2")

        # ---- a3: DDE structure features ------------------------------------------------------
        topic = getattr(batch, "topic_one_hot", None)
        if topic is None:
            raise ValueError("topic_one_hot is required for DDE-based structure features.")
        if not torch.is_tensor(topic):
            topic = torch.as_tensor(topic, dtype=torch.float32)
        topic = topic.to(device=dev, dtype=torch.float32)
        if topic.dim() == 1:
            topic = topic.unsqueeze(-1)
        if topic.dim() != 2:
            raise ValueError(f"topic_one_hot must be 2D (N, C), got shape {tuple(topic.shape)}")
        if topic.size(0) != N:
            raise ValueError(f"topic_one_hot first dim {topic.size(0)} != num_nodes {N}")
        if topic.size(-1) < self.num_topics:
            raise ValueError(
                f"topic_one_hot feature dim {topic.size(-1)} < num_topics={self.num_topics}; "
                "rebuild g_retrieval caches or update configs/build_retrieval_pipeline.yaml."
            )
        topic = topic.contiguous()
        rf, rr = self.dde.num_rounds, self.dde.num_reverse_rounds
        ns_dim = 2 * (1 + rf + rr)
        node_struct = torch.empty((N, ns_dim), dtype=torch.float32, device=dev)
        dde_ws = self._workspace(max(int(L.evi_dde_workspace_bytes(N)), 256 * d * 4), dev)

        def dde(h32, t32, n_edges, fwd_rounds, rev_rounds, out):
            st_ = L.evi_dde_forward(h32.data_ptr(), t32.data_ptr(), n_edges, N, topic.data_ptr(), topic.size(1),
                                    fwd_rounds, rev_rounds, out.data_ptr(), dde_ws.data_ptr(), dde_ws.numel(), stream)
            _lib.check(st_, "evi_dde_forward")

        rev = getattr(batch, "reverse_edge_index", None)
        if rev is not None and torch.is_tensor(rev):
            # retriever.py:547-550 / graph.py:47-51: the reverse rounds propagate over the given edge list (source ->
            # target) instead of the flipped one. Two launches of the same kernel (forward rounds only) and a column
            # stitch into the (x0, f1..fRf, r1..rRr) layout; the dataset never sets this attribute.
            rev = rev.to(device=dev, dtype=torch.long)
            if rev.dim() != 2 or rev.size(0) != 2:
                raise ValueError(f"reverse_edge_index must have shape [2, E], got {tuple(rev.shape)}")
            if rev.numel() and (int(rev.min()) < 0 or int(rev.max()) >= N):
                raise ValueError("reverse_edge_index out of range for node embeddings.")
            part_f = torch.empty((N, 2 * (1 + rf)), dtype=torch.float32, device=dev)
            part_r = torch.empty((N, 2 * (1 + rr)), dtype=torch.float32, device=dev)
            dde(head32, tail32, E, rf, 0, part_f)
            rh, rt = rev[0].to(torch.int32).contiguous(), rev[1].to(torch.int32).contiguous()
            dde(rh, rt, int(rh.numel()), rr, 0, part_r)
            ns3 = node_struct.view(N, 2, 1 + rf + rr)
            ns3[:, :, : 1 + rf] = part_f.view(N, 2, 1 + rf)
            ns3[:, :, 1 + rf:] = part_r.view(N, 2, 1 + rr)[:, :, 1:]
        else:
            dde_ev = self.dde_events  # optional (begin, end) torch events around the propagation (bench.py)
            if dde_ev is not None:
                dde_ev[0].record()
            dde(head32, tail32, E, rf, rr, node_struct)
            if dde_ev is not None:
                dde_ev[1].record()
        node_uv = torch.empty((N, 2 * d), dtype=torch.float32, device=dev)
        _lib.check(L.evi_struct_project(node_struct.data_ptr(), N, ns_dim, p["ws_t"].data_ptr(), p["bs"].data_ptr(),
                                        node_uv.data_ptr(), stream), "evi_struct_project")

        direction = {"bidirectional": _lib.DIR_BIDIRECTIONAL, "forward": _lib.DIR_FORWARD,
                     "backward": _lib.DIR_BACKWARD}[self.direction_mode]
        # batch centre of the structure context: the operand kernel stores fp16(sc - c), and W_sc c (fp32 weights)
        # rides in the bias of the pair rows
        sc_center, pair_bias = None, p["b1"]
        if self.center_structure_context:
            sc_center = torch.empty(d, dtype=torch.float32, device=dev)
            pair_bias = torch.empty(self.hidden_dim, dtype=torch.float32, device=dev)
            _lib.check(L.evi_sc_center(head32.data_ptr(), tail32.data_ptr(), E, node_uv.data_ptr(), N,
                                       p["ls_g"].data_ptr(), p["ls_b"].data_ptr(), direction, p["w_sc32"].data_ptr(),
                                       p["b1"].data_ptr(), dde_ws.data_ptr(), dde_ws.numel(), sc_center.data_ptr(),
                                       pair_bias.data_ptr(), stream), "evi_sc_center")
        pair_rq = torch.empty((cap, self.hidden_dim), dtype=torch.float32, device=dev)
        self._project(rc16, p["w1"], pair_bias, none4, out32=pair_rq, p=p, rows_dev=num_pairs, col0=2 * d, in_dim=d,
                      x_lo=d if strict else -1, w_lo=3 * d if strict else -1)
        del rc16

        # ---- a5/a6: edge scoring -------------------------------------------------------------
        logits = torch.empty(E, dtype=torch.float32, device=dev)
        logits_fwd = torch.empty(E, dtype=torch.float32, device=dev) if direction != _lib.DIR_BACKWARD else None
        logits_bwd = torch.empty(E, dtype=torch.float32, device=dev) if direction != _lib.DIR_FORWARD else None
        feats = torch.empty((E, self.hidden_dim), dtype=torch.float32, device=dev) if return_features else None
        # chunking bounds the workspace: `pipeline_chunks` chunks per batch (never below 16k edges per chunk),
        # capped by max_workspace_bytes
        prec = _lib.PRECISION_STRICT if strict else _lib.PRECISION_FAST
        per_edge = int(L.evi_score_workspace_bytes2(1 << 20, direction, int(return_features), prec)) / float(1 << 20)
        chunk_cap = max(1024, int(self.max_workspace_bytes / per_edge) // 1024 * 1024)
        chunk = -(-E // self.pipeline_chunks)
        chunk = min(chunk_cap, max(int(os.environ.get("EVI_MIN_CHUNK", "16384")), (chunk + 1023) // 1024 * 1024))
        if chunk >= E:
            chunk = E
        per_chunk = int(L.evi_score_workspace_bytes2(chunk, direction, int(return_features), prec))
        # two workspace halves + a second stream: the GEMMs of chunk c overlap the HBM-bound stages of chunks c-1 / c+1
        overlap = self.overlap_streams and chunk < E and not return_features
        ws = self._score_workspace(per_chunk * (2 if overlap else 1), dev)
        a = _lib.ScoreArgs()
        if overlap:
            if self._aux_stream is None or self._aux_stream.device != dev:
                self._aux_stream = torch.cuda.Stream(device=dev)
            a.aux_stream = self._aux_stream.cuda_stream
        a.emb_dim, a.hidden_dim, a.direction, a.operand_format = d, self.hidden_dim, direction, fmt
        a.want_features, a.num_sms = int(return_features), p["num_sms"]
        a.num_nodes, a.num_rel_rows, a.num_questions, a.flags = N, R, B, flags.data_ptr()
        a.num_pair_rows = cap
        a.node_repr16, a.rel_repr16 = node_repr.data_ptr(), rel_repr.data_ptr()
        a.node_repr_ld, a.rel_repr_ld = int(node_repr.stride(0)), int(rel_repr.stride(0))
        a.num_rep_rows = rep_rows
        a.head_rep32, a.tail_rep32 = _lib.ptr(head_rep32), _lib.ptr(tail_rep32)
        a.sc_center = _lib.ptr(sc_center)
        a.precision = prec
        a.gate_bias32, a.node_uv32 = gate_bias.data_ptr(), node_uv.data_ptr()
        a.node_p32, a.pair_rq32 = node_p.data_ptr(), pair_rq.data_ptr()
        a.struct_ln_g, a.struct_ln_b = p["ls_g"].data_ptr(), p["ls_b"].data_ptr()
        a.nav_w, a.nav_b = p["wn"].data_ptr(), p["bn"]
        a.w1_main16, a.w1_dist_col = p["w1"].data_ptr(), p["w1d"].data_ptr()
        a.w1_ld, a.w1_lo_off = int(p["w1"].stride(0)), 3 * d
        a.ln1_g, a.ln1_b, a.head_v, a.head_c0 = p["l1_g"].data_ptr(), p["l1_b"].data_ptr(), p["v"].data_ptr(), p["c0"]
        a.w2_16, a.w2_bias, a.head_w, a.head_b = p["w2"].data_ptr(), p["b2"].data_ptr(), p["wo"].data_ptr(), p["bo"]
        a.w2_ld = int(p["w2"].stride(0))
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        a.num_edges, a.chunk_edges = E, chunk
        a.head32, a.tail32, a.qid32, a.rel_slot32 = head32.data_ptr(), tail32.data_ptr(), qid32.data_ptr(), rel_slot32.data_ptr()
        a.pair_slot32 = pair_slot32.data_ptr()
        a.logits = logits.data_ptr()
        a.logits_fwd = logits_fwd.data_ptr() if logits_fwd is not None else None
        a.logits_bwd = logits_bwd.data_ptr() if logits_bwd is not None else None
        a.edge_embeddings = feats.data_ptr() if feats is not None else None
        self.last_chunk_edges = min(chunk, E)
        self.last_num_chunks = -(-E // max(1, min(chunk, E)))
        events = self.gemm_events  # (torch.cuda.Event begin, end) around one W_sc GEMM launch, or None
        a.ev_gemm_begin = events[0].cuda_event if events is not None else None
        a.ev_gemm_end = events[1].cuda_event if events is not None else None
        tail_events = self.tail_events  # the same for the tail kernel of that chunk
        a.ev_tail_begin = tail_events[0].cuda_event if tail_events is not None else None
        a.ev_tail_end = tail_events[1].cuda_event if tail_events is not None else None
        _lib.check(L.evi_score_edges(C.byref(a), stream), "evi_score_edges")
        self._pending_flags = True
        # device-side row counts of the last forward (distinct embedding ids, (question, relation) pairs) for callers
        # that account for executed work after a synchronisation (bench.py); None / N / E when not deduplicated
        self.last_counts = {"nodes": N, "edges": E, "relation_rows": R, "questions": B,
                            "entity_rows": 0 if use_ptable else (rows_dev if rows_dev is not None else N),
                            "projected_table_rows": rep_rows if use_ptable else 0,
                            "pair_rows": num_pairs if use_tables else E}

        # ---- a7: hide-and-seek bias (eval only when apply_in_eval; retriever.py:247-256) ----
        logits_ref = logits_fwd if logits_fwd is not None else logits_bwd
        edge_bias = self._compute_hide_seek_bias(batch, edge_index=edge_index, logits=logits_ref)
        if edge_bias is not None:
            # the same bias on both directions leaves the softmax weights unchanged
            logits = logits + edge_bias
            logits_fwd = logits_fwd + edge_bias if logits_fwd is not None else None
            logits_bwd = logits_bwd + edge_bias if logits_bwd is not None else None
        if self.direction_mode == "forward":
            logits = logits_fwd
        elif self.direction_mode == "backward":
            logits = logits_bwd

        if self.validate:
            self.check_flags()
        output = RetrieverOutput(
            logits=logits,
            query_ids=query_ids,
            relation_ids=edge_attr,
            logits_fwd=logits_fwd,
            logits_bwd=logits_bwd,
            edge_embeddings=feats,
        )
        return output, feats

    # ---- a7 helpers: same semantics as retriever.py:291-367 -----------------------------------
    def _init_hide_seek(self, hide_seek_cfg: Optional[Dict[str, Any]]) -> None:
        cfg = hide_seek_cfg or {}
        self.hide_seek_enabled = bool(cfg.get("enabled", False))
        self.hide_seek_apply_in_eval = bool(cfg.get("apply_in_eval", False))
        self.hide_seek_p_near = float(cfg.get("p_near", 0.0))
        self.hide_seek_p_far = float(cfg.get("p_far", 0.0))
        self.hide_seek_bias_near = float(cfg.get("bias_near", 0.0))
        self.hide_seek_bias_far = float(cfg.get("bias_far", 0.0))
        for name, prob in {"p_near": self.hide_seek_p_near, "p_far": self.hide_seek_p_far}.items():
            if prob < 0.0 or prob > 1.0:
                raise ValueError(f"hide_seek_cfg.{name} must be in [0, 1], got {prob}")
        for name, bias in {"bias_near": self.hide_seek_bias_near, "bias_far": self.hide_seek_bias_far}.items():
            if bias > 0.0:
                raise ValueError(f"hide_seek_cfg.{name} must be <= 0 (penalty), got {bias}")

    def _should_apply_hide_seek(self) -> bool:
        if not self.hide_seek_enabled:
            return False
        if self.training:
            return True
        return self.hide_seek_apply_in_eval

    def _compute_hide_seek_bias(self, batch: Any, *, edge_index: torch.Tensor, logits: torch.Tensor):
        if not self._should_apply_hide_seek():
            return None
        edge_is_near = getattr(batch, "edge_is_near", None)
        if edge_is_near is None:
            q_local = getattr(batch, "q_local_indices", None)
            a_local = getattr(batch, "a_local_indices", None)
            if q_local is None or a_local is None:
                raise ValueError("Batch missing q_local_indices/a_local_indices required for hide-and-seek.")
            num_nodes = getattr(batch, "num_nodes", None)
            if num_nodes is None:
                raise ValueError("Batch missing num_nodes required for hide-and-seek.")
            node_mask = torch.zeros(int(num_nodes), device=edge_index.device, dtype=torch.bool)
            qa = torch.cat([torch.as_tensor(q_local).view(-1), torch.as_tensor(a_local).view(-1)]).to(edge_index.device)
            if qa.numel() > 0:
                node_mask[qa.long()] = True  # compute_qa_edge_mask, graph_utils.py:107-153
            near_mask = node_mask[edge_index[0]] | node_mask[edge_index[1]]
        else:
            near_mask = torch.as_tensor(edge_is_near).to(device=edge_index.device, dtype=torch.bool)
            if near_mask.numel() != edge_index.size(1):
                raise ValueError(f"edge_is_near length mismatch: {near_mask.numel()} vs edges {edge_index.size(1)}")
        if near_mask.numel() == 0:
            return None
        if self.hide_seek_p_near <= 0.0 and self.hide_seek_p_far <= 0.0:
            return None
        if self.hide_seek_bias_near == 0.0 and self.hide_seek_bias_far == 0.0:
            return None
        drop_prob = torch.where(near_mask, logits.new_tensor(self.hide_seek_p_near), logits.new_tensor(self.hide_seek_p_far))
        drop_mask = torch.rand_like(drop_prob) < drop_prob
        bias_values = torch.where(near_mask, logits.new_tensor(self.hide_seek_bias_near),
                                  logits.new_tensor(self.hide_seek_bias_far))
        return torch.where(drop_mask, bias_values, logits.new_zeros(drop_prob.shape))

    def _empty_output(self, *, head_idx: torch.Tensor, return_features: bool):
        """retriever.py:383-401."""
        param = next(self.parameters(), None)
        dtype = param.dtype if param is not None else torch.float32
        device = head_idx.device
        empty_logits = torch.empty(0, device=device, dtype=dtype)
        features = torch.empty((0, self.hidden_dim), device=device, dtype=dtype)
        output = RetrieverOutput(
            logits=empty_logits, query_ids=head_idx.new_empty(0), relation_ids=None,
            logits_fwd=empty_logits, logits_bwd=empty_logits, edge_embeddings=features,
        )
        return output, (features if return_features else None)

    @staticmethod
    def _normalize_core_mode(mode: str) -> str:
        mode_clean = str(mode or "").strip().lower()
        if mode_clean in {"geometry", "structured", "full"}:
            return "geometry"
        raise ValueError(f"core_mode must be 'geometry' for DDE-based retriever, got {mode!r}")

    @staticmethod
    def _normalize_direction_mode(mode: str) -> str:
        mode_clean = str(mode or "").strip().lower()
        if mode_clean in {"bidirectional", "forward", "backward"}:
            return mode_clean
        raise ValueError(
            "direction_mode must be one of {'bidirectional', 'forward', 'backward'}, "
            f"got {mode!r}."
        )


__all__ = ["Retriever", "RetrieverOutput"]
