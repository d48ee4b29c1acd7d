"""RetrieverLoss on the device.

Drop-in for ``src/losses/retriever_loss.py``: same constructor keywords and validation (:32-66), same
``forward(output, targets, training_step, *, edge_batch, num_graphs, edge_is_near, path_edge_indices)`` contract,
same ``LossOutput(loss, components, metrics)``. The ~20 scatter launches and dozen ``.item()`` syncs of the
reference become three launches (``evi_retriever_loss``) and one read-back of 12 scalars. ``loss`` is differentiable
with respect to ``output.logits`` (custom autograd function whose backward is the gradient the kernel already
wrote), so a training step that computes logits with an autograd-capable module keeps working; the fused scorer
itself is inference-only.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import lib as _lib

NUM_STATS, NUM_SUMMARY = 12, 12


@dataclass
class LossOutput:
    loss: torch.Tensor
    components: Dict[str, float]
    metrics: Dict[str, float]


@_lib.device_guard
def _launch(logits, targets, edge_ptr, qid32, is_near, cfg, want_grad: bool):
    L = _lib.load()
    dev = logits.device
    E, B = int(logits.numel()), int(edge_ptr.numel() - 1)
    stats = torch.empty((B, NUM_STATS), dtype=torch.float32, device=dev)
    summary = torch.empty(NUM_SUMMARY, dtype=torch.float32, device=dev)
    grad = torch.empty(E, dtype=torch.float32, device=dev) if want_grad else None
    st = L.evi_retriever_loss(logits.data_ptr(), targets.data_ptr(), edge_ptr.data_ptr(), _lib.ptr(qid32), E, B,
                              _lib.ptr(is_near), cfg["w_near"], cfg["w_bridge"], cfg["temperature"], cfg["infonce_weight"],
                              cfg["bce_weight"], stats.data_ptr(), summary.data_ptr(), _lib.ptr(grad), _lib.stream_handle())
    _lib.check(st, "evi_retriever_loss")
    return stats, summary, grad


class _LossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, edge_ptr, qid32, is_near, cfg):
        stats, summary, grad = _launch(logits.detach().contiguous(), targets, edge_ptr, qid32, is_near, cfg,
                                       want_grad=logits.requires_grad)
        ctx.grad_logits = grad
        ctx.mark_non_differentiable(summary, stats)
        return summary[11].clone(), summary, stats

    @staticmethod
    def backward(ctx, g_loss, _g_summary, _g_stats):
        return (ctx.grad_logits * g_loss if ctx.grad_logits is not None else None), None, None, None, None, None


class RetrieverLoss(nn.Module):
    """Composite loss: multi-positive InfoNCE (+ optional BCE), triple-only supervision (retriever_loss.py:29-66)."""

    def __init__(self, *, path_weight: float = 0.0, path_warmup_steps: int = 0, infonce_temperature: float = 1.0,
                 infonce_weight: float = 1.0, bce_weight: float = 0.0, edge_weight_near: float = 1.0,
                 edge_weight_bridge: float = 1.0) -> None:
        super().__init__()
        self.path_weight = float(path_weight)
        if self.path_weight != 0.0:
            raise ValueError("RetrieverLoss forbids path supervision; set path_weight=0 and keep path_edge_indices unset.")
        self.path_warmup_steps = int(path_warmup_steps)
        if self.path_warmup_steps != 0:
            raise ValueError("RetrieverLoss forbids path warmup; set path_warmup_steps=0.")
        self.infonce_temperature = float(infonce_temperature)
        if self.infonce_temperature <= 0.0:
            raise ValueError(f"infonce_temperature must be positive, got {self.infonce_temperature}")
        self.infonce_weight = float(infonce_weight)
        self.bce_weight = float(bce_weight)
        if self.infonce_weight < 0.0 or self.bce_weight < 0.0:
            raise ValueError("infonce_weight and bce_weight must be non-negative.")
        if self.infonce_weight == 0.0 and self.bce_weight == 0.0:
            raise ValueError("RetrieverLoss requires at least one non-zero loss weight.")
        self.edge_weight_near = float(edge_weight_near)
        self.edge_weight_bridge = float(edge_weight_bridge)
        if self.edge_weight_near <= 0.0 or self.edge_weight_bridge <= 0.0:
            raise ValueError("edge_weight_near and edge_weight_bridge must be positive.")

    @property
    def requires_edge_is_near(self) -> bool:
        return self.edge_weight_near != 1.0 or self.edge_weight_bridge != 1.0

    def forward(self, output, targets: torch.Tensor, training_step: int = 0, *, edge_batch: Optional[torch.Tensor] = None,
                num_graphs: Optional[int] = None, edge_is_near: Optional[torch.Tensor] = None,
                path_edge_indices: Optional[torch.Tensor] = None) -> LossOutput:
        logits = getattr(output, "logits", None)
        if logits is None:
            raise ValueError("RetrieverLoss requires output.logits.")
        if edge_batch is None:
            raise ValueError("RetrieverLoss requires edge_batch to define per-graph groups.")
        if path_edge_indices is not None:
            raise ValueError("RetrieverLoss forbids path_edge_indices; retriever is triple-only.")
        logits = logits.view(-1)
        if logits.device.type != "cuda":
            raise RuntimeError("RetrieverLoss runs on CUDA tensors only (no CPU fallback).")
        dev = logits.device
        targets = targets.view(-1).to(device=dev, dtype=torch.float32).contiguous()
        edge_batch = edge_batch.view(-1).to(device=dev, dtype=torch.long)
        if logits.numel() == 0:
            raise ValueError("RetrieverLoss received empty logits/targets; check dataset filtering.")
        if logits.numel() != targets.numel() or logits.numel() != edge_batch.numel():
            raise ValueError(
                f"logits/targets/edge_batch shape mismatch: {logits.shape} vs {targets.shape} vs {edge_batch.shape}")
        if num_graphs is None:
            num_graphs = int(edge_batch.max().item()) + 1
        num_graphs = int(num_graphs)
        if num_graphs <= 0:
            raise ValueError(f"num_graphs must be positive, got {num_graphs}")
        if bool((edge_batch[1:] < edge_batch[:-1]).any()):
            raise ValueError("RetrieverLoss expects edges grouped by question (PyG collate order).")
        is_near = None
        if self.requires_edge_is_near:
            if edge_is_near is None:
                raise ValueError("RetrieverLoss requires edge_is_near when edge weights are enabled.")
            is_near = edge_is_near.to(device=dev, dtype=torch.bool).view(-1)
            if is_near.numel() != logits.numel():
                raise ValueError(f"edge_is_near length mismatch: {is_near.numel()} vs logits {logits.numel()}")
            is_near = is_near.to(torch.uint8).contiguous()
        edge_ptr = torch.zeros(num_graphs + 1, dtype=torch.long, device=dev)
        edge_ptr[1:] = torch.cumsum(torch.bincount(edge_batch, minlength=num_graphs)[:num_graphs], 0)
        qid32 = edge_batch.to(torch.int32).contiguous()
        cfg = {"w_near": self.edge_weight_near, "w_bridge": self.edge_weight_bridge, "temperature": self.infonce_temperature,
               "infonce_weight": self.infonce_weight, "bce_weight": self.bce_weight}
        loss, summary, _ = _LossFn.apply(logits.to(torch.float32), targets, edge_ptr, qid32, is_near, cfg)
        s = summary.tolist()  # the one read-back
        infonce_metrics = {"infonce_pos_edges": s[4], "infonce_neg_edges": s[5], "infonce_graphs": s[2]}
        if s[4] > 0 and s[5] > 0:
            infonce_metrics.update(infonce_graphs_no_pos=s[6], infonce_graphs_no_neg=s[7])
        bce_metrics = {"bce_graphs": s[3], "bce_edges": s[10]} if self.bce_weight > 0.0 else {"bce_graphs": 0.0, "bce_edges": 0.0}
        return LossOutput(
            loss=loss,
            components={"infonce": s[0], "infonce_weight": self.infonce_weight, "bce": s[1], "bce_weight": self.bce_weight,
                        "path": 0.0, "path_weight": 0.0},
            metrics={"pos_prob": s[8], "neg_prob": s[9], "separation": s[8] - s[9], **infonce_metrics, **bce_metrics,
                     "path_graphs": 0.0})


__all__ = ["RetrieverLoss", "LossOutput"]
