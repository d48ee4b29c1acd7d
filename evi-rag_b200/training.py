"""Training path of the drop-in ``Retriever`` ((f) rank 4): forward with autograd so that
``RetrieverModule.training_step`` (src/models/retriever_module.py:290-334) can use the same ``_target_`` as eval.

What runs where. Every dense contraction of the reference's training forward AND of its backward — ``x W^T``,
``dY W`` and ``dY^T X`` for entity / relation / query projections, q_gate, q_bias, ``state_net.0`` (K = 3072) and
``state_net.4`` — goes through ``tc_matmul``: the tcgen05 GEMM of ``csrc/gemm2_tcgen05.cuh`` over ``[hi | lo]`` fp16
operand pairs (three tensor passes, fp32 accumulation; operands are scaled by a power of two per tensor so small
gradients do not underflow fp16), i.e. fp32-grade products on the tensor cores. The DDE structure features (no
parameters) come from ``evi_dde_forward``. The element-wise pieces (tanh, gates, LayerNorm, GELU, dropout, the
20-wide ``struct_proj.0``, the two width-1 heads, gathers and their scatter-add gradients, the softmax combine) are
torch CUDA ops under autograd: this path follows the reference's per-edge formulation (retriever.py:195-289,
453-495) literally and is not fused — training throughput is not the benchmarked hot path; what it buys is
gradients that match the reference's (tests/test_gpu_parity.py::test_training_gradients_match_reference) from the
same module that serves eval. Dropout and the hide-and-seek draw consume the torch CUDA generator in the reference's
order (forward direction, backward direction, hide-and-seek)."""
from __future__ import annotations

import ctypes as C
from typing import Any, Optional

import torch
import torch.nn.functional as F

from . import lib as _lib

_ACT_NONE = (C.c_int * 8)(*([_lib.ACT_NONE] * 8))


def _split16(x: torch.Tensor, k_pad: int) -> torch.Tensor:
    """[rows, k] fp32 -> [rows, 2 k_pad] fp16 rows [hi | lo] (zero padded), the operand layout of gemm::SplitK."""
    rows, k = x.shape
    out = torch.zeros((rows, 2 * k_pad), dtype=torch.float16, device=x.device)
    hi = x.to(torch.float16)
    out[:, :k] = hi
    out[:, k_pad:k_pad + k] = (x - hi.float()).to(torch.float16)
    return out


def _pow2_scale(x: torch.Tensor) -> torch.Tensor:
    """A power of two s (device scalar, no host sync) with max|x s| in [2^10, 2^11): the [hi | lo] halves of the
    scaled tensor stay inside fp16's normal range whatever the magnitude of x (gradients can be ~1e-8)."""
    m = x.detach().abs().amax().clamp_min(1e-30)
    return torch.exp2(10.0 - torch.floor(torch.log2(m)))


_K_CHUNK = 4096


@torch.no_grad()
def tc_matmul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``a [M, K] @ b [N, K]^T`` in fp32-grade precision on the tcgen05 GEMM (three passes over [hi | lo] pairs).
    The tensor core's fp32 accumulator loses a little on every addition (measured 2.8e-5 of max|result| at K = 5000),
    so a long contraction — dW = dY^T X has K = number of edges — is cut into K blocks of 4096 whose partial results
    are added in fp32 on the CUDA cores."""
    if a.device.type != "cuda" or b.device != a.device:
        raise RuntimeError("tc_matmul runs on CUDA tensors of one device (no CPU fallback).")
    M, K = a.shape
    N = b.size(0)
    if b.size(1) != K:
        raise ValueError(f"inner dimensions differ: {tuple(a.shape)} x {tuple(b.shape)}^T")
    if M == 0 or N == 0:
        return torch.zeros((M, N), dtype=torch.float32, device=a.device)
    if K > _K_CHUNK:
        out = None
        for k0 in range(0, K, _K_CHUNK):
            part = tc_matmul(a[:, k0:k0 + _K_CHUNK].contiguous(), b[:, k0:k0 + _K_CHUNK].contiguous())
            out = part if out is None else out.add_(part)
        return out
    L = _lib.load()
    k_pad = (K + 63) // 64 * 64
    a = a.to(torch.float32)
    b = b.to(torch.float32)
    sa, sb = _pow2_scale(a), _pow2_scale(b)
    a16 = _split16(a * sa, k_pad)
    out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    num_sms = torch.cuda.get_device_properties(a.device).multi_processor_count
    with torch.cuda.device(a.device):
        stream = _lib.stream_handle(a.device)
        for n0 in range(0, N, 2048):  # evi_project_split takes up to 8 column blocks of 256
            n1 = min(N, n0 + 2048)
            n_pad = (n1 - n0 + 255) // 256 * 256
            b16 = torch.zeros((n_pad, 2 * k_pad), dtype=torch.float16, device=a.device)
            b16[: n1 - n0] = _split16(b[n0:n1] * sb, k_pad)
            blk = out[:, n0:n1] if n_pad == n1 - n0 and n0 == 0 and n1 == N else \
                torch.empty((M, n_pad), dtype=torch.float32, device=a.device)
            st = L.evi_project_split(a16.data_ptr(), 2 * k_pad, k_pad, M, None, b16.data_ptr(), 2 * k_pad, k_pad, n_pad,
                                     k_pad, None, _ACT_NONE, None, 0, -1, blk.data_ptr(), _lib.FMT_F16, num_sms, stream)
            _lib.check(st, "evi_project_split")
            if blk.data_ptr() != out.data_ptr():
                out[:, n0:n1] = blk[:, : n1 - n0]
    return out / (sa * sb)


class _TcLinear(torch.autograd.Function):
    """``y = x W^T + b`` with all three products (y, dX = dY W, dW = dY^T X) on the tensor cores."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        y = tc_matmul(x, weight)
        return y + bias if bias is not None else y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dy = dy.contiguous()
        dx = tc_matmul(dy, weight.t().contiguous()) if ctx.needs_input_grad[0] else None
        dw = tc_matmul(dy.t().contiguous(), x.t().contiguous()) if ctx.needs_input_grad[1] else None
        db = dy.sum(dim=0) if ctx.has_bias and ctx.needs_input_grad[2] else None
        return dx, dw, db


def tc_linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    return _TcLinear.apply(x, weight, bias)


def _lin(x: torch.Tensor, layer: torch.nn.Linear) -> torch.Tensor:
    return tc_linear(x, layer.weight, layer.bias)


def _score_direction(model: Any, head_repr, r_ctx, tail_repr, s20, want_features: bool):
    """``_score_edges`` + ``_encode_structure`` (retriever.py:453-495) for one direction, under autograd."""
    d = model.emb_dim
    sp0, sp_ln = model.struct_proj[0], model.struct_proj[1]
    sc = F.gelu(F.layer_norm(F.linear(s20, sp0.weight, sp0.bias), (d,), sp_ln.weight, sp_ln.bias, sp_ln.eps))
    nav = torch.sigmoid(F.linear(sc, model.struct_gate_net[0].weight, model.struct_gate_net[0].bias))
    iv = head_repr * r_ctx * tail_repr * nav
    err = head_repr + r_ctx - tail_repr
    dist = -torch.norm(err, p=2, dim=-1, keepdim=True)
    sn0, sn_ln, drop, sn4 = model.state_net[0], model.state_net[1], model.state_net[3], model.state_net[4]
    # state_net.0 on [iv | sc | err | dist]: the 3D-wide block on the tensor cores, the dist column as a rank-1 term
    z = tc_linear(torch.cat([iv, sc, err], dim=-1), sn0.weight[:, : 3 * d], sn0.bias) + dist * sn0.weight[:, 3 * d]
    g = F.dropout(F.gelu(F.layer_norm(z, (z.size(-1),), sn_ln.weight, sn_ln.bias, sn_ln.eps)), p=drop.p, training=True)
    feat = _lin(g, sn4)
    logit = F.linear(feat, model.score_head.weight, model.score_head.bias).squeeze(-1)
    return logit, (feat if want_features else None)


def forward_train(model: Any, batch: Any, return_features: bool):
    """``Retriever._forward_impl`` in training mode (retriever.py:195-289). Returns ``(RetrieverOutput, features)``."""
    from . import graph as G
    from .retriever import RetrieverOutput

    edge_index = getattr(batch, "edge_index", None)
    if edge_index is None:
        raise ValueError("Batch missing edge_index required for scoring.")
    dev = model._device()
    if dev.type != "cuda":
        raise RuntimeError("evi_rag_b200.Retriever runs on CUDA only (no CPU fallback); call .cuda() first.")
    edge_index = edge_index.to(device=dev)
    if edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError(f"edge_index must have shape [2, E], got {tuple(edge_index.shape)}")
    if edge_index.size(1) == 0:
        return model._empty_output(head_idx=edge_index[0], return_features=return_features)
    edge_index = edge_index.to(torch.long).contiguous()
    question_emb = getattr(batch, "question_emb", None)
    node_embedding_ids = getattr(batch, "node_embedding_ids", None)
    edge_attr = getattr(batch, "edge_attr", None)
    if question_emb is None or node_embedding_ids is None or edge_attr is None:
        raise ValueError("Batch must provide question_emb, node_embedding_ids, and edge_attr.")
    if getattr(batch, "reverse_edge_index", None) is not None:
        raise ValueError("reverse_edge_index is supported by the inference path only.")
    node_ptr = getattr(batch, "ptr", None)
    if node_ptr is None:
        raise ValueError("Batch missing ptr required for edge_batch computation.")
    with torch.cuda.device(dev):
        with torch.no_grad():  # segmentation (validated like compute_edge_batch) and the parameter-free DDE features
            gb = G.GraphBatch.from_edge_index(edge_index, node_ptr, device=dev, validate=True)
            topic = getattr(batch, "topic_one_hot", None)
            if topic is None:
                raise ValueError("topic_one_hot is required for DDE-based structure features.")
            topic = torch.as_tensor(topic, dtype=torch.float32).to(dev)
            if topic.dim() != 2 or topic.size(0) != gb.num_nodes or topic.size(-1) < model.num_topics:
                raise ValueError(f"topic_one_hot must be [num_nodes, >= {model.num_topics}], got {tuple(topic.shape)}")
            ns = G.dde_node_features(gb, topic.contiguous(), model.dde.num_rounds, model.dde.num_reverse_rounds)
        query_ids = gb.edge_batch
        if getattr(batch, "edge_batch", None) is None:
            batch.edge_batch = query_ids  # retriever.py:612
        head, tail = edge_index[0], edge_index[1]
        ids = torch.as_tensor(node_embedding_ids, dtype=torch.long).to(dev)
        rel_ids = torch.as_tensor(edge_attr, dtype=torch.long).to(dev)
        d = model.emb_dim

        node_rows = getattr(batch, "node_embeddings", None)
        edge_rows = getattr(batch, "edge_embeddings", None)
        tables = model._tables
        if node_rows is None or edge_rows is None:
            if tables is None:
                raise ValueError("Batch must provide node_embeddings and edge_embeddings.")
            full = lambda t, i: (t[i, :d].float() + t[i, d:].float()) if tables["split"] else t[i].float()
            uniq, inv = torch.unique(ids, return_inverse=True)  # one projected row per distinct embedding id
            x_nodes = full(tables["entity"], uniq)
            uniq_r, inv_r = torch.unique(rel_ids, return_inverse=True)
            x_rel = full(tables["relation"], uniq_r)
        else:
            uniq, inv = ids, None
            x_nodes = node_rows.to(device=dev, dtype=torch.float32)
            x_rel, inv_r = edge_rows.to(device=dev, dtype=torch.float32), None
        # rows of non-text entities are the learned embedding (retriever.py:499-506)
        x_nodes = torch.where((uniq == 0).unsqueeze(-1), model.non_text_entity_emb.weight[0].unsqueeze(0), x_nodes)
        node_repr = torch.tanh(_lin(x_nodes, model.entity_proj.network[0]))
        rel_repr = torch.tanh(_lin(x_rel, model.relation_proj.network[0]))
        q_proj = torch.tanh(_lin(question_emb.to(device=dev, dtype=torch.float32), model.query_proj.network[0]))
        gate_q = torch.sigmoid(_lin(q_proj, model.q_gate[0]))
        bias_q = torch.tanh(_lin(q_proj, model.q_bias[0]))
        if inv is not None:
            node_repr = node_repr[inv]
        h, t = node_repr[head], node_repr[tail]
        r = rel_repr[inv_r] if inv_r is not None else rel_repr
        r_ctx = r * gate_q[query_ids] + bias_q[query_ids]  # retriever.py:464
        s_f = torch.cat([ns[head], ns[tail]], dim=-1)  # retriever.py:509-517
        s_b = torch.cat([ns[tail], ns[head]], dim=-1)
        mode = model.direction_mode
        lf = lb = ff = fb = None
        if mode in ("forward", "bidirectional"):
            lf, ff = _score_direction(model, h, r_ctx, t, s_f, return_features)
        if mode in ("backward", "bidirectional"):
            lb, fb = _score_direction(model, t, r_ctx, h, s_b, return_features)
        edge_bias = model._compute_hide_seek_bias(batch, edge_index=edge_index, logits=lf if lf is not None else lb)
        if edge_bias is not None:
            lf = lf + edge_bias if lf is not None else None
            lb = lb + edge_bias if lb is not None else None
        if mode == "bidirectional":  # retriever.py:369-381
            st = torch.stack([lf, lb], dim=0)
            w = torch.softmax(st, dim=0)
            logits = (w * st).sum(dim=0)
            feats = w[0].unsqueeze(-1) * ff + w[1].unsqueeze(-1) * fb if return_features else None
        elif mode == "forward":
            logits, feats = lf, ff
        else:
            logits, feats = lb, fb
    out = RetrieverOutput(logits=logits, query_ids=query_ids, relation_ids=edge_attr, logits_fwd=lf, logits_bwd=lb,
                          edge_embeddings=feats)
    return out, feats


__all__ = ["tc_matmul", "tc_linear", "forward_train"]
