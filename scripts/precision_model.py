#!/usr/bin/env python
"""CPU emulation of the rounding points of the CUDA scoring path (16-bit tensor-core operands, fp32
accumulation) against the fp32 oracle, to price storage-format choices before spending GPU time.

    python scripts/precision_model.py [--questions 4] [--edges 1500]

Every variant rounds what the kernels round (embedding rows, weights, projected rows, operand rows m / sc,
r_ctx for the pair GEMM) and optionally more (U/V, P/Rq, the partial products). Reported: norm-wise
relative error of the logits and max abs / max|ref| — the two numbers the GPU parity tests bound.
Test infrastructure only (imports the oracle)."""
import argparse
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import evi_oracle as orc  # noqa: E402
from evi_rag_b200.synthetic import generate_questions, tiny_spec  # noqa: E402
from evi_rag_b200.weights import make_state_dict  # noqa: E402


def rnd(x, dt):
    return x.to(dt).float() if dt is not None else x


def quant_rows_i16(x, block=256):
    """W_sc sc rows as stored by the EpiQ16Tma epilogue: per (row, 256-column block) scale = max|x| / 32767 and
    round-to-nearest 16-bit integers; the tail folds the dequantisation offset into one fma, which adds one fp32
    rounding at 2^8 x the block maximum (modelled by the second term)."""
    r, c = x.shape
    xb = x.view(r, c // block, block)
    mx = xb.abs().amax(-1, keepdim=True)
    scale = torch.where(mx > 0, mx / 32767.0, torch.ones_like(mx))
    q = torch.round(xb / scale)
    fold = (torch.rand_like(xb) - 0.5) * (256.0 * mx * 2.0 ** -23)
    return (q * scale + fold).view(r, c)


def gpu_model(sd, qb, ent, rel, op=torch.float16, uv=None, prq=None, part=None, sites=None, mm=None, sq=False):
    """op: operand dtype; uv / prq / part: extra rounding of U,V / P,Rq / W_iv m, W_sc sc (None = fp32)."""
    t = torch.from_numpy
    W = lambda n: sd[n].float()
    _op = op
    class _S:  # per-site operand dtype: sites = set of names kept in fp32 (ablation of the error sources)
        def __getattr__(self, k):
            return None if sites and k in sites else _op
    S_ = _S()
    w16 = lambda n: rnd(W(n), S_.w)
    lin16 = lambda x, n: x @ w16(n + ".weight").t() + W(n + ".bias")
    ids = t(qb.node_embedding_ids)
    x_nodes = rnd(t(ent[qb.node_embedding_ids]), S_.x)
    x_nodes[ids == 0] = rnd(W("non_text_entity_emb.weight")[0], S_.x)
    node_repr = rnd(torch.tanh(lin16(x_nodes, "entity_proj.network.0")), S_.node)
    rel_repr = rnd(torch.tanh(lin16(rnd(t(rel), S_.x), "relation_proj.network.0")), S_.rel)
    qp = rnd(torch.tanh(lin16(rnd(t(qb.question_emb), S_.x), "query_proj.network.0")), S_.q)
    gate = torch.sigmoid(lin16(qp, "q_gate.0"))
    bias = torch.tanh(lin16(qp, "q_bias.0"))
    ns = orc.dde_node_features(t(qb.topic_one_hot), t(qb.edge_index).long(), 2, 2)
    ws = W("struct_proj.0.weight")
    k = ns.size(1)
    U = rnd(ns @ ws[:, :k].t() + W("struct_proj.0.bias"), uv)
    V = rnd(ns @ ws[:, k:].t(), uv)
    w1 = W("state_net.0.weight")
    d = 1024
    w_iv, w_sc, w_err, w_d = rnd(w1[:, :d], S_.w1), rnd(w1[:, d:2 * d], S_.w1), rnd(w1[:, 2 * d:3 * d], S_.w1), w1[:, 3 * d]
    P = rnd(node_repr @ w_err.t(), prq)
    h_i, t_i = t(qb.edge_index[0]).long(), t(qb.edge_index[1]).long()
    q_i = torch.repeat_interleave(torch.arange(qb.num_graphs), t(np.diff(qb.edge_ptr)))
    r_i = t(qb.edge_attr).long()
    h, tt = node_repr[h_i], node_repr[t_i]
    rc = rel_repr[r_i] * gate[q_i] + bias[q_i]
    Rq = rnd(rnd(rc, S_.rc) @ w_err.t() + W("state_net.0.bias"), prq)
    m = rnd(h * rc * tt, S_.m)
    Mm = rnd(rnd(m @ w_iv.t(), part), mm)  # mm: W_iv m alone stored in 16 bits
    ln = lambda x, n: F.layer_norm(x, (d,), W(n + ".weight"), W(n + ".bias"), 1e-5)
    v = (W("state_net.4.weight").double().t() @ W("score_head.weight").double()[0]).float()
    c0 = float(W("score_head.weight").double()[0] @ W("state_net.4.bias").double() + W("score_head.bias").double()[0])
    out = []
    for a, b, sign in ((h_i, t_i, 1.0), (t_i, h_i, -1.0)):
        sc = F.gelu(ln(U[a] + V[b], "struct_proj.1"))
        nav = torch.sigmoid(sc @ W("struct_gate_net.0.weight")[0] + W("struct_gate_net.0.bias")[0])
        S = rnd(rnd(sc, S_.sc) @ w_sc.t(), part)
        if sq:
            S = quant_rows_i16(S)
        dist = -torch.norm(rc + sign * (h - tt), dim=-1)
        z = nav[:, None] * Mm + S + sign * (P[h_i] - P[t_i]) + Rq + dist[:, None] * w_d[None]
        g = F.gelu(ln(z, "state_net.1"))
        out.append(g @ v + c0)
    st = torch.stack(out)
    return (torch.softmax(st, 0) * st).sum(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--questions", type=int, default=4)
    ap.add_argument("--edges", type=int, default=1500)
    args = ap.parse_args()
    torch.manual_seed(0)
    spec = tiny_spec(args.questions, args.edges, seed=11, num_relations=200, num_entities=3000)
    qb = generate_questions(spec)
    rng = np.random.default_rng(0)
    ent = (rng.standard_normal((spec.num_entities + 1, 1024)) * 0.5).astype(np.float32)
    ent[0] = 0
    rel = (rng.standard_normal((spec.num_relations, 1024)) * 0.5).astype(np.float32)
    for trained in (True, False):
        sd = make_state_dict(3, trained_like=trained)
        t = torch.from_numpy
        ref = orc.retriever_forward(sd, edge_index=t(qb.edge_index), ptr=t(qb.ptr), question_emb=t(qb.question_emb),
                                    node_embeddings=t(ent[qb.node_embedding_ids]), node_embedding_ids=t(qb.node_embedding_ids),
                                    edge_embeddings=t(rel[qb.edge_attr]), topic_one_hot=t(qb.topic_one_hot)).logits.double()
        print(f"trained_like={trained} E={qb.num_edges} max|ref|={ref.abs().max():.3f}")
        variants = [
            ("fp16 operands (shipped)", dict(op=torch.float16)),
            ("fp16 + U/V fp16", dict(op=torch.float16, uv=torch.float16)),
            ("fp16 + P/Rq fp16", dict(op=torch.float16, prq=torch.float16)),
            ("fp16 + partial products fp16", dict(op=torch.float16, part=torch.float16)),
            ("fp16 + U/V + P/Rq fp16", dict(op=torch.float16, uv=torch.float16, prq=torch.float16)),
            ("fp16 + W_iv m stored fp16", dict(op=torch.float16, mm=torch.float16)),
            ("fp16 + W_iv m fp16 + W_sc sc int16/block (shipped)", dict(op=torch.float16, mm=torch.float16, sq=True)),
            ("fp16 + W_iv m stored bf16", dict(op=torch.float16, mm=torch.bfloat16)),
            ("bf16 operands", dict(op=torch.bfloat16)),
            ("fp16, fp32 kept: proj weights", dict(op=torch.float16, sites={"w"})),
            ("fp16, fp32 kept: state_net.0 weights", dict(op=torch.float16, sites={"w1"})),
            ("fp16, fp32 kept: input rows x", dict(op=torch.float16, sites={"x"})),
            ("fp16, fp32 kept: node_repr", dict(op=torch.float16, sites={"node"})),
            ("fp16, fp32 kept: rel_repr, q", dict(op=torch.float16, sites={"rel", "q"})),
            ("fp16, fp32 kept: m", dict(op=torch.float16, sites={"m"})),
            ("fp16, fp32 kept: sc", dict(op=torch.float16, sites={"sc"})),
            ("fp16, fp32 kept: rc", dict(op=torch.float16, sites={"rc"})),
            ("fp32 everywhere (restructuring only)", dict(op=None)),
        ]
        for name, kw in variants:
            got = gpu_model(sd, qb, ent, rel, **kw).double()
            rel_l2 = float(torch.linalg.norm(got - ref) / torch.linalg.norm(ref))
            mx = float((got - ref).abs().max() / ref.abs().max())
            print(f"  {name:40s} rel_l2={rel_l2:.3e}  max_abs/max_ref={mx:.3e}")


if __name__ == "__main__":
    main()
