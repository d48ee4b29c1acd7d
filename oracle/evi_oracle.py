"""CPU ORACLE — test infrastructure only, never the product path.

A plain numpy / torch-fp32 restatement of the reference's retrieval hot path
(Martin1007Wang/EVI-RAG). Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module, and only as the checker or
as the timed CPU baseline. Nothing under ``evi-rag_b200/`` imports it.

Pinning: the reference ships no tests, golden vectors or fixtures. This
restatement is pinned against the reference's own code imported in the build container
(``oracle/make_golden.py`` -> ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` re-checks
the oracle against those vectors on every run). The one third-party piece, torch_geometric's
``MessagePassing(aggr="mean")`` (not vendored, not version-pinned by the reference:
absent from requirements.txt), is restated from its published semantics in
``dde_node_features`` — parity for that piece is "unpinned" (see DESIGN.md).

Each function cites the reference file:line it follows.
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

LN_EPS = 1e-5  # nn.LayerNorm default, retriever.py:162,177


# ----------------------------------------------------------------------------------------------
# a1: edge -> question segmentation                        src/utils/graph_utils.py:49-104
# ----------------------------------------------------------------------------------------------
def compute_edge_batch(edge_index: np.ndarray, ptr: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """``bucketize(edge_index[0], ptr[1:], right=True)`` + validation + ``edge_ptr`` cumsum."""
    edge_index = np.asarray(edge_index, dtype=np.int64)
    ptr = np.asarray(ptr, dtype=np.int64)
    num_graphs = ptr.shape[0] - 1
    if edge_index.ndim != 2 or edge_index.shape[0] != 2:
        raise ValueError(f"edge_index must have shape [2, E], got {tuple(edge_index.shape)}")
    eb = np.searchsorted(ptr[1:], edge_index[0], side="right").astype(np.int64)
    tb = np.searchsorted(ptr[1:], edge_index[1], side="right").astype(np.int64)
    if eb.size > 0:
        if eb.min() < 0 or eb.max() >= num_graphs:
            raise ValueError("edge_batch contains out-of-range indices")
    if not np.array_equal(eb, tb):
        raise ValueError("edge_index crosses graph boundaries; head/tail graph assignments differ.")
    if eb.size > 1 and not bool(np.all(eb[:-1] <= eb[1:])):
        raise ValueError("edge_batch is not non-decreasing along the flattened edge list")
    counts = np.bincount(eb, minlength=num_graphs).astype(np.int64)
    edge_ptr = np.zeros(num_graphs + 1, dtype=np.int64)
    np.cumsum(counts, out=edge_ptr[1:])
    return eb, edge_ptr


# ----------------------------------------------------------------------------------------------
# a3: DDE = rounds of directional mean propagation        src/models/components/graph.py:13-74
# ----------------------------------------------------------------------------------------------
def _mean_propagate(x: torch.Tensor, src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    """PyG ``MessagePassing(aggr='mean', flow='source_to_target')``: messages go
    ``edge_index[0] -> edge_index[1]``; mean over incoming edges with multiplicity; nodes with no
    incoming edge get 0 (graph.py:16-23 + torch_geometric semantics)."""
    out = torch.zeros_like(x)
    out.index_add_(0, dst, x[src])
    deg = torch.zeros(x.size(0), dtype=x.dtype)
    deg.index_add_(0, dst, torch.ones(dst.numel(), dtype=x.dtype))
    return out / deg.clamp(min=1).unsqueeze(-1)


def dde_node_features(
    topic_one_hot: torch.Tensor,
    edge_index: torch.Tensor,
    num_rounds: int = 2,
    num_reverse_rounds: int = 2,
    num_topics: int = 2,
    reverse_edge_index: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Node structure features ``[N, C*(1+R_f+R_r)]`` laid out per topic class
    ``[c0:(x0,f1..,r1..), c1:(...)]`` (retriever.py:519-553, graph.py:41-74). ``reverse_edge_index`` replaces the
    flipped edge list of the reverse rounds (retriever.py:547-550)."""
    x0 = topic_one_hot.to(torch.float32)[:, :num_topics]
    feats = [x0]
    h = x0
    for _ in range(num_rounds):
        h = _mean_propagate(h, edge_index[0], edge_index[1])
        feats.append(h)
    h = x0
    rsrc, rdst = (edge_index[1], edge_index[0]) if reverse_edge_index is None else \
        (reverse_edge_index[0].long(), reverse_edge_index[1].long())  # edge_index.flip(0), graph.py:47-48
    for _ in range(num_reverse_rounds):
        h = _mean_propagate(h, rsrc, rdst)
        feats.append(h)
    stacked = torch.stack(feats, dim=-1)
    return stacked.reshape(stacked.size(0), -1)


# ----------------------------------------------------------------------------------------------
# a4-a6: the scorer                                        src/models/components/retriever.py
# ----------------------------------------------------------------------------------------------
@dataclass
class OracleOutput:
    logits: torch.Tensor
    query_ids: torch.Tensor
    logits_fwd: Optional[torch.Tensor]
    logits_bwd: Optional[torch.Tensor]
    edge_embeddings: Optional[torch.Tensor]


def _lin(x, sd, name):
    return F.linear(x, sd[f"{name}.weight"], sd[f"{name}.bias"])


def _score_direction(sd, q, h, r, t, s20, want_features: bool, dropout_p: float = 0.0):
    """``_score_edges`` + ``_encode_structure`` (retriever.py:453-495). ``dropout_p`` > 0: the training-mode
    ``nn.Dropout`` of ``state_net`` (retriever.py:175-181), drawn from the torch generator of the tensors' device."""
    r_ctx = r * torch.sigmoid(_lin(q, sd, "q_gate.0")) + torch.tanh(_lin(q, sd, "q_bias.0"))
    sc = F.gelu(
        F.layer_norm(_lin(s20, sd, "struct_proj.0"), (sd["struct_proj.1.weight"].numel(),),
                     sd["struct_proj.1.weight"], sd["struct_proj.1.bias"], LN_EPS)
    )
    nav = torch.sigmoid(_lin(sc, sd, "struct_gate_net.0"))
    iv = h * r_ctx * t * nav
    err = h + r_ctx - t
    dist = -torch.norm(err, p=2, dim=-1, keepdim=True)
    comb = torch.cat([iv, sc, err, dist], dim=-1)
    z = _lin(comb, sd, "state_net.0")
    g = F.gelu(F.layer_norm(z, (z.size(-1),), sd["state_net.1.weight"], sd["state_net.1.bias"], LN_EPS))
    if dropout_p > 0.0:
        g = F.dropout(g, p=dropout_p, training=True)
    feat = _lin(g, sd, "state_net.4")
    logit = _lin(feat, sd, "score_head").squeeze(-1)
    return logit, (feat if want_features else None)


@torch.no_grad()
def retriever_forward(
    sd: Dict[str, torch.Tensor],
    *,
    edge_index: torch.Tensor,
    ptr: torch.Tensor,
    question_emb: torch.Tensor,
    node_embeddings: torch.Tensor,
    node_embedding_ids: torch.Tensor,
    edge_embeddings: torch.Tensor,
    topic_one_hot: torch.Tensor,
    direction_mode: str = "bidirectional",
    num_rounds: int = 2,
    num_reverse_rounds: int = 2,
    want_features: bool = False,
    edge_chunk: int = 32768,
    reverse_edge_index: Optional[torch.Tensor] = None,
) -> OracleOutput:
    """``Retriever._forward_impl`` in eval mode (retriever.py:195-289), fp32, same op order as the
    reference (per-edge relation / gate / bias projections included). Edges are processed in
    chunks only to bound memory; per-edge math is independent so results do not change."""
    sd = {k: v.to(torch.float32) if v.is_floating_point() else v for k, v in sd.items()}
    head, tail = edge_index[0].long(), edge_index[1].long()
    E = head.numel()
    eb, _ = compute_edge_batch(edge_index.numpy(), ptr.numpy())
    query_ids = torch.from_numpy(eb)
    if E == 0:
        z = torch.empty(0)
        return OracleOutput(z, query_ids, z, z, torch.empty(0, sd["state_net.4.weight"].size(0)))
    q_proj = torch.tanh(_lin(question_emb.float(), sd, "query_proj.network.0"))
    node_repr = torch.tanh(_lin(node_embeddings.float(), sd, "entity_proj.network.0"))
    non_text = torch.tanh(_lin(sd["non_text_entity_emb.weight"], sd, "entity_proj.network.0"))[0]
    mask = node_embedding_ids.long() == 0  # retriever.py:499-506
    if bool(mask.any()):
        node_repr = torch.where(mask.unsqueeze(-1), non_text.unsqueeze(0), node_repr)
    ns = dde_node_features(topic_one_hot, edge_index.long(), num_rounds, num_reverse_rounds,
                           reverse_edge_index=reverse_edge_index)

    lf_all, lb_all, lo_all, fe_all = [], [], [], []
    for s in range(0, E, edge_chunk):
        sl = slice(s, min(E, s + edge_chunk))
        hh, tt = head[sl], tail[sl]
        q = q_proj[query_ids[sl]]
        h, t = node_repr[hh], node_repr[tt]
        r = torch.tanh(_lin(edge_embeddings[sl].float(), sd, "relation_proj.network.0"))
        s_f = torch.cat([ns[hh], ns[tt]], dim=-1)  # retriever.py:509-517
        s_b = torch.cat([ns[tt], ns[hh]], dim=-1)
        lf = lb = ff = fb = None
        if direction_mode in ("forward", "bidirectional"):
            lf, ff = _score_direction(sd, q, h, r, t, s_f, want_features)
        if direction_mode in ("backward", "bidirectional"):
            lb, fb = _score_direction(sd, q, t, r, h, s_b, want_features)
        if direction_mode == "bidirectional":  # retriever.py:369-381
            st = torch.stack([lf, lb], dim=0)
            w = torch.softmax(st, dim=0)
            lo = (w * st).sum(dim=0)
            fe = w[0].unsqueeze(-1) * ff + w[1].unsqueeze(-1) * fb if want_features else None
        elif direction_mode == "forward":
            lo, fe = lf, ff
        else:
            lo, fe = lb, fb
        lo_all.append(lo)
        lf_all.append(lf)
        lb_all.append(lb)
        fe_all.append(fe)
    cat = lambda xs: torch.cat(xs) if xs[0] is not None else None
    return OracleOutput(cat(lo_all), query_ids, cat(lf_all), cat(lb_all), cat(fe_all))


# ----------------------------------------------------------------------------------------------
# a10-a12: ragged top-k
# ----------------------------------------------------------------------------------------------
def segmented_topk(scores: np.ndarray, edge_ptr: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Per question: ``k_top = min(E_q, k)``; ``topk(scores_q, k_top, largest, sorted)``
    (retriever_metrics.py:134-157, retriever_topk_edge_writer.py:295-303). Tie rule = the
    stable-argsort rule of g_agent_builder.py:640-652: descending score, ascending index.
    Returns ``values [B,k] f32`` (pad -inf), ``indices [B,k] i32`` (question-local, pad -1),
    ``counts [B] i32``. NaN scores order as torch.topk does: NaN is the largest value."""
    scores = np.asarray(scores, dtype=np.float32)
    B = len(edge_ptr) - 1
    vals = np.full((B, k), -np.inf, dtype=np.float32)
    idx = np.full((B, k), -1, dtype=np.int32)
    cnt = np.zeros(B, dtype=np.int32)
    for b in range(B):
        s = scores[int(edge_ptr[b]) : int(edge_ptr[b + 1])]
        kt = min(len(s), k)
        cnt[b] = kt
        if kt == 0:
            continue
        key = np.where(np.isnan(s), np.float32(np.inf), s)
        nan_first = np.isnan(s)
        order = np.lexsort((np.arange(len(s)), -key.astype(np.float64), ~nan_first))[:kt]
        vals[b, :kt] = s[order]
        idx[b, :kt] = order
    return vals, idx, cnt


def sigmoid_f32(x: np.ndarray) -> np.ndarray:
    """``torch.sigmoid`` on fp32 (retriever_topk_edge_writer.py:203)."""
    return torch.sigmoid(torch.from_numpy(np.asarray(x, dtype=np.float32))).numpy()


def writer_payload(
    logits, logits_fwd, logits_bwd, edge_index, edge_attr, labels, node_global_ids, edge_ptr, k
):
    """Writer-flavour selection (retriever_topk_edge_writer.py:201-248, 295-328): sigmoid scores,
    per-question top-k, gathered ``head/rel/tail`` global ids, label, fwd/bwd logits; rank = pos+1."""
    scores = sigmoid_f32(logits)
    vals, idx, cnt = segmented_topk(scores, edge_ptr, k)
    B = len(edge_ptr) - 1
    out = {
        "score": vals,
        "index": idx,
        "count": cnt,
        "head_id": np.full((B, k), -1, np.int64),
        "tail_id": np.full((B, k), -1, np.int64),
        "relation_id": np.full((B, k), -1, np.int64),
        "label": np.zeros((B, k), np.float32),
        "logit_fwd": np.zeros((B, k), np.float32),
        "logit_bwd": np.zeros((B, k), np.float32),
    }
    for b in range(B):
        kt = int(cnt[b])
        g = idx[b, :kt].astype(np.int64) + int(edge_ptr[b])
        out["head_id"][b, :kt] = node_global_ids[edge_index[0, g]]
        out["tail_id"][b, :kt] = node_global_ids[edge_index[1, g]]
        out["relation_id"][b, :kt] = edge_attr[g]
        out["label"][b, :kt] = labels[g]
        out["logit_fwd"][b, :kt] = logits_fwd[g]
        out["logit_bwd"][b, :kt] = logits_bwd[g]
    return out


def recall_at_k(scores, labels, edge_ptr, k_values: Sequence[int]) -> Dict[int, float]:
    """``EdgeRecallAtK`` (retriever_metrics.py:117-170): mean over questions of
    hits@k / max(pos, eps); questions with zero edges are skipped, zero positives count as 0."""
    max_k = max(k_values)
    _, idx, cnt = segmented_topk(scores, edge_ptr, max_k)
    sums = {int(k): 0.0 for k in k_values}
    n = 0
    for b in range(len(edge_ptr) - 1):
        kt = int(cnt[b])
        lab = np.asarray(labels[int(edge_ptr[b]) : int(edge_ptr[b + 1])], dtype=np.float32)
        if lab.size == 0:
            continue
        n += 1
        top = lab[idx[b, :kt]]
        cum = np.cumsum(top)
        denom = max(float(lab.sum()), 1.0)  # _COUNT_EPS = 1.0 (retriever_metrics.py:12,148)
        for k in k_values:
            ke = min(int(k), kt)
            sums[int(k)] += float(cum[ke - 1] / denom) if ke > 0 else 0.0
    return {k: (v / n if n else 0.0) for k, v in sums.items()}


def qa_edge_mask(edge_index: np.ndarray, num_nodes: int, q_local_indices, a_local_indices) -> np.ndarray:
    """``compute_qa_edge_mask`` (src/utils/graph_utils.py:107-153): edges touching a seed or answer node."""
    node_mask = np.zeros(int(num_nodes), dtype=bool)
    qa = np.concatenate([np.asarray(q_local_indices, dtype=np.int64).ravel(), np.asarray(a_local_indices, dtype=np.int64).ravel()])
    if qa.size:
        node_mask[qa] = True
    return node_mask[edge_index[0]] | node_mask[edge_index[1]]


def bridge_recall_at_k(scores, labels, edge_index, edge_ptr, num_nodes, q_local_indices, a_local_indices,
                       k_values: Sequence[int]) -> Dict[int, float]:
    """``BridgeEdgeRecallAtK`` (retriever_metrics.py:173-268): near edges are dropped before the per-graph
    top-k; graphs without a bridge positive are skipped."""
    keep = ~qa_edge_mask(edge_index, num_nodes, q_local_indices, a_local_indices)
    sums = {int(k): 0.0 for k in k_values}
    n = 0
    max_k = max(k_values)
    for b in range(len(edge_ptr) - 1):
        sl = slice(int(edge_ptr[b]), int(edge_ptr[b + 1]))
        sc = np.asarray(scores[sl])[keep[sl]]
        lab = (np.asarray(labels[sl])[keep[sl]] > 0.5).astype(np.float32)
        if sc.size == 0 or lab.sum() <= 0:
            continue
        _, idx, cnt = segmented_topk(sc, np.asarray([0, sc.size]), max_k)
        kt = int(cnt[0])
        cum = np.cumsum(lab[idx[0, :kt]])
        for k in k_values:
            ke = min(int(k), kt)
            sums[int(k)] += float(cum[ke - 1] / max(float(lab.sum()), 1.0)) if ke > 0 else 0.0
        n += 1
    return {k: (v / max(n, 1)) for k, v in sums.items()}


def score_margin(scores, labels, edge_ptr) -> float:
    """``ScoreMargin`` (retriever_metrics.py:331-395): mean over graphs with a positive and a negative of
    min(pos scores) - max(neg scores)."""
    tot, n = 0.0, 0
    for b in range(len(edge_ptr) - 1):
        sl = slice(int(edge_ptr[b]), int(edge_ptr[b + 1]))
        sc, lab = np.asarray(scores[sl], dtype=np.float32), np.asarray(labels[sl]) > 0.5
        if sc.size == 0 or not lab.any() or lab.all():
            continue
        tot += float(sc[lab].min() - sc[~lab].max())
        n += 1
    return tot / max(n, 1)


def reachability_at_k(scores, edge_index, edge_ptr, node_ptr, q_local_indices, q_ptr, a_local_indices, a_ptr,
                      k_values: Sequence[int]) -> Dict[int, float]:
    """``AnswerReachability`` (src/metrics/reachability.py:129-179, 200-228, 329-381): union-find over the first
    min(k, k_top) top-scored edges; a graph is valid when it has edges, in-range seeds and in-range answers."""
    ks = sorted({int(k) for k in k_values if int(k) > 0})
    hits = {k: 0.0 for k in ks}
    valid = 0
    for g in range(len(edge_ptr) - 1):
        e0, e1 = int(edge_ptr[g]), int(edge_ptr[g + 1])
        if e1 <= e0:
            continue
        n0, n1 = int(node_ptr[g]), int(node_ptr[g + 1])
        nn = n1 - n0
        qn = np.asarray(q_local_indices[int(q_ptr[g]):int(q_ptr[g + 1])], dtype=np.int64)
        an = np.asarray(a_local_indices[int(a_ptr[g]):int(a_ptr[g + 1])], dtype=np.int64)
        ql = qn[(qn >= n0) & (qn < n1)] - n0
        al = an[(an >= n0) & (an < n1)] - n0
        if nn <= 0 or ql.size == 0 or al.size == 0:
            continue
        _, idx, cnt = segmented_topk(np.asarray(scores[e0:e1]), np.asarray([0, e1 - e0]), max(ks))
        kt = int(cnt[0])
        parent = list(range(nn))

        def find(x):
            while parent[x] != x:
                parent[x] = parent[parent[x]]
                x = parent[x]
            return x

        def reachable():
            roots = {find(int(s)) for s in ql}
            return any(find(int(a)) in roots for a in al)

        k_check = sorted({min(k, kt) for k in ks})
        reach_map, nxt = {}, 0
        for j in range(kt):
            u, v = int(edge_index[0, e0 + idx[0, j]]) - n0, int(edge_index[1, e0 + idx[0, j]]) - n0
            if 0 <= u < nn and 0 <= v < nn:
                ru, rv = find(u), find(v)
                if ru != rv:
                    parent[rv] = ru
            while nxt < len(k_check) and j + 1 >= k_check[nxt]:
                reach_map[k_check[nxt]] = reachable()
                nxt += 1
        valid += 1
        for k in ks:
            if reach_map[min(k, kt)]:
                hits[k] += 1.0
    return {k: hits[k] / max(valid, 1) for k in ks}


# ----------------------------------------------------------------------------------------------
# a13-a15: BFS hop distances and shortest-path edge labels  scripts/build_retrieval_pipeline.py
# ----------------------------------------------------------------------------------------------
DIST_UNREACHABLE = -1  # build_retrieval_pipeline.py:360


def build_undirected_adjacency(num_nodes: int, edge_src, edge_dst) -> List[List[int]]:
    """build_retrieval_pipeline.py:570-586: both directions, self loop once, drop out-of-range."""
    adj: List[List[int]] = [[] for _ in range(num_nodes)]
    for u, v in zip(edge_src, edge_dst):
        u, v = int(u), int(v)
        if u < 0 or v < 0 or u >= num_nodes or v >= num_nodes:
            continue
        adj[u].append(v)
        if u != v:
            adj[v].append(u)
    for nb in adj:
        nb.sort()
    return adj


def build_directed_adjacency(num_nodes: int, edge_src, edge_dst) -> List[List[int]]:
    """build_retrieval_pipeline.py:589-603."""
    adj: List[List[int]] = [[] for _ in range(num_nodes)]
    for u, v in zip(edge_src, edge_dst):
        u, v = int(u), int(v)
        if u < 0 or v < 0 or u >= num_nodes or v >= num_nodes:
            continue
        adj[u].append(v)
    for nb in adj:
        nb.sort()
    return adj


def bfs_dist(num_nodes: int, adjacency, sources) -> List[int]:
    """build_retrieval_pipeline.py:610-631: multi-source BFS, unreachable = -1."""
    dist = [DIST_UNREACHABLE] * num_nodes
    if num_nodes <= 0:
        return dist
    dq: deque = deque()
    for s in sources:
        s = int(s)
        if 0 <= s < num_nodes and dist[s] < 0:
            dist[s] = 0
            dq.append(s)
    while dq:
        u = dq.popleft()
        du = dist[u] + 1
        for v in adjacency[u]:
            if dist[v] >= 0:
                continue
            dist[v] = du
            dq.append(v)
    return dist


def bfs_dist_np(num_nodes: int, edge_src, edge_dst, sources, *, directed: bool = False,
                max_depth: int = -1) -> np.ndarray:
    """Level-synchronous numpy BFS: same distances as ``bfs_dist`` over
    ``build_(un)directed_adjacency`` (hop distance does not depend on visiting order).
    ``max_depth >= 0`` caps the search (dist > cap -> -1; SURVEY.md §8d config 2 variant)."""
    dist = np.full(num_nodes, DIST_UNREACHABLE, dtype=np.int32)
    if num_nodes <= 0:
        return dist
    u = np.asarray(edge_src, dtype=np.int64)
    v = np.asarray(edge_dst, dtype=np.int64)
    ok = (u >= 0) & (v >= 0) & (u < num_nodes) & (v < num_nodes)
    u, v = u[ok], v[ok]
    if not directed:
        u, v = np.concatenate([u, v]), np.concatenate([v, u])
    src = np.asarray([int(s) for s in sources if 0 <= int(s) < num_nodes], dtype=np.int64)
    dist[src] = 0
    frontier = np.zeros(num_nodes, dtype=bool)
    frontier[src] = True
    level = 0
    while frontier.any() and (max_depth < 0 or level < max_depth):
        level += 1
        cand = v[frontier[u]]
        nxt = np.zeros(num_nodes, dtype=bool)
        nxt[cand] = True
        nxt &= dist < 0
        dist[nxt] = level
        frontier = nxt
    return dist


def shortest_path_union_mask_by_pair(num_nodes, edge_src, edge_dst, sources, targets, *, directed=False):
    """build_retrieval_pipeline.py:691-752 (undirected) and :755-815 (qa_directed). One BFS per
    unique in-range seed and answer; for each reachable (s, a) pair the edges lying on some
    shortest path; returns ``(mask, pair_start, pair_answer, pair_edge_local_ids,
    pair_edge_counts, pair_shortest_lengths)`` as python lists like the reference."""
    num_edges = len(edge_src)
    empty = ([False] * num_edges, [], [], [], [], [])
    if num_nodes <= 0 or num_edges == 0 or len(sources) == 0 or len(targets) == 0:
        return empty
    es = np.asarray(edge_src, dtype=np.int64)
    ed = np.asarray(edge_dst, dtype=np.int64)
    valid = np.nonzero((es >= 0) & (ed >= 0) & (es < num_nodes) & (ed < num_nodes))[0]
    if valid.size == 0:
        return empty
    esv, edv = es[valid], ed[valid]
    starts = sorted({int(s) for s in sources if 0 <= int(s) < num_nodes})
    answers = sorted({int(t) for t in targets if 0 <= int(t) < num_nodes})
    if not starts or not answers:
        return empty
    if directed:
        adj = build_directed_adjacency(num_nodes, edge_src, edge_dst)
        radj = build_directed_adjacency(num_nodes, edge_dst, edge_src)
    else:
        adj = radj = build_undirected_adjacency(num_nodes, edge_src, edge_dst)
    d_s = {s: np.asarray(bfs_dist(num_nodes, adj, [s]), dtype=np.int64) for s in starts}
    d_a = {a: np.asarray(bfs_dist(num_nodes, radj, [a]), dtype=np.int64) for a in answers}
    mask = np.zeros(num_edges, dtype=bool)
    p_s, p_a, p_ids, p_cnt, p_len = [], [], [], [], []
    for s in starts:
        ds = d_s[s]
        for a in answers:
            da = d_a[a]
            dsa = int(ds[a])
            if dsa < 0:
                continue
            p_s.append(s)
            p_a.append(a)
            p_len.append(dsa)
            m = (ds[esv] >= 0) & (da[edv] >= 0) & (ds[esv] + 1 + da[edv] == dsa)
            if not directed:
                m |= (ds[edv] >= 0) & (da[esv] >= 0) & (ds[edv] + 1 + da[esv] == dsa)
            ids = valid[np.nonzero(m)[0]]
            if ids.size:
                mask[ids] = True
                p_ids.extend(ids.tolist())
            p_cnt.append(int(ids.size))
    return mask.tolist(), p_s, p_a, p_ids, p_cnt, p_len


def has_connectivity(graph, seeds, answers, *, path_mode: str = "undirected") -> bool:
    """build_retrieval_pipeline.py:946-979: first-seen local ids, multi-source BFS, any answer reached."""
    if not graph or not seeds or not answers:
        return False
    if path_mode not in ("undirected", "qa_directed"):
        raise ValueError(f"Unsupported path_mode: {path_mode}. Expected one of ('undirected', 'qa_directed').")
    node_index: Dict[str, int] = {}
    es, ed = [], []
    for h, _, t in graph:
        es.append(node_index.setdefault(h, len(node_index)))
        ed.append(node_index.setdefault(t, len(node_index)))
    s_ids = [node_index[s] for s in seeds if s in node_index]
    a_ids = [node_index[a] for a in answers if a in node_index]
    if not s_ids or not a_ids:
        return False
    n = len(node_index)
    adj = build_directed_adjacency(n, es, ed) if path_mode == "qa_directed" else build_undirected_adjacency(n, es, ed)
    dist = bfs_dist(n, adj, s_ids)
    return any(dist[a] >= 0 for a in a_ids)


def labels_for_batch(qb, *, directed: bool = False) -> np.ndarray:
    """Per-question shortest-path-union positive mask over a ragged batch (the ``labels`` field
    the writer reads, built by build_graph via a14). ``qb`` is a synthetic ``QuestionBatch``."""
    labels = np.zeros(qb.num_edges, dtype=np.float32)
    for b in range(qb.num_graphs):
        n0, n1 = int(qb.ptr[b]), int(qb.ptr[b + 1])
        e0, e1 = int(qb.edge_ptr[b]), int(qb.edge_ptr[b + 1])
        s = qb.q_local_indices[int(qb.q_ptr[b]) : int(qb.q_ptr[b + 1])] - n0
        a = qb.a_local_indices[int(qb.a_ptr[b]) : int(qb.a_ptr[b + 1])] - n0
        m = shortest_path_union_mask_by_pair(
            n1 - n0, qb.edge_index[0, e0:e1] - n0, qb.edge_index[1, e0:e1] - n0, s, a, directed=directed
        )[0]
        labels[e0:e1] = np.asarray(m, dtype=np.float32)
    return labels


# ----------------------------------------------------------------------------------------------
# a12: g_agent selection                     src/data/components/g_agent_builder.py:148-724
# ----------------------------------------------------------------------------------------------
GAGENT_PROB_EPS = np.float32(1e-6)  # g_agent_builder.py:20


def node_softmax_logit(scores, heads, tails, num_nodes: int) -> np.ndarray:
    """``_node_softmax_logit`` (g_agent_builder.py:594-626): logit of the mean of the softmax over edges
    sharing the head and the softmax over edges sharing the tail, clamped to [1e-6, 1-1e-6]. fp32,
    sums in ascending edge order (CPU ``scatter_add_``)."""
    s = np.asarray(scores, dtype=np.float32)
    if s.size == 0:
        return s
    h = np.asarray(heads, dtype=np.int64)
    t = np.asarray(tails, dtype=np.int64)

    def side(idx):
        mx = np.full(num_nodes, -np.inf, dtype=np.float32)
        np.maximum.at(mx, idx, s)
        ex = np.exp(s - mx[idx]).astype(np.float32)
        sm = np.zeros(num_nodes, dtype=np.float32)
        np.add.at(sm, idx, ex)
        return (ex / np.maximum(sm[idx], GAGENT_PROB_EPS)).astype(np.float32)

    prob = ((side(h) + side(t)) * np.float32(0.5)).astype(np.float32)
    prob = np.clip(prob, GAGENT_PROB_EPS, np.float32(1.0) - GAGENT_PROB_EPS).astype(np.float32)
    return (np.log(prob) - np.log1p(-prob)).astype(np.float32)


def gagent_select_topk(scores, edge_top_k: int) -> np.ndarray:
    """``_select_topk_edges`` (g_agent_builder.py:640-652): every edge (index order) when E <= k, else the
    first k of a stable descending argsort."""
    s = np.asarray(scores, dtype=np.float32).reshape(-1)
    if s.size <= edge_top_k:
        return np.arange(s.size, dtype=np.int64)
    return np.argsort(-s.astype(np.float64), kind="stable")[:edge_top_k].astype(np.int64)


def gagent_start_edge_budget(deg, keep_ratio: float, min_edges: int, max_edges: Optional[int]) -> np.ndarray:
    """k_s = min(deg, clamp(ceil(deg * ratio) in fp32, min_edges, max_edges)) (g_agent_builder.py:676-686)."""
    deg = np.asarray(deg, dtype=np.int64)
    k = np.ceil(deg.astype(np.float32) * np.float32(keep_ratio)).astype(np.int64)
    if min_edges > 0:
        k = np.maximum(k, min_edges)
    if max_edges is not None:
        k = np.zeros_like(k) if max_edges == 0 else np.minimum(k, max_edges)
    return np.minimum(k, deg)


def gagent_select_start_edges(heads, tails, scores, start_nodes, num_nodes: int, keep_ratio: float = 0.25,
                              min_edges: int = 1, max_edges: Optional[int] = None) -> np.ndarray:
    """``_select_start_edges`` (g_agent_builder.py:654-724): for every unique start node the k_s best incident
    entries, the incident list being (edges with head == s, in edge order) then (edges with tail == s),
    ranked by a stable descending sort; returns the sorted unique edge ids."""
    h = np.asarray(heads, dtype=np.int64).reshape(-1)
    t = np.asarray(tails, dtype=np.int64).reshape(-1)
    s = np.asarray(scores, dtype=np.float32).reshape(-1)
    starts = np.unique(np.asarray(start_nodes, dtype=np.int64))
    if starts.size == 0 or s.size == 0:
        return np.empty(0, dtype=np.int64)
    deg = np.bincount(h, minlength=num_nodes) + np.bincount(t, minlength=num_nodes)
    budget = gagent_start_edge_budget(deg[starts], keep_ratio, min_edges, max_edges)
    picked: List[int] = []
    for node, k_s in zip(starts.tolist(), budget.tolist()):
        if k_s <= 0:
            continue
        inc_edges = np.concatenate([np.nonzero(h == node)[0], np.nonzero(t == node)[0]])
        order = np.argsort(-s[inc_edges].astype(np.float64), kind="stable")[:k_s]
        picked.extend(inc_edges[order].tolist())
    return np.unique(np.asarray(picked, dtype=np.int64))


def gagent_build_sample(*, heads, tails, relations, labels, scores, node_global_ids, node_embedding_ids,
                        start_entity_ids, answer_entity_ids, edge_top_k: int = 50, score_mode: str = "node_softmax",
                        start_keep_ratio: float = 0.25, start_min_edges: int = 1,
                        start_max_edges: Optional[int] = None, allow_empty_answer: bool = False,
                        selection_scores=None):
    """``GAgentBuilder._build_and_add_sample`` steps A, C, D (g_agent_builder.py:236-470) for one question:
    selection on the normalised scores, union with the start edges, (h, r, t) dedup in first-seen order
    with max raw score / label, re-indexing on the sorted unique entity ids, re-normalised scores, start /
    answer locals. ``heads`` / ``tails`` are question-local node indices; ``scores`` are the calibrated
    logits; ``selection_scores`` (tests only) replaces the normalised scores of step A so that a selection can
    be checked bit-exactly on scores computed elsewhere. Returns ``None`` where the reference drops the sample, else a dict of arrays."""
    h = np.asarray(heads, dtype=np.int64)
    t = np.asarray(tails, dtype=np.int64)
    r = np.asarray(relations, dtype=np.int64)
    raw = np.asarray(scores, dtype=np.float32)
    lab = np.asarray(labels, dtype=np.float32)
    gids = np.asarray(node_global_ids, dtype=np.int64)
    emb = np.asarray(node_embedding_ids, dtype=np.int64)
    starts_e = np.asarray(start_entity_ids, dtype=np.int64).reshape(-1)
    answers_e = np.asarray(answer_entity_ids, dtype=np.int64).reshape(-1)
    N = int(gids.size)
    if h.size == 0:
        return None
    if start_max_edges is None:
        start_max_edges = edge_top_k  # GAgentSettings.__post_init__, g_agent_builder.py:71-72
    norm = (lambda s_, h_, t_, n_: s_) if score_mode == "logits" else node_softmax_logit
    start_locals = np.nonzero(np.isin(gids, starts_e))[0]
    if start_locals.size == 0:
        raise ValueError("Start entities missing from retrieval graph")
    sel_scores = norm(raw, h, t, N) if selection_scores is None else np.asarray(selection_scores, dtype=np.float32)
    top = gagent_select_topk(sel_scores, edge_top_k)
    st = gagent_select_start_edges(h, t, sel_scores, start_locals, N, start_keep_ratio, start_min_edges,
                                   start_max_edges)
    env = np.unique(np.concatenate([top, st]))
    # (h, r, t) dedup, first-seen order, max aggregation (g_agent_builder.py:338-354)
    agg: Dict[Tuple[int, int, int], List[float]] = {}
    for e in env.tolist():
        key = (int(gids[h[e]]), int(r[e]), int(gids[t[e]]))
        cur = agg.get(key)
        if cur is None:
            agg[key] = [float(raw[e]), float(lab[e])]
        else:
            cur[0] = max(cur[0], float(raw[e]))
            cur[1] = max(cur[1], float(lab[e]))
    triples = list(agg.keys())
    eh = np.asarray([k[0] for k in triples], dtype=np.int64)
    er = np.asarray([k[1] for k in triples], dtype=np.int64)
    et = np.asarray([k[2] for k in triples], dtype=np.int64)
    es = np.asarray([agg[k][0] for k in triples], dtype=np.float32)
    el = np.asarray([agg[k][1] for k in triples], dtype=np.float32)
    node_entity_ids = np.unique(np.concatenate([eh, et]))
    lookup = {int(g): int(e_) for g, e_ in zip(gids.tolist(), emb.tolist())}
    node_emb = np.asarray([lookup[int(g)] for g in node_entity_ids.tolist()], dtype=np.int64)
    hl = np.searchsorted(node_entity_ids, eh).astype(np.int64)
    tl = np.searchsorted(node_entity_ids, et).astype(np.int64)
    es_norm = norm(es, hl, tl, int(node_entity_ids.size))

    def ordered_locals(ids):
        out, seen = [], set()
        for g in ids.tolist():
            p = int(np.searchsorted(node_entity_ids, g))
            if p < node_entity_ids.size and node_entity_ids[p] == g and p not in seen:
                seen.add(p)
                out.append(p)
        return np.asarray(out, dtype=np.int64)

    start_node_locals = ordered_locals(starts_e)
    if start_node_locals.size == 0:
        return None
    seen_a, ordered_a = set(), []
    for a in answers_e.tolist():
        if a not in seen_a:
            seen_a.add(a)
            ordered_a.append(a)
    answers_u = np.asarray(ordered_a, dtype=np.int64)
    # the reference does not dedup answer locals beyond the id dedup (g_agent_builder.py:423-428)
    answer_node_locals = ordered_locals(answers_u)
    dummy = answer_node_locals.size == 0
    if dummy:
        if not allow_empty_answer:
            return None
        el = np.zeros_like(el)
    return {
        "env_edge_ids": env, "edge_heads_global": eh, "edge_relations": er, "edge_tails_global": et,
        "edge_scores": es_norm, "edge_labels": el, "edge_head_locals": hl, "edge_tail_locals": tl,
        "node_entity_ids": node_entity_ids, "node_embedding_ids": node_emb,
        "start_node_locals": start_node_locals, "answer_entity_ids": answers_u,
        "answer_node_locals": answer_node_locals, "is_dummy_agent": dummy,
        "is_answer_reachable": not dummy,
    }


# ----------------------------------------------------------------------------------------------
# a16: deterministic single shortest path          scripts/build_retrieval_pipeline.py:453-530
# ----------------------------------------------------------------------------------------------
def shortest_path_single(num_nodes: int, edge_src, edge_dst, sources, targets) -> Tuple[List[int], List[int]]:
    """FIFO BFS from the ascending unique in-range sources over adjacency lists sorted by (neighbour, edge
    idx); best target = min distance then min id; the path follows parent / parent_edge back to a source.
    Returns ``(edges, nodes)``; ``([], [])`` without a path, ``([], [t])`` when the target is a source."""
    if len(sources) == 0 or len(targets) == 0 or num_nodes <= 0:
        return [], []
    adj: List[List[Tuple[int, int]]] = [[] for _ in range(num_nodes)]
    for idx, (u, v) in enumerate(zip(edge_src, edge_dst)):
        u, v = int(u), int(v)
        if 0 <= u < num_nodes and 0 <= v < num_nodes:
            adj[u].append((v, idx))
            if u != v:
                adj[v].append((u, idx))
    for a in adj:
        a.sort()
    src_u = sorted({int(s) for s in sources if 0 <= int(s) < num_nodes})
    tgt_u = sorted({int(t) for t in targets if 0 <= int(t) < num_nodes})
    if not src_u or not tgt_u:
        return [], []
    dist = [-1] * num_nodes
    parent = [-1] * num_nodes
    parent_edge = [-1] * num_nodes
    q = deque()
    for s in src_u:
        dist[s] = 0
        q.append(s)
    while q:
        cur = q.popleft()
        for nb, e in adj[cur]:
            if dist[nb] == -1:
                dist[nb] = dist[cur] + 1
                parent[nb], parent_edge[nb] = cur, e
                q.append(nb)
    reach = [(dist[t], t) for t in tgt_u if dist[t] >= 0]
    if not reach:
        return [], []
    _, cur = min(reach)
    nodes, edges = [cur], []
    while dist[cur] != 0:
        edges.append(parent_edge[cur])
        cur = parent[cur]
        nodes.append(cur)
    return edges[::-1], nodes[::-1]


# ----------------------------------------------------------------------------------------------
# config 4: single-graph retrieval (no reference code path; SURVEY.md §8(d) row 4 defines it as the
# composition of _bfs_dist and the reference scorer on the extracted subgraph)
# ----------------------------------------------------------------------------------------------
def pool_candidates(heads, tails, num_nodes: int, seeds, *, depth: int = 2, cap: int = 20000):
    """One query: capped multi-source BFS from ``seeds`` over the undirected pool, candidates = edges with both
    endpoints reached, the ``cap`` smallest by ``(dist_h + dist_t, edge id)``, returned in ascending edge id.
    Returns ``(edge_ids [C] i64, dist [num_nodes] i32)``."""
    h = np.asarray(heads, dtype=np.int64)
    t = np.asarray(tails, dtype=np.int64)
    dist = bfs_dist_np(num_nodes, h, t, seeds, max_depth=depth)
    ok = (dist[h] >= 0) & (dist[t] >= 0)
    eid = np.nonzero(ok)[0]
    prio = dist[h[eid]].astype(np.int64) + dist[t[eid]].astype(np.int64)
    order = np.lexsort((eid, prio))[:cap]
    return np.sort(eid[order]), dist


def pool_subgraph_batch(heads, rels, tails, num_nodes: int, seeds_per_query, *, depth: int = 2, cap: int = 20000,
                        node_embedding_ids=None):
    """Candidate subgraphs of a query batch in the reference's batch layout (nodes = ascending global ids)."""
    h_all, r_all, t_all = (np.asarray(x, dtype=np.int64) for x in (heads, rels, tails))
    ei, ea, eg, ng, topic, ptr, eptr = [], [], [], [], [], [0], [0]
    for seeds in seeds_per_query:
        eid, dist = pool_candidates(h_all, t_all, num_nodes, seeds, depth=depth, cap=cap)
        nodes = np.unique(np.concatenate([h_all[eid], t_all[eid]])) if eid.size else np.empty(0, np.int64)
        hl = np.searchsorted(nodes, h_all[eid]) + ptr[-1]
        tl = np.searchsorted(nodes, t_all[eid]) + ptr[-1]
        ei.append(np.stack([hl, tl]).astype(np.int64))
        ea.append(r_all[eid])
        eg.append(eid)
        ng.append(nodes)
        oh = np.zeros((nodes.size, 2), dtype=np.float32)
        oh[np.arange(nodes.size), (dist[nodes] == 0).astype(np.int64)] = 1.0
        topic.append(oh)
        ptr.append(ptr[-1] + nodes.size)
        eptr.append(eptr[-1] + eid.size)
    node_global = np.concatenate(ng) if ng else np.empty(0, np.int64)
    emb = node_global if node_embedding_ids is None else np.asarray(node_embedding_ids, dtype=np.int64)[node_global]
    return {
        "edge_index": np.concatenate(ei, axis=1) if ei else np.zeros((2, 0), np.int64),
        "edge_attr": np.concatenate(ea) if ea else np.empty(0, np.int64),
        "edge_global_id": np.concatenate(eg) if eg else np.empty(0, np.int64),
        "node_global_ids": node_global, "node_embedding_ids": emb,
        "topic_one_hot": np.concatenate(topic) if topic else np.zeros((0, 2), np.float32),
        "ptr": np.asarray(ptr, dtype=np.int64), "edge_ptr": np.asarray(eptr, dtype=np.int64),
    }


# ----------------------------------------------------------------------------------------------
# (f) rank 4: RetrieverLoss                                   src/losses/retriever_loss.py:72-297
# ----------------------------------------------------------------------------------------------
def retriever_loss(logits, targets, edge_batch, num_graphs: int, *, infonce_temperature: float = 1.0,
                   infonce_weight: float = 1.0, bce_weight: float = 0.0, edge_is_near=None,
                   edge_weight_near: float = 1.0, edge_weight_bridge: float = 1.0):
    """Multi-positive InfoNCE per question (``logsumexp(all) - logsumexp(positives)``, mean over questions with at
    least one positive and one negative, :72-146) plus optional BCE (:148-178), and the analytic gradient with
    respect to the logits. float64. Returns ``(total, grad [E], info dict)``."""
    x = np.asarray(logits, dtype=np.float64)
    y = np.asarray(targets, dtype=np.float64)
    g = np.asarray(edge_batch, dtype=np.int64)
    weighted = edge_weight_near != 1.0 or edge_weight_bridge != 1.0
    w = np.where(np.asarray(edge_is_near, dtype=bool), edge_weight_near, edge_weight_bridge) if weighted else np.ones_like(x)
    pos = y > 0.5
    s = x / infonce_temperature + (np.log(np.maximum(w, 1e-6)) if weighted else 0.0)
    grad = np.zeros_like(x)
    inf_terms, bce_terms, inf_valid, bce_valid = [], [], [], []
    for q in range(num_graphs):
        m = g == q
        if not m.any():
            continue
        sq, pq = s[m], pos[m]
        if pq.any() and (~pq).any():
            lse_all = np.logaddexp.reduce(sq)
            lse_pos = np.logaddexp.reduce(sq[pq])
            inf_terms.append(lse_all - lse_pos)
            inf_valid.append(q)
        denom = max(w[m].sum(), 1e-6) if weighted else float(m.sum())
        if (w[m].sum() > 0) if weighted else True:
            xb, yb = x[m], y[m]
            bce_terms.append((w[m] * (np.maximum(xb, 0) - xb * yb + np.log1p(np.exp(-np.abs(xb))))).sum() / denom)
            bce_valid.append((q, denom))
    any_inf = pos.any() and (~pos).any() and len(inf_terms) > 0
    l_inf = float(np.mean(inf_terms)) if any_inf else 0.0
    l_bce = float(np.mean(bce_terms)) if (bce_weight > 0 and bce_terms) else 0.0
    if any_inf and infonce_weight > 0:
        for q in inf_valid:
            m = g == q
            sq, pq = s[m], pos[m]
            d = np.exp(sq - np.logaddexp.reduce(sq))
            d[pq] -= np.exp(sq[pq] - np.logaddexp.reduce(sq[pq]))
            grad[m] += infonce_weight * d / (infonce_temperature * len(inf_valid))
    if bce_weight > 0 and bce_terms:
        for q, denom in bce_valid:
            m = g == q
            grad[m] += bce_weight * w[m] * (1 / (1 + np.exp(-x[m])) - y[m]) / (denom * len(bce_valid))
    sig = 1 / (1 + np.exp(-x))
    info = {"infonce": l_inf, "bce": l_bce, "infonce_graphs": float(len(inf_valid)) if any_inf else 0.0,
            "pos_prob": float(sig[pos].mean()) if pos.any() else 0.0,
            "neg_prob": float(sig[~pos].mean()) if (~pos).any() else 0.0}
    return infonce_weight * l_inf + bce_weight * l_bce, grad, info


# ----------------------------------------------------------------------------------------------
# a7: hide-and-seek bias (src/models/components/retriever.py:307-367)
# ----------------------------------------------------------------------------------------------
def hide_seek_bias(edge_index: torch.Tensor, like: torch.Tensor, *, p_near: float, p_far: float, bias_near: float,
                   bias_far: float, num_nodes: Optional[int] = None, q_local_indices=None, a_local_indices=None,
                   edge_is_near=None) -> Optional[torch.Tensor]:
    """Restatement of ``_compute_hide_seek_bias`` for an enabled module: near mask from ``edge_is_near`` or from the
    q/a local indices (``compute_qa_edge_mask``, graph_utils.py:107-153); one uniform draw per edge from the torch
    generator of ``like``'s device (``torch.rand_like``, retriever.py:362); edges whose draw falls below p_near / p_far
    receive bias_near / bias_far. Returns None in the reference's early-exit cases (:352-359)."""
    dev = edge_index.device
    if edge_is_near is None:
        mask = torch.zeros(int(num_nodes), dtype=torch.bool, device=dev)
        qa = torch.cat([torch.as_tensor(q_local_indices, dtype=torch.long, device=dev).view(-1),
                        torch.as_tensor(a_local_indices, dtype=torch.long, device=dev).view(-1)])
        if qa.numel():
            mask[qa] = True
        near = mask[edge_index[0]] | mask[edge_index[1]]
    else:
        near = torch.as_tensor(edge_is_near).to(device=dev, dtype=torch.bool)
    if near.numel() == 0 or (p_near <= 0.0 and p_far <= 0.0) or (bias_near == 0.0 and bias_far == 0.0):
        return None
    drop_prob = torch.where(near, like.new_tensor(p_near), like.new_tensor(p_far))
    drop = torch.rand_like(drop_prob) < drop_prob
    values = torch.where(near, like.new_tensor(bias_near), like.new_tensor(bias_far))
    return torch.where(drop, values, like.new_zeros(drop_prob.shape))


# ----------------------------------------------------------------------------------------------
# (f) rank 4: the training-mode forward (src/models/components/retriever.py:195-289 with self.training)
# ----------------------------------------------------------------------------------------------
def retriever_train_forward(sd: Dict[str, torch.Tensor], *, edge_index, ptr, question_emb, node_embeddings,
                            node_embedding_ids, edge_embeddings, topic_one_hot, direction_mode: str = "bidirectional",
                            num_rounds: int = 2, num_reverse_rounds: int = 2, dropout_p: float = 0.1,
                            hide_seek: Optional[Dict] = None) -> OracleOutput:
    """The reference's forward under autograd (``sd`` holds the parameters, some requiring grad), on the device of
    its tensors, fp32. Dropout (forward direction first, then backward) and the hide-and-seek draw consume the torch
    generator in the reference's order, so a seeded run reproduces the reference's training logits. ``hide_seek`` =
    keyword arguments of ``hide_seek_bias`` (an enabled hide_seek_cfg), or None."""
    dev = sd["score_head.weight"].device
    ei = edge_index.to(dev).long()
    head, tail = ei[0], ei[1]
    eb, _ = compute_edge_batch(edge_index.cpu().numpy(), ptr.cpu().numpy())
    query_ids = torch.from_numpy(eb).to(dev)
    f32 = lambda x: x.to(device=dev, dtype=torch.float32)
    q_proj = torch.tanh(_lin(f32(question_emb), sd, "query_proj.network.0"))
    node_repr = torch.tanh(_lin(f32(node_embeddings), sd, "entity_proj.network.0"))
    non_text = torch.tanh(_lin(sd["non_text_entity_emb.weight"], sd, "entity_proj.network.0"))[0]
    mask = node_embedding_ids.to(dev).long() == 0
    node_repr = torch.where(mask.unsqueeze(-1), non_text.unsqueeze(0), node_repr)
    with torch.no_grad():
        ns = dde_node_features(topic_one_hot.cpu(), edge_index.cpu().long(), num_rounds, num_reverse_rounds).to(dev)
    q = q_proj[query_ids]
    h, t = node_repr[head], node_repr[tail]
    r = torch.tanh(_lin(f32(edge_embeddings), sd, "relation_proj.network.0"))
    s_f, s_b = torch.cat([ns[head], ns[tail]], dim=-1), torch.cat([ns[tail], ns[head]], dim=-1)
    lf = lb = None
    if direction_mode in ("forward", "bidirectional"):
        lf, _ = _score_direction(sd, q, h, r, t, s_f, False, dropout_p)
    if direction_mode in ("backward", "bidirectional"):
        lb, _ = _score_direction(sd, q, t, r, h, s_b, False, dropout_p)
    if hide_seek is not None:
        bias = hide_seek_bias(ei, lf if lf is not None else lb, **hide_seek)
        if bias is not None:
            lf = lf + bias if lf is not None else None
            lb = lb + bias if lb is not None else None
    if direction_mode == "bidirectional":
        st = torch.stack([lf, lb], dim=0)
        lo = (torch.softmax(st, dim=0) * st).sum(dim=0)
    else:
        lo = lf if lf is not None else lb
    return OracleOutput(lo, query_ids, lf, lb, None)
