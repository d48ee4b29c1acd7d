"""GPU parity of the g_agent selection against the golden samples produced by the unmodified
reference ``GAgentBuilder`` and against the CPU oracle. Integer outputs (kept triples, locals, entity ids) are
bit-exact. Scores are logit(p) of an fp32 probability: near p -> 1 one ulp of p (6e-8) moves the logit by up to
6e-8 / (1 - p), so a score passes when |d logit| <= 1e-4 or |d p| <= 4e-7 (a few fp32 ulps of p; expf / logf
rounding and summation order are the only sources of difference)."""
import os

import numpy as np
import pytest
import torch

from oracle import evi_oracle as orc
from evi_rag_b200 import gagent as GA
from evi_rag_b200.synthetic import generate_questions, tiny_spec

pytestmark = pytest.mark.gpu

SCORE_ATOL = 1e-4
PROB_ATOL = 4e-7


def score_error(got, ref):
    """Worst score error under the rule above, scaled so that <= 1 passes."""
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    if got.size == 0:
        return 0.0
    d_logit = np.abs(got - ref) / SCORE_ATOL
    d_prob = np.abs(1 / (1 + np.exp(-got)) - 1 / (1 + np.exp(-ref))) / PROB_ATOL
    return float(np.minimum(d_logit, d_prob).max())


SETTINGS = {
    "default": dict(edge_top_k=50),
    "logits_k20": dict(edge_top_k=20, score_mode="logits", start_keep_ratio=0.5, start_min_edges=2, start_max_edges=5),
    "dummy_ok": dict(edge_top_k=50, allow_empty_answer=True, score_temperature=2.0, score_bias=0.5),
    "no_start_edges": dict(edge_top_k=64, start_max_edges=0, start_min_edges=0),
}
INT_FIELDS = ["edge_relations", "edge_head_locals", "edge_tail_locals", "node_entity_ids", "node_embedding_ids",
              "start_node_locals", "answer_entity_ids", "answer_node_locals"]


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def _ragged_lists(flat, ptr):
    return [flat[int(ptr[i]):int(ptr[i + 1])].tolist() for i in range(len(ptr) - 1)]


@pytest.mark.parametrize("tag", list(SETTINGS))
def test_gagent_builder_matches_reference_golden(dev, golden_dir, tag):
    import types

    g = np.load(os.path.join(golden_dir, "gagent.npz"), allow_pickle=False)
    B = len(g["ptr"]) - 1
    starts = _ragged_lists(g["start_entity_ids"], g["start_ptr"])
    answers = _ragged_lists(g["answer_entity_ids"], g["answer_ptr"])

    class Store:
        def load_sample(self, sid):
            i = int(sid[1:])
            return {"question_emb": np.zeros(4, np.float32), "question": f"question {i}",
                    "seed_entity_ids": starts[i], "answer_entity_ids": answers[i]}

    batch = types.SimpleNamespace(
        ptr=torch.from_numpy(g["ptr"]), edge_index=torch.from_numpy(g["edge_index"]),
        edge_attr=torch.from_numpy(g["edge_attr"]), labels=torch.from_numpy(g["labels"]),
        node_global_ids=torch.from_numpy(g["node_global_ids"]),
        node_embedding_ids=torch.from_numpy(g["node_embedding_ids"]), sample_id=[f"s{i}" for i in range(B)])
    qid = torch.from_numpy(np.repeat(np.arange(B), np.diff(g["edge_ptr"])))
    out = types.SimpleNamespace(logits=torch.from_numpy(g["logits"]).to(dev), query_ids=qid.to(dev))
    builder = GA.GAgentBuilder(GA.GAgentSettings(**SETTINGS[tag]), embedding_store=Store())
    builder.process_batch(batch, out)
    kept = np.zeros(B, dtype=bool)
    for s in builder.samples:
        i = int(s.sample_id[1:])
        kept[i] = True
        for f in INT_FIELDS:
            assert np.array_equal(getattr(s, f).numpy(), g[f"{tag}_q{i}_{f}"]), (tag, i, f)
        assert score_error(s.edge_scores.numpy(), g[f"{tag}_q{i}_edge_scores"]) <= 1.0
        assert np.array_equal(s.edge_labels.numpy(), g[f"{tag}_q{i}_edge_labels"])
        assert bool(s.is_dummy_agent) == bool(g[f"{tag}_q{i}_is_dummy_agent"])
        assert s.edge_head_locals.dtype == torch.long and s.node_entity_ids.dtype == torch.long
    assert np.array_equal(kept, g[f"{tag}_kept"])
    assert builder.stats["retrieval_failed"] == int(g[f"{tag}_retrieval_failed"])
    rec = builder._sample_to_record(builder.samples[0])
    assert set(rec) >= {"sample_id", "edge_scores", "pair_edge_counts", "is_dummy_agent", "gt_path_exists"}


def test_node_softmax_helper_golden_and_deterministic(dev, golden_dir):
    g = np.load(os.path.join(golden_dir, "gagent.npz"), allow_pickle=False)
    e0, e1, n0, n1 = int(g["edge_ptr"][2]), int(g["edge_ptr"][3]), int(g["ptr"][2]), int(g["ptr"][3])
    h = torch.from_numpy(g["edge_index"][0, e0:e1] - n0).to(dev)
    t = torch.from_numpy(g["edge_index"][1, e0:e1] - n0).to(dev)
    s = torch.from_numpy(g["logits"][e0:e1]).to(dev)
    cal, ns = GA.node_softmax_logit(s, h, t, n1 - n0)
    assert torch.equal(cal, s)
    assert score_error(ns.cpu().numpy(), g["helper_node_softmax"]) <= 1.0
    for _ in range(3):  # fixed-point sums: bit-stable run to run
        assert torch.equal(GA.node_softmax_logit(s, h, t, n1 - n0)[1], ns)
    cal2, lg = GA.node_softmax_logit(s, h, t, n1 - n0, temperature=2.0, bias=0.5, score_mode="logits")
    ref = torch.from_numpy(g["logits"][e0:e1]) / 2.0 + 0.5
    assert torch.equal(cal2.cpu(), ref) and torch.equal(lg.cpu(), ref)


@pytest.mark.parametrize("mode,k", [("node_softmax", 50), ("logits", 200), ("node_softmax", 500)])
def test_gagent_select_batch_vs_oracle(dev, mode, k):
    """WebQSP-shaped questions; the oracle's selection runs on the GPU's selection scores so every integer
    output must be bit-exact, and the GPU selection scores are checked against the oracle within SCORE_ATOL."""
    spec = tiny_spec(12, 1500, seed=31 + k, num_relations=40, num_entities=4000)
    qb = generate_questions(spec)
    rng = np.random.default_rng(k)
    E, B = qb.num_edges, qb.num_graphs
    logits = (rng.standard_normal(E) * 2.5).astype(np.float32)
    logits[rng.integers(0, E, E // 10)] = 1.25  # exact ties
    labels = (rng.random(E) < 0.02).astype(np.float32)
    starts, answers = [], []
    for b in range(B):
        sq = qb.q_local_indices[qb.q_ptr[b]:qb.q_ptr[b + 1]]
        aq = qb.a_local_indices[qb.a_ptr[b]:qb.a_ptr[b + 1]]
        starts.append(qb.node_global_ids[sq].tolist() + [10 ** 9 + b])
        answers.append(qb.node_global_ids[aq].tolist() + [10 ** 9 + 7] + qb.node_global_ids[aq][:1].tolist())
    cfg = GA.GAgentSettings(edge_top_k=k, score_mode=mode, start_keep_ratio=0.1, start_min_edges=3, start_max_edges=40,
                            allow_empty_answer=True)
    res = GA.select_batch(
        logits=torch.from_numpy(logits).to(dev), edge_index=torch.from_numpy(qb.edge_index), edge_ptr=torch.from_numpy(qb.edge_ptr),
        ptr=torch.from_numpy(qb.ptr), edge_attr=torch.from_numpy(qb.edge_attr), labels=torch.from_numpy(labels),
        node_global_ids=torch.from_numpy(qb.node_global_ids), node_embedding_ids=torch.from_numpy(qb.node_embedding_ids),
        start_entity_ids=starts, answer_entity_ids=answers, settings=cfg)
    sel = res["selection_scores"].cpu().numpy()
    worst = 0.0
    for b in range(B):
        e0, e1, n0, n1 = int(qb.edge_ptr[b]), int(qb.edge_ptr[b + 1]), int(qb.ptr[b]), int(qb.ptr[b + 1])
        h, t = qb.edge_index[0, e0:e1] - n0, qb.edge_index[1, e0:e1] - n0
        if mode == "node_softmax":
            ref_sel = orc.node_softmax_logit(logits[e0:e1], h, t, n1 - n0)
            worst = max(worst, score_error(sel[e0:e1], ref_sel))
        else:
            assert np.array_equal(sel[e0:e1], logits[e0:e1])
        ref = orc.gagent_build_sample(
            heads=h, tails=t, relations=qb.edge_attr[e0:e1], labels=labels[e0:e1], scores=logits[e0:e1],
            node_global_ids=qb.node_global_ids[n0:n1], node_embedding_ids=qb.node_embedding_ids[n0:n1],
            start_entity_ids=starts[b], answer_entity_ids=answers[b], edge_top_k=k, score_mode=mode,
            start_keep_ratio=0.1, start_min_edges=3, start_max_edges=40, allow_empty_answer=True,
            selection_scores=sel[e0:e1])
        st = int(res["status"][b])
        assert st & ~16 == 0, (b, st)
        assert ref is not None
        m, nu = int(res["num_triples"][b]), int(res["num_nodes"][b])
        assert m == len(ref["edge_relations"]) and nu == len(ref["node_entity_ids"])
        assert np.array_equal(res["edge_heads_global"][b, :m].numpy(), ref["edge_heads_global"])
        assert np.array_equal(res["edge_tails_global"][b, :m].numpy(), ref["edge_tails_global"])
        assert np.array_equal(res["edge_relations"][b, :m].numpy(), ref["edge_relations"])
        assert np.array_equal(res["edge_head_locals"][b, :m].numpy(), ref["edge_head_locals"])
        assert np.array_equal(res["edge_tail_locals"][b, :m].numpy(), ref["edge_tail_locals"])
        assert np.array_equal(res["node_entity_ids"][b, :nu].numpy(), ref["node_entity_ids"])
        assert np.array_equal(res["node_embedding_ids"][b, :nu].numpy(), ref["node_embedding_ids"])
        assert np.array_equal(res["edge_labels"][b, :m].numpy(), ref["edge_labels"]) or ref["is_dummy_agent"]
        assert score_error(res["edge_scores"][b, :m].numpy(), ref["edge_scores"]) <= 1.0
        nsl, nal, nau = int(res["num_start_locals"][b]), int(res["num_answer_locals"][b]), int(res["num_answers"][b])
        assert np.array_equal(res["start_node_locals"][b, :nsl].numpy(), ref["start_node_locals"])
        assert np.array_equal(res["answer_node_locals"][b, :nal].numpy(), ref["answer_node_locals"])
        assert np.array_equal(res["answer_entity_ids"][b, :nau].numpy(), ref["answer_entity_ids"])
        assert bool(st & 16) == bool(ref["is_dummy_agent"])
    print(f"[parity] gagent {mode} k={k}: worst selection-score error = {worst:.3f} of the tolerance")
    assert worst <= 1.0


def test_gagent_errors(dev):
    spec = tiny_spec(2, 200, seed=3, num_relations=8, num_entities=500)
    qb = generate_questions(spec)
    kw = dict(logits=torch.zeros(qb.num_edges, device=dev), edge_index=torch.from_numpy(qb.edge_index),
              edge_ptr=torch.from_numpy(qb.edge_ptr), ptr=torch.from_numpy(qb.ptr), edge_attr=torch.from_numpy(qb.edge_attr),
              labels=None, node_global_ids=torch.from_numpy(qb.node_global_ids),
              node_embedding_ids=torch.from_numpy(qb.node_embedding_ids))
    res = GA.select_batch(**kw, start_entity_ids=[[-5], [int(qb.node_global_ids[qb.ptr[1]])]], answer_entity_ids=[[1], [2]],
                          settings=GA.GAgentSettings())
    assert int(res["status"][0]) & 2  # start entities missing from the retrieval graph
    assert int(res["status"][1]) & 2 == 0
    with pytest.raises(RuntimeError):
        GA.select_batch(**{**kw, "logits": torch.zeros(qb.num_edges)}, start_entity_ids=[[1], [2]],
                        answer_entity_ids=[[1], [2]], settings=GA.GAgentSettings())
    with pytest.raises(ValueError):
        GA.GAgentSettings(edge_top_k=0)
    with pytest.raises(ValueError):
        GA.GAgentSettings(start_min_edges=9, start_max_edges=3)


def test_gagent_question_without_edges_is_skipped(dev):
    """g_agent_builder.py:205-207: a question with no edges is skipped, the others are unaffected."""
    import types

    ei = torch.tensor([[0, 1, 5, 6], [1, 2, 6, 5]])
    ptr = torch.tensor([0, 3, 5, 7])             # q1 has nodes 3..4 and no edges
    batch = types.SimpleNamespace(ptr=ptr, edge_index=ei, edge_attr=torch.tensor([0, 1, 2, 2]), labels=torch.tensor([1., 0, 0, 1]),
                                  node_global_ids=torch.arange(100, 107), node_embedding_ids=torch.arange(7),
                                  sample_id=["a", "b", "c"])

    class Store:
        def load_sample(self, sid):
            return {"question_emb": [0.0, 1.0], "question": sid, "seed_entity_ids": {"a": [100], "c": [105]}[sid],
                    "answer_entity_ids": {"a": [102], "c": [106]}[sid]}

    out = types.SimpleNamespace(logits=torch.tensor([0.5, -1.0, 2.0, 0.1], device=dev), query_ids=torch.tensor([0, 0, 2, 2]))
    builder = GA.GAgentBuilder(GA.GAgentSettings(edge_top_k=8), embedding_store=Store())
    builder.process_batch(batch, out)
    assert [s.sample_id for s in builder.samples] == ["a", "c"]
    a, c = builder.samples
    assert a.node_entity_ids.tolist() == [100, 101, 102] and a.edge_head_locals.tolist() == [0, 1]
    assert c.node_entity_ids.tolist() == [105, 106] and c.edge_relations.tolist() == [2, 2]
    assert c.start_node_locals.tolist() == [0] and c.answer_node_locals.tolist() == [1] and c.question_emb.shape == (1, 2)
    builder.process_batch(batch, types.SimpleNamespace(logits=torch.empty(0, device=dev), query_ids=torch.empty(0)))
    assert len(builder.samples) == 2
