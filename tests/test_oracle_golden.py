"""The oracle against the golden vectors produced by the unmodified reference
(oracle/make_golden.py). CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import evi_oracle as orc
from evi_rag_b200.weights import make_state_dict


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


@pytest.mark.parametrize("tag", ["d1024_init", "d1024_trained", "d64_r31", "d64_r04"])
def test_scorer_matches_reference(golden_dir, tag):
    g = _load(golden_dir, f"scorer_{tag}.npz")
    dim, seed, trained, rf, rr = [int(x) for x in g["meta"]]
    sd = make_state_dict(seed, dim, dim, rf, rr, trained_like=bool(trained))
    rel = torch.from_numpy(g["relation_table"])
    for mode in ("bidirectional", "forward", "backward"):
        out = orc.retriever_forward(
            sd,
            edge_index=torch.from_numpy(g["edge_index"]),
            ptr=torch.from_numpy(g["ptr"]),
            question_emb=torch.from_numpy(g["question_emb"]),
            node_embeddings=torch.from_numpy(g["entity_rows"]),
            node_embedding_ids=torch.from_numpy(g["node_embedding_ids"]),
            edge_embeddings=rel[torch.from_numpy(g["edge_attr"])],
            topic_one_hot=torch.from_numpy(g["topic_one_hot"]),
            direction_mode=mode, num_rounds=rf, num_reverse_rounds=rr, want_features=True,
        )
        scale = max(1.0, float(np.abs(g[f"{mode}.logits"]).max()))
        np.testing.assert_allclose(out.logits.numpy(), g[f"{mode}.logits"], atol=2e-5 * scale, rtol=0)
        assert np.array_equal(out.query_ids.numpy(), g[f"{mode}.query_ids"])
        if mode != "backward":
            np.testing.assert_allclose(out.logits_fwd.numpy(), g[f"{mode}.logits_fwd"], atol=2e-5 * scale, rtol=0)
        if mode != "forward":
            np.testing.assert_allclose(out.logits_bwd.numpy(), g[f"{mode}.logits_bwd"], atol=2e-5 * scale, rtol=0)
        fe = out.edge_embeddings.numpy()
        np.testing.assert_allclose(fe[:4], g[f"{mode}.feat_head"], atol=2e-5, rtol=1e-5)
        np.testing.assert_allclose(np.abs(fe).sum(1), g[f"{mode}.feat_rowabs"], rtol=1e-5)
    ns = orc.dde_node_features(
        torch.from_numpy(g["topic_one_hot"]), torch.from_numpy(g["edge_index"]), rf, rr
    )
    np.testing.assert_allclose(ns.numpy(), g["node_struct"], atol=1e-7, rtol=0)


def test_dde_corner_cases(golden_dir):
    g = _load(golden_dir, "dde.npz")
    for rf, rr in [(2, 2), (4, 4), (1, 0), (0, 0)]:
        ns = orc.dde_node_features(torch.from_numpy(g["topic"]), torch.from_numpy(g["edge_index"]), rf, rr)
        np.testing.assert_allclose(ns.numpy(), g[f"ns_{rf}_{rr}"], atol=1e-7, rtol=0)
    for rf, rr in [(2, 2), (1, 3)]:
        ns = orc.dde_node_features(torch.from_numpy(g["topic"]), torch.from_numpy(g["edge_index"]), rf, rr,
                                   reverse_edge_index=torch.from_numpy(g["reverse_edge_index"]))
        np.testing.assert_allclose(ns.numpy(), g[f"ns_rev_{rf}_{rr}"], atol=1e-7, rtol=0)


def test_edge_batch(golden_dir):
    g = _load(golden_dir, "edge_batch.npz")
    eb, ep = orc.compute_edge_batch(g["edge_index"], g["ptr"])
    assert np.array_equal(eb, g["edge_batch"]) and np.array_equal(ep, g["edge_ptr"])
    for tag, frag in (("cross", "crosses graph boundaries"), ("order", "not non-decreasing"), ("range", "out-of-range")):
        assert frag in str(g[f"err_{tag}"])  # what the reference said
        with pytest.raises(ValueError, match=frag):
            orc.compute_edge_batch(g[f"bad_{tag}"], g["ptr"])


def test_bfs_and_shortest_path_labels(golden_dir):
    g = _load(golden_dir, "bfs.npz")
    for ci in range(int(g["num_cases"])):
        n = int(g[f"c{ci}_n"])
        src, dst = g[f"c{ci}_src"], g[f"c{ci}_dst"]
        sources, targets = g[f"c{ci}_sources"].tolist(), g[f"c{ci}_targets"].tolist()
        du = orc.bfs_dist(n, orc.build_undirected_adjacency(n, src, dst), sources)
        dd = orc.bfs_dist(n, orc.build_directed_adjacency(n, src, dst), sources)
        assert np.array_equal(np.array(du, dtype=np.int64), g[f"c{ci}_dist_u"])
        assert np.array_equal(np.array(dd, dtype=np.int64), g[f"c{ci}_dist_d"])
        assert np.array_equal(orc.bfs_dist_np(n, src, dst, sources), g[f"c{ci}_dist_u"])
        assert np.array_equal(orc.bfs_dist_np(n, src, dst, sources, directed=True), g[f"c{ci}_dist_d"])
        for mode in ("undirected", "qa_directed"):
            res = orc.shortest_path_union_mask_by_pair(
                n, src.tolist(), dst.tolist(), sources, targets, directed=(mode == "qa_directed")
            )
            pre = f"c{ci}_{mode}"
            for got, key in zip(res, ("mask", "ps", "pa", "ids", "cnt", "len")):
                assert np.array_equal(np.asarray(got), g[f"{pre}_{key}"]), (ci, mode, key)
    triples = [("a", "r1", "b"), ("b", "r2", "c"), ("d", "r3", "c"), ("e", "r1", "f")]
    cases = [
        (["a"], ["c"], "undirected"), (["a"], ["d"], "undirected"), (["a"], ["d"], "qa_directed"),
        (["a"], ["f"], "undirected"), (["zz"], ["c"], "undirected"), (["c"], ["a"], "qa_directed"),
        (["a", "e"], ["f"], "qa_directed"),
    ]
    got = [orc.has_connectivity(triples, s, a, path_mode=m) for s, a, m in cases]
    assert got == g["has_connectivity"].tolist()


def test_bfs_depth_cap():
    rng = np.random.default_rng(0)
    n, e = 200, 300
    src, dst = rng.integers(0, n, e), rng.integers(0, n, e)
    full = orc.bfs_dist_np(n, src, dst, [0, 5])
    cap = orc.bfs_dist_np(n, src, dst, [0, 5], max_depth=2)
    assert np.array_equal(cap, np.where((full >= 0) & (full <= 2), full, -1))


def test_segmented_topk_matches_torch_topk_and_stable_argsort():
    rng = np.random.default_rng(1)
    edge_ptr = np.array([0, 0, 7, 300, 301, 1500])
    scores = rng.standard_normal(1500).astype(np.float32)
    scores[10:40] = scores[10]  # ties
    vals, idx, cnt = orc.segmented_topk(scores, edge_ptr, 100)
    for b in range(5):
        s = torch.from_numpy(scores[edge_ptr[b]:edge_ptr[b + 1]])
        kt = min(len(s), 100)
        assert cnt[b] == kt
        if kt == 0:
            continue
        tv = torch.topk(s, kt, largest=True, sorted=True).values.numpy()
        np.testing.assert_array_equal(vals[b, :kt], tv)
        order = torch.argsort(s, descending=True, stable=True)[:kt].numpy()  # g_agent_builder.py:651
        np.testing.assert_array_equal(idx[b, :kt], order)
        assert (idx[b, kt:] == -1).all()


GAGENT_SETTINGS = {
    "default": dict(edge_top_k=50),
    "logits_k20": dict(edge_top_k=20, score_mode="logits", start_keep_ratio=0.5, start_min_edges=2, start_max_edges=5),
    "dummy_ok": dict(edge_top_k=50, allow_empty_answer=True, score_temperature=2.0, score_bias=0.5),
    "no_start_edges": dict(edge_top_k=64, start_max_edges=0, start_min_edges=0),
}
GAGENT_INT_FIELDS = ["edge_relations", "edge_head_locals", "edge_tail_locals", "node_entity_ids", "node_embedding_ids",
                     "start_node_locals", "answer_entity_ids", "answer_node_locals"]


def gagent_question(g, i):
    e0, e1 = int(g["edge_ptr"][i]), int(g["edge_ptr"][i + 1])
    n0, n1 = int(g["ptr"][i]), int(g["ptr"][i + 1])
    return dict(
        heads=g["edge_index"][0, e0:e1] - n0, tails=g["edge_index"][1, e0:e1] - n0, relations=g["edge_attr"][e0:e1],
        labels=g["labels"][e0:e1], scores=g["logits"][e0:e1], node_global_ids=g["node_global_ids"][n0:n1],
        node_embedding_ids=g["node_embedding_ids"][n0:n1],
        start_entity_ids=g["start_entity_ids"][int(g["start_ptr"][i]):int(g["start_ptr"][i + 1])],
        answer_entity_ids=g["answer_entity_ids"][int(g["answer_ptr"][i]):int(g["answer_ptr"][i + 1])])


def test_gagent_helpers_match_reference(golden_dir):
    g = _load(golden_dir, "gagent.npz")
    q = gagent_question(g, 2)
    n = len(q["node_global_ids"])
    ns = orc.node_softmax_logit(q["scores"], q["heads"], q["tails"], n)
    np.testing.assert_allclose(ns, g["helper_node_softmax"], atol=2e-5, rtol=1e-5)
    assert np.array_equal(orc.gagent_select_topk(q["scores"], 37), g["helper_topk"])
    st = orc.gagent_select_start_edges(q["heads"], q["tails"], q["scores"], g["helper_starts"], n, 0.25, 1, 40)
    assert np.array_equal(st, g["helper_start_edges"])


@pytest.mark.parametrize("tag", list(GAGENT_SETTINGS))
def test_gagent_samples_match_reference(golden_dir, tag):
    g = _load(golden_dir, "gagent.npz")
    cfg = dict(GAGENT_SETTINGS[tag])
    temp, bias = cfg.pop("score_temperature", 1.0), cfg.pop("score_bias", 0.0)
    failed = 0
    for i in range(len(g["ptr"]) - 1):
        q = gagent_question(g, i)
        if temp != 1.0 or bias != 0.0:  # g_agent_builder.py:164-165
            q["scores"] = (torch.from_numpy(q["scores"]) / float(temp) + float(bias)).numpy()
        out = orc.gagent_build_sample(**q, **cfg)
        assert (out is not None) == bool(g[f"{tag}_kept"][i])
        if out is None:
            failed += 1
            continue
        for f in GAGENT_INT_FIELDS:
            assert np.array_equal(out[f], g[f"{tag}_q{i}_{f}"]), (tag, i, f)
        np.testing.assert_allclose(out["edge_scores"], g[f"{tag}_q{i}_edge_scores"], atol=2e-5, rtol=1e-5)
        assert np.array_equal(out["edge_labels"], g[f"{tag}_q{i}_edge_labels"])
        assert bool(out["is_dummy_agent"]) == bool(g[f"{tag}_q{i}_is_dummy_agent"])
    assert failed == int(g[f"{tag}_retrieval_failed"])


def test_shortest_path_single_matches_reference(golden_dir):
    g = _load(golden_dir, "bfs.npz")
    for c in range(int(g["num_cases"])):
        e, n = orc.shortest_path_single(int(g[f"c{c}_n"]), g[f"c{c}_src"], g[f"c{c}_dst"], g[f"c{c}_sources"].tolist(),
                                        g[f"c{c}_targets"].tolist())
        assert e == g[f"c{c}_single_edges"].tolist() and n == g[f"c{c}_single_nodes"].tolist(), c


def test_writer_payload_matches_reference_artifact(golden_dir):
    g = _load(golden_dir, "writer_inputs.npz")
    art = torch.load(os.path.join(golden_dir, "writer_artifact.pt"), weights_only=False)
    ks = art["payload"]["settings"]["topk_values"]
    assert ks == [1, 5, 20]
    out = orc.writer_payload(g["logits"], g["logits_fwd"], g["logits_bwd"], g["edge_index"], g["edge_attr"], g["labels"],
                             g["node_global_ids"], g["edge_ptr"], max(ks))
    for b, smp in enumerate(art["payload"]["samples"]):
        assert smp["sample_id"] == f"q{b}"
        assert smp["answer_entity_ids"] == g["answer_entity_ids"][g["answer_ptr"][b]:g["answer_ptr"][b + 1]].tolist()
        for k in ks:
            trip = smp["triplets_by_k"][k]
            assert len(trip) == min(k, int(out["count"][b]))
            for j, t in enumerate(trip):
                assert (t["head_entity_id"], t["relation_id"], t["tail_entity_id"], t["rank"]) == (
                    int(out["head_id"][b, j]), int(out["relation_id"][b, j]), int(out["tail_id"][b, j]), j + 1)
                assert abs(t["score"] - float(out["score"][b, j])) <= 1e-7 and t["label"] == float(out["label"][b, j])
                assert t["logit_fwd"] == float(out["logit_fwd"][b, j]) and t["logit_bwd"] == float(out["logit_bwd"][b, j])


LOSS_CASES = {
    "infonce": dict(),
    "temp_bce": dict(infonce_temperature=0.5, infonce_weight=0.7, bce_weight=0.3),
    "weighted": dict(bce_weight=1.0, edge_weight_near=2.0, edge_weight_bridge=0.5),
    "bce_only": dict(infonce_weight=0.0, bce_weight=1.0),
}


@pytest.mark.parametrize("tag", list(LOSS_CASES))
def test_retriever_loss_matches_reference(golden_dir, tag):
    g = _load(golden_dir, "loss.npz")
    total, grad, info = orc.retriever_loss(g["logits"], g["targets"], g["edge_batch"], 6, edge_is_near=g["edge_is_near"],
                                           **LOSS_CASES[tag])
    assert abs(total - float(g[f"{tag}_loss"])) <= 2e-5 * max(1.0, abs(total))
    np.testing.assert_allclose(grad, g[f"{tag}_grad"], atol=2e-6, rtol=1e-4)
    for k in ("infonce", "bce", "infonce_graphs", "pos_prob", "neg_prob"):
        assert abs(info[k] - float(g[f"{tag}_m_{k}"])) <= 2e-5 * max(1.0, abs(info[k])), k


def test_hide_seek_bias_matches_reference(golden_dir):
    """a7: the oracle's restatement of `_compute_hide_seek_bias` draws the same CPU stream as the reference after the
    same seed and lands on the same bias vector, for both ways of naming the near edges."""
    import torch

    g = np.load(os.path.join(golden_dir, "hide_seek.npz"))
    p_near, p_far, bias_near, bias_far = [float(x) for x in g["cfg"]]
    kw = dict(p_near=p_near, p_far=p_far, bias_near=bias_near, bias_far=bias_far)
    ei = torch.from_numpy(g["edge_index"])
    like = torch.zeros(ei.size(1))
    torch.manual_seed(int(g["seed"]))
    got = orc.hide_seek_bias(ei, like, num_nodes=int(g["num_nodes"]), q_local_indices=g["q_local_indices"],
                             a_local_indices=g["a_local_indices"], **kw)
    assert np.array_equal(got.numpy(), g["bias_qa"])
    torch.manual_seed(int(g["seed"]))
    got = orc.hide_seek_bias(ei, like, edge_is_near=g["edge_is_near"], **kw)
    assert np.array_equal(got.numpy(), g["bias_mask"])
    assert set(np.unique(g["bias_qa"]).tolist()) == {bias_near, bias_far, 0.0}
    assert orc.hide_seek_bias(ei, like, edge_is_near=g["edge_is_near"], p_near=0.0, p_far=0.0, bias_near=-1.0,
                              bias_far=-1.0) is None


def _training_case(golden_dir):
    import types

    g = np.load(os.path.join(golden_dir, "training_d64.npz"))
    dim, seed, rng_seed = [int(x) for x in g["meta"]]
    t = torch.from_numpy
    inputs = dict(edge_index=t(g["edge_index"]), ptr=t(g["ptr"]), question_emb=t(g["question_emb"]),
                  node_embeddings=t(g["entity_rows"]), node_embedding_ids=t(g["node_embedding_ids"]),
                  edge_embeddings=t(g["relation_table"][g["edge_attr"]]), topic_one_hot=t(g["topic_one_hot"]))
    hs = dict(p_near=0.6, p_far=0.15, bias_near=-2.5, bias_far=-0.75, num_nodes=int(g["ptr"][-1]),
              q_local_indices=g["q_local_indices"], a_local_indices=g["a_local_indices"])
    return g, dim, seed, rng_seed, inputs, hs


def training_loss(out, n_edges):
    rng = np.random.default_rng(2024)  # oracle/make_golden.py::training_loss
    c = [torch.from_numpy(rng.standard_normal(n_edges).astype(np.float32)).to(out.logits.device) for _ in range(3)]
    return (out.logits * c[0]).sum() + 0.5 * (out.logits_fwd * c[1]).sum() + 0.25 * (out.logits_bwd * c[2]).sum()


@pytest.mark.parametrize("tag", ["p0", "drop"])
def test_training_forward_and_gradients_match_reference(golden_dir, tag):
    """(f) rank 4: the oracle's training-mode forward under autograd against the reference in train() mode on the CPU:
    logits and every parameter's gradient; the dropout + hide-and-seek case pins the generator consumption order."""
    g, dim, seed, rng_seed, inputs, hs = _training_case(golden_dir)
    sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() else v)
          for k, v in make_state_dict(seed, dim, dim, 2, 2, trained_like=True).items()}
    torch.manual_seed(rng_seed)
    out = orc.retriever_train_forward(sd, **inputs, dropout_p=0.0 if tag == "p0" else 0.1,
                                      hide_seek=None if tag == "p0" else hs)
    for name in ("logits", "logits_fwd", "logits_bwd"):
        ref = g[f"{tag}.{name}"]
        assert np.abs(getattr(out, name).detach().numpy() - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max()), name
    training_loss(out, inputs["edge_index"].shape[1]).backward()
    for k, v in sd.items():
        if not v.is_floating_point():
            continue
        ref = g[f"{tag}.grad.{k}"]
        got = v.grad.numpy() if v.grad is not None else np.zeros_like(ref)
        assert np.abs(got - ref).max() <= 2e-4 * max(1e-3, np.abs(ref).max()), (k, np.abs(got - ref).max(), np.abs(ref).max())
