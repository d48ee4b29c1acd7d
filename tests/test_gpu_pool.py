"""GPU parity of the single-graph retrieval over a sharded triple pool (BASELINE config 4, miniature): candidate
sets bit-exact against the oracle's composition of the capped BFS and the (dist_h + dist_t, edge id) cut, for 1, 2
and 3 shards on one device; subgraph structure equal up to node numbering; scores within the score tolerance and
top-k edge ids equal to the oracle's top-k of the GPU scores."""
import numpy as np
import pytest
import torch

from oracle import evi_oracle as orc
from evi_rag_b200.kg_pool import PoolRetriever, TriplePoolShard, partition_pool_by_head
from evi_rag_b200.retriever import Retriever
from evi_rag_b200.weights import make_state_dict

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def make_pool(seed: int, num_nodes: int, num_edges: int, num_rel: int):
    rng = np.random.default_rng(seed)
    # power-law endpoints (hubs), self loops and duplicate triples included
    w = 1.0 / np.arange(1, num_nodes + 1) ** 0.7
    w /= w.sum()
    perm = rng.permutation(num_nodes)
    h = perm[rng.choice(num_nodes, num_edges, p=w)]
    t = perm[rng.choice(num_nodes, num_edges, p=w)]
    h[:20] = t[:20]
    r = rng.integers(0, num_rel, num_edges)
    h[100:120], t[100:120], r[100:120] = h[200:220], t[200:220], r[200:220]
    return h.astype(np.int64), r.astype(np.int64), t.astype(np.int64)


def make_shards(h, r, t, num_nodes, world, dev):
    parts = partition_pool_by_head(h, num_nodes, world)
    eid = np.arange(h.size, dtype=np.int64)
    shards = []
    for lo, hi in parts:
        m = (h >= lo) & (h < hi)
        shards.append(TriplePoolShard.from_triples(h[m], r[m], t[m], eid[m], num_nodes, dev))
    assert sum(s.num_triples for s in shards) == h.size
    return shards


def make_queries(seed, num_nodes, Q):
    rng = np.random.default_rng(seed)
    seeds, ptr = [], [0]
    for q in range(Q):
        n = int(rng.integers(1, 3))
        s = rng.integers(0, num_nodes, n).tolist()
        if q == 1:
            s = s + [s[0], -3, num_nodes + 5]  # duplicate and out-of-range seeds
        seeds += s
        ptr.append(len(seeds))
    return np.asarray(seeds, np.int64), np.asarray(ptr, np.int64)


@pytest.mark.parametrize("world", [1, 2, 3])
@pytest.mark.parametrize("cap", [64, 1500])
def test_pool_candidates_match_oracle(dev, world, cap):
    num_nodes, num_edges = 3000, 9000
    h, r, t = make_pool(7, num_nodes, num_edges, 30)
    shards = make_shards(h, r, t, num_nodes, world, dev)
    seeds, sp = make_queries(11, num_nodes, 12)
    emb_ids = torch.arange(num_nodes) * 3 % 1000
    pr = PoolRetriever(shards, model=None, node_embedding_ids=emb_ids, cap=cap, top_k=50, depth=2, table_capacity=1 << 10)
    b = pr.extract(seeds, sp)
    assert pr.table_capacity > 1 << 10  # the visited set outgrew the first table: re-sized and repeated
    ref = orc.pool_subgraph_batch(h, r, t, num_nodes, [seeds[sp[q]:sp[q + 1]].tolist() for q in range(12)], depth=2,
                                  cap=cap, node_embedding_ids=emb_ids.numpy())
    assert np.array_equal(b.edge_ptr.cpu().numpy(), ref["edge_ptr"])
    assert np.array_equal(b.edge_global_id.cpu().numpy(), ref["edge_global_id"])  # same candidates, same order
    assert np.array_equal(b.edge_attr.cpu().numpy(), ref["edge_attr"])
    assert np.array_equal(b.ptr.cpu().numpy(), ref["ptr"])
    # node numbering is arbitrary: compare through the global ids
    ng = b.node_global_ids.cpu().numpy()
    ei = b.edge_index.cpu().numpy()
    assert np.array_equal(ng[ei[0]], h[ref["edge_global_id"]]) and np.array_equal(ng[ei[1]], t[ref["edge_global_id"]])
    for q in range(12):
        n0, n1 = ref["ptr"][q], ref["ptr"][q + 1]
        assert np.array_equal(np.sort(ng[n0:n1]), ref["node_global_ids"][n0:n1])
        assert ei[:, ref["edge_ptr"][q]:ref["edge_ptr"][q + 1]].min(initial=n0) >= n0
        assert ei[:, ref["edge_ptr"][q]:ref["edge_ptr"][q + 1]].max(initial=n0) < max(n1, n0 + 1)
    order = np.argsort(ng + np.repeat(np.arange(12), np.diff(ref["ptr"])) * (num_nodes + 1), kind="stable")
    assert np.array_equal(ng[order], ref["node_global_ids"])
    assert np.array_equal(b.node_embedding_ids.cpu().numpy()[order], ref["node_embedding_ids"])
    assert np.array_equal(b.topic_one_hot.cpu().numpy()[order], ref["topic_one_hot"])
    assert (np.diff(ref["edge_ptr"]) == cap).any() or cap > 1000  # the cap binds on the small-cap case


def test_pool_retrieve_scores_and_topk(dev):
    num_nodes, num_edges, num_rel, Q, k = 2500, 8000, 24, 6, 40
    h, r, t = make_pool(3, num_nodes, num_edges, num_rel)
    rng = np.random.default_rng(0)
    ent = (rng.standard_normal((601, 1024)) * 0.5).astype(np.float32)
    ent[0] = 0
    rel = (rng.standard_normal((num_rel, 1024)) * 0.5).astype(np.float32)
    emb_ids = torch.from_numpy(rng.integers(0, 601, num_nodes))
    sd = make_state_dict(5, trained_like=True)
    model = Retriever(emb_dim=1024, hidden_dim=1024).eval()
    model.load_state_dict(sd, strict=True)
    model.to(dev)
    model.attach_embedding_tables(torch.from_numpy(ent), torch.from_numpy(rel))
    seeds, sp = make_queries(21, num_nodes, Q)
    qemb = torch.from_numpy((rng.standard_normal((Q, 1024)) * 0.5).astype(np.float32))
    results = []
    for world in (1, 2):
        pr = PoolRetriever(make_shards(h, r, t, num_nodes, world, dev), model, node_embedding_ids=emb_ids, cap=600,
                           top_k=k, depth=2, table_capacity=1 << 18)
        results.append(pr.retrieve(qemb, seeds, sp))
    res = results[0]
    assert torch.equal(res["edge_id"], results[1]["edge_id"])  # shard count does not change the answer
    ref_b = orc.pool_subgraph_batch(h, r, t, num_nodes, [seeds[sp[q]:sp[q + 1]].tolist() for q in range(Q)], depth=2,
                                    cap=600, node_embedding_ids=emb_ids.numpy())
    ref = orc.retriever_forward(
        sd, edge_index=torch.from_numpy(ref_b["edge_index"]), ptr=torch.from_numpy(ref_b["ptr"]), question_emb=qemb,
        node_embeddings=torch.from_numpy(ent[ref_b["node_embedding_ids"]]),
        node_embedding_ids=torch.from_numpy(ref_b["node_embedding_ids"]),
        edge_embeddings=torch.from_numpy(rel[ref_b["edge_attr"]]), topic_one_hot=torch.from_numpy(ref_b["topic_one_hot"]))
    got = res["logits_local"].cpu().numpy()
    refl = ref.logits.numpy()
    rel_l2 = np.linalg.norm(got - refl) / np.linalg.norm(refl)
    print(f"[parity] pool retrieval: rel_l2={rel_l2:.3e}")
    assert rel_l2 <= 1.0e-3  # the score contract (tests/test_gpu_parity.py REL_L2)
    _, idx, cnt = orc.segmented_topk(got, ref_b["edge_ptr"], k)
    for q in range(Q):
        kt = int(cnt[q])
        assert int(res["count"][q]) == kt
        want = ref_b["edge_global_id"][ref_b["edge_ptr"][q] + idx[q, :kt]]
        assert np.array_equal(res["edge_id"][q, :kt].cpu().numpy(), want)
        assert (res["edge_id"][q, kt:] == -1).all()


def test_pool_edge_cases(dev):
    """A query without valid seeds yields an empty subgraph; depth 0 keeps only seed-to-seed edges; an isolated seed
    yields no candidates; counts and pointers stay consistent."""
    num_nodes = 50
    h = np.array([0, 1, 2, 3, 3, 7], dtype=np.int64)
    t = np.array([1, 2, 3, 3, 0, 8], dtype=np.int64)
    r = np.arange(6, dtype=np.int64)
    shard = TriplePoolShard.from_triples(h, r, t, np.arange(6), num_nodes, dev)
    seeds = np.array([-1, 99, 0, 1, 40, 3], dtype=np.int64)      # q0: invalid only, q1: {0, 1}, q2: isolated, q3: {3}
    sp = np.array([0, 2, 4, 5, 6], dtype=np.int64)
    for depth in (0, 2):
        pr = PoolRetriever([shard], model=None, cap=16, top_k=4, depth=depth, table_capacity=1 << 10)
        b = pr.extract(seeds, sp)
        ref = orc.pool_subgraph_batch(h, r, t, num_nodes, [seeds[sp[q]:sp[q + 1]].tolist() for q in range(4)], depth=depth,
                                      cap=16)
        assert np.array_equal(b.edge_ptr.cpu().numpy(), ref["edge_ptr"])
        assert np.array_equal(b.edge_global_id.cpu().numpy(), ref["edge_global_id"])
        assert np.array_equal(b.ptr.cpu().numpy(), ref["ptr"])
        assert int(b.edge_ptr[1]) == 0 and int(b.edge_ptr[3]) == int(b.edge_ptr[2])
    assert ref["edge_global_id"].tolist()[:1] == [0]  # depth 2 from {0, 1} starts with edge 0
    with pytest.raises(ValueError):
        PoolRetriever([shard], model=None, cap=0)
    with pytest.raises(ValueError):
        PoolRetriever([shard], model=None, table_capacity=1000)
    with pytest.raises(ValueError):
        TriplePoolShard.from_triples([0], [0], [60], [0], num_nodes, dev)
