"""Seeded synthetic retrieval workloads (SURVEY.md §8d restatement of BASELINE.json configs).

Everything is produced on the host with numpy so that the oracle (CPU) and the CUDA path
see bit-identical inputs. Only ids travel to the device; embedding rows live in device-resident
tables (reference layout: ``scripts/text_encode_utils.py:60-67,125-134`` -> fp32
``[max_id+1, 1024]`` with row 0 = zeros for non-text entities).

Shapes follow the reference batch schema read by ``Retriever._prepare_edge_inputs``
(``src/models/components/retriever.py:403-451``): ``edge_index`` holds batch-global node ids,
``ptr`` is the node prefix sum, ``edge_attr`` the relation id per triple.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

EMB_DIM = 1024
_PERM_MULT = 1000003  # prime > any N_q: (r * mult + off) % N_q is a bijection on [0, N_q)


@dataclass
class WorkloadSpec:
    """One BASELINE.json config restated as generator parameters."""

    name: str
    num_questions: int
    median_edges: int = 5000
    sigma: float = 0.4
    min_edges: int = 500
    max_edges: int = 20000
    edges_per_node: float = 3.3
    num_relations: int = 6000
    num_entities: int = 200_000
    non_text_frac: float = 0.35
    unreachable_frac: float = 0.05
    hub_exponent: float = 0.8
    top_k: int = 100
    seed: int = 1234


WEBQSP_CPU = WorkloadSpec("webqsp_1k_cpu", 1000, seed=1234 + 0)
WEBQSP_10K = WorkloadSpec("webqsp_10k", 10_000, seed=1234 + 1)
CWQ_50K = WorkloadSpec(
    "cwq_50k", 50_000, median_edges=20000, min_edges=2000, max_edges=80000, seed=1234 + 2
)


@dataclass
class QuestionBatch:
    """Ragged batch of question subgraphs in the reference's PyG ``Batch`` field names."""

    edge_index: np.ndarray  # [2, E] int64, batch-global node ids
    edge_attr: np.ndarray  # [E] int64 relation ids
    ptr: np.ndarray  # [B+1] int64 node prefix
    edge_ptr: np.ndarray  # [B+1] int64 edge prefix (== _slice_dict["edge_index"])
    node_embedding_ids: np.ndarray  # [N] int64 (0 = non-text)
    node_global_ids: np.ndarray  # [N] int64 entity ids
    topic_one_hot: np.ndarray  # [N, 2] float32 one-hot(is_seed)
    question_emb: np.ndarray  # [B, D] float32
    q_local_indices: np.ndarray  # seeds, batch-global node ids
    q_ptr: np.ndarray  # [B+1]
    a_local_indices: np.ndarray  # answers, batch-global node ids
    a_ptr: np.ndarray  # [B+1]
    extras: Dict[str, np.ndarray] = field(default_factory=dict)

    @property
    def num_graphs(self) -> int:
        return int(self.ptr.shape[0] - 1)

    @property
    def num_nodes(self) -> int:
        return int(self.ptr[-1])

    @property
    def num_edges(self) -> int:
        return int(self.edge_index.shape[1])

    def slice(self, q0: int, q1: int) -> "QuestionBatch":
        """Questions [q0, q1) re-based to their own batch-global ids."""
        n0, n1 = int(self.ptr[q0]), int(self.ptr[q1])
        e0, e1 = int(self.edge_ptr[q0]), int(self.edge_ptr[q1])
        s0, s1 = int(self.q_ptr[q0]), int(self.q_ptr[q1])
        a0, a1 = int(self.a_ptr[q0]), int(self.a_ptr[q1])
        return QuestionBatch(
            edge_index=self.edge_index[:, e0:e1] - n0,
            edge_attr=self.edge_attr[e0:e1],
            ptr=self.ptr[q0 : q1 + 1] - n0,
            edge_ptr=self.edge_ptr[q0 : q1 + 1] - e0,
            node_embedding_ids=self.node_embedding_ids[n0:n1],
            node_global_ids=self.node_global_ids[n0:n1],
            topic_one_hot=self.topic_one_hot[n0:n1],
            question_emb=self.question_emb[q0:q1],
            q_local_indices=self.q_local_indices[s0:s1] - n0,
            q_ptr=self.q_ptr[q0 : q1 + 1] - s0,
            a_local_indices=self.a_local_indices[a0:a1] - n0,
            a_ptr=self.a_ptr[q0 : q1 + 1] - a0,
        )


def make_embedding_tables(
    spec: WorkloadSpec, *, dim: int = EMB_DIM, dtype=np.float32
) -> tuple[np.ndarray, np.ndarray]:
    """Entity table ``[T+1, D]`` (row 0 zeros) and relation table ``[R, D]``, N(0, 0.5^2)."""
    rng = np.random.default_rng(spec.seed ^ 0x5EED7AB1)
    ent = rng.standard_normal((spec.num_entities + 1, dim), dtype=np.float32) * np.float32(0.5)
    ent[0] = 0.0
    rel = rng.standard_normal((spec.num_relations, dim), dtype=np.float32) * np.float32(0.5)
    return ent.astype(dtype, copy=False), rel.astype(dtype, copy=False)


def _zipf_relation_ids(rng: np.random.Generator, n: int, num_relations: int, a: float = 1.1) -> np.ndarray:
    ranks = np.arange(1, num_relations + 1, dtype=np.float64)
    w = ranks ** (-a)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    u = rng.random(n)
    return np.searchsorted(cdf, u, side="left").astype(np.int64)


def generate_questions(
    spec: WorkloadSpec,
    *,
    q_start: int = 0,
    q_count: Optional[int] = None,
    dim: int = EMB_DIM,
) -> QuestionBatch:
    """Generate questions ``[q_start, q_start+q_count)`` of ``spec`` (each question's stream is
    keyed by its global index, so any shard of the workload is reproducible on its own)."""
    if q_count is None:
        q_count = spec.num_questions - q_start
    q_count = int(q_count)
    # Per-question sizes come from one stream over the whole workload so shards agree.
    size_rng = np.random.default_rng(spec.seed)
    raw = np.exp(size_rng.normal(np.log(spec.median_edges), spec.sigma, size=spec.num_questions))
    e_target_all = np.clip(np.rint(raw), spec.min_edges, spec.max_edges).astype(np.int64)
    n_seeds_all = size_rng.integers(1, 3, size=spec.num_questions)
    n_answers_all = size_rng.integers(1, 4, size=spec.num_questions)
    unreachable_all = size_rng.random(spec.num_questions) < spec.unreachable_frac

    e_target = e_target_all[q_start : q_start + q_count]
    n_seeds = n_seeds_all[q_start : q_start + q_count]
    n_answers = n_answers_all[q_start : q_start + q_count]
    unreachable = unreachable_all[q_start : q_start + q_count]
    n_core = np.maximum(16, (e_target / spec.edges_per_node).astype(np.int64))
    # unreachable questions get one extra isolated node that serves as the answer
    n_q = n_core + unreachable.astype(np.int64)

    rng = np.random.default_rng([spec.seed, q_start, q_count])
    total_e = int(e_target.sum())
    qid = np.repeat(np.arange(q_count, dtype=np.int64), e_target)
    n_core_e = n_core[qid]
    expo = 1.0 / (1.0 - spec.hub_exponent)
    # rank r with P(r) ~ r^-hub_exponent  <=>  r = floor(N * U^(1/(1-a)))
    rank_h = np.minimum((n_core_e * rng.random(total_e) ** expo).astype(np.int64), n_core_e - 1)
    rank_t = np.minimum((n_core_e * rng.random(total_e) ** expo).astype(np.int64), n_core_e - 1)
    off = rng.integers(0, 1 << 30, size=q_count, dtype=np.int64)
    off_e = off[qid]
    h = (rank_h * _PERM_MULT + off_e) % n_core_e
    t = (rank_t * _PERM_MULT + off_e) % n_core_e
    r = _zipf_relation_ids(rng, total_e, spec.num_relations)
    keep = h != t  # build_graph removes self loops (build_retrieval_pipeline.py:1484-1489)
    qid, h, t, r = qid[keep], h[keep], t[keep], r[keep]
    n_max = int(n_core.max()) + 2
    key = ((qid * n_max + h) * n_max + t) * spec.num_relations + r
    _, first = np.unique(key, return_index=True)  # drop duplicate (h, r, t)
    first.sort()
    qid, h, t, r = qid[first], h[first], t[first], r[first]

    e_cnt = np.bincount(qid, minlength=q_count).astype(np.int64)
    edge_ptr = np.zeros(q_count + 1, dtype=np.int64)
    np.cumsum(e_cnt, out=edge_ptr[1:])
    ptr = np.zeros(q_count + 1, dtype=np.int64)
    np.cumsum(n_q, out=ptr[1:])
    node_base_e = ptr[:-1][qid]
    edge_index = np.stack([h + node_base_e, t + node_base_e]).astype(np.int64)

    total_n = int(ptr[-1])
    node_q = np.repeat(np.arange(q_count, dtype=np.int64), n_q)
    node_local = np.arange(total_n, dtype=np.int64) - ptr[:-1][node_q]
    # seeds = hub ranks 0..n_seeds-1, answers = ranks 2..2+n_answers-1 (or the isolated node)
    q_ptr = np.zeros(q_count + 1, dtype=np.int64)
    np.cumsum(n_seeds, out=q_ptr[1:])
    sq = np.repeat(np.arange(q_count, dtype=np.int64), n_seeds)
    s_rank = np.arange(int(q_ptr[-1]), dtype=np.int64) - q_ptr[:-1][sq]
    q_local = (s_rank * _PERM_MULT + off[sq]) % n_core[sq] + ptr[:-1][sq]
    n_ans_eff = np.where(unreachable, 1, n_answers).astype(np.int64)
    a_ptr = np.zeros(q_count + 1, dtype=np.int64)
    np.cumsum(n_ans_eff, out=a_ptr[1:])
    aq = np.repeat(np.arange(q_count, dtype=np.int64), n_ans_eff)
    a_rank = np.arange(int(a_ptr[-1]), dtype=np.int64) - a_ptr[:-1][aq] + 2
    a_local = np.where(
        unreachable[aq],
        n_core[aq],  # the isolated extra node
        (a_rank * _PERM_MULT + off[aq]) % n_core[aq],
    ) + ptr[:-1][aq]

    is_seed = np.zeros(total_n, dtype=bool)
    is_seed[q_local] = True
    topic = np.zeros((total_n, 2), dtype=np.float32)
    topic[np.arange(total_n), is_seed.astype(np.int64)] = 1.0

    node_global = rng.integers(1, spec.num_entities + 1, size=total_n, dtype=np.int64)
    non_text = rng.random(total_n) < spec.non_text_frac
    node_emb_ids = np.where(non_text, 0, node_global).astype(np.int64)
    q_emb_rng = np.random.default_rng([spec.seed, 0xC0FFEE, q_start, q_count])
    question_emb = q_emb_rng.standard_normal((q_count, dim), dtype=np.float32) * np.float32(0.5)
    del node_local
    return QuestionBatch(
        edge_index=edge_index,
        edge_attr=r.astype(np.int64),
        ptr=ptr,
        edge_ptr=edge_ptr,
        node_embedding_ids=node_emb_ids,
        node_global_ids=node_global,
        topic_one_hot=topic,
        question_emb=question_emb,
        q_local_indices=q_local.astype(np.int64),
        q_ptr=q_ptr,
        a_local_indices=a_local.astype(np.int64),
        a_ptr=a_ptr,
    )


def tiny_spec(num_questions: int = 4, median_edges: int = 300, seed: int = 7, **kw) -> WorkloadSpec:
    """Small workload for tests / smoke: same generator, reduced sizes."""
    base = dict(
        name=f"tiny_{num_questions}x{median_edges}",
        num_questions=num_questions,
        median_edges=median_edges,
        min_edges=max(8, median_edges // 8),
        max_edges=median_edges * 4,
        num_relations=64,
        num_entities=512,
        seed=seed,
    )
    base.update(kw)
    return WorkloadSpec(**base)
