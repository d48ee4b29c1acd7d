"""CPU-only checks: the C-ABI library loads and exports every declared symbol, the host module
mirrors the reference's constructor / state_dict contract, and the product path refuses to run
without CUDA (no fallback)."""
import os
import re
import types

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from evi_rag_b200 import lib

    header = open(os.path.join(ROOT, "include", "evi_rag_b200.h"), encoding="utf-8").read()
    declared = set(re.findall(r"\b(evi_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    L = lib.load()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/evi_rag_b200.h but not exported"
    assert declared == set(lib.PROTOTYPES), declared ^ set(lib.PROTOTYPES)
    assert L.evi_abi_version() == 5
    limit = re.search(r"#define EVI_DDE_STATIC_MAX_NODES (\d+)", header)
    assert limit and L.evi_dde_static_max_nodes() == int(limit.group(1)) == 262144  # the static-scratch entry's stated limit
    assert L.evi_score_workspace_bytes(1000, lib.DIR_BIDIRECTIONAL, 0) >= 1000 * 1024 * (2 + 4 + 4 + 8)


def test_score_args_struct_layout_matches_header():
    from evi_rag_b200 import lib

    header = open(os.path.join(ROOT, "include", "evi_rag_b200.h"), encoding="utf-8").read()
    body = header[header.index("typedef struct evi_score_args {"):header.index("} evi_score_args;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for stmt in body.split("{", 1)[1].split(";"):
        stmt = stmt.strip()
        if not stmt:
            continue
        for part in stmt.split(","):
            names.append(re.findall(r"([A-Za-z_][A-Za-z0-9_]*)\s*$", part.strip())[0])
    assert names == [f[0] for f in lib.ScoreArgs._fields_]


def test_state_dict_contract_matches_reference_keys():
    from evi_rag_b200.retriever import Retriever
    from evi_rag_b200.weights import make_state_dict, state_dict_shapes

    m = Retriever(emb_dim=1024, hidden_dim=1024, dde_cfg={"num_rounds": 2, "num_reverse_rounds": 2})
    sd = m.state_dict()
    expect = dict(state_dict_shapes())
    expect["parity_meta"] = (4,)
    assert {k: tuple(v.shape) for k, v in sd.items()} == expect  # SURVEY §8b key list
    assert sd["parity_meta"].tolist() == [1, 2, 2, 2]
    m.load_state_dict(make_state_dict(1), strict=True)
    assert sum(p.numel() for p in m.parameters()) == 9_474_050  # reference model size


def test_constructor_validation_like_reference():
    from evi_rag_b200.retriever import Retriever

    with pytest.raises(ValueError, match="topic_pe must be enabled"):
        Retriever(8, 8, topic_pe=False)
    with pytest.raises(ValueError, match="num_topics must be 2"):
        Retriever(8, 8, num_topics=3)
    with pytest.raises(ValueError, match="core_mode must be 'geometry'"):
        Retriever(8, 8, core_mode="dot")
    with pytest.raises(ValueError, match="direction_mode must be one of"):
        Retriever(8, 8, direction_mode="both")
    with pytest.raises(ValueError, match="at most 4 rounds"):
        Retriever(8, 8, dde_cfg={"num_rounds": 5})
    with pytest.raises(ValueError, match="must be in \\[0, 1\\]"):
        Retriever(8, 8, hide_seek_cfg={"p_near": 1.5})
    with pytest.raises(ValueError, match="must be <= 0"):
        Retriever(8, 8, hide_seek_cfg={"bias_far": 0.5})
    m = Retriever(8, 8, some_future_kwarg=1)  # extra kwargs are swallowed (retriever.py:116)
    assert m.direction_mode == "bidirectional"


def test_no_cpu_fallback():
    from evi_rag_b200.retriever import Retriever
    from evi_rag_b200.topk import segmented_topk

    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    m = Retriever(1024, 1024).eval()
    b = types.SimpleNamespace(edge_index=torch.tensor([[0], [1]]), ptr=torch.tensor([0, 2]),
                              question_emb=torch.zeros(1, 1024), node_embedding_ids=torch.tensor([1, 2]),
                              edge_attr=torch.tensor([0]), node_embeddings=torch.zeros(2, 1024),
                              edge_embeddings=torch.zeros(1, 1024), topic_one_hot=torch.zeros(2, 2))
    with pytest.raises(RuntimeError, match="CUDA only"):
        m(b)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        segmented_topk(torch.zeros(4), torch.tensor([0, 4]), 2)


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "evi-rag_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), encoding="utf-8").read()
                assert "oracle" not in text.replace("oracle/", "").lower() or f == "weights.py" or \
                    "import oracle" not in text and "from oracle" not in text, f


def test_synthetic_generator_is_deterministic_and_shardable():
    from evi_rag_b200.synthetic import generate_questions, tiny_spec
    from oracle import evi_oracle as orc

    spec = tiny_spec(12, 300, seed=3)
    a = generate_questions(spec)
    b = generate_questions(spec)
    assert np.array_equal(a.edge_index, b.edge_index) and np.array_equal(a.question_emb, b.question_emb)
    eb, ep = orc.compute_edge_batch(a.edge_index, a.ptr)  # valid PyG-style batch
    assert np.array_equal(ep, a.edge_ptr)
    assert (a.edge_index[0] != a.edge_index[1]).all()
    key = (eb * 10**6 + a.edge_index[0]) * 10**6 + a.edge_index[1]
    assert len(np.unique(key * 100 + a.edge_attr)) == a.num_edges  # no duplicate (h, r, t)
    s = a.slice(3, 7)
    assert s.num_graphs == 4 and s.ptr[0] == 0 and s.edge_index.min() >= 0 and s.edge_index.max() < s.num_nodes
    assert set(np.unique(a.topic_one_hot.sum(1))) == {1.0}


def test_writer_and_gagent_host_validation(tmp_path):
    from evi_rag_b200.gagent import GAgentSettings
    from evi_rag_b200.writer import RetrieverTopKEdgeWriter, normalize_k_values

    assert normalize_k_values([5, "1", 20, 5, -3, "x"]) == [1, 5, 20]
    assert normalize_k_values(None, default=(1, 5, 10)) == [1, 5, 10]
    with pytest.raises(ValueError):
        RetrieverTopKEdgeWriter(output_dir=tmp_path, artifact_name=" ")
    with pytest.raises(ValueError):
        RetrieverTopKEdgeWriter(output_dir=tmp_path, schema_version=0)
    w = RetrieverTopKEdgeWriter(output_dir=tmp_path, topk_values=[10, 5])
    assert w.topk_values == [5, 10] and w._max_topk == 10
    (tmp_path / "test.pt").write_bytes(b"x")
    with pytest.raises(FileExistsError):
        RetrieverTopKEdgeWriter(output_dir=tmp_path, overwrite=False).on_predict_start(None, None)
    s = GAgentSettings(edge_top_k=7)
    assert s.start_max_edges == 7 and s.to_metadata()["output_path"].endswith("g_agent_samples.pt")
    for bad in (dict(max_hops=-1), dict(score_temperature=0.0), dict(start_keep_ratio=1.5), dict(score_mode="x")):
        with pytest.raises(ValueError):
            GAgentSettings(**bad)


def test_native_record_builder_equals_python_loop(tmp_path):
    """csrc/records.c (the writer's plain-C record builder) against the Python loop it replaces: equal records for
    ragged counts, optional logit columns and text maps."""
    import numpy as np
    import torch

    from evi_rag_b200 import writer as W

    if W._evi_records is None:
        pytest.skip("_evi_records has not been built (./build.sh)")
    B, K = 37, 24
    rng = np.random.default_rng(0)
    f32 = lambda: torch.from_numpy(rng.standard_normal((B, K)).astype(np.float32))
    counts = rng.integers(0, K + 1, B).astype(np.int32)
    counts[3], counts[5] = 0, K
    host = {"score": f32(), "count": torch.from_numpy(counts), "head_id": torch.from_numpy(rng.integers(0, 1 << 40, (B, K))),
            "tail_id": torch.from_numpy(rng.integers(0, 1 << 40, (B, K))),
            "relation_id": torch.from_numpy(rng.integers(0, 50, (B, K))), "label": f32(), "logit_fwd": f32(),
            "logit_bwd": f32()}

    class Done:
        def synchronize(self):
            pass

    for has_fwd, has_bwd, maps in ((True, True, False), (False, True, True), (False, False, False)):
        item = W._PendingBatch()
        item.host, item.done, item.num_graphs = host, Done(), B
        item.sample_ids, item.questions = [f"s{i}" for i in range(B)], [f"q{i}" for i in range(B)]
        item.answers, item.answer_ptr = list(range(2 * B)), list(range(0, 2 * B + 1, 2))
        item.has_fwd, item.has_bwd = has_fwd, has_bwd
        w = W.RetrieverTopKEdgeWriter(output_dir=tmp_path, topk_values=[3, 1, 50])
        if maps:
            w._relation_map = {i: f"rel{i}" for i in range(0, 50, 3)}
            w._entity_map = {int(host["head_id"][0, 0]): "an entity"}
        native = w._materialise(item)
        w.use_native_records = False
        python = w._materialise(item)
        assert native == python and len(native) == int((counts > 0).sum())
        edge = native[0]["triplets_by_k"][3][0]
        assert list(edge)[:9] == list(W._EDGE_KEYS) and ("logit_fwd" in edge) == has_fwd and ("logit_bwd" in edge) == has_bwd
        assert native[0]["triplets_by_k"][1][0] is not native[0]["triplets_by_k"][3][0]  # every k has its own dicts
