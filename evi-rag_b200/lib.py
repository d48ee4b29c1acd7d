"""ctypes binding of ``libevi_rag_b200.so`` (the C ABI declared in ``include/evi_rag_b200.h``).

There is no CPU fallback: if the shared library is missing every compute call raises
``RuntimeError`` telling the user to build it (``./build.sh`` or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
# EVI_LIB_PATH: another build of the same library (A/B of compile-time choices on one GPU box)
LIB_PATH = os.environ.get("EVI_LIB_PATH") or os.path.join(_HERE, "libevi_rag_b200.so")

EVI_OK = 0
_STATUS = {1: "invalid argument", 2: "CUDA error", 3: "workspace too small", 4: "unsupported shape",
           5: "size exceeds the fixed scratch of a *_static entry point (use the workspace form)"}
EVI_ERR_STATIC_LIMIT = 5

FLAG_OUT_OF_RANGE, FLAG_CROSS_GRAPH, FLAG_NOT_MONOTONE = 1, 2, 4
FMT_F16, FMT_BF16 = 0, 1
DIR_BIDIRECTIONAL, DIR_FORWARD, DIR_BACKWARD = 0, 1, 2
ACT_NONE, ACT_TANH, ACT_SIGMOID = 0, 1, 2
TOPK_RAW, TOPK_SIGMOID = 0, 1
PRECISION_FAST, PRECISION_STRICT = 0, 1

_vp, _i, _ll, _f, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t


class ScoreArgs(C.Structure):
    """``evi_score_args`` (include/evi_rag_b200.h)."""

    _fields_ = [
        ("num_edges", _ll), ("emb_dim", _i), ("hidden_dim", _i), ("direction", _i),
        ("operand_format", _i), ("want_features", _i), ("num_sms", _i),
        ("num_nodes", _i), ("num_rel_rows", _i), ("num_questions", _i), ("num_pair_rows", _i), ("flags", _vp),
        ("head32", _vp), ("tail32", _vp), ("qid32", _vp), ("rel_slot32", _vp), ("pair_slot32", _vp),
        ("node_repr16", _vp), ("rel_repr16", _vp), ("gate_bias32", _vp), ("node_uv32", _vp),
        ("node_p32", _vp), ("pair_rq32", _vp),
        ("struct_ln_g", _vp), ("struct_ln_b", _vp),
        ("nav_w", _vp), ("nav_b", _f),
        ("w1_main16", _vp), ("w1_dist_col", _vp), ("ln1_g", _vp), ("ln1_b", _vp),
        ("head_v", _vp), ("head_c0", _f), ("w2_16", _vp), ("w2_bias", _vp), ("head_w", _vp),
        ("head_b", _f),
        ("logits", _vp), ("logits_fwd", _vp), ("logits_bwd", _vp), ("edge_embeddings", _vp),
        ("workspace", _vp), ("workspace_bytes", _sz),
        ("ev_gemm_begin", _vp), ("ev_gemm_end", _vp),
        ("chunk_edges", _ll),
        ("aux_stream", _vp),
        ("precision", _i), ("node_repr_ld", _i), ("rel_repr_ld", _i), ("w1_ld", _i), ("w1_lo_off", _i),
        ("num_rep_rows", _i), ("head_rep32", _vp), ("tail_rep32", _vp), ("sc_center", _vp), ("w2_ld", _i),
        ("ev_tail_begin", _vp), ("ev_tail_end", _vp),
    ]


class GAgentArgs(C.Structure):
    """``evi_gagent_args`` (include/evi_rag_b200.h)."""

    _fields_ = [(n, _i) for n in ("num_graphs", "cap", "score_mode", "topk_stride", "max_starts", "max_answers")] + [
        (n, _vp) for n in (
            "raw_scores", "labels", "head32", "tail32", "edge_attr", "edge_ptr", "node_ptr", "node_global_ids",
            "node_embedding_ids", "start_entity_ids", "start_ptr", "answer_entity_ids", "answer_ptr", "topk_indices",
            "topk_counts", "mark", "status", "num_triples", "num_nodes_out", "num_start_locals", "num_answers",
            "num_answer_locals", "env_edge_ids", "edge_heads_global", "edge_tails_global", "edge_relations",
            "edge_scores", "edge_labels", "edge_head_locals", "edge_tail_locals", "node_entity_ids",
            "node_embedding_ids_out", "node_batch_index", "start_node_locals", "answer_ids_unique",
            "answer_node_locals")]


class PoolShard(C.Structure):
    """``evi_pool_shard`` (include/evi_rag_b200.h)."""

    _fields_ = [(n, _vp) for n in ("fwd_ptr", "fwd_other", "fwd_rel", "fwd_eid", "rev_ptr", "rev_other")] + [
        ("num_nodes", _i)]


class PoolOverflow(C.Structure):
    """``evi_pool_overflow`` (include/evi_rag_b200.h)."""

    _fields_ = [("items", _vp), ("count", _vp), ("capacity", C.c_ulonglong)]


class PoolTable(C.Structure):
    """``evi_pool_table`` (include/evi_rag_b200.h)."""

    _fields_ = [("keys", _vp), ("local_id", _vp), ("capacity", C.c_ulonglong)]


POOL_TABLE_FULL, POOL_VISITED_FULL, POOL_MAX_CAP = 1, 2, 32768
GAGENT_LOGITS, GAGENT_NODE_SOFTMAX, GAGENT_MAX_CAP = 0, 1, 2048
GAGENT_EMPTY, GAGENT_NO_START_NODE, GAGENT_OVERFLOW, GAGENT_NO_START_LOCAL, GAGENT_NO_ANSWER = 1, 2, 4, 8, 16

# name -> (restype, argtypes); every symbol include/evi_rag_b200.h declares
PROTOTYPES = {
    "evi_abi_version": (_i, []),
    "evi_edge_batch": (_i, [_vp, _ll, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "evi_dde_workspace_bytes": (_sz, [_i]),
    "evi_dde_forward": (_i, [_vp, _vp, _ll, _i, _vp, _i, _i, _i, _vp, _vp, _sz, _vp]),
    "evi_dde_static_max_nodes": (_i, []),
    "evi_dde_forward_static": (_i, [_vp, _vp, _ll, _i, _vp, _i, _i, _i, _vp, _vp]),
    "evi_build_csr": (_i, [_vp, _vp, _ll, _i, _i, _vp, _vp, _vp, _vp]),
    "evi_bfs_multi_source": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "evi_bfs_edge_sweep": (_i, [_vp, _vp, _i] + [_vp] * 7 + [_i, _i, _i, _vp, _vp]),
    "evi_shortest_path_pairs": (_i, [_vp, _vp, _i] + [_vp] * 9 + [_i] + [_vp] * 4 + [_vp]),
    "evi_shortest_path_single": (_i, [_vp, _vp, _i] + [_vp] * 8 + [_i] + [_vp] * 5 + [_i] + [_vp] * 3 + [_vp]),
    "evi_gather_rows": (_i, [_vp, _i, _ll, _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _vp]),
    "evi_gather_rows2": (_i, [_vp, _i, _ll, _i, _vp, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _i, _vp, _vp]),
    "evi_unique_rows": (_i, [_vp, _i, _vp, _ll, _vp, _vp, _vp, _vp, _vp]),
    "evi_edge_rep_index": (_i, [_vp, _vp, _ll, _vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "evi_project_split": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp]),
    "evi_sc_center": (_i, [_vp, _vp, _ll, _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _sz, _vp, _vp, _vp]),
    "evi_pair_context2": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _vp, _i, _vp, _i, _i, _i, _vp, _vp]),
    "evi_score_workspace_bytes2": (_sz, [_ll, _i, _i, _i]),
    "evi_struct_project": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp]),
    "evi_convert_f32": (_i, [_vp, _vp, _ll, _i, _vp]),
    "evi_project": (_i, [_vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "evi_relation_pairs": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "evi_pair_context": (_i, [_vp, _vp, _vp, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _vp]),
    "evi_score_workspace_bytes": (_sz, [_ll, _i, _i]),
    "evi_score_edges": (_i, [C.POINTER(ScoreArgs), _vp]),
    "evi_segmented_topk": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "evi_topk_gather_payload": (_i, [_vp, _vp, _vp, _i, _i, _vp, _ll] + [_vp] * 11 + [_vp]),
    "evi_topk_metrics": (_i, [_vp, _vp, _vp, _i, _i, _vp, _i] + [_vp] * 16 + [_vp]),
    "evi_node_softmax_workspace_bytes": (_sz, [_i]),
    "evi_node_softmax_logit": (_i, [_vp, _vp, _vp, _ll, _i, _i, _f, _f, _vp, _sz, _vp, _vp, _vp]),
    "evi_gagent_start_edges": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _f, _i, _i, _vp, _vp, _vp]),
    "evi_gagent_build": (_i, [C.POINTER(GAgentArgs), _vp]),
    "evi_retriever_loss": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _vp, _f, _f, _f, _f, _f, _vp, _vp, _vp, _vp]),
    "evi_pool_reset": (_i, [C.POINTER(PoolTable), _vp]),
    "evi_pool_insert": (_i, [C.POINTER(PoolTable), _vp, _ll, _i, _vp, _vp, C.c_ulonglong, _vp, _vp]),
    "evi_pool_expand": (_i, [C.POINTER(PoolTable), C.POINTER(PoolShard), _vp, C.c_ulonglong, C.c_ulonglong, _i, _vp,
                             C.c_ulonglong, _vp, _vp, C.c_ulonglong, _vp, C.POINTER(PoolOverflow), _i, _vp]),
    "evi_pool_candidates": (_i, [C.POINTER(PoolTable), C.POINTER(PoolShard), _vp, C.c_ulonglong, _i, _i, _i] +
                            [_vp] * 8 + [C.POINTER(PoolOverflow), _i, _vp]),
    "evi_pool_select": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i] + [_vp] * 8 + [_vp]),
    "evi_pool_relabel": (_i, [C.POINTER(PoolTable), _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "evi_pool_pack": (_i, [C.POINTER(PoolTable)] + [_vp] * 4 + [_i, _i, _i] + [_vp] * 5 + [_ll] + [_vp] * 6 + [_vp]),
}

_lib: Optional[C.CDLL] = None

# kernels launched per C-ABI call (bench.py reports the total as `gpu_launches`)
_KERNELS_PER_CALL = {
    "evi_edge_batch": lambda a: (2 if a[8] else 1) if a[1] > 0 else 1,
    "evi_dde_forward": lambda a: 1 + (2 if a[2] > 0 else 1) * max(a[6], a[7]),
    "evi_dde_forward_static": lambda a: 1 + (2 if a[2] > 0 else 1) * max(a[6], a[7]),
    "evi_build_csr": lambda a: 3 if a[2] > 0 else 1,
    "evi_bfs_multi_source": lambda a: 1,
    "evi_bfs_edge_sweep": lambda a: 1,
    "evi_shortest_path_pairs": lambda a: 1,
    "evi_shortest_path_single": lambda a: 1,
    "evi_gather_rows": lambda a: 1,
    "evi_gather_rows2": lambda a: 1,
    "evi_unique_rows": lambda a: 3 if a[1] > 0 else 0,
    "evi_edge_rep_index": lambda a: 1 if a[2] > 0 else 0,
    "evi_project_split": lambda a: 1,
    "evi_sc_center": lambda a: 3 if a[2] > 0 else 0,
    "evi_pair_context2": lambda a: 1,
    "evi_struct_project": lambda a: 1,
    "evi_convert_f32": lambda a: 1,
    "evi_project": lambda a: 1,
    "evi_relation_pairs": lambda a: 1,
    "evi_pair_context": lambda a: 1,
    "evi_score_edges": lambda a: (6 if a[0]._obj.want_features else 4) * _num_chunks(a[0]._obj),  # same kernels in either mode
    "evi_segmented_topk": lambda a: 1,
    "evi_topk_gather_payload": lambda a: 1,
    "evi_topk_metrics": lambda a: 1,
    "evi_node_softmax_logit": lambda a: 1 if a[5] == 0 else 3,
    "evi_gagent_start_edges": lambda a: 1,
    "evi_gagent_build": lambda a: 1,
    "evi_retriever_loss": lambda a: 3 if a[14] else 2,
    "evi_pool_insert": lambda a: 1 if a[2] > 0 else 0,
    "evi_pool_expand": lambda a: (2 if a[12] else 1) if a[4] > a[3] else 0,
    "evi_pool_candidates": lambda a: ((2 if a[15] else 1) if a[3] > 0 else 0) + (0 if a[4] else 1),
    "evi_pool_select": lambda a: 1,
    "evi_pool_relabel": lambda a: 2,
    "evi_pool_pack": lambda a: 1,
}


def _num_chunks(a) -> int:
    if a.chunk_edges <= 0 or a.chunk_edges >= a.num_edges:
        return 1
    return int((a.num_edges + a.chunk_edges - 1) // a.chunk_edges)


LAUNCHES = {"kernels": 0, "calls": 0}
# EVI_DEBUG_SYNC=1: synchronize around every C-ABI call so a device fault names the call that made it
_DEBUG_SYNC = os.environ.get("EVI_DEBUG_SYNC", "0") == "1"


class _Counted:
    """CDLL facade that counts kernel launches per call."""

    def __init__(self, cdll):
        self._cdll = cdll
        for name in PROTOTYPES:
            fn = getattr(cdll, name)
            counter = _KERNELS_PER_CALL.get(name)
            setattr(self, name, self._wrap(fn, counter) if counter else fn)

    @staticmethod
    def _wrap(fn, counter):
        def call(*args):
            LAUNCHES["kernels"] += counter(args)
            LAUNCHES["calls"] += 1
            if not _DEBUG_SYNC:
                return fn(*args)
            import torch

            torch.cuda.synchronize()  # a fault raised here belongs to work queued before this call
            st = fn(*args)
            try:
                torch.cuda.synchronize()
            except Exception as exc:  # noqa: BLE001
                raise RuntimeError(f"device fault inside {fn.__name__}: {exc}") from exc
            return st

        return call


def load() -> C.CDLL:
    """Load the library once and attach prototypes. Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: the CUDA extension is not built (run ./build.sh or "
            "`python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = _Counted(lib)
    return _lib


def check(status: int, what: str) -> None:
    if status != EVI_OK:
        raise RuntimeError(f"{what} failed: {_STATUS.get(status, 'status ' + str(status))}")


def ptr(t) -> Optional[int]:
    """Device pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


_GUARD_ATTRS = ("head32", "values", "logits", "fwd_ptr")  # GraphBatch, TopK, RetrieverOutput, TriplePoolShard


def _cuda_device_of(obj):
    import torch

    if torch.is_tensor(obj):
        return obj.device if obj.is_cuda else None
    if isinstance(obj, torch.device):
        return obj if obj.type == "cuda" else None
    for name in _GUARD_ATTRS:
        t = getattr(obj, name, None)
        if torch.is_tensor(t) and t.is_cuda:
            return t.device
    return None


def device_guard(fn):
    """Run ``fn`` with the CUDA device of its first device-carrying argument (a CUDA tensor, a ``device=`` keyword,
    or one of this package's tensor holders) made current. The C ABI launches kernels on the calling thread's current
    device and ``stream_handle()`` returns that device's current stream, so a call on tensors of cuda:1 while cuda:0
    is current would otherwise launch on the wrong device (the reference's plain-torch code works from any device)."""
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        import torch

        dev = None
        explicit = kwargs.get("device")
        if explicit is not None:
            d = torch.device(explicit)
            dev = d if d.type == "cuda" else None
        if dev is None:
            for a in list(args) + list(kwargs.values()):
                dev = _cuda_device_of(a)
                if dev is not None:
                    break
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)

    return wrapper


def stream_handle(device=None) -> int:
    """The torch current stream of ``device`` (default: the current device). Callers that take a device also make
    it current (``torch.cuda.device``) around their C-ABI calls: kernels launch on the calling thread's device."""
    import torch

    return torch.cuda.current_stream(device).cuda_stream
