"""evi_dde_forward_static: the DDE propagation with its scratch in a file-scope device array (no workspace argument,
no allocation inside the call), checked through the C ABI against evi_dde_forward on the same inputs.

Collected last on purpose: the entry point has no counterpart in the reference
(include/evi_rag_b200.h). Passes alone in a fresh process and after the
other GPU tests."""
import numpy as np
import pytest
import torch

from evi_rag_b200.synthetic import generate_questions, tiny_spec

pytestmark = pytest.mark.gpu


def _inputs(dev, num_questions, median_edges, seed):
    qb = generate_questions(tiny_spec(num_questions, median_edges, seed=seed))
    head = torch.from_numpy(qb.edge_index[0].astype(np.int32)).to(dev)
    tail = torch.from_numpy(qb.edge_index[1].astype(np.int32)).to(dev)
    topic = torch.from_numpy(qb.topic_one_hot).to(device=dev, dtype=torch.float32).contiguous()
    return head, tail, topic, int(qb.num_nodes)


def _workspace_form(L, lib, head, tail, topic, n, rf, rr):
    out = torch.empty((n, 2 * (1 + rf + rr)), dtype=torch.float32, device=head.device)
    ws = torch.empty(int(L.evi_dde_workspace_bytes(n)), dtype=torch.uint8, device=head.device)
    lib.check(L.evi_dde_forward(head.data_ptr(), tail.data_ptr(), head.numel(), n, topic.data_ptr(), topic.size(1),
                                rf, rr, out.data_ptr(), ws.data_ptr(), ws.numel(), lib.stream_handle()),
              "evi_dde_forward")
    return out


def test_dde_forward_static_matches_workspace_form_twice():
    from evi_rag_b200 import lib

    L = lib.load()
    dev = torch.device("cuda", 0)
    limit = int(L.evi_dde_static_max_nodes())
    assert limit == 262144
    with torch.cuda.device(dev):
        # two calls in a row on DIFFERENT graphs, the larger first: the second call finds the first one's accumulators,
        # degrees and per-round states in the shared array and must not see them
        cases = [_inputs(dev, 12, 1500, seed=91) + (2, 2), _inputs(dev, 5, 400, seed=92) + (3, 1)]
        expected = [_workspace_form(L, lib, *c) for c in cases]
        outs = [torch.full((c[3], 2 * (1 + c[4] + c[5])), float("nan"), dtype=torch.float32, device=dev) for c in cases]
        torch.cuda.synchronize(dev)
        for i, ((head, tail, topic, n, rf, rr), out) in enumerate(zip(cases, outs)):
            allocated = torch.cuda.memory_allocated(dev)
            free_before = torch.cuda.mem_get_info(dev)[0]
            st = L.evi_dde_forward_static(head.data_ptr(), tail.data_ptr(), head.numel(), n, topic.data_ptr(),
                                          topic.size(1), rf, rr, out.data_ptr(), lib.stream_handle())
            lib.check(st, "evi_dde_forward_static")
            torch.cuda.synchronize(dev)
            assert torch.cuda.memory_allocated(dev) == allocated  # nothing allocated through torch
            if i == 1:  # the first call may load the kernels' module; the second allocates nothing anywhere
                assert torch.cuda.mem_get_info(dev)[0] >= free_before
        for out, ref in zip(outs, expected):
            # integer fixed-point accumulation: the two entry points give the same bits (the parity tests' DDE rule)
            assert torch.equal(out, ref)
            assert float((out - ref).abs().max()) <= 1e-6
        # and once more on the first graph: stale contents of the smaller second call are cleared as well
        head, tail, topic, n, rf, rr = cases[0]
        again = torch.empty_like(outs[0])
        lib.check(L.evi_dde_forward_static(head.data_ptr(), tail.data_ptr(), head.numel(), n, topic.data_ptr(),
                                           topic.size(1), rf, rr, again.data_ptr(), lib.stream_handle()),
                  "evi_dde_forward_static")
        torch.cuda.synchronize(dev)
        assert torch.equal(again, expected[0])


def test_dde_forward_static_rejects_sizes_beyond_its_scratch():
    from evi_rag_b200 import lib

    L = lib.load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        head, tail, topic, n, = _inputs(dev, 3, 200, seed=93)
        out = torch.zeros((n, 10), dtype=torch.float32, device=dev)
        limit = int(L.evi_dde_static_max_nodes())
        # the size check comes first: nothing is launched, so the short buffers are never touched
        st = L.evi_dde_forward_static(head.data_ptr(), tail.data_ptr(), head.numel(), limit + 1, topic.data_ptr(),
                                      topic.size(1), 2, 2, out.data_ptr(), lib.stream_handle())
        assert st == lib.EVI_ERR_STATIC_LIMIT
        with pytest.raises(RuntimeError, match="fixed scratch"):
            lib.check(st, "evi_dde_forward_static")
        torch.cuda.synchronize(dev)
        assert float(out.abs().max()) == 0.0
