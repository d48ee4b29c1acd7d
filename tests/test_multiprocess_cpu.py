"""World-size-2 tests of the multi-GPU host logic on the CPU (gloo): the question partition, the
fixed-shape gather of the selected triples (ranks hold different question counts) and the metric
all-reduce. The data path has no collective — every rank scores its own questions — so what is
exercised here is exactly what `bench.py --gpus N` runs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from evi_rag_b200.sharding import allreduce_sums, gather_topk, partition_by_edges


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_topk(q0: int, q1: int, k: int):
    """Deterministic per-question results so the gathered arrays can be checked against a closed form."""
    q = torch.arange(q0, q1, dtype=torch.int64)
    j = torch.arange(k, dtype=torch.int64)
    index = ((q[:, None] * 7 + j[None, :] * 3) % 1000).to(torch.int32)
    score = (q[:, None].double() * 0.5 - j[None, :].double() * 0.01).float()
    count = torch.clamp(q % (k + 3), max=k).to(torch.int32)
    head = q[:, None] * 1000 + j[None, :]
    return {"index": index, "score": score, "count": count, "head_id": head}


def _worker(rank: int, world: int, port: int, edge_counts, k: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        parts = partition_by_edges(edge_counts, world)
        q0, q1 = parts[rank]
        local = _fake_topk(q0, q1, k)
        got = gather_topk(local, q1 - q0)
        full = _fake_topk(0, len(edge_counts), k)
        known = gather_topk(local, q1 - q0, counts=[b - a for a, b in parts])  # no size exchange
        for name in full:
            assert got[name].dtype == full[name].dtype, name
            assert torch.equal(got[name], full[name]), (rank, name)
            assert torch.equal(known[name], full[name]), (rank, name)
        with pytest.raises(ValueError):
            gather_topk(local, q1 - q0, counts=[1] * (world + 1))
        # metric accumulators: hits / totals summed over ranks (dist_reduce_fx="sum")
        acc = torch.tensor([float(q1 - q0), float(sum(edge_counts[q0:q1]))], dtype=torch.float64)
        tot = allreduce_sums(acc)
        assert tot.tolist() == [float(len(edge_counts)), float(sum(edge_counts))]
        assert acc.tolist() == [float(q1 - q0), float(sum(edge_counts[q0:q1]))]  # input untouched
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_topk_and_metric_sums_over_gloo(tmp_path, world):
    rng = np.random.default_rng(5)
    edge_counts = rng.integers(1, 9000, size=23).tolist()
    edge_counts[4] = 0  # a question without edges still occupies a row
    mp.spawn(_worker, args=(world, _free_port(), edge_counts, 6, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_partition_by_edges_is_contiguous_and_balanced():
    rng = np.random.default_rng(0)
    counts = np.clip(np.rint(np.exp(rng.normal(np.log(5000), 0.4, size=1000))), 500, 20000).astype(int)
    for world in (1, 2, 4, 8):
        parts = partition_by_edges(counts, world)
        assert parts[0][0] == 0 and parts[-1][1] == len(counts)
        assert all(parts[r][1] == parts[r + 1][0] for r in range(world - 1))
        loads = [int(counts[a:b].sum()) for a, b in parts]
        assert max(loads) - min(loads) <= 2 * counts.max(), loads  # within one question of the ideal cut
    # more ranks than questions: trailing ranks get empty shards, nothing is lost
    parts = partition_by_edges([10, 20], 4)
    assert sum(b - a for a, b in parts) == 2 and all(b >= a for a, b in parts)
    with pytest.raises(ValueError):
        partition_by_edges([1, 2, 3], 0)


def _worker_empty_rank(rank: int, world: int, port: int, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        parts = [(0, 2), (2, 2), (2, 5)]  # the middle rank owns no question (config 4: fewer sub-batches than ranks)
        q0, q1 = parts[rank]
        got = gather_topk(_fake_topk(q0, q1, 4), q1 - q0, counts=[b - a for a, b in parts])
        full = _fake_topk(0, 5, 4)
        for name in full:
            assert torch.equal(got[name], full[name]), (rank, name)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_gather_topk_with_an_empty_rank(tmp_path):
    mp.spawn(_worker_empty_rank, args=(3, _free_port(), str(tmp_path)), nprocs=3, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(3))


def test_pool_sub_batch_layout_is_world_independent():
    """Config 4: scoring sub-batches are cut on the whole query list and ranks own whole sub-batches, so a query is
    scored in the same sub-batch whatever the world size (the N-rank scores then equal the 1-rank scores bit for bit)."""
    import types

    from evi_rag_b200.kg_pool import PoolRetriever

    rng = np.random.default_rng(3)
    counts = np.minimum(rng.pareto(1.2, size=500) * 150, 20000).astype(int)  # heavy tail: a few capped queries
    counts[7] = 0
    ep = [0] + np.cumsum(counts).tolist()
    stub = types.SimpleNamespace(score_edges_per_batch=4096)
    cuts = PoolRetriever._sub_batch_cuts(stub, ep, len(counts))
    assert cuts[0] == 0 and cuts[-1] == len(counts) and all(b > a for a, b in zip(cuts[:-1], cuts[1:]))
    for a, b in zip(cuts[:-1], cuts[1:]):  # a sub-batch exceeds the budget only when one query alone does
        assert ep[b] - ep[a] <= 4096 or b == a + 1
    sub_edges = [ep[b] - ep[a] for a, b in zip(cuts[:-1], cuts[1:])]
    for world in (2, 4, 8):
        sparts = partition_by_edges(sub_edges, world)
        owned = [c for s0, s1 in sparts for c in cuts[s0:s1]] + [cuts[-1]]
        assert owned == cuts  # every rank's cuts are the global cuts inside its range
        loads = [sum(sub_edges[s0:s1]) for s0, s1 in sparts]
        assert max(loads) <= sum(loads) / world + max(sub_edges)
