#!/usr/bin/env python
"""Config-4 check and measurement under torch.distributed (one rank per GPU, NCCL):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/pool_dist_check.py [--triples 2000000 --nodes 600000 --queries 64 --cap 20000 --k 500]

Every rank builds its head-range shard of the same seeded pool; the N-rank result (frontier exchange, candidate
merge, query-sharded scoring, gathered top-k) is compared on rank 0 with a single-shard run over the whole pool.
Prints one JSON line with the parity verdict and device timings of the extraction and scoring stages."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from evi_rag_b200.kg_pool import PoolRetriever, TriplePoolShard, partition_pool_by_head  # noqa: E402
from evi_rag_b200.retriever import Retriever  # noqa: E402
from evi_rag_b200.weights import make_state_dict  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--triples", type=int, default=2_000_000)
    ap.add_argument("--nodes", type=int, default=600_000)
    ap.add_argument("--relations", type=int, default=2000)
    ap.add_argument("--entities", type=int, default=50_000)
    ap.add_argument("--queries", type=int, default=64)
    ap.add_argument("--cap", type=int, default=20000)
    ap.add_argument("--k", type=int, default=500)
    ap.add_argument("--table", type=int, default=1 << 22, help="initial visited-table slots (grows x4 on overflow)")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--hub-seeds", action="store_true", help="draw seeds in proportion to degree (popular entities)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rng = np.random.default_rng(4)
    w = 1.0 / np.arange(1, args.nodes + 1) ** 0.6  # hubs: degree ~ rank^-0.6
    w /= w.sum()
    perm = rng.permutation(args.nodes)
    h = perm[rng.choice(args.nodes, args.triples, p=w)].astype(np.int64)
    t = rng.integers(0, args.nodes, args.triples).astype(np.int64)
    r = rng.integers(0, args.relations, args.triples).astype(np.int64)
    eid = np.arange(args.triples, dtype=np.int64)
    emb_ids = torch.from_numpy(rng.integers(0, args.entities + 1, args.nodes))
    ent = (rng.standard_normal((args.entities + 1, 1024)) * 0.5).astype(np.float32)
    ent[0] = 0
    rel = (rng.standard_normal((args.relations, 1024)) * 0.5).astype(np.float32)
    sd = make_state_dict(0, trained_like=True)
    model = Retriever(emb_dim=1024, hidden_dim=1024).eval()
    model.load_state_dict(sd, strict=True)
    model.to(dev)
    model.attach_embedding_tables(torch.from_numpy(ent), torch.from_numpy(rel))
    if args.hub_seeds:
        seeds = h[rng.integers(0, args.triples, args.queries)].astype(np.int64)  # endpoint of a random triple
    else:
        seeds = rng.integers(0, args.nodes, args.queries).astype(np.int64)  # uniform: mostly low-degree entities
    sp = np.arange(args.queries + 1, dtype=np.int64)
    qemb = torch.from_numpy((rng.standard_normal((args.queries, 1024)) * 0.5).astype(np.float32))

    lo, hi = partition_pool_by_head(h, args.nodes, world)[rank]
    m = (h >= lo) & (h < hi)
    shard = TriplePoolShard.from_triples(h[m], r[m], t[m], eid[m], args.nodes, dev)
    pr = PoolRetriever([shard], model, node_embedding_ids=emb_ids, cap=args.cap, top_k=args.k, depth=2,
                       table_capacity=args.table, group=None)
    times = {"extract_ms": [], "score_ms": []}
    for _ in range(args.reps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        b = pr.extract(seeds, sp)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        res = pr.score(b, qemb)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        times["extract_ms"].append((t1 - t0) * 1e3)
        times["score_ms"].append((t2 - t1) * 1e3)
    # adjacency entries the traversal reads on this shard: both rows of every pair expanded (levels 0..depth-1), and
    # the head row of every visited pair twice (count pass + emit pass); each entry costs 4 B of adjacency + an
    # 8 B hash probe (+ 12 B of relation / edge id when emitted)
    stats = pr.traversal_stats()
    out = {"world": world, "triples": args.triples, "nodes": args.nodes, "queries": args.queries, "cap": args.cap, "k": args.k,
           "candidates": int(b.edge_ptr[-1]), "table_slots": pr.table_capacity, "extract_ms": min(times["extract_ms"]), "score_ms": min(times["score_ms"]), **stats}
    entries = stats["expand_entries"] + stats["candidate_entries"]
    out["traversal_entries_per_s"] = entries / (out["extract_ms"] * 1e-3)
    out["traversal_algorithmic_GBps"] = entries * 12 / (out["extract_ms"] * 1e-3) / 1e9
    if world > 1:
        if rank == 0:  # the same queries over the whole pool on one shard
            dist_group_world = pr.world
            full = TriplePoolShard.from_triples(h, r, t, eid, args.nodes, dev)
            single = PoolRetriever([full], model, node_embedding_ids=emb_ids, cap=args.cap, top_k=args.k, depth=2,
                                   table_capacity=args.table)
            single.world, single.rank = 1, 0  # no collective on this check path
            ref = single.retrieve(qemb, seeds, sp)
            out["parity_edge_ids_equal"] = bool(torch.equal(ref["edge_id"], res["edge_id"]))
            out["parity_candidates_equal"] = bool(torch.equal(ref["batch"].edge_global_id, b.edge_global_id))
            out["max_score_diff"] = float((ref["score"] - res["score"]).nan_to_num(posinf=0, neginf=0).abs().max())
            assert dist_group_world == world
        dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
