#!/usr/bin/env python
"""Device-resident throughput of the BASELINE configs that bench.py does not time (bench.py times config[1]):
config 3 (CWQ-shaped, 4-hop BFS, ~20k triples per question) and config 5 (mask ablation sweep: DDE on / off,
one top-k at k = 500 whose prefixes are the k in {50, 100, 200, 500} lists, like the writer's triplets_by_k).
One JSON line per case; CUDA events around K steps after W warm-up steps, fresh batch every step.

    python scripts/bench_configs.py [--steps 6 --warmup 2] > gpurun_out/<tag>_configs.jsonl
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from evi_rag_b200.pipeline import RetrievalPipeline, device_batch, host_batch  # noqa: E402
from evi_rag_b200.retriever import Retriever  # noqa: E402
from evi_rag_b200.synthetic import CWQ_50K, WEBQSP_10K, generate_questions, make_embedding_tables  # noqa: E402
from evi_rag_b200.weights import make_state_dict  # noqa: E402


def run_case(name, spec, qb_size, top_k, depth, dde_on, steps, warmup, dev, model):
    pipe = RetrievalPipeline(model, top_k=top_k, bfs_max_depth=depth, writer_scores=True, validate=False)
    batches = []
    for s in range(steps + warmup):
        hb = host_batch(generate_questions(spec, q_start=s * qb_size, q_count=qb_size))
        if not dde_on:  # _forward_with_topic_ablation (retriever_module.py:459-467): zeroed topic_one_hot
            hb.topic_one_hot = torch.zeros_like(hb.topic_one_hot)
        batches.append(device_batch(hb, dev))
    torch.cuda.synchronize()
    for b in batches[:warmup]:
        pipe.run_device(b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for b in batches[warmup:]:
        pipe.run_device(b)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    edges = sum(int(b.edge_index.size(1)) for b in batches[warmup:])
    print(json.dumps({"case": name, "questions_per_step": qb_size, "top_k": top_k, "bfs_depth": depth, "dde": dde_on,
                      "steps": steps, "ms_per_step": ms / steps, "questions_per_s": steps * qb_size / (ms * 1e-3),
                      "edges_per_s": edges / (ms * 1e-3), "edges_per_step": edges / steps}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--sweep", action="store_true",
                    help="config 5 as separate top-k launches at every k in {50, 100, 200, 500}, DDE on and off")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sd = make_state_dict(0, trained_like=True)
    model = Retriever(emb_dim=1024, hidden_dim=1024).eval()
    model.load_state_dict(sd, strict=True)
    model.to(dev)
    ent, rel = make_embedding_tables(WEBQSP_10K)
    model.attach_embedding_tables(torch.from_numpy(ent), torch.from_numpy(rel))
    run_case("config3_cwq_4hop", CWQ_50K, 32, 100, 4, True, args.steps, args.warmup, dev, model)
    for dde_on in (True, False):
        run_case(f"config5_k500_dde_{'on' if dde_on else 'off'}", WEBQSP_10K, 128, 500, 2, dde_on, args.steps, args.warmup,
                 dev, model)
    run_case("config2_k100_reference_point", WEBQSP_10K, 128, 100, 2, True, args.steps, args.warmup, dev, model)
    if args.sweep:
        for k in (50, 100, 200, 500):
            for dde_on in (True, False):
                run_case(f"config5_sweep_k{k}_dde_{'on' if dde_on else 'off'}", WEBQSP_10K, 128, k, 2, dde_on, args.steps,
                         args.warmup, dev, model)


if __name__ == "__main__":
    main()
