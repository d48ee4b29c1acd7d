#!/usr/bin/env python
"""bench.py — questions/sec of the retrieval hot path (triple scoring + BFS/DDE + top-k).

    python bench.py --gpus 1 --steps 20 --warmup 3                 # this repo's CUDA path
    python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 # the reference's CPU algorithm

One "step" = one pass of the hot path over one batch of synthetic questions of BASELINE.json
config[1] ("WebQSP-shaped 10k questions, 2-hop BFS from seed entities + triple scoring, top-100"):
edge segmentation -> 2-hop multi-source BFS -> DDE -> projections -> tcgen05 scoring -> ragged
top-100 + writer gather. Every step takes a fresh batch (per-step working set of several GB, far
beyond the 126 MB L2). Timed with CUDA events on the launching stream, barrier + synchronize on
both sides, max over ranks. Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from evi_rag_b200.synthetic import WEBQSP_10K, generate_questions, make_embedding_tables  # noqa: E402
from evi_rag_b200.weights import make_state_dict  # noqa: E402

F_SC_PER_EDGE_DIR = 2 * 1024 * 1024              # W_sc sc: the per-edge, per-direction block of state_net.0
F_EDGE = 16_871_424                              # SURVEY §8d, bidirectional scoring per edge
F_NODE = 2_097_152
TOP_K = 100
BFS_DEPTH = 2
# epilogue of the W_sc sc launch (csrc/scorer_kernels.cu, ChunkPlan): 16-bit block fixed-point rows unless an
# experiment knob keeps the fp32 rows
W_SC_EPILOGUE = "EpiF32Tma" if any(os.environ.get(k) for k in ("EVI_S_F32", "EVI_MM_F32", "EVI_GEMM_1SM")) else "EpiQ16Tma"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--questions-per-step", type=int, default=128)
    ap.add_argument("--ref-questions-per-step", type=int, default=4)
    ap.add_argument("--cpu-sample-questions", type=int, default=160,
                    help="questions of the cpu_baseline sample (batches of 32 like configs/data/retriever.yaml:8; ~15 s)")
    ap.add_argument("--operand-dtype", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pipeline-chunks", type=int, default=4,
                    help="edge chunks per batch (bounds the scoring workspace)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle restatement of the reference algorithm on the host cores
# ------------------------------------------------------------------------------------------------
CPU_STAGE_SECONDS = {"gather+scoring": 0.0, "topk+payload": 0.0, "adjacency+bfs": 0.0}


def cpu_reference_step(qb, sd, ent, rel, k=TOP_K):
    """One pass of the reference's path on the CPU for batch ``qb``: CPU index_select of the
    embedding rows (embedding_store.py:101-150), Retriever.forward in fp32 (same op order as the
    reference, per-edge projections included), writer-style sigmoid + per-graph top-k + gathers,
    and the per-question adjacency build + multi-source BFS of the dataset builder. Stage times
    accumulate in CPU_STAGE_SECONDS (CPU baseline timing plan steps 3-5)."""
    from oracle import evi_oracle as orc

    t = torch.from_numpy
    t0 = time.perf_counter()
    node_rows = ent.index_select(0, t(qb.node_embedding_ids))
    edge_rows = rel.index_select(0, t(qb.edge_attr))
    out = orc.retriever_forward(sd, edge_index=t(qb.edge_index), ptr=t(qb.ptr), question_emb=t(qb.question_emb),
                                node_embeddings=node_rows, node_embedding_ids=t(qb.node_embedding_ids),
                                edge_embeddings=edge_rows, topic_one_hot=t(qb.topic_one_hot))
    t1 = time.perf_counter()
    payload = orc.writer_payload(out.logits.numpy(), out.logits_fwd.numpy(), out.logits_bwd.numpy(), qb.edge_index,
                                 qb.edge_attr, np.zeros(qb.num_edges, np.float32), qb.node_global_ids, qb.edge_ptr, k)
    t2 = time.perf_counter()
    for b in range(qb.num_graphs):
        n0, n1, e0, e1 = qb.ptr[b], qb.ptr[b + 1], qb.edge_ptr[b], qb.edge_ptr[b + 1]
        adj = orc.build_undirected_adjacency(int(n1 - n0), qb.edge_index[0, e0:e1] - n0, qb.edge_index[1, e0:e1] - n0)
        orc.bfs_dist(int(n1 - n0), adj, qb.q_local_indices[qb.q_ptr[b]:qb.q_ptr[b + 1]] - n0)
    t3 = time.perf_counter()
    CPU_STAGE_SECONDS["gather+scoring"] += t1 - t0
    CPU_STAGE_SECONDS["topk+payload"] += t2 - t1
    CPU_STAGE_SECONDS["adjacency+bfs"] += t3 - t2
    return payload


def time_cpu(spec, sd, ent, rel, questions_per_step, steps, warmup):
    torch.set_num_threads(os.cpu_count() or 1)
    ent_t, rel_t = torch.from_numpy(ent), torch.from_numpy(rel)
    batches = [generate_questions(spec, q_start=(i * questions_per_step) % spec.num_questions,
                                  q_count=questions_per_step) for i in range(steps + warmup)]
    for qb in batches[:warmup]:
        cpu_reference_step(qb, sd, ent_t, rel_t)
    for key in CPU_STAGE_SECONDS:
        CPU_STAGE_SECONDS[key] = 0.0
    t0 = time.perf_counter()
    for qb in batches[warmup:]:
        cpu_reference_step(qb, sd, ent_t, rel_t)
    dt = time.perf_counter() - t0
    return steps * questions_per_step / dt, dt


def cpu_stage_qps(num_questions):
    """questions/s of each CPU stage alone over the timed sample."""
    return {key: (num_questions / sec if sec > 0 else None) for key, sec in CPU_STAGE_SECONDS.items()}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    spec = WEBQSP_10K
    sd = make_state_dict(0, trained_like=True)
    ent, rel = make_embedding_tables(spec)
    qps, dt = time_cpu(spec, sd, ent, rel, args.ref_questions_per_step, args.steps, max(1, args.warmup))
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": "questions/sec (triple scoring + top-k)", "value": qps,
        "unit": "questions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.ref_questions_per_step), "top_k": TOP_K, "bfs_depth": BFS_DEPTH},
        "cpu_baseline": {"value": qps, "unit": "questions/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {args.ref_questions_per_step} questions of the same workload; "
                                   "oracle port of the reference (pure-Python reference cannot travel), torch fp32, "
                                   f"{cores} threads",
                         "stage_questions_per_s": cpu_stage_qps(args.steps * args.ref_questions_per_step)},
        "e2e": {"value": qps, "unit": "questions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_name(qps):
    return (f"webqsp_10k (BASELINE config[1]): 10k WebQSP-shaped questions, E_q~LogNormal(ln5000,0.4), N_q=E_q/3.3, "
            f"1024-d, 2-hop BFS + DDE(2+2) + bidirectional scoring + top-100; {qps} questions per step")


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(suffix=".csv")
        self.idx = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "25", "-i", str(self.idx)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for ln in open(self.path):
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1415.7), d.get("hbm_gbs", 6553.3), "MEASURED_PEAKS.json (sustained cuBLAS bf16)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md: ~1.4 PF sustained)"


def roofline_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the roofline kernel, from the committed
    `ncu --set full` capture (profiles/roofline_traffic.json, written by scripts/summarize_ncu.py); null if absent."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if not os.path.exists(p):
        return None
    try:
        return float(json.load(open(p))["dram_bytes_per_launch"])
    except Exception:
        return None


def run_ours(args):
    import torch.distributed as dist

    from evi_rag_b200 import lib as evilib
    from evi_rag_b200.pipeline import RetrievalPipeline, device_batch, h2d_bytes, host_batch
    from evi_rag_b200.retriever import Retriever
    from evi_rag_b200.sharding import gather_topk

    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    evilib.load()  # fails loudly if the extension is missing

    spec = WEBQSP_10K
    QB = args.questions_per_step
    sd = make_state_dict(0, trained_like=True)
    model = Retriever(emb_dim=1024, hidden_dim=1024, operand_dtype=args.operand_dtype,
                      pipeline_chunks=args.pipeline_chunks).eval()
    model.load_state_dict(sd, strict=True)
    model.to(dev)
    ent, rel = make_embedding_tables(spec)
    model.attach_embedding_tables(torch.from_numpy(ent), torch.from_numpy(rel))
    pipe = RetrievalPipeline(model, top_k=TOP_K, bfs_max_depth=BFS_DEPTH, writer_scores=True, validate=False)

    n_steps = args.warmup + args.steps
    n_global_batches = (spec.num_questions + QB - 1) // QB
    host_batches = []
    for s in range(n_steps):
        gb_idx = (s * world + rank) % n_global_batches  # rank r walks its own batches: weak scaling
        q0 = gb_idx * QB
        host_batches.append(host_batch(generate_questions(spec, q_start=q0, q_count=min(QB, spec.num_questions - q0))))
    edges_timed = sum(int(hb.edge_index.size(1)) for hb in host_batches[args.warmup:])
    nodes_timed = sum(int(hb.ptr_host[-1]) for hb in host_batches[args.warmup:])
    questions_timed = sum(hb.num_graphs for hb in host_batches[args.warmup:])

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_counts(s):
        """Questions every rank holds at step s (the schedule above is deterministic, so no size exchange)."""
        out = []
        for r in range(world):
            q0 = ((s * world + r) % n_global_batches) * QB
            out.append(min(QB, spec.num_questions - q0))
        return out

    def exchange(res, nq, s):
        if world > 1:  # gather the selected triples on every rank: one packed all-gather over NCCL, no host sync
            gather_topk({k: res[k] for k in ("score", "index", "count", "head_id", "tail_id", "relation_id")}, nq,
                        counts=step_counts(s))

    sampler = ClockSampler(local_rank)
    # ---------------- device-resident: inputs already in HBM --------------------------------
    dev_batches = [device_batch(hb, dev) for hb in host_batches]
    torch.cuda.synchronize(dev)
    gemm_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for a, b in gemm_ev:  # torch creates the cudaEvent_t lazily; record once so .cuda_event is a live handle
        a.record(); b.record()
    for i in range(args.warmup):
        exchange(pipe.run_device(dev_batches[i]), dev_batches[i].num_graphs, i)
    sync_all()
    if rank == 0:
        sampler.start()
    evilib.LAUNCHES["kernels"] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    chunk_edges, chunk_counts = [], []
    for i in range(args.steps):
        model.gemm_events = gemm_ev[i]
        b = dev_batches[args.warmup + i]
        res = pipe.run_device(b)
        exchange(res, b.num_graphs, args.warmup + i)
        chunk_edges.append(model.last_chunk_edges)
        chunk_counts.append(model.last_num_chunks)
    e1.record()
    sync_all()
    model.gemm_events = None
    launches = evilib.LAUNCHES["kernels"]
    flags_word = int(res["flags"].item())
    ms_dev = e0.elapsed_time(e1)
    gemm_ms = [a.elapsed_time(b) for a, b in gemm_ev]
    del dev_batches
    # ---------------- end to end: host buffers in, host results out -------------------------
    for r in pipe.run_stream(host_batches[:args.warmup], dev):  # also sizes the rotating host / device buffers
        pass
    sync_all()
    d2h = RetrievalPipeline.d2h_bytes(r)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.perf_counter()
    e2.record()
    # the public streaming call: H2D of step i+1 overlaps the kernels of step i; every copy is inside the region
    timed = host_batches[args.warmup:]
    nq_iter = iter([(hb.num_graphs, args.warmup + i) for i, hb in enumerate(timed)])
    post = (lambda res: exchange(res, *next(nq_iter))) if world > 1 else None
    for r in pipe.run_stream(timed, dev, post_device=post):
        pass
    e3.record()
    sync_all()
    wall_e2e = time.perf_counter() - t_wall0
    ms_e2e = max(e2.elapsed_time(e3), 0.0)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms_dev, ms_e2e, wall_e2e * 1e3], dtype=torch.float64, device=dev)
    cnt = torch.tensor([questions_timed, edges_timed, nodes_timed, launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_dev, ms_e2e, wall_ms = [float(x) for x in t.tolist()]
    q_all, e_all, n_all, launches_all = [float(x) for x in cnt.tolist()]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak_tf, peak_hbm, peak_src = measured_peaks()
    gemm_flops = [2 * ce * F_SC_PER_EDGE_DIR for ce in chunk_edges]  # both directions in one launch
    achieved_tf = float(np.mean([f / (ms * 1e-3) / 1e12 for f, ms in zip(gemm_flops, gemm_ms)]))
    step_flops = e_all * F_EDGE + n_all * F_NODE
    line = {
        "metric": "questions/sec (triple scoring + top-k)",
        "value": q_all / (ms_dev * 1e-3),
        "unit": "questions/s",
        "n_gpus": world,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "f16 operands / f32 accumulate" if args.operand_dtype == "fp16" else "bf16 operands / f32 accumulate",
        "data": "synthetic",
        "config": {
            "workload": workload_name(QB), "top_k": TOP_K, "bfs_depth": BFS_DEPTH, "parallelism": f"questions sharded x{world}",
            "l2": "fresh batch every step; per-step working set (operand rows + activations) is several GB >> 126 MB L2",
            "edges_per_step": e_all / args.steps / world, "validation_flags": flags_word,
            "intermediates": ("W_iv m rows fp16, W_sc sc rows 16-bit block fixed point (one fp32 scale per row and 256 columns), "
                              "everything else fp32; scores within 1e-3 of the fp32 reference (tests/test_gpu_parity.py)"
                              if W_SC_EPILOGUE == "EpiQ16Tma" else "fp32 partial products (experiment knob set)"),
        },
        "e2e": {"value": q_all / (ms_e2e * 1e-3), "unit": "questions/s",
                "h2d_bytes_per_step": h2d_bytes(host_batches[args.warmup]), "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": wall_ms / args.steps},
        "gpu_launches": int(launches_all),
        "clocks": clocks,
        "roofline": {
            "bound": "tensor", "kernel": f"gemm2::gemm2_kernel<{W_SC_EPILOGUE}> (W_sc block of state_net.0: "
                                         "[2E_chunk,1024]x[1024,1024]; flops counted = executed by that launch)",
            "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
            "traffic": roofline_traffic(), "peak_source": peak_src,
            "ms_per_launch": float(np.mean(gemm_ms)), "launches_per_step": float(np.mean(chunk_counts)),
            # one W_sc launch per step carries the events (a full-sized chunk); the step runs one such launch per chunk
            "share_of_step": float(np.sum([ms * c for ms, c in zip(gemm_ms, chunk_counts)])) / ms_dev,
            "step_algorithmic_tflops": step_flops / (ms_dev * 1e-3) / 1e12 / world,
            "step_frac_of_peak": step_flops / (ms_dev * 1e-3) / 1e12 / world / peak_tf,
        },
    }
    if world == 1 and not args.no_cpu_baseline:
        per = min(32, max(1, args.cpu_sample_questions))
        n_steps = max(1, args.cpu_sample_questions // per)
        qps, dt = time_cpu(spec, sd, ent, rel, per, n_steps, 1)
        line["cpu_baseline"] = {"value": qps, "unit": "questions/s", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"{n_steps * per} questions of the same workload in batches of {per} after one "
                                          f"warm-up batch; oracle port of the reference path (torch fp32, "
                                          f"{os.cpu_count()} threads), {dt:.1f} s",
                                "stage_questions_per_s": cpu_stage_qps(n_steps * per)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
