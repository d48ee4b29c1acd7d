#!/usr/bin/env python
"""bench.py — questions/sec of the retrieval hot path (triple scoring + BFS/DDE + top-k).

    python bench.py --gpus 1 --steps 20 --warmup 3                 # this repo's CUDA path (config 2 + other_configs)
    python bench.py --config 3 --gpus 8 --steps 6 --warmup 3       # CWQ-shaped strong scaling as the main line
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
import gc
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

from evi_rag_b200.synthetic import CWQ_50K, WEBQSP_10K, generate_questions, make_embedding_tables  # noqa: E402
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
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 5],
                    help="BASELINE.json config the main line reports: 2 = WebQSP-shaped 10k (the headline), 3 = CWQ-shaped "
                         "4-hop strong scaling, 5 = mask ablation sweep (k = 500, DDE on). Config 4 (100M-triple pool) "
                         "has its own driver: scripts/pool_scale.py")
    ap.add_argument("--no-other-configs", action="store_true",
                    help="skip the compact other_configs array (configs 3 and 5, a few steps each) of the default run")
    ap.add_argument("--precision", default="fast", choices=["fast", "strict"])
    ap.add_argument("--entity-rows", default="auto", choices=["auto", "batch", "table"],
                    help="entity projection + P once per distinct id of every batch, or once per row of the device-resident "
                         "table inside each timed region (auto: table when the projected rows fit 16 GB)")
    ap.add_argument("--cwq-questions-per-step", type=int, default=128,
                    help="config 3: questions of the fixed global batch split over the ranks by cumulative edge count")
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


def time_cpu(spec, sd, ent, rel, questions_per_step, steps, warmup, k=TOP_K):
    torch.set_num_threads(os.cpu_count() or 1)
    ent_t, rel_t = torch.from_numpy(ent), torch.from_numpy(rel)
    batches = [generate_questions(spec, q_start=(i * questions_per_step) % spec.num_questions,
                                  q_count=questions_per_step) for i in range(steps + warmup)]
    for qb in batches[:warmup]:
        cpu_reference_step(qb, sd, ent_t, rel_t, k)
    for key in CPU_STAGE_SECONDS:
        CPU_STAGE_SECONDS[key] = 0.0
    t0 = time.perf_counter()
    for qb in batches[warmup:]:
        cpu_reference_step(qb, sd, ent_t, rel_t, k)
    dt = time.perf_counter() - t0
    return steps * questions_per_step / dt, dt


def cpu_stage_qps(num_questions):
    """questions/s of each CPU stage alone over the timed sample."""
    return {key: (num_questions / sec if sec > 0 else None) for key, sec in CPU_STAGE_SECONDS.items()}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    wl = workload(args, args.config)
    spec = wl["spec"]
    per = args.ref_questions_per_step if wl["id"] != 3 else 1  # a bounded sample per step: CWQ questions are 4x larger
    sd = make_state_dict(0, trained_like=True)
    ent, rel = make_embedding_tables(WEBQSP_10K)
    qps, dt = time_cpu(spec, sd, ent, rel, per, args.steps, max(1, args.warmup), k=wl["k"])
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": "questions/sec (triple scoring + top-k)", "value": qps,
        "unit": "questions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "baseline_config": wl["id"], "top_k": wl["k"], "bfs_depth": wl["depth"],
                   "questions_per_step": per},
        "cpu_baseline": {"value": qps, "unit": "questions/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {per} questions of the same workload; "
                                   "oracle port of the reference (pure-Python reference cannot travel), torch fp32, "
                                   f"{cores} threads",
                         "stage_questions_per_s": cpu_stage_qps(args.steps * per)},
        "e2e": {"value": qps, "unit": "questions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_name(qps=None):
    """The same string in both arms (the driver compares it); the batch size is config.questions_per_step."""
    return ("webqsp_10k (BASELINE config[1]): 10k WebQSP-shaped questions, E_q~LogNormal(ln5000,0.4), N_q=E_q/3.3, "
            "1024-d, 2-hop BFS + DDE(2+2) + bidirectional scoring + top-100")


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
    """(sustained bf16 TFLOP/s, burst bf16 TFLOP/s, HBM GB/s, source). The fallback is the pool's last driver-written
    MEASURED_PEAKS.json (2026-11-04: 1401.5 sustained / 1663.7 burst / 6450.9 GB/s)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return (d.get("bf16_tflops_sustained", 1401.5), d.get("bf16_tflops", 1663.7), d.get("hbm_gbs", 6450.9),
                "MEASURED_PEAKS.json (cuBLAS bf16 sustained / burst, copy bandwidth)")
    return 1401.5, 1663.7, 6450.9, "fallback: constants of the pool's MEASURED_PEAKS.json of 2026-11-04"


def roofline_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the roofline kernels, from the committed
    `ncu --set full` capture (profiles/roofline_traffic.json, written by scripts/summarize_ncu.py): STATIC numbers of the
    capture named there, not of this run. Returns (tensor-kernel bytes, tail-kernel bytes, label)."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if not os.path.exists(p):
        return None, None, None
    try:
        d = json.load(open(p))
        return (float(d["dram_bytes_per_launch"]), float(d["tail_dram_bytes_per_launch"]) if "tail_dram_bytes_per_launch" in d
                else None, f"static, from profiles/roofline_traffic.json ({d.get('capture', 'ncu --set full')})")
    except Exception:
        return None, None, None


def workload(args, cfg):
    """BASELINE.json configs as generator parameters."""
    if cfg == 2:
        return dict(id=2, spec=WEBQSP_10K, qps=args.questions_per_step, k=TOP_K, depth=BFS_DEPTH, dde=True, scaling="weak",
                    name=workload_name(args.questions_per_step))
    if cfg == 5:
        return dict(id=5, spec=WEBQSP_10K, qps=args.questions_per_step, k=500, depth=BFS_DEPTH, dde=True, scaling="weak",
                    name=("mask ablation sweep (BASELINE config[4]): config[1] inputs, one top-k at k=500 whose prefixes are "
                          "the k in {50,100,200,500} lists (retriever_topk_edge_writer.py:324-327), DDE on"))
    return dict(id=3, spec=CWQ_50K, qps=args.cwq_questions_per_step, k=TOP_K, depth=4, dde=True, scaling="strong",
                name=("cwq_50k (BASELINE config[2]): CWQ-shaped questions, E_q~LogNormal(ln20000,0.4), N_q=E_q/3.3, 4-hop "
                      "BFS + DDE(2+2) + bidirectional scoring + top-100; a FIXED global batch per step split over the ranks "
                      "by cumulative edge count (sharding.partition_by_edges)"))


def build_host_batches(wl, n_steps, rank, world, pin=True):
    """Host batches of this rank for every step, the per-rank question counts of every step (known to all ranks: no
    size exchange at run time) and, for the strong-scaling split, max/mean edges per rank."""
    from evi_rag_b200.pipeline import host_batch

    spec, QB = wl["spec"], wl["qps"]
    key = (spec.name, QB, wl["scaling"], n_steps, rank, world)
    if key in _QB_CACHE:  # the k / DDE grid of config 5 walks the same questions
        batches, counts, imbalance = _QB_CACHE[key]
    else:
        batches, counts, imbalance = _generate_batches(wl, n_steps, rank, world)
        _QB_CACHE.clear()
        _QB_CACHE[key] = (batches, counts, imbalance)
    hbs = []
    for qb in batches:
        hb = host_batch(qb, pin=pin)
        if not wl["dde"]:  # _forward_with_topic_ablation (retriever_module.py:459-467): zeroed topic_one_hot
            hb.topic_one_hot = torch.zeros_like(hb.topic_one_hot)
            if pin:
                hb.topic_one_hot = hb.topic_one_hot.pin_memory()
        hbs.append(hb)
    return hbs, counts, (float(np.mean(imbalance)) if imbalance else None)


_QB_CACHE = {}


def _generate_batches(wl, n_steps, rank, world):
    from evi_rag_b200.sharding import partition_by_edges

    spec, QB = wl["spec"], wl["qps"]
    batches, counts, imbalance = [], [], []
    if wl["scaling"] == "weak":
        n_global = (spec.num_questions + QB - 1) // QB
        for s in range(n_steps):
            per_rank = []
            for r in range(world):  # rank r walks its own batches
                q0 = ((s * world + r) % n_global) * QB
                per_rank.append((q0, min(QB, spec.num_questions - q0)))
            q0, n = per_rank[rank]
            batches.append(generate_questions(spec, q_start=q0, q_count=n))
            counts.append([c for _, c in per_rank])
    else:
        if QB < world:
            raise SystemExit(f"config 3 needs at least one question per rank: {QB} questions for {world} ranks")
        n_global = spec.num_questions // QB
        for s in range(n_steps):
            whole = generate_questions(spec, q_start=(s % n_global) * QB, q_count=QB)
            ecount = np.diff(whole.edge_ptr)
            parts = partition_by_edges(ecount, world)
            per_rank_edges = [int(ecount[a:b].sum()) for a, b in parts]
            imbalance.append(max(per_rank_edges) / max(1.0, float(np.mean(per_rank_edges))))
            a, b = parts[rank]
            batches.append(whole.slice(a, b))
            counts.append([b_ - a_ for a_, b_ in parts])
    return batches, counts, imbalance


class Exchanger:
    """The one data-path collective: every rank receives every rank's selected triples (sharding.gather_topk, one packed
    all-gather). It runs on a side stream behind the step's last kernel, so the next step's kernels start at once;
    the step's tensors stay referenced until two more steps have been queued (allocator safety across streams)."""

    KEYS = ("score", "index", "count", "head_id", "tail_id", "relation_id")

    def __init__(self, world, dev):
        self.world, self.dev = world, dev
        self.side = torch.cuda.Stream(device=dev) if world > 1 else None
        self.keep = []

    def __call__(self, res, nq, counts):
        if self.world == 1:
            return
        from evi_rag_b200.sharding import gather_topk

        compute = torch.cuda.current_stream(self.dev)
        self.side.wait_stream(compute)
        with torch.cuda.stream(self.side):
            out = gather_topk({k: res[k] for k in self.KEYS}, nq, counts=counts)
        self.keep.append((res, out))
        del self.keep[:-3]

    def join(self):
        if self.side is not None:
            torch.cuda.current_stream(self.dev).wait_stream(self.side)


def args_precision_strict(args) -> bool:
    return args.precision == "strict"


class Bench:
    """One process per GPU: the model with its device-resident tables, and the timing loops."""

    def __init__(self, args):
        import torch.distributed as dist

        from evi_rag_b200 import lib as evilib
        from evi_rag_b200.retriever import Retriever

        self.args, self.dist, self.evilib = args, dist, evilib
        self.rank, self.local_rank, self.world = dist_env()
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        evilib.load()  # fails loudly if the extension is missing
        self.sd = make_state_dict(0, trained_like=True)
        self.model = Retriever(emb_dim=1024, hidden_dim=1024, operand_dtype=args.operand_dtype, precision=args.precision,
                               pipeline_chunks=args.pipeline_chunks, validate=False,
                               project_entity_table={"auto": None, "batch": False, "table": True}[args.entity_rows]).eval()
        self.model.load_state_dict(self.sd, strict=True)
        self.model.to(self.dev)
        # WEBQSP_10K and CWQ_50K share the table shapes (200k entities, 6000 relations)
        self.ent, self.rel = make_embedding_tables(WEBQSP_10K)
        self.model.attach_embedding_tables(torch.from_numpy(self.ent), torch.from_numpy(self.rel))
        self.exchange = Exchanger(self.world, self.dev)

    def sync_all(self):
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize(self.dev)

    def measure(self, wl, steps, warmup, full):
        """Device-resident timing of `steps` steps of workload `wl` (inputs in HBM when the region starts), and with
        `full` the end-to-end timing through RetrievalPipeline.run_stream (pinned host ids in, host results out)."""
        from evi_rag_b200.pipeline import RetrievalPipeline, device_batch, h2d_bytes

        dev, model, ex = self.dev, self.model, self.exchange
        pipe = RetrievalPipeline(model, top_k=wl["k"], bfs_max_depth=wl["depth"], writer_scores=True, validate=False)
        host_batches, counts, imbalance = build_host_batches(wl, warmup + steps, self.rank, self.world)
        timed = host_batches[warmup:]
        m = dict(imbalance=imbalance, questions=sum(hb.num_graphs for hb in timed),
                 edges=sum(int(hb.edge_index.size(1)) for hb in timed), nodes=sum(int(hb.ptr_host[-1]) for hb in timed),
                 questions_all_ranks=sum(sum(c) for c in counts[warmup:]), h2d=h2d_bytes(timed[0]))
        dev_batches = [device_batch(hb, dev) for hb in host_batches]
        torch.cuda.synchronize(dev)
        mk = lambda: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        gemm_ev, tail_ev = [mk() for _ in range(steps)], [mk() for _ in range(steps)]
        stage_ev = {name: [mk() for _ in range(steps)] for name in ("bfs", "dde", "topk")}
        for a, b in gemm_ev + tail_ev + [e for evs in stage_ev.values() for e in evs]:  # torch creates the cudaEvent_t lazily; record once so .cuda_event is live
            a.record(); b.record()
        for i in range(warmup):
            ex(pipe.run_device(dev_batches[i]), dev_batches[i].num_graphs, counts[i])
        # one more untimed pass over the LARGEST timed batch: the caching allocator then owns blocks big enough for
        # every timed step (a cudaMalloc inside the region is a host-side stall of milliseconds), and the collector
        # stays off while the region is timed
        big = max(range(warmup, warmup + steps), key=lambda i: int(host_batches[i].edge_index.size(1)))
        ex(pipe.run_device(dev_batches[big]), dev_batches[big].num_graphs, counts[big])
        ex.join()
        self.sync_all()
        gc.collect()
        gc.disable()
        self.evilib.LAUNCHES["kernels"] = 0
        e0, e1 = mk()
        chunk_edges, chunk_counts = [], []
        model.invalidate_projected_table()  # the per-table entity projection is paid INSIDE the timed region
        t_host0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            model.gemm_events, model.tail_events = (gemm_ev[i], tail_ev[i]) if full else (None, None)
            model.dde_events = stage_ev["dde"][i] if full else None
            pipe.stage_events = {"bfs": stage_ev["bfs"][i], "topk": stage_ev["topk"][i]} if full else None
            b = dev_batches[warmup + i]
            res = pipe.run_device(b)
            ex(res, b.num_graphs, counts[warmup + i])
            chunk_edges.append(model.last_chunk_edges)
            chunk_counts.append(model.last_num_chunks)
        ex.join()
        e1.record()
        m["enqueue_ms"] = (time.perf_counter() - t_host0) * 1e3  # host time to queue the region (no synchronisation)
        self.sync_all()
        gc.enable()
        model.gemm_events = model.tail_events = model.dde_events = None
        pipe.stage_events = None
        lc = model.last_counts  # of the last timed step: tensor FLOPs the step really executed
        to_int = lambda v: int(v.item()) if torch.is_tensor(v) else int(v)
        if lc is not None:
            rows, pairs, e_last = to_int(lc["entity_rows"]), to_int(lc["pair_rows"]), lc["edges"]
            strict = args_precision_strict(self.args)
            per_edge = (9 if strict else 3) * e_last                      # W_iv m, W_sc sc_f, W_sc sc_b (x3 passes in strict)
            # entity projection and P, three passes each: per distinct id of the step, or (projected-table mode) the
            # whole table once per timed region, shared by its steps
            per_id = 3 * rows * 2 + 3 * 2 * lc.get("projected_table_rows", 0) / float(steps)
            small = 3 * (lc["relation_rows"] + 3 * lc["questions"])       # relation / query / gate|bias projections
            m["executed_flops_last_step"] = 2.0 * 1024 * 1024 * (per_edge + per_id + (3 if strict else 1) * pairs + small)
            m["last_step_rows"] = {"entity_rows": rows, "pair_rows": pairs, "nodes": lc["nodes"], "edges": e_last,
                                   "projected_table_rows": lc.get("projected_table_rows", 0)}
        m["launches"] = self.evilib.LAUNCHES["kernels"]
        m["flags"] = int(res["flags"].item())
        m["ms_dev"] = e0.elapsed_time(e1)
        if full:
            m["gemm_ms"] = [a.elapsed_time(b) for a, b in gemm_ev]
            m["tail_ms"] = [a.elapsed_time(b) for a, b in tail_ev]
            m["stage_ms"] = {name: [a.elapsed_time(b) for a, b in evs] for name, evs in stage_ev.items()}
        m["chunk_edges"], m["chunk_counts"] = chunk_edges, chunk_counts
        del dev_batches
        if not full:
            return m
        # ---------------- end to end: host buffers in, host results out -------------------------
        for r in pipe.run_stream(host_batches[:warmup], dev):  # also sizes the rotating host / device buffers
            pass
        self.sync_all()
        m["d2h"] = RetrievalPipeline.d2h_bytes(r)
        e2, e3 = mk()
        model.invalidate_projected_table()
        t_wall0 = time.perf_counter()
        e2.record()
        # the public streaming call: H2D of step i+1 overlaps the kernels of step i; every copy is inside the region
        it = iter([(hb.num_graphs, counts[warmup + i]) for i, hb in enumerate(timed)])
        post = (lambda res_: ex(res_, *next(it))) if self.world > 1 else None
        for r in pipe.run_stream(timed, dev, post_device=post):
            pass
        ex.join()
        e3.record()
        self.sync_all()
        m["wall_e2e_ms"] = (time.perf_counter() - t_wall0) * 1e3
        m["ms_e2e"] = max(e2.elapsed_time(e3), 0.0)
        m["host_batches"] = host_batches
        return m

    def reduce(self, m, keys_max, keys_sum):
        """max / sum over ranks of the named entries (timings: max; counters: sum)."""
        t = torch.tensor([m[k] for k in keys_max], dtype=torch.float64, device=self.dev)
        c = torch.tensor([m[k] for k in keys_sum], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            self.dist.all_reduce(c, op=self.dist.ReduceOp.SUM)
        out = dict(m)
        out.update({k: float(v) for k, v in zip(keys_max, t.tolist())})
        out.update({k: float(v) for k, v in zip(keys_sum, c.tolist())})
        return out


def measure_dropin(bench, wl, host_batches, warmup):
    """End to end through the REFERENCE's call sites: `Retriever.forward(batch)` followed by
    `RetrieverTopKEdgeWriter.on_test_batch_end` (retriever_topk_edge_writer.py:175-184; records :321-349) per batch,
    host ids in / Python records out. The H2D copies, the forward, the writer's top-k + payload + D2H and the record
    building are all inside the timed region; only the final torch.save of the artifact is outside (reported)."""
    import shutil
    import types

    from evi_rag_b200.pipeline import device_batch
    from evi_rag_b200.writer import RetrieverTopKEdgeWriter

    dev, model = bench.dev, bench.model
    out_dir = tempfile.mkdtemp(prefix="evi_dropin_")
    writer = RetrieverTopKEdgeWriter(output_dir=out_dir, split="test", topk_values=[10, wl["k"]])

    def as_reference_batch(hb, i):
        b = device_batch(hb, dev)  # ids only: the tables are device-resident
        b.labels = torch.zeros(int(hb.edge_index.size(1)), dtype=torch.float32, device=dev)
        b.answer_entity_ids = torch.zeros(hb.num_graphs, dtype=torch.long)
        b.answer_entity_ids_ptr = torch.arange(hb.num_graphs + 1, dtype=torch.long)
        b.sample_id = [f"q{i}_{j}" for j in range(hb.num_graphs)]
        # a PyG Batch carries its per-graph edge offsets (the reference slices by them, retriever_topk_edge_writer.py:389-395)
        b._slice_dict = {"edge_index": torch.from_numpy(hb.edge_ptr_host)}
        return b

    def run(batches, base):
        for i, hb in enumerate(batches):
            b = as_reference_batch(hb, base + i)
            writer.on_test_batch_end(None, None, model(b), b, base + i)
        writer._drain(0)  # the last batches' copies have landed; the selected triples are host columns now

    writer.on_test_start()
    run(host_batches[:warmup], 0)
    torch.cuda.synchronize(dev)
    writer.on_test_start()  # drop the warm-up batches
    timed = host_batches[warmup:]
    gc.collect()
    gc.disable()  # the region allocates ~14k dicts per batch: keep the cyclic collector out of the timing
    model.invalidate_projected_table()
    t0 = time.perf_counter()
    run(timed, warmup)  # with materialise="batch" the records of batch i-1 are built while the device runs batch i
    torch.cuda.synchronize(dev)
    t_build0 = time.perf_counter()
    records = list(writer._records) + [r for item in writer._columnar for r in writer._materialise(item)]
    t1 = time.perf_counter()
    gc.enable()
    dt, t_build_end = t1 - t0, t1 - t_build0
    model._pending_flags = True
    model.check_flags()  # one read-back for the whole run instead of one synchronisation per forward
    t2 = time.perf_counter()
    writer.on_test_end(types.SimpleNamespace(global_rank=0))  # torch.save + manifest (and the records again in "end" mode)
    t_end = time.perf_counter() - t2
    n_q = sum(hb.num_graphs for hb in timed)
    shutil.rmtree(out_dir, ignore_errors=True)
    return {"value": n_q / dt, "unit": "questions/s", "ms_per_step": 1e3 * dt / len(timed), "records": len(records),
            "materialise": writer.materialise, "record_builder": "csrc/records.c" if writer.materialise == "batch" else "python",
            "record_build_after_loop_s": t_build_end, "artifact_write_s": t_end,
            "path": "Retriever.forward(batch) + RetrieverTopKEdgeWriter.on_test_batch_end per batch: host ids in, the "
                    "reference's per-question records (triplets_by_k dicts) out, ALL inside the timed region (wall clock, "
                    "synchronised at both ends); only the final torch.save of the artifact is outside "
                    "(artifact_write_s); validate=False with one flag read-back at the end"}


def measure_precision(bench, wl, questions=8):
    """Scores of the CUDA path against the oracle on a slice of THIS workload (bench tables, bench weights): the
    numbers behind config.precision. Runs in the cpu_baseline leg (the one place bench.py may execute oracle/)."""
    import types

    from oracle import evi_oracle as orc

    qb = generate_questions(wl["spec"], q_start=1024, q_count=questions)
    t = torch.from_numpy
    ent, rel = torch.from_numpy(bench.ent), torch.from_numpy(bench.rel)
    ref = orc.retriever_forward(bench.sd, edge_index=t(qb.edge_index), ptr=t(qb.ptr), question_emb=t(qb.question_emb),
                                node_embeddings=ent.index_select(0, t(qb.node_embedding_ids)),
                                node_embedding_ids=t(qb.node_embedding_ids),
                                edge_embeddings=rel.index_select(0, t(qb.edge_attr)), topic_one_hot=t(qb.topic_one_hot))
    b = types.SimpleNamespace(edge_index=t(qb.edge_index), ptr=t(qb.ptr), question_emb=t(qb.question_emb),
                              node_embedding_ids=t(qb.node_embedding_ids), edge_attr=t(qb.edge_attr),
                              topic_one_hot=t(qb.topic_one_hot))
    got = bench.model(b).logits.double().cpu()
    want = ref.logits.double()
    return {"rel_l2": float(torch.linalg.norm(got - want) / torch.linalg.norm(want)),
            "max_rel": float((got - want).abs().max() / want.abs().max()),
            "sample": f"{questions} questions / {qb.num_edges} edges of this workload against the fp32 oracle"}


def compact_case(bench, wl, steps, warmup, tag):
    """A short device-resident measurement of another BASELINE config for the other_configs array."""
    m = bench.measure(wl, steps, warmup, full=False)
    r = bench.reduce(m, ["ms_dev"], ["questions", "edges"])
    out = {"config": wl["id"], "case": tag, "scaling": wl["scaling"], "top_k": wl["k"], "bfs_depth": wl["depth"],
           "dde": wl["dde"], "steps": steps, "warmup": warmup, "ms_per_step": r["ms_dev"] / steps,
           "questions_per_s": r["questions"] / (r["ms_dev"] * 1e-3), "edges_per_s": r["edges"] / (r["ms_dev"] * 1e-3),
           "questions_per_step_all_ranks": r["questions"] / steps}
    if m["imbalance"] is not None:
        out["imbalance_max_over_mean_edges"] = m["imbalance"]
    return out


def other_configs(bench, args):
    """Configs 3 and 5 at a handful of steps each, visible to the driver in the default run. Every rank takes part
    (the gather of the selected triples is a collective)."""
    out = []
    steps, warmup = 3, 1
    out.append(compact_case(bench, workload(args, 3), steps, warmup, "config3 cwq 4-hop, fixed global batch"))
    base5 = workload(args, 5)
    for dde in (True, False):  # retriever_module.py:459-467: DDE off = topic_one_hot zeroed
        for k in (50, 100, 200, 500):  # run_retriever_mask_ablation.sh:113-118
            wl = dict(base5, k=k, dde=dde)
            out.append(compact_case(bench, wl, steps, warmup, f"config5 k={k} dde={'on' if dde else 'off'}"))
    return out


def run_ours(args):
    bench = Bench(args)
    rank, world, dist = bench.rank, bench.world, bench.dist
    wl = workload(args, args.config)
    sampler = ClockSampler(bench.local_rank)
    if rank == 0:
        sampler.start()
    m = bench.measure(wl, args.steps, args.warmup, full=True)
    clocks = sampler.stop() if rank == 0 else None
    r = bench.reduce(m, ["ms_dev", "ms_e2e", "wall_e2e_ms", "enqueue_ms"], ["questions", "edges", "nodes", "launches"])
    dropin = None
    if world == 1:  # the reference's eval is one process (src/eval.py:32-77)
        dropin = measure_dropin(bench, wl, m["host_batches"], args.warmup)
    others = None
    if args.config == 2 and not args.no_other_configs:
        others = other_configs(bench, args)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ms_dev, ms_e2e = r["ms_dev"], r["ms_e2e"]
    q_all, e_all, n_all = r["questions"], r["edges"], r["nodes"]
    peak_tf, peak_burst, peak_hbm, peak_src = measured_peaks()
    gemm_ms, tail_ms, chunk_edges, chunk_counts = m["gemm_ms"], m["tail_ms"], m["chunk_edges"], m["chunk_counts"]
    gemm_flops = [2 * ce * F_SC_PER_EDGE_DIR for ce in chunk_edges]  # both directions in one launch
    achieved_tf = float(np.mean([f / (ms * 1e-3) / 1e12 for f, ms in zip(gemm_flops, gemm_ms)]))
    step_flops = e_all * F_EDGE + n_all * F_NODE
    tensor_traffic, tail_traffic, traffic_label = roofline_traffic()
    # the worst non-tensor kernel: the tail streams 6 KB per edge from HBM (fp16 W_iv m row + two u16 W_sc sc rows, block
    # scales, scalars, ids, logits); the three fp32 rows it gathers per edge (P[h], P[t], Rq[pair], 12 KB) are per-node /
    # per-pair rows that mostly hit L2 and are not counted
    tail_bytes = [ce * (6 * 1024 + 32 + 16 + 12 + 12) for ce in chunk_edges]
    tail_gbs = float(np.mean([b / (ms * 1e-3) / 1e9 for b, ms in zip(tail_bytes, tail_ms)]))
    fast_rows = args.precision == "fast" and W_SC_EPILOGUE == "EpiQ16Tma"
    line = {
        "metric": "questions/sec (triple scoring + top-k)",
        "value": q_all / (ms_dev * 1e-3),
        "unit": "questions/s",
        "n_gpus": world,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True,
        "scaling": wl["scaling"],
        "vs_baseline": None,
        "dtype": "f16 operands / f32 accumulate" if args.operand_dtype == "fp16" else "bf16 operands / f32 accumulate",
        "data": "synthetic",
        "config": {
            "workload": wl["name"], "baseline_config": wl["id"], "top_k": wl["k"], "bfs_depth": wl["depth"],
            "questions_per_step": wl["qps"],
            "parallelism": f"questions sharded x{world}",
            "l2": "fresh batch every step; per-step working set (operand rows + activations) is several GB >> 126 MB L2",
            "edges_per_step": e_all / args.steps / world, "validation_flags": m["flags"],
            "precision_mode": args.precision,
            "entity_rows": ("tanh(W_e x) and W_err node_repr computed once per row of the device-resident entity table "
                            f"({m['last_step_rows']['projected_table_rows']} rows) INSIDE every timed region (the projected "
                            "table is invalidated when the region starts), not once per batch"
                            if m.get("last_step_rows", {}).get("projected_table_rows") else
                            "projected once per distinct embedding id of each batch"),
            "intermediates": ("per-node / per-relation / per-question contractions over [hi | lo] fp16 pairs (three tensor "
                              "passes); per-edge operands fp16 with the structure context centred on its batch mean; W_iv m "
                              "rows fp16, W_sc sc rows 16-bit block fixed point; everything else fp32"
                              if fast_rows else "strict mode or an experiment knob: fp32 partial-product rows"),
        },
        "e2e": {"value": q_all / (ms_e2e * 1e-3), "unit": "questions/s",
                "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": r["wall_e2e_ms"] / args.steps,
                "path": "RetrievalPipeline.run_stream: pinned host ids in, selected triples + BFS distances out"},
        "gpu_launches": int(r["launches"]),
        "host_enqueue_ms_per_step": r["enqueue_ms"] / args.steps,
        "clocks": clocks,
        "roofline": {
            "bound": "tensor", "kernel": f"gemm2::gemm2_kernel<{W_SC_EPILOGUE}> (W_sc block of state_net.0: "
                                         "[2E_chunk,1024]x[1024,1024]; flops counted = executed by that launch)",
            "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
            "frac_of_burst": achieved_tf / peak_burst, "peak_burst": peak_burst,
            "traffic": tensor_traffic, "traffic_source": traffic_label, "peak_source": peak_src,
            "ms_per_launch": float(np.mean(gemm_ms)), "launches_per_step": float(np.mean(chunk_counts)),
            # one W_sc launch per step carries the events (a full-sized chunk); the step runs one such launch per chunk
            "share_of_step": float(np.sum([ms * c for ms, c in zip(gemm_ms, chunk_counts)])) / ms_dev,
            "step_algorithmic_tflops": step_flops / (ms_dev * 1e-3) / 1e12 / world,
            "step_frac_of_peak": step_flops / (ms_dev * 1e-3) / 1e12 / world / peak_tf,
            # what the tensor cores really executed in rank 0's last timed step (the linearity split of DESIGN.md §4 halves
            # the per-edge work; the [hi | lo] passes add to the per-id work), over the average step time
            "step_executed_tflops": m.get("executed_flops_last_step", 0.0) / (ms_dev / args.steps * 1e-3) / 1e12,
            "step_executed_frac_of_peak": m.get("executed_flops_last_step", 0.0) / (ms_dev / args.steps * 1e-3) / 1e12 / peak_tf,
            "last_step_rows": m.get("last_step_rows"),
        },
        "roofline_non_tensor": {
            "bound": "hbm", "kernel": "evi::score_tail_bulk_kernel (z assembly + LayerNorm + GELU + head dot + combine)",
            "achieved": tail_gbs, "peak": peak_hbm, "unit": "GB/s", "frac": tail_gbs / peak_hbm,
            "bytes_per_edge": 6 * 1024 + 72, "traffic": tail_traffic, "traffic_source": traffic_label,
            "ms_per_launch": float(np.mean(tail_ms)),
            "share_of_step": float(np.sum([ms * c for ms, c in zip(tail_ms, chunk_counts)])) / ms_dev,
            "note": "issue-bound fp32 code, not bandwidth-bound (profiles/*_kernels_ncu.md): the fraction is reported, "
                    "not claimed as a roofline hit",
        },
    }
    # the graph / selection stages against the HBM roofline on SURVEY.md §8(d)'s algorithmic bytes (rank 0's steps). Their
    # working sets (5k edges per question) sit in L2 / shared memory, so the fractions are low by construction.
    e_step, n_step, q_step = m["edges"] / args.steps, m["nodes"] / args.steps, m["questions"] / args.steps
    stage_bytes = {"bfs": 8 * e_step + 12 * n_step,                      # adjacency entries + dist + frontier, per BFS
                   "dde": 96 * e_step + 48 * n_step,                     # 4 rounds x 24 B per edge + features + degrees
                   "topk": 4 * e_step + (8 + 40) * wl["k"] * q_step}     # scores read + [B, k] results + writer payload
    line["stage_rooflines"] = {
        name: {"bound": "hbm", "ms_per_step": float(np.mean(ms)), "algorithmic_bytes_per_step": stage_bytes[name],
               "achieved": stage_bytes[name] / (float(np.mean(ms)) * 1e-3) / 1e9, "unit": "GB/s", "peak": peak_hbm,
               "frac": stage_bytes[name] / (float(np.mean(ms)) * 1e-3) / 1e9 / peak_hbm}
        for name, ms in m["stage_ms"].items()}
    if m["imbalance"] is not None:
        line["config"]["imbalance_max_over_mean_edges"] = m["imbalance"]
    if dropin is not None:
        line["e2e_dropin"] = dropin
    if others is not None:
        line["other_configs"] = others
    if world == 1 and not args.no_cpu_baseline:
        per = min(32, max(1, args.cpu_sample_questions)) if wl["id"] != 3 else 4
        n_steps = max(1, (args.cpu_sample_questions if wl["id"] != 3 else 16) // per)
        qps, dt = time_cpu(wl["spec"], bench.sd, bench.ent, bench.rel, per, n_steps, 1, k=wl["k"])
        line["cpu_baseline"] = {"value": qps, "unit": "questions/s", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"{n_steps * per} questions of the same workload in batches of {per} after one "
                                          f"warm-up batch; oracle port of the reference path (torch fp32, "
                                          f"{os.cpu_count()} threads), {dt:.1f} s",
                                "stage_questions_per_s": cpu_stage_qps(n_steps * per)}
        line["config"]["precision"] = measure_precision(bench, wl, questions=8 if wl["id"] != 3 else 2)
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
