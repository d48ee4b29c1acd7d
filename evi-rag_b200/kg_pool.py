"""Single-graph retrieval over a sharded triple pool (BASELINE config 4).

The reference only knows per-question subgraphs (AGENTS.md:47); its dataset build cuts them out of the knowledge
graph offline. This module does that cut on the device, per query batch, from one large triple pool:

1. depth-capped multi-source BFS from the query's seed entities over the undirected pool — the hop-distance
   semantics of ``_bfs_dist`` (scripts/build_retrieval_pipeline.py:610-631);
2. candidate triples = edges whose two endpoints were reached, capped at ``cap`` by smallest
   ``(dist_head + dist_tail, edge id)``;
3. the candidates of a query, in ascending edge id, become one question subgraph (``topic_one_hot`` =
   one-hot(is seed)) and go through the same ``Retriever`` kernels and ragged top-k as every other config.

Sharding: the pool is edge-partitioned by contiguous head-node range; every shard indexes its own
triples by head and by tail (global node ids). A BFS level is "expand the frontier over my triples" followed by an
exchange of the newly reached (query, node) pairs, so all shards hold the same visited set; the candidate cut is
"my `cap` best" per shard followed by an exchange and the same selection over the union. Scoring shards by query
(the subgraphs are small once cut). Shards can be several objects on one device (tests) or one per rank of a
``torch.distributed`` group (NCCL on the GPU box).

Index construction (``TriplePoolShard.from_triples``) is a one-off load-time sort, like the reference's offline
parquet build; the per-query path below is kernels from ``csrc/pool_kernels.cu`` plus sizes read back at batch
boundaries.
"""
from __future__ import annotations

import ctypes as C
import time
import types
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import lib as _lib
from .sharding import gather_topk, partition_by_edges
from .topk import segmented_topk

_EMPTY = -1  # 0xFFFF... as int64: padding of exchange buffers


def _next_pow2(x: int) -> int:
    p = 1
    while p < x:
        p <<= 1
    return p


def partition_pool_by_head(heads: np.ndarray, num_nodes: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous head-node ranges ``[(lo, hi)] * world`` holding about ``E / world`` triples each."""
    deg = np.bincount(np.asarray(heads, dtype=np.int64), minlength=num_nodes)
    return partition_by_edges(deg, world)


@dataclass
class TriplePoolShard:
    """Device-resident index of one shard's triples by head (fwd) and by tail (rev), global node ids."""

    fwd_ptr: torch.Tensor    # [num_nodes + 1] i32
    fwd_other: torch.Tensor  # [M] i32 tail
    fwd_rel: torch.Tensor    # [M] i32
    fwd_eid: torch.Tensor    # [M] i64 global edge id
    rev_ptr: torch.Tensor    # [num_nodes + 1] i32
    rev_other: torch.Tensor  # [M] i32 head
    num_nodes: int

    @classmethod
    def from_triples(cls, heads, rels, tails, edge_ids, num_nodes: int, device) -> "TriplePoolShard":
        dev = torch.device(device)
        put = lambda x: (x if torch.is_tensor(x) else torch.as_tensor(np.asarray(x))).to(device=dev, dtype=torch.long)
        h, t, r, e = put(heads), put(tails), put(rels), put(edge_ids)
        if h.numel() and (int(h.min()) < 0 or int(t.min()) < 0 or int(h.max()) >= num_nodes or int(t.max()) >= num_nodes):
            raise ValueError("triple endpoints must lie in [0, num_nodes)")

        def rows(key):
            order = torch.argsort(key, stable=True)
            ptr = torch.zeros(num_nodes + 1, dtype=torch.long, device=dev)
            ptr[1:] = torch.cumsum(torch.bincount(key, minlength=num_nodes), 0)
            return order, ptr.to(torch.int32)

        fo, fptr = rows(h)
        ro, rptr = rows(t)
        pad = lambda x, dt: (x if x.numel() else torch.zeros(1, dtype=x.dtype, device=dev)).to(dt).contiguous()
        return cls(fptr.contiguous(), pad(t[fo], torch.int32), pad(r[fo], torch.int32), pad(e[fo], torch.long),
                   rptr.contiguous(), pad(h[ro], torch.int32), int(num_nodes))

    @property
    def num_triples(self) -> int:
        return int(self.fwd_ptr[-1])

    def c_struct(self) -> _lib.PoolShard:
        s = _lib.PoolShard()
        s.fwd_ptr, s.fwd_other, s.fwd_rel = self.fwd_ptr.data_ptr(), self.fwd_other.data_ptr(), self.fwd_rel.data_ptr()
        s.fwd_eid, s.rev_ptr, s.rev_other = self.fwd_eid.data_ptr(), self.rev_ptr.data_ptr(), self.rev_other.data_ptr()
        s.num_nodes = self.num_nodes
        return s


class _ShardState:
    """Per-shard scratch of one query batch: the visited hash table and the visited-pair list."""

    def __init__(self, shard: TriplePoolShard, table_capacity: int, visited_cap: int, dev):
        self.shard = shard
        self.keys = torch.empty(table_capacity, dtype=torch.long, device=dev)  # dist << 58 | query << 32 | node
        self.local_id = torch.empty(table_capacity, dtype=torch.int32, device=dev)
        self.visited = torch.empty(visited_cap, dtype=torch.long, device=dev)
        self.visited_count = torch.zeros(1, dtype=torch.long, device=dev)
        self.fresh = torch.empty(visited_cap, dtype=torch.long, device=dev)
        self.fresh_count = torch.zeros(1, dtype=torch.long, device=dev)
        self.flags = torch.zeros(1, dtype=torch.int32, device=dev)
        self.table = _lib.PoolTable()
        self.table.keys = self.keys.data_ptr()
        self.table.local_id, self.table.capacity = self.local_id.data_ptr(), table_capacity
        self.cshard = shard.c_struct()
        self.bounds = [0]
        # work list for the segments of long adjacency rows (include/evi_rag_b200.h: evi_pool_overflow)
        self.ovf_items = torch.empty(1 << 20, dtype=torch.long, device=dev)
        self.ovf_count = torch.zeros(1, dtype=torch.long, device=dev)
        self.overflow = _lib.PoolOverflow()
        self.overflow.items, self.overflow.count = self.ovf_items.data_ptr(), self.ovf_count.data_ptr()
        self.overflow.capacity = self.ovf_items.numel()


class PoolRetriever:
    """Query batch -> per-query candidate subgraphs -> scores -> top-k global edge ids."""

    def __init__(self, shards: Sequence[TriplePoolShard], model: Any, *, node_embedding_ids: Optional[torch.Tensor] = None,
                 cap: int = 20000, top_k: int = 500, depth: int = 2, table_capacity: int = 1 << 22,
                 max_table_capacity: int = 1 << 30, group=None, num_sms: int = 148,
                 score_edges_per_batch: int = 1 << 18) -> None:
        if not shards:
            raise ValueError("at least one shard is required")
        if not (0 < cap <= _lib.POOL_MAX_CAP):
            raise ValueError(f"cap must be in [1, {_lib.POOL_MAX_CAP}], got {cap}")
        if not (0 <= depth <= 60):
            raise ValueError(f"depth must be in [0, 60], got {depth}")
        if table_capacity & (table_capacity - 1):
            raise ValueError("table_capacity must be a power of two")
        self.shards = list(shards)
        self.model = model
        self.dev = self.shards[0].fwd_ptr.device
        if self.dev.type != "cuda":
            raise RuntimeError("PoolRetriever runs on CUDA tensors only (no CPU fallback).")
        self.cap, self.top_k, self.depth = int(cap), int(top_k), int(depth)
        # the table starts small (32 MB of keys at 2^22 slots: L2-resident probes) and is re-sized x4 whenever a batch
        # fills half of it; the size that worked is kept for the next batch
        self.table_capacity = int(table_capacity)
        self.max_table_capacity = max(int(max_table_capacity), self.table_capacity)
        self.node_embedding_ids = None if node_embedding_ids is None else \
            node_embedding_ids.to(device=self.dev, dtype=torch.long).contiguous()
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        if self.world > 1 and len(self.shards) != 1:
            raise ValueError("with torch.distributed every rank holds exactly one shard")
        self.num_sms = int(num_sms)
        # scoring walks the slab's queries in sub-batches of about this many candidate edges (the scorer's per-node
        # rows cost ~20 KB per node: a 64k-query slab does not fit as one batch)
        self.score_edges_per_batch = max(1, int(score_edges_per_batch))
        self.timings: Dict[str, float] = {}
        self.split_long_rows = True  # queue the segments of long adjacency rows for a second launch (evi_pool_overflow)
        self.profile = False  # True: synchronise around the frontier exchange and accumulate timings["exchange_s"]
        self.L = _lib.load()

    # ------------------------------------------------------------------ exchange helpers
    def _gather_var(self, t: torch.Tensor, n: int) -> List[torch.Tensor]:
        """All ranks' ``t[:n]`` (different n per rank): sizes first, then one padded all-gather."""
        sizes = [torch.zeros(1, dtype=torch.long, device=self.dev) for _ in range(self.world)]
        dist.all_gather(sizes, torch.tensor([n], dtype=torch.long, device=self.dev), group=self.group)
        sizes = [int(s.item()) for s in sizes]
        m = max(max(sizes), 1)
        buf = torch.full((m,), _EMPTY, dtype=t.dtype, device=self.dev)
        buf[:n] = t[:n]
        out = torch.empty(self.world * m, dtype=t.dtype, device=self.dev)
        dist.all_gather_into_tensor(out, buf, group=self.group)
        return [out[r * m: r * m + sizes[r]] for r in range(self.world)]

    def _traverse(self, seed_pairs: torch.Tensor) -> Optional[List[_ShardState]]:
        """Depth-capped BFS of the whole batch on every local shard with the per-level frontier exchange. Returns the
        shard states, or None when the visited set outgrew the table (the caller re-sizes and repeats)."""
        L, dev, stream = self.L, self.dev, _lib.stream_handle()
        visited_cap = self.table_capacity // 4  # load factor <= 0.25 keeps linear-probe chains and CAS retries short
        states = [_ShardState(s, self.table_capacity, visited_cap, dev) for s in self.shards]
        multi = len(states) > 1 or self.world > 1
        for st in states:
            _lib.check(L.evi_pool_reset(C.byref(st.table), stream), "evi_pool_reset")
            _lib.check(L.evi_pool_insert(C.byref(st.table), seed_pairs.data_ptr(), int(seed_pairs.numel()), 0,
                                         st.visited.data_ptr(), st.visited_count.data_ptr(), visited_cap,
                                         st.flags.data_ptr(), stream), "evi_pool_insert")
            st.bounds.append(min(int(st.visited_count.item()), visited_cap))
        if self._overflowed(states):
            return None
        for level in range(self.depth):
            for st in states:
                st.fresh_count.zero_()
                _lib.check(L.evi_pool_expand(
                    C.byref(st.table), C.byref(st.cshard), st.visited.data_ptr(), st.bounds[level], st.bounds[level + 1],
                    level, st.visited_count.data_ptr(), visited_cap, st.fresh.data_ptr() if multi else None,
                    st.fresh_count.data_ptr() if multi else None, visited_cap, st.flags.data_ptr(),
                    C.byref(st.overflow) if self.split_long_rows else None, self.num_sms, stream), "evi_pool_expand")
            if multi:  # frontier exchange: every shard learns the pairs the others reached at this level
                if self.profile:
                    torch.cuda.synchronize(dev)
                    t_x0 = time.perf_counter()
                lists = [st.fresh[: min(int(st.fresh_count.item()), visited_cap)] for st in states]
                if self.world > 1:
                    remote = self._gather_var(lists[0], int(lists[0].numel()))
                    others_of = [[remote[r] for r in range(self.world) if r != self.rank]]
                else:
                    others_of = [[lists[j] for j in range(len(states)) if j != i] for i in range(len(states))]
                for st, others in zip(states, others_of):
                    for pairs in others:
                        if pairs.numel():
                            pairs = pairs.contiguous()
                            _lib.check(L.evi_pool_insert(C.byref(st.table), pairs.data_ptr(), int(pairs.numel()), level + 1,
                                                         st.visited.data_ptr(), st.visited_count.data_ptr(),
                                                         visited_cap, st.flags.data_ptr(), stream), "evi_pool_insert")
                if self.profile:
                    torch.cuda.synchronize(dev)
                    self.timings["exchange_s"] = self.timings.get("exchange_s", 0.0) + time.perf_counter() - t_x0
            for st in states:
                st.bounds.append(min(int(st.visited_count.item()), visited_cap))
            if self._overflowed(states):
                return None
        return states

    def _overflowed(self, states: List[_ShardState]) -> bool:
        """True when any shard (on any rank) filled its table beyond half or overran its visited list."""
        word = torch.stack([st.flags for st in states]).max().view(1)
        if self.world > 1:
            dist.all_reduce(word, op=dist.ReduceOp.MAX, group=self.group)
        return int(word.item()) != 0

    # ------------------------------------------------------------------ steps 1-2: BFS + candidate cut
    def extract(self, seeds: torch.Tensor, seed_ptr: torch.Tensor) -> types.SimpleNamespace:
        with torch.cuda.device(self.dev):  # the C ABI launches on the current device
            return self._extract_on_device(seeds, seed_ptr)

    def _extract_on_device(self, seeds: torch.Tensor, seed_ptr: torch.Tensor) -> types.SimpleNamespace:
        """``seeds`` = global node ids grouped by ``seed_ptr [Q+1]``. Returns the subgraph batch (device tensors under
        the reference's PyG attribute names) plus ``edge_global_id [E]`` and ``edge_ptr [Q+1]``."""
        dev = self.dev
        seeds = torch.as_tensor(seeds, dtype=torch.long, device=dev).view(-1)
        seed_ptr = torch.as_tensor(seed_ptr, dtype=torch.long, device=dev).view(-1)
        Q = int(seed_ptr.numel() - 1)
        if Q >= (1 << 26):
            raise ValueError("at most 2^26 queries per batch (query id bits of the visited-table key)")
        num_nodes = self.shards[0].num_nodes
        qid = torch.repeat_interleave(torch.arange(Q, device=dev), seed_ptr[1:] - seed_ptr[:-1])
        ok = (seeds >= 0) & (seeds < num_nodes)  # out-of-range sources are ignored (:616-618)
        seed_pairs = torch.where(ok, (qid << 32) | seeds.clamp(min=0), torch.full_like(seeds, _EMPTY)).contiguous()
        tick = self._tick
        tick(None)
        while True:
            states = self._traverse(seed_pairs)
            if states is not None:
                break
            if self.table_capacity >= self.max_table_capacity:
                raise RuntimeError(f"visited set does not fit {self.max_table_capacity} table slots: use smaller query "
                                   "batches or raise max_table_capacity")
            self.table_capacity = min(self.table_capacity * 4, self.max_table_capacity)
        multi = len(states) > 1 or self.world > 1
        self._last_state = states[0]  # kept for traversal_stats()
        tick("traverse_s")
        raw = [self._shard_candidates(st, Q) for st in states]
        tick("candidates_s")
        # Row stride of the fixed-shape [Q, stride] selection arrays: the batch's largest per-query candidate count
        # (power of two), not `cap` — uniform-seed queries keep a few hundred candidates against cap = 20000, and
        # [Q, cap] rows of a 64k-query slab would take tens of GB. Every count fits its row, so the selection is the
        # same; one read-back (and one MAX all-reduce across ranks) per slab.
        most = torch.stack([r[4] for r in raw]).max().view(1)
        if self.world > 1:
            dist.all_reduce(most, op=dist.ReduceOp.MAX, group=self.group)
        stride = min(self.cap, _next_pow2(max(int(most.item()), 1)))
        sels = [self._select(key, ht, rel, off, Q, stride) for key, ht, rel, off, _ in raw]
        tick("select_s")
        if self.world > 1:
            b = self._owner_merge_pack(states[0], sels[0], Q, stride)
            tick("merge_pack_s")
            return b
        if multi:
            n_sh = len(states)
            out_stride = min(self.cap, n_sh * stride)
            sel = self._merge(sels, Q, stride, out_stride)
            stride = out_stride
            tick("merge_s")
        else:
            sel = sels[0]
        b = self._pack(states[0], sel, Q, stride)
        tick("pack_s")
        return b

    def _tick(self, name: Optional[str]) -> None:
        """profile mode: synchronise and add the time since the previous tick to timings[name]"""
        if not self.profile:
            return
        torch.cuda.synchronize(self.dev)
        now = time.perf_counter()
        if name is not None:
            self.timings[name] = self.timings.get(name, 0.0) + now - self._t_last
        self._t_last = now

    def traversal_stats(self) -> Dict[str, int]:
        """Adjacency entries shard 0 read for the last batch (bandwidth accounting, not on the timed path): both
        rows of every expanded pair, and the head row of every visited pair twice (count pass + emit pass)."""
        st = self._last_state
        nodes = st.visited[: st.bounds[-1]] & 0xFFFFFFFF
        fdeg = (st.shard.fwd_ptr[nodes + 1] - st.shard.fwd_ptr[nodes]).long()
        rdeg = (st.shard.rev_ptr[nodes + 1] - st.shard.rev_ptr[nodes]).long()
        n_exp = st.bounds[self.depth] if self.depth < len(st.bounds) else st.bounds[-1]
        return {"visited_pairs": int(st.bounds[-1]),
                "expand_entries": int((fdeg[:n_exp] + rdeg[:n_exp]).sum()),
                "candidate_entries": int(2 * fdeg.sum())}

    def _select(self, key, ht, rel, seg_off, Q: int, cap: int):
        """``cap`` = rows' stride and the number of keys kept per query (``self.cap``, or the batch's largest count)."""
        dev = self.dev
        sort_len = _next_pow2(cap)
        u64 = lambda *s: torch.empty(s, dtype=torch.long, device=dev)
        stage_key, stage_ht, stage_rel = u64(Q, cap), u64(Q, cap), torch.empty((Q, cap), dtype=torch.int32, device=dev)
        out = {"key": u64(Q, cap), "ht": u64(Q, cap), "rel": torch.empty((Q, cap), dtype=torch.int32, device=dev),
               "count": torch.empty(Q, dtype=torch.int32, device=dev)}
        sort_ws = u64(Q, sort_len)
        _lib.check(self.L.evi_pool_select(key.data_ptr(), ht.data_ptr(), rel.data_ptr(), seg_off.data_ptr(), Q, cap, sort_len,
                                          stage_key.data_ptr(), stage_ht.data_ptr(), stage_rel.data_ptr(), sort_ws.data_ptr(),
                                          out["key"].data_ptr(), out["ht"].data_ptr(), out["rel"].data_ptr(),
                                          out["count"].data_ptr(), _lib.stream_handle()), "evi_pool_select")
        return out

    def _shard_candidates(self, st: _ShardState, Q: int):
        L, dev, stream = self.L, self.dev, _lib.stream_handle()
        ovf = C.byref(st.overflow) if self.split_long_rows else None
        nv = st.bounds[-1]
        count = torch.empty((Q, 8), dtype=torch.int32, device=dev)
        limit = torch.empty(Q, dtype=torch.int32, device=dev)
        emit_count = torch.empty(Q, dtype=torch.int32, device=dev)
        _lib.check(L.evi_pool_candidates(C.byref(st.table), C.byref(st.cshard), st.visited.data_ptr(), nv, 0, Q, self.cap,
                                         count.data_ptr(), limit.data_ptr(), emit_count.data_ptr(), None, None, None, None,
                                         None, ovf, self.num_sms, stream), "evi_pool_candidates")
        off = torch.zeros(Q + 1, dtype=torch.long, device=dev)
        off[1:] = torch.cumsum(emit_count.long(), 0)
        total = max(int(off[-1].item()), 1)
        key = torch.empty(total, dtype=torch.long, device=dev)
        ht = torch.empty(total, dtype=torch.long, device=dev)
        rel = torch.empty(total, dtype=torch.int32, device=dev)
        cursor = torch.zeros(Q, dtype=torch.int32, device=dev)
        _lib.check(L.evi_pool_candidates(C.byref(st.table), C.byref(st.cshard), st.visited.data_ptr(), nv, 1, Q, self.cap,
                                         count.data_ptr(), limit.data_ptr(), emit_count.data_ptr(), off.data_ptr(),
                                         cursor.data_ptr(), key.data_ptr(), ht.data_ptr(), rel.data_ptr(), ovf, self.num_sms,
                                         stream), "evi_pool_candidates")
        return key, ht, rel, off, emit_count.max()

    def _compact(self, sel: Dict[str, torch.Tensor], cap: int) -> Dict[str, torch.Tensor]:
        """The valid entries of [Q, cap] selection rows as flat lists grouped by query."""
        valid = torch.arange(cap, device=self.dev).unsqueeze(0) < sel["count"].unsqueeze(1)
        return {name: sel[name][valid] for name in ("key", "ht", "rel")}

    def _union(self, flats: List[Dict[str, torch.Tensor]], cnt: torch.Tensor, nq: int):
        """Per query the concatenation of every shard's survivors. ``flats[s]`` = shard s's entries grouped by query,
        ``cnt [S, nq]`` their per-query counts. Returns the ragged union (key, ht, rel) and its offsets [nq + 1]."""
        dev = self.dev
        seg_off = torch.zeros(nq + 1, dtype=torch.long, device=dev)
        seg_off[1:] = torch.cumsum(cnt.sum(dim=0), 0)
        before = torch.cumsum(cnt, dim=0) - cnt  # entries of earlier shards inside the query's segment
        total = sum(int(f["key"].numel()) for f in flats)
        cat = {"key": torch.empty(max(total, 1), dtype=torch.long, device=dev),
               "ht": torch.empty(max(total, 1), dtype=torch.long, device=dev),
               "rel": torch.empty(max(total, 1), dtype=torch.int32, device=dev)}
        qs = torch.arange(nq, device=dev)
        for r, f in enumerate(flats):
            n_r = int(f["key"].numel())
            if n_r == 0:
                continue
            qid = torch.repeat_interleave(qs, cnt[r], output_size=n_r)
            start_r = torch.cumsum(cnt[r], 0) - cnt[r]  # segment starts inside shard r's list
            dest = seg_off[qid] + before[r][qid] + (torch.arange(n_r, device=dev) - start_r[qid])
            for name in ("key", "ht", "rel"):
                cat[name][dest] = f[name]
        return cat, seg_off

    def _merge(self, sels: List[Dict[str, torch.Tensor]], Q: int, cap: int, out_cap: int):
        """Several shards on one device: the ``out_cap`` smallest keys per query over the union of their survivors."""
        flats = [self._compact(s, cap) for s in sels]
        cnt = torch.stack([s["count"].long() for s in sels])
        cat, seg_off = self._union(flats, cnt, Q)
        return self._select(cat["key"], cat["ht"], cat["rel"], seg_off, Q, out_cap)

    def _sub_batch_cuts(self, ep: List[int], Q: int) -> List[int]:
        """Query indices where the scoring sub-batches start: runs of whole queries of about ``score_edges_per_batch``
        candidate edges, cut on the WHOLE query list whatever the world size (a query is always scored inside the same
        sub-batch, so N-rank scores equal 1-rank scores bit for bit: the batch centre depends on the batch)."""
        cuts = [0]
        while cuts[-1] < Q:
            a = cuts[-1]
            c = a + 1
            while c < Q and ep[c + 1] - ep[a] <= self.score_edges_per_batch:
                c += 1
            cuts.append(c)
        return cuts

    def _owner_merge_pack(self, st: _ShardState, sel: Dict[str, torch.Tensor], Q: int, cap: int) -> types.SimpleNamespace:
        """One shard per rank: every rank becomes the OWNER of a contiguous run of the batch's queries (whole scoring
        sub-batches, balanced by candidate edges) and only merges, relabels, packs and scores those. The per-query
        survivor counts of every shard are all-gathered first — keys are unique per query, so the merged count is
        min(sum of counts, cap) and every rank can lay out the sub-batches without doing the merge — then ONE uneven
        all-to-all per array moves each shard's survivors to their owners."""
        dev, W = self.dev, self.world
        flat = self._compact(sel, cap)
        counts = sel["count"].long().contiguous()
        cnt_all = torch.empty((W, Q), dtype=torch.long, device=dev)
        dist.all_gather_into_tensor(cnt_all, counts, group=self.group)
        final = cnt_all.sum(dim=0).clamp(max=self.cap)
        ep = [0] + torch.cumsum(final, 0).tolist()  # read-back 1 of 2 per slab: the sub-batch layout is a host decision
        cuts_all = self._sub_batch_cuts(ep, Q)
        sparts = partition_by_edges([ep[c] - ep[a] for a, c in zip(cuts_all[:-1], cuts_all[1:])], W)
        parts = [(cuts_all[s0], cuts_all[s1]) for s0, s1 in sparts]
        s0, s1 = sparts[self.rank]
        q0, q1 = parts[self.rank]
        nq = q1 - q0
        # split sizes: what I send to every owner, what every shard sends me
        csum = torch.zeros(Q + 1, dtype=torch.long, device=dev)
        csum[1:] = torch.cumsum(counts, 0)
        bounds = torch.tensor([p[0] for p in parts] + [parts[-1][1]], dtype=torch.long, device=dev)
        sizes = torch.cat([csum[bounds], cnt_all[:, q0:q1].sum(dim=1)]).tolist()  # read-back 2 of 2
        send = [sizes[i + 1] - sizes[i] for i in range(W)]
        recv = [int(x) for x in sizes[W + 1:]]
        got = {}
        for name in ("key", "ht", "rel"):
            out = torch.empty(max(sum(recv), 1), dtype=flat[name].dtype, device=dev)
            dist.all_to_all_single(out[: sum(recv)], flat[name].contiguous(), output_split_sizes=recv,
                                   input_split_sizes=send, group=self.group)
            got[name] = out
        offs = [0]
        for x in recv:
            offs.append(offs[-1] + x)
        flats = [{name: got[name][offs[r]: offs[r + 1]] for name in got} for r in range(W)]
        b = types.SimpleNamespace()
        if nq > 0:
            cat, seg_off = self._union(flats, cnt_all[:, q0:q1].contiguous(), nq)
            most = int(final[q0:q1].max().item()) if nq else 1
            stride = min(self.cap, _next_pow2(max(most, 1)))
            merged = self._select(cat["key"], cat["ht"], cat["rel"], seg_off, nq, stride)
            b = self._pack(st, merged, nq, stride, query_base=q0)
        else:
            z = lambda *shape, dt=torch.long: torch.zeros(shape, dtype=dt, device=dev)
            b.edge_index, b.edge_attr, b.edge_global_id = z(2, 0), z(0), z(0)
            b.node_global_ids, b.node_embedding_ids, b.topic_one_hot = z(0), z(0), z(0, 2, dt=torch.float32)
            b.ptr, b.edge_ptr, b.num_graphs = z(1), z(1), 0
        b.owner_parts, b.owner_range, b.num_queries_total = parts, (q0, q1), Q
        b.owner_cuts = [c - q0 for c in cuts_all[s0:s1 + 1]] if nq > 0 else [0]
        return b

    def _pack(self, st: _ShardState, sel, Q: int, cap: int, query_base: int = 0) -> types.SimpleNamespace:
        L, dev, stream = self.L, self.dev, _lib.stream_handle()
        node_count = torch.empty(Q, dtype=torch.int32, device=dev)
        node_global = torch.empty((Q, 2 * cap), dtype=torch.int32, device=dev)
        node_is_seed = torch.empty((Q, 2 * cap), dtype=torch.uint8, device=dev)
        _lib.check(L.evi_pool_relabel(C.byref(st.table), sel["ht"].data_ptr(), sel["count"].data_ptr(), Q, cap,
                                      int(query_base), node_count.data_ptr(), node_global.data_ptr(), node_is_seed.data_ptr(), stream),
                   "evi_pool_relabel")
        edge_ptr = torch.zeros(Q + 1, dtype=torch.long, device=dev)
        edge_ptr[1:] = torch.cumsum(sel["count"].long(), 0)
        node_ptr = torch.zeros(Q + 1, dtype=torch.long, device=dev)
        node_ptr[1:] = torch.cumsum(node_count.long(), 0)
        sizes = torch.stack([edge_ptr[-1], node_ptr[-1]]).tolist()  # one read-back to size the batch arrays
        E, N = int(sizes[0]), int(sizes[1])
        b = types.SimpleNamespace()
        b.edge_index = torch.zeros((2, max(E, 1)), dtype=torch.long, device=dev)
        b.edge_attr = torch.zeros(max(E, 1), dtype=torch.long, device=dev)
        b.edge_global_id = torch.zeros(max(E, 1), dtype=torch.long, device=dev)
        b.node_global_ids = torch.zeros(max(N, 1), dtype=torch.long, device=dev)
        b.node_embedding_ids = torch.zeros(max(N, 1), dtype=torch.long, device=dev)
        b.topic_one_hot = torch.zeros((max(N, 1), 2), dtype=torch.float32, device=dev)
        _lib.check(L.evi_pool_pack(C.byref(st.table), sel["key"].data_ptr(), sel["ht"].data_ptr(), sel["rel"].data_ptr(),
                                   sel["count"].data_ptr(), Q, cap, int(query_base), edge_ptr.data_ptr(), node_ptr.data_ptr(),
                                   node_global.data_ptr(), node_is_seed.data_ptr(), _lib.ptr(self.node_embedding_ids),
                                   max(E, 1), b.edge_index.data_ptr(), b.edge_attr.data_ptr(), b.edge_global_id.data_ptr(),
                                   b.node_global_ids.data_ptr(), b.node_embedding_ids.data_ptr(),
                                   b.topic_one_hot.data_ptr(), stream), "evi_pool_pack")
        b.edge_index, b.edge_attr, b.edge_global_id = b.edge_index[:, :E], b.edge_attr[:E], b.edge_global_id[:E]
        b.node_global_ids, b.node_embedding_ids, b.topic_one_hot = b.node_global_ids[:N], b.node_embedding_ids[:N], b.topic_one_hot[:N]
        b.ptr, b.edge_ptr, b.num_graphs = node_ptr, edge_ptr, Q
        b.candidate_keys, b.candidate_count = sel["key"], sel["count"]
        return b

    # ------------------------------------------------------------------ step 3: score + top-k
    @staticmethod
    def _slice(b, q0: int, q1: int, question_emb: torch.Tensor, ep, npt) -> types.SimpleNamespace:
        """Questions [q0, q1) of a subgraph batch, re-based to their own batch-global node ids. ``ep`` / ``npt`` are
        host copies of ``b.edge_ptr`` / ``b.ptr``."""
        e0, e1, n0, n1 = ep[q0], ep[q1], npt[q0], npt[q1]
        s = types.SimpleNamespace()
        s.edge_index = (b.edge_index[:, e0:e1] - n0).contiguous()
        s.edge_attr, s.edge_global_id = b.edge_attr[e0:e1].contiguous(), b.edge_global_id[e0:e1]
        s.node_embedding_ids, s.topic_one_hot = b.node_embedding_ids[n0:n1].contiguous(), b.topic_one_hot[n0:n1].contiguous()
        s.ptr, s.edge_ptr = (b.ptr[q0:q1 + 1] - n0).contiguous(), (b.edge_ptr[q0:q1 + 1] - e0).contiguous()
        s.question_emb, s.num_graphs = question_emb[q0:q1].contiguous(), q1 - q0
        return s

    def score(self, b: types.SimpleNamespace, question_emb: torch.Tensor) -> Dict[str, torch.Tensor]:
        with torch.cuda.device(self.dev):  # the C ABI launches on the current device
            return self._score_on_device(b, question_emb)

    def _score_on_device(self, b: types.SimpleNamespace, question_emb: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Bidirectional scores of every candidate and the per-query top-k as global edge ids, in sub-batches of about
        ``score_edges_per_batch`` candidate edges. With one shard per rank ``b`` holds the rank's own query range
        (``_owner_merge_pack``): each rank scores its queries and the ``[Q, k]`` results are gathered on every rank.
        ``question_emb`` always covers the whole batch."""
        dev, k = self.dev, self.top_k
        question_emb = question_emb.to(device=dev, dtype=torch.float32)
        nq = b.num_graphs
        ep, npt = b.edge_ptr.tolist(), b.ptr.tolist()  # one read-back per slab: sub-batch slicing is a host decision
        owner = getattr(b, "owner_range", None)
        if owner is not None:
            q0, cuts = owner[0], b.owner_cuts
            question_emb = question_emb[q0:owner[1]]
        else:
            cuts = self._sub_batch_cuts(ep, nq)
        edge_id = torch.full((nq, k), -1, dtype=torch.long, device=dev)
        score = torch.full((nq, k), float("-inf"), device=dev)
        count = torch.zeros(nq, dtype=torch.int32, device=dev)
        logits = torch.empty(ep[nq], dtype=torch.float32, device=dev)
        ar_k = torch.arange(k, device=dev).unsqueeze(0)
        for a, c in zip(cuts[:-1], cuts[1:]):
            if ep[c] == ep[a]:
                continue  # no candidates in this range
            sb = self._slice(b, a, c, question_emb, ep, npt)
            out = self.model(sb)
            top = segmented_topk(out.logits, sb.edge_ptr, k)
            gidx = (top.indices.long() + sb.edge_ptr[:-1].unsqueeze(1)).clamp(min=0, max=sb.edge_global_id.numel() - 1)
            valid = ar_k < top.counts.unsqueeze(1)
            edge_id[a:c] = torch.where(valid, sb.edge_global_id[gidx], torch.full_like(gidx, -1))
            score[a:c], count[a:c] = top.values, top.counts
            logits[ep[a]:ep[c]] = out.logits
        res = {"edge_id": edge_id, "score": score, "count": count}
        if owner is not None:
            res = gather_topk(res, nq, group=self.group, counts=[q1_ - q0_ for q0_, q1_ in b.owner_parts])
        res["logits_local"] = logits
        return res

    def retrieve(self, question_emb: torch.Tensor, seeds, seed_ptr) -> Dict[str, torch.Tensor]:
        """``edge_id [Q, k]`` (global triple ids, -1 padded), ``score [Q, k]``, ``count [Q]``, plus the candidate
        subgraph batch under ``"batch"``."""
        b = self.extract(seeds, seed_ptr)
        res = self.score(b, question_emb)
        res["batch"] = b
        return res


__all__ = ["TriplePoolShard", "PoolRetriever", "partition_pool_by_head"]
