// Single-graph retrieval over a sharded triple pool (BASELINE config 4).
// The reference only has per-question subgraphs (AGENTS.md:47); this stage produces them on the device from one
// large knowledge graph: per query, a depth-capped multi-source BFS from the seed entities
// (scripts/build_retrieval_pipeline.py:610-631 semantics) and the candidate triples = edges whose two endpoints
// were reached, capped by smallest (dist_h + dist_t, edge id). The extracted subgraphs then go through the same
// scoring + ragged top-k kernels as every other config.
//
// Data layout. A shard owns a subset of the triples, indexed twice by GLOBAL node id: rows by head (fwd) and rows
// by tail (rev), each row a contiguous run of (other endpoint, relation, global edge id), so one warp reads a row
// with coalesced 4-byte loads. The visited set of a query batch is one open-addressing hash table keyed by
// (query << 32 | node) with the hop distance packed in the same word: HBM-bound integer work, no tensor cores.
//
//   evi_pool_expand      warp per frontier pair: insert every neighbour (both directions) at dist level+1
//   evi_pool_insert      thread per pair: merge the pairs other shards discovered (the frontier exchange)
//   evi_pool_candidates  warp per visited pair: count / emit edges whose other endpoint is visited too
//   evi_pool_select      CTA per query: radix-select the `cap` smallest (priority, edge id) keys, sort by edge id
//   evi_pool_relabel     local node ids per query through the same hash table, then the PyG-style batch arrays
#include "common.cuh"
#include "evi_rag_b200.h"
#include "select.cuh"

namespace evi {

constexpr unsigned long long kPoolEmpty = ~0ull;
constexpr int kPoolEidBits = 40;
constexpr unsigned long long kPoolEidMask = (1ull << kPoolEidBits) - 1ull;

// One 8-byte word per slot: hop distance in the top 6 bits, (query << 32 | node) below, so a lookup is ONE random
// access (a separate distance array costs a second DRAM sector per probe: 146 B per scanned entry measured).
constexpr int kPoolDistShift = 58;
constexpr unsigned long long kPoolKeyMask = (1ull << kPoolDistShift) - 1ull;

struct PoolTable {
  unsigned long long* keys;  // [capacity] dist << 58 | query << 32 | node, kPoolEmpty when free
  int* local_id;             // [capacity] scratch of evi_pool_relabel
  unsigned mask;             // capacity - 1 (capacity is a power of two)
};

EVI_DEVINL unsigned long long pool_mix(unsigned long long x) {  // splitmix64 finaliser
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27; x *= 0x94d049bb133111ebull;
  x ^= x >> 31;
  return x;
}
// 1: inserted, 0: already present, -1: table full
EVI_DEVINL int pool_insert(const PoolTable& t, unsigned long long key, unsigned char d, unsigned* slot_out) {
  unsigned slot = static_cast<unsigned>(pool_mix(key)) & t.mask;
  for (unsigned probe = 0; probe <= t.mask; ++probe) {
    const unsigned long long prev =
        atomicCAS(&t.keys[slot], kPoolEmpty, key | (static_cast<unsigned long long>(d) << kPoolDistShift));
    if (prev == kPoolEmpty) {
      if (slot_out) *slot_out = slot;
      return 1;
    }
    if ((prev & kPoolKeyMask) == key) {
      if (slot_out) *slot_out = slot;
      return 0;
    }
    slot = (slot + 1) & t.mask;
  }
  return -1;
}
// slot of `key` (or -1); *dist receives its hop distance
EVI_DEVINL int pool_find(const PoolTable& t, unsigned long long key, int* dist = nullptr) {
  unsigned slot = static_cast<unsigned>(pool_mix(key)) & t.mask;
  for (unsigned probe = 0; probe <= t.mask; ++probe) {
    const unsigned long long cur = t.keys[slot];
    if (cur == kPoolEmpty) return -1;
    if ((cur & kPoolKeyMask) == key) {
      if (dist) *dist = static_cast<int>(cur >> kPoolDistShift);
      return static_cast<int>(slot);
    }
    slot = (slot + 1) & t.mask;
  }
  return -1;
}

struct PoolShard {
  const int* fwd_ptr;        // [num_nodes + 1] rows by head
  const int* fwd_other;      // tail of each entry
  const int* fwd_rel;
  const long long* fwd_eid;  // global edge id
  const int* rev_ptr;        // [num_nodes + 1] rows by tail
  const int* rev_other;      // head of each entry
};

__global__ void pool_insert_kernel(PoolTable t, const unsigned long long* __restrict__ pairs, long long n,
                                   int dist_value, unsigned long long* __restrict__ visited,
                                   unsigned long long* __restrict__ visited_count, unsigned long long visited_cap,
                                   unsigned* __restrict__ flags) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = pairs[i];
  if (key == kPoolEmpty) return;  // padding of the exchange buffer
  const int r = pool_insert(t, key, static_cast<unsigned char>(dist_value), nullptr);
  if (r < 0) atomicOr(flags, EVI_POOL_TABLE_FULL);
  if (r == 1) {
    const unsigned long long pos = atomicAdd(visited_count, 1ull);
    if (pos < visited_cap) visited[pos] = key; else atomicOr(flags, EVI_POOL_VISITED_FULL);
  }
}

// Long rows. A warp scans a row 32 entries at a time with one hash probe per entry; a hub row of 40k entries is 1300
// dependent iterations (milliseconds), and one such warp sets the duration of the whole launch. Rows longer than
// kPoolSeg entries are therefore cut into segments: the warp that meets the row keeps segment 0 and appends the others
// to an overflow list (evi_pool_overflow) that a second launch spreads over all warps. Segments that do not fit the
// list are scanned by the first warp as before, so the result never depends on the list's capacity.
constexpr int kPoolSeg = 1024;
struct PoolOverflow {
  unsigned long long* items;  // pair index << 17 | side << 16 | segment
  unsigned long long* count;
  unsigned long long capacity;
};
// reserve list slots for segments 1 .. nseg-1 of a long row; returns the first segment the caller must scan itself
EVI_DEVINL int overflow_push(const PoolOverflow& o, unsigned long long pair, int side, int nseg, int lane) {
  if (o.items == nullptr || nseg <= 1 || nseg > 65535) return 1;
  unsigned long long pos = 0;
  if (lane == 0) pos = atomicAdd(o.count, static_cast<unsigned long long>(nseg - 1));
  pos = __shfl_sync(0xffffffffu, pos, 0);
  const unsigned long long room = pos < o.capacity ? o.capacity - pos : 0ull;
  for (int sgm = 1 + lane; sgm < nseg; sgm += 32)
    if (static_cast<unsigned long long>(sgm - 1) < room)
      o.items[pos + sgm - 1] = (pair << 17) | (static_cast<unsigned long long>(side) << 16) | static_cast<unsigned>(sgm);
  const unsigned long long queued = room < static_cast<unsigned long long>(nseg - 1) ? room : nseg - 1;
  return static_cast<int>(queued) + 1;  // segments [queued + 1, nseg) stay with the caller
}

// insert the neighbours listed in entries [i0, i1) of a row of `other` for the query in qbits
EVI_DEVINL void expand_scan(PoolTable t, const int* __restrict__ other, int i0, int i1, unsigned long long qbits, int level,
                            int lane, unsigned long long* __restrict__ visited,
                            unsigned long long* __restrict__ visited_count, unsigned long long visited_cap,
                            unsigned long long* __restrict__ fresh, unsigned long long* __restrict__ fresh_count,
                            unsigned long long fresh_cap, unsigned* __restrict__ flags) {
  for (int i = i0 + lane; i < i1; i += 32) {
    const unsigned long long nk = qbits | static_cast<unsigned>(other[i]);
    const int r = pool_insert(t, nk, static_cast<unsigned char>(level + 1), nullptr);
    if (r < 0) atomicOr(flags, EVI_POOL_TABLE_FULL);
    if (r == 1) {
      const unsigned long long pos = atomicAdd(visited_count, 1ull);
      if (pos < visited_cap) visited[pos] = nk; else atomicOr(flags, EVI_POOL_VISITED_FULL);
      if (fresh != nullptr) {
        const unsigned long long fp = atomicAdd(fresh_count, 1ull);
        if (fp < fresh_cap) fresh[fp] = nk; else atomicOr(flags, EVI_POOL_VISITED_FULL);
      }
    }
  }
}

// frontier = visited[begin, end); discoveries are inserted at level + 1 and appended to visited (own list) and to
// fresh (what the other shards must be told about). OVERFLOW = false: one warp per frontier pair (segment 0 of long
// rows); OVERFLOW = true: one warp per queued segment.
template <bool OVERFLOW>
__global__ void __launch_bounds__(256)
pool_expand_kernel(PoolTable t, PoolShard s, unsigned long long* __restrict__ visited, unsigned long long begin,
                   unsigned long long end, int level, unsigned long long* __restrict__ visited_count,
                   unsigned long long visited_cap, unsigned long long* __restrict__ fresh,
                   unsigned long long* __restrict__ fresh_count, unsigned long long fresh_cap,
                   unsigned* __restrict__ flags, PoolOverflow ovf) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x) >> 5;
  const unsigned long long nwarps = (gridDim.x * static_cast<unsigned long long>(blockDim.x)) >> 5;
  if constexpr (OVERFLOW) {
    const unsigned long long n = *ovf.count < ovf.capacity ? *ovf.count : ovf.capacity;
    for (unsigned long long it = warp; it < n; it += nwarps) {
      const unsigned long long item = ovf.items[it];
      const unsigned long long key = visited[item >> 17];
      const int side = static_cast<int>((item >> 16) & 1ull), sgm = static_cast<int>(item & 0xFFFFull);
      const int u = static_cast<int>(key & 0xFFFFFFFFull);
      const int* ptr = side == 0 ? s.fwd_ptr : s.rev_ptr;
      const int r0 = ptr[u], r1 = ptr[u + 1];
      const int i0 = r0 + sgm * kPoolSeg, i1 = i0 + kPoolSeg < r1 ? i0 + kPoolSeg : r1;
      expand_scan(t, side == 0 ? s.fwd_other : s.rev_other, i0, i1, key & 0xFFFFFFFF00000000ull, level, lane, visited,
                  visited_count, visited_cap, fresh, fresh_count, fresh_cap, flags);
    }
  } else {
  for (unsigned long long p = begin + warp; p < end; p += nwarps) {
    const unsigned long long key = visited[p];
    const unsigned long long qbits = key & 0xFFFFFFFF00000000ull;
    const int u = static_cast<int>(key & 0xFFFFFFFFull);
#pragma unroll
    for (int side = 0; side < 2; ++side) {
      const int* ptr = side == 0 ? s.fwd_ptr : s.rev_ptr;
      const int* other = side == 0 ? s.fwd_other : s.rev_other;
      const int r0 = ptr[u], r1 = ptr[u + 1];
      const int nseg = (r1 - r0 + kPoolSeg - 1) / kPoolSeg;
      const int mine = overflow_push(ovf, p, side, nseg, lane);  // segments [mine, nseg) stay here, like segment 0
      expand_scan(t, other, r0, r0 + kPoolSeg < r1 ? r0 + kPoolSeg : r1, qbits, level, lane, visited, visited_count,
                  visited_cap, fresh, fresh_count, fresh_cap, flags);
      if (mine < nseg)
        expand_scan(t, other, r0 + mine * kPoolSeg, r1, qbits, level, lane, visited, visited_count, visited_cap, fresh,
                    fresh_count, fresh_cap, flags);
    }
  }
  }
}

// Candidates of this shard: every local edge (u -> v) whose endpoints are both visited for the query, found from
// the head side so that each edge is seen once. Priority class = dist_u + dist_v (clamped to kPoolPrioBins - 1).
//   EMIT = false: cand_count[q, class] += 1 (warp-reduced per class);
//   EMIT = true : only classes <= class_limit[q] (the class in which the running count reaches `cap`; everything
//                 above can never survive the cut) are written: key = (dist_u + dist_v) << 40 | edge id, (u, v) and
//                 the relation at cand_off[q] + cursor[q] (one atomicAdd per warp iteration, ballot-ranked lanes).
constexpr int kPoolPrioBins = 8;

// entries [i0, i1) of the head row of u for query q (du = hop distance of u): count or emit the edges whose tail is visited
template <bool EMIT>
EVI_DEVINL void candidates_scan(PoolTable t, PoolShard s, unsigned long long qbits, int q, int u, int du, int limit, int i0,
                                int i1, int lane, int* __restrict__ cand_count, const long long* __restrict__ cand_off,
                                int* __restrict__ cursor, unsigned long long* __restrict__ cand_key,
                                unsigned long long* __restrict__ cand_ht, int* __restrict__ cand_rel) {
  int found[kPoolPrioBins];
#pragma unroll
  for (int b = 0; b < kPoolPrioBins; ++b) found[b] = 0;
  for (int base = i0; base < i1; base += 32) {
    const int i = base + lane;
    int slot = -1, v = 0, dv = 0;
    if (i < i1) {
      v = s.fwd_other[i];
      slot = pool_find(t, qbits | static_cast<unsigned>(v), &dv);
    }
    const int prio = slot >= 0 ? du + dv : 0;
    const int cls = prio < kPoolPrioBins - 1 ? prio : kPoolPrioBins - 1;
    if (!EMIT) {
      if (slot >= 0) {
#pragma unroll
        for (int b = 0; b < kPoolPrioBins; ++b) found[b] += cls == b ? 1 : 0;
      }
    } else {
      const bool take = slot >= 0 && cls <= limit;
      const unsigned m = __ballot_sync(0xffffffffu, take);
      if (m != 0u) {
        long long base_pos = 0;
        if (lane == 0) base_pos = cand_off[q] + atomicAdd(&cursor[q], __popc(m));
        base_pos = __shfl_sync(0xffffffffu, base_pos, 0);
        if (take) {
          const long long pos = base_pos + __popc(m & ((1u << lane) - 1u));
          cand_key[pos] = (static_cast<unsigned long long>(prio) << kPoolEidBits) |
                          static_cast<unsigned long long>(s.fwd_eid[i]);
          cand_ht[pos] = (static_cast<unsigned long long>(static_cast<unsigned>(u)) << 32) | static_cast<unsigned>(v);
          cand_rel[pos] = s.fwd_rel[i];
        }
      }
    }
  }
  if (!EMIT) {
#pragma unroll
    for (int b = 0; b < kPoolPrioBins; ++b) {
      const int c = __reduce_add_sync(0xffffffffu, found[b]);
      if (lane == 0 && c) atomicAdd(&cand_count[q * kPoolPrioBins + b], c);
    }
  }
}

// OVERFLOW = false: one warp per visited pair (segment 0 of long rows); OVERFLOW = true: one warp per queued segment
template <bool EMIT, bool OVERFLOW>
__global__ void __launch_bounds__(256)
pool_candidates_kernel(PoolTable t, PoolShard s, const unsigned long long* __restrict__ visited,
                       unsigned long long num_visited, int* __restrict__ cand_count,
                       const int* __restrict__ class_limit, const long long* __restrict__ cand_off,
                       int* __restrict__ cursor, unsigned long long* __restrict__ cand_key,
                       unsigned long long* __restrict__ cand_ht, int* __restrict__ cand_rel, PoolOverflow ovf) {
  const int lane = threadIdx.x & 31;
  const unsigned long long warp = (blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x) >> 5;
  const unsigned long long nwarps = (gridDim.x * static_cast<unsigned long long>(blockDim.x)) >> 5;
  const unsigned long long n_items =
      OVERFLOW ? (*ovf.count < ovf.capacity ? *ovf.count : ovf.capacity) : num_visited;
  for (unsigned long long it = warp; it < n_items; it += nwarps) {
    const unsigned long long item = OVERFLOW ? ovf.items[it] : 0ull;
    const unsigned long long p = OVERFLOW ? item >> 17 : it;
    const unsigned long long key = visited[p];
    const unsigned long long qbits = key & 0xFFFFFFFF00000000ull;
    const int q = static_cast<int>(key >> 32);
    const int u = static_cast<int>(key & 0xFFFFFFFFull);
    const int r0 = s.fwd_ptr[u], r1 = s.fwd_ptr[u + 1];
    if (r0 == r1) continue;
    int du = 0;
    pool_find(t, key, &du);
    const int limit = EMIT ? class_limit[q] : 0;
    if (EMIT && du > limit) continue;  // dist_v >= 0: every edge of this row is above the limit
    if (OVERFLOW) {
      const int i0 = r0 + static_cast<int>(item & 0xFFFFull) * kPoolSeg;
      candidates_scan<EMIT>(t, s, qbits, q, u, du, limit, i0, i0 + kPoolSeg < r1 ? i0 + kPoolSeg : r1, lane, cand_count,
                            cand_off, cursor, cand_key, cand_ht, cand_rel);
    } else {
      const int nseg = (r1 - r0 + kPoolSeg - 1) / kPoolSeg;
      const int mine = overflow_push(ovf, p, 0, nseg, lane);  // segments [mine, nseg) stay here, like segment 0
      candidates_scan<EMIT>(t, s, qbits, q, u, du, limit, r0, r0 + kPoolSeg < r1 ? r0 + kPoolSeg : r1, lane, cand_count,
                            cand_off, cursor, cand_key, cand_ht, cand_rel);
      if (mine < nseg)
        candidates_scan<EMIT>(t, s, qbits, q, u, du, limit, r0 + mine * kPoolSeg, r1, lane, cand_count, cand_off, cursor,
                              cand_key, cand_ht, cand_rel);
    }
  }
}

// per query: the class in which the running count reaches cap, and how many candidates will be emitted
__global__ void pool_boundary_kernel(const int* __restrict__ cand_count, int num_queries, int cap,
                                     int* __restrict__ class_limit, int* __restrict__ emit_count) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= num_queries) return;
  long long run = 0;
  int limit = kPoolPrioBins - 1;
  for (int b = 0; b < kPoolPrioBins; ++b) {
    run += cand_count[q * kPoolPrioBins + b];
    if (run >= cap) { limit = b; break; }
  }
  class_limit[q] = limit;
  emit_count[q] = static_cast<int>(run);
}

}  // namespace evi

namespace evi {

// ---------------------------------------------------------------------------------------------
// Per query: keep the `cap` smallest keys of segment [seg_off[q], seg_off[q+1]) and return them ordered by edge
// id. Keys are unique inside a segment (one per edge), so the 6-pass LSB..MSB-free radix search below (MSB
// first, 8 bits per pass over the 48 significant bits) yields a threshold with exactly `cap` keys at or under it.
// The survivors are staged unordered, then (edge id << 16 | staging slot) is bitonic-sorted in the query's
// scratch row and the payload is permuted into the outputs. Used twice: on each shard's own candidates, and on
// the concatenation of the shards' survivors (the merge after the exchange).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTopkThreads)
pool_select_kernel(const unsigned long long* __restrict__ cand_key, const unsigned long long* __restrict__ cand_ht,
                   const int* __restrict__ cand_rel, const long long* __restrict__ seg_off, int cap, int sort_len,
                   unsigned long long* __restrict__ stage_key, unsigned long long* __restrict__ stage_ht,
                   int* __restrict__ stage_rel, unsigned long long* __restrict__ sort_ws,
                   unsigned long long* __restrict__ out_key, unsigned long long* __restrict__ out_ht,
                   int* __restrict__ out_rel, int* __restrict__ out_count) {
  __shared__ int hist[256];
  __shared__ int warp_tot[kTopkThreads / 32];
  __shared__ unsigned long long sel_prefix;
  __shared__ int sel_need, cursor;
  const int q = blockIdx.x, tid = threadIdx.x;
  const long long s0 = seg_off[q];
  const long long n = seg_off[q + 1] - s0;
  const unsigned long long* key = cand_key + s0;
  const size_t ob = static_cast<size_t>(q) * cap;
  // all-ones keys are padding (the merge pass concatenates fixed-size rows): count the real ones first
  if (tid == 0) cursor = 0;
  __syncthreads();
  {
    int c = 0;
    for (long long i = tid; i < n; i += kTopkThreads) c += key[i] != kPoolEmpty ? 1 : 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((tid & 31) == 0 && c) atomicAdd(&cursor, c);
  }
  __syncthreads();
  const long long n_valid = cursor;
  const int kt = static_cast<int>(n_valid < cap ? n_valid : cap);
  __syncthreads();
  if (tid == 0) out_count[q] = kt, cursor = 0, sel_prefix = 0ull, sel_need = kt;
  __syncthreads();
  if (kt == 0) return;
  unsigned long long threshold = kPoolEmpty - 1ull;  // everything real
  if (n_valid > cap) {
    unsigned long long mask = 0ull;
    for (int pass = 0; pass < 6; ++pass) {
      const int shift = 40 - 8 * pass;
      hist[tid] = 0;
      __syncthreads();
      const unsigned long long prefix = sel_prefix;
      const int need = sel_need;
      for (long long base = 0; base < n; base += kTopkThreads) {
        const long long i = base + tid;
        unsigned long long k = 0ull;
        const bool in = i < n && (k = key[i]) != kPoolEmpty && (k & mask) == prefix;
        const unsigned active = __ballot_sync(0xffffffffu, in);
        if (in) {
          const int bin = static_cast<int>((k >> shift) & 255ull);
          const unsigned peers = __match_any_sync(active, bin);
          if ((tid & 31) == __ffs(peers) - 1) atomicAdd(&hist[bin], __popc(peers));
        }
      }
      __syncthreads();
      const int own = hist[tid];  // ascending: thread t owns bin t
      int total;
      const int incl = block_inclusive_scan(own, warp_tot, &total);
      if (incl >= need && incl - own < need) {
        sel_prefix = prefix | (static_cast<unsigned long long>(tid) << shift);
        sel_need = need - (incl - own);
      }
      mask |= 255ull << shift;
      __syncthreads();
    }
    threshold = sel_prefix;
  }
  for (long long i = tid; i < n; i += kTopkThreads) {
    const unsigned long long k = key[i];
    if (k <= threshold) {
      const int j = atomicAdd(&cursor, 1);
      stage_key[ob + j] = k;
      stage_ht[ob + j] = cand_ht[s0 + i];
      stage_rel[ob + j] = cand_rel[s0 + i];
    }
  }
  __syncthreads();
  unsigned long long* sb = sort_ws + static_cast<size_t>(q) * sort_len;
  int P = 1;
  while (P < kt) P <<= 1;
  for (int j = tid; j < P; j += kTopkThreads)
    sb[j] = j < kt ? (((stage_key[ob + j] & kPoolEidMask) << 16) | static_cast<unsigned long long>(j)) : ~0ull;
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < P; i += kTopkThreads) {
        const int j = i ^ stride;
        if (j > i) {
          const unsigned long long x = sb[i], y = sb[j];
          const bool ascending = (i & size) == 0;
          if (ascending ? (x > y) : (x < y)) sb[i] = y, sb[j] = x;
        }
      }
      __syncthreads();
    }
  }
  for (int j = tid; j < kt; j += kTopkThreads) {
    const int src = static_cast<int>(sb[j] & 0xFFFFull);
    out_key[ob + j] = stage_key[ob + src];
    out_ht[ob + j] = stage_ht[ob + src];
    out_rel[ob + j] = stage_rel[ob + src];
  }
}

// ---------------------------------------------------------------------------------------------
// Local node ids and the PyG-style batch arrays of the selected subgraphs.
//   mark:   every endpoint of a selected edge flags its table slot (idempotent stores)
//   assign: the first thread to claim a flagged slot takes the next local id of its query and writes the node row
//           (numbering depends on thread order; nothing downstream does: DDE sums are order-independent integers,
//           projections are per row, results are reported by global edge id)
//   pack:   edge_index in batch-global node ids, relation ids, global edge ids
// ---------------------------------------------------------------------------------------------
__global__ void pool_mark_kernel(PoolTable t, const unsigned long long* __restrict__ sel_ht,
                                 const int* __restrict__ sel_count, int cap, long long total, int query_base) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int q = static_cast<int>(i / cap), j = static_cast<int>(i % cap);
  if (j >= sel_count[q]) return;
  // row q of the selection arrays is query (query_base + q) of the batch the visited table was built for
  const unsigned long long ht = sel_ht[i], qbits = static_cast<unsigned long long>(q + query_base) << 32;
  t.local_id[pool_find(t, qbits | (ht >> 32))] = -2;
  t.local_id[pool_find(t, qbits | (ht & 0xFFFFFFFFull))] = -2;
}

__global__ void pool_assign_kernel(PoolTable t, const unsigned long long* __restrict__ sel_ht,
                                   const int* __restrict__ sel_count, int cap, long long total,
                                   int* __restrict__ node_count, int* __restrict__ node_global /*[Q, 2 cap]*/,
                                   unsigned char* __restrict__ node_is_seed, int query_base) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int q = static_cast<int>(i / cap), j = static_cast<int>(i % cap);
  if (j >= sel_count[q]) return;
  const unsigned long long ht = sel_ht[i], qbits = static_cast<unsigned long long>(q + query_base) << 32;
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    const unsigned node = static_cast<unsigned>(side == 0 ? (ht >> 32) : (ht & 0xFFFFFFFFull));
    int dn = 0;
    const int slot = pool_find(t, qbits | node, &dn);
    if (atomicCAS(&t.local_id[slot], -2, -3) == -2) {
      const int id = atomicAdd(&node_count[q], 1);
      t.local_id[slot] = id;
      node_global[static_cast<size_t>(q) * 2 * cap + id] = static_cast<int>(node);
      node_is_seed[static_cast<size_t>(q) * 2 * cap + id] = dn == 0 ? 1 : 0;
    }
  }
}

__global__ void pool_pack_kernel(PoolTable t, const unsigned long long* __restrict__ sel_key,
                                 const unsigned long long* __restrict__ sel_ht, const int* __restrict__ sel_rel,
                                 const int* __restrict__ sel_count, int cap, long long total,
                                 const long long* __restrict__ edge_ptr, const long long* __restrict__ node_ptr,
                                 const int* __restrict__ node_global, const unsigned char* __restrict__ node_is_seed,
                                 const long long* __restrict__ node_embedding_table, long long num_edges_out,
                                 long long* __restrict__ edge_index /*[2, E]*/, long long* __restrict__ edge_attr,
                                 long long* __restrict__ edge_global_id, long long* __restrict__ node_global_ids,
                                 long long* __restrict__ node_embedding_ids, float* __restrict__ topic_one_hot,
                                 int query_base) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= total) return;
  const int q = static_cast<int>(i / cap), j = static_cast<int>(i % cap);
  const long long nb = node_ptr[q];
  const int nn = static_cast<int>(node_ptr[q + 1] - nb);
  // nodes of query q: the first 2*cap slots of its row hold nn entries; thread j (and j + cap) copies them
  for (int v = j; v < nn; v += cap) {
    const int g = node_global[static_cast<size_t>(q) * 2 * cap + v];
    node_global_ids[nb + v] = g;
    node_embedding_ids[nb + v] = node_embedding_table ? node_embedding_table[g] : g;
    const bool seed = node_is_seed[static_cast<size_t>(q) * 2 * cap + v] != 0;
    topic_one_hot[2 * (nb + v)] = seed ? 0.f : 1.f;   // one_hot(is_seed): column 1 marks a seed
    topic_one_hot[2 * (nb + v) + 1] = seed ? 1.f : 0.f;
  }
  if (j >= sel_count[q]) return;
  const unsigned long long ht = sel_ht[i], qbits = static_cast<unsigned long long>(q + query_base) << 32;
  const long long e = edge_ptr[q] + j;
  edge_index[e] = nb + t.local_id[pool_find(t, qbits | (ht >> 32))];
  edge_index[num_edges_out + e] = nb + t.local_id[pool_find(t, qbits | (ht & 0xFFFFFFFFull))];
  edge_attr[e] = sel_rel[i];
  edge_global_id[e] = static_cast<long long>(sel_key[i] & kPoolEidMask);
}

}  // namespace evi

using namespace evi;

namespace {

inline bool table_ok(const evi_pool_table* t) {
  return t != nullptr && t->keys != nullptr && t->local_id != nullptr && t->capacity >= 2 &&
         (t->capacity & (t->capacity - 1)) == 0 && t->capacity <= (1ull << 32);
}
inline PoolTable as_table(const evi_pool_table* t) {
  return PoolTable{t->keys, t->local_id, static_cast<unsigned>(t->capacity - 1)};
}
inline bool shard_ok(const evi_pool_shard* s) {
  return s != nullptr && s->fwd_ptr != nullptr && s->rev_ptr != nullptr && s->num_nodes > 0;
}
inline PoolShard as_shard(const evi_pool_shard* s) {
  return PoolShard{s->fwd_ptr, s->fwd_other, s->fwd_rel, s->fwd_eid, s->rev_ptr, s->rev_other};
}
inline int sweep_grid(int num_sms) { return (num_sms > 0 ? num_sms : 148) * 8; }  // 8 CTAs of 8 warps per SM
inline PoolOverflow as_overflow(const evi_pool_overflow* o) {
  if (o == nullptr || o->items == nullptr || o->count == nullptr || o->capacity == 0) return PoolOverflow{nullptr, nullptr, 0};
  return PoolOverflow{o->items, o->count, o->capacity};
}

}  // namespace

extern "C" {

int evi_pool_reset(const evi_pool_table* table, cudaStream_t stream) {
  if (!table_ok(table)) return kErrInvalidArg;
  if (cudaMemsetAsync(table->keys, 0xFF, sizeof(unsigned long long) * table->capacity, stream) != cudaSuccess ||
      cudaMemsetAsync(table->local_id, 0xFF, sizeof(int) * table->capacity, stream) != cudaSuccess)
    return kErrCuda;
  return kOk;
}

int evi_pool_insert(const evi_pool_table* table, const unsigned long long* pairs, long long num_pairs, int dist_value,
                    unsigned long long* visited, unsigned long long* visited_count, unsigned long long visited_cap,
                    unsigned* flags, cudaStream_t stream) {
  if (num_pairs <= 0) return kOk;
  if (!table_ok(table) || pairs == nullptr || visited == nullptr || visited_count == nullptr || flags == nullptr ||
      dist_value < 0 || dist_value > 62)
    return kErrInvalidArg;
  pool_insert_kernel<<<static_cast<int>((num_pairs + 255) / 256), 256, 0, stream>>>(
      as_table(table), pairs, num_pairs, dist_value, visited, visited_count, visited_cap, flags);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_pool_expand(const evi_pool_table* table, const evi_pool_shard* shard, unsigned long long* visited,
                    unsigned long long begin, unsigned long long end, int level, unsigned long long* visited_count,
                    unsigned long long visited_cap, unsigned long long* fresh, unsigned long long* fresh_count,
                    unsigned long long fresh_cap, unsigned* flags, const evi_pool_overflow* overflow, int num_sms,
                    cudaStream_t stream) {
  if (end <= begin) return kOk;
  if (!table_ok(table) || !shard_ok(shard) || visited == nullptr || visited_count == nullptr || flags == nullptr ||
      level < 0 || level > 61 || (fresh != nullptr && fresh_count == nullptr))
    return kErrInvalidArg;
  const unsigned long long warps = end - begin;
  const int grid = static_cast<int>(warps * 32 / 256 + 1 < static_cast<unsigned long long>(sweep_grid(num_sms))
                                        ? warps * 32 / 256 + 1 : sweep_grid(num_sms));
  const PoolOverflow ovf = as_overflow(overflow);
  if (ovf.items != nullptr && cudaMemsetAsync(ovf.count, 0, sizeof(unsigned long long), stream) != cudaSuccess)
    return kErrCuda;
  pool_expand_kernel<false><<<grid, 256, 0, stream>>>(as_table(table), as_shard(shard), visited, begin, end, level,
                                                      visited_count, visited_cap, fresh, fresh_count, fresh_cap, flags,
                                                      ovf);
  if (ovf.items != nullptr)  // the queued segments of long rows, spread over all warps
    pool_expand_kernel<true><<<sweep_grid(num_sms), 256, 0, stream>>>(as_table(table), as_shard(shard), visited, begin,
                                                                     end, level, visited_count, visited_cap, fresh,
                                                                     fresh_count, fresh_cap, flags, ovf);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_pool_candidates(const evi_pool_table* table, const evi_pool_shard* shard, const unsigned long long* visited,
                        unsigned long long num_visited, int emit, int num_queries, int cap, int* cand_count,
                        int* class_limit, int* emit_count, const long long* cand_off, int* cursor,
                        unsigned long long* cand_key, unsigned long long* cand_ht, int* cand_rel,
                        const evi_pool_overflow* overflow, int num_sms, cudaStream_t stream) {
  if (!table_ok(table) || !shard_ok(shard) || visited == nullptr || num_queries <= 0 || cap <= 0 ||
      cand_count == nullptr || class_limit == nullptr || emit_count == nullptr)
    return kErrInvalidArg;
  const int grid = static_cast<int>(num_visited * 32 / 256 + 1 < static_cast<unsigned long long>(sweep_grid(num_sms))
                                        ? num_visited * 32 / 256 + 1 : sweep_grid(num_sms));
  const PoolOverflow ovf = as_overflow(overflow);
  if (ovf.items != nullptr && cudaMemsetAsync(ovf.count, 0, sizeof(unsigned long long), stream) != cudaSuccess)
    return kErrCuda;
  const int ogrid = sweep_grid(num_sms);
  if (!emit) {
    if (cudaMemsetAsync(cand_count, 0, sizeof(int) * num_queries * kPoolPrioBins, stream) != cudaSuccess) return kErrCuda;
    if (num_visited > 0) {
      pool_candidates_kernel<false, false><<<grid, 256, 0, stream>>>(as_table(table), as_shard(shard), visited,
                                                                     num_visited, cand_count, nullptr, nullptr, nullptr,
                                                                     nullptr, nullptr, nullptr, ovf);
      if (ovf.items != nullptr)
        pool_candidates_kernel<false, true><<<ogrid, 256, 0, stream>>>(as_table(table), as_shard(shard), visited,
                                                                       num_visited, cand_count, nullptr, nullptr, nullptr,
                                                                       nullptr, nullptr, nullptr, ovf);
    }
    pool_boundary_kernel<<<(num_queries + 127) / 128, 128, 0, stream>>>(cand_count, num_queries, cap, class_limit,
                                                                       emit_count);
  } else {
    if (cand_off == nullptr || cursor == nullptr || cand_key == nullptr || cand_ht == nullptr || cand_rel == nullptr)
      return kErrInvalidArg;
    if (num_visited > 0) {
      pool_candidates_kernel<true, false><<<grid, 256, 0, stream>>>(as_table(table), as_shard(shard), visited, num_visited,
                                                                    nullptr, class_limit, cand_off, cursor, cand_key,
                                                                    cand_ht, cand_rel, ovf);
      if (ovf.items != nullptr)
        pool_candidates_kernel<true, true><<<ogrid, 256, 0, stream>>>(as_table(table), as_shard(shard), visited,
                                                                      num_visited, nullptr, class_limit, cand_off, cursor,
                                                                      cand_key, cand_ht, cand_rel, ovf);
    }
  }
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_pool_select(const unsigned long long* cand_key, const unsigned long long* cand_ht, const int* cand_rel,
                    const long long* seg_off, int num_queries, int cap, int sort_len, unsigned long long* stage_key,
                    unsigned long long* stage_ht, int* stage_rel, unsigned long long* sort_ws,
                    unsigned long long* out_key, unsigned long long* out_ht, int* out_rel, int* out_count,
                    cudaStream_t stream) {
  if (num_queries <= 0) return kOk;
  if (cap <= 0 || cap > EVI_POOL_MAX_CAP || sort_len < cap || (sort_len & (sort_len - 1)) != 0 || seg_off == nullptr ||
      stage_key == nullptr || stage_ht == nullptr || stage_rel == nullptr || sort_ws == nullptr || out_key == nullptr ||
      out_ht == nullptr || out_rel == nullptr || out_count == nullptr)
    return kErrInvalidArg;
  pool_select_kernel<<<num_queries, kTopkThreads, 0, stream>>>(cand_key, cand_ht, cand_rel, seg_off, cap, sort_len,
                                                              stage_key, stage_ht, stage_rel, sort_ws, out_key, out_ht,
                                                              out_rel, out_count);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_pool_relabel(const evi_pool_table* table, const unsigned long long* sel_ht, const int* sel_count,
                     int num_queries, int cap, int query_base, int* node_count, int* node_global,
                     unsigned char* node_is_seed, cudaStream_t stream) {
  if (num_queries <= 0) return kOk;
  if (!table_ok(table) || sel_ht == nullptr || sel_count == nullptr || cap <= 0 || node_count == nullptr ||
      node_global == nullptr || node_is_seed == nullptr)
    return kErrInvalidArg;
  const long long total = static_cast<long long>(num_queries) * cap;
  const int grid = static_cast<int>((total + 255) / 256);
  if (cudaMemsetAsync(node_count, 0, sizeof(int) * num_queries, stream) != cudaSuccess) return kErrCuda;
  pool_mark_kernel<<<grid, 256, 0, stream>>>(as_table(table), sel_ht, sel_count, cap, total, query_base);
  pool_assign_kernel<<<grid, 256, 0, stream>>>(as_table(table), sel_ht, sel_count, cap, total, node_count, node_global,
                                               node_is_seed, query_base);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_pool_pack(const evi_pool_table* table, const unsigned long long* sel_key, const unsigned long long* sel_ht,
                  const int* sel_rel, const int* sel_count, int num_queries, int cap, int query_base,
                  const long long* edge_ptr,
                  const long long* node_ptr, const int* node_global, const unsigned char* node_is_seed,
                  const long long* node_embedding_table, long long num_edges_out, long long* edge_index,
                  long long* edge_attr, long long* edge_global_id, long long* node_global_ids,
                  long long* node_embedding_ids, float* topic_one_hot, cudaStream_t stream) {
  if (num_queries <= 0) return kOk;
  if (!table_ok(table) || sel_key == nullptr || sel_ht == nullptr || sel_rel == nullptr || sel_count == nullptr ||
      cap <= 0 || edge_ptr == nullptr || node_ptr == nullptr || node_global == nullptr || node_is_seed == nullptr ||
      edge_index == nullptr || edge_attr == nullptr || edge_global_id == nullptr || node_global_ids == nullptr ||
      node_embedding_ids == nullptr || topic_one_hot == nullptr)
    return kErrInvalidArg;
  const long long total = static_cast<long long>(num_queries) * cap;
  pool_pack_kernel<<<static_cast<int>((total + 255) / 256), 256, 0, stream>>>(
      as_table(table), sel_key, sel_ht, sel_rel, sel_count, cap, total, edge_ptr, node_ptr, node_global, node_is_seed,
      node_embedding_table, num_edges_out, edge_index, edge_attr, edge_global_id, node_global_ids, node_embedding_ids,
      topic_one_hot, query_base);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // extern "C"
