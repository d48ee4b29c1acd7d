// Graph kernels of the retrieval hot path (HBM/L2-bound integer work; no tensor cores):
//   evi_edge_batch            compute_edge_batch            src/utils/graph_utils.py:49-104
//   evi_dde_forward           DDE / PEConv mean propagation src/models/components/graph.py:13-74
//   evi_build_csr             adjacency builders            scripts/build_retrieval_pipeline.py:570-603
//   evi_bfs_multi_source      _bfs_dist                     scripts/build_retrieval_pipeline.py:610-631
//   evi_shortest_path_pairs   _select_shortest_edges_*      scripts/build_retrieval_pipeline.py:650-688
// All entry points take raw device pointers + sizes + a stream and return an int status.
#include "common.cuh"
#include "evi_rag_b200.h"

namespace evi {

static inline int check_launch() { return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda; }
static inline int cdiv(long long a, int b) { return static_cast<int>((a + b - 1) / b); }

// ---------------------------------------------------------------------------------------------
// a1: edge -> question id, validation, edge_ptr
// ---------------------------------------------------------------------------------------------
// graph id of node v = #{g : ptr[g+1] <= v}  ==  bucketize(v, ptr[1:], right=True)
EVI_DEVINL int bucketize_right(const long long* __restrict__ ptr1, int n, long long v) {
  int lo = 0, hi = n;  // first index with ptr1[idx] > v
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(ptr1 + mid) <= v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void edge_batch_kernel(const long long* __restrict__ edge_index, long long num_edges,
                                  const long long* __restrict__ ptr, int num_graphs,
                                  long long* __restrict__ edge_batch64, int* __restrict__ qid32,
                                  int* __restrict__ head32, int* __restrict__ tail32,
                                  long long* __restrict__ edge_ptr, unsigned* __restrict__ flags) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const long long total_nodes = __ldg(ptr + num_graphs);
  const long long h = __ldg(edge_index + e);
  const long long t = __ldg(edge_index + num_edges + e);
  unsigned f = 0;
  if (h < 0 || t < 0 || h >= total_nodes || t >= total_nodes) f |= kFlagOutOfRange;
  const int gh = bucketize_right(ptr + 1, num_graphs, h);
  const int gt = bucketize_right(ptr + 1, num_graphs, t);
  if (gh >= num_graphs) f |= kFlagOutOfRange;
  if (gh != gt) f |= kFlagCrossGraph;
  int gprev = -1;
  if (e > 0) {
    gprev = bucketize_right(ptr + 1, num_graphs, __ldg(edge_index + e - 1));
    if (gprev > gh) f |= kFlagNotMonotone;
  }
  if (edge_batch64) edge_batch64[e] = gh;
  if (qid32) qid32[e] = gh;
  if (head32) head32[e] = static_cast<int>(h);
  if (tail32) tail32[e] = static_cast<int>(t);
  // A cross-graph edge still opens its head's graph: only an out-of-range or out-of-order head leaves edge_ptr alone
  // (entries nobody writes keep the host-side prefill num_edges = an empty range; consumers run before the flag
  // word is read back, so every offset must stay inside [0, num_edges]).
  if (edge_ptr != nullptr && (f & (kFlagOutOfRange | kFlagNotMonotone)) == 0) {
    // edges are grouped by graph: edge e opens every graph in (gprev, gh]
    for (int g = gprev + 1; g <= gh; ++g) edge_ptr[g] = e;
    if (e == num_edges - 1)
      for (int g = gh + 1; g <= num_graphs; ++g) edge_ptr[g] = num_edges;
  }
  if (f) atomicOr(flags, f);
}

__global__ void fill_i64_kernel(long long* p, int n, long long v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---------------------------------------------------------------------------------------------
// a3: DDE. Per round: scatter fixed-point contributions along edges (order-independent integer
// atomics -> run-to-run deterministic), then normalise by in-degree. Forward rounds use
// head->tail, reverse rounds tail->head (graph.py:47-51); both advance in the same launch.
// ---------------------------------------------------------------------------------------------
constexpr float kFixScale = 4294967296.0f;  // 2^32

__global__ void dde_scatter_kernel(const int* __restrict__ head, const int* __restrict__ tail,
                                   long long num_edges, int num_nodes, const float* __restrict__ x_fwd,
                                   const float* __restrict__ x_rev, int ld_x_fwd, int ld_x_rev,
                                   int do_fwd, int do_rev, int count_deg,
                                   unsigned long long* __restrict__ acc,  // [N,4]: f0 f1 r0 r1
                                   int* __restrict__ deg) {               // [N,2]: in, out
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const int h = __ldg(head + e), t = __ldg(tail + e);
  if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
      static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes))
    return;  // flagged by evi_edge_batch; never touch memory outside the batch
  if (do_fwd) {
    const float a = __ldg(x_fwd + static_cast<size_t>(h) * ld_x_fwd);
    const float b = __ldg(x_fwd + static_cast<size_t>(h) * ld_x_fwd + 1);
    if (a != 0.f) atomicAdd(acc + 4ull * t + 0, static_cast<unsigned long long>(__float2ll_rn(a * kFixScale)));
    if (b != 0.f) atomicAdd(acc + 4ull * t + 1, static_cast<unsigned long long>(__float2ll_rn(b * kFixScale)));
  }
  if (do_rev) {
    const float a = __ldg(x_rev + static_cast<size_t>(t) * ld_x_rev);
    const float b = __ldg(x_rev + static_cast<size_t>(t) * ld_x_rev + 1);
    if (a != 0.f) atomicAdd(acc + 4ull * h + 2, static_cast<unsigned long long>(__float2ll_rn(a * kFixScale)));
    if (b != 0.f) atomicAdd(acc + 4ull * h + 3, static_cast<unsigned long long>(__float2ll_rn(b * kFixScale)));
  }
  if (count_deg) {
    atomicAdd(deg + 2ull * t + 0, 1);
    atomicAdd(deg + 2ull * h + 1, 1);
  }
}

// node_struct layout [N, 2*L], L = 1 + Rf + Rr: per class c: (x0, f1..fRf, r1..rRr) (retriever.py:551-552)
__global__ void dde_normalise_kernel(int num_nodes, int round, int rf, int rr,
                                     unsigned long long* __restrict__ acc, const int* __restrict__ deg,
                                     float* __restrict__ x_fwd, float* __restrict__ x_rev,
                                     float* __restrict__ node_struct) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= num_nodes) return;
  const int L = 1 + rf + rr;
  const double inv_in = 1.0 / (static_cast<double>(max(deg[2 * n + 0], 1)) * 4294967296.0);
  const double inv_out = 1.0 / (static_cast<double>(max(deg[2 * n + 1], 1)) * 4294967296.0);
  if (round < rf) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float v = static_cast<float>(static_cast<double>(static_cast<long long>(acc[4ull * n + c])) * inv_in);
      x_fwd[2 * n + c] = v;
      node_struct[static_cast<size_t>(n) * 2 * L + c * L + 1 + round] = v;
      acc[4ull * n + c] = 0ull;
    }
  }
  if (round < rr) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float v = static_cast<float>(static_cast<double>(static_cast<long long>(acc[4ull * n + 2 + c])) * inv_out);
      x_rev[2 * n + c] = v;
      node_struct[static_cast<size_t>(n) * 2 * L + c * L + 1 + rf + round] = v;
      acc[4ull * n + 2 + c] = 0ull;
    }
  }
}

__global__ void dde_init_kernel(int num_nodes, const float* __restrict__ topic, int ld_topic, int rf, int rr,
                                float* __restrict__ node_struct) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= num_nodes) return;
  const int L = 1 + rf + rr;
  node_struct[static_cast<size_t>(n) * 2 * L + 0] = topic[static_cast<size_t>(n) * ld_topic + 0];
  node_struct[static_cast<size_t>(n) * 2 * L + L] = topic[static_cast<size_t>(n) * ld_topic + 1];
}

// ---------------------------------------------------------------------------------------------
// CSR adjacency (undirected / forward / reverse), one pass each: degree count, scan, fill
// ---------------------------------------------------------------------------------------------
// mode 0: undirected (both directions, self loop once)  1: head->tail  2: tail->head
__global__ void csr_count_kernel(const int* __restrict__ head, const int* __restrict__ tail, long long num_edges,
                                 int num_nodes, int mode, int* __restrict__ counts) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const int u = __ldg(head + e), v = __ldg(tail + e);
  if (u < 0 || v < 0 || u >= num_nodes || v >= num_nodes) return;  // dropped like the reference (:579-580)
  if (mode == 0) {
    atomicAdd(counts + u, 1);
    if (u != v) atomicAdd(counts + v, 1);
  } else if (mode == 1) {
    atomicAdd(counts + u, 1);
  } else {
    atomicAdd(counts + v, 1);
  }
}

// single-CTA exclusive scan: row_ptr[0..n] from counts[0..n-1]; also resets counts to 0 (fill cursor)
__global__ void csr_scan_kernel(int* __restrict__ counts, int n, int* __restrict__ row_ptr) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const int c = i < n ? counts[i] : 0;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (w == 0) {
      int t = lane < nw ? warp_tot[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += y;
      }
      warp_tot[lane] = t;  // inclusive over warps
    }
    __syncthreads();
    const int warp_off = w == 0 ? 0 : warp_tot[w - 1];
    const int excl = carry + warp_off + incl - c;
    if (i < n) {
      row_ptr[i] = excl;
      counts[i] = 0;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = excl + c;
    __syncthreads();
  }
  if (threadIdx.x == 0) row_ptr[n] = carry;
}

__global__ void csr_fill_kernel(const int* __restrict__ head, const int* __restrict__ tail, long long num_edges,
                                int num_nodes, int mode, const int* __restrict__ row_ptr,
                                int* __restrict__ cursor, int* __restrict__ adj) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const int u = __ldg(head + e), v = __ldg(tail + e);
  if (u < 0 || v < 0 || u >= num_nodes || v >= num_nodes) return;
  if (mode == 0) {
    adj[row_ptr[u] + atomicAdd(cursor + u, 1)] = v;
    if (u != v) adj[row_ptr[v] + atomicAdd(cursor + v, 1)] = u;
  } else if (mode == 1) {
    adj[row_ptr[u] + atomicAdd(cursor + u, 1)] = v;
  } else {
    adj[row_ptr[v] + atomicAdd(cursor + v, 1)] = u;
  }
}

// ---------------------------------------------------------------------------------------------
// a13: multi-source BFS, one CTA per job, frontier-parallel with 8-lane groups per frontier node
// (coalesced adjacency-row reads). A job = (node range of one question, list of sources).
// dist is the output [job][N_job] laid out at job_dist_off; queues ping-pong in workspace.
// ---------------------------------------------------------------------------------------------
constexpr int kBfsThreads = 256;
constexpr int kGroup = 8;

__global__ void __launch_bounds__(kBfsThreads)
bfs_jobs_kernel(const int* __restrict__ row_ptr, const int* __restrict__ adj,
                const int* __restrict__ job_node_begin, const int* __restrict__ job_node_end,
                const long long* __restrict__ job_dist_off, const int* __restrict__ job_src_ptr,
                const int* __restrict__ sources, int sources_are_global, int max_depth,
                int* __restrict__ dist_out, int* __restrict__ queue_ws) {
  const int job = blockIdx.x;
  const int n0 = job_node_begin[job], n1 = job_node_end[job];
  const int nn = n1 - n0;
  int* dist = dist_out + job_dist_off[job];
  int* q_cur = queue_ws + 2 * job_dist_off[job];
  int* q_nxt = q_cur + nn;
  __shared__ int cur_size, nxt_size;
  for (int i = threadIdx.x; i < nn; i += blockDim.x) dist[i] = -1;
  if (threadIdx.x == 0) cur_size = 0, nxt_size = 0;
  __syncthreads();
  // sources: in-range, first occurrence only (build_retrieval_pipeline.py:617-621)
  for (int i = job_src_ptr[job] + threadIdx.x; i < job_src_ptr[job + 1]; i += blockDim.x) {
    const int s = sources[i] - (sources_are_global ? n0 : 0);
    if (s >= 0 && s < nn && atomicCAS(dist + s, -1, 0) == -1) q_cur[atomicAdd(&cur_size, 1)] = s;
  }
  __syncthreads();
  const int group = threadIdx.x / kGroup, gl = threadIdx.x % kGroup, ngroups = blockDim.x / kGroup;
  int level = 0;
  while (cur_size > 0 && (max_depth < 0 || level < max_depth)) {
    const int cs = cur_size;
    for (int i = group; i < cs; i += ngroups) {
      const int u = q_cur[i] + n0;
      const int r1 = __ldg(row_ptr + u + 1);
      for (int j = __ldg(row_ptr + u) + gl; j < r1; j += kGroup) {
        const int v = __ldg(adj + j) - n0;
        if (dist[v] == -1 && atomicCAS(dist + v, -1, level + 1) == -1) q_nxt[atomicAdd(&nxt_size, 1)] = v;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) cur_size = nxt_size, nxt_size = 0;
    int* tmp = q_cur; q_cur = q_nxt; q_nxt = tmp;
    ++level;
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// a13, depth-capped flavour: level-synchronous sweeps over the question's edge list (no CSR needed).
// One CTA per job; dist lives in shared memory; each level reads the job's edges once, coalesced
// (8 B/edge/level). Meant for the small hop caps of retrieval (2-hop / 4-hop subgraphs): the cost is
// O(levels * E_q), so uncapped BFS over long-diameter graphs should use the CSR frontier kernel above.
// A node reached at level L+1 is written by every thread that discovers it with the same value, so
// the result does not depend on thread order.
// ---------------------------------------------------------------------------------------------
constexpr int kSweepThreads = 1024;

__global__ void __launch_bounds__(kSweepThreads)
bfs_sweep_kernel(const int* __restrict__ head, const int* __restrict__ tail,
                 const long long* __restrict__ job_edge_begin, const long long* __restrict__ job_edge_end,
                 const int* __restrict__ job_node_begin, const int* __restrict__ job_node_end,
                 const long long* __restrict__ job_dist_off, const int* __restrict__ job_src_ptr,
                 const int* __restrict__ sources, int sources_are_global, int directed, int max_depth,
                 int smem_nodes, int* __restrict__ dist_out) {
  extern __shared__ int sdist[];
  const int job = blockIdx.x;
  const int n0 = job_node_begin[job], nn = job_node_end[job] - n0;
  const long long e0 = job_edge_begin[job], e1 = job_edge_end[job];
  int* gdist = dist_out + job_dist_off[job];
  int* dist = nn <= smem_nodes ? sdist : gdist;  // oversized jobs work in global memory
  for (int i = threadIdx.x; i < nn; i += blockDim.x) dist[i] = -1;
  __syncthreads();
  for (int i = job_src_ptr[job] + threadIdx.x; i < job_src_ptr[job + 1]; i += blockDim.x) {
    const int s = sources[i] - (sources_are_global ? n0 : 0);
    if (s >= 0 && s < nn) dist[s] = 0;
  }
  __syncthreads();
  for (int level = 0; max_depth < 0 || level < max_depth; ++level) {
    int changed = 0;
    for (long long e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
      const int u = __ldg(head + e) - n0, v = __ldg(tail + e) - n0;
      if (static_cast<unsigned>(u) >= static_cast<unsigned>(nn) || static_cast<unsigned>(v) >= static_cast<unsigned>(nn))
        continue;  // dropped like the reference adjacency builders (:579-580)
      const int du = dist[u], dv = dist[v];
      if (du == level && dv == -1) dist[v] = level + 1, changed = 1;
      if (!directed && dv == level && du == -1) dist[u] = level + 1, changed = 1;
    }
    if (!__syncthreads_or(changed)) break;
  }
  if (dist != gdist)
    for (int i = threadIdx.x; i < nn; i += blockDim.x) gdist[i] = dist[i];
}

// ---------------------------------------------------------------------------------------------
// a14: per (start, answer) pair: edges on some shortest path. One CTA per pair.
//   undirected: ds[u]+1+da[v]==dsa  or  ds[v]+1+da[u]==dsa ; directed: first test only.
// Writes the OR-mask over pairs, the per-pair edge membership bytes and counts / lengths.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
shortest_pairs_kernel(const int* __restrict__ head, const int* __restrict__ tail,
                      const int* __restrict__ pair_edge_begin, const int* __restrict__ pair_edge_end,
                      const int* __restrict__ pair_node_begin, const int* __restrict__ pair_num_nodes,
                      const long long* __restrict__ pair_ds_off, const long long* __restrict__ pair_da_off,
                      const int* __restrict__ pair_answer_local, const long long* __restrict__ pair_out_off,
                      const int* __restrict__ dist, int directed, unsigned char* __restrict__ mask,
                      unsigned char* __restrict__ pair_member, int* __restrict__ pair_count,
                      int* __restrict__ pair_len) {
  const int p = blockIdx.x;
  const int* ds = dist + pair_ds_off[p];
  const int* da = dist + pair_da_off[p];
  const int nn = pair_num_nodes[p], n0 = pair_node_begin[p];
  const int dsa = ds[pair_answer_local[p]];
  __shared__ int total;
  if (threadIdx.x == 0) total = 0;
  __syncthreads();
  if (dsa < 0) {  // unreachable pair: skipped by the reference (:733-734)
    if (threadIdx.x == 0) pair_count[p] = -1, pair_len[p] = -1;
    return;
  }
  int local = 0;
  unsigned char* member = pair_member + pair_out_off[p];
  const int e0 = pair_edge_begin[p], e1 = pair_edge_end[p];
  for (int e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
    const int u = __ldg(head + e) - n0, v = __ldg(tail + e) - n0;
    bool keep = false;
    if (u >= 0 && v >= 0 && u < nn && v < nn) {
      const int dsu = ds[u], dav = da[v];
      keep = dsu >= 0 && dav >= 0 && dsu + 1 + dav == dsa;
      if (!directed && !keep) {
        const int dsv = ds[v], dau = da[u];
        keep = dsv >= 0 && dau >= 0 && dsv + 1 + dau == dsa;
      }
    }
    member[e - e0] = keep ? 1 : 0;
    if (keep) mask[e] = 1, ++local;
  }
  local = __reduce_add_sync(0xffffffffu, local);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(&total, local);
  __syncthreads();
  if (threadIdx.x == 0) pair_count[p] = total, pair_len[p] = dsa;
}

// ---------------------------------------------------------------------------------------------
// a16: deterministic single shortest path (_shortest_path_single, build_retrieval_pipeline.py:453-530).
// The reference runs a FIFO BFS over adjacency lists sorted by (neighbour, edge idx) from the ascending
// unique sources, so parent[v] is the earliest-queued neighbour one level up and parent_edge[v] the smallest
// edge idx between them; the queue position of v is the rank of (queue position of parent, v) inside its level.
// Level-synchronous restatement, one CTA per job: a sweep over the job's edges does
// atomicMin(key[v], queue_pos[u] << 32 | edge idx) for frontier u; the new level is then ranked by sorting
// (parent queue position, node id) (bitonic, in the job's scratch). Level 0 ranks are the ordered compaction of
// the source ids. The path is read back from the best target (min dist, then min id) through key[].
// ---------------------------------------------------------------------------------------------
constexpr int kSpsThreads = 256;

__global__ void __launch_bounds__(kSpsThreads)
shortest_path_single_kernel(const int* __restrict__ head, const int* __restrict__ tail,
                            const long long* __restrict__ job_edge_begin, const long long* __restrict__ job_edge_end,
                            const int* __restrict__ job_node_begin, const int* __restrict__ job_node_end,
                            const int* __restrict__ job_src_ptr, const int* __restrict__ sources,
                            const int* __restrict__ job_tgt_ptr, const int* __restrict__ targets, int ids_are_global,
                            const long long* __restrict__ job_sort_off, int* __restrict__ dist_ws,
                            int* __restrict__ order_ws, unsigned long long* __restrict__ key_ws,
                            unsigned long long* __restrict__ sort_ws, int max_len, int* __restrict__ path_len,
                            int* __restrict__ path_edges, int* __restrict__ path_nodes) {
  __shared__ int counter;
  __shared__ int warp_tot[kSpsThreads / 32];
  const int job = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int n0 = job_node_begin[job], nn = job_node_end[job] - n0;
  const long long e0 = job_edge_begin[job];
  const int ne = static_cast<int>(job_edge_end[job] - e0);
  int* d = dist_ws + n0;
  int* ord = order_ws + n0;
  unsigned long long* key = key_ws + n0;
  unsigned long long* sb = sort_ws + job_sort_off[job];
  const int* hq = head + e0;
  const int* tq = tail + e0;
  const int base_id = ids_are_global ? n0 : 0;
  for (int i = tid; i < nn; i += kSpsThreads) d[i] = -1, key[i] = ~0ull;
  __syncthreads();
  for (int i = job_src_ptr[job] + tid; i < job_src_ptr[job + 1]; i += kSpsThreads) {
    const int sidx = sources[i] - base_id;
    if (sidx >= 0 && sidx < nn) d[sidx] = 0;
  }
  __syncthreads();
  // queue positions of level 0 = rank of the source ids (sorted(set(sources)), :481)
  int running = 0;
  for (int b0 = 0; b0 < nn; b0 += kSpsThreads) {
    const int i = b0 + tid;
    const int on = (i < nn && d[i] == 0) ? 1 : 0;
    int incl = on;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    int off = 0, tot = 0;
#pragma unroll
    for (int j = 0; j < kSpsThreads / 32; ++j) {
      if (j < w) off += warp_tot[j];
      tot += warp_tot[j];
    }
    if (on) ord[i] = running + off + incl - 1;
    running += tot;
    __syncthreads();
  }
  const int num_src = running;
  for (int level = 0; num_src > 0; ++level) {
    for (int e = tid; e < ne; e += kSpsThreads) {
      const int u = hq[e] - n0, v = tq[e] - n0;
      if (static_cast<unsigned>(u) >= static_cast<unsigned>(nn) || static_cast<unsigned>(v) >= static_cast<unsigned>(nn) ||
          u == v)
        continue;  // out-of-range endpoints are dropped (:466); a self loop never discovers a node
      const int du = d[u], dv = d[v];
      if (du == level && (dv == -1 || dv == level + 1)) {
        atomicMin(&key[v], (static_cast<unsigned long long>(ord[u]) << 32) | static_cast<unsigned>(e));
        if (dv == -1) d[v] = level + 1;
      }
      if (dv == level && (du == -1 || du == level + 1)) {
        atomicMin(&key[u], (static_cast<unsigned long long>(ord[v]) << 32) | static_cast<unsigned>(e));
        if (du == -1) d[u] = level + 1;
      }
    }
    if (tid == 0) counter = 0;
    __syncthreads();
    for (int i = tid; i < nn; i += kSpsThreads)
      if (d[i] == level + 1) sb[atomicAdd(&counter, 1)] = (key[i] & 0xFFFFFFFF00000000ull) | static_cast<unsigned>(i);
    __syncthreads();
    const int F = counter;
    if (F == 0) break;
    int P = 1;
    while (P < F) P <<= 1;
    for (int i = F + tid; i < P; i += kSpsThreads) sb[i] = ~0ull;
    __syncthreads();
    for (int size = 2; size <= P; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        for (int i = tid; i < P; i += kSpsThreads) {
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
    for (int r = tid; r < F; r += kSpsThreads) ord[static_cast<int>(sb[r] & 0xFFFFFFFFull)] = r;
    __syncthreads();
  }
  if (tid != 0) return;
  int best = -1, best_d = 0;
  if (num_src > 0)
    for (int i = job_tgt_ptr[job]; i < job_tgt_ptr[job + 1]; ++i) {
      const int t = targets[i] - base_id;
      if (t < 0 || t >= nn || d[t] < 0) continue;
      if (best < 0 || d[t] < best_d || (d[t] == best_d && t < best)) best = t, best_d = d[t];
    }
  if (best < 0) {
    path_len[job] = -1;
    return;
  }
  if (best_d > max_len) {
    path_len[job] = -2;
    return;
  }
  int* pe = path_edges + static_cast<size_t>(job) * max_len;
  int* pn = path_nodes + static_cast<size_t>(job) * (max_len + 1);
  int cur = best;
  pn[best_d] = cur;
  for (int i = best_d - 1; i >= 0; --i) {
    const int e = static_cast<int>(key[cur] & 0xFFFFFFFFull);
    const int u = hq[e] - n0, v = tq[e] - n0;
    const int prev = u == cur ? v : u;
    pe[i] = e;
    pn[i] = prev;
    cur = prev;
  }
  path_len[job] = best_d;
}

}  // namespace evi

using namespace evi;

extern "C" {

int evi_edge_batch(const long long* edge_index, long long num_edges, const long long* ptr, int num_graphs,
                   long long* edge_batch64, int* qid32, int* head32, int* tail32, long long* edge_ptr,
                   unsigned* flags, cudaStream_t stream) {
  if (num_graphs <= 0 || num_edges < 0 || ptr == nullptr || flags == nullptr) return kErrInvalidArg;
  if (num_edges == 0) {  // flags is OR-ed into, never cleared: the caller zeroes it
    if (edge_ptr) fill_i64_kernel<<<cdiv(num_graphs + 1, 256), 256, 0, stream>>>(edge_ptr, num_graphs + 1, 0);
    return check_launch();
  }
  if (edge_index == nullptr) return kErrInvalidArg;
  if (edge_ptr) fill_i64_kernel<<<cdiv(num_graphs + 1, 256), 256, 0, stream>>>(edge_ptr, num_graphs + 1, num_edges);
  edge_batch_kernel<<<cdiv(num_edges, 256), 256, 0, stream>>>(edge_index, num_edges, ptr, num_graphs,
                                                              edge_batch64, qid32, head32, tail32,
                                                              edge_ptr, flags);
  return check_launch();
}

size_t evi_dde_workspace_bytes(int num_nodes) {
  // acc [N,4] u64 + deg [N,2] i32 + x_fwd [N,2] f32 + x_rev [N,2] f32
  return static_cast<size_t>(num_nodes) * (32 + 8 + 8 + 8) + 256;
}

int evi_dde_forward(const int* head32, const int* tail32, long long num_edges, int num_nodes,
                    const float* topic_one_hot, int ld_topic, int num_rounds, int num_reverse_rounds,
                    float* node_struct, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  if (num_nodes <= 0) return kOk;
  if (num_rounds < 0 || num_rounds > 4 || num_reverse_rounds < 0 || num_reverse_rounds > 4 || ld_topic < 2)
    return kErrInvalidArg;  // graph.py:33-37
  if (workspace_bytes < evi_dde_workspace_bytes(num_nodes)) return kErrWorkspace;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  unsigned long long* acc = reinterpret_cast<unsigned long long*>(ws);
  int* deg = reinterpret_cast<int*>(ws + static_cast<size_t>(num_nodes) * 32);
  float* x_fwd = reinterpret_cast<float*>(ws + static_cast<size_t>(num_nodes) * 40);
  float* x_rev = reinterpret_cast<float*>(ws + static_cast<size_t>(num_nodes) * 48);
  if (cudaMemsetAsync(ws, 0, static_cast<size_t>(num_nodes) * 40, stream) != cudaSuccess) return kErrCuda;
  dde_init_kernel<<<cdiv(num_nodes, 256), 256, 0, stream>>>(num_nodes, topic_one_hot, ld_topic, num_rounds,
                                                            num_reverse_rounds, node_struct);
  const int rounds = num_rounds > num_reverse_rounds ? num_rounds : num_reverse_rounds;
  for (int r = 0; r < rounds; ++r) {
    if (num_edges > 0) {
      const float* xf = r == 0 ? topic_one_hot : x_fwd;
      const float* xr = r == 0 ? topic_one_hot : x_rev;
      const int ld = r == 0 ? ld_topic : 2;
      dde_scatter_kernel<<<cdiv(num_edges, 256), 256, 0, stream>>>(head32, tail32, num_edges, num_nodes, xf, xr, ld, ld,
                                                                  r < num_rounds, r < num_reverse_rounds,
                                                                  r == 0, acc, deg);
    }
    dde_normalise_kernel<<<cdiv(num_nodes, 256), 256, 0, stream>>>(num_nodes, r, num_rounds, num_reverse_rounds,
                                                                  acc, deg, x_fwd, x_rev, node_struct);
  }
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// evi_dde_forward for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the same kernels with the propagation scratch (accumulators, degrees, per-round states) in a
// file-scope device array instead of a caller-supplied workspace. No counterpart in the
// reference.
//   * the array is sized for EVI_DDE_STATIC_MAX_NODES nodes (56 bytes per node + 256 = 14 MiB); every GPU that
//     loads the library holds it until the process exits;
//   * in host code `g_dde_static_scratch` is a host-side stand-in, not a device address: the address the kernels get
//     comes from cudaGetSymbolAddress, taken once per device with that device current;
//   * every call shares the array, which keeps the previous call's contents: the accumulators and degrees are cleared
//     on the stream at the start of each call (the per-round states are written before they are read); two calls that
//     overlap on one GPU (different streams) race on it.
// ---------------------------------------------------------------------------------------------
constexpr size_t kDdeStaticBytes = static_cast<size_t>(EVI_DDE_STATIC_MAX_NODES) * 56 + 256;
static __device__ __align__(256) unsigned char g_dde_static_scratch[kDdeStaticBytes];
constexpr int kMaxDevices = 64;
static std::atomic<void*> g_dde_static_addr[kMaxDevices];

int evi_dde_static_max_nodes(void) { return EVI_DDE_STATIC_MAX_NODES; }

int evi_dde_forward_static(const int* head32, const int* tail32, long long num_edges, int num_nodes,
                           const float* topic_one_hot, int ld_topic, int num_rounds, int num_reverse_rounds,
                           float* node_struct, cudaStream_t stream) {
  if (num_nodes <= 0) return kOk;
  if (num_nodes > EVI_DDE_STATIC_MAX_NODES) return kErrStaticLimit;  // beyond the static scratch: use evi_dde_forward
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return kErrCuda;
  if (dev < 0 || dev >= kMaxDevices) return kErrUnsupported;
  void* scratch = g_dde_static_addr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {  // once per device, with that device current (concurrent first calls get the same address)
    if (cudaGetSymbolAddress(&scratch, g_dde_static_scratch) != cudaSuccess || scratch == nullptr) return kErrCuda;
    g_dde_static_addr[dev].store(scratch, std::memory_order_release);
  }
  // evi_dde_forward clears acc / deg with cudaMemsetAsync on `stream` and allocates nothing
  return evi_dde_forward(head32, tail32, num_edges, num_nodes, topic_one_hot, ld_topic, num_rounds, num_reverse_rounds,
                         node_struct, scratch, kDdeStaticBytes, stream);
}

int evi_build_csr(const int* head32, const int* tail32, long long num_edges, int num_nodes, int mode,
                  int* row_ptr, int* adj, int* scratch_counts, cudaStream_t stream) {
  if (num_nodes < 0 || mode < 0 || mode > 2) return kErrInvalidArg;
  if (cudaMemsetAsync(scratch_counts, 0, sizeof(int) * static_cast<size_t>(num_nodes > 0 ? num_nodes : 1), stream) !=
      cudaSuccess)
    return kErrCuda;
  if (num_edges > 0)
    csr_count_kernel<<<cdiv(num_edges, 256), 256, 0, stream>>>(head32, tail32, num_edges, num_nodes, mode,
                                                               scratch_counts);
  csr_scan_kernel<<<1, 1024, 0, stream>>>(scratch_counts, num_nodes, row_ptr);
  if (num_edges > 0)
    csr_fill_kernel<<<cdiv(num_edges, 256), 256, 0, stream>>>(head32, tail32, num_edges, num_nodes, mode, row_ptr,
                                                              scratch_counts, adj);
  return check_launch();
}

int evi_bfs_multi_source(const int* row_ptr, const int* adj, int num_jobs, const int* job_node_begin,
                         const int* job_node_end, const long long* job_dist_off, const int* job_src_ptr,
                         const int* sources, int sources_are_global, int max_depth, int* dist_out,
                         int* queue_workspace, cudaStream_t stream) {
  if (num_jobs <= 0) return kOk;
  bfs_jobs_kernel<<<num_jobs, kBfsThreads, 0, stream>>>(row_ptr, adj, job_node_begin, job_node_end, job_dist_off,
                                                        job_src_ptr, sources, sources_are_global, max_depth,
                                                        dist_out, queue_workspace);
  return check_launch();
}

int evi_bfs_edge_sweep(const int* head32, const int* tail32, int num_jobs, const long long* job_edge_begin,
                       const long long* job_edge_end, const int* job_node_begin, const int* job_node_end,
                       const long long* job_dist_off, const int* job_src_ptr, const int* sources,
                       int sources_are_global, int directed, int max_depth, int* dist_out, cudaStream_t stream) {
  if (num_jobs <= 0) return kOk;
  constexpr int kSmemNodes = 40 * 1024;  // 160 KB of dist per CTA
  static std::atomic<unsigned long long> smem_ready{0};
  if (!ensure_dynamic_smem(bfs_sweep_kernel, kSmemNodes * 4, smem_ready)) return kErrCuda;
  bfs_sweep_kernel<<<num_jobs, kSweepThreads, kSmemNodes * 4, stream>>>(
      head32, tail32, job_edge_begin, job_edge_end, job_node_begin, job_node_end, job_dist_off, job_src_ptr, sources,
      sources_are_global, directed, max_depth, kSmemNodes, dist_out);
  return check_launch();
}

int evi_shortest_path_pairs(const int* head32, const int* tail32, int num_pairs, const int* pair_edge_begin,
                            const int* pair_edge_end, const int* pair_node_begin, const int* pair_num_nodes,
                            const long long* pair_ds_off, const long long* pair_da_off,
                            const int* pair_answer_local, const long long* pair_out_off, const int* dist,
                            int directed, unsigned char* mask, unsigned char* pair_member, int* pair_count,
                            int* pair_len, cudaStream_t stream) {
  if (num_pairs <= 0) return kOk;
  shortest_pairs_kernel<<<num_pairs, 256, 0, stream>>>(head32, tail32, pair_edge_begin, pair_edge_end,
                                                       pair_node_begin, pair_num_nodes, pair_ds_off, pair_da_off,
                                                       pair_answer_local, pair_out_off, dist, directed, mask,
                                                       pair_member, pair_count, pair_len);
  return check_launch();
}

int evi_shortest_path_single(const int* head32, const int* tail32, int num_jobs, const long long* job_edge_begin,
                             const long long* job_edge_end, const int* job_node_begin, const int* job_node_end,
                             const int* job_src_ptr, const int* sources, const int* job_tgt_ptr, const int* targets,
                             int ids_are_global, const long long* job_sort_off, int* dist_ws, int* order_ws,
                             unsigned long long* key_ws, unsigned long long* sort_ws, int max_len, int* path_len,
                             int* path_edges, int* path_nodes, cudaStream_t stream) {
  if (num_jobs <= 0) return kOk;
  if (max_len < 0 || job_edge_begin == nullptr || job_node_begin == nullptr || job_src_ptr == nullptr ||
      job_tgt_ptr == nullptr || job_sort_off == nullptr || dist_ws == nullptr || order_ws == nullptr ||
      key_ws == nullptr || sort_ws == nullptr || path_len == nullptr || path_edges == nullptr || path_nodes == nullptr)
    return kErrInvalidArg;
  shortest_path_single_kernel<<<num_jobs, kSpsThreads, 0, stream>>>(
      head32, tail32, job_edge_begin, job_edge_end, job_node_begin, job_node_end, job_src_ptr, sources, job_tgt_ptr,
      targets, ids_are_global, job_sort_off, dist_ws, order_ws, key_ws, sort_ws, max_len, path_len, path_edges,
      path_nodes);
  return check_launch();
}

}  // extern "C"
