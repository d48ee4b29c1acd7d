// g_agent selection on the device (src/data/components/g_agent_builder.py).
// The reference slices the batch into questions and, per question on the CPU: node-softmax scores, a stable
// top-k, per-start-node top incident edges, a Python-dict (h, r, t) dedup, re-indexing through Python dicts.
// Here:
//   evi_node_softmax_logit  whole batch at once (node ids are batch-global, so no segmentation is needed):
//                           atomicMax on order-preserving keys, then order-independent 2^-40 fixed-point sums
//   evi_gagent_start_edges  one CTA per question: start nodes by id membership, per start node the k_s best
//                           incident entries by a radix threshold over the virtual (heads | tails) list
//   evi_gagent_build        one CTA per question: union with the top-k, ordered compaction, first-seen dedup
//                           with max aggregation, sorted unique entity ids, locals by binary search,
//                           re-normalised scores (sums in ascending edge order like the CPU scatter_add)
// All integer outputs are bit-exact against the reference; scores differ by expf/logf rounding only.
#include <math.h>

#include "common.cuh"
#include "evi_rag_b200.h"
#include "select.cuh"

namespace evi {

constexpr float kProbEps = 1e-6f;                 // g_agent_builder.py:20
constexpr float kFixScale = 1099511627776.0f;     // 2^40
constexpr double kFixInv = 1.0 / 1099511627776.0;
constexpr int kMaxStartNodes = 64;

EVI_DEVINL float key_to_float(uint32_t k) {
  const uint32_t b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
  return __uint_as_float(b);
}
EVI_DEVINL float prob_to_logit(float prob) {
  prob = fminf(fmaxf(prob, kProbEps), 1.0f - kProbEps);
  return logf(prob) - log1pf(-prob);
}

struct NodeSoftmaxWs {  // per node, zero-initialised
  unsigned long long sum_head, sum_tail;
  uint32_t max_head, max_tail;
};

__global__ void node_softmax_max_kernel(const float* __restrict__ scores, const int* __restrict__ head,
                                        const int* __restrict__ tail, long long num_edges, float temperature,
                                        float bias, int calibrate, NodeSoftmaxWs* __restrict__ ws,
                                        float* __restrict__ calibrated) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  float s = scores[e];
  if (calibrate) s = s / temperature + bias;  // g_agent_builder.py:164-165
  if (calibrated) calibrated[e] = s;
  if (ws == nullptr) return;
  const uint32_t k = order_key(s);
  atomicMax(&ws[head[e]].max_head, k);
  atomicMax(&ws[tail[e]].max_tail, k);
}

__global__ void node_softmax_sum_kernel(const float* __restrict__ cal, const int* __restrict__ head,
                                        const int* __restrict__ tail, long long num_edges,
                                        NodeSoftmaxWs* __restrict__ ws) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const float s = cal[e];
  const int h = head[e], t = tail[e];
  const float eh = expf(s - key_to_float(ws[h].max_head));
  const float et = expf(s - key_to_float(ws[t].max_tail));
  atomicAdd(&ws[h].sum_head, __float2ull_rn(eh * kFixScale));
  atomicAdd(&ws[t].sum_tail, __float2ull_rn(et * kFixScale));
}

__global__ void node_softmax_final_kernel(const float* __restrict__ cal, const int* __restrict__ head,
                                          const int* __restrict__ tail, long long num_edges,
                                          const NodeSoftmaxWs* __restrict__ ws, float* __restrict__ out) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const float s = cal[e];
  const NodeSoftmaxWs wh = ws[head[e]], wt = ws[tail[e]];
  const float eh = expf(s - key_to_float(wh.max_head));
  const float et = expf(s - key_to_float(wt.max_tail));
  const float sh = static_cast<float>(static_cast<double>(wh.sum_head) * kFixInv);
  const float st = static_cast<float>(static_cast<double>(wt.sum_tail) * kFixInv);
  const float prob = (eh / fmaxf(sh, kProbEps) + et / fmaxf(st, kProbEps)) * 0.5f;
  out[e] = prob_to_logit(prob);
}

// ---------------------------------------------------------------------------------------------
// start edges (g_agent_builder.py:654-724)
// ---------------------------------------------------------------------------------------------
EVI_DEVINL int start_budget(int deg, float ratio, int min_edges, int max_edges) {
  int k = static_cast<int>(ceilf(static_cast<float>(deg) * ratio));
  if (min_edges > 0) k = max(k, min_edges);
  if (max_edges >= 0) k = max_edges == 0 ? 0 : min(k, max_edges);
  return min(k, deg);
}

__global__ void __launch_bounds__(kTopkThreads)
gagent_start_edges_kernel(const float* __restrict__ sel_scores, const int* __restrict__ head,
                          const int* __restrict__ tail, const long long* __restrict__ edge_ptr,
                          const long long* __restrict__ node_ptr, const long long* __restrict__ node_global_ids,
                          const long long* __restrict__ start_entity_ids, const long long* __restrict__ start_ptr,
                          float keep_ratio, int min_edges, int max_edges, unsigned char* __restrict__ mark,
                          int* __restrict__ status) {
  __shared__ RadixScratch sc;
  __shared__ int start_nodes[kMaxStartNodes];
  __shared__ int n_start, deg_sh;
  const int b = blockIdx.x, tid = threadIdx.x;
  const long long e0 = edge_ptr[b];
  const int n = static_cast<int>(edge_ptr[b + 1] - e0);
  const long long n0 = node_ptr[b];
  const int nn = static_cast<int>(node_ptr[b + 1] - n0);
  const long long s0 = start_ptr[b], s1 = start_ptr[b + 1];
  if (tid == 0) n_start = 0;
  for (int i = tid; i < n; i += kTopkThreads) mark[e0 + i] = 0;
  __syncthreads();
  // start_mask = isin(node_global_ids, start_entity_ids) (:283-286); order of the start nodes is irrelevant here
  for (int i = tid; i < nn; i += kTopkThreads) {
    const long long g = node_global_ids[n0 + i];
    bool hit = false;
    for (long long j = s0; j < s1 && !hit; ++j) hit = start_entity_ids[j] == g;
    if (hit) {
      const int slot = atomicAdd(&n_start, 1);
      if (slot < kMaxStartNodes) start_nodes[slot] = i;
    }
  }
  __syncthreads();
  const int ns = n_start;
  if (tid == 0) {
    int st = 0;
    if (n > 0 && ns == 0) st |= EVI_GAGENT_NO_START_NODE;
    if (ns > kMaxStartNodes) st |= EVI_GAGENT_OVERFLOW;
    status[b] = st;
  }
  if (n == 0 || ns == 0 || ns > kMaxStartNodes) return;
  const float* s = sel_scores + e0;
  const int* hq = head + e0;
  const int* tq = tail + e0;
  for (int si = 0; si < ns; ++si) {
    const int node = static_cast<int>(n0) + start_nodes[si];
    if (tid == 0) deg_sh = 0;
    __syncthreads();
    int d = 0;
    for (int i = tid; i < n; i += kTopkThreads) d += (hq[i] == node ? 1 : 0) + (tq[i] == node ? 1 : 0);
    d = __reduce_add_sync(0xffffffffu, d);
    if ((tid & 31) == 0 && d) atomicAdd(&deg_sh, d);
    __syncthreads();
    const int deg = deg_sh;
    const int ks = start_budget(deg, keep_ratio, min_edges, max_edges);
    __syncthreads();
    if (ks <= 0) continue;
    if (ks == deg) {
      for (int i = tid; i < n; i += kTopkThreads)
        if (hq[i] == node || tq[i] == node) mark[e0 + i] = 1;
      continue;
    }
    // incident list = (edges with head == node, edge order) then (edges with tail == node): virtual index v
    auto key_at = [&](int v, uint32_t* key) {
      const int e = v < n ? v : v - n;
      const bool in = (v < n ? hq[e] : tq[e]) == node;
      if (in) *key = order_key(s[e]);
      return in;
    };
    uint32_t kth;
    int need;
    block_radix_threshold(2 * n, ks, key_at, &sc, &kth, &need);
    int running = 0;
    for (int base = 0; base < 2 * n; base += kTopkThreads) {
      const int v = base + tid;
      uint32_t key = 0u;
      const bool in = v < 2 * n && key_at(v, &key);
      const int e = v < n ? v : v - n;
      if (in && key > kth) mark[e0 + e] = 1;
      if (running < need) {  // uniform across the CTA
        const bool tie = in && key == kth;
        int total;
        const int incl = block_inclusive_scan(tie ? 1 : 0, sc.warp_tot, &total);
        if (tie && running + incl - 1 < need) mark[e0 + e] = 1;
        running += total;
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// union, dedup, re-index (g_agent_builder.py:318-470)
// ---------------------------------------------------------------------------------------------
struct GAgentSmem {
  long long* gh;       // [cap] head entity id
  long long* gt;       // [cap]
  long long* rel;      // [cap]
  long long* nkey;     // [2 cap pow2] node sort keys, then the unique ids
  int* npay;           // [2 cap pow2] batch node index
  int* env;            // [cap] question-local edge id
  int* hl;             // [cap]
  int* tl;             // [cap]
  float* raw;          // [cap]
  float* lab;          // [cap]
  float* tmp;          // [cap]
  unsigned char* keep; // [cap]
};

EVI_DEVINL int lower_bound_ll(const long long* a, int n, long long x) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(kTopkThreads)
gagent_build_kernel(evi_gagent_args a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int warp_tot[kTopkThreads / 32];
  const int cap = a.cap;
  int P2 = 1;
  while (P2 < 2 * cap) P2 <<= 1;
  GAgentSmem S;
  {
    unsigned char* p = smem_raw;
    S.gh = reinterpret_cast<long long*>(p); p += sizeof(long long) * cap;
    S.gt = reinterpret_cast<long long*>(p); p += sizeof(long long) * cap;
    S.rel = reinterpret_cast<long long*>(p); p += sizeof(long long) * cap;
    S.nkey = reinterpret_cast<long long*>(p); p += sizeof(long long) * P2;
    S.npay = reinterpret_cast<int*>(p); p += sizeof(int) * P2;
    S.env = reinterpret_cast<int*>(p); p += sizeof(int) * cap;
    S.hl = reinterpret_cast<int*>(p); p += sizeof(int) * cap;
    S.tl = reinterpret_cast<int*>(p); p += sizeof(int) * cap;
    S.raw = reinterpret_cast<float*>(p); p += sizeof(float) * cap;
    S.lab = reinterpret_cast<float*>(p); p += sizeof(float) * cap;
    S.tmp = reinterpret_cast<float*>(p); p += sizeof(float) * cap;
    S.keep = p;
  }
  const int b = blockIdx.x, tid = threadIdx.x;
  const long long e0 = a.edge_ptr[b];
  const int n = static_cast<int>(a.edge_ptr[b + 1] - e0);
  unsigned char* mark = a.mark + e0;
  int status = a.status[b];
  const size_t ob = static_cast<size_t>(b) * cap;       // edge-array row
  const size_t nb = static_cast<size_t>(b) * 2 * cap;   // node-array row

  auto finish = [&](int st, int n_tri, int n_nodes, int n_sl, int n_au, int n_al) {
    if (tid == 0) {
      a.status[b] = st;
      a.num_triples[b] = n_tri;
      a.num_nodes_out[b] = n_nodes;
      a.num_start_locals[b] = n_sl;
      a.num_answers[b] = n_au;
      a.num_answer_locals[b] = n_al;
    }
  };
  if (n == 0 || (status & (EVI_GAGENT_NO_START_NODE | EVI_GAGENT_OVERFLOW))) {
    finish(status | (n == 0 ? EVI_GAGENT_EMPTY : 0), 0, 0, 0, 0, 0);
    return;
  }
  // A. union with the global top-k (:298-316)
  const int kt = a.topk_counts[b];
  for (int j = tid; j < kt; j += kTopkThreads) mark[a.topk_indices[static_cast<size_t>(b) * a.topk_stride + j]] = 1;
  __syncthreads();
  // ordered compaction = torch.unique(cat(...), sorted=True)
  int running = 0;
  for (int base = 0; base < n; base += kTopkThreads) {
    const int i = base + tid;
    const bool on = i < n && mark[i] != 0;
    int total;
    const int incl = block_inclusive_scan(on ? 1 : 0, warp_tot, &total);
    const int pos = running + incl - 1;
    if (on && pos < cap) S.env[pos] = i;
    running += total;
  }
  if (running > cap) {
    finish(status | EVI_GAGENT_OVERFLOW, 0, 0, 0, 0, 0);
    return;
  }
  const int ne = running;
  __syncthreads();
  // C. (h, r, t) dedup in first-seen order, max score / label (:318-354)
  for (int i = tid; i < ne; i += kTopkThreads) {
    const long long e = e0 + S.env[i];
    S.hl[i] = a.head32[e];
    S.tl[i] = a.tail32[e];
    S.gh[i] = a.node_global_ids[a.head32[e]];
    S.gt[i] = a.node_global_ids[a.tail32[e]];
    S.rel[i] = a.edge_attr[e];
    S.raw[i] = a.raw_scores[e];
    S.lab[i] = a.labels ? a.labels[e] : 0.f;
  }
  __syncthreads();
  for (int i = tid; i < ne; i += kTopkThreads) {
    const long long h = S.gh[i], t = S.gt[i], r = S.rel[i];
    bool first = true;
    for (int j = 0; j < i; ++j)
      if (S.gh[j] == h && S.gt[j] == t && S.rel[j] == r) { first = false; break; }
    float sc = S.raw[i], lb = S.lab[i];
    if (first)
      for (int j = i + 1; j < ne; ++j)
        if (S.gh[j] == h && S.gt[j] == t && S.rel[j] == r) sc = fmaxf(sc, S.raw[j]), lb = fmaxf(lb, S.lab[j]);
    S.keep[i] = first ? 1 : 0;
    S.tmp[i] = sc;
    // label max goes through registers: lab[i] is still read by other threads, so stage it in the env slot's
    // spare array after the barrier below
    a.edge_labels[ob + i] = lb;  // staged at the uncompacted position; compacted below
  }
  __syncthreads();
  // compact kept triples (order preserved). Positions only move down, so in-place is safe chunk by chunk
  // after reading into registers.
  running = 0;
  for (int base = 0; base < ne; base += kTopkThreads) {
    const int i = base + tid;
    const bool on = i < ne && S.keep[i] != 0;
    long long h = 0, t = 0, r = 0;
    int hl = 0, tl = 0, ev = 0;
    float sc = 0.f, lb = 0.f;
    if (on) h = S.gh[i], t = S.gt[i], r = S.rel[i], hl = S.hl[i], tl = S.tl[i], ev = S.env[i], sc = S.tmp[i],
            lb = a.edge_labels[ob + i];
    int total;
    const int incl = block_inclusive_scan(on ? 1 : 0, warp_tot, &total);  // barriers inside: reads precede writes
    const int pos = running + incl - 1;
    if (on) S.gh[pos] = h, S.gt[pos] = t, S.rel[pos] = r, S.hl[pos] = hl, S.tl[pos] = tl, S.env[pos] = ev,
            S.raw[pos] = sc, S.lab[pos] = lb;
    running += total;
    __syncthreads();
  }
  const int m = running;
  // sorted unique entity ids (:364) with the batch node index as payload (for node_embedding_ids, :366-375)
  int P = 1;
  while (P < 2 * m) P <<= 1;
  for (int i = tid; i < P; i += kTopkThreads) {
    long long key = 0x7FFFFFFFFFFFFFFFll;
    int pay = -1;
    if (i < m) key = S.gh[i], pay = S.hl[i];
    else if (i < 2 * m) key = S.gt[i - m], pay = S.tl[i - m];
    S.nkey[i] = key;
    S.npay[i] = pay;
  }
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < P; i += kTopkThreads) {
        const int j = i ^ stride;
        if (j > i) {
          const long long x = S.nkey[i], y = S.nkey[j];
          const bool ascending = (i & size) == 0;
          if (ascending ? (x > y) : (x < y)) {
            S.nkey[i] = y, S.nkey[j] = x;
            const int px = S.npay[i];
            S.npay[i] = S.npay[j], S.npay[j] = px;
          }
        }
      }
      __syncthreads();
    }
  }
  running = 0;
  for (int base = 0; base < 2 * m; base += kTopkThreads) {
    const int i = base + tid;
    const bool on = i < 2 * m && (i == 0 || S.nkey[i] != S.nkey[i - 1]);
    const long long key = on ? S.nkey[i] : 0;
    const int pay = on ? S.npay[i] : 0;
    int total;
    const int incl = block_inclusive_scan(on ? 1 : 0, warp_tot, &total);
    const int pos = running + incl - 1;
    if (on) S.nkey[pos] = key, S.npay[pos] = pay;
    running += total;
    __syncthreads();
  }
  const int nu = running;
  // embedding_lookup is a dict filled over all nodes of the question (:366-367): on a repeated global id the
  // last node wins, so the payload of the sort is only a hint and the question's node range is scanned
  {
    const long long n0 = a.node_ptr[b];
    const int nn = static_cast<int>(a.node_ptr[b + 1] - n0);
    for (int i = tid; i < nu; i += kTopkThreads) {
      const long long key = S.nkey[i];
      int last = S.npay[i];
      for (int j = nn - 1; j >= 0; --j)
        if (a.node_global_ids[n0 + j] == key) { last = static_cast<int>(n0) + j; break; }
      a.node_entity_ids[nb + i] = key;
      a.node_embedding_ids_out[nb + i] = a.node_embedding_ids[last];
      a.node_batch_index[nb + i] = last;
    }
  }
  for (int i = tid; i < m; i += kTopkThreads) {
    S.hl[i] = lower_bound_ll(S.nkey, nu, S.gh[i]);
    S.tl[i] = lower_bound_ll(S.nkey, nu, S.gt[i]);
  }
  __syncthreads();
  // re-normalised scores on the deduplicated graph (:380-385); sums run in ascending edge order
  if (a.score_mode == EVI_GAGENT_NODE_SOFTMAX) {
    for (int i = tid; i < m; i += kTopkThreads) {
      const int h = S.hl[i], t = S.tl[i];
      const float s = S.raw[i];
      float mh = -INFINITY, mt = -INFINITY;
      for (int j = 0; j < m; ++j) {
        if (S.hl[j] == h) mh = fmaxf(mh, S.raw[j]);
        if (S.tl[j] == t) mt = fmaxf(mt, S.raw[j]);
      }
      float sh = 0.f, st = 0.f;
      for (int j = 0; j < m; ++j) {
        if (S.hl[j] == h) sh += expf(S.raw[j] - mh);
        if (S.tl[j] == t) st += expf(S.raw[j] - mt);
      }
      const float prob = (expf(s - mh) / fmaxf(sh, kProbEps) + expf(s - mt) / fmaxf(st, kProbEps)) * 0.5f;
      S.tmp[i] = prob_to_logit(prob);
    }
  } else {
    for (int i = tid; i < m; i += kTopkThreads) S.tmp[i] = S.raw[i];
  }
  __syncthreads();
  for (int i = tid; i < m; i += kTopkThreads) {
    a.env_edge_ids[ob + i] = S.env[i];
    a.edge_heads_global[ob + i] = S.gh[i];
    a.edge_tails_global[ob + i] = S.gt[i];
    a.edge_relations[ob + i] = S.rel[i];
    a.edge_scores[ob + i] = S.tmp[i];
    a.edge_labels[ob + i] = S.lab[i];
    a.edge_head_locals[ob + i] = S.hl[i];
    a.edge_tail_locals[ob + i] = S.tl[i];
  }
  // D. anchors (:387-428): start ids in their given order, answers deduplicated in order
  if (tid == 0) {
    const long long s0 = a.start_ptr[b], s1 = a.start_ptr[b + 1];
    int* sl = a.start_node_locals + static_cast<size_t>(b) * a.max_starts;
    int nsl = 0;
    for (long long i = s0; i < s1; ++i) {
      const int p = lower_bound_ll(S.nkey, nu, a.start_entity_ids[i]);
      if (p >= nu || S.nkey[p] != a.start_entity_ids[i]) continue;
      bool seen = false;
      for (int j = 0; j < nsl && !seen; ++j) seen = sl[j] == p;
      if (!seen && nsl < a.max_starts) sl[nsl++] = p;
    }
    const long long a0 = a.answer_ptr[b], a1 = a.answer_ptr[b + 1];
    long long* au = a.answer_ids_unique + static_cast<size_t>(b) * a.max_answers;
    int* al = a.answer_node_locals + static_cast<size_t>(b) * a.max_answers;
    int nau = 0, nal = 0;
    for (long long i = a0; i < a1; ++i) {
      const long long g = a.answer_entity_ids[i];
      bool seen = false;
      for (int j = 0; j < nau && !seen; ++j) seen = au[j] == g;
      if (seen || nau >= a.max_answers) continue;
      au[nau++] = g;
      const int p = lower_bound_ll(S.nkey, nu, g);
      if (p < nu && S.nkey[p] == g) al[nal++] = p;
    }
    int st = status;
    if (nsl == 0) st |= EVI_GAGENT_NO_START_LOCAL;
    if (nal == 0) st |= EVI_GAGENT_NO_ANSWER;
    a.status[b] = st;
    a.num_triples[b] = m;
    a.num_nodes_out[b] = nu;
    a.num_start_locals[b] = nsl;
    a.num_answers[b] = nau;
    a.num_answer_locals[b] = nal;
  }
}

}  // namespace evi

using namespace evi;

extern "C" {

size_t evi_node_softmax_workspace_bytes(int num_nodes) {
  return sizeof(NodeSoftmaxWs) * static_cast<size_t>(num_nodes > 0 ? num_nodes : 0);
}

int evi_node_softmax_logit(const float* scores, const int* head32, const int* tail32, long long num_edges,
                           int num_nodes, int score_mode, float temperature, float bias, void* workspace,
                           size_t workspace_bytes, float* calibrated, float* out, cudaStream_t stream) {
  if (num_edges <= 0) return kOk;
  if (scores == nullptr || calibrated == nullptr || out == nullptr || !(temperature > 0.f)) return kErrInvalidArg;
  const int calibrate = (temperature != 1.0f || bias != 0.0f) ? 1 : 0;
  const int grid = static_cast<int>((num_edges + 255) / 256);
  if (score_mode == EVI_GAGENT_LOGITS) {
    node_softmax_max_kernel<<<grid, 256, 0, stream>>>(scores, nullptr, nullptr, num_edges, temperature, bias, calibrate,
                                                      nullptr, calibrated);
    if (cudaMemcpyAsync(out, calibrated, sizeof(float) * num_edges, cudaMemcpyDeviceToDevice, stream) != cudaSuccess)
      return kErrCuda;
    return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
  }
  if (score_mode != EVI_GAGENT_NODE_SOFTMAX || head32 == nullptr || tail32 == nullptr || workspace == nullptr)
    return kErrInvalidArg;
  if (workspace_bytes < evi_node_softmax_workspace_bytes(num_nodes)) return kErrWorkspace;
  NodeSoftmaxWs* ws = static_cast<NodeSoftmaxWs*>(workspace);
  if (cudaMemsetAsync(ws, 0, sizeof(NodeSoftmaxWs) * static_cast<size_t>(num_nodes), stream) != cudaSuccess)
    return kErrCuda;
  node_softmax_max_kernel<<<grid, 256, 0, stream>>>(scores, head32, tail32, num_edges, temperature, bias, calibrate, ws,
                                                    calibrated);
  node_softmax_sum_kernel<<<grid, 256, 0, stream>>>(calibrated, head32, tail32, num_edges, ws);
  node_softmax_final_kernel<<<grid, 256, 0, stream>>>(calibrated, head32, tail32, num_edges, ws, out);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_gagent_start_edges(const float* sel_scores, const int* head32, const int* tail32, const long long* edge_ptr,
                           const long long* node_ptr, int num_graphs, const long long* node_global_ids,
                           const long long* start_entity_ids, const long long* start_ptr, float start_keep_ratio,
                           int start_min_edges, int start_max_edges, unsigned char* mark, int* status,
                           cudaStream_t stream) {
  if (num_graphs <= 0) return kOk;
  if (sel_scores == nullptr || head32 == nullptr || tail32 == nullptr || edge_ptr == nullptr || node_ptr == nullptr ||
      node_global_ids == nullptr || start_entity_ids == nullptr || start_ptr == nullptr || mark == nullptr ||
      status == nullptr || !(start_keep_ratio >= 0.f && start_keep_ratio <= 1.f) || start_min_edges < 0)
    return kErrInvalidArg;
  gagent_start_edges_kernel<<<num_graphs, kTopkThreads, 0, stream>>>(
      sel_scores, head32, tail32, edge_ptr, node_ptr, node_global_ids, start_entity_ids, start_ptr, start_keep_ratio,
      start_min_edges, start_max_edges, mark, status);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

static size_t gagent_smem_bytes(int cap) {
  size_t P2 = 1;
  while (P2 < 2 * static_cast<size_t>(cap)) P2 <<= 1;
  return static_cast<size_t>(cap) * (3 * 8 + 3 * 4 + 3 * 4 + 1) + P2 * (8 + 4) + 16;
}

int evi_gagent_build(const evi_gagent_args* args, cudaStream_t stream) {
  if (args == nullptr) return kErrInvalidArg;
  const evi_gagent_args& a = *args;
  if (a.num_graphs <= 0) return kOk;
  if (a.cap <= 0 || a.cap > EVI_GAGENT_MAX_CAP || a.max_starts <= 0 || a.max_answers <= 0 || a.topk_stride <= 0)
    return kErrInvalidArg;
  if (a.score_mode != EVI_GAGENT_LOGITS && a.score_mode != EVI_GAGENT_NODE_SOFTMAX) return kErrInvalidArg;
  const void* required[] = {a.raw_scores, a.head32, a.tail32, a.edge_attr, a.edge_ptr, a.node_ptr, a.node_global_ids,
                            a.node_embedding_ids, a.start_entity_ids, a.start_ptr, a.answer_entity_ids, a.answer_ptr,
                            a.topk_indices, a.topk_counts, a.mark, a.status, a.num_triples, a.num_nodes_out,
                            a.num_start_locals, a.num_answers, a.num_answer_locals, a.env_edge_ids,
                            a.edge_heads_global, a.edge_tails_global, a.edge_relations, a.edge_scores, a.edge_labels,
                            a.edge_head_locals, a.edge_tail_locals, a.node_entity_ids, a.node_embedding_ids_out,
                            a.node_batch_index, a.start_node_locals, a.answer_ids_unique, a.answer_node_locals};
  for (const void* p : required)
    if (p == nullptr) return kErrInvalidArg;
  const size_t smem = gagent_smem_bytes(a.cap);
  if (cudaFuncSetAttribute(gagent_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) !=
      cudaSuccess)
    return kErrCuda;
  gagent_build_kernel<<<a.num_graphs, kTopkThreads, smem, stream>>>(a);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // extern "C"
