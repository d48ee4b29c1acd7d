// Ragged per-question top-k. One CTA per question:
//   1. MSB-first radix select (4 x 8-bit passes, warp-aggregated shared-memory histograms) finds
//      the k-th largest order-preserving key and how many ties at that key are needed;
//   2. everything above the threshold is collected unordered, ties are collected in ascending
//      index order (ballot prefix scan) so the result is deterministic;
//   3. the <= 1024 candidates are bitonic-sorted on (key desc, index asc).
// Scores are read straight from global memory each pass: a question's scores (E_q * 4 B, ~20 KB
// for WebQSP-shaped inputs) stay L1/L2 resident, so HBM sees them once.
#include <math.h>

#include "common.cuh"
#include "evi_rag_b200.h"
#include "select.cuh"

namespace evi {

__global__ void __launch_bounds__(kTopkThreads)
segmented_topk_kernel(const float* __restrict__ scores, const long long* __restrict__ edge_ptr, int k,
                      int transform, float* __restrict__ values, int* __restrict__ indices,
                      int* __restrict__ counts) {
  __shared__ unsigned long long cand[kTopkMaxK];
  __shared__ int hist[256];
  __shared__ int warp_tot[kTopkThreads / 32];
  __shared__ uint32_t sel_prefix;
  __shared__ int sel_need, cursor;

  const int b = blockIdx.x;
  const long long e0 = edge_ptr[b];
  const int n = static_cast<int>(edge_ptr[b + 1] - e0);
  const int kt = n < k ? n : k;
  const float* s = scores + e0;
  const int tid = threadIdx.x;
  if (tid == 0) counts[b] = kt;
  for (int j = kt + tid; j < k; j += kTopkThreads) {
    values[static_cast<size_t>(b) * k + j] = -INFINITY;
    indices[static_cast<size_t>(b) * k + j] = -1;
  }
  if (kt == 0) return;

  int P = 1;
  while (P < kt) P <<= 1;
  for (int j = kt + tid; j < P; j += kTopkThreads) cand[j] = 0ull;

  if (n <= k) {
    for (int i = tid; i < n; i += kTopkThreads)
      cand[i] = (static_cast<unsigned long long>(order_key(apply_transform(s[i], transform))) << 32) |
                static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<uint32_t>(i));
  } else {
    if (tid == 0) sel_prefix = 0u, sel_need = kt, cursor = 0;
    __syncthreads();
    uint32_t mask = 0u;
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      hist[tid] = 0;
      __syncthreads();
      const uint32_t prefix = sel_prefix;
      const int need = sel_need;  // read before this pass's winner overwrites it
      for (int base = 0; base < n; base += kTopkThreads) {
        const int i = base + tid;
        uint32_t key = 0u;
        bool in = false;
        if (i < n) {
          key = order_key(apply_transform(s[i], transform));
          in = (key & mask) == prefix;
        }
        const unsigned active = __ballot_sync(0xffffffffu, in);
        if (in) {
          const int bin = static_cast<int>((key >> shift) & 255u);
          const unsigned peers = __match_any_sync(active, bin);
          if ((tid & 31) == __ffs(peers) - 1) atomicAdd(&hist[bin], __popc(peers));
        }
      }
      __syncthreads();
      // thread t owns bin 255-t: inclusive scan over t == count of keys in bins >= 255-t
      const int own = hist[255 - tid];
      int total;
      const int incl = block_inclusive_scan(own, warp_tot, &total);
      if (incl >= need && incl - own < need) {
        sel_prefix = prefix | (static_cast<uint32_t>(255 - tid) << shift);
        sel_need = need - (incl - own);
      }
      mask |= 255u << shift;
      __syncthreads();
    }
    const uint32_t kth = sel_prefix;
    const int need = sel_need;       // ties at kth to take, >= 1
    const int n_greater = kt - need;
    for (int i = tid; i < n; i += kTopkThreads) {
      const uint32_t key = order_key(apply_transform(s[i], transform));
      if (key > kth)
        cand[atomicAdd(&cursor, 1)] = (static_cast<unsigned long long>(key) << 32) |
                                      static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<uint32_t>(i));
    }
    int running = 0;
    for (int base = 0; base < n && running < need; base += kTopkThreads) {
      const int i = base + tid;
      const bool tie = i < n && order_key(apply_transform(s[i], transform)) == kth;
      int total;
      const int incl = block_inclusive_scan(tie ? 1 : 0, warp_tot, &total);
      const int rank = running + incl - 1;
      if (tie && rank < need)
        cand[n_greater + rank] = (static_cast<unsigned long long>(kth) << 32) |
                                 static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<uint32_t>(i));
      running += total;
    }
  }
  __syncthreads();
  // bitonic sort, descending on the 64-bit composite
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < P; i += kTopkThreads) {
        const int j = i ^ stride;
        if (j > i) {
          const unsigned long long a = cand[i], c = cand[j];
          const bool descending = (i & size) == 0;
          if (descending ? (a < c) : (a > c)) cand[i] = c, cand[j] = a;
        }
      }
      __syncthreads();
    }
  }
  for (int j = tid; j < kt; j += kTopkThreads) {
    const int idx = static_cast<int>(0xFFFFFFFFu - static_cast<uint32_t>(cand[j] & 0xFFFFFFFFull));
    indices[static_cast<size_t>(b) * k + j] = idx;
    values[static_cast<size_t>(b) * k + j] = apply_transform(s[idx], transform);
  }
}

__global__ void topk_payload_kernel(const int* __restrict__ indices, const int* __restrict__ counts,
                                    const long long* __restrict__ edge_ptr, int num_graphs, int k,
                                    const long long* __restrict__ edge_index, long long num_edges,
                                    const long long* __restrict__ node_global_ids,
                                    const long long* __restrict__ edge_attr, const float* __restrict__ labels,
                                    const float* __restrict__ logits_fwd, const float* __restrict__ logits_bwd,
                                    long long* __restrict__ head_ids, long long* __restrict__ tail_ids,
                                    long long* __restrict__ relation_ids, float* __restrict__ out_labels,
                                    float* __restrict__ out_lf, float* __restrict__ out_lb) {
  const long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (t >= static_cast<long long>(num_graphs) * k) return;
  const int b = static_cast<int>(t / k), j = static_cast<int>(t % k);
  const bool valid = j < counts[b];
  const long long e = valid ? edge_ptr[b] + indices[t] : 0;
  if (head_ids) head_ids[t] = valid ? node_global_ids[edge_index[e]] : -1;
  if (tail_ids) tail_ids[t] = valid ? node_global_ids[edge_index[num_edges + e]] : -1;
  if (relation_ids) relation_ids[t] = valid ? edge_attr[e] : -1;
  if (out_labels) out_labels[t] = valid && labels ? labels[e] : 0.f;
  if (out_lf) out_lf[t] = valid && logits_fwd ? logits_fwd[e] : 0.f;
  if (out_lb) out_lb[t] = valid && logits_bwd ? logits_bwd[e] : 0.f;
}

// ---------------------------------------------------------------------------------------------
// Metric consumers of one ragged top-k, one CTA per question, from the selected
// indices instead of one torch.topk + several host syncs per graph and per metric:
//   hits[b, i]   = positives among the first min(k_i, k_top) selected edges     retriever_metrics.py:134-157
//   pos[b]       = positives among all edges of the question (label > 0.5)      retriever_metrics.py:43,142
//   reach[b, i]  = some answer node is connected to some seed node by the first min(k_i, k_top)
//                  selected edges (union-find)                                   reachability.py:329-381
//   margin[b]    = min positive score - max negative score                       retriever_metrics.py:372-386
// k_values must be ascending (normalize_k_values, src/utils/metrics.py:25-40).
// ---------------------------------------------------------------------------------------------
constexpr int kMetricThreads = 128;
constexpr int kMaxMetricK = 16;

struct MetricKs {
  int v[kMaxMetricK];
  int n;
};

EVI_DEVINL int uf_find(int* parent, int x) {
  while (parent[x] != x) {
    parent[x] = parent[parent[x]];  // path halving (reachability.py:300-304)
    x = parent[x];
  }
  return x;
}

__global__ void __launch_bounds__(kMetricThreads)
topk_metrics_kernel(const int* __restrict__ indices, const int* __restrict__ counts,
                    const long long* __restrict__ edge_ptr, int k, MetricKs ks,
                    const float* __restrict__ labels, const float* __restrict__ scores,
                    const int* __restrict__ head, const int* __restrict__ tail, const long long* __restrict__ node_ptr,
                    const long long* __restrict__ seeds, const long long* __restrict__ seed_ptr,
                    const long long* __restrict__ answers, const long long* __restrict__ answer_ptr,
                    int* __restrict__ uf_parent,  // [total nodes] scratch
                    float* __restrict__ hits, float* __restrict__ pos, unsigned char* __restrict__ reach,
                    unsigned char* __restrict__ reach_valid, float* __restrict__ margin,
                    unsigned char* __restrict__ margin_valid) {
  __shared__ float red_a[kMetricThreads / 32], red_b[kMetricThreads / 32];
  __shared__ int red_i[kMetricThreads / 32], s_num_pos;
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const long long e0 = edge_ptr[b];
  const int ne = static_cast<int>(edge_ptr[b + 1] - e0);
  const int kt = counts[b];
  const int* idx = indices + static_cast<size_t>(b) * k;

  if (labels != nullptr) {
    // positives of the whole question, and hits at each k
    int p = 0;  // counted as integers: exact for any question size
    for (int i = tid; i < ne; i += kMetricThreads) p += labels[e0 + i] > 0.5f ? 1 : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
    if (lane == 0) red_i[w] = p;
    __syncthreads();
    if (tid == 0) {
      int t = 0;
      for (int i = 0; i < kMetricThreads / 32; ++i) t += red_i[i];
      s_num_pos = t;
      pos[b] = static_cast<float>(t);
    }
    __syncthreads();
    for (int ki = 0; ki < ks.n; ++ki) {
      const int ke = ks.v[ki] < kt ? ks.v[ki] : kt;
      int h = 0;
      for (int j = tid; j < ke; j += kMetricThreads) h += labels[e0 + idx[j]] > 0.5f ? 1 : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
      if (lane == 0) red_i[w] = h;
      __syncthreads();
      if (tid == 0) {
        int t = 0;
        for (int i = 0; i < kMetricThreads / 32; ++i) t += red_i[i];
        hits[static_cast<size_t>(b) * ks.n + ki] = static_cast<float>(t);
      }
      __syncthreads();
    }
    if (scores != nullptr && margin != nullptr) {
      float mn = INFINITY, mx = -INFINITY;  // min over positives, max over negatives
      for (int i = tid; i < ne; i += kMetricThreads) {
        const float s = scores[e0 + i];
        if (labels[e0 + i] > 0.5f) mn = fminf(mn, s); else mx = fmaxf(mx, s);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      }
      if (lane == 0) red_a[w] = mn, red_b[w] = mx;
      __syncthreads();
      if (tid == 0) {
        for (int i = 1; i < kMetricThreads / 32; ++i) mn = fminf(mn, red_a[i]), mx = fmaxf(mx, red_b[i]);
        // validity from label presence like the reference (labels.any() / (~labels).any(), retriever_metrics.py:372-386),
        // not from the sentinels: a question whose only positive scores +inf is still a valid question
        const bool ok = s_num_pos > 0 && ne - s_num_pos > 0;
        margin[b] = ok ? mn - mx : 0.f;
        margin_valid[b] = ok ? 1 : 0;
      }
      __syncthreads();
    }
  }

  if (reach == nullptr) return;
  const long long n0 = node_ptr[b];
  const int nn = static_cast<int>(node_ptr[b + 1] - n0);
  int* parent = uf_parent + n0;
  for (int i = tid; i < nn; i += kMetricThreads) parent[i] = i;
  __syncthreads();
  if (tid != 0) return;
  // seeds / answers inside this question's node range (reachability.py:160-165)
  const long long s0 = seed_ptr[b], s1 = seed_ptr[b + 1], a0 = answer_ptr[b], a1 = answer_ptr[b + 1];
  int n_seed = 0, n_ans = 0;
  for (long long i = s0; i < s1; ++i) n_seed += (seeds[i] >= n0 && seeds[i] < n0 + nn) ? 1 : 0;
  for (long long i = a0; i < a1; ++i) n_ans += (answers[i] >= n0 && answers[i] < n0 + nn) ? 1 : 0;
  const bool valid = ne > 0 && nn > 0 && n_seed > 0 && n_ans > 0;
  reach_valid[b] = valid ? 1 : 0;
  if (!valid) {
    for (int ki = 0; ki < ks.n; ++ki) reach[static_cast<size_t>(b) * ks.n + ki] = 0;
    return;
  }
  auto reachable = [&]() {
    for (long long a = a0; a < a1; ++a) {
      if (answers[a] < n0 || answers[a] >= n0 + nn) continue;
      const int ra = uf_find(parent, static_cast<int>(answers[a] - n0));
      for (long long s = s0; s < s1; ++s) {
        if (seeds[s] < n0 || seeds[s] >= n0 + nn) continue;
        if (uf_find(parent, static_cast<int>(seeds[s] - n0)) == ra) return true;
      }
    }
    return false;
  };
  int next = 0;
  bool cur = false;
  for (int j = 0; j < kt && next < ks.n; ++j) {
    const long long e = e0 + idx[j];
    const int u = head[e] - static_cast<int>(n0), v = tail[e] - static_cast<int>(n0);
    if (u >= 0 && v >= 0 && u < nn && v < nn) {
      const int ru = uf_find(parent, u), rv = uf_find(parent, v);
      if (ru != rv) parent[rv] = ru;
    }
    bool fresh = false;
    while (next < ks.n && j + 1 >= (ks.v[next] < kt ? ks.v[next] : kt)) {
      if (!fresh) cur = reachable(), fresh = true;
      reach[static_cast<size_t>(b) * ks.n + next] = cur ? 1 : 0;
      ++next;
    }
  }
  for (; next < ks.n; ++next) reach[static_cast<size_t>(b) * ks.n + next] = cur ? 1 : 0;  // kt == 0 cannot happen when valid
}

}  // namespace evi

using namespace evi;

extern "C" {

int evi_abi_version(void) { return 5; }

int evi_segmented_topk(const float* scores, const long long* edge_ptr, int num_graphs, int k, int transform,
                       float* values, int* indices, int* counts, cudaStream_t stream) {
  if (num_graphs <= 0) return kOk;
  if (k <= 0 || k > kTopkMaxK || edge_ptr == nullptr || values == nullptr || indices == nullptr ||
      counts == nullptr)
    return kErrInvalidArg;
  if (transform != EVI_TOPK_RAW && transform != EVI_TOPK_SIGMOID) return kErrInvalidArg;
  segmented_topk_kernel<<<num_graphs, kTopkThreads, 0, stream>>>(scores, edge_ptr, k, transform, values, indices,
                                                                 counts);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_topk_gather_payload(const int* indices, const int* counts, const long long* edge_ptr, int num_graphs,
                            int k, const long long* edge_index, long long num_edges,
                            const long long* node_global_ids, const long long* edge_attr, const float* labels,
                            const float* logits_fwd, const float* logits_bwd, long long* head_ids,
                            long long* tail_ids, long long* relation_ids, float* out_labels,
                            float* out_logits_fwd, float* out_logits_bwd, cudaStream_t stream) {
  if (num_graphs <= 0 || k <= 0) return kOk;
  const long long total = static_cast<long long>(num_graphs) * k;
  topk_payload_kernel<<<static_cast<int>((total + 255) / 256), 256, 0, stream>>>(
      indices, counts, edge_ptr, num_graphs, k, edge_index, num_edges, node_global_ids, edge_attr, labels,
      logits_fwd, logits_bwd, head_ids, tail_ids, relation_ids, out_labels, out_logits_fwd, out_logits_bwd);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

int evi_topk_metrics(const int* indices, const int* counts, const long long* edge_ptr, int num_graphs, int k,
                     const int* k_values_host, int num_k, const float* labels, const float* scores, const int* head32,
                     const int* tail32, const long long* node_ptr, const long long* seeds, const long long* seed_ptr,
                     const long long* answers, const long long* answer_ptr, int* uf_parent_scratch, float* hits,
                     float* pos, unsigned char* reach, unsigned char* reach_valid, float* margin,
                     unsigned char* margin_valid, cudaStream_t stream) {
  if (num_graphs <= 0) return kOk;
  if (num_k <= 0 || num_k > kMaxMetricK || k <= 0 || indices == nullptr || counts == nullptr || edge_ptr == nullptr ||
      k_values_host == nullptr)
    return kErrInvalidArg;
  if (labels != nullptr && (hits == nullptr || pos == nullptr)) return kErrInvalidArg;
  if (margin != nullptr && (scores == nullptr || labels == nullptr || margin_valid == nullptr)) return kErrInvalidArg;
  if (reach != nullptr && (head32 == nullptr || tail32 == nullptr || node_ptr == nullptr || seeds == nullptr ||
                           seed_ptr == nullptr || answers == nullptr || answer_ptr == nullptr ||
                           uf_parent_scratch == nullptr || reach_valid == nullptr))
    return kErrInvalidArg;
  MetricKs ks;
  ks.n = num_k;
  for (int i = 0; i < num_k; ++i) {
    if (k_values_host[i] <= 0 || (i > 0 && k_values_host[i] <= k_values_host[i - 1])) return kErrInvalidArg;
    ks.v[i] = k_values_host[i];
  }
  topk_metrics_kernel<<<num_graphs, kMetricThreads, 0, stream>>>(indices, counts, edge_ptr, k, ks, labels, scores, head32,
                                                              tail32, node_ptr, seeds, seed_ptr, answers, answer_ptr,
                                                              uf_parent_scratch, hits, pos, reach, reach_valid, margin,
                                                              margin_valid);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // extern "C"
