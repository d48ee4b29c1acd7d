// RetrieverLoss on the device (src/losses/retriever_loss.py:72-178): segmented multi-positive
// InfoNCE + optional BCE over the edges of each question, the separation metrics, and d(loss)/d(logits).
// The reference runs ~20 scatter launches and a dozen `.item()` syncs per step; here:
//   loss_stats_kernel    one CTA per question: max / sum-exp over all edges and over positives, counts, BCE sums,
//                        sigmoid sums (fixed reduction tree: deterministic)
//   loss_reduce_kernel   one CTA: means over the valid questions -> 12 scalars
//   loss_grad_kernel     thread per edge: (softmax_all - [pos] softmax_pos) / T / n_valid + w (sigmoid - y) / denom
// HBM-bound: 8 B/edge in (+4 B/edge weight flag), 4 B/edge out for the gradient.
#include <math.h>

#include "common.cuh"
#include "evi_rag_b200.h"

namespace evi {

constexpr int kLossThreads = 256;
constexpr float kMinEdgeWeight = 1e-6f;   // retriever_loss.py:19
constexpr float kPosThreshold = 0.5f;     // :16

enum LossStat { kLseAll = 0, kLsePos, kPosCount, kEdgeCount, kBceSum, kWeightSum, kSigPos, kSigNeg, kMaxAll, kSumAll,
                kMaxPos, kSumPos, kNumStats };

struct BlockReduce {
  float buf[kLossThreads / 32];
  EVI_DEVINL float sum(float v) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kLossThreads / 32; ++i) t += buf[i];
    return t;
  }
  EVI_DEVINL float max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = -INFINITY;
#pragma unroll
    for (int i = 0; i < kLossThreads / 32; ++i) t = fmaxf(t, buf[i]);
    return t;
  }
};

EVI_DEVINL float edge_weight(const unsigned char* is_near, long long e, float w_near, float w_bridge) {
  return is_near == nullptr ? 1.0f : (is_near[e] ? w_near : w_bridge);
}

__global__ void __launch_bounds__(kLossThreads)
loss_stats_kernel(const float* __restrict__ logits, const float* __restrict__ targets,
                  const long long* __restrict__ edge_ptr, const unsigned char* __restrict__ is_near, float w_near,
                  float w_bridge, float temperature, float* __restrict__ stats) {
  __shared__ BlockReduce red;
  const int q = blockIdx.x, tid = threadIdx.x;
  const long long e0 = edge_ptr[q], e1 = edge_ptr[q + 1];
  float mx_all = -INFINITY, mx_pos = -INFINITY;
  for (long long e = e0 + tid; e < e1; e += kLossThreads) {
    float s = logits[e] / temperature;
    if (is_near) s += logf(fmaxf(edge_weight(is_near, e, w_near, w_bridge), kMinEdgeWeight));
    mx_all = fmaxf(mx_all, s);
    if (targets[e] > kPosThreshold) mx_pos = fmaxf(mx_pos, s);
  }
  mx_all = red.max(mx_all);
  mx_pos = red.max(mx_pos);
  float sum_all = 0.f, sum_pos = 0.f, n_pos = 0.f, bce = 0.f, wsum = 0.f, sig_pos = 0.f, sig_neg = 0.f;
  for (long long e = e0 + tid; e < e1; e += kLossThreads) {
    const float x = logits[e], y = targets[e];
    const float w = edge_weight(is_near, e, w_near, w_bridge);
    float s = x / temperature;
    if (is_near) s += logf(fmaxf(w, kMinEdgeWeight));
    const bool pos = y > kPosThreshold;
    sum_all += expf(s - mx_all);
    if (pos) sum_pos += expf(s - mx_pos), n_pos += 1.f;
    bce += w * (fmaxf(x, 0.f) - x * y + log1pf(expf(-fabsf(x))));  // binary_cross_entropy_with_logits
    wsum += w;
    const float p = 1.0f / (1.0f + expf(-x));
    if (pos) sig_pos += p; else sig_neg += p;
  }
  sum_all = red.sum(sum_all);
  sum_pos = red.sum(sum_pos);
  n_pos = red.sum(n_pos);
  bce = red.sum(bce);
  wsum = red.sum(wsum);
  sig_pos = red.sum(sig_pos);
  sig_neg = red.sum(sig_neg);
  if (tid == 0) {
    float* o = stats + static_cast<size_t>(q) * kNumStats;
    o[kLseAll] = mx_all + logf(fmaxf(sum_all, 1e-12f));
    o[kLsePos] = mx_pos + logf(fmaxf(sum_pos, 1e-12f));
    o[kPosCount] = n_pos;
    o[kEdgeCount] = static_cast<float>(e1 - e0);
    o[kBceSum] = bce;
    o[kWeightSum] = wsum;
    o[kSigPos] = sig_pos;
    o[kSigNeg] = sig_neg;
    o[kMaxAll] = mx_all;
    o[kSumAll] = sum_all;
    o[kMaxPos] = mx_pos;
    o[kSumPos] = sum_pos;
  }
}

// summary[0..11]: infonce, bce, n_valid_infonce, n_valid_bce, pos_edges, neg_edges, graphs_no_pos, graphs_no_neg,
//                 pos_prob, neg_prob, total edges, total loss
__global__ void __launch_bounds__(kLossThreads)
loss_reduce_kernel(const float* __restrict__ stats, int num_graphs, int weighted, float infonce_weight,
                   float bce_weight, float* __restrict__ summary) {
  __shared__ BlockReduce red;
  const int tid = threadIdx.x;
  float inf = 0.f, nvi = 0.f, bce = 0.f, nvb = 0.f, pe = 0.f, ne = 0.f, nop = 0.f, non = 0.f, sp = 0.f, sn = 0.f;
  for (int q = tid; q < num_graphs; q += kLossThreads) {
    const float* s = stats + static_cast<size_t>(q) * kNumStats;
    const float pos = s[kPosCount], neg = s[kEdgeCount] - s[kPosCount];
    if (pos > 0.f && neg > 0.f) inf += s[kLseAll] - s[kLsePos], nvi += 1.f;
    const bool vb = weighted ? s[kWeightSum] > 0.f : s[kEdgeCount] > 0.f;
    if (vb) bce += s[kBceSum] / (weighted ? fmaxf(s[kWeightSum], kMinEdgeWeight) : s[kEdgeCount]), nvb += 1.f;
    pe += pos, ne += neg;
    nop += pos == 0.f ? 1.f : 0.f;
    non += neg == 0.f ? 1.f : 0.f;
    sp += s[kSigPos], sn += s[kSigNeg];
  }
  inf = red.sum(inf); nvi = red.sum(nvi); bce = red.sum(bce); nvb = red.sum(nvb); pe = red.sum(pe);
  ne = red.sum(ne); nop = red.sum(nop); non = red.sum(non); sp = red.sum(sp); sn = red.sum(sn);
  if (tid == 0) {
    const bool any = pe > 0.f && ne > 0.f && nvi > 0.f;  // early exits of _infonce_loss (:84-91, :131-138)
    const float l_inf = any ? inf / nvi : 0.f;
    const float l_bce = (bce_weight > 0.f && nvb > 0.f) ? bce / nvb : 0.f;
    summary[0] = l_inf;
    summary[1] = l_bce;
    summary[2] = any ? nvi : 0.f;
    summary[3] = bce_weight > 0.f ? nvb : 0.f;
    summary[4] = pe;
    summary[5] = ne;
    summary[6] = nop;
    summary[7] = non;
    summary[8] = pe > 0.f ? sp / pe : 0.f;
    summary[9] = ne > 0.f ? sn / ne : 0.f;
    summary[10] = pe + ne;
    summary[11] = infonce_weight * l_inf + bce_weight * l_bce;
  }
}

__global__ void loss_grad_kernel(const float* __restrict__ logits, const float* __restrict__ targets,
                                 const int* __restrict__ qid, long long num_edges,
                                 const unsigned char* __restrict__ is_near, float w_near, float w_bridge,
                                 float temperature, int weighted, float infonce_weight, float bce_weight,
                                 const float* __restrict__ stats, const float* __restrict__ summary,
                                 float* __restrict__ grad) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  const float* s = stats + static_cast<size_t>(qid[e]) * kNumStats;
  const float x = logits[e], y = targets[e];
  const float w = edge_weight(is_near, e, w_near, w_bridge);
  float g = 0.f;
  const float pos_n = s[kPosCount], neg_n = s[kEdgeCount] - s[kPosCount];
  if (infonce_weight > 0.f && summary[2] > 0.f && pos_n > 0.f && neg_n > 0.f) {
    float sc = x / temperature;
    if (is_near) sc += logf(fmaxf(w, kMinEdgeWeight));
    float d = expf(sc - s[kMaxAll]) / s[kSumAll];
    if (y > kPosThreshold) d -= expf(sc - s[kMaxPos]) / s[kSumPos];
    g += infonce_weight * d / (temperature * summary[2]);
  }
  if (bce_weight > 0.f && summary[3] > 0.f) {
    const bool vb = weighted ? s[kWeightSum] > 0.f : s[kEdgeCount] > 0.f;
    if (vb) {
      const float denom = weighted ? fmaxf(s[kWeightSum], kMinEdgeWeight) : s[kEdgeCount];
      g += bce_weight * w * (1.0f / (1.0f + expf(-x)) - y) / (denom * summary[3]);
    }
  }
  grad[e] = g;
}

}  // namespace evi

using namespace evi;

extern "C" {

int evi_retriever_loss(const float* logits, const float* targets, const long long* edge_ptr, const int* qid32,
                       long long num_edges, int num_graphs, const unsigned char* edge_is_near, float edge_weight_near,
                       float edge_weight_bridge, float infonce_temperature, float infonce_weight, float bce_weight,
                       float* stats, float* summary, float* grad_logits, cudaStream_t stream) {
  if (num_graphs <= 0 || num_edges <= 0) return kErrInvalidArg;
  if (logits == nullptr || targets == nullptr || edge_ptr == nullptr || stats == nullptr || summary == nullptr ||
      !(infonce_temperature > 0.f) || infonce_weight < 0.f || bce_weight < 0.f ||
      (grad_logits != nullptr && qid32 == nullptr))
    return kErrInvalidArg;
  const int weighted = edge_is_near != nullptr ? 1 : 0;
  loss_stats_kernel<<<num_graphs, kLossThreads, 0, stream>>>(logits, targets, edge_ptr, edge_is_near, edge_weight_near,
                                                            edge_weight_bridge, infonce_temperature, stats);
  loss_reduce_kernel<<<1, kLossThreads, 0, stream>>>(stats, num_graphs, weighted, infonce_weight, bce_weight, summary);
  if (grad_logits != nullptr)
    loss_grad_kernel<<<static_cast<int>((num_edges + 255) / 256), 256, 0, stream>>>(
        logits, targets, qid32, num_edges, edge_is_near, edge_weight_near, edge_weight_bridge, infonce_temperature,
        weighted, infonce_weight, bce_weight, stats, summary, grad_logits);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // extern "C"
