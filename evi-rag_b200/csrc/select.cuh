// Block-level selection helpers shared by the ragged top-k kernel and the g_agent selection kernels:
// order-preserving float keys, a 256-thread inclusive scan, and the MSB-first radix threshold search.
#pragma once

#include <math.h>

#include "common.cuh"
#include "evi_rag_b200.h"

namespace evi {

constexpr int kTopkThreads = 256;
constexpr int kTopkMaxK = 1024;

EVI_DEVINL float apply_transform(float x, int transform) {
  return transform == EVI_TOPK_SIGMOID ? 1.0f / (1.0f + expf(-x)) : x;
}
// order-preserving float -> uint key; NaN sorts first like torch.topk, -0 == +0
EVI_DEVINL uint32_t order_key(float x) {
  if (x != x) return 0xFFFFFFFFu;
  const uint32_t b = __float_as_uint(x + 0.0f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// block-wide inclusive scan of one int per thread (blockDim == kTopkThreads)
EVI_DEVINL int block_inclusive_scan(int v, int* warp_tot /*[8]*/, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  if (lane == 31) warp_tot[w] = incl;
  __syncthreads();
  int off = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < kTopkThreads / 32; ++i) {
    const int t = warp_tot[i];
    if (i < w) off += t;
    tot += t;
  }
  __syncthreads();
  *total = tot;
  return incl + off;
}


// MSB-first radix search (4 x 8-bit passes) for the kt-th largest key among the entries of a virtual list
// of length n for which key_at(i, &key) is true. All kTopkThreads threads of the CTA call it. On return
// *kth = that key and *need = how many entries equal to it belong to the top kt (>= 1); entries with a
// larger key number kt - need. Requires 1 <= kt <= (number of entries).
struct RadixScratch {
  int hist[256];
  int warp_tot[kTopkThreads / 32];
  uint32_t sel_prefix;
  int sel_need;
};

template <class KeyAt>
EVI_DEVINL void block_radix_threshold(int n, int kt, KeyAt key_at, RadixScratch* sc, uint32_t* kth, int* need_out) {
  const int tid = threadIdx.x;
  if (tid == 0) sc->sel_prefix = 0u, sc->sel_need = kt;
  __syncthreads();
  uint32_t mask = 0u;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = 24 - 8 * pass;
    sc->hist[tid] = 0;
    __syncthreads();
    const uint32_t prefix = sc->sel_prefix;
    const int need = sc->sel_need;
    for (int base = 0; base < n; base += kTopkThreads) {
      const int i = base + tid;
      uint32_t key = 0u;
      bool in = false;
      if (i < n) in = key_at(i, &key) && (key & mask) == prefix;
      const unsigned active = __ballot_sync(0xffffffffu, in);
      if (in) {
        const int bin = static_cast<int>((key >> shift) & 255u);
        const unsigned peers = __match_any_sync(active, bin);
        if ((tid & 31) == __ffs(peers) - 1) atomicAdd(&sc->hist[bin], __popc(peers));
      }
    }
    __syncthreads();
    const int own = sc->hist[255 - tid];
    int total;
    const int incl = block_inclusive_scan(own, sc->warp_tot, &total);
    if (incl >= need && incl - own < need) {
      sc->sel_prefix = prefix | (static_cast<uint32_t>(255 - tid) << shift);
      sc->sel_need = need - (incl - own);
    }
    mask |= 255u << shift;
    __syncthreads();
  }
  *kth = sc->sel_prefix;
  *need_out = sc->sel_need;
  __syncthreads();
}

}  // namespace evi
