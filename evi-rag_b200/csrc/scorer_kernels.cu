// Triple scoring: Retriever._prepare_edge_inputs / _score_edges /
// _combine_directional_outputs of src/models/components/retriever.py:403-495, 369-381.
//
// Dense contractions run on tcgen05 (gemm_tcgen05.cuh); the per-edge elementwise work that builds
// the GEMM operand rows ([h*r*t*nav | GELU(LN(struct)) | h+r-t]) and the LayerNorm/GELU/score-head
// tail run in the warp-per-edge kernels below with 128-bit loads of the 1024-d rows.
#include "gemm_tcgen05.cuh"
#include "evi_rag_b200.h"

namespace evi {

constexpr int D = 1024;        // emb_dim == hidden_dim (configs/model/retriever_module.yaml:10-11)
constexpr int kMaxStruct = 36; // 2*2*(1+4+4)
constexpr float kLnEps = 1e-5f;

static inline int check_launch() { return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda; }

// ---------------------------------------------------------------------------------------------
// Row gather / convert: fp32 or 16-bit source rows -> dense 16-bit GEMM operand.
// ids == nullptr: identity. Rows with sub_mask_ids[row] == 0 take `sub_row` (fp32) instead
// (the non-text entity embedding, retriever.py:499-506 — tanh(W x + b) is applied by the GEMM
// epilogue, so substituting the *input* row is equivalent to substituting the projected row).
// ---------------------------------------------------------------------------------------------
template <int FMT, bool SRC_F32>
__global__ void gather_rows_kernel(const void* __restrict__ src, long long src_rows,
                                   const long long* __restrict__ ids, const long long* __restrict__ sub_mask_ids,
                                   const float* __restrict__ sub_row, int num_rows,
                                   uint16_t* __restrict__ dst, unsigned* __restrict__ flags) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= num_rows) return;
  const int lane = threadIdx.x & 31;
  long long sid = ids ? __ldg(ids + row) : row;
  const bool substitute = sub_mask_ids != nullptr && __ldg(sub_mask_ids + row) == 0;
  if (!substitute && (sid < 0 || sid >= src_rows)) {
    if (lane == 0) atomicOr(flags, kFlagOutOfRange);
    sid = 0;
  }
  uint4* out = reinterpret_cast<uint4*>(dst + static_cast<size_t>(row) * D);
#pragma unroll
  for (int i = 0; i < D / 256; ++i) {
    const int col = i * 256 + lane * 8;
    float v[8];
    if (substitute || SRC_F32) {
      const float* s = substitute ? sub_row : reinterpret_cast<const float*>(src) + static_cast<size_t>(sid) * D;
      const float4 a = __ldg(reinterpret_cast<const float4*>(s + col));
      const float4 b = __ldg(reinterpret_cast<const float4*>(s + col + 4));
      v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
      uint4 w;
      w.x = Op16<FMT>::pack2(v[0], v[1]);
      w.y = Op16<FMT>::pack2(v[2], v[3]);
      w.z = Op16<FMT>::pack2(v[4], v[5]);
      w.w = Op16<FMT>::pack2(v[6], v[7]);
      out[i * 32 + lane] = w;
    } else {
      out[i * 32 + lane] =
          ldg_nc_v4(reinterpret_cast<const uint16_t*>(src) + static_cast<size_t>(sid) * D + col);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Edge operand kernel: one warp per edge, both directions.
//   s_f = [ns[h], ns[t]], s_b = [ns[t], ns[h]]                          retriever.py:509-517
//   sc  = GELU(LN(W_s s + b_s)) ; nav = sigmoid(w_n . sc + b_n)          retriever.py:486-495
//   r_ctx = rel * gate[q] + bias[q]                                      retriever.py:464
//   iv  = h * r_ctx * t * nav ; err = h + r_ctx - t ; dist = -||err||    retriever.py:472-478
// Writes operand rows A_dir[e] = [iv | sc | err] (16-bit, 3*D) and dist_dir[e].
// Lane l owns columns {256*i + 8*l .. +7}, i = 0..3  (128-bit loads / stores).
// ---------------------------------------------------------------------------------------------
struct StructParams {
  const float* ws_t;   // [S, D] struct_proj.0.weight transposed
  const float* bs;     // [D]
  const float* ln_g;   // [D] struct_proj.1.weight
  const float* ln_b;   // [D]
  const float* wn;     // [D] struct_gate_net.0.weight
  float bn;
  int s_dim;           // 2 * node_struct width
};

template <int FMT>
EVI_DEVINL void load_row16(const uint16_t* __restrict__ base, int lane, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 w = *reinterpret_cast<const uint4*>(base + i * 256 + lane * 8);
    const float2 a = Op16<FMT>::unpack2(w.x), b = Op16<FMT>::unpack2(w.y), c = Op16<FMT>::unpack2(w.z),
                 d = Op16<FMT>::unpack2(w.w);
    v[i * 8 + 0] = a.x, v[i * 8 + 1] = a.y, v[i * 8 + 2] = b.x, v[i * 8 + 3] = b.y;
    v[i * 8 + 4] = c.x, v[i * 8 + 5] = c.y, v[i * 8 + 6] = d.x, v[i * 8 + 7] = d.y;
  }
}
EVI_DEVINL void load_row32(const float* __restrict__ base, int lane, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(base + i * 256 + lane * 8));
    const float4 b = __ldg(reinterpret_cast<const float4*>(base + i * 256 + lane * 8 + 4));
    v[i * 8 + 0] = a.x, v[i * 8 + 1] = a.y, v[i * 8 + 2] = a.z, v[i * 8 + 3] = a.w;
    v[i * 8 + 4] = b.x, v[i * 8 + 5] = b.y, v[i * 8 + 6] = b.z, v[i * 8 + 7] = b.w;
  }
}
template <int FMT>
EVI_DEVINL void store_row16(uint16_t* __restrict__ base, int lane, const float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 w;
    w.x = Op16<FMT>::pack2(v[i * 8 + 0], v[i * 8 + 1]);
    w.y = Op16<FMT>::pack2(v[i * 8 + 2], v[i * 8 + 3]);
    w.z = Op16<FMT>::pack2(v[i * 8 + 4], v[i * 8 + 5]);
    w.w = Op16<FMT>::pack2(v[i * 8 + 6], v[i * 8 + 7]);
    *reinterpret_cast<uint4*>(base + i * 256 + lane * 8) = w;
  }
}

// struct context for one direction; returns nav. `x` receives sc.
EVI_DEVINL float struct_context(const StructParams& sp, const float* __restrict__ s_ws_t /*smem [S][D]*/,
                                const float* __restrict__ s_vec /*smem: bs, g, b, wn each [D]*/,
                                const float* __restrict__ sa, const float* __restrict__ sb, int half, int lane,
                                float (&x)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 b0 = *reinterpret_cast<const float4*>(s_vec + i * 256 + lane * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(s_vec + i * 256 + lane * 8 + 4);
    x[i * 8 + 0] = b0.x, x[i * 8 + 1] = b0.y, x[i * 8 + 2] = b0.z, x[i * 8 + 3] = b0.w;
    x[i * 8 + 4] = b1.x, x[i * 8 + 5] = b1.y, x[i * 8 + 6] = b1.z, x[i * 8 + 7] = b1.w;
  }
  for (int k = 0; k < 2 * half; ++k) {
    const float s = k < half ? sa[k] : sb[k - half];  // warp-uniform
    if (s == 0.f) continue;                            // most DDE features are exactly zero
    const float* w = s_ws_t + k * D;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 w0 = *reinterpret_cast<const float4*>(w + i * 256 + lane * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(w + i * 256 + lane * 8 + 4);
      x[i * 8 + 0] = fmaf(s, w0.x, x[i * 8 + 0]), x[i * 8 + 1] = fmaf(s, w0.y, x[i * 8 + 1]);
      x[i * 8 + 2] = fmaf(s, w0.z, x[i * 8 + 2]), x[i * 8 + 3] = fmaf(s, w0.w, x[i * 8 + 3]);
      x[i * 8 + 4] = fmaf(s, w1.x, x[i * 8 + 4]), x[i * 8 + 5] = fmaf(s, w1.y, x[i * 8 + 5]);
      x[i * 8 + 6] = fmaf(s, w1.z, x[i * 8 + 6]), x[i * 8 + 7] = fmaf(s, w1.w, x[i * 8 + 7]);
    }
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) sum += x[j];
  const float mean = warp_sum(sum) * (1.0f / D);
  float var = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float d = x[j] - mean;
    var = fmaf(d, d, var);
  }
  const float rstd = rsqrtf(warp_sum(var) * (1.0f / D) + kLnEps);
  float nav = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int col = i * 256 + lane * 8 + j;
      const float y = gelu_erf(fmaf((x[i * 8 + j] - mean) * rstd, s_vec[D + col], s_vec[2 * D + col]));
      x[i * 8 + j] = y;
      nav = fmaf(y, s_vec[3 * D + col], nav);
    }
  }
  return 1.0f / (1.0f + expf(-(warp_sum(nav) + sp.bn)));
}

constexpr int kEdgeWarps = 8;

template <int FMT>
__global__ void __launch_bounds__(kEdgeWarps * 32)
edge_operand_kernel(const int* __restrict__ head, const int* __restrict__ tail, const int* __restrict__ qid,
                    const int* __restrict__ rel_slot, long long num_edges,
                    const uint16_t* __restrict__ node_repr,   // [N, D]
                    const uint16_t* __restrict__ rel_repr,    // [R, D]
                    const float* __restrict__ gate_bias,      // [B, 2D]: sigmoid gate | tanh bias
                    const float* __restrict__ node_struct, int ns_dim, StructParams sp,
                    int do_fwd, int do_bwd, int num_nodes, int num_rel_rows, int num_questions,
                    unsigned* __restrict__ flags,
                    uint16_t* __restrict__ a_fwd, uint16_t* __restrict__ a_bwd,  // [E, 3D]
                    float* __restrict__ dist_fwd, float* __restrict__ dist_bwd) {
  extern __shared__ float smem_f[];
  float* s_ws_t = smem_f;                       // [S][D]
  float* s_vec = smem_f + sp.s_dim * D;         // bs | ln_g | ln_b | wn
  float* s_ns = s_vec + 4 * D;                  // per warp: 2 * ns_dim
  for (int i = threadIdx.x; i < sp.s_dim * D; i += blockDim.x) s_ws_t[i] = sp.ws_t[i];
  for (int i = threadIdx.x; i < D; i += blockDim.x) {
    s_vec[i] = sp.bs[i];
    s_vec[D + i] = sp.ln_g[i];
    s_vec[2 * D + i] = sp.ln_b[i];
    s_vec[3 * D + i] = sp.wn[i];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* my_ns = s_ns + warp * 2 * kMaxStruct;
  for (long long e = blockIdx.x * static_cast<long long>(kEdgeWarps) + warp; e < num_edges;
       e += static_cast<long long>(gridDim.x) * kEdgeWarps) {
    int h = __ldg(head + e), t = __ldg(tail + e), q = __ldg(qid + e), rs = __ldg(rel_slot + e);
    if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(q) >= static_cast<unsigned>(num_questions) ||
        static_cast<unsigned>(rs) >= static_cast<unsigned>(num_rel_rows)) {
      if (lane == 0) atomicOr(flags, kFlagOutOfRange);  // keep going on row 0: no wild reads
      h = t = q = rs = 0;
    }
    __syncwarp();
    if (lane < ns_dim) {
      my_ns[lane] = __ldg(node_struct + static_cast<size_t>(h) * ns_dim + lane);
      my_ns[kMaxStruct + lane] = __ldg(node_struct + static_cast<size_t>(t) * ns_dim + lane);
    }
    __syncwarp();
    float hv[32], tv[32], rv[32], x[32];
    load_row16<FMT>(node_repr + static_cast<size_t>(h) * D, lane, hv);
    load_row16<FMT>(node_repr + static_cast<size_t>(t) * D, lane, tv);
    load_row16<FMT>(rel_repr + static_cast<size_t>(rs) * D, lane, rv);
    {
      float gv[32];
      load_row32(gate_bias + static_cast<size_t>(q) * 2 * D, lane, gv);
#pragma unroll
      for (int j = 0; j < 32; ++j) rv[j] *= gv[j];
      load_row32(gate_bias + static_cast<size_t>(q) * 2 * D + D, lane, gv);
#pragma unroll
      for (int j = 0; j < 32; ++j) rv[j] += gv[j];  // rv = r_ctx
    }
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
      if (dir == 0 ? !do_fwd : !do_bwd) continue;
      const float* sa = dir == 0 ? my_ns : my_ns + kMaxStruct;
      const float* sb = dir == 0 ? my_ns + kMaxStruct : my_ns;
      uint16_t* arow = (dir == 0 ? a_fwd : a_bwd) + static_cast<size_t>(e) * 3 * D;
      const float nav = struct_context(sp, s_ws_t, s_vec, sa, sb, ns_dim, lane, x);
      store_row16<FMT>(arow + D, lane, x);
      float n2 = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float hh = dir == 0 ? hv[j] : tv[j], tt = dir == 0 ? tv[j] : hv[j];
        const float er = hh + rv[j] - tt;
        n2 = fmaf(er, er, n2);
        x[j] = er;
      }
      store_row16<FMT>(arow + 2 * D, lane, x);
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = hv[j] * rv[j] * tv[j] * nav;
      store_row16<FMT>(arow, lane, x);
      n2 = warp_sum(n2);
      if (lane == 0) (dir == 0 ? dist_fwd : dist_bwd)[e] = -sqrtf(n2);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Tail: LayerNorm (exact fp32 stats from the GEMM epilogue) + GELU + collapsed score head.
//   logit = w_o . (W_2 g + b_2) + b_o = (W_2^T w_o) . g + (w_o . b_2 + b_o)      retriever.py:482-483
// and the fwd/bwd softmax combine (retriever.py:369-381). One warp per edge.
// Optionally writes g (16-bit) for the feature path (state_net.4 as a second GEMM).
// ---------------------------------------------------------------------------------------------
struct TailParams {
  const float* ln_g;  // state_net.1.weight
  const float* ln_b;
  const float* v;     // W_2^T w_o
  float c0;           // w_o . b_2 + b_o
};

template <int FMT>
EVI_DEVINL float tail_one(const float* __restrict__ zrow, const float2* __restrict__ st, int n_blocks,
                          const TailParams& tp, int lane, uint16_t* __restrict__ g_out) {
  float sum = 0.f, sumsq = 0.f;
  for (int i = 0; i < n_blocks; ++i) {
    const float2 s = __ldg(st + i);
    sum += s.x, sumsq += s.y;
  }
  const float mean = sum * (1.0f / D);
  const float var = fmaxf(sumsq * (1.0f / D) - mean * mean, 0.f);
  const float rstd = rsqrtf(var + kLnEps);
  float z[32];
  load_row32(zrow, lane, z);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int col = i * 256 + lane * 8;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(tp.ln_g + col));
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(tp.ln_g + col + 4));
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(tp.ln_b + col));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(tp.ln_b + col + 4));
    const float4 v0 = __ldg(reinterpret_cast<const float4*>(tp.v + col));
    const float4 v1 = __ldg(reinterpret_cast<const float4*>(tp.v + col + 4));
    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float y = gelu_erf(fmaf((z[i * 8 + j] - mean) * rstd, gg[j], bb[j]));
      z[i * 8 + j] = y;
      acc = fmaf(y, vv[j], acc);
    }
  }
  if (g_out != nullptr) store_row16<FMT>(g_out, lane, z);
  return warp_sum(acc) + tp.c0;
}

template <int FMT>
__global__ void __launch_bounds__(256)
score_tail_kernel(const float* __restrict__ z_fwd, const float* __restrict__ z_bwd,
                  const float2* __restrict__ st_fwd, const float2* __restrict__ st_bwd, int n_blocks,
                  long long num_edges, TailParams tp, uint16_t* __restrict__ g_fwd, uint16_t* __restrict__ g_bwd,
                  float* __restrict__ logits, float* __restrict__ logits_fwd, float* __restrict__ logits_bwd) {
  const int lane = threadIdx.x & 31;
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (e >= num_edges) return;
  float lf = 0.f, lb = 0.f;
  const bool has_f = z_fwd != nullptr, has_b = z_bwd != nullptr;
  if (has_f)
    lf = tail_one<FMT>(z_fwd + static_cast<size_t>(e) * D, st_fwd + static_cast<size_t>(e) * n_blocks, n_blocks, tp,
                       lane, g_fwd ? g_fwd + static_cast<size_t>(e) * D : nullptr);
  if (has_b)
    lb = tail_one<FMT>(z_bwd + static_cast<size_t>(e) * D, st_bwd + static_cast<size_t>(e) * n_blocks, n_blocks, tp,
                       lane, g_bwd ? g_bwd + static_cast<size_t>(e) * D : nullptr);
  if (lane == 0) {
    if (has_f && logits_fwd) logits_fwd[e] = lf;
    if (has_b && logits_bwd) logits_bwd[e] = lb;
    if (logits) {
      float out;
      if (has_f && has_b) {  // softmax([lf, lb]) weighted sum
        const float m = fmaxf(lf, lb);
        const float ef = expf(lf - m), eb = expf(lb - m);
        const float inv = 1.0f / (ef + eb);
        out = (ef * inv) * lf + (eb * inv) * lb;
      } else {
        out = has_f ? lf : lb;
      }
      logits[e] = out;
    }
  }
}

// Feature path: logits from the materialised fp32 features (score_head on state_net output) and
// the softmax-weighted feature blend (retriever.py:377-380). One warp per edge.
__global__ void __launch_bounds__(256)
feature_blend_kernel(const float* __restrict__ feat_fwd, const float* __restrict__ feat_bwd,
                     const float* __restrict__ w_o, float b_o, long long num_edges,
                     float* __restrict__ edge_emb, float* __restrict__ logits, float* __restrict__ logits_fwd,
                     float* __restrict__ logits_bwd) {
  const int lane = threadIdx.x & 31;
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (e >= num_edges) return;
  float ff[32], fb[32], w[32];
  load_row32(w_o, lane, w);
  float lf = 0.f, lb = 0.f;
  if (feat_fwd) {
    load_row32(feat_fwd + static_cast<size_t>(e) * D, lane, ff);
#pragma unroll
    for (int j = 0; j < 32; ++j) lf = fmaf(ff[j], w[j], lf);
    lf = warp_sum(lf) + b_o;
  }
  if (feat_bwd) {
    load_row32(feat_bwd + static_cast<size_t>(e) * D, lane, fb);
#pragma unroll
    for (int j = 0; j < 32; ++j) lb = fmaf(fb[j], w[j], lb);
    lb = warp_sum(lb) + b_o;
  }
  float wf = 1.f, wb = 0.f, out = lf;
  if (feat_fwd && feat_bwd) {
    const float m = fmaxf(lf, lb);
    const float ef = expf(lf - m), eb = expf(lb - m);
    const float inv = 1.0f / (ef + eb);
    wf = ef * inv, wb = eb * inv;
    out = wf * lf + wb * lb;
  } else if (!feat_fwd) {
    wf = 0.f, wb = 1.f, out = lb;
  }
  if (edge_emb != nullptr) {
    float* o = edge_emb + static_cast<size_t>(e) * D;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float r[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        r[j] = (feat_fwd ? wf * ff[i * 8 + j] : 0.f) + (feat_bwd ? wb * fb[i * 8 + j] : 0.f);
      *reinterpret_cast<float4*>(o + i * 256 + lane * 8) = make_float4(r[0], r[1], r[2], r[3]);
      *reinterpret_cast<float4*>(o + i * 256 + lane * 8 + 4) = make_float4(r[4], r[5], r[6], r[7]);
    }
  }
  if (lane == 0) {
    if (logits) logits[e] = out;
    if (logits_fwd && feat_fwd) logits_fwd[e] = lf;
    if (logits_bwd && feat_bwd) logits_bwd[e] = lb;
  }
}

template <int FMT>
__global__ void convert_f32_to_16_kernel(const float* __restrict__ src, uint16_t* __restrict__ dst, long long n8) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n8) return;
  const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i);
  const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
  uint4 w;
  w.x = Op16<FMT>::pack2(a.x, a.y);
  w.y = Op16<FMT>::pack2(a.z, a.w);
  w.z = Op16<FMT>::pack2(b.x, b.y);
  w.w = Op16<FMT>::pack2(b.z, b.w);
  reinterpret_cast<uint4*>(dst)[i] = w;
}

// -------------------------------------------------------------------------------------------
// host-side dispatch on operand format
// -------------------------------------------------------------------------------------------
template <int FMT>
static int gather_rows_impl(const void* src, int src_is_f32, long long src_rows, const long long* ids,
                            const long long* sub_mask_ids, const float* sub_row, int num_rows, void* dst,
                            unsigned* flags, cudaStream_t stream) {
  if (num_rows <= 0) return kOk;
  const int warps = 8;
  const int grid = (num_rows + warps - 1) / warps;
  if (src_is_f32)
    gather_rows_kernel<FMT, true><<<grid, warps * 32, 0, stream>>>(src, src_rows, ids, sub_mask_ids, sub_row,
                                                                  num_rows, static_cast<uint16_t*>(dst), flags);
  else
    gather_rows_kernel<FMT, false><<<grid, warps * 32, 0, stream>>>(src, src_rows, ids, sub_mask_ids, sub_row,
                                                                   num_rows, static_cast<uint16_t*>(dst), flags);
  return check_launch();
}

template <int FMT>
static int linear_act_impl(const void* x16, int m, const void* w16, int n, int k, const float* bias,
                           const int* act_per_block, void* out16, float* out32, int num_sms,
                           cudaStream_t stream) {
  gemm::EpiBiasAct::Params p;
  p.bias = bias, p.out16 = out16, p.out32 = out32, p.ld16 = n, p.ld32 = n;
  for (int i = 0; i < 8; ++i) p.act[i] = act_per_block[i];
  return gemm::launch<FMT, gemm::EpiBiasAct>(x16, k, w16, k, m, n, k, p, num_sms, stream);
}

template <int FMT>
static int score_edges_impl(const evi_score_args& a, cudaStream_t stream) {
  const long long E = a.num_edges;
  if (E <= 0) return kOk;
  const int do_f = a.direction != EVI_DIR_BACKWARD, do_b = a.direction != EVI_DIR_FORWARD;
  const int ndir = do_f + do_b;
  const int n_blocks = D / gemm::BN;
  // workspace carve-up: A [ndir*E, 3D] 16-bit | z [ndir*E, D] 16-bit | stats [ndir*E, n_blocks] float2 |
  //                     (z is fp32) dist [ndir*E] f32 | (features) g [ndir*E, D] 16-bit | feat [ndir*E, D] f32
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    uint8_t* p = ws + off;
    off += (bytes + 255) & ~static_cast<size_t>(255);
    return p;
  };
  const size_t rows = static_cast<size_t>(ndir) * E;
  uint16_t* A = reinterpret_cast<uint16_t*>(take(rows * 3 * D * 2));
  float* Z = reinterpret_cast<float*>(take(rows * D * sizeof(float)));
  float2* ST = reinterpret_cast<float2*>(take(rows * n_blocks * sizeof(float2)));
  float* DIST = reinterpret_cast<float*>(take(rows * sizeof(float)));
  uint16_t* G = nullptr;
  float* FEAT = nullptr;
  if (a.want_features) {
    G = reinterpret_cast<uint16_t*>(take(rows * D * 2));
    FEAT = reinterpret_cast<float*>(take(rows * D * sizeof(float)));
  }
  if (off > a.workspace_bytes) return kErrWorkspace;
  uint16_t* A_f = A;
  uint16_t* A_b = A + (do_f ? static_cast<size_t>(E) * 3 * D : 0);
  float* dist_f = DIST;
  float* dist_b = DIST + (do_f ? E : 0);

  StructParams sp;
  sp.ws_t = a.struct_w_t, sp.bs = a.struct_b, sp.ln_g = a.struct_ln_g, sp.ln_b = a.struct_ln_b;
  sp.wn = a.nav_w, sp.bn = a.nav_b, sp.s_dim = 2 * a.ns_dim;
  if (a.ns_dim > kMaxStruct / 2 || a.ns_dim < 2) return kErrInvalidArg;
  if (a.flags == nullptr || a.num_nodes <= 0 || a.num_rel_rows <= 0 || a.num_questions <= 0) return kErrInvalidArg;
  const size_t smem = (static_cast<size_t>(sp.s_dim) * D + 4 * D + kEdgeWarps * 2 * kMaxStruct) * sizeof(float);
  auto ek = edge_operand_kernel<FMT>;
  if (cudaFuncSetAttribute(ek, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return kErrCuda;
  long long want = (E + kEdgeWarps - 1) / kEdgeWarps;
  const int grid = static_cast<int>(want < 4LL * a.num_sms ? want : 4LL * a.num_sms);
  ek<<<grid, kEdgeWarps * 32, smem, stream>>>(a.head32, a.tail32, a.qid32, a.rel_slot32, E,
                                             static_cast<const uint16_t*>(a.node_repr16),
                                             static_cast<const uint16_t*>(a.rel_repr16), a.gate_bias32,
                                             a.node_struct, a.ns_dim, sp, do_f, do_b, a.num_nodes, a.num_rel_rows,
                                             a.num_questions, a.flags, A_f, A_b, dist_f, dist_b);
  if (check_launch() != kOk) return kErrCuda;

  gemm::EpiPreNorm::Params zp;
  zp.bias = a.w1_bias, zp.w_dist = a.w1_dist_col, zp.dist = DIST, zp.z32 = Z, zp.stats = ST;
  zp.ldz = D, zp.n_blocks = n_blocks;
  if (a.ev_gemm_begin) cudaEventRecord(static_cast<cudaEvent_t>(a.ev_gemm_begin), stream);
  int rc = gemm::launch<FMT, gemm::EpiPreNorm>(A, 3 * D, a.w1_main16, 3 * D, static_cast<int>(rows), D, 3 * D, zp,
                                               a.num_sms, stream);
  if (a.ev_gemm_end) cudaEventRecord(static_cast<cudaEvent_t>(a.ev_gemm_end), stream);
  if (rc != kOk) return rc;

  TailParams tp;
  tp.ln_g = a.ln1_g, tp.ln_b = a.ln1_b, tp.v = a.head_v, tp.c0 = a.head_c0;
  const float* z_f = do_f ? Z : nullptr;
  const float* z_b = do_b ? Z + (do_f ? static_cast<size_t>(E) * D : 0) : nullptr;
  const float2* st_f = ST;
  const float2* st_b = ST + (do_f ? static_cast<size_t>(E) * n_blocks : 0);
  uint16_t* g_f = G;
  uint16_t* g_b = G ? G + (do_f ? static_cast<size_t>(E) * D : 0) : nullptr;
  const int tgrid = static_cast<int>((E + 7) / 8);
  if (!a.want_features) {
    score_tail_kernel<FMT><<<tgrid, 256, 0, stream>>>(z_f, z_b, st_f, st_b, n_blocks, E, tp, nullptr, nullptr,
                                                      a.logits, a.logits_fwd, a.logits_bwd);
    return check_launch();
  }
  score_tail_kernel<FMT><<<tgrid, 256, 0, stream>>>(z_f, z_b, st_f, st_b, n_blocks, E, tp, do_f ? g_f : nullptr,
                                                    do_b ? g_b : nullptr, nullptr, nullptr, nullptr);
  if (check_launch() != kOk) return kErrCuda;
  gemm::EpiBiasF32::Params fp;
  fp.bias = a.w2_bias, fp.out32 = FEAT, fp.ld32 = D;
  rc = gemm::launch<FMT, gemm::EpiBiasF32>(G, D, a.w2_16, D, static_cast<int>(rows), D, D, fp, a.num_sms, stream);
  if (rc != kOk) return rc;
  const float* f_f = do_f ? FEAT : nullptr;
  const float* f_b = do_b ? FEAT + (do_f ? static_cast<size_t>(E) * D : 0) : nullptr;
  feature_blend_kernel<<<tgrid, 256, 0, stream>>>(f_f, f_b, a.head_w, a.head_b, E, a.edge_embeddings, a.logits,
                                                  a.logits_fwd, a.logits_bwd);
  return check_launch();
}

}  // namespace evi

using namespace evi;

extern "C" {

size_t evi_score_workspace_bytes(long long num_edges, int direction, int want_features) {
  const size_t ndir = direction == EVI_DIR_BIDIRECTIONAL ? 2 : 1;
  const size_t rows = ndir * static_cast<size_t>(num_edges);
  auto al = [](size_t b) { return (b + 255) & ~static_cast<size_t>(255); };
  size_t total = al(rows * 3 * D * 2) + al(rows * D * sizeof(float)) + al(rows * (D / gemm::BN) * sizeof(float2)) +
                 al(rows * sizeof(float));
  if (want_features) total += al(rows * D * 2) + al(rows * D * sizeof(float));
  return total + 256;
}

int evi_gather_rows(const void* src, int src_is_f32, long long src_rows, const long long* ids,
                    const long long* sub_mask_ids, const float* sub_row, int num_rows, int dim, void* dst16,
                    int operand_format, unsigned* flags, cudaStream_t stream) {
  if (dim != D) return kErrUnsupported;
  if (operand_format == kFmtF16)
    return gather_rows_impl<kFmtF16>(src, src_is_f32, src_rows, ids, sub_mask_ids, sub_row, num_rows, dst16, flags,
                                     stream);
  if (operand_format == kFmtBF16)
    return gather_rows_impl<kFmtBF16>(src, src_is_f32, src_rows, ids, sub_mask_ids, sub_row, num_rows, dst16, flags,
                                      stream);
  return kErrInvalidArg;
}

int evi_convert_f32(const float* src, void* dst16, long long count, int operand_format, cudaStream_t stream) {
  if (count <= 0) return kOk;
  if (count % 8 != 0) return kErrInvalidArg;
  const long long n8 = count / 8;
  const int grid = static_cast<int>((n8 + 255) / 256);
  if (operand_format == kFmtF16)
    convert_f32_to_16_kernel<kFmtF16><<<grid, 256, 0, stream>>>(src, static_cast<uint16_t*>(dst16), n8);
  else if (operand_format == kFmtBF16)
    convert_f32_to_16_kernel<kFmtBF16><<<grid, 256, 0, stream>>>(src, static_cast<uint16_t*>(dst16), n8);
  else
    return kErrInvalidArg;
  return check_launch();
}

int evi_project(const void* x16, int num_rows, const void* w16, int out_dim, int in_dim, const float* bias,
                const int* act_per_block, void* out16, float* out32, int operand_format, int num_sms,
                cudaStream_t stream) {
  if (out_dim % gemm::BN != 0 || out_dim / gemm::BN > 8 || in_dim % gemm::BK != 0) return kErrUnsupported;
  if (operand_format == kFmtF16)
    return linear_act_impl<kFmtF16>(x16, num_rows, w16, out_dim, in_dim, bias, act_per_block, out16, out32, num_sms,
                                    stream);
  if (operand_format == kFmtBF16)
    return linear_act_impl<kFmtBF16>(x16, num_rows, w16, out_dim, in_dim, bias, act_per_block, out16, out32,
                                     num_sms, stream);
  return kErrInvalidArg;
}

int evi_score_edges(const evi_score_args* args, cudaStream_t stream) {
  if (args == nullptr) return kErrInvalidArg;
  if (args->emb_dim != D || args->hidden_dim != D) return kErrUnsupported;
  if (args->operand_format == kFmtF16) return score_edges_impl<kFmtF16>(*args, stream);
  if (args->operand_format == kFmtBF16) return score_edges_impl<kFmtBF16>(*args, stream);
  return kErrInvalidArg;
}

}  // extern "C"
