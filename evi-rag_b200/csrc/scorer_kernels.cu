// Triple scoring: Retriever._prepare_edge_inputs / _score_edges /
// _combine_directional_outputs of src/models/components/retriever.py:403-495, 369-381.
//
// Dense contractions run on tcgen05 (gemm_tcgen05.cuh); the per-edge elementwise work that builds
// the GEMM operand rows ([h*r*t*nav | GELU(LN(struct)) | h+r-t]) and the LayerNorm/GELU/score-head
// tail run in the warp-per-edge kernels below with 128-bit loads of the 1024-d rows.
#include <cstdio>
#include <cstdlib>

#include "gemm_tcgen05.cuh"
#include "gemm2_tcgen05.cuh"
#include "evi_rag_b200.h"

namespace evi {

constexpr int D = 1024;        // emb_dim == hidden_dim (configs/model/retriever_module.yaml:10-11)
constexpr int kMaxNodeStruct = 18;  // 2*(1+4+4)
constexpr float kLnEps = 1e-5f;

static inline int check_launch() { return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda; }

// fp32-output GEMM: the CTA-pair kernel (256x256 tiles, tcgen05 cta_group::2) unless EVI_GEMM_1SM is set
template <int FMT>
static int gemm_f32(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                    const gemm::EpiF32Tma::Params& ep, int num_sms, cudaStream_t stream, const int* m_dev, float* out32,
                    int ldc, const gemm::SplitK& sk = gemm::make_split(-1, -1)) {
  static const bool one_sm = getenv("EVI_GEMM_1SM") != nullptr;
  if (one_sm || num_sms < 2)
    return gemm::launch<FMT, gemm::EpiF32Tma>(A, lda, B, ldb, M, N, K, ep, num_sms, stream, m_dev, out32, ldc, sk);
  return gemm2::launch<FMT, gemm::EpiF32Tma>(A, lda, B, ldb, M, N, K, ep, num_sms, stream, m_dev, out32, ldc, sk);
}

// ---------------------------------------------------------------------------------------------
// Row gather / convert: fp32 or 16-bit source rows -> dense 16-bit GEMM operand rows.
// ids == nullptr: identity. Rows with sub_mask_ids[row] == 0 take `sub_row` (fp32) instead
// (the non-text entity embedding, retriever.py:499-506 — tanh(W x + b) is applied by the GEMM
// epilogue, so substituting the *input* row is equivalent to substituting the projected row).
// SPLIT: destination rows are [hi | lo] (lo = x - hi at column D, gemm::SplitK); a 16-bit source then holds
// [hi | lo] rows as well, an fp32 source (or the substitute row) is split on the fly.
// rows_dev (optional): device-side row count, e.g. the number of distinct embedding ids of a batch.
// ---------------------------------------------------------------------------------------------
template <int FMT, bool SRC_F32, bool SPLIT>
__global__ void gather_rows_kernel(const void* __restrict__ src, long long src_rows, int src_ld,
                                   const long long* __restrict__ ids, const long long* __restrict__ sub_mask_ids,
                                   const float* __restrict__ sub_row, int num_rows, const int* __restrict__ rows_dev,
                                   uint16_t* __restrict__ dst, int dst_ld, unsigned* __restrict__ flags) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (rows_dev != nullptr) num_rows = min(num_rows, __ldg(rows_dev));
  if (row >= num_rows) return;
  const int lane = threadIdx.x & 31;
  long long sid = ids ? __ldg(ids + row) : row;
  const bool substitute = sub_mask_ids != nullptr && __ldg(sub_mask_ids + row) == 0;
  if (!substitute && (sid < 0 || sid >= src_rows)) {
    if (lane == 0) atomicOr(flags, kFlagOutOfRange);
    sid = 0;
  }
  uint16_t* out = dst + static_cast<size_t>(row) * dst_ld;
#pragma unroll
  for (int i = 0; i < D / 256; ++i) {
    const int col = i * 256 + lane * 8;
    if (substitute || SRC_F32) {
      const float* sp = substitute ? sub_row : reinterpret_cast<const float*>(src) + static_cast<size_t>(sid) * src_ld;
      const float4 a = __ldg(reinterpret_cast<const float4*>(sp + col));
      const float4 b = __ldg(reinterpret_cast<const float4*>(sp + col + 4));
      const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      uint4 w;
      w.x = Op16<FMT>::pack2(v[0], v[1]);
      w.y = Op16<FMT>::pack2(v[2], v[3]);
      w.z = Op16<FMT>::pack2(v[4], v[5]);
      w.w = Op16<FMT>::pack2(v[6], v[7]);
      *reinterpret_cast<uint4*>(out + col) = w;
      if (SPLIT) {
        const float2 h0 = Op16<FMT>::unpack2(w.x), h1 = Op16<FMT>::unpack2(w.y), h2 = Op16<FMT>::unpack2(w.z),
                     h3 = Op16<FMT>::unpack2(w.w);
        uint4 l;
        l.x = Op16<FMT>::pack2(v[0] - h0.x, v[1] - h0.y);
        l.y = Op16<FMT>::pack2(v[2] - h1.x, v[3] - h1.y);
        l.z = Op16<FMT>::pack2(v[4] - h2.x, v[5] - h2.y);
        l.w = Op16<FMT>::pack2(v[6] - h3.x, v[7] - h3.y);
        *reinterpret_cast<uint4*>(out + D + col) = l;
      }
    } else {
      const uint16_t* sp = reinterpret_cast<const uint16_t*>(src) + static_cast<size_t>(sid) * src_ld;
      *reinterpret_cast<uint4*>(out + col) = ldg_nc_v4(sp + col);
      if (SPLIT) *reinterpret_cast<uint4*>(out + D + col) = ldg_nc_v4(sp + D + col);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Per-node half of struct_proj.0 (retriever.py:486-495 applied to retriever.py:509-517):
//   W_s [s_fwd ; s_bwd-half] splits by node:  W_s [ns[a], ns[b]] + b_s = U[a] + V[b]
//   U[n] = W_s[:, :ns] ns[n] + b_s,   V[n] = W_s[:, ns:] ns[n]            -> uv [N, 2D] fp32
// so the per-edge kernel adds two gathered rows instead of 2*ns FMAs per element.
// ---------------------------------------------------------------------------------------------
constexpr int kUvNodes = 8;  // nodes per CTA: every weight row fetched from L1 feeds 8 nodes

__global__ void __launch_bounds__(128)
node_struct_proj_kernel(const float* __restrict__ node_struct, int num_nodes, int ns_dim,
                        const float* __restrict__ ws_t /*[2*ns_dim, D]*/, const float* __restrict__ bs,
                        float* __restrict__ uv) {
  __shared__ float s_ns[kUvNodes][kMaxNodeStruct];
  const int n0 = blockIdx.x * kUvNodes;
  for (int i = threadIdx.x; i < kUvNodes * ns_dim; i += blockDim.x) {
    const int n = n0 + i / ns_dim;
    s_ns[i / ns_dim][i % ns_dim] = n < num_nodes ? __ldg(node_struct + static_cast<size_t>(n) * ns_dim + i % ns_dim) : 0.f;
  }
  __syncthreads();
  const int c = threadIdx.x * 8;  // 8 columns of U and the same 8 of V
  float u[kUvNodes][8], v[kUvNodes][8];
  {
    float b[8];
    ldg_v8_keep(bs + c, b);
#pragma unroll
    for (int n = 0; n < kUvNodes; ++n)
#pragma unroll
      for (int j = 0; j < 8; ++j) u[n][j] = b[j], v[n][j] = 0.f;
  }
  for (int k = 0; k < ns_dim; ++k) {
    float wu[8], wv[8];
    ldg_v8_keep(ws_t + static_cast<size_t>(k) * D + c, wu);
    ldg_v8_keep(ws_t + static_cast<size_t>(ns_dim + k) * D + c, wv);
#pragma unroll
    for (int n = 0; n < kUvNodes; ++n) {
      const float sv = s_ns[n][k];  // broadcast
#pragma unroll
      for (int j = 0; j < 8; ++j) u[n][j] = fmaf(sv, wu[j], u[n][j]), v[n][j] = fmaf(sv, wv[j], v[n][j]);
    }
  }
#pragma unroll
  for (int n = 0; n < kUvNodes; ++n) {
    if (n0 + n >= num_nodes) break;
    float* o = uv + static_cast<size_t>(n0 + n) * 2 * D + c;
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(o), "f"(u[n][0]), "f"(u[n][1]),
                 "f"(u[n][2]), "f"(u[n][3]), "f"(u[n][4]), "f"(u[n][5]), "f"(u[n][6]), "f"(u[n][7])
                 : "memory");
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(o + D), "f"(v[n][0]), "f"(v[n][1]),
                 "f"(v[n][2]), "f"(v[n][3]), "f"(v[n][4]), "f"(v[n][5]), "f"(v[n][6]), "f"(v[n][7])
                 : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Edge operand kernel: one warp per edge, both directions.
//   x_f = U[h] + V[t], x_b = U[t] + V[h]                                 (struct_proj.0, see above)
//   sc  = GELU(LN(x)) ; nav = sigmoid(w_n . sc + b_n)                    retriever.py:486-495
//   r_ctx = rel * gate[q] + bias[q]                                      retriever.py:464
//   iv  = h * r_ctx * t * nav ; err = h + r_ctx - t ; dist = -||err||    retriever.py:472-478
// Writes operand rows A_dir[e] = [iv | sc | err] (16-bit, 3*D) and dist_dir[e].
// Lane l owns columns {256*i + 8*l .. +7}, i = 0..3  (128-bit loads / stores).
// ---------------------------------------------------------------------------------------------
struct StructParams {
  const float* ln_g;    // [D] struct_proj.1.weight
  const float* ln_b;    // [D]
  const float* wn;      // [D] struct_gate_net.0.weight
  float bn;
  // [D] or null: a batch-constant row c subtracted from sc before it is rounded to 16 bits. W_sc sc = W_sc (sc - c) +
  // W_sc c, and the caller adds W_sc c (fp32 weights) to the pair bias: the rounding error of the operand scales with
  // |sc - c| instead of |sc| (most edges of a batch have similar structure contexts), at no tensor cost.
  const float* center;
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
    float t[8];
    ldg_v8_stream(base + i * 256 + lane * 8, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[i * 8 + j] = t[j];
  }
}
// one [hi | lo] row (2D elements): the 16-bit value at base[c] and its residual at base[D + c] — the operand of the
// three-pass state_net.4 contraction of the features path
template <int FMT>
EVI_DEVINL void store_row16(uint16_t* __restrict__ base, int lane, const float2 (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 w;
    w.x = Op16<FMT>::pack2(v[i * 4 + 0].x, v[i * 4 + 0].y);
    w.y = Op16<FMT>::pack2(v[i * 4 + 1].x, v[i * 4 + 1].y);
    w.z = Op16<FMT>::pack2(v[i * 4 + 2].x, v[i * 4 + 2].y);
    w.w = Op16<FMT>::pack2(v[i * 4 + 3].x, v[i * 4 + 3].y);
    *reinterpret_cast<uint4*>(base + i * 256 + lane * 8) = w;
    uint4 l;
    l.x = Op16<FMT>::pack2(sub2(v[i * 4 + 0], Op16<FMT>::unpack2(w.x)));
    l.y = Op16<FMT>::pack2(sub2(v[i * 4 + 1], Op16<FMT>::unpack2(w.y)));
    l.z = Op16<FMT>::pack2(sub2(v[i * 4 + 2], Op16<FMT>::unpack2(w.z)));
    l.w = Op16<FMT>::pack2(sub2(v[i * 4 + 3], Op16<FMT>::unpack2(w.w)));
    *reinterpret_cast<uint4*>(base + D + i * 256 + lane * 8) = l;
  }
}
template <int FMT>
EVI_DEVINL void unpack8(const uint4& w, float (&v)[8]) {
  const float2 a = Op16<FMT>::unpack2(w.x), b = Op16<FMT>::unpack2(w.y), c = Op16<FMT>::unpack2(w.z),
               d = Op16<FMT>::unpack2(w.w);
  v[0] = a.x, v[1] = a.y, v[2] = b.x, v[3] = b.y, v[4] = c.x, v[5] = c.y, v[6] = d.x, v[7] = d.y;
}
template <int FMT>
EVI_DEVINL void unpack8(const uint4& w, float2 (&v)[4]) {
  v[0] = Op16<FMT>::unpack2(w.x), v[1] = Op16<FMT>::unpack2(w.y), v[2] = Op16<FMT>::unpack2(w.z),
  v[3] = Op16<FMT>::unpack2(w.w);
}
template <int FMT>
EVI_DEVINL uint4 pack8(const float (&v)[8]) {
  uint4 w;
  w.x = Op16<FMT>::pack2(v[0], v[1]);
  w.y = Op16<FMT>::pack2(v[2], v[3]);
  w.z = Op16<FMT>::pack2(v[4], v[5]);
  w.w = Op16<FMT>::pack2(v[6], v[7]);
  return w;
}
// streaming 128-bit store: operand rows are written once and read once by the GEMM's TMA
EVI_DEVINL void st_cs_v4(void* p, const uint4& w) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w)
               : "memory");
}

// one 8-column group of an operand row: the 16-bit value at row[c], and with SPLIT its residual at row[D + c]
template <int FMT, bool SPLIT>
EVI_DEVINL void store_operand8(uint16_t* __restrict__ row, int c, const float (&v)[8]) {
  const uint4 w = pack8<FMT>(v);
  st_cs_v4(row + c, w);
  if constexpr (SPLIT) {
    float h[8], l[8];
    unpack8<FMT>(w, h);
#pragma unroll
    for (int j = 0; j < 8; ++j) l[j] = v[j] - h[j];
    st_cs_v4(row + D + c, pack8<FMT>(l));
  }
}
// struct context of one direction: writes sc (16-bit) to `sc_out`, returns nav.
// ua / vb: the U row of the first node and the V row of the second.
#ifndef EVI_SC_GELU
#define EVI_SC_GELU gelu_fast5  // sc is rounded to 16 bits on the spot; -DEVI_SC_GELU=gelu_fast for the degree-7 fit
#endif
template <bool SPLIT>
EVI_DEVINL float sc_gelu(float x) {
  if constexpr (SPLIT) return gelu_fast(x);  // [hi | lo] rows keep ~22 bits: the degree-7 fit (4.8e-7)
  else return EVI_SC_GELU(x);
}
template <int FMT, bool SPLIT>
EVI_DEVINL float struct_context(const float* __restrict__ ua, const float* __restrict__ vb,
                                const StructParams& sp, int lane, uint16_t* __restrict__ sc_out) {
  float x[32];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float u[8], v[8];
    ldg_v8_stream(ua + c, u);
    ldg_v8_stream(vb + c, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[i * 8 + j] = u[j] + v[j];
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) sum += x[j];
  const float mean = warp_sum(sum) * (1.0f / D);
  float var = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    x[j] -= mean;  // centred in place: the normalisation below is one multiply
    var = fmaf(x[j], x[j], var);
  }
  const float rstd = rsqrtf(warp_sum(var) * (1.0f / D) + kLnEps);
  float nav = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float g[8], b[8], w[8], y[8];
    ldg_v8_keep(sp.ln_g + c, g);  // 12 KB of parameters: L1-resident, no shared memory (a CTA with static
    ldg_v8_keep(sp.ln_b + c, b);  // shared memory does not fit beside the GEMM CTA's carve-out)
    ldg_v8_keep(sp.wn + c, w);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      y[j] = sc_gelu<SPLIT>(fmaf(x[i * 8 + j] * rstd, g[j], b[j]));
      nav = fmaf(y[j], w[j], nav);
    }
    if (sp.center != nullptr) {
      float cen[8];
      ldg_v8_keep(sp.center + c, cen);
#pragma unroll
      for (int j = 0; j < 8; ++j) y[j] -= cen[j];
    }
    store_operand8<FMT, SPLIT>(sc_out, c, y);
  }
  return 1.0f / (1.0f + __expf(-(warp_sum(nav) + sp.bn)));
}

// Both directions of one edge in one pass over the parameter vectors: x_f = U[h] + V[t], x_b = U[t] + V[h].
// ln_g / ln_b / w_n are read once per edge instead of once per direction (the kernel is bound by L1 throughput).
template <int FMT, bool SPLIT>
EVI_DEVINL void struct_context_pair(const float* __restrict__ uv_h, const float* __restrict__ uv_t,
                                    const StructParams& sp, int lane, uint16_t* __restrict__ sc_f,
                                    uint16_t* __restrict__ sc_b, float& nav_f, float& nav_b) {
  float xf[32], xb[32];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float uh[8], vt[8], ut[8], vh[8];
    ldg_v8_stream(uv_h + c, uh);
    ldg_v8_stream(uv_t + D + c, vt);
    ldg_v8_stream(uv_t + c, ut);
    ldg_v8_stream(uv_h + D + c, vh);
#pragma unroll
    for (int j = 0; j < 8; ++j) xf[i * 8 + j] = uh[j] + vt[j], xb[i * 8 + j] = ut[j] + vh[j];
  }
  float sum_f = 0.f, sum_b = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) sum_f += xf[j], sum_b += xb[j];
  const float mean_f = warp_sum(sum_f) * (1.0f / D), mean_b = warp_sum(sum_b) * (1.0f / D);
  float var_f = 0.f, var_b = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    xf[j] -= mean_f, xb[j] -= mean_b;  // centred in place: the normalisation below is one multiply
    var_f = fmaf(xf[j], xf[j], var_f);
    var_b = fmaf(xb[j], xb[j], var_b);
  }
  const float rstd_f = rsqrtf(warp_sum(var_f) * (1.0f / D) + kLnEps);
  const float rstd_b = rsqrtf(warp_sum(var_b) * (1.0f / D) + kLnEps);
  float acc_f = 0.f, acc_b = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float g[8], b[8], w[8], yf[8], yb[8];
    ldg_v8_keep(sp.ln_g + c, g);
    ldg_v8_keep(sp.ln_b + c, b);
    ldg_v8_keep(sp.wn + c, w);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      yf[j] = sc_gelu<SPLIT>(fmaf(xf[i * 8 + j] * rstd_f, g[j], b[j]));
      yb[j] = sc_gelu<SPLIT>(fmaf(xb[i * 8 + j] * rstd_b, g[j], b[j]));
      acc_f = fmaf(yf[j], w[j], acc_f);
      acc_b = fmaf(yb[j], w[j], acc_b);
    }
    if (sp.center != nullptr) {
      float cen[8];
      ldg_v8_keep(sp.center + c, cen);
#pragma unroll
      for (int j = 0; j < 8; ++j) yf[j] -= cen[j], yb[j] -= cen[j];
    }
    store_operand8<FMT, SPLIT>(sc_f, c, yf);
    store_operand8<FMT, SPLIT>(sc_b, c, yb);
  }
  nav_f = 1.0f / (1.0f + __expf(-(warp_sum(acc_f) + sp.bn)));
  nav_b = 1.0f / (1.0f + __expf(-(warp_sum(acc_b) + sp.bn)));
}

constexpr int kEdgeWarps = 8;
#ifndef EVI_OPERAND_CTAS
#define EVI_OPERAND_CTAS 2  // 128 registers: both directions' rows stay in registers (struct_context_pair)
#endif

// Operand rows of one chunk of edges (the inputs of the two per-edge contractions of state_net.0):
//   m16  [E, ld]    h * r_ctx * t            (shared by both directions: iv_dir = nav_dir * m)
//   sc16 [nd*E, ld] struct context (minus the batch centre), forward rows first, then backward rows
//   scal [E, 4]     nav_f, nav_b, dist_f, dist_b
// ld = D, or 2D with SPLIT ([hi | lo] rows for the three-pass contraction of the strict mode; the node / relation
// rows are then read as hi + lo too).
// head / tail index the per-node rows U | V; head_rep / tail_rep the rows of node_repr (one per distinct embedding id).
template <int FMT, bool BIDIR = false, bool SPLIT = false>
__global__ void __launch_bounds__(kEdgeWarps * 32, EVI_OPERAND_CTAS)
edge_operand_kernel(const int* __restrict__ head, const int* __restrict__ tail, const int* __restrict__ head_rep,
                    const int* __restrict__ tail_rep, const int* __restrict__ qid,
                    const int* __restrict__ rel_slot, long long num_edges,
                    const uint16_t* __restrict__ node_repr, int node_ld,   // [rep rows, node_ld]
                    const uint16_t* __restrict__ rel_repr, int rel_ld,     // [R, rel_ld]
                    const float* __restrict__ gate_bias,      // [B, 2D]: sigmoid gate | tanh bias
                    const float* __restrict__ node_uv,        // [N, 2D]: U | V (node_struct_proj_kernel)
                    StructParams sp, int do_fwd, int do_bwd, int num_nodes, int num_rep_rows, int num_rel_rows,
                    int num_questions, unsigned* __restrict__ flags,
                    uint16_t* __restrict__ m16, uint16_t* __restrict__ sc_fwd, uint16_t* __restrict__ sc_bwd,
                    float4* __restrict__ scal) {
  constexpr int kRowLd = SPLIT ? 2 * D : D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long stride = static_cast<long long>(gridDim.x) * kEdgeWarps;
  for (long long e = blockIdx.x * static_cast<long long>(kEdgeWarps) + warp; e < num_edges; e += stride) {
    int h = __ldg(head + e), t = __ldg(tail + e), q = __ldg(qid + e), rs = __ldg(rel_slot + e);
    int hr = __ldg(head_rep + e), tr = __ldg(tail_rep + e);
    if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(hr) >= static_cast<unsigned>(num_rep_rows) ||
        static_cast<unsigned>(tr) >= static_cast<unsigned>(num_rep_rows) ||
        static_cast<unsigned>(q) >= static_cast<unsigned>(num_questions) ||
        static_cast<unsigned>(rs) >= static_cast<unsigned>(num_rel_rows)) {
      if (lane == 0) atomicOr(flags, kFlagOutOfRange);  // keep going on row 0: no wild reads
      h = t = q = rs = hr = tr = 0;
    }
    const float* uv_h = node_uv + static_cast<size_t>(h) * 2 * D;
    const float* uv_t = node_uv + static_cast<size_t>(t) * 2 * D;
    float nav_f = 0.f, nav_b = 0.f;
    if (BIDIR || (do_fwd && do_bwd))  // BIDIR: both directions, known at compile time (the default mode)
      struct_context_pair<FMT, SPLIT>(uv_h, uv_t, sp, lane, sc_fwd + static_cast<size_t>(e) * kRowLd,
                                      sc_bwd + static_cast<size_t>(e) * kRowLd, nav_f, nav_b);
    else if (do_fwd)
      nav_f = struct_context<FMT, SPLIT>(uv_h, uv_t + D, sp, lane, sc_fwd + static_cast<size_t>(e) * kRowLd);
    else if (do_bwd)
      nav_b = struct_context<FMT, SPLIT>(uv_t, uv_h + D, sp, lane, sc_bwd + static_cast<size_t>(e) * kRowLd);

    const uint16_t* hp = node_repr + static_cast<size_t>(hr) * node_ld;
    const uint16_t* tp = node_repr + static_cast<size_t>(tr) * node_ld;
    const uint16_t* rp = rel_repr + static_cast<size_t>(rs) * rel_ld;
    const float* gp = gate_bias + static_cast<size_t>(q) * 2 * D;
    uint16_t* mrow = m16 + static_cast<size_t>(e) * kRowLd;
    float n2f = 0.f, n2b = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = i * 256 + lane * 8;
      float hv[8], tv[8], rv[8], g[8], b[8];
      unpack8<FMT>(ldg_nc_v4(hp + c), hv);
      unpack8<FMT>(ldg_nc_v4(tp + c), tv);
      unpack8<FMT>(ldg_nc_v4(rp + c), rv);
      if constexpr (SPLIT) {  // the residual halves of the projected rows
        float lo[8];
        unpack8<FMT>(ldg_nc_v4(hp + D + c), lo);
#pragma unroll
        for (int j = 0; j < 8; ++j) hv[j] += lo[j];
        unpack8<FMT>(ldg_nc_v4(tp + D + c), lo);
#pragma unroll
        for (int j = 0; j < 8; ++j) tv[j] += lo[j];
        unpack8<FMT>(ldg_nc_v4(rp + D + c), lo);
#pragma unroll
        for (int j = 0; j < 8; ++j) rv[j] += lo[j];
      }
      ldg_v8_keep(gp + c, g);
      ldg_v8_keep(gp + D + c, b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        rv[j] = fmaf(rv[j], g[j], b[j]);  // r_ctx (retriever.py:464)
        const float d = hv[j] - tv[j];
        const float ef = rv[j] + d, eb = rv[j] - d;  // err of either direction (retriever.py:477)
        n2f = fmaf(ef, ef, n2f);
        n2b = fmaf(eb, eb, n2b);
        rv[j] = hv[j] * rv[j] * tv[j];  // symmetric DistMult term (retriever.py:472)
      }
      store_operand8<FMT, SPLIT>(mrow, c, rv);
    }
    n2f = warp_sum(n2f);
    n2b = warp_sum(n2b);
    if (lane == 0) scal[e] = make_float4(nav_f, nav_b, -sqrtf(n2f), -sqrtf(n2b));
  }
}

// ---------------------------------------------------------------------------------------------
// Batch centre of the structure context (StructParams::center): the mean of sc over a fixed-stride sample of the
// batch's edges (both directions). Any row would do for correctness — W_sc c is added back exactly — the mean is
// the one that makes |sc - c| smallest. Deterministic: per-warp register sums, fixed-order reductions, no atomics.
//   sc_center_partial_kernel: grid x 8 warps, partial [grid, D]
//   sc_center_finish_kernel : center [D] = sum of partials / count
//   sc_center_bias_kernel   : bias_out [H] = b1 + W_sc (fp32) center     (folded into Rq's bias)
// ---------------------------------------------------------------------------------------------
constexpr int kCenterCtas = 256;
constexpr int kCenterSample = 4096;  // edges sampled per batch (each contributes its forward and backward rows)

__global__ void __launch_bounds__(256)
sc_center_partial_kernel(const int* __restrict__ head, const int* __restrict__ tail, long long num_edges,
                         long long sample, const float* __restrict__ node_uv, int num_nodes,
                         const float* __restrict__ ln_g, const float* __restrict__ ln_b, int do_fwd, int do_bwd,
                         float* __restrict__ partial) {
  __shared__ float red[8][D];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = 0.f;
  const long long total_warps = static_cast<long long>(gridDim.x) * 8;
  for (long long i = blockIdx.x * 8LL + warp; i < sample; i += total_warps) {
    const long long e = sample >= num_edges ? i : (i * num_edges) / sample;
    int h = __ldg(head + e), t = __ldg(tail + e);
    if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes))
      h = t = 0;  // flagged by the operand kernel
    for (int dir = 0; dir < 2; ++dir) {
      if (dir == 0 ? !do_fwd : !do_bwd) continue;
      const float* ua = node_uv + static_cast<size_t>(dir == 0 ? h : t) * 2 * D;
      const float* vb = node_uv + static_cast<size_t>(dir == 0 ? t : h) * 2 * D + D;
      float x[32];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int c = k * 256 + lane * 8;
        float u[8], v[8];
        ldg_v8_keep(ua + c, u);
        ldg_v8_keep(vb + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) x[k * 8 + j] = u[j] + v[j];
      }
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) sum += x[j];
      const float mean = warp_sum(sum) * (1.0f / D);
      float var = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        x[j] -= mean;
        var = fmaf(x[j], x[j], var);
      }
      const float rstd = rsqrtf(warp_sum(var) * (1.0f / D) + kLnEps);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int c = k * 256 + lane * 8;
        float g[8], b[8];
        ldg_v8_keep(ln_g + c, g);
        ldg_v8_keep(ln_b + c, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[k * 8 + j] += gelu_fast(fmaf(x[k * 8 + j] * rstd, g[j], b[j]));
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp][k * 256 + lane * 8 + j] = acc[k * 8 + j];
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += red[w][c];
    partial[static_cast<size_t>(blockIdx.x) * D + c] = v;
  }
}

__global__ void __launch_bounds__(256)
sc_center_finish_kernel(const float* __restrict__ partial, int num_partials, float inv_count, float* __restrict__ center) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  float v = 0.f;
  for (int b = 0; b < num_partials; ++b) v += partial[static_cast<size_t>(b) * D + c];
  center[c] = v * inv_count;
}

// one warp per output row j: bias_out[j] = b1[j] + sum_k w_sc32[j, k] * center[k]
__global__ void __launch_bounds__(256)
sc_center_bias_kernel(const float* __restrict__ w_sc32, const float* __restrict__ center, const float* __restrict__ b1,
                      float* __restrict__ bias_out) {
  const int j = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (j >= D) return;
  const float* wr = w_sc32 + static_cast<size_t>(j) * D;
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = k * 256 + lane * 8;
    float w[8], cv[8];
    ldg_v8_stream(wr + c, w);
    ldg_v8_keep(center + c, cv);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc = fmaf(w[i], cv[i], acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) bias_out[j] = (b1 != nullptr ? __ldg(b1 + j) : 0.f) + acc;
}

// ---------------------------------------------------------------------------------------------
// Distinct embedding ids of a batch. tanh(W_e x + b) and P = W_err node_repr depend only on the embedding id
// (retriever.py:497-507): 35 % of a WebQSP-shaped batch's nodes are non-text (one shared row) and the text ids of
// 128 questions repeat, so the projection and the P contraction run once per distinct id.
//   slot_map [table_rows] int32, all -1 between calls (claim -> lookup -> reset inside one call).
//   Slot numbering follows an atomic counter (arbitrary); rows are computed independently, so nothing depends on it.
// ---------------------------------------------------------------------------------------------
__global__ void unique_claim_kernel(const long long* __restrict__ ids, int n, int* __restrict__ slot_map,
                                    long long table_rows, long long* __restrict__ uniq_ids,
                                    int* __restrict__ num_unique, unsigned* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  long long id = __ldg(ids + i);
  if (id < 0 || id >= table_rows) {
    atomicOr(flags, kFlagOutOfRange);
    id = 0;
  }
  // a third of a WebQSP-shaped batch shares id 0: a plain (volatile) look first, so that only the first arrivals
  // reach the atomic instead of tens of thousands of CAS operations serialising on one address
  if (*reinterpret_cast<volatile int*>(slot_map + id) == -1 && atomicCAS(slot_map + id, -1, -2) == -1) {
    const int s = atomicAdd(num_unique, 1);
    uniq_ids[s] = id;
    slot_map[id] = s;  // read by the next kernel only
  }
}
__global__ void unique_lookup_kernel(const long long* __restrict__ ids, int n, const int* __restrict__ slot_map,
                                     long long table_rows, int* __restrict__ node_slot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  long long id = __ldg(ids + i);
  if (id < 0 || id >= table_rows) id = 0;
  node_slot[i] = slot_map[id];
}
__global__ void unique_reset_kernel(const long long* __restrict__ uniq_ids, const int* __restrict__ num_unique, int n,
                                    int* __restrict__ slot_map) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || i >= __ldg(num_unique)) return;
  slot_map[uniq_ids[i]] = -1;
}
// per-edge rows of node_repr / P (through node_slot) and the int32 relation row
__global__ void edge_rep_index_kernel(const int* __restrict__ head, const int* __restrict__ tail, long long num_edges,
                                      const int* __restrict__ node_slot, int num_nodes,
                                      const long long* __restrict__ edge_attr, int* __restrict__ head_rep,
                                      int* __restrict__ tail_rep, int* __restrict__ rel_slot) {
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (e >= num_edges) return;
  if (node_slot != nullptr) {
    const int h = __ldg(head + e), t = __ldg(tail + e);
    const bool ok = static_cast<unsigned>(h) < static_cast<unsigned>(num_nodes) &&
                    static_cast<unsigned>(t) < static_cast<unsigned>(num_nodes);  // flagged by the operand kernel
    head_rep[e] = ok ? __ldg(node_slot + h) : 0;
    tail_rep[e] = ok ? __ldg(node_slot + t) : 0;
  }
  if (edge_attr != nullptr) {
    const long long r = __ldg(edge_attr + e);
    rel_slot[e] = r < 0 || r > 0x7fffffffLL ? -1 : static_cast<int>(r);  // -1: flagged as out of range downstream
  }
}

// ---------------------------------------------------------------------------------------------
// Tail: assemble the state_net.0 pre-activation of both directions from its linear pieces,
//   z_dir = nav_dir * (W_iv m) + W_sc sc_dir +- (P[h] - P[t]) + Rq[pair] + dist_dir * w_dist
//   P[n] = W_err node_repr[n],  Rq[pair] = W_err r_ctx[pair] + b_1          (retriever.py:477-482, linear in err)
// then LayerNorm (two-pass, fp32, in registers) + GELU + the collapsed score head
//   logit = w_o . (W_2 g + b_2) + b_o = (W_2^T w_o) . g + (w_o . b_2 + b_o)      retriever.py:482-483
// and the fwd/bwd softmax combine (retriever.py:369-381). One warp per edge.
// Optionally writes g (16-bit) for the feature path (state_net.4 as a second GEMM).
// ---------------------------------------------------------------------------------------------
struct TailParams {
  const float* w_dist;  // [H] state_net.0.weight[:, 3D]
  const float* ln_g;    // state_net.1.weight
  const float* ln_b;
  const float* v;       // W_2^T w_o
  float c0;             // w_o . b_2 + b_o
};

// Which columns of 256-column group i of a 1024-wide row a lane owns (the tail is elementwise per column, so any
// mapping works as long as every row and parameter vector is read with the same one):
//   SPLIT4 = false: 8 consecutive columns, c = 256 i + 8 lane — one 256-bit global load per group;
//   SPLIT4 = true : two runs of 4, c = 256 i + 4 lane and c + 128 — the 32 lanes of a 128-bit shared-memory load then
//                   cover 512 contiguous bytes. With 8 consecutive floats per lane a quarter-warp's two 128-bit loads put
//                   lanes k and k + 4 on the same banks: ncu counted 15.6M excess shared wavefronts of 39.1M in the tail
//                   (profiles/r04e_kernels_ncu.md, memory_l1_wavefronts_shared vs _ideal).
template <bool SPLIT4>
struct TailCols {
  static EVI_DEVINL int col(int i, int lane) { return i * 256 + lane * (SPLIT4 ? 4 : 8); }
  static constexpr int kSecond = SPLIT4 ? 128 : 4;  // offset of the lane's second run of four columns
  // eight fp32 values of a shared-memory row as four (even, odd) pairs
  static EVI_DEVINL void lds_f32(const float* __restrict__ row, int i, int lane, float2 (&v)[4]) {
    const int c = col(i, lane);
    const float4 a = *reinterpret_cast<const float4*>(row + c), b = *reinterpret_cast<const float4*>(row + c + kSecond);
    v[0] = make_float2(a.x, a.y), v[1] = make_float2(a.z, a.w), v[2] = make_float2(b.x, b.y), v[3] = make_float2(b.z, b.w);
  }
  // the same eight columns of a global fp32 parameter vector (L1-resident)
  static EVI_DEVINL void ldg_f32(const float* __restrict__ row, int i, int lane, float2 (&v)[4]) {
    const int c = col(i, lane);
    if constexpr (!SPLIT4) {
      ldg_v8_keep(row + c, v);
    } else {
      const float4 a = __ldg(reinterpret_cast<const float4*>(row + c));
      const float4 b = __ldg(reinterpret_cast<const float4*>(row + c + kSecond));
      v[0] = make_float2(a.x, a.y), v[1] = make_float2(a.z, a.w), v[2] = make_float2(b.x, b.y), v[3] = make_float2(b.z, b.w);
    }
  }
  // the same eight columns of a 16-bit shared-memory row, packed in column order
  static EVI_DEVINL uint4 lds_16(const uint16_t* __restrict__ row, int i, int lane) {
    const int c = col(i, lane);
    if constexpr (!SPLIT4) {
      return *reinterpret_cast<const uint4*>(row + c);
    } else {
      const uint2 a = *reinterpret_cast<const uint2*>(row + c), b = *reinterpret_cast<const uint2*>(row + c + kSecond);
      return make_uint4(a.x, a.y, b.x, b.y);
    }
  }
};

// LayerNorm statistics of one pre-activation row spread over the warp: rstd and -mean * rstd
// (the lane's 32 columns as 16 (even, odd) pairs, see common.cuh)
EVI_DEVINL void row_norm_stats(const float2 (&z)[16], float& rstd, float& shift) {
  float2 sum = z[0];
#pragma unroll
  for (int j = 1; j < 16; ++j) sum = add2(sum, z[j]);
  const float mean = warp_sum(hsum2(sum)) * (1.0f / D);
  float2 var = bc2(0.f);
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float2 d = sub2(z[j], bc2(mean));
    var = fma2(d, d, var);
  }
  rstd = rsqrtf(warp_sum(hsum2(var)) * (1.0f / D) + kLnEps);
  shift = -mean * rstd;
}

// LayerNorm + GELU + collapsed score head of one direction (z is overwritten with g)
template <int FMT>
EVI_DEVINL float tail_finish(float2 (&z)[16], const TailParams& tp, int lane, uint16_t* __restrict__ g_out) {
  float rstd, shift;
  row_norm_stats(z, rstd, shift);
  float2 acc = bc2(0.f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int col = i * 256 + lane * 8;
    float2 gg[4], bb[4], vv[4];
    ldg_v8_keep(tp.ln_g + col, gg);
    ldg_v8_keep(tp.ln_b + col, bb);
    ldg_v8_keep(tp.v + col, vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 y = gelu_fast(fma2(fma2(z[i * 4 + j], bc2(rstd), bc2(shift)), gg[j], bb[j]));
      z[i * 4 + j] = y;
      acc = fma2(y, vv[j], acc);
    }
  }
  if (g_out != nullptr) store_row16<FMT>(g_out, lane, z);
  return warp_sum(hsum2(acc)) + tp.c0;
}

// Both directions of one edge in one pass over the parameter vectors (ln_g, ln_b, v are read once per edge)
// FEATS = false: the GELU outputs are not stored (no feature path), so they need not be kept in registers
template <int FMT, bool FEATS = true, bool SPLIT4 = false>
EVI_DEVINL void tail_finish_pair(float2 (&zf)[16], float2 (&zb)[16], const TailParams& tp, int lane,
                                 uint16_t* __restrict__ g_fwd, uint16_t* __restrict__ g_bwd, float& lf, float& lb) {
  static_assert(!(FEATS && SPLIT4), "feature rows are stored with 8 consecutive columns per lane");
  float rstd_f, shift_f, rstd_b, shift_b;
  row_norm_stats(zf, rstd_f, shift_f);
  row_norm_stats(zb, rstd_b, shift_b);
  float2 acc_f = bc2(0.f), acc_b = bc2(0.f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 gg[4], bb[4], vv[4];
    TailCols<SPLIT4>::ldg_f32(tp.ln_g, i, lane, gg);
    TailCols<SPLIT4>::ldg_f32(tp.ln_b, i, lane, bb);
    TailCols<SPLIT4>::ldg_f32(tp.v, i, lane, vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 yf = gelu_fast(fma2(fma2(zf[i * 4 + j], bc2(rstd_f), bc2(shift_f)), gg[j], bb[j]));
      const float2 yb = gelu_fast(fma2(fma2(zb[i * 4 + j], bc2(rstd_b), bc2(shift_b)), gg[j], bb[j]));
      if constexpr (FEATS) zf[i * 4 + j] = yf, zb[i * 4 + j] = yb;
      acc_f = fma2(yf, vv[j], acc_f);
      acc_b = fma2(yb, vv[j], acc_b);
    }
  }
  if constexpr (FEATS) {
    if (g_fwd != nullptr) store_row16<FMT>(g_fwd, lane, zf);
    if (g_bwd != nullptr) store_row16<FMT>(g_bwd, lane, zb);
  }
  lf = warp_sum(hsum2(acc_f)) + tp.c0;
  lb = warp_sum(hsum2(acc_b)) + tp.c0;
}

// 16-bit block fixed-point rows of W_sc sc (gemm::EpiQ16Tma): eight u16 -> floats 2^23 + u (the u16 spliced under
// the exponent of 2^23 by one PRMT each). value = (2^23 + u - kQ16Bias) * scale; callers fold -kQ16Bias * scale into an
// fma addend they already have.
constexpr float kQ16Bias = 8388608.0f + 32768.0f;
EVI_DEVINL void splice_u16x8(const uint4& q, float2 (&fb)[4]) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int j = 0; j < 4; ++j)
    fb[j] = make_float2(__uint_as_float(__byte_perm(w[j], 0x4B000000u, 0x7610)),
                        __uint_as_float(__byte_perm(w[j], 0x4B000000u, 0x7632)));
}

template <int FMT, bool MM16, bool SQ16>
__global__ void __launch_bounds__(256)
score_tail_kernel(const void* __restrict__ mm,                                   // [E, D]  W_iv m (fp32, or fp16 if MM16)
                  const void* __restrict__ s_fwd, const void* __restrict__ s_bwd,  // [E, D]  W_sc sc_dir (or null)
                  const float4* __restrict__ s_scale_fwd, const float4* __restrict__ s_scale_bwd,  // SQ16 block scales
                  const float4* __restrict__ scal, const int* __restrict__ head, const int* __restrict__ tail,
                  const int* __restrict__ pair_slot, const float* __restrict__ node_p, const float* __restrict__ pair_rq,
                  int num_nodes, int pair_rows, long long num_edges, TailParams tp, uint16_t* __restrict__ g_fwd,
                  uint16_t* __restrict__ g_bwd, float* __restrict__ logits, float* __restrict__ logits_fwd,
                  float* __restrict__ logits_bwd) {
  const int lane = threadIdx.x & 31;
  const long long e = blockIdx.x * static_cast<long long>(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (e >= num_edges) return;
  int h = __ldg(head + e), t = __ldg(tail + e), ps = __ldg(pair_slot + e);
  if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
      static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes) ||
      static_cast<unsigned>(ps) >= static_cast<unsigned>(pair_rows))
    h = t = ps = 0;  // flagged by the operand kernel / pair kernel
  const float4 sc = __ldg(scal + e);
  const bool has_f = s_fwd != nullptr, has_b = s_bwd != nullptr;
  const float* mrow = static_cast<const float*>(mm) + static_cast<size_t>(e) * D;
  const uint16_t* mrow16 = static_cast<const uint16_t*>(mm) + static_cast<size_t>(e) * D;
  const float* ph = node_p + static_cast<size_t>(h) * D;
  const float* pt = node_p + static_cast<size_t>(t) * D;
  const float* rq = pair_rq + static_cast<size_t>(ps) * D;
  float2 zf[16], zb[16];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float2 m[4], a[4], b[4], r[4], wd[4], d[4], dw_f[4], dw_b[4];
    if constexpr (MM16) unpack8<kFmtF16>(ldg_nc_v4(mrow16 + c), m);
    else ldg_v8_stream(mrow + c, m);
    ldg_v8_keep(ph + c, a);
    ldg_v8_keep(pt + c, b);
    ldg_v8_keep(rq + c, r);
    ldg_v8_keep(tp.w_dist + c, wd);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      d[j] = sub2(a[j], b[j]);
      dw_f[j] = fma2(bc2(sc.z), wd[j], r[j]);
      dw_b[j] = fma2(bc2(sc.w), wd[j], r[j]);
    }
    if constexpr (SQ16) {
      if (has_f) {
        float2 fb[4];
        splice_u16x8(ldg_nc_v4(static_cast<const uint16_t*>(s_fwd) + static_cast<size_t>(e) * D + c), fb);
        const float qs = __ldg(reinterpret_cast<const float*>(s_scale_fwd + e) + i), off = -kQ16Bias * qs;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          zf[i * 4 + j] = fma2(fb[j], bc2(qs), add2(add2(fma2(bc2(sc.x), m[j], bc2(off)), d[j]), dw_f[j]));
      }
      if (has_b) {
        float2 fb[4];
        splice_u16x8(ldg_nc_v4(static_cast<const uint16_t*>(s_bwd) + static_cast<size_t>(e) * D + c), fb);
        const float qs = __ldg(reinterpret_cast<const float*>(s_scale_bwd + e) + i), off = -kQ16Bias * qs;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          zb[i * 4 + j] = fma2(fb[j], bc2(qs), add2(sub2(fma2(bc2(sc.y), m[j], bc2(off)), d[j]), dw_b[j]));
      }
    } else {
      if (has_f) {
        float2 sv[4];
        ldg_v8_stream(static_cast<const float*>(s_fwd) + static_cast<size_t>(e) * D + c, sv);
#pragma unroll
        for (int j = 0; j < 4; ++j) zf[i * 4 + j] = add2(add2(fma2(bc2(sc.x), m[j], sv[j]), d[j]), dw_f[j]);
      }
      if (has_b) {
        float2 sv[4];
        ldg_v8_stream(static_cast<const float*>(s_bwd) + static_cast<size_t>(e) * D + c, sv);
#pragma unroll
        for (int j = 0; j < 4; ++j) zb[i * 4 + j] = add2(sub2(fma2(bc2(sc.y), m[j], sv[j]), d[j]), dw_b[j]);
      }
    }
  }
  float lf = 0.f, lb = 0.f;
  if (has_f && has_b)
    tail_finish_pair<FMT>(zf, zb, tp, lane, g_fwd ? g_fwd + static_cast<size_t>(e) * 2 * D : nullptr,
                          g_bwd ? g_bwd + static_cast<size_t>(e) * 2 * D : nullptr, lf, lb);
  else if (has_f) lf = tail_finish<FMT>(zf, tp, lane, g_fwd ? g_fwd + static_cast<size_t>(e) * 2 * D : nullptr);
  else if (has_b) lb = tail_finish<FMT>(zb, tp, lane, g_bwd ? g_bwd + static_cast<size_t>(e) * 2 * D : nullptr);
  if (lane == 0) {
    if (has_f && logits_fwd) logits_fwd[e] = lf;
    if (has_b && logits_bwd) logits_bwd[e] = lb;
    if (logits) {
      float out;
      if (has_f && has_b) {  // softmax([lf, lb]) weighted sum
        const float mx = fmaxf(lf, lb);
        const float ef = expf(lf - mx), eb = expf(lb - mx);
        const float inv = 1.0f / (ef + eb);
        out = (ef * inv) * lf + (eb * inv) * lb;
      } else {
        out = has_f ? lf : lb;
      }
      logits[e] = out;
    }
  }
}

// Same tail, fed by the TMA engine: persistent CTAs of 16 warps around a ring of eight 24 KB shared-memory
// buffers. 1-D bulk copies fill a buffer with the six rows of one edge (W_iv m, W_sc sc_f, W_sc sc_b in fp32 or 16 bits,
// P[h], P[t], Rq[pair]); warps b and b+8 take turns on buffer b: the warp that has just moved its edge
// from the buffer into registers hands the buffer back to the TMA engine for its partner's next edge and
// then does LayerNorm / GELU / the head dot in registers. 192 KB stay in flight per SM without spending
// registers or LSU issue slots on it (the register-fed version above waits on the long scoreboard 57% of
// the time at 46% of DRAM peak). Each warp has its own full-barrier, so every barrier is waited on by one
// warp through consecutive phases.
constexpr int kTailWarps = 16;
constexpr int kTailBufs = 8;
constexpr int kTailRowBytes = D * 4;
constexpr int kTailBufBytes = 6 * kTailRowBytes;
constexpr int kTailSmemBytes = kTailBufs * kTailBufBytes + kTailWarps * 8;

// SQ16: the W_sc sc rows arrive as 16-bit block fixed point (gemm::EpiQ16Tma): u16 values with one fp32 scale per
// (row, 256-column block) in s_scale_*[edge * 4 + block]; value = (u - 32768) * scale. The u16 is spliced under the
// exponent of 2^23 (one PRMT), and the constant -(2^23 + 32768) * scale rides in an existing fma addend.
template <int FMT, bool MM16, bool SQ16, bool BIDIR = false>
__global__ void __launch_bounds__(kTailWarps * 32, 1)
score_tail_bulk_kernel(const void* __restrict__ mm, const void* __restrict__ s_fwd, const void* __restrict__ s_bwd,
                       const float4* __restrict__ s_scale_fwd, const float4* __restrict__ s_scale_bwd,
                       const float4* __restrict__ scal, const int* __restrict__ head, const int* __restrict__ tail,
                       const int* __restrict__ pair_slot, const float* __restrict__ node_p,
                       const float* __restrict__ pair_rq, int num_nodes, int pair_rows, long long num_edges,
                       TailParams tp, uint16_t* __restrict__ g_fwd, uint16_t* __restrict__ g_bwd,
                       float* __restrict__ logits, float* __restrict__ logits_fwd, float* __restrict__ logits_bwd) {
  extern __shared__ __align__(128) uint8_t tail_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* buf = reinterpret_cast<float*>(tail_smem + (warp % kTailBufs) * kTailBufBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail_smem + kTailBufs * kTailBufBytes);
  uint64_t* my_bar = bars + warp;
  uint64_t* partner_bar = bars + (warp + kTailBufs) % kTailWarps;
  // BIDIR: both directions present and no feature rows wanted, known at compile time (the default mode)
  const bool has_f = BIDIR || s_fwd != nullptr, has_b = BIDIR || s_bwd != nullptr;
  constexpr uint32_t kMmRowBytes = MM16 ? D * 2 : kTailRowBytes;  // W_iv m rows are fp16 when MM16
  constexpr uint32_t kSRowBytes = SQ16 ? D * 2 : kTailRowBytes;   // W_sc sc rows are 16-bit fixed point when SQ16
  const uint32_t tx_bytes = static_cast<uint32_t>(3 * kTailRowBytes) +
                            static_cast<uint32_t>((has_f ? 1 : 0) + (has_b ? 1 : 0)) * kSRowBytes + kMmRowBytes;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kTailWarps; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  // CTA-local sequence s -> edge blockIdx.x + s * gridDim.x; warp w takes s = w, w + 16, ...; s uses buffer s % 8
  const long long cta_stride = gridDim.x;
  auto edge_of = [&](long long s) { return blockIdx.x + s * cta_stride; };
  auto load_ids = [&](long long edge, int& h, int& t, int& ps) {
    h = __ldg(head + edge), t = __ldg(tail + edge), ps = __ldg(pair_slot + edge);
    if (static_cast<unsigned>(h) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(t) >= static_cast<unsigned>(num_nodes) ||
        static_cast<unsigned>(ps) >= static_cast<unsigned>(pair_rows))
      h = t = ps = 0;  // flagged by the operand / pair kernels
  };
  auto issue = [&](uint64_t* bar, long long edge, int h, int t, int ps) {  // one lane
    mbar_arrive_expect_tx(bar, tx_bytes);
    bulk_load_1d(buf + 0 * D, static_cast<const uint8_t*>(mm) + static_cast<size_t>(edge) * kMmRowBytes, kMmRowBytes, bar);
    if (has_f)
      bulk_load_1d(buf + 1 * D, static_cast<const uint8_t*>(s_fwd) + static_cast<size_t>(edge) * kSRowBytes, kSRowBytes, bar);
    if (has_b)
      bulk_load_1d(buf + 2 * D, static_cast<const uint8_t*>(s_bwd) + static_cast<size_t>(edge) * kSRowBytes, kSRowBytes, bar);
    bulk_load_1d(buf + 3 * D, node_p + static_cast<size_t>(h) * D, kTailRowBytes, bar);
    bulk_load_1d(buf + 4 * D, node_p + static_cast<size_t>(t) * D, kTailRowBytes, bar);
    bulk_load_1d(buf + 5 * D, pair_rq + static_cast<size_t>(ps) * D, kTailRowBytes, bar);
  };

  long long s = warp;
  if (warp < kTailBufs && edge_of(s) < num_edges) {  // the first eight edges: each buffer's first owner loads its own
    int h, t, ps;
    load_ids(edge_of(s), h, t, ps);
    if (lane == 0) issue(my_bar, edge_of(s), h, t, ps);
  }
  uint32_t phase = 0;
  for (; edge_of(s) < num_edges; s += kTailWarps) {
    const long long e = edge_of(s);
    const long long e_part = edge_of(s + kTailBufs);  // the partner warp's next edge: same buffer, loaded by this warp
    int hn = 0, tn = 0, pn = 0;
    if (e_part < num_edges) load_ids(e_part, hn, tn, pn);  // in flight while this edge's rows land
    const float4 sc = __ldg(scal + e);
    float qs_lane = 0.f;  // lanes 0-3: forward block scales, lanes 4-7: backward (one register instead of eight)
    if constexpr (SQ16) {
      if (lane < 4 && has_f) qs_lane = __ldg(reinterpret_cast<const float*>(s_scale_fwd + e) + lane);
      if (lane >= 4 && lane < 8 && has_b) qs_lane = __ldg(reinterpret_cast<const float*>(s_scale_bwd + e) + (lane - 4));
    }
    mbar_wait(my_bar, phase);
    phase ^= 1u;
    using Cols = TailCols<BIDIR>;  // the default mode reads shared memory without bank conflicts (see TailCols)
    float2 zf[16], zb[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 wd[4], m[4], a[4], b[4], r[4], d[4], dw_f[4], dw_b[4];
      Cols::ldg_f32(tp.w_dist, i, lane, wd);
      if constexpr (MM16) unpack8<kFmtF16>(Cols::lds_16(reinterpret_cast<const uint16_t*>(buf), i, lane), m);
      else Cols::lds_f32(buf, i, lane, m);
      Cols::lds_f32(buf + 3 * D, i, lane, a);
      Cols::lds_f32(buf + 4 * D, i, lane, b);
      Cols::lds_f32(buf + 5 * D, i, lane, r);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        d[j] = sub2(a[j], b[j]);
        dw_f[j] = fma2(bc2(sc.z), wd[j], r[j]);
        dw_b[j] = fma2(bc2(sc.w), wd[j], r[j]);
      }
      if constexpr (SQ16) {
        // rows of an absent direction are never loaded: their slots hold stale bytes and the results are unused
        const float qf = __shfl_sync(0xffffffffu, qs_lane, i), qb = __shfl_sync(0xffffffffu, qs_lane, 4 + i);
        const float off_f = -kQ16Bias * qf, off_b = -kQ16Bias * qb;
        float2 ff[4], fbk[4];
        splice_u16x8(Cols::lds_16(reinterpret_cast<const uint16_t*>(buf + 1 * D), i, lane), ff);
        splice_u16x8(Cols::lds_16(reinterpret_cast<const uint16_t*>(buf + 2 * D), i, lane), fbk);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          zf[i * 4 + j] = fma2(ff[j], bc2(qf), add2(add2(fma2(bc2(sc.x), m[j], bc2(off_f)), d[j]), dw_f[j]));
          zb[i * 4 + j] = fma2(fbk[j], bc2(qb), add2(sub2(fma2(bc2(sc.y), m[j], bc2(off_b)), d[j]), dw_b[j]));
        }
      } else {
        if (has_f) {
          float2 sv[4];
          Cols::lds_f32(buf + 1 * D, i, lane, sv);
#pragma unroll
          for (int j = 0; j < 4; ++j) zf[i * 4 + j] = add2(add2(fma2(bc2(sc.x), m[j], sv[j]), d[j]), dw_f[j]);
        }
        if (has_b) {
          float2 sv[4];
          Cols::lds_f32(buf + 2 * D, i, lane, sv);
#pragma unroll
          for (int j = 0; j < 4; ++j) zb[i * 4 + j] = add2(sub2(fma2(bc2(sc.y), m[j], sv[j]), d[j]), dw_b[j]);
        }
      }
    }
    // every lane has moved its part of the buffer into registers: hand the buffer to the partner's next edge
    __syncwarp();
    if (e_part < num_edges && lane == 0) {
      // the generic-proxy reads of this buffer above are ordered before the async-proxy writes of the next bulk copies;
      // only shared memory is involved (the global rows were written by earlier launches), so the shared::cta form of
      // the fence is enough (the state-space-wide form showed up as 0.7 warps per scheduler stalled on membar)
      fence_proxy_async_smem();
      issue(partner_bar, e_part, hn, tn, pn);
    }
    float lf = 0.f, lb = 0.f;
    if constexpr (BIDIR)  // both directions, logits only
      tail_finish_pair<FMT, false, true>(zf, zb, tp, lane, nullptr, nullptr, lf, lb);
    else if (has_f && has_b)
      tail_finish_pair<FMT>(zf, zb, tp, lane, g_fwd ? g_fwd + static_cast<size_t>(e) * 2 * D : nullptr,
                            g_bwd ? g_bwd + static_cast<size_t>(e) * 2 * D : nullptr, lf, lb);
    else if (has_f) lf = tail_finish<FMT>(zf, tp, lane, g_fwd ? g_fwd + static_cast<size_t>(e) * 2 * D : nullptr);
    else if (has_b) lb = tail_finish<FMT>(zb, tp, lane, g_bwd ? g_bwd + static_cast<size_t>(e) * 2 * D : nullptr);
    if (lane == 0) {
      if (has_f && logits_fwd) logits_fwd[e] = lf;
      if (has_b && logits_bwd) logits_bwd[e] = lb;
      if (logits) {
        float out;
        if (has_f && has_b) {  // softmax([lf, lb]) weighted sum
          const float mx = fmaxf(lf, lb);
          const float ef = expf(lf - mx), eb = expf(lb - mx);
          const float inv = 1.0f / (ef + eb);
          out = (ef * inv) * lf + (eb * inv) * lb;
        } else {
          out = has_f ? lf : lb;
        }
        logits[e] = out;
      }
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

// ---------------------------------------------------------------------------------------------
// Distinct (question, relation) pairs of a batch. r_ctx = rel * gate[q] + bias[q] (retriever.py:464)
// depends only on the pair, so W_err r_ctx is computed once per pair instead of once per edge and
// direction. One CTA per question: relation bitmap in shared memory -> rank of each relation id
// inside the question -> pair slot = base (one atomicAdd per question) + rank. Slot numbering
// across questions depends on CTA order; nothing downstream depends on the numbering.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
relation_pairs_kernel(const int* __restrict__ rel, const long long* __restrict__ edge_ptr, int num_relations,
                      int words, int capacity, int* __restrict__ pair_slot, int* __restrict__ pair_q,
                      int* __restrict__ pair_rel, int* __restrict__ num_pairs, unsigned* __restrict__ flags) {
  extern __shared__ uint32_t bm[];       // [words] bitmap, then [words] exclusive popcount prefix
  uint32_t* prefix = bm + words;
  __shared__ int warp_tot[8];
  __shared__ int s_carry, s_base;
  const int q = blockIdx.x, tid = threadIdx.x;
  const long long e0 = edge_ptr[q], e1 = edge_ptr[q + 1];
  for (int i = tid; i < words; i += blockDim.x) bm[i] = 0u;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (long long e = e0 + tid; e < e1; e += blockDim.x) {
    const int r = __ldg(rel + e);
    if (static_cast<unsigned>(r) >= static_cast<unsigned>(num_relations)) {
      atomicOr(flags, kFlagOutOfRange);
      continue;
    }
    atomicOr(&bm[r >> 5], 1u << (r & 31));
  }
  __syncthreads();
  const int lane = tid & 31, w = tid >> 5;
  for (int base = 0; base < words; base += blockDim.x) {
    const int i = base + tid;
    const int c = i < words ? __popc(bm[i]) : 0;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    int off = s_carry;
    for (int j = 0; j < w; ++j) off += warp_tot[j];
    if (i < words) prefix[i] = static_cast<uint32_t>(off + incl - c);
    __syncthreads();
    if (tid == blockDim.x - 1) s_carry = off + incl;
    __syncthreads();
  }
  if (tid == 0) {
    const int count = s_carry;
    int b = atomicAdd(num_pairs, count);
    if (b + count > capacity) {
      atomicOr(flags, kFlagOutOfRange);
      b = 0;  // never write outside the pair arrays; the flag makes the host raise
    }
    s_base = b;
  }
  __syncthreads();
  const int pbase = s_base;
  const bool fits = pbase + s_carry <= capacity;
  for (long long e = e0 + tid; e < e1; e += blockDim.x) {
    const int r = __ldg(rel + e);
    int slot = 0;
    if (static_cast<unsigned>(r) < static_cast<unsigned>(num_relations) && fits)
      slot = pbase + static_cast<int>(prefix[r >> 5]) + __popc(bm[r >> 5] & ((1u << (r & 31)) - 1u));
    pair_slot[e] = slot;
  }
  if (fits) {
    for (int i = tid; i < words; i += blockDim.x) {
      uint32_t bits = bm[i];
      int slot = pbase + static_cast<int>(prefix[i]);
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        pair_q[slot] = q;
        pair_rel[slot] = i * 32 + b;
        ++slot;
      }
    }
  }
}

// rc16[slot] = fp16/bf16(rel_repr[pair_rel] * gate[pair_q] + bias[pair_q]); one warp per pair slot.
// SPLIT: rel_repr rows are [hi | lo] and so are the rc rows written (strict mode)
template <int FMT, bool SPLIT>
__global__ void __launch_bounds__(256)
pair_context_kernel(const int* __restrict__ pair_q, const int* __restrict__ pair_rel, const int* __restrict__ num_pairs,
                    int capacity, const uint16_t* __restrict__ rel_repr, int rel_ld, int num_rel_rows,
                    const float* __restrict__ gate_bias, int num_questions, uint16_t* __restrict__ rc16, int rc_ld,
                    unsigned* __restrict__ flags) {
  const int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int n = min(__ldg(num_pairs), capacity);
  if (slot >= n) return;
  const int lane = threadIdx.x & 31;
  int q = __ldg(pair_q + slot), r = __ldg(pair_rel + slot);
  if (static_cast<unsigned>(q) >= static_cast<unsigned>(num_questions) ||
      static_cast<unsigned>(r) >= static_cast<unsigned>(num_rel_rows)) {
    if (lane == 0) atomicOr(flags, kFlagOutOfRange);
    q = r = 0;
  }
  const uint16_t* rp = rel_repr + static_cast<size_t>(r) * rel_ld;
  const float* gp = gate_bias + static_cast<size_t>(q) * 2 * D;
  uint16_t* out = rc16 + static_cast<size_t>(slot) * rc_ld;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = i * 256 + lane * 8;
    float rv[8], g[8], b[8];
    unpack8<FMT>(ldg_nc_v4(rp + c), rv);
    if constexpr (SPLIT) {
      float lo[8];
      unpack8<FMT>(ldg_nc_v4(rp + D + c), lo);
#pragma unroll
      for (int j = 0; j < 8; ++j) rv[j] += lo[j];
    }
    ldg_v8_keep(gp + c, g);
    ldg_v8_keep(gp + D + c, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) rv[j] = fmaf(rv[j], g[j], b[j]);
    const uint4 w = pack8<FMT>(rv);
    *reinterpret_cast<uint4*>(out + c) = w;
    if constexpr (SPLIT) {
      float h[8];
      unpack8<FMT>(w, h);
#pragma unroll
      for (int j = 0; j < 8; ++j) rv[j] -= h[j];
      *reinterpret_cast<uint4*>(out + D + c) = pack8<FMT>(rv);
    }
  }
}

// -------------------------------------------------------------------------------------------
// host-side dispatch on operand format
// -------------------------------------------------------------------------------------------
template <int FMT>
static int gather_rows_impl(const void* src, int src_is_f32, long long src_rows, int src_ld, const long long* ids,
                            const long long* sub_mask_ids, const float* sub_row, int num_rows, const int* rows_dev,
                            void* dst, int dst_ld, int split, unsigned* flags, cudaStream_t stream) {
  if (num_rows <= 0) return kOk;
  if (src_ld < (split && !src_is_f32 ? 2 * D : D) || dst_ld < (split ? 2 * D : D) || dst_ld % 8 != 0 ||
      src_ld % (src_is_f32 ? 4 : 8) != 0)
    return kErrInvalidArg;
  const int warps = 8;
  const int grid = (num_rows + warps - 1) / warps;
  auto go = [&](auto kern) {
    kern<<<grid, warps * 32, 0, stream>>>(src, src_rows, src_ld, ids, sub_mask_ids, sub_row, num_rows, rows_dev,
                                          static_cast<uint16_t*>(dst), dst_ld, flags);
  };
  if (src_is_f32) {
    if (split) go(gather_rows_kernel<FMT, true, true>);
    else go(gather_rows_kernel<FMT, true, false>);
  } else {
    if (split) go(gather_rows_kernel<FMT, false, true>);
    else go(gather_rows_kernel<FMT, false, false>);
  }
  return check_launch();
}

// out = act(x W^T + b) with optional [hi | lo] halves of x (x_lo_off), W (w_lo_off) and of the 16-bit output (out_lo_off)
template <int FMT>
static int linear_act_impl(const void* x16, int ld_x, int x_lo_off, int m, const int* m_dev, const void* w16, int ldw,
                           int w_lo_off, int n, int k, const float* bias, const int* act_per_block, void* out16, int ld16,
                           int out_lo_off, float* out32, int num_sms, cudaStream_t stream) {
  const gemm::SplitK sk = gemm::make_split(x_lo_off, w_lo_off);
  if (out16 == nullptr && out32 != nullptr) {  // fp32 only: staged TMA-store epilogue
    gemm::EpiF32Tma::Params t;
    t.bias = bias;
    for (int i = 0; i < 8; ++i) t.act[i] = act_per_block[i];
    return gemm_f32<FMT>(x16, ld_x, w16, ldw, m, n, k, t, num_sms, stream, m_dev, out32, n, sk);
  }
  static const bool one_sm = getenv("EVI_GEMM_1SM") != nullptr;
  if (out32 == nullptr && out_lo_off < 0 && !one_sm && num_sms >= 2 && n % 64 == 0 && ld16 % 8 == 0 &&
      (reinterpret_cast<uintptr_t>(out16) & 15) == 0) {
    gemm::EpiBiasAct16Tma::Params t;  // 16-bit rows only: staged TMA store instead of per-thread row stores
    t.bias = bias;
    for (int i = 0; i < 8; ++i) t.act[i] = act_per_block[i];
    return gemm2::launch<FMT, gemm::EpiBiasAct16Tma>(x16, ld_x, w16, ldw, m, n, k, t, num_sms, stream, m_dev, out16, ld16,
                                                      sk);
  }
  if (out_lo_off >= 0 && (out_lo_off % 8 != 0 || out_lo_off + n > ld16)) return kErrInvalidArg;
  static const bool hilo_direct = getenv("EVI_HILO_DIRECT") != nullptr;  // A/B knob: the per-thread row stores below
  if (out32 == nullptr && out_lo_off >= n && out_lo_off % 64 == 0 && !one_sm && !hilo_direct && num_sms >= 2 &&
      n % 64 == 0 && ld16 % 8 == 0 && (reinterpret_cast<uintptr_t>(out16) & 15) == 0) {
    gemm::EpiBiasActHiLoTma::Params t;  // [hi | lo] rows: both halves staged and TMA-stored per 32 x 64 box
    t.bias = bias, t.lo_off = out_lo_off;
    for (int i = 0; i < 8; ++i) t.act[i] = act_per_block[i];
    return gemm2::launch<FMT, gemm::EpiBiasActHiLoTma>(x16, ld_x, w16, ldw, m, n, k, t, num_sms, stream, m_dev, out16,
                                                        ld16, sk);
  }
  gemm::EpiBiasAct::Params p;
  p.bias = bias, p.out16 = out16, p.out32 = out32, p.ld16 = ld16, p.ld32 = n, p.lo_off = out_lo_off;
  for (int i = 0; i < 8; ++i) p.act[i] = act_per_block[i];
  if (one_sm || num_sms < 2)
    return gemm::launch<FMT, gemm::EpiBiasAct>(x16, ld_x, w16, ldw, m, n, k, p, num_sms, stream, m_dev, nullptr, 0, sk);
  return gemm2::launch<FMT, gemm::EpiBiasAct>(x16, ld_x, w16, ldw, m, n, k, p, num_sms, stream, m_dev, nullptr, 0, sk);
}

// workspace of one chunk of `E` edges:
//   m16 [E, D] 16-bit | sc16 [nd*E, D] 16-bit | mm [E, D] f32 | s [nd*E, D] f32 | scal [E] float4
//   (the mm / s regions keep their fp32 size; by default they hold fp16 rows, and u16 block fixed-point rows followed
//   by the [nd*E, 4] block scales: ChunkPlan::mm16 / sq16)
//   (features) g16 [nd*E, D] 16-bit | feat [nd*E, D] f32
static size_t chunk_workspace_bytes(long long num_edges, int direction, int want_features, int strict) {
  const size_t nd = direction == EVI_DIR_BIDIRECTIONAL ? 2 : 1;
  const size_t E = static_cast<size_t>(num_edges);
  const size_t op_ld = strict ? 2 * D : D;  // strict: [hi | lo] operand rows
  auto al = [](size_t b) { return (b + 255) & ~static_cast<size_t>(255); };
  size_t total = al(E * op_ld * 2) + al(nd * E * op_ld * 2) + al(E * D * sizeof(float)) + al(nd * E * D * sizeof(float)) +
                 al(E * sizeof(float4));
  if (want_features) total += al(nd * E * 2 * D * 2) + al(nd * E * D * sizeof(float));  // g rows are [hi | lo]
  return total + 256;
}

// One chunk of edges = three stages that can be enqueued separately (the pipelined mode overlaps the tensor-bound
// stage of chunk c with the HBM-bound stages of its neighbours):
//   operand  edge_operand_kernel -> M16, SC16, SCAL
//   gemm     W_iv m -> MM, W_sc sc -> S
//   tail     z assembly, LN, GELU, v-dot, softmax combine -> logits (or the features path)
template <int FMT>
struct ChunkPlan {
  const evi_score_args& a;
  long long e0, E;
  int do_f, do_b;
  size_t nd, n;
  uint16_t *M16, *SC16, *G;
  float *MM, *S, *FEAT;  // MM holds fp16 rows when mm16 (the buffer keeps its fp32 size)
  float4* SCAL;
  bool ok, mm16, sq16, strict;
  size_t op_ld;  // row pitch of the operand rows: D, or 2D in the strict mode ([hi | lo])
  int w1_ld, w1_lo;

  ChunkPlan(const evi_score_args& args, long long e0_, long long E_, uint8_t* ws, size_t ws_bytes)
      : a(args), e0(e0_), E(E_), G(nullptr), FEAT(nullptr) {
    do_f = a.direction != EVI_DIR_BACKWARD, do_b = a.direction != EVI_DIR_FORWARD;
    nd = static_cast<size_t>(do_f + do_b);
    n = static_cast<size_t>(E);
    strict = a.precision == EVI_PRECISION_STRICT;
    op_ld = strict ? 2 * D : D;
    w1_ld = a.w1_ld > 0 ? a.w1_ld : 3 * D;
    w1_lo = a.w1_lo_off;
    size_t off = 0;
    auto take = [&](size_t bytes) {
      uint8_t* p = ws + off;
      off += (bytes + 255) & ~static_cast<size_t>(255);
      return p;
    };
    M16 = reinterpret_cast<uint16_t*>(take(n * op_ld * 2));
    SC16 = reinterpret_cast<uint16_t*>(take(nd * n * op_ld * 2));
    MM = reinterpret_cast<float*>(take(n * D * sizeof(float)));
    S = reinterpret_cast<float*>(take(nd * n * D * sizeof(float)));
    SCAL = reinterpret_cast<float4*>(take(n * sizeof(float4)));
    if (a.want_features) {
      G = reinterpret_cast<uint16_t*>(take(nd * n * 2 * D * 2));
      FEAT = reinterpret_cast<float*>(take(nd * n * D * sizeof(float)));
    }
    ok = off <= ws_bytes && (!strict || (w1_lo > 0 && w1_lo + 3 * D <= w1_ld));
    // W_iv m in fp16 needs the CTA-pair kernel's 16-bit TMA-store epilogue
    static const bool one_sm = getenv("EVI_GEMM_1SM") != nullptr, mm32 = getenv("EVI_MM_F32") != nullptr;
    mm16 = !one_sm && !mm32 && a.num_sms >= 2 && !strict;  // strict keeps fp32 rows of both partial products
    // W_sc sc as 16-bit block fixed point: the CTA-pair kernel's quantising epilogue; both tail kernels read it. The S
    // buffer keeps its fp32 size: u16 rows in the first half, the [nd * n, 4] scales behind them.
    static const bool s32 = getenv("EVI_S_F32") != nullptr;
    sq16 = mm16 && !s32;
  }
  uint8_t* s_rows(int dir_slot) const {  // dir_slot: 0 = first direction present, 1 = second
    return reinterpret_cast<uint8_t*>(S) + static_cast<size_t>(dir_slot) * n * D * (sq16 ? 2 : 4);
  }
  float* s_scales(int dir_slot) const {
    return reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(S) + nd * n * D * 2) + static_cast<size_t>(dir_slot) * n * 4;
  }

  int operand(cudaStream_t stream) const {
    uint16_t* sc_f = SC16;
    uint16_t* sc_b = SC16 + (do_f ? n * op_ld : 0);
    StructParams sp;
    sp.ln_g = a.struct_ln_g, sp.ln_b = a.struct_ln_b, sp.wn = a.nav_w, sp.bn = a.nav_b, sp.center = a.sc_center;
    const long long want = (E + kEdgeWarps - 1) / kEdgeWarps;
    // resident CTAs only: one sweep over the chunk in edge order, so the node rows of a question (~15 MB) stay in
    // L2 while its edges are processed; a larger grid makes later waves sweep the chunk again from DRAM
    const long long resident = static_cast<long long>(EVI_OPERAND_CTAS) * a.num_sms;
    const int grid = static_cast<int>(want < resident ? want : resident);
    auto kern = strict            ? edge_operand_kernel<FMT, false, true>
                : do_f && do_b ? edge_operand_kernel<FMT, true, false>
                               : edge_operand_kernel<FMT, false, false>;
    const int* head_rep = a.head_rep32 != nullptr ? a.head_rep32 : a.head32;
    const int* tail_rep = a.tail_rep32 != nullptr ? a.tail_rep32 : a.tail32;
    const int rep_rows = a.num_rep_rows > 0 ? a.num_rep_rows : a.num_nodes;
    kern<<<grid, kEdgeWarps * 32, 0, stream>>>(
        a.head32 + e0, a.tail32 + e0, head_rep + e0, tail_rep + e0, a.qid32 + e0, a.rel_slot32 + e0, E,
        static_cast<const uint16_t*>(a.node_repr16), a.node_repr_ld > 0 ? a.node_repr_ld : D,
        static_cast<const uint16_t*>(a.rel_repr16), a.rel_repr_ld > 0 ? a.rel_repr_ld : D, a.gate_bias32, a.node_uv32,
        sp, do_f, do_b, a.num_nodes, rep_rows, a.num_rel_rows, a.num_questions, a.flags, M16, sc_f, sc_b, SCAL);
    return check_launch();
  }

  // state_net.0 split by input block (weight columns [iv | sc | err | dist], retriever.py:470-481):
  //   W_iv m -> MM (once per edge, shared by both directions),  W_sc sc_dir -> S (once per edge and direction)
  int gemms(cudaStream_t stream, int gemm_sms, cudaEvent_t time_begin, cudaEvent_t time_end) const {
    const uint16_t* w1 = static_cast<const uint16_t*>(a.w1_main16);
    gemm::EpiF32Tma::Params plain;
    plain.bias = nullptr;
    for (int i = 0; i < 8; ++i) plain.act[i] = gemm::kActNone;
    // strict: [hi | lo] operand rows against [hi | lo] weight blocks, three passes into the same accumulator
    const gemm::SplitK sk = strict ? gemm::make_split(D, w1_lo) : gemm::make_split(-1, -1);
    const int lda = static_cast<int>(op_ld);
    int rc = mm16 ? gemm2::launch<FMT, gemm::EpiF16Tma>(M16, lda, w1, w1_ld, static_cast<int>(E), D, D,
                                                         gemm::EpiF16Tma::Params{}, gemm_sms, stream, nullptr, MM, D)
                  : gemm_f32<FMT>(M16, lda, w1, w1_ld, static_cast<int>(E), D, D, plain, gemm_sms, stream, nullptr, MM, D,
                                  sk);
    if (rc != kOk) return rc;
    if (time_begin) cudaEventRecord(time_begin, stream);
    if (sq16) {
      gemm::EpiQ16Tma::Params qp;
      qp.scale = s_scales(0), qp.ld_scale = D / 256;
      rc = gemm2::launch<FMT, gemm::EpiQ16Tma>(SC16, lda, w1 + D, w1_ld, static_cast<int>(nd * n), D, D, qp, gemm_sms,
                                               stream, nullptr, S, D);
    } else {
      rc = gemm_f32<FMT>(SC16, lda, w1 + D, w1_ld, static_cast<int>(nd * n), D, D, plain, gemm_sms, stream, nullptr, S, D,
                         sk);
    }
    if (time_end) cudaEventRecord(time_end, stream);
    return rc;
  }

  // tail_mode 0: TMA ring (persistent, 192 KB of shared memory per CTA); 1: register-fed, `tail_threads` per CTA
  // (no shared memory, so its CTAs fit beside a resident GEMM CTA)
  int tail(cudaStream_t stream, int tail_mode, int tail_threads, cudaEvent_t time_begin = nullptr,
           cudaEvent_t time_end = nullptr) const {
    float* logits = a.logits ? a.logits + e0 : nullptr;
    float* logits_fwd = a.logits_fwd ? a.logits_fwd + e0 : nullptr;
    float* logits_bwd = a.logits_bwd ? a.logits_bwd + e0 : nullptr;
    float* edge_emb = a.edge_embeddings ? a.edge_embeddings + static_cast<size_t>(e0) * D : nullptr;
    TailParams tp;
    tp.w_dist = a.w1_dist_col, tp.ln_g = a.ln1_g, tp.ln_b = a.ln1_b, tp.v = a.head_v, tp.c0 = a.head_c0;
    const void* sq_f = do_f ? s_rows(0) : nullptr;
    const void* sq_b = do_b ? s_rows(do_f ? 1 : 0) : nullptr;
    const float4* sqs_f = reinterpret_cast<const float4*>(s_scales(0));
    const float4* sqs_b = reinterpret_cast<const float4*>(s_scales(do_f ? 1 : 0));
    uint16_t* g_f = G;
    uint16_t* g_b = G ? G + (do_f ? n * 2 * D : 0) : nullptr;
    const bool feats = a.want_features != 0;
    // the tail reads node ids only to fetch P rows, which exist once per distinct embedding id
    const int* p_head = (a.head_rep32 != nullptr ? a.head_rep32 : a.head32) + e0;
    const int* p_tail = (a.tail_rep32 != nullptr ? a.tail_rep32 : a.tail32) + e0;
    const int p_rows = a.num_rep_rows > 0 ? a.num_rep_rows : a.num_nodes;
    if (time_begin) cudaEventRecord(time_begin, stream);
    if (tail_mode == 1) {
      const int warps = tail_threads / 32;
      const int tgrid = static_cast<int>((E + warps - 1) / warps);
      auto kern = sq16   ? score_tail_kernel<FMT, true, true>
                  : mm16 ? score_tail_kernel<FMT, true, false>
                         : score_tail_kernel<FMT, false, false>;
      kern<<<tgrid, tail_threads, 0, stream>>>(
          MM, sq_f, sq_b, sqs_f, sqs_b, SCAL, p_head, p_tail, a.pair_slot32 + e0, a.node_p32, a.pair_rq32, p_rows,
          a.num_pair_rows, E, tp, feats && do_f ? g_f : nullptr, feats && do_b ? g_b : nullptr, feats ? nullptr : logits,
          feats ? nullptr : logits_fwd, feats ? nullptr : logits_bwd);
    } else {
      const bool bidir = sq16 && do_f && do_b && !feats;  // the compile-time variant: both directions, logits only
      auto kern = bidir  ? score_tail_bulk_kernel<FMT, true, true, true>
                  : sq16 ? score_tail_bulk_kernel<FMT, true, true>
                  : mm16 ? score_tail_bulk_kernel<FMT, true, false>
                         : score_tail_bulk_kernel<FMT, false, false>;
      static std::atomic<unsigned long long> smem_ready[4] = {{0}, {0}, {0}, {0}};  // per FMT instantiation, per variant
      if (!ensure_dynamic_smem(kern, kTailSmemBytes, smem_ready[bidir ? 3 : sq16 ? 2 : mm16 ? 1 : 0])) return kErrCuda;
      const long long want_ctas = (E + kTailWarps - 1) / kTailWarps;
      const int bgrid = static_cast<int>(want_ctas < a.num_sms ? want_ctas : a.num_sms);  // one persistent CTA per SM
      kern<<<bgrid, kTailWarps * 32, kTailSmemBytes, stream>>>(
          MM, sq_f, sq_b, sqs_f, sqs_b, SCAL, p_head, p_tail, a.pair_slot32 + e0, a.node_p32, a.pair_rq32, p_rows,
          a.num_pair_rows, E, tp, feats && do_f ? g_f : nullptr, feats && do_b ? g_b : nullptr, feats ? nullptr : logits,
          feats ? nullptr : logits_fwd, feats ? nullptr : logits_bwd);
    }
    if (time_end) cudaEventRecord(time_end, stream);
    if (check_launch() != kOk) return kErrCuda;
    if (!feats) return kOk;
    // feature path: state_net.4 as a second GEMM, logits from the fp32 features, softmax-weighted blend
    gemm::EpiBiasF32::Params fp;
    fp.bias = a.w2_bias, fp.out32 = FEAT, fp.ld32 = D;
    // state_net.4 over [hi | lo] rows of g and of W_2 (a.w2_ld >= 2D: lo block at column D): fp32-grade features
    const bool w2_split = a.w2_ld >= 2 * D;
    int rc = gemm::launch<FMT, gemm::EpiBiasF32>(G, 2 * D, a.w2_16, w2_split ? a.w2_ld : D, static_cast<int>(nd * n), D, D,
                                                 fp, a.num_sms, stream, nullptr, nullptr, 0,
                                                 gemm::make_split(D, w2_split ? D : -1));
    if (rc != kOk) return rc;
    const float* f_f = do_f ? FEAT : nullptr;
    const float* f_b = do_b ? FEAT + (do_f ? n * D : 0) : nullptr;
    const int tgrid = static_cast<int>((E + 7) / 8);
    feature_blend_kernel<<<tgrid, 256, 0, stream>>>(f_f, f_b, a.head_w, a.head_b, E, edge_emb, logits, logits_fwd,
                                                    logits_bwd);
    return check_launch();
  }
};

static int env_int(const char* name, int fallback) {
  const char* v = getenv(name);
  return v != nullptr && *v != '\0' ? atoi(v) : fallback;
}

template <int FMT>
static int score_edges_impl(const evi_score_args& a, cudaStream_t stream) {
  const long long E = a.num_edges;
  if (E <= 0) return kOk;
  if (a.flags == nullptr || a.node_uv32 == nullptr || a.node_p32 == nullptr || a.pair_rq32 == nullptr ||
      a.pair_slot32 == nullptr || a.num_nodes <= 0 || a.num_rel_rows <= 0 || a.num_questions <= 0 ||
      a.num_pair_rows <= 0)
    return kErrInvalidArg;
  const long long chunk = a.chunk_edges > 0 && a.chunk_edges < E ? a.chunk_edges : E;
  const long long nchunks = (E + chunk - 1) / chunk;
  const size_t per = chunk_workspace_bytes(chunk, a.direction, a.want_features, a.precision == EVI_PRECISION_STRICT);
  if (a.workspace_bytes < per) return kErrWorkspace;
  uint8_t* ws = static_cast<uint8_t*>(a.workspace);
  // experiment knobs (A/B on the GPU box): tail kernel flavour / CTA size, SMs given to the GEMM
  static const int tail_mode_serial = env_int("EVI_TAIL_LEGACY", 0) ? 1 : 0;
  static const int tail_mode_pipe = env_int("EVI_PIPE_TAIL", 1);
  static const int tail_threads = env_int("EVI_TAIL_THREADS", 128);
  static const int gemm_sms_env = env_int("EVI_GEMM_SMS", 0);
  const int gemm_sms = gemm_sms_env > 0 && gemm_sms_env < a.num_sms ? gemm_sms_env : a.num_sms;
  cudaStream_t aux = static_cast<cudaStream_t>(a.aux_stream);
  const bool pipelined = aux != nullptr && aux != stream && nchunks >= 2 && a.workspace_bytes >= 2 * per &&
                         !a.want_features;
  if (!pipelined) {
    for (long long c = 0; c < nchunks; ++c) {
      const bool timed = c == (nchunks > 1 ? 1 : 0);
      const long long e0 = c * chunk, n = E - e0 < chunk ? E - e0 : chunk;
      ChunkPlan<FMT> plan(a, e0, n, ws, per);
      if (!plan.ok) return kErrWorkspace;
      int rc = plan.operand(stream);
      if (rc != kOk) return rc;
      rc = plan.gemms(stream, gemm_sms, timed ? static_cast<cudaEvent_t>(a.ev_gemm_begin) : nullptr,
                      timed ? static_cast<cudaEvent_t>(a.ev_gemm_end) : nullptr);
      if (rc != kOk) return rc;
      rc = plan.tail(stream, tail_mode_serial, 256, timed ? static_cast<cudaEvent_t>(a.ev_tail_begin) : nullptr,
                     timed ? static_cast<cudaEvent_t>(a.ev_tail_end) : nullptr);
      if (rc != kOk) return rc;
    }
    return kOk;
  }
  // Pipelined over two streams and two workspace halves: the GEMMs of chunk c (caller's stream, tensor-bound) run
  // while the aux stream does the tail of chunk c-1 and the operand rows of chunk c+1 (HBM-bound; their CTAs use no
  // shared memory and fit beside the resident GEMM CTAs). Aux-stream order: op0, op1, tail0, op2, tail1, ...
  // so op(c) follows tail(c-2) (same workspace half) and gemm(c) waits for op(c), which follows tail(c-2).
  const int n_ev = static_cast<int>(2 * nchunks + 2);
  cudaEvent_t evs[2 * 64 + 2];
  if (nchunks > 64) return kErrUnsupported;
  for (int i = 0; i < n_ev; ++i)
    if (cudaEventCreateWithFlags(&evs[i], cudaEventDisableTiming) != cudaSuccess) return kErrCuda;
  cudaEvent_t* ev_op = evs;
  cudaEvent_t* ev_g = evs + nchunks;
  cudaEvent_t ev_start = evs[2 * nchunks], ev_end = evs[2 * nchunks + 1];
  auto plan_of = [&](long long c) {
    const long long e0 = c * chunk, n = E - e0 < chunk ? E - e0 : chunk;
    return ChunkPlan<FMT>(a, e0, n, ws + (c & 1) * per, per);
  };
  int rc = kOk;
  cudaEventRecord(ev_start, stream);
  cudaStreamWaitEvent(aux, ev_start, 0);
  rc = plan_of(0).operand(aux);
  cudaEventRecord(ev_op[0], aux);
  for (long long c = 0; c < nchunks && rc == kOk; ++c) {
    const bool timed = c == 1;
    cudaStreamWaitEvent(stream, ev_op[c], 0);
    rc = plan_of(c).gemms(stream, gemm_sms, timed ? static_cast<cudaEvent_t>(a.ev_gemm_begin) : nullptr,
                          timed ? static_cast<cudaEvent_t>(a.ev_gemm_end) : nullptr);
    cudaEventRecord(ev_g[c], stream);
    if (rc == kOk && c >= 1) {
      cudaStreamWaitEvent(aux, ev_g[c - 1], 0);
      rc = plan_of(c - 1).tail(aux, tail_mode_pipe, tail_threads);
    }
    if (rc == kOk && c + 1 < nchunks) {
      rc = plan_of(c + 1).operand(aux);
      cudaEventRecord(ev_op[c + 1], aux);
    }
  }
  if (rc == kOk) {
    cudaStreamWaitEvent(aux, ev_g[nchunks - 1], 0);
    rc = plan_of(nchunks - 1).tail(aux, tail_mode_pipe, tail_threads);
  }
  cudaEventRecord(ev_end, aux);
  cudaStreamWaitEvent(stream, ev_end, 0);  // the caller's stream owns the results again
  for (int i = 0; i < n_ev; ++i) cudaEventDestroy(evs[i]);
  return rc;
}

}  // namespace evi

using namespace evi;

extern "C" {

size_t evi_score_workspace_bytes(long long chunk_edges, int direction, int want_features) {
  return chunk_workspace_bytes(chunk_edges, direction, want_features, 0);
}
size_t evi_score_workspace_bytes2(long long chunk_edges, int direction, int want_features, int precision) {
  return chunk_workspace_bytes(chunk_edges, direction, want_features, precision == EVI_PRECISION_STRICT);
}

int evi_gather_rows2(const void* src, int src_is_f32, long long src_rows, int src_ld, const long long* ids,
                     const long long* sub_mask_ids, const float* sub_row, int num_rows, const int* num_rows_dev, int dim,
                     void* dst16, int dst_ld, int split, int operand_format, unsigned* flags, cudaStream_t stream) {
  if (dim != D) return kErrUnsupported;
  if (operand_format == kFmtF16)
    return gather_rows_impl<kFmtF16>(src, src_is_f32, src_rows, src_ld, ids, sub_mask_ids, sub_row, num_rows,
                                     num_rows_dev, dst16, dst_ld, split, flags, stream);
  if (operand_format == kFmtBF16)
    return gather_rows_impl<kFmtBF16>(src, src_is_f32, src_rows, src_ld, ids, sub_mask_ids, sub_row, num_rows,
                                      num_rows_dev, dst16, dst_ld, split, flags, stream);
  return kErrInvalidArg;
}

int evi_gather_rows(const void* src, int src_is_f32, long long src_rows, const long long* ids,
                    const long long* sub_mask_ids, const float* sub_row, int num_rows, int dim, void* dst16,
                    int operand_format, unsigned* flags, cudaStream_t stream) {
  return evi_gather_rows2(src, src_is_f32, src_rows, dim, ids, sub_mask_ids, sub_row, num_rows, nullptr, dim, dst16, dim,
                          0, operand_format, flags, stream);
}

int evi_struct_project(const float* node_struct, int num_nodes, int ns_dim, const float* struct_w_t,
                       const float* struct_b, float* node_uv32, cudaStream_t stream) {
  if (num_nodes <= 0) return kOk;
  if (ns_dim < 2 || ns_dim > kMaxNodeStruct || node_struct == nullptr || struct_w_t == nullptr ||
      struct_b == nullptr || node_uv32 == nullptr)
    return kErrInvalidArg;
  node_struct_proj_kernel<<<(num_nodes + kUvNodes - 1) / kUvNodes, 128, 0, stream>>>(node_struct, num_nodes, ns_dim, struct_w_t,
                                                                   struct_b, node_uv32);
  return check_launch();
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

int evi_project_split(const void* x16, int ld_x, int x_lo_off, int num_rows, const int* num_rows_dev, const void* w16,
                      int ld_w, int w_lo_off, int out_dim, int in_dim, const float* bias, const int* act_per_block,
                      void* out16, int ld_out16, int out_lo_off, float* out32, int operand_format, int num_sms,
                      cudaStream_t stream) {
  if (out_dim % gemm::BN != 0 || out_dim / gemm::BN > 8 || in_dim % gemm::BK != 0 || ld_w < in_dim || ld_x < in_dim)
    return kErrUnsupported;
  if (out16 != nullptr && ld_out16 < out_dim) return kErrInvalidArg;
  if (operand_format == kFmtF16)
    return linear_act_impl<kFmtF16>(x16, ld_x, x_lo_off, num_rows, num_rows_dev, w16, ld_w, w_lo_off, out_dim, in_dim,
                                    bias, act_per_block, out16, ld_out16, out_lo_off, out32, num_sms, stream);
  if (operand_format == kFmtBF16)
    return linear_act_impl<kFmtBF16>(x16, ld_x, x_lo_off, num_rows, num_rows_dev, w16, ld_w, w_lo_off, out_dim, in_dim,
                                     bias, act_per_block, out16, ld_out16, out_lo_off, out32, num_sms, stream);
  return kErrInvalidArg;
}

int evi_project(const void* x16, int num_rows, const int* num_rows_dev, const void* w16, int ld_w, int out_dim,
                int in_dim, const float* bias, const int* act_per_block, void* out16, float* out32,
                int operand_format, int num_sms, cudaStream_t stream) {
  return evi_project_split(x16, in_dim, -1, num_rows, num_rows_dev, w16, ld_w, -1, out_dim, in_dim, bias, act_per_block,
                           out16, out_dim, -1, out32, operand_format, num_sms, stream);
}

int evi_sc_center(const int* head32, const int* tail32, long long num_edges, const float* node_uv32, int num_nodes,
                  const float* struct_ln_g, const float* struct_ln_b, int direction, const float* w_sc32,
                  const float* b1, float* workspace, size_t workspace_bytes, float* center, float* bias_out,
                  cudaStream_t stream) {
  if (num_edges <= 0) return kOk;
  if (head32 == nullptr || tail32 == nullptr || node_uv32 == nullptr || struct_ln_g == nullptr || struct_ln_b == nullptr ||
      w_sc32 == nullptr || workspace == nullptr || center == nullptr || bias_out == nullptr || num_nodes <= 0)
    return kErrInvalidArg;
  if (workspace_bytes < static_cast<size_t>(kCenterCtas) * D * sizeof(float)) return kErrWorkspace;
  const int do_f = direction != EVI_DIR_BACKWARD, do_b = direction != EVI_DIR_FORWARD;
  const long long sample = num_edges < kCenterSample ? num_edges : kCenterSample;
  const int grid = static_cast<int>((sample + 7) / 8 < kCenterCtas ? (sample + 7) / 8 : kCenterCtas);
  sc_center_partial_kernel<<<grid, 256, 0, stream>>>(head32, tail32, num_edges, sample, node_uv32, num_nodes, struct_ln_g,
                                                     struct_ln_b, do_f, do_b, workspace);
  sc_center_finish_kernel<<<D / 256, 256, 0, stream>>>(workspace, grid, 1.0f / static_cast<float>(sample * (do_f + do_b)),
                                                       center);
  sc_center_bias_kernel<<<D / 8, 256, 0, stream>>>(w_sc32, center, b1, bias_out);
  return check_launch();
}

int evi_unique_rows(const long long* ids, int num_rows, int* slot_map, long long table_rows, long long* uniq_ids,
                    int* node_slot, int* num_unique, unsigned* flags, cudaStream_t stream) {
  if (num_rows <= 0) return kOk;
  if (ids == nullptr || slot_map == nullptr || uniq_ids == nullptr || node_slot == nullptr || num_unique == nullptr ||
      flags == nullptr || table_rows <= 0)
    return kErrInvalidArg;
  if (cudaMemsetAsync(num_unique, 0, sizeof(int), stream) != cudaSuccess) return kErrCuda;
  const int grid = (num_rows + 255) / 256;
  unique_claim_kernel<<<grid, 256, 0, stream>>>(ids, num_rows, slot_map, table_rows, uniq_ids, num_unique, flags);
  unique_lookup_kernel<<<grid, 256, 0, stream>>>(ids, num_rows, slot_map, table_rows, node_slot);
  unique_reset_kernel<<<grid, 256, 0, stream>>>(uniq_ids, num_unique, num_rows, slot_map);
  return check_launch();
}

int evi_edge_rep_index(const int* head32, const int* tail32, long long num_edges, const int* node_slot, int num_nodes,
                       const long long* edge_attr, int* head_rep32, int* tail_rep32, int* rel_slot32,
                       cudaStream_t stream) {
  if (num_edges <= 0) return kOk;
  if (node_slot != nullptr && (head32 == nullptr || tail32 == nullptr || head_rep32 == nullptr || tail_rep32 == nullptr))
    return kErrInvalidArg;
  if (edge_attr != nullptr && rel_slot32 == nullptr) return kErrInvalidArg;
  edge_rep_index_kernel<<<static_cast<int>((num_edges + 255) / 256), 256, 0, stream>>>(
      head32, tail32, num_edges, node_slot, num_nodes, edge_attr, head_rep32, tail_rep32, rel_slot32);
  return check_launch();
}

int evi_relation_pairs(const int* rel32, const long long* edge_ptr, int num_questions, int num_relations,
                       int capacity, int* pair_slot32, int* pair_q32, int* pair_rel32, int* num_pairs,
                       unsigned* flags, cudaStream_t stream) {
  if (num_questions <= 0) return kOk;
  if (num_relations <= 0 || capacity <= 0 || rel32 == nullptr || edge_ptr == nullptr || pair_slot32 == nullptr ||
      pair_q32 == nullptr || pair_rel32 == nullptr || num_pairs == nullptr || flags == nullptr)
    return kErrInvalidArg;
  const int words = (num_relations + 31) / 32;
  const size_t smem = static_cast<size_t>(words) * 8;
  if (smem > 200 * 1024) return kErrUnsupported;  // > 819k relation ids
  static std::atomic<unsigned long long> smem_ready{0};
  if (smem > 48 * 1024 && !ensure_dynamic_smem(relation_pairs_kernel, 200 * 1024, smem_ready)) return kErrCuda;
  if (cudaMemsetAsync(num_pairs, 0, sizeof(int), stream) != cudaSuccess) return kErrCuda;
  relation_pairs_kernel<<<num_questions, 256, smem, stream>>>(rel32, edge_ptr, num_relations, words, capacity,
                                                              pair_slot32, pair_q32, pair_rel32, num_pairs, flags);
  return check_launch();
}

int evi_pair_context2(const int* pair_q32, const int* pair_rel32, const int* num_pairs, int capacity,
                      const void* rel_repr16, int rel_ld, int num_rel_rows, const float* gate_bias32, int num_questions,
                      void* rc16, int rc_ld, int split, int operand_format, unsigned* flags, cudaStream_t stream) {
  if (capacity <= 0) return kOk;
  if (rel_ld < (split ? 2 * D : D) || rc_ld < (split ? 2 * D : D) || rel_ld % 8 != 0 || rc_ld % 8 != 0)
    return kErrInvalidArg;
  const int grid = (capacity + 7) / 8;
  auto go = [&](auto kern) {
    kern<<<grid, 256, 0, stream>>>(pair_q32, pair_rel32, num_pairs, capacity, static_cast<const uint16_t*>(rel_repr16),
                                   rel_ld, num_rel_rows, gate_bias32, num_questions, static_cast<uint16_t*>(rc16), rc_ld,
                                   flags);
  };
  if (operand_format == kFmtF16) {
    if (split) go(pair_context_kernel<kFmtF16, true>);
    else go(pair_context_kernel<kFmtF16, false>);
  } else if (operand_format == kFmtBF16) {
    if (split) go(pair_context_kernel<kFmtBF16, true>);
    else go(pair_context_kernel<kFmtBF16, false>);
  } else {
    return kErrInvalidArg;
  }
  return check_launch();
}

int evi_pair_context(const int* pair_q32, const int* pair_rel32, const int* num_pairs, int capacity,
                     const void* rel_repr16, int num_rel_rows, const float* gate_bias32, int num_questions,
                     void* rc16, int operand_format, unsigned* flags, cudaStream_t stream) {
  return evi_pair_context2(pair_q32, pair_rel32, num_pairs, capacity, rel_repr16, D, num_rel_rows, gate_bias32,
                           num_questions, rc16, D, 0, operand_format, flags, stream);
}

int evi_score_edges(const evi_score_args* args, cudaStream_t stream) {
  if (args == nullptr) return kErrInvalidArg;
  if (args->emb_dim != D || args->hidden_dim != D) return kErrUnsupported;
  if (args->operand_format == kFmtF16) return score_edges_impl<kFmtF16>(*args, stream);
  if (args->operand_format == kFmtBF16) return score_edges_impl<kFmtBF16>(*args, stream);
  return kErrInvalidArg;
}

}  // extern "C"
