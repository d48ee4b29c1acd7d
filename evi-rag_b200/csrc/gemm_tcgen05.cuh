// Persistent warp-specialised tcgen05 GEMM for sm_100a:  D[M,N] = A[M,K] * B[N,K]^T  (+ epilogue)
//
//  * A and B are 16-bit (fp16 or bf16), K-major (row-major [rows, K]); nn.Linear weights are
//    already [N, K], activations are [M, K]  ->  no transposes anywhere.
//  * TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) stages 128x64 A tiles and 256x64 B tiles into a
//    4-deep shared-memory ring; one elected thread issues tcgen05.mma (M=128, N=256, K=16) into a
//    double-buffered fp32 accumulator in TMEM (2 x 256 columns); four epilogue warps drain the
//    accumulator with tcgen05.ld while the next tile's MMAs run.
//  * Grid = min(#tiles, #SMs); tiles are walked n-fastest so the CTAs that share an A row-block
//    run at the same time and hit it in L2.
//
// The epilogue is a functor: Epi::apply(params, row, col0, acc[32]) is called once per row per
// 32-column chunk; Epi::finish_row(params, row, n_blk, state) once per row per tile.
#pragma once

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched at run time)

#include "common.cuh"

namespace evi {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 x 16-bit = 128 B = one swizzle row
constexpr int STAGES = 4;
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB
constexpr int B_STAGE_BYTES = BN * BK * 2;  // 32 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int ACC_STAGES = 2;
constexpr int TMEM_COLS = ACC_STAGES * BN;  // 512
constexpr int NUM_THREADS = 256;            // warps 0-3: TMA / MMA / TMEM-alloc / spare, 4-7: epilogue
// fp32 epilogues stage 32x32 sub-tiles (one 128-byte swizzle row per accumulator row) in shared memory
// and write them with TMA stores: 2 buffers x 4 KB per epilogue warp
constexpr int EPI_STAGE_BYTES = 32 * 128;
constexpr int EPI_SMEM_BYTES = 4 * 2 * EPI_STAGE_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_SMEM_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

struct TileCoord {
  int m_blk, n_blk;
};

// K-block schedule over split operands. A 16-bit operand row may carry a second half `lo = x - fp16(x)` at a column
// offset of the same matrix ([hi | lo] rows); the contraction then runs up to three passes over K into the same fp32
// accumulator:  A_hi B_hi  (+ A_lo B_hi)  (+ A_hi B_lo)   -- the lo * lo term is below fp32 rounding.
// Two fp16 halves carry ~22 significant bits, so a split contraction is as good as an fp32 one
// (tests/test_gpu_parity.py::test_project_split_matches_fp32); lo halves of small weights are fp16 subnormals, which
// the tensor core takes as they are.
struct SplitK {
  int passes;    // 1..3
  int a_off[3];  // K coordinate (elements) where pass p starts in A's rows
  int b_off[3];
};
inline SplitK make_split(int a_lo_off, int b_lo_off) {  // negative offset: that operand has no lo half
  SplitK s;
  s.passes = 1;
  for (int i = 0; i < 3; ++i) s.a_off[i] = 0, s.b_off[i] = 0;
  if (a_lo_off >= 0) s.a_off[s.passes] = a_lo_off, s.b_off[s.passes] = 0, ++s.passes;
  if (b_lo_off >= 0) s.a_off[s.passes] = 0, s.b_off[s.passes] = b_lo_off, ++s.passes;
  return s;
}

template <int FMT, class Epi>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
            const __grid_constant__ CUtensorMap tmap_c, int M, int N, int K, const int* __restrict__ m_dev,
            typename Epi::Params ep, const SplitK sk) {
  // optional device-side row count (e.g. the number of distinct (question, relation) pairs, known
  // only on the device): rows beyond it are neither loaded nor written
  if (m_dev != nullptr) {
    const int md = __ldg(m_dev);
    M = md < M ? md : M;
  }
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* epi_stage = smem + STAGES * STAGE_BYTES;  // 1024-byte aligned (STAGE_BYTES is a multiple of 1024)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + EPI_SMEM_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + ACC_STAGES);

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m_blocks = (M + BM - 1) / BM;
  const int num_n_blocks = (N + BN - 1) / BN;
  const int num_tiles = num_m_blocks * num_n_blocks;
  const int k_blocks_per_pass = (K + BK - 1) / BK;
  const int num_k_blocks = k_blocks_per_pass * sk.passes;  // split operands: up to three passes over K

  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (Epi::kTmaStore) tma_prefetch_desc(&tmap_c);
  }
  if (warp_idx == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < ACC_STAGES; ++i) {
      mbar_init(&tmem_full_bar[i], 1);
      mbar_init(&tmem_empty_bar[i], 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp_idx == 2) {
    tmem_alloc(tmem_base_slot, TMEM_COLS);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp_idx == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / num_n_blocks, n_blk = tile % num_n_blocks;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          const int pass = kb / k_blocks_per_pass, kk = (kb - pass * k_blocks_per_pass) * BK;
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_STAGE_BYTES;
          mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
          tma_load_2d(sa, &tmap_a, &full_bar[stage], sk.a_off[pass] + kk, m_blk * BM);
          tma_load_2d(sb, &tmap_b, &full_bar[stage], sk.b_off[pass] + kk, n_blk * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = umma_idesc_f16(FMT, BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u);
      tc_fence_after_sync();
      const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
      for (int kb = 0; kb < num_k_blocks; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after_sync();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint64_t da = umma_smem_desc_sw128(sa);
          const uint64_t db = umma_smem_desc_sw128(sa + A_STAGE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // +32 bytes per K=16 step inside the 128-byte swizzle row (start address is >> 4)
            umma_f16_ss(tmem_d, da + static_cast<uint64_t>(k * 2), db + static_cast<uint64_t>(k * 2),
                        idesc, static_cast<uint32_t>((kb | k) != 0));
          }
          umma_commit(&empty_bar[stage]);                               // smem slot reusable
          if (kb == num_k_blocks - 1) umma_commit(&tmem_full_bar[acc]);  // accumulator ready
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1u;
        }
      }
      if (++acc == ACC_STAGES) {
        acc = 0;
        acc_phase ^= 1u;
      }
    }
  } else if (warp_idx >= 4) {
    // ===================== epilogue (TMEM -> registers -> global) =====================
    const int quarter = warp_idx & 3;  // TMEM lane quarter this warp may read
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / num_n_blocks, n_blk = tile % num_n_blocks;
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after_sync();
      const int row = m_blk * BM + quarter * 32 + lane;
      const uint32_t taddr =
          tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(acc * BN);
      typename Epi::RowState st;
      Epi::init_row(st);
      static_assert(!Epi::kOut16, "the 16-bit TMA-store epilogue exists in the CTA-pair kernel only");
      if constexpr (Epi::kTmaStore) {
        // accumulator row -> registers -> epilogue math -> swizzled 32x32 fp32 sub-tile in shared memory ->
        // one TMA store per sub-tile (full 128-byte lines, clipped at the matrix edge by the tensor map)
        uint8_t* my_stage = epi_stage + quarter * 2 * EPI_STAGE_BYTES;
        const int row0 = m_blk * BM + quarter * 32;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + static_cast<uint32_t>(c * 32), r);
          uint8_t* buf = my_stage + (c & 1) * EPI_STAGE_BYTES;
          if (lane == 0) bulk_wait_group_read<1>();  // the store issued two sub-tiles ago has left `buf`
          __syncwarp();
          tmem_ld_wait();
          const int col0 = n_blk * BN + c * 32;
          Epi::template transform<FMT>(ep, st, row, col0, r);
          uint8_t* dst = buf + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) =
                make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (row0 < M && col0 < N) tma_store_2d(&tmap_c, buf, col0, row0);
            bulk_commit_group();  // always (an empty group is fine): keeps the buffer accounting of wait_group exact
          }
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + static_cast<uint32_t>(c * 32), r);
          tmem_ld_wait();
          const int col0 = n_blk * BN + c * 32;
          if (row < M && col0 < N) Epi::template apply<FMT>(ep, st, row, col0, r);
        }
        if (row < M) Epi::finish_row(ep, st, row, n_blk);
      }
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
      if (++acc == ACC_STAGES) {
        acc = 0;
        acc_phase ^= 1u;
      }
    }
    if (Epi::kTmaStore && lane == 0) bulk_wait_group<0>();  // every store of this warp has been performed
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp_idx == 2) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// --------------------------------------------------------------------------------------------
// Epilogues
// --------------------------------------------------------------------------------------------
enum Act : int { kActNone = 0, kActTanh = 1, kActSigmoid = 2 };

// out = act(acc + bias[n]); activation chosen per 256-column block (lets [W_gate ; W_bias] run as
// one N=2048 GEMM). Writes a 16-bit copy (next GEMM operand / gather source) and/or fp32.
struct EpiBiasAct {
  static constexpr bool kTmaStore = false;
  static constexpr bool kOut16 = false;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;   // [N]
    void* out16;         // [M, ld16] 16-bit or null
    float* out32;        // [M, ld32] or null
    int ld16, ld32;
    int act[8];          // per n-block activation
    int lo_off;          // >= 0: out16 rows are [hi | lo], lo = out - hi stored at this column offset; < 0: hi only
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void apply(const Params& p, RowState&, int row, int col0, uint32_t (&r)[32]) {
    const int act = p.act[(col0 / BN) & 7];
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = p.bias != nullptr ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
      v[i + 0] = __uint_as_float(r[i + 0]) + b.x;
      v[i + 1] = __uint_as_float(r[i + 1]) + b.y;
      v[i + 2] = __uint_as_float(r[i + 2]) + b.z;
      v[i + 3] = __uint_as_float(r[i + 3]) + b.w;
    }
    if (act == kActTanh) {
      if (p.out32 == nullptr && p.lo_off < 0) {
        // 16-bit output only (the per-node / per-relation projections): tanh_fast is good to ~1e-6, far below the
        // 2^-12 rounding of the store, and keeps the epilogue shorter than the tile's MMA time
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = tanh_fast(v[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = tanhf(v[i]);
      }
    } else if (act == kActSigmoid) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 1.0f / (1.0f + expf(-v[i]));
    }
    if (p.out32 != nullptr) {
      float4* o = reinterpret_cast<float4*>(p.out32 + static_cast<size_t>(row) * p.ld32 + col0);
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    if (p.out16 != nullptr) {
      uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out16) +
                                          static_cast<size_t>(row) * p.ld16 + col0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 w;
        w.x = Op16<FMT>::pack2(v[8 * i + 0], v[8 * i + 1]);
        w.y = Op16<FMT>::pack2(v[8 * i + 2], v[8 * i + 3]);
        w.z = Op16<FMT>::pack2(v[8 * i + 4], v[8 * i + 5]);
        w.w = Op16<FMT>::pack2(v[8 * i + 6], v[8 * i + 7]);
        o[i] = w;
        if (p.lo_off >= 0) {  // residual half: the next contraction reads [hi | lo] rows (SplitK)
          const float2 h0 = Op16<FMT>::unpack2(w.x), h1 = Op16<FMT>::unpack2(w.y), h2 = Op16<FMT>::unpack2(w.z),
                       h3 = Op16<FMT>::unpack2(w.w);
          uint4 l;
          l.x = Op16<FMT>::pack2(v[8 * i + 0] - h0.x, v[8 * i + 1] - h0.y);
          l.y = Op16<FMT>::pack2(v[8 * i + 2] - h1.x, v[8 * i + 3] - h1.y);
          l.z = Op16<FMT>::pack2(v[8 * i + 4] - h2.x, v[8 * i + 5] - h2.y);
          l.w = Op16<FMT>::pack2(v[8 * i + 6] - h3.x, v[8 * i + 7] - h3.y);
          o[i + p.lo_off / 8] = l;
        }
      }
    }
  }
  static EVI_DEVINL void finish_row(const Params&, RowState&, int, int) {}
};

// state_net.0 pre-activation: z = acc + dist[m] * w_dist[n] + b1[n]  (the 3073rd input column of
// retriever.py:481 folded in as a rank-1 update), stored fp32 (a 16-bit z costs ~2.5e-4 of the
// 1e-3 score budget), plus fp32 LayerNorm partial sums (sum, sum of squares) per row per
// 256-column block.
struct EpiPreNorm {
  static constexpr bool kTmaStore = false;
  static constexpr bool kOut16 = false;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;    // [N] state_net.0.bias
    const float* w_dist;  // [N] state_net.0.weight[:, 3D]
    const float* dist;    // [M]  -||err||
    float* z32;           // [M, ldz]
    float2* stats;        // [M, N / BN]
    int ldz, n_blocks;
  };
  struct RowState {
    float sum, sumsq, dist;
  };
  static EVI_DEVINL void init_row(RowState& s) { s.sum = 0.f, s.sumsq = 0.f, s.dist = 0.f; }
  template <int FMT>
  static EVI_DEVINL void apply(const Params& p, RowState& s, int row, int col0, uint32_t (&r)[32]) {
    const float d = __ldg(p.dist + row);
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i));
      const float4 w = __ldg(reinterpret_cast<const float4*>(p.w_dist + col0 + i));
      v[i + 0] = __uint_as_float(r[i + 0]) + fmaf(d, w.x, b.x);
      v[i + 1] = __uint_as_float(r[i + 1]) + fmaf(d, w.y, b.y);
      v[i + 2] = __uint_as_float(r[i + 2]) + fmaf(d, w.z, b.z);
      v[i + 3] = __uint_as_float(r[i + 3]) + fmaf(d, w.w, b.w);
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      s.sum += v[i];
      s.sumsq = fmaf(v[i], v[i], s.sumsq);
    }
    float4* o = reinterpret_cast<float4*>(p.z32 + static_cast<size_t>(row) * p.ldz + col0);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  static EVI_DEVINL void finish_row(const Params& p, RowState& s, int row, int n_blk) {
    p.stats[static_cast<size_t>(row) * p.n_blocks + n_blk] = make_float2(s.sum, s.sumsq);
  }
};

// state_net.4 output: feat = acc + b2 (fp32, retriever.py:180) — only when edge features are requested
struct EpiBiasF32 {
  static constexpr bool kTmaStore = false;
  static constexpr bool kOut16 = false;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;
    float* out32;
    int ld32;
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void apply(const Params& p, RowState&, int row, int col0, uint32_t (&r)[32]) {
    float4* o = reinterpret_cast<float4*>(p.out32 + static_cast<size_t>(row) * p.ld32 + col0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 4 * i));
      o[i] = make_float4(__uint_as_float(r[4 * i]) + b.x, __uint_as_float(r[4 * i + 1]) + b.y,
                         __uint_as_float(r[4 * i + 2]) + b.z, __uint_as_float(r[4 * i + 3]) + b.w);
    }
  }
  static EVI_DEVINL void finish_row(const Params&, RowState&, int, int) {}
};

// Plain accumulator rows stored as fp16 through the same staged TMA-store path (two 32-column sub-tiles share one
// 128-byte staging line, one store per 32 x 64 box). Used for W_iv m (CTA-pair kernel only): the term is one addend
// of the state_net.0 pre-activation and its 2^-12 rounding is far below the operand rounding already in z
// (scripts/precision_model.py), while the 16-bit rows halve what this launch writes and what the tail re-reads.
struct EpiF16Tma {
  static constexpr bool kTmaStore = true;
  static constexpr bool kOut16 = true;
  static constexpr int kOutFmt = kFmtF16;  // always fp16 (-1: the operand format)
  struct Params {};
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void transform(const Params&, RowState&, int, int, uint32_t (&)[32]) {}
};

// W_sc sc rows as 16-bit block fixed point (CTA-pair kernel only). fp16 rows would add a 2^-12 relative rounding to
// the largest addend of the state_net.0 pre-activation (scripts/precision_model.py: 6.4e-4 -> 7.3e-4 on the logits);
// one fp32 scale per (row, 256-column tile) and round-to-nearest 16-bit integers keep 15 bits against the tile
// maximum (6.41e-4 -> 6.44e-4) while the launch writes, and the tail re-reads, half the bytes of fp32 rows.
// The epilogue thread owns one accumulator row of the tile: a first pass over its TMEM columns finds max|acc|, the
// second quantises u = rn(acc * 32767 / max) + 32768 through the 1.5 * 2^23 magic add (the low 16 bits of the sum)
// and goes through the 16-bit staged TMA store. scale[row * ld_scale + n_blk] = max / 32767 (0 when the tile row is 0).
struct EpiQ16Tma {
  static constexpr bool kTmaStore = true;
  static constexpr bool kOut16 = true;
  static constexpr bool kQuant = true;
  static constexpr int kOutFmt = kFmtF16;  // 16-bit boxes; the tensor map only moves bytes
  struct Params {
    float* scale;  // [M, ld_scale]
    int ld_scale;  // number of 256-column tiles per row
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void transform(const Params&, RowState&, int, int, uint32_t (&)[32]) {}
};
constexpr float kQ16Magic = 12582912.0f + 32768.0f;  // 1.5 * 2^23 + the unsigned bias

template <class E, class = void>
struct epi_is_quant {
  static constexpr bool value = false;
};
template <class E>
struct epi_is_quant<E, decltype(void(E::kQuant))> {
  static constexpr bool value = E::kQuant;
};

// out16 = act(acc + bias[n]) in the operand format through the same 16-bit staged TMA store: the per-node /
// per-relation / per-question projections whose only consumer reads 16-bit rows (EpiBiasAct's per-thread row
// stores made those launches epilogue-bound).
struct EpiBiasAct16Tma {
  static constexpr bool kTmaStore = true;
  static constexpr bool kOut16 = true;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;  // [N] or null
    int act[8];         // per n-block activation
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void transform(const Params& p, RowState&, int, int col0, uint32_t (&r)[32]) {
    const int act = p.act[(col0 / BN) & 7];
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = p.bias != nullptr ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
      v[i + 0] = __uint_as_float(r[i + 0]) + b.x;
      v[i + 1] = __uint_as_float(r[i + 1]) + b.y;
      v[i + 2] = __uint_as_float(r[i + 2]) + b.z;
      v[i + 3] = __uint_as_float(r[i + 3]) + b.w;
    }
    if (act == kActTanh) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = tanh_fast(v[i]);  // ~1e-6, below the 16-bit rounding of the store
    } else if (act == kActSigmoid) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 1.0f / (1.0f + expf(-v[i]));
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
  }
};

// [hi | lo] 16-bit rows through the staged TMA store (CTA-pair kernel only): out = act(acc + bias[n]) with exact
// tanhf / expf, hi = 16-bit(out) at column n, lo = 16-bit(out - hi) at column lo_off + n of the same row. The rows are
// the A operand of a later split contraction (gemm::SplitK): entity rows for P = W_err node_repr, projected question
// rows for gate | bias. Both halves of a 32 x 64 box are staged and stored together.
struct EpiBiasActHiLoTma {
  static constexpr bool kTmaStore = true;
  static constexpr bool kOut16 = true;
  static constexpr bool kHiLo = true;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;  // [N] or null
    int act[8];         // per n-block activation
    int lo_off;         // column offset of the lo half (multiple of 64)
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void transform(const Params& p, RowState&, int, int col0, uint32_t (&r)[32]) {
    const int act = p.act[(col0 / BN) & 7];
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = p.bias != nullptr ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
      v[i + 0] = __uint_as_float(r[i + 0]) + b.x;
      v[i + 1] = __uint_as_float(r[i + 1]) + b.y;
      v[i + 2] = __uint_as_float(r[i + 2]) + b.z;
      v[i + 3] = __uint_as_float(r[i + 3]) + b.w;
    }
    if (act == kActTanh) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = tanhf(v[i]);  // the lo half keeps ~22 bits: no fast approximation here
    } else if (act == kActSigmoid) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 1.0f / (1.0f + expf(-v[i]));
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
  }
};

template <class E, class = void>
struct epi_is_hilo {
  static constexpr bool value = false;
};
template <class E>
struct epi_is_hilo<E, decltype(void(E::kHiLo))> {
  static constexpr bool value = E::kHiLo;
};

// fp32 output through the staged TMA store: out = act(acc + bias[n]) with bias optional and the
// activation chosen per 256-column block. Covers the partial products W_iv m / W_sc sc of state_net.0
// (no bias, no activation), P = W_err node_repr, Rq = W_err r_ctx + b_1 and the gate|bias projection.
struct EpiF32Tma {
  static constexpr bool kTmaStore = true;
  static constexpr bool kOut16 = false;
  static constexpr int kOutFmt = -1;
  struct Params {
    const float* bias;  // [N] or null
    int act[8];         // per n-block activation
  };
  struct RowState {};
  static EVI_DEVINL void init_row(RowState&) {}
  template <int FMT>
  static EVI_DEVINL void transform(const Params& p, RowState&, int, int col0, uint32_t (&r)[32]) {
    const int act = p.act[(col0 / BN) & 7];
    if (p.bias == nullptr && act == kActNone) return;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 b = p.bias != nullptr ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + i))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
      v[i + 0] = __uint_as_float(r[i + 0]) + b.x;
      v[i + 1] = __uint_as_float(r[i + 1]) + b.y;
      v[i + 2] = __uint_as_float(r[i + 2]) + b.z;
      v[i + 3] = __uint_as_float(r[i + 3]) + b.w;
    }
    if (act == kActTanh) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = tanhf(v[i]);
    } else if (act == kActSigmoid) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 1.0f / (1.0f + expf(-v[i]));
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
  }
};

// --------------------------------------------------------------------------------------------
// Host side: tensor maps + launch
// --------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// row-major [rows, cols] fp32 output matrix with row pitch ld (elements); box = 32 rows x 32 cols (128 B)
inline bool make_tmap_out_f32(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 4};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// row-major [rows, cols] 16-bit output matrix with row pitch ld (elements); box = 32 rows x 64 cols (128 B)
inline bool make_tmap_out_16(CUtensorMap* map, int fmt, const void* base, uint64_t rows, uint64_t cols, uint64_t ld) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, fmt == kFmtBF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                  const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// row-major [rows, cols] 16-bit matrix with row pitch ld (elements); box = box_rows x 64 cols
inline bool make_tmap_2d(CUtensorMap* map, int fmt, const void* base, uint64_t rows, uint64_t cols,
                         uint64_t ld, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(BK), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, fmt == kFmtBF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                  2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// columns a tensor map must address from the operand's base pointer: K, or the end of the last split pass
inline int split_width(const SplitK& sk, const int (&off)[3], int K) {
  int w = K;
  for (int i = 0; i < sk.passes; ++i) w = off[i] + K > w ? off[i] + K : w;
  return w;
}

template <int FMT, class Epi>
inline int launch(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                  const typename Epi::Params& ep, int num_sms, cudaStream_t stream, const int* m_dev = nullptr,
                  float* out32 = nullptr, int ldc = 0, const SplitK& sk = make_split(-1, -1)) {
  if (M <= 0) return kOk;
  if (K % 8 != 0 || lda % 8 != 0 || ldb % 8 != 0 || N % 32 != 0) return kErrInvalidArg;
  if (sk.passes > 1 && K % BK != 0) return kErrInvalidArg;  // a pass must end on a k-block boundary
  const int wa = split_width(sk, sk.a_off, K), wb = split_width(sk, sk.b_off, K);
  if (wa > lda || wb > ldb) return kErrInvalidArg;
  CUtensorMap ta, tb;
  if (!make_tmap_2d(&ta, FMT, A, static_cast<uint64_t>(M), static_cast<uint64_t>(wa),
                    static_cast<uint64_t>(lda), BM))
    return kErrCuda;
  if (!make_tmap_2d(&tb, FMT, B, static_cast<uint64_t>(N), static_cast<uint64_t>(wb),
                    static_cast<uint64_t>(ldb), BN))
    return kErrCuda;
  CUtensorMap tc = ta;  // placeholder when the epilogue stores by itself
  if (Epi::kTmaStore) {
    if (out32 == nullptr || ldc % 4 != 0 || (reinterpret_cast<uintptr_t>(out32) & 15) != 0) return kErrInvalidArg;
    if (!make_tmap_out_f32(&tc, out32, static_cast<uint64_t>(M), static_cast<uint64_t>(N), static_cast<uint64_t>(ldc)))
      return kErrCuda;
  }
  auto kern = gemm_kernel<FMT, Epi>;
  static std::atomic<unsigned long long> smem_ready{0};  // per template instantiation
  if (!ensure_dynamic_smem(kern, SMEM_BYTES, smem_ready)) return kErrCuda;
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int grid = tiles < num_sms ? tiles : num_sms;
  kern<<<grid, NUM_THREADS, SMEM_BYTES, stream>>>(ta, tb, tc, M, N, K, m_dev, ep, sk);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // namespace gemm
}  // namespace evi
