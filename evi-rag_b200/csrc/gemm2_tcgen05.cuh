// CTA-pair tcgen05 GEMM for sm_100a:  D[M,N] = A[M,K] * B[N,K]^T  with 256x256 tiles computed by two CTAs
// (the two SMs of a TPC, a 2-CTA cluster) through tcgen05.mma.cta_group::2.
//
// Why a second kernel: in the one-CTA kernel (gemm_tcgen05.cuh) every 128x256 tile streams 48 KB of
// operands per 64-wide k-block through shared memory for 512 tensor-pipe cycles (96 B/cycle per SM); the
// pair splits the B tile between the two CTAs (each loads 128 of its 256 rows, the tensor cores read both
// halves), so each SM streams 32 KB per k-block (64 B/cycle) and the ring holds 6 stages instead of 4.
//
//  * each CTA: TMA producer warp (its 128 rows of A, its 128 rows of B, completion on the leader's full
//    barrier), epilogue warps 4-7 (its 128 accumulator rows: TMEM lanes 0-127, 2 x 256 columns double-buffered);
//  * leader CTA (cluster rank 0) only: one thread issues the MMAs for the pair and commits them to the
//    empty / tmem-full barriers of both CTAs (multicast commit);
//  * the epilogue warps of both CTAs arrive on the leader's tmem-empty barrier (remote mbarrier arrive).
// Epilogues are the functors of gemm_tcgen05.cuh.
#pragma once

#include "gemm_tcgen05.cuh"

namespace evi {
namespace gemm2 {

constexpr int BM = 256;       // per pair
constexpr int BM_CTA = 128;   // rows of A / accumulator rows per CTA
constexpr int BN = 256;
constexpr int BN_CTA = 128;   // rows of B loaded per CTA
constexpr int BK = 64;
constexpr int STAGES = 6;
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM_CTA * BK * 2;  // 16 KB
constexpr int B_STAGE_BYTES = BN_CTA * BK * 2;  // 16 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int ACC_STAGES = 2;
constexpr int TMEM_COLS = ACC_STAGES * BN;  // 512
constexpr int NUM_THREADS = 256;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + gemm::EPI_SMEM_BYTES + 1024 + 256;

EVI_DEVINL uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
EVI_DEVINL void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
EVI_DEVINL uint32_t map_to_cta(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
EVI_DEVINL void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// bounded wait: a protocol bug turns into a trap (an error the host sees) instead of a hung GPU
EVI_DEVINL void mbar_wait_guarded(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  for (uint32_t spins = 0;; ++spins) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if (spins > 2000u) __trap();  // each failed try_wait already suspends for up to ~10 ms
  }
}
EVI_DEVINL void tmem_alloc_pair(uint32_t* smem_result, uint32_t ncols) {  // whole warp, same warp id in both CTAs
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
EVI_DEVINL void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// this CTA's half of a pair operand; the transaction bytes are credited to `leader_bar` (shared::cluster address)
EVI_DEVINL void tma_load_2d_pair(void* smem_dst, const void* tmap, uint32_t leader_bar, int c_inner, int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(leader_bar), "r"(c_inner), "r"(c_outer)
      : "memory");
}
EVI_DEVINL void umma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this shared-memory offset in both CTAs once the pair's MMAs issued so far are done
EVI_DEVINL void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}

template <int FMT, class Epi>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
             const __grid_constant__ CUtensorMap tmap_c, int M, int N, int K, const int* __restrict__ m_dev,
             typename Epi::Params ep, const gemm::SplitK sk) {
  if (m_dev != nullptr) {
    const int md = __ldg(m_dev);
    M = md < M ? md : M;
  }
  extern __shared__ uint8_t smem_raw[];
  // identical offsets in both CTAs (same kernel image, same dynamic-shared base): the pair's MMA addresses the
  // peer's operand tiles by the leader's offsets
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* epi_stage = smem + STAGES * STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + gemm::EPI_SMEM_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint64_t* tmem_empty_bar = tmem_full_bar + ACC_STAGES;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + ACC_STAGES);

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_m_blocks = (M + BM - 1) / BM;
  const int num_n_blocks = (N + BN - 1) / BN;
  const int num_tiles = num_m_blocks * num_n_blocks;
  const int k_blocks_per_pass = (K + BK - 1) / BK;
  const int num_k_blocks = k_blocks_per_pass * sk.passes;  // split operands: up to three passes over K (gemm::SplitK)

  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (Epi::kTmaStore) tma_prefetch_desc(&tmap_c);
  }
  if (warp_idx == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);   // the leader's producer arms it; both CTAs' TMA bytes complete it
      mbar_init(&empty_bar[i], 1);  // one multicast commit from the leader's MMA thread
    }
    for (int i = 0; i < ACC_STAGES; ++i) {
      mbar_init(&tmem_full_bar[i], 1);   // multicast commit
      mbar_init(&tmem_empty_bar[i], 8);  // 4 epilogue warps x 2 CTAs (the leader's copy is the one waited on)
    }
    fence_barrier_init();
  }
  if (warp_idx == 2) tmem_alloc_pair(tmem_base_slot, TMEM_COLS);
  tc_fence_before_sync();
  cluster_sync_all();  // barriers of both CTAs are initialised before anyone signals across the pair
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;

  if (warp_idx == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        const int m_blk = tile / num_n_blocks, n_blk = tile % num_n_blocks;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          const int pass = kb / k_blocks_per_pass, kk = (kb - pass * k_blocks_per_pass) * BK;
          mbar_wait_guarded(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_STAGE_BYTES;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          const uint32_t lbar = map_to_cta(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_pair(sa, &tmap_a, lbar, sk.a_off[pass] + kk, m_blk * BM + static_cast<int>(rank) * BM_CTA);
          tma_load_2d_pair(sb, &tmap_b, lbar, sk.b_off[pass] + kk, n_blk * BN + static_cast<int>(rank) * BN_CTA);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_f16(FMT, BM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        mbar_wait_guarded(&tmem_empty_bar[acc], acc_phase ^ 1u);
        tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait_guarded(&full_bar[stage], phase);
          tc_fence_after_sync();
          if (lane == 0) {
            const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
            const uint64_t da = umma_smem_desc_sw128(sa);
            const uint64_t db = umma_smem_desc_sw128(sa + A_STAGE_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_f16_ss_pair(tmem_d, da + static_cast<uint64_t>(k * 2), db + static_cast<uint64_t>(k * 2), idesc,
                               static_cast<uint32_t>((kb | k) != 0));
            umma_commit_pair(&empty_bar[stage]);                               // slot reusable in both CTAs
            if (kb == num_k_blocks - 1) umma_commit_pair(&tmem_full_bar[acc]);  // accumulators ready in both CTAs
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
    }
  } else if (warp_idx >= 4) {
    // ===================== epilogue (both CTAs; own 128 accumulator rows) =====================
    const int quarter = warp_idx & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      const int m_blk = tile / num_n_blocks, n_blk = tile % num_n_blocks;
      mbar_wait_guarded(&tmem_full_bar[acc], acc_phase);
      tc_fence_after_sync();
      const int row0 = m_blk * BM + static_cast<int>(rank) * BM_CTA + quarter * 32;
      const int row = row0 + lane;
      const uint32_t taddr =
          tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(acc * BN);
      typename Epi::RowState st;
      Epi::init_row(st);
      if constexpr (Epi::kTmaStore && Epi::kOut16) {
        // fp16 rows: sub-tiles 2i and 2i+1 fill the two 64-byte halves of one staging line, one store per 32 x 64 box
        uint8_t* my_stage = epi_stage + quarter * 2 * gemm::EPI_STAGE_BYTES;
        constexpr bool kQuant = gemm::epi_is_quant<Epi>::value;
        constexpr bool kHiLo = gemm::epi_is_hilo<Epi>::value;  // both staging buffers per box: hi rows and lo rows
        float q_inv = 0.f;
        if constexpr (kQuant) {
          // first pass over this thread's accumulator row of the tile: max |acc| -> scale of the (row, tile) block
          float mx = 0.f;
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + static_cast<uint32_t>(c * 32), r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) mx = fmaxf(mx, fabsf(__uint_as_float(r[j])));
          }
          // an all-zero tile row gets scale 0 (not 1: the reader's folded offset -(2^23 + 32768) * scale must vanish
          // with the values, or its fp32 rounding would be an absolute error); a non-finite one gets NaN
          const bool live = mx > 0.f && mx < 3.0e38f;
          q_inv = live ? 32767.0f / mx : 0.f;
          if (row < M && n_blk < ep.ld_scale)
            ep.scale[static_cast<size_t>(row) * ep.ld_scale + n_blk] = live ? mx * (1.0f / 32767.0f) : mx * 0.0f;
        }
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + static_cast<uint32_t>(c * 32), r);
          uint8_t* buf = my_stage + (kHiLo ? 0 : ((c >> 1) & 1)) * gemm::EPI_STAGE_BYTES;
          if ((c & 1) == 0) {
            // the store issued two boxes ago has left `buf` (hi / lo: the previous box's two stores have left both)
            if (lane == 0) {
              if constexpr (kHiLo) bulk_wait_group_read<0>();
              else bulk_wait_group_read<1>();
            }
            __syncwarp();
          }
          tmem_ld_wait();
          const int col0 = n_blk * BN + c * 32;
          Epi::template transform<FMT>(ep, st, row, col0, r);
          uint8_t* dst = buf + lane * 128;
          const int unit0 = (c & 1) * 4;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 w;
            if constexpr (kQuant) {
              // u = rn(acc * q_inv) + 32768 sits in the low 16 bits of acc * q_inv + (1.5 * 2^23 + 32768)
              uint32_t u[8];
#pragma unroll
              for (int t = 0; t < 8; ++t)
                u[t] = __float_as_uint(fmaf(__uint_as_float(r[8 * j + t]), q_inv, gemm::kQ16Magic));
              w.x = __byte_perm(u[0], u[1], 0x5410);
              w.y = __byte_perm(u[2], u[3], 0x5410);
              w.z = __byte_perm(u[4], u[5], 0x5410);
              w.w = __byte_perm(u[6], u[7], 0x5410);
            } else {
              using Out = Op16<(Epi::kOutFmt < 0 ? FMT : Epi::kOutFmt)>;
              w.x = Out::pack2(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
              w.y = Out::pack2(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
              w.z = Out::pack2(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
              w.w = Out::pack2(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
            }
            *reinterpret_cast<uint4*>(dst + (((unit0 + j) ^ (lane & 7)) << 4)) = w;
            if constexpr (kHiLo) {  // residual half into the second staging buffer, same swizzled position
              using Out = Op16<(Epi::kOutFmt < 0 ? FMT : Epi::kOutFmt)>;
              const float2 h0 = Out::unpack2(w.x), h1 = Out::unpack2(w.y), h2 = Out::unpack2(w.z), h3 = Out::unpack2(w.w);
              uint4 l;
              l.x = Out::pack2(__uint_as_float(r[8 * j + 0]) - h0.x, __uint_as_float(r[8 * j + 1]) - h0.y);
              l.y = Out::pack2(__uint_as_float(r[8 * j + 2]) - h1.x, __uint_as_float(r[8 * j + 3]) - h1.y);
              l.z = Out::pack2(__uint_as_float(r[8 * j + 4]) - h2.x, __uint_as_float(r[8 * j + 5]) - h2.y);
              l.w = Out::pack2(__uint_as_float(r[8 * j + 6]) - h3.x, __uint_as_float(r[8 * j + 7]) - h3.y);
              *reinterpret_cast<uint4*>(dst + gemm::EPI_STAGE_BYTES + (((unit0 + j) ^ (lane & 7)) << 4)) = l;
            }
          }
          if (c & 1) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (row0 < M && col0 - 32 < N) {
                tma_store_2d(&tmap_c, buf, col0 - 32, row0);
                if constexpr (kHiLo) tma_store_2d(&tmap_c, buf + gemm::EPI_STAGE_BYTES, ep.lo_off + col0 - 32, row0);
              }
              bulk_commit_group();
            }
          }
        }
      } else if constexpr (Epi::kTmaStore) {
        uint8_t* my_stage = epi_stage + quarter * 2 * gemm::EPI_STAGE_BYTES;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + static_cast<uint32_t>(c * 32), r);
          uint8_t* buf = my_stage + (c & 1) * gemm::EPI_STAGE_BYTES;
          if (lane == 0) bulk_wait_group_read<1>();
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
            bulk_commit_group();
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
      if (lane == 0) mbar_arrive_remote(map_to_cta(smem_u32(&tmem_empty_bar[acc]), 0));
      if (++acc == ACC_STAGES) {
        acc = 0;
        acc_phase ^= 1u;
      }
    }
    if (Epi::kTmaStore && lane == 0) bulk_wait_group<0>();
  }

  // nobody leaves while the peer may still read this CTA's operand tiles or signal its barriers
  tc_fence_before_sync();
  cluster_sync_all();
  if (warp_idx == 2) {
    tc_fence_after_sync();
    tmem_dealloc_pair(tmem_base, TMEM_COLS);
  }
}

template <int FMT, class Epi>
inline int launch(const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                  const typename Epi::Params& ep, int num_sms, cudaStream_t stream, const int* m_dev = nullptr,
                  void* out = nullptr, int ldc = 0,  // out: fp32 rows, or 16-bit rows when Epi::kOut16
                  const gemm::SplitK& sk = gemm::make_split(-1, -1)) {
  if (M <= 0) return kOk;
  if (K % 8 != 0 || lda % 8 != 0 || ldb % 8 != 0 || N % 32 != 0) return kErrInvalidArg;
  if (sk.passes > 1 && K % BK != 0) return kErrInvalidArg;
  const int wa = gemm::split_width(sk, sk.a_off, K), wb = gemm::split_width(sk, sk.b_off, K);
  if (wa > lda || wb > ldb) return kErrInvalidArg;
  CUtensorMap ta, tb;
  if (!gemm::make_tmap_2d(&ta, FMT, A, static_cast<uint64_t>(M), static_cast<uint64_t>(wa), static_cast<uint64_t>(lda),
                          BM_CTA))
    return kErrCuda;
  if (!gemm::make_tmap_2d(&tb, FMT, B, static_cast<uint64_t>(N), static_cast<uint64_t>(wb), static_cast<uint64_t>(ldb),
                          BN_CTA))
    return kErrCuda;
  CUtensorMap tc = ta;
  if (Epi::kTmaStore) {
    if (out == nullptr || (reinterpret_cast<uintptr_t>(out) & 15) != 0) return kErrInvalidArg;
    if (Epi::kOut16) {
      if (ldc % 8 != 0 || N % 64 != 0) return kErrInvalidArg;
      uint64_t out_cols = static_cast<uint64_t>(N);
      if constexpr (gemm::epi_is_hilo<Epi>::value) {
        if (ep.lo_off % 64 != 0 || ep.lo_off < N || ep.lo_off + N > ldc) return kErrInvalidArg;
        out_cols = static_cast<uint64_t>(ep.lo_off) + static_cast<uint64_t>(N);
      }
      if (!gemm::make_tmap_out_16(&tc, Epi::kOutFmt < 0 ? FMT : Epi::kOutFmt, out, static_cast<uint64_t>(M), out_cols,
                                  static_cast<uint64_t>(ldc)))
        return kErrCuda;
    } else {
      if (ldc % 4 != 0) return kErrInvalidArg;
      if (!gemm::make_tmap_out_f32(&tc, out, static_cast<uint64_t>(M), static_cast<uint64_t>(N),
                                   static_cast<uint64_t>(ldc)))
        return kErrCuda;
    }
  }
  auto kern = gemm2_kernel<FMT, Epi>;
  static std::atomic<unsigned long long> smem_ready{0};  // per template instantiation
  if (!ensure_dynamic_smem(kern, SMEM_BYTES, smem_ready)) return kErrCuda;
  const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int pairs_max = num_sms / 2;
  const int pairs = tiles < pairs_max ? tiles : pairs_max;
  kern<<<2 * pairs, NUM_THREADS, SMEM_BYTES, stream>>>(ta, tb, tc, M, N, K, m_dev, ep, sk);
  return cudaGetLastError() == cudaSuccess ? kOk : kErrCuda;
}

}  // namespace gemm2
}  // namespace evi
