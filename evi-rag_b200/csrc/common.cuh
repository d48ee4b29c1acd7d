// Shared device helpers for the EVI-RAG retrieval kernels (sm_100a only).
// Thin inline-PTX wrappers for mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (MMA / TMEM) plus
// the small math helpers every kernel shares. No library dependencies.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <atomic>

#define EVI_DEVINL __device__ __forceinline__

namespace evi {

// ---------------------------------------------------------------------------------------------
// status codes returned through the C ABI (include/evi_rag_b200.h mirrors these)
// ---------------------------------------------------------------------------------------------
enum Status : int {
  kOk = 0,
  kErrInvalidArg = 1,
  kErrCuda = 2,
  kErrWorkspace = 3,
  kErrUnsupported = 4,
  kErrStaticLimit = 5,  // a size beyond the fixed scratch of a *_static entry point
};

// bits of the device-side validation word written by evi_edge_batch (graph_utils.py:58-99)
enum EdgeFlag : unsigned {
  kFlagOutOfRange = 1u,      // "edge_batch contains out-of-range indices" / edge_index out of range
  kFlagCrossGraph = 2u,      // "edge_index crosses graph boundaries"
  kFlagNotMonotone = 4u,     // "edge_batch is not non-decreasing"
};

// ---------------------------------------------------------------------------------------------
// generic helpers
// ---------------------------------------------------------------------------------------------
// Opt a kernel into `bytes` of dynamic shared memory once per device: the attribute is per device (a process that
// drives several GPUs needs it on each) and `done` is one bit per device ordinal, safe under concurrent callers.
template <class Kernel>
inline bool ensure_dynamic_smem(Kernel kern, int bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  const unsigned long long bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return true;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return false;
  done.fetch_or(bit, std::memory_order_release);
  return true;
}

EVI_DEVINL uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
EVI_DEVINL uint32_t lane_id() { return threadIdx.x & 31u; }

EVI_DEVINL bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

EVI_DEVINL float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// erf GELU (nn.GELU() default, retriever.py:163,178), branch-free:
//   gelu(x) = max(x, 0) - |x| * (0.5 erfc(|x| / sqrt 2)),   0.5 erfc(a / sqrt 2) = 2^-q(a)
// q is a degree-7 weighted least-squares fit of 1 - log2(erfc(a / sqrt 2)) at Chebyshev nodes of [0, 5.94]
// (scripts/fit_gelu.py). Beyond the fitted range q keeps growing (q(5.94) = 29.4, leading coefficient > 0),
// so 2^-q underflows towards 0 on its own and no clamp is needed.
// max |gelu_fast - gelu_erf| = 4.8e-7 over [-12, 12] in fp32; 7 FMA + MUFU.EX2 + FMNMX + FMA.
EVI_DEVINL float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
EVI_DEVINL float gelu_fast(float x) {
  const float ax = fabsf(x);
  float q = 1.8662889260667725e-06f;
  q = fmaf(q, ax, -6.2319784774445e-05f);
  q = fmaf(q, ax, 0.0009371304186061025f);
  q = fmaf(q, ax, -0.008538476191461086f);
  q = fmaf(q, ax, 0.05403471365571022f);
  q = fmaf(q, ax, 0.45837363600730896f);
  q = fmaf(q, ax, 1.1512972116470337f);
  q = fmaf(q, ax, 0.9999892115592957f);
  return fmaf(-ax, ex2_approx(-q), fmaxf(x, 0.f));
}
// Degree-5 fit of the same q (scripts/fit_gelu.py with DEG = 5): max |error| = 2.2e-5, two FMAs fewer. For values
// that are rounded to 16 bits right away (the structure context rows sc, |sc| ~ 0.4: half an fp16 ulp is 1e-4);
// q(5.94) = 29.4 and the leading coefficient is positive, so no clamp here either.
EVI_DEVINL float gelu_fast5(float x) {
  const float ax = fabsf(x);
  float q = 0.00023532388149760664f;
  q = fmaf(q, ax, -0.00492762541398406f);
  q = fmaf(q, ax, 0.04495760053396225f);
  q = fmaf(q, ax, 0.4690363109111786f);
  q = fmaf(q, ax, 1.1463907957077026f);
  q = fmaf(q, ax, 1.0005117654800415f);
  return fmaf(-ax, ex2_approx(-q), fmaxf(x, 0.f));
}
// tanh(x) = sign(x) (1 - e) / (1 + e), e = exp(-2|x|): 2 MUFU + 5 ALU, relative error ~1e-6 away from 0 and
// absolute error < 1e-7 near 0 (used where the result is rounded to 16 bits anyway)
EVI_DEVINL float tanh_fast(float x) {
  const float e = ex2_approx(-2.8853900817779268f * fabsf(x));
  return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}
// Packed fp32 pairs (sm_100 FFMA2 / FADD2 / FMUL2: two IEEE fp32 operations per instruction, each half bit-identical
// to the scalar form; a scalar or immediate operand is broadcast by the instruction, |x| and -x are operand modifiers).
// An FFMA2 holds the issue port for two cycles (scripts/probes/ffma2_probe.cu, profiles/r01y_gemm_vs_cublas.md), so it
// buys no arithmetic throughput; what it buys is a 28 % shorter instruction stream. Measured A/B on one box
// (profiles/r04a_f32x2_ab.md): score_tail_bulk_kernel -7 % (kept: the tail keeps a lane's columns as (even, odd)
// pairs), edge_operand_kernel +17 % (not kept: that kernel stays scalar).
// -DEVI_NO_F32X2 expands the same helpers to scalar operations.
EVI_DEVINL float2 bc2(float a) { return make_float2(a, a); }
EVI_DEVINL float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
EVI_DEVINL float2 abs2(float2 a) { return make_float2(fabsf(a.x), fabsf(a.y)); }
#ifndef EVI_NO_F32X2
EVI_DEVINL float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
EVI_DEVINL float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
EVI_DEVINL float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
#else
EVI_DEVINL float2 fma2(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
EVI_DEVINL float2 add2(float2 a, float2 b) { return make_float2(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y)); }
EVI_DEVINL float2 mul2(float2 a, float2 b) { return make_float2(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)); }
#endif
EVI_DEVINL float2 sub2(float2 a, float2 b) { return add2(a, neg2(b)); }
EVI_DEVINL float hsum2(float2 a) { return a.x + a.y; }
// gelu_fast / gelu_fast5 on a pair: 8 (6) FFMA2 + 2 MUFU.EX2 + 2 FMNMX for two values
EVI_DEVINL float2 gelu_fast(float2 x) {
  const float2 ax = abs2(x);
  float2 q = bc2(1.8662889260667725e-06f);
  q = fma2(q, ax, bc2(-6.2319784774445e-05f));
  q = fma2(q, ax, bc2(0.0009371304186061025f));
  q = fma2(q, ax, bc2(-0.008538476191461086f));
  q = fma2(q, ax, bc2(0.05403471365571022f));
  q = fma2(q, ax, bc2(0.45837363600730896f));
  q = fma2(q, ax, bc2(1.1512972116470337f));
  q = fma2(q, ax, bc2(0.9999892115592957f));
  return fma2(neg2(ax), make_float2(ex2_approx(-q.x), ex2_approx(-q.y)), make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)));
}
EVI_DEVINL float2 gelu_fast5(float2 x) {
  const float2 ax = abs2(x);
  float2 q = bc2(0.00023532388149760664f);
  q = fma2(q, ax, bc2(-0.00492762541398406f));
  q = fma2(q, ax, bc2(0.04495760053396225f));
  q = fma2(q, ax, bc2(0.4690363109111786f));
  q = fma2(q, ax, bc2(1.1463907957077026f));
  q = fma2(q, ax, bc2(1.0005117654800415f));
  return fma2(neg2(ax), make_float2(ex2_approx(-q.x), ex2_approx(-q.y)), make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)));
}
EVI_DEVINL float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
EVI_DEVINL float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

// 16-bit operand formats for the tensor-core path. Both run tcgen05 kind::f16 at the same rate.
enum OperandFormat : int { kFmtF16 = 0, kFmtBF16 = 1 };

template <int FMT> struct Op16;
template <> struct Op16<kFmtF16> {
  using T = __half;
  using T2 = __half2;
  static EVI_DEVINL T from_float(float v) { return __float2half_rn(v); }
  static EVI_DEVINL float to_float(T v) { return __half2float(v); }
  static EVI_DEVINL uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static EVI_DEVINL uint32_t pack2(float2 v) { return pack2(v.x, v.y); }
  static EVI_DEVINL float2 unpack2(uint32_t u) {
    return __half22float2(*reinterpret_cast<__half2*>(&u));
  }
};
template <> struct Op16<kFmtBF16> {
  using T = __nv_bfloat16;
  using T2 = __nv_bfloat162;
  static EVI_DEVINL T from_float(float v) { return __float2bfloat16_rn(v); }
  static EVI_DEVINL float to_float(T v) { return __bfloat162float(v); }
  static EVI_DEVINL uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  static EVI_DEVINL uint32_t pack2(float2 v) { return pack2(v.x, v.y); }
  static EVI_DEVINL float2 unpack2(uint32_t u) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
  }
};

// 128-bit read-only streaming load (embedding rows, edge lists)
EVI_DEVINL uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// 256-bit loads (sm_100: LDG.E.256): one lane reads 8 consecutive floats, a warp 1 KB contiguous.
// `_stream`: read-once rows (no L1 allocation); `_keep`: rows re-read by neighbouring warps.
EVI_DEVINL void ldg_v8_stream(const float* p, float (&v)[8]) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p));
}
EVI_DEVINL void ldg_v8_keep(const float* p, float (&v)[8]) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "l"(p));
}

// the same loads into four (even, odd) column pairs
EVI_DEVINL void ldg_v8_stream(const float* p, float2 (&v)[4]) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0].x), "=f"(v[0].y), "=f"(v[1].x), "=f"(v[1].y), "=f"(v[2].x), "=f"(v[2].y), "=f"(v[3].x),
                 "=f"(v[3].y)
               : "l"(p));
}
EVI_DEVINL void ldg_v8_keep(const float* p, float2 (&v)[4]) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0].x), "=f"(v[0].y), "=f"(v[1].x), "=f"(v[1].y), "=f"(v[2].x), "=f"(v[2].y), "=f"(v[3].x),
                 "=f"(v[3].y)
               : "l"(p));
}
// ---------------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------------
EVI_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
EVI_DEVINL void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
EVI_DEVINL void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
EVI_DEVINL void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
EVI_DEVINL void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
EVI_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// TMA: 2-D tiled bulk tensor load, completion on an mbarrier
// ---------------------------------------------------------------------------------------------
EVI_DEVINL void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
EVI_DEVINL void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar, int c_inner,
                            int c_outer) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer)
      : "memory");
}

// 1-D bulk copy global -> shared (the TMA engine moves `bytes`, a multiple of 16, both addresses 16-byte
// aligned); completion is signalled on an mbarrier through its transaction count
EVI_DEVINL void bulk_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
EVI_DEVINL void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// 2-D tiled bulk tensor store shared -> global (bulk async-group completion). The box is clipped to
// the tensor map's bounds, so partial tiles need no masking.
EVI_DEVINL void tma_store_2d(const void* tmap, const void* smem_src, int c_inner, int c_outer) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tmap),
               "r"(smem_u32(smem_src)), "r"(c_inner), "r"(c_outer)
               : "memory");
}
EVI_DEVINL void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
EVI_DEVINL void bulk_wait_group_read() {  // the sources of all but the N newest groups have been read
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
EVI_DEVINL void bulk_wait_group() {  // all but the N newest groups are complete (writes performed)
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05: TMEM allocation, MMA issue, commit, TMEM loads
// ---------------------------------------------------------------------------------------------
EVI_DEVINL void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
EVI_DEVINL void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
EVI_DEVINL void tc_fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
EVI_DEVINL void tc_fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T ; one thread issues for the CTA
EVI_DEVINL void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once all previously issued MMAs of this thread have completed
EVI_DEVINL void umma_commit(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
          smem_u32(bar))
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp reads lane (warp%4)*32+t
EVI_DEVINL void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
EVI_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile in the canonical 128-byte-swizzle layout (what TMA SWIZZLE_128B writes):
// rows of 128 bytes, 8-row groups 1024 bytes apart. Field layout: cute/arch/mma_sm100_desc.hpp.
EVI_DEVINL uint64_t umma_smem_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);        // start address  [0,14)
  d |= static_cast<uint64_t>(1) << 16;                           // LBO (ignored for SW128 K-major)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;                   // SBO = 1024 B   [32,46)
  d |= static_cast<uint64_t>(1) << 46;                           // descriptor version (sm_100)
  d |= static_cast<uint64_t>(2) << 61;                           // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: fp32 accumulate, A/B K-major, 16-bit format `fmt`
__host__ __device__ constexpr uint32_t umma_idesc_f16(int fmt, int m, int n) {
  return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

}  // namespace evi
