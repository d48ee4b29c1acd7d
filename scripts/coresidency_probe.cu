// Probe: which resource stops a small kernel from being placed beside a 1-CTA/SM big-shared-memory kernel?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 scripts/coresidency_probe.cu -o scripts/bin/coresidency_probe
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ void spin(long long cycles) {
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {}
}
__global__ void big_kernel(long long cycles, int use_tmem, int* sink) {
  extern __shared__ uint8_t smem[];
  __shared__ uint32_t slot;
  uint32_t taddr = 0;
  if (use_tmem && threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  taddr = slot;
  smem[threadIdx.x] = (uint8_t)threadIdx.x;
  spin(cycles);
  if (sink && smem[(threadIdx.x + 1) % blockDim.x] == 255 && cycles < 0) sink[0] = 1;
  __syncthreads();
  if (use_tmem && threadIdx.x < 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(512) : "memory");
}
__global__ void small_kernel(long long cycles, int* sink) {
  spin(cycles);
  if (sink && cycles < 0) sink[0] = 2;
}
static float run(int mode, int big_smem, int use_tmem, int small_carveout, cudaStream_t sa, cudaStream_t sb,
                 long long cyc) {
  cudaEvent_t e0, e1, ea, eb;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&ea); cudaEventCreate(&eb);
  cudaFuncSetAttribute(big_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, big_smem);
  cudaFuncSetAttribute(small_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, small_carveout);
  cudaDeviceSynchronize();
  cudaEventRecord(e0, 0);
  cudaStreamWaitEvent(sa, e0, 0); cudaStreamWaitEvent(sb, e0, 0);
  if (mode & 1) big_kernel<<<148, 256, big_smem, sa>>>(cyc, use_tmem, nullptr);
  if (mode & 2) small_kernel<<<148 * 4, 256, 0, sb>>>(cyc / 4, nullptr);
  cudaEventRecord(ea, sa); cudaEventRecord(eb, sb);
  cudaStreamWaitEvent(0, ea, 0); cudaStreamWaitEvent(0, eb, 0);
  cudaEventRecord(e1, 0);
  cudaDeviceSynchronize();
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}
int main() {
  cudaStream_t sa, sb;
  int lo, hi; cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStreamCreateWithPriority(&sa, cudaStreamNonBlocking, hi);
  cudaStreamCreateWithPriority(&sb, cudaStreamNonBlocking, lo);
  const long long cyc = 2000000;  // ~1 ms
  run(3, 100 * 1024, 0, -1, sa, sb, cyc);
  printf("big alone %.3f ms, small alone %.3f ms (4 CTAs/SM of quarter length: same)\n",
         run(1, 197 * 1024, 0, -1, sa, sb, cyc), run(2, 0, 0, -1, sa, sb, cyc));
  for (int smem : {48 * 1024, 100 * 1024, 160 * 1024, 197 * 1024})
    for (int tm : {0, 1})
      for (int carve : {-1, 100})
        printf("big smem %3d KB tmem %d | small carveout %3d : both %.3f ms\n", smem / 1024, tm, carve,
               run(3, smem, tm, carve, sa, sb, cyc));
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
