// FFMA vs FFMA2 (fma.rn.f32x2) issue / pipe throughput on sm_100a: 8 independent chains per thread, 1024 threads per
// CTA, one CTA per SM-slot. Prints FMA/clk/SM for both forms. Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long pk(float a, float b) {
  unsigned long long r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk(unsigned long long v, float& a, float& b) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
constexpr int kIters = 4096;
__global__ void scalar_kernel(float* out, float s) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 0.5f);
  }
  float r = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
__global__ void packed_kernel(float* out, float s) {
  unsigned long long a[8];
  const unsigned long long ss = pk(s, s), h = pk(0.5f, 0.5f);
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = pk(threadIdx.x * 1e-3f + 2 * i, threadIdx.x * 1e-3f + 2 * i + 1);
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = fma2(a[i], ss, h);
  }
  float r = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float x, y;
    upk(a[i], x, y);
    r += x + y;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
// mixed: 1 packed FMA stream + independent integer work competing for issue slots
__global__ void mixed_kernel(float* out, float s, int packed) {
  float a[16];
  unsigned u[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
#pragma unroll
  for (int i = 0; i < 8; ++i) u[i] = threadIdx.x + i;
  for (int it = 0; it < kIters; ++it) {
    if (packed) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        unsigned long long v = fma2(pk(a[2 * i], a[2 * i + 1]), pk(s, s), pk(0.5f, 0.5f));
        upk(v, a[2 * i], a[2 * i + 1]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, 0.5f);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = (u[i] ^ (u[i] >> 3)) + 0x9e3779b9u;  // 3 ALU ops each
  }
  float r = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += a[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) r += u[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
  int sms = 0, khz = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  float* out;
  cudaMalloc(&out, sizeof(float) * sms * 2 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  auto run = [&](const char* name, auto launch, double fma_per_thread) {
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fma = fma_per_thread * 1024.0 * 2 * sms * 5;
    printf("%-28s %8.3f ms  %7.1f GFMA/s  %6.1f FMA/clk/SM at nominal %d MHz\n", name, ms / 5, fma / (ms * 1e6),
           fma / (ms * 1e-3) / sms / (khz * 1e3), khz / 1000);
  };
  run("FFMA  (16 chains)", [&] { scalar_kernel<<<2 * sms, 1024>>>(out, 0.999f); }, 16.0 * kIters);
  run("FFMA2 (8 packed chains)", [&] { packed_kernel<<<2 * sms, 1024>>>(out, 0.999f); }, 16.0 * kIters);
  run("FFMA  + 24 ALU ops", [&] { mixed_kernel<<<2 * sms, 1024>>>(out, 0.999f, 0); }, 16.0 * kIters);
  run("FFMA2 + 24 ALU ops", [&] { mixed_kernel<<<2 * sms, 1024>>>(out, 0.999f, 1); }, 16.0 * kIters);
  printf("cuda status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
