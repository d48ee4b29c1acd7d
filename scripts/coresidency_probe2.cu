// Probe 2: is a register-heavy spinner placed beside the real tcgen05 GEMM (evi_project)?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I include scripts/coresidency_probe2.cu -o scripts/bin/coresidency_probe2 -L evi-rag_b200 -levi_rag_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../evi-rag_b200'
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include "evi_rag_b200.h"

template <int NREG>
__global__ void __launch_bounds__(128) spinner(long long cycles, float* sink) {
  float r[NREG];
#pragma unroll
  for (int i = 0; i < NREG; ++i) r[i] = threadIdx.x * 0.5f + i;
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {
#pragma unroll
    for (int i = 0; i < NREG; ++i) r[i] = r[i] * 1.0001f + 0.5f;
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < NREG; ++i) s += r[i];
  if (s == 12345.678f) sink[0] = s;
}
int main() {
  const int M = 400000;
  void *x, *w, *out; float* bias; float* sink;
  cudaMalloc(&x, (size_t)M * 1024 * 2); cudaMalloc(&w, 1024 * 1024 * 2); cudaMalloc(&out, (size_t)M * 1024 * 2);
  cudaMalloc(&bias, 4096); cudaMalloc(&sink, 4);
  cudaMemset(x, 0, (size_t)M * 1024 * 2); cudaMemset(w, 0, 1024 * 1024 * 2); cudaMemset(bias, 0, 4096);
  int acts[8] = {0};
  cudaStream_t sa, sb; int lo, hi; cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStreamCreateWithPriority(&sa, cudaStreamNonBlocking, hi);
  cudaStreamCreateWithPriority(&sb, cudaStreamNonBlocking, lo);
  cudaEvent_t e0, e1, ea, eb; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&ea); cudaEventCreate(&eb);
  auto run = [&](int mode, int which) {
    cudaDeviceSynchronize();
    cudaEventRecord(e0, 0);
    cudaStreamWaitEvent(sa, e0, 0); cudaStreamWaitEvent(sb, e0, 0);
    if (mode & 1) evi_project(x, M, w, 1024, 1024, bias, acts, out, nullptr, EVI_FMT_F16, 148, sa);
    if (mode & 2) {
      if (which == 0) spinner<8><<<148, 128, 0, sb>>>(1200000, sink);
      else spinner<100><<<148, 128, 0, sb>>>(1200000, sink);
    }
    cudaEventRecord(ea, sa); cudaEventRecord(eb, sb);
    cudaStreamWaitEvent(0, ea, 0); cudaStreamWaitEvent(0, eb, 0);
    cudaEventRecord(e1, 0);
    cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  };
  run(3, 0); run(3, 1);
  printf("gemm alone %.3f | spinner8 alone %.3f | spinner100 alone %.3f\n", run(1, 0), run(2, 0), run(2, 1));
  printf("gemm + spinner8  : %.3f ms\n", run(3, 0));
  printf("gemm + spinner100: %.3f ms\n", run(3, 1));
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
