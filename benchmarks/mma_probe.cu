// Micro-probe: issue rate of the LEGACY tensor path (mma.sync m16n8k16 bf16/f16, m16n8k32 s8) on B200, per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_probe mma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
template <int KIND>
__global__ void __launch_bounds__(512) k(int iters, float* out) {
  float d[4][4] = {};
  int di[4][4] = {};
  uint32_t a = threadIdx.x * 0x01010101u, b = 0x3c003c00u;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (KIND == 0)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%5}, {%0,%1,%2,%3};"
                     : "+f"(d[j][0]), "+f"(d[j][1]), "+f"(d[j][2]), "+f"(d[j][3]) : "r"(a), "r"(b));
      else if (KIND == 1)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%5}, {%0,%1,%2,%3};"
                     : "+f"(d[j][0]), "+f"(d[j][1]), "+f"(d[j][2]), "+f"(d[j][3]) : "r"(a), "r"(b));
      else
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%5}, {%0,%1,%2,%3};"
                     : "+r"(di[j][0]), "+r"(di[j][1]), "+r"(di[j][2]), "+r"(di[j][3]) : "r"(a), "r"(b));
    }
  }
  float s = 0;
  for (int j = 0; j < 4; ++j) for (int e = 0; e < 4; ++e) s += d[j][e] + di[j][e];
  if (s == 123.f) out[0] = s;
}
int main() {
  float* out; cudaMalloc(&out, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  const char* names[3] = {"mma.sync m16n8k16 bf16", "mma.sync m16n8k16 f16", "mma.sync m16n8k32 s8"};
  for (int kind = 0; kind < 3; ++kind) for (int warps : {4, 8, 16}) {
    auto launch = [&] { if (kind == 0) k<0><<<148, warps * 32>>>(iters, out); else if (kind == 1) k<1><<<148, warps * 32>>>(iters, out); else k<2><<<148, warps * 32>>>(iters, out); };
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double n = 148.0 * warps * iters * 4;
    printf("%-26s warps/SM=%2d: %7.2f G mma/s = %5.2f clk per mma per SMSP @1.9GHz  (%s)\n", names[kind], warps,
           n / (ms * 1e-3) / 1e9, 1.9e9 * 592 / (n / (ms * 1e-3)), cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
