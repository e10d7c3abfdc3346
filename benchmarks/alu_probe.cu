// Micro-probe: reciprocal throughput (clocks per warp instruction per SMSP) of the integer ops the int4 dequant is
// built from, alone and mixed, on B200.  16 warps per SM (4 per SMSP), 8 independent chains per thread.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o alu_probe alu_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#define OP_LOP3(r) asm volatile("lop3.b32 %0, %0, 0x000f000f, 0x64006400, 0xea;" : "+r"(r))
#define OP_SHF(r) asm volatile("shr.u32 %0, %0, 8;" : "+r"(r))
#define OP_MULHI(r) asm volatile("mul.hi.u32 %0, %0, 0x01000000;" : "+r"(r))
#define OP_IMAD(r) asm volatile("mad.lo.u32 %0, %0, 0x10001, %0;" : "+r"(r))
#define OP_PRMT(r) asm volatile("prmt.b32 %0, %0, 0x64646464, 0x4140;" : "+r"(r))
#define OP_HADD2(r) asm volatile("add.f16x2 %0, %0, %0;" : "+r"(r))
#define OP_FFMA(r) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(*(float*)&r))

template <int KIND>
__global__ void __launch_bounds__(512) k(int iters, uint32_t* out) {
  uint32_t r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r[j] = threadIdx.x * 2654435761u + j;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (KIND == 0) OP_LOP3(r[j]);
      if (KIND == 1) OP_SHF(r[j]);
      if (KIND == 2) OP_MULHI(r[j]);
      if (KIND == 3) OP_IMAD(r[j]);
      if (KIND == 4) OP_PRMT(r[j]);
      if (KIND == 5) OP_HADD2(r[j]);
      if (KIND == 6) OP_FFMA(r[j]);
      if (KIND == 7) { if (j & 1) OP_MULHI(r[j]); else OP_LOP3(r[j]); }          // 4 LOP3 + 4 IMAD.HI
      if (KIND == 8) { if (j % 5 == 4) OP_MULHI(r[j]); else OP_LOP3(r[j]); }      // 7 LOP3 + 1 IMAD.HI (j = 4)
      if (KIND == 9) { if (j % 5 == 4) OP_SHF(r[j]); else OP_LOP3(r[j]); }        // 7 LOP3 + 1 SHF
      if (KIND == 10) { if (j & 1) OP_HADD2(r[j]); else OP_LOP3(r[j]); }          // 4 LOP3 + 4 HADD2
      if (KIND == 11) { if (j & 1) OP_FFMA(r[j]); else OP_LOP3(r[j]); }           // 4 LOP3 + 4 FFMA
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s ^= r[j];
  if (s == 0x12345) out[0] = s;
}

int main() {
  uint32_t* out;
  cudaMalloc(&out, 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000;
  const char* names[12] = {"LOP3", "SHF", "IMAD.HI", "IMAD", "PRMT", "HADD2", "FFMA", "4 LOP3 + 4 IMAD.HI",
                           "7 LOP3 + 1 IMAD.HI", "7 LOP3 + 1 SHF", "4 LOP3 + 4 HADD2", "4 LOP3 + 4 FFMA"};
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  for (int kind = 0; kind < 12; ++kind) {
    auto launch = [&] {
      switch (kind) {
#define C(K) case K: k<K><<<148, 512>>>(iters, out); break;
        C(0) C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8) C(9) C(10) C(11)
#undef C
      }
    };
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr_per_smsp = 4.0 * iters * 8;  // 4 warps per SMSP
    printf("%-22s %6.2f clk per warp instruction per SMSP (at %d MHz nominal; %s)\n", names[kind],
           ms * 1e-3 * clk_khz * 1e3 / warp_instr_per_smsp, clk_khz / 1000, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
