// Timeline probe of the tcgen05 int4 decode GEMV (csrc/gemv_int4_tc.cu compiled with AO_TC_TRACE): per-stage clock
// stamps of the producer, one dequant warp, the MMA thread and one epilogue warp of CTA 0.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I include -o benchmarks/gemv_tc_trace benchmarks/gemv_tc_trace.cu
#ifndef NO_TRACE
#define AO_TC_TRACE 1
#endif
#include "../include/ao_b200.h"
#include "../torchao_fork_b200/csrc/gemv_int4_tc.cu"

#include <cstdio>

int main(int argc, char** argv) {
  const bool pdl_flag = getenv("PDL") != nullptr;
  const int64_t N = argc > 1 ? atoll(argv[1]) : 128256, K = argc > 2 ? atoll(argv[2]) : 4096;
  const int G = 128, M = argc > 3 ? atoi(argv[3]) : 1;
  const int copies = argc > 4 ? atoi(argv[4]) : 4;
  const int64_t Np = (N + 127) / 128 * 128, Kp = K;
  uint8_t* w;
  uint32_t* sz;
  __nv_bfloat16 *x, *y;
  const size_t wbytes = (size_t)Np * Kp / 2, szbytes = (size_t)Np * (Kp / G) * 4;
  cudaMalloc(&w, wbytes * copies);
  cudaMalloc(&sz, szbytes * copies);
  cudaMalloc(&x, M * K * 2);
  cudaMalloc(&y, M * Np * 2);
  cudaMemset(w, 0x5a, wbytes * copies);
  cudaMemset(sz, 0x3c, szbytes * copies);
  cudaMemset(x, 0x3c, M * K * 2);
  void* ws;
  const int64_t wsb = ao::int4_gemv_tc_workspace_bytes(Np);
  cudaMalloc(&ws, wsb);
  cudaMemset(ws, 0, wsb);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int rc = 0;
  for (int it = 0; it < 2 * copies; ++it)
    rc |= ao::int4_gemv_tc(x, M, K, K, w + (it % copies) * wbytes, sz + (it % copies) * (szbytes / 4), Np, Kp, G, y, N, Np, nullptr,
                           0.f, nullptr, 0, 0, ws, wsb, pdl_flag, 0);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int it = 0; it < 3 * copies; ++it)
    rc |= ao::int4_gemv_tc(x, M, K, K, w + (it % copies) * wbytes, sz + (it % copies) * (szbytes / 4), Np, Kp, G, y, N, Np, nullptr,
                           0.f, nullptr, 0, 0, ws, wsb, pdl_flag, 0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double us = ms * 1e3 / (3 * copies);
  printf("rc=%d err=%s  %.2f us per launch, %.0f GB/s\n", rc, cudaGetErrorString(cudaGetLastError()), us,
         (wbytes + szbytes) / us * 1e-3);
#ifdef AO_TC_TRACE
  static long long tr[6][ao::kTcMmaWarp + 1][ao::kTcTraceIts];
  cudaMemcpyFromSymbol(tr, ao::ao_tc_trace, sizeof tr);
  const int PW = ao::kTcProducerWarp, MW = ao::kTcMmaWarp, EW = ao::kTcDequantWarps, DW = 0;
  const long long t0 = tr[5][DW][0];
  printf("entry->after bar2: dequant w0 %lld, mma %lld, producer %lld; staging done (w0) %lld\n", tr[5][DW][1] - t0,
         tr[5][MW][1] - t0, tr[5][PW][1] - t0, tr[5][DW][2] - t0);
  printf("%3s | %8s %8s | %8s %8s %8s %8s | %8s %8s %8s | %8s %8s %8s\n", "it", "P.wait", "P.issued", "D.begin", "D.full",
         "D.aempty", "D.done", "M.begin", "M.afull", "M.commit", "E.begin", "E.full", "E.done");
  for (int it = 0; it < 40; ++it)
    printf("%3d | %8lld %8lld | %8lld %8lld %8lld %8lld | %8lld %8lld %8lld | %8lld %8lld %8lld\n", it, tr[0][PW][it] - t0,
           tr[1][PW][it] - t0, tr[0][DW][it] - t0, tr[3][DW][it] - t0, tr[1][DW][it] - t0, tr[2][DW][it] - t0, tr[0][MW][it] - t0,
           tr[1][MW][it] - t0, tr[2][MW][it] - t0, tr[0][EW][it] - t0, tr[1][EW][it] - t0, tr[2][EW][it] - t0);
  printf("MMA thread, stage 4: clocks before/after each MMA issue (relative to the first)\n");
  for (int j = 0; j < 16; ++j) printf("  j=%2d before %6lld after %6lld\n", j, tr[3][MW][j] - tr[3][MW][0], tr[4][MW][j] - tr[3][MW][0]);
#endif
  return 0;
}
