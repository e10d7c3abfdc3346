// Timeline probe of the tcgen05 prefill dequant-GEMM (csrc/gemm_int4_tc.cu compiled with AO_GM_TRACE): per-stage clock
// stamps of the producer, one dequant warp of each set, the MMA warp and one epilogue warp of CTA (0, 0).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I include -o benchmarks/gemm_trace benchmarks/gemm_trace.cu
#ifndef NO_TRACE
#define AO_GM_TRACE 1
#endif
#include "../include/ao_b200.h"
#include "../torchao_fork_b200/csrc/gemm_int4_tc.cu"

#include <cstdio>

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atoll(argv[1]) : 4096, K = argc > 2 ? atoll(argv[2]) : 4096, M = argc > 3 ? atoll(argv[3]) : 4096;
  const int G = 128;
  const int64_t Np = (N + 127) / 128 * 128, Kp = K;
  uint8_t* w;
  uint32_t* sz;
  __nv_bfloat16 *x, *y;
  const size_t wbytes = (size_t)Np * Kp / 2, szbytes = (size_t)Np * (Kp / G) * 4;
  cudaMalloc(&w, wbytes);
  cudaMalloc(&sz, szbytes);
  cudaMalloc(&x, M * K * 2);
  cudaMalloc(&y, M * Np * 2);
  cudaMemset(w, 0x5a, wbytes);
  cudaMemset(sz, 0x3c, szbytes);
  cudaMemset(x, 0x3c, M * K * 2);
  void* ws;
  const int64_t wsb = ao::int4_gemm_tc_workspace_bytes(M, Np, Kp);
  cudaMalloc(&ws, wsb);
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventCreate(&e2);
  int rc = 0;
  for (int it = 0; it < 3; ++it) rc |= ao::int4_gemm_tc(x, M, K, K, w, sz, Np, Kp, G, y, N, Np, ws, wsb, 0);
  cudaDeviceSynchronize();
  const int iters = 10;
  cudaEventRecord(e0);
  for (int it = 0; it < iters; ++it) rc |= ao::int4_gemm_tc(x, M, K, K, w, sz, Np, Kp, G, y, N, Np, ws, wsb, 0);
  cudaEventRecord(e1);
  const int NT = ao::gm_tile_rows(M);
  const int64_t Mpad = (M + NT - 1) / NT * NT, items = Mpad * (Kp / 64);
  for (int it = 0; it < iters; ++it)
    ao::x_permute_kernel<<<(unsigned)((items + 255) / 256), 256>>>(x, M, K, K, Mpad, (int)(Kp / 64), NT, (__nv_bfloat16*)ws);
  cudaEventRecord(e2);
  cudaDeviceSynchronize();
  float ms, ms2;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventElapsedTime(&ms2, e1, e2);
  const double us = ms * 1e3 / iters, us_p = ms2 * 1e3 / iters;
  printf("rc=%d err=%s  %.1f us per call (permute alone %.1f us), %.1f TFLOP/s (GEMM alone %.1f)\n", rc,
         cudaGetErrorString(cudaGetLastError()), us, us_p, 2.0 * M * N * K / us * 1e-6, 2.0 * M * N * K / (us - us_p) * 1e-6);
#ifdef AO_GM_TRACE
  static long long tr[4][ao::kGmMmaWarp + 1][ao::kGmTraceIts];
  cudaMemcpyFromSymbol(tr, ao::ao_gm_trace, sizeof tr);
  const int PW = ao::kGmProducerWarp, MW = ao::kGmMmaWarp, D0 = ao::kGmEpiWarps, D1 = ao::kGmEpiWarps + 8;
  const long long t0 = tr[0][PW][0];
  printf("%3s | %8s %8s | %8s %8s %8s %8s | %8s %8s %8s %8s\n", "it", "P.begin", "P.empty", "D.begin", "D.full", "D.deq", "D.stored",
         "M.begin", "M.afull", "M.full", "M.issued");
  for (int it = 0; it < 64 && it < Kp / 64; ++it) {
    const int DW = (it & 1) ? D1 : D0;
    printf("%3d | %8lld %8lld | %8lld %8lld %8lld %8lld | %8lld %8lld %8lld %8lld\n", it, tr[0][PW][it] - t0, tr[1][PW][it] - t0,
           tr[0][DW][it] - t0, tr[1][DW][it] - t0, tr[2][DW][it] - t0, tr[3][DW][it] - t0, tr[0][MW][it] - t0, tr[1][MW][it] - t0,
           tr[2][MW][it] - t0, tr[3][MW][it] - t0);
  }
  printf("epilogue warp 0: begin %lld, d_full %lld, done %lld\n", tr[0][0][0] - t0, tr[1][0][0] - t0, tr[2][0][0] - t0);
#endif
  return 0;
}
