// Timeline probe of the fused int4 decode GEMV: compiles the product kernel with AO_GEMV_TRACE, runs the Llama-3-8B
// output projection shape (128256 x 4096, G = 128, M = 1) and prints, for one CTA, when the producer issued each
// stage, how long every consumer warp waited for it and how long it worked on it.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -o benchmarks/gemv_trace benchmarks/gemv_trace.cu
#ifndef NO_TRACE
#define AO_GEMV_TRACE 1
#endif
#include "../include/ao_b200.h"
#include "../torchao_fork_b200/csrc/gemv_int4_fused.cu"

#include <cstdio>
#include <vector>

int main(int argc, char** argv) {
  const bool pdl_flag = getenv("PDL") != nullptr;
  const int64_t N = argc > 1 ? atoll(argv[1]) : 128256, K = argc > 2 ? atoll(argv[2]) : 4096;
  const int G = 128, M = argc > 3 ? atoi(argv[3]) : 1;
  const int max_stages = argc > 4 ? atoi(argv[4]) : 16;
  const int64_t Np = (N + 127) / 128 * 128, Kp = K;
  const int copies = argc > 5 ? atoi(argv[5]) : 4;  // rotate over distinct weight buffers (> L2); 1 + small N = L2-resident
  uint8_t* w;
  uint32_t* sz;
  __nv_bfloat16 *x, *y;
  const size_t wbytes = (size_t)Np * Kp / 2, szbytes = (size_t)Np * (Kp / G) * 4;
  cudaMalloc(&w, wbytes * copies);
  cudaMalloc(&sz, szbytes * copies);
  cudaMalloc(&x, M * K * 2);
  cudaMalloc(&y, M * Np * 2);
  cudaMemset(w, 0x5a, wbytes * copies);
  cudaMemset(sz, 0x3c, szbytes * copies);  // bf16 0x3c3c ~ 0.0115
  cudaMemset(x, 0x3c, M * K * 2);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int rc = 0;
  for (int it = 0; it < 3 * copies; ++it)
    rc |= ao::int4_gemv_fused(x, M, K, K, w + (it % copies) * wbytes, sz + (it % copies) * (szbytes / 4), Np, Kp, G, y, N, Np,
                              nullptr, 0.f, nullptr, 0, 0, max_stages, pdl_flag, 0);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int it = 0; it < 5 * copies; ++it)
    rc |= ao::int4_gemv_fused(x, M, K, K, w + (it % copies) * wbytes, sz + (it % copies) * (szbytes / 4), Np, Kp, G, y, N, Np,
                              nullptr, 0.f, nullptr, 0, 0, max_stages, pdl_flag, 0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double us = ms * 1e3 / (5 * copies);
  printf("rc=%d err=%s  %.2f us per launch, %.0f GB/s\n", rc, cudaGetErrorString(cudaGetLastError()), us,
         (wbytes + szbytes) / us * 1e-3);

#ifdef AO_GEMV_TRACE
  static long long tr[ao::kTraceKinds][ao::kFConsumers + 2][ao::kTraceIts];
  cudaMemcpyFromSymbol(tr, ao::ao_trace, sizeof tr);
  const int PW = ao::kFConsumers;  // producer row
  const int n_chunks = (int)((Kp / 64 + ao::kStageTiles - 1) / ao::kStageTiles);
  const int its = 96 < ao::kTraceIts ? 96 : ao::kTraceIts;
  const long long t0 = tr[1][PW][0];
  printf("clock stamps relative to the producer's first issue (SM clocks)\n");
  printf("%4s %9s %9s | %9s %9s %9s | per-warp: wait(full) clocks / work clocks\n", "it", "p.wait", "p.issue", "first.rdy",
         "last.rdy", "last.done");
  double sum_wait = 0, sum_work = 0, sum_pwait = 0;
  int cnt = 0;
  for (int it = 0; it < its; ++it) {
    long long first_rdy = 1LL << 62, last_rdy = 0, last_done = 0;
    for (int w_ = 0; w_ < ao::kFConsumers; ++w_) {
      first_rdy = std::min(first_rdy, tr[1][w_][it]);
      last_rdy = std::max(last_rdy, tr[1][w_][it]);
      last_done = std::max(last_done, tr[2][w_][it]);
    }
    printf("%4d %9lld %9lld | %9lld %9lld %9lld |", it, tr[0][PW][it] - t0, tr[1][PW][it] - t0, first_rdy - t0,
           last_rdy - t0, last_done - t0);
    for (int w_ = 0; w_ < ao::kFConsumers; w_ += 5)
      printf(" w%-2d %5lld/%-5lld", w_, tr[1][w_][it] - tr[0][w_][it], tr[2][w_][it] - tr[1][w_][it]);
    printf("\n");
    if (it >= 16) {
      for (int w_ = 0; w_ < ao::kFConsumers; ++w_) {
        sum_wait += tr[1][w_][it] - tr[0][w_][it];
        sum_work += tr[2][w_][it] - tr[1][w_][it];
      }
      sum_pwait += tr[1][PW][it] - tr[0][PW][it];
      ++cnt;
    }
  }
  printf("steady state (it >= 16): consumer wait(full) %.0f clk, work %.0f clk per warp-stage; producer wait(empty) %.0f clk\n",
         sum_wait / (cnt * ao::kFConsumers), sum_work / (cnt * ao::kFConsumers), sum_pwait / cnt);
  printf("stage period: %.0f clk\n", (double)(tr[1][PW][its - 1] - tr[1][PW][16]) / (its - 1 - 16));
  printf("prologue phases of warp 0 / warp 7 (clocks after kernel entry): after-init-sync, after-griddep-wait, x loaded+bar, row scale+bar, rescale+bar, X sums+bar, first stage ready\n");
  for (int w_ : {0, 7, 15})
    printf("  w%-2d: %lld %lld %lld %lld %lld %lld | first rdy %lld, producer first issue %lld\n", w_, tr[5][w_][1] - tr[5][w_][0],
           tr[5][w_][2] - tr[5][w_][0], tr[5][w_][3] - tr[5][w_][0], tr[5][w_][4] - tr[5][w_][0], tr[5][w_][5] - tr[5][w_][0],
           tr[5][w_][6] - tr[5][w_][0], tr[1][w_][0] - tr[5][w_][0], t0 - tr[5][w_][0]);
  printf("per-warp ready / done stamps (relative), its 40..47\n");
  for (int it = 40; it < 48; ++it) {
    printf("it %d rdy :", it);
    for (int w_ = 0; w_ < ao::kFConsumers; ++w_) printf(" %6lld", tr[1][w_][it] - t0);
    printf("\nit %d done:", it);
    for (int w_ = 0; w_ < ao::kFConsumers; ++w_) printf(" %6lld", tr[2][w_][it] - t0);
    printf("\n");
    if (it & 1) {
      printf("strip %d red_free-passed:", it / 2);
      for (int w_ = 0; w_ < ao::kFConsumers; ++w_) printf(" %6lld", tr[3][w_][it / 2] - t0);
      printf("\n");
    }
  }
  printf("strip hand-off (strip li): red_free wait end (w0) / reducer start / reducer end, relative\n");
  for (int li = 2; li < 12; ++li) {
    const int rw = li % ao::kFConsumers;
    printf("  li %2d: stage-done(w%d) %9lld  reducer start %9lld  end %9lld\n", li, rw,
           tr[2][rw][li * n_chunks + n_chunks - 1] - t0, tr[4][PW + 1][li] - t0, tr[5][PW + 1][li] - t0);
  }
#endif
  return 0;
}
