// Floor of a programmatic-dependent-launch chain on B200: each kernel lets its successor launch at once, waits for its
// predecessor, does one dependent global load -> store (grid of `ctas` CTAs x 512 threads), exits.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o benchmarks/pdl_floor benchmarks/pdl_floor.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512) link(const float* in, float* out, int work) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  float v = in[threadIdx.x];
  for (int i = 0; i < work; ++i) v = v * 1.0001f + 0.5f;
  if (blockIdx.x == 0) out[threadIdx.x] = v;
}
int main() {
  float *a, *b;
  cudaMalloc(&a, 4096);
  cudaMalloc(&b, 4096);
  cudaMemset(a, 0, 4096);
  cudaStream_t s;
  cudaStreamCreate(&s);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int pdl = 0; pdl < 2; ++pdl)
    for (int ctas : {1, 148, 296}) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(ctas);
      cfg.blockDim = dim3(512);
      cfg.stream = s;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = pdl;
      const int n = 64;
      cudaGraph_t g;
      cudaGraphExec_t ge;
      cudaStreamBeginCapture(s, cudaStreamCaptureModeGlobal);
      for (int i = 0; i < n; ++i) cudaLaunchKernelEx(&cfg, link, (const float*)((i & 1) ? b : a), (i & 1) ? a : b, 0);
      cudaStreamEndCapture(s, &g);
      cudaGraphInstantiate(&ge, g, 0);
      for (int i = 0; i < 3; ++i) cudaGraphLaunch(ge, s);
      cudaStreamSynchronize(s);
      cudaEventRecord(e0, s);
      for (int i = 0; i < 10; ++i) cudaGraphLaunch(ge, s);
      cudaEventRecord(e1, s);
      cudaStreamSynchronize(s);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("pdl=%d ctas=%3d: %.2f us per dependent kernel (%s)\n", pdl, ctas, ms * 1e3 / (10 * n), cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
