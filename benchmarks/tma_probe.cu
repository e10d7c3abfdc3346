// Micro-probe: how fast can ONE CTA per SM stream HBM into shared memory on B200, by mechanism?
//   mode 0: cp.async.bulk (1-D TMA bulk copy), `nprod` producer threads each owning a ring of `stages` x `chunk` bytes
//   mode 1: cp.async.cg 16 B (LDGSTS) issued by `nprod` producer warps, completion by cp.async.mbarrier.arrive.noinc
//   mode 2: plain ld.global.nc.v4 by all threads (register path), xor-reduced so the loads are not dead
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_probe tma_probe.cu ; run: ./tma_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}"
                 : "=r"(ok) : "r"(s32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk(void* d, const void* s, uint32_t n, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(d)),
               "l"(s), "r"(n), "r"(s32(b)) : "memory");
}

// mode 0 / 1.  blockDim = 32 * (nprod + 1): warps 0..nprod-1 produce, warp nprod consumes (wait full -> arrive empty).
__global__ void __launch_bounds__(1024) ring_kernel(const uint8_t* src, size_t bytes_per_cta, int chunk, int stages,
                                                    int nprod, int mode, unsigned* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // [nprod][stages][2]
  uint8_t* ring = smem + 1024;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < nprod * stages; ++i) {
      mbar_init(&bars[2 * i], mode == 0 ? 1 : 32);
      mbar_init(&bars[2 * i + 1], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const size_t per_prod = bytes_per_cta / nprod;
  const int n_it = (int)(per_prod / chunk);
  if (warp < nprod) {
    const uint8_t* s = src + (size_t)blockIdx.x * bytes_per_cta + (size_t)warp * per_prod;
    uint8_t* rg = ring + (size_t)warp * stages * chunk;
    uint64_t* b = bars + 2 * warp * stages;
    int slot = 0;
    uint32_t par = 1;
    for (int i = 0; i < n_it; ++i) {
      if (mode == 0) {
        if (lane == 0) {
          mbar_wait(&b[2 * slot + 1], par);
          mbar_expect(&b[2 * slot], chunk);
          bulk(rg + (size_t)slot * chunk, s + (size_t)i * chunk, chunk, &b[2 * slot]);
        }
      } else {
        mbar_wait(&b[2 * slot + 1], par);
        for (int o = lane * 16; o < chunk; o += 512)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s32(rg + (size_t)slot * chunk + o)),
                       "l"(s + (size_t)i * chunk + o) : "memory");
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(s32(&b[2 * slot])) : "memory");
      }
      if (++slot == stages) { slot = 0; par ^= 1; }
    }
  } else if (warp == nprod) {
    unsigned acc = 0;
    for (int pr = lane; pr < nprod; pr += 32) {}  // (consumer handles all producers round-robin below)
    int slot = 0;
    uint32_t par = 0;
    for (int i = 0; i < n_it; ++i) {
      for (int pr = 0; pr < nprod; ++pr) {
        uint64_t* b = bars + 2 * pr * stages;
        mbar_wait(&b[2 * slot], par);
        acc ^= *reinterpret_cast<const unsigned*>(ring + ((size_t)pr * stages + slot) * chunk + lane * 4);
        __syncwarp();
        if (lane == 0) mbar_arrive(&b[2 * slot + 1]);
      }
      if (++slot == stages) { slot = 0; par ^= 1; }
    }
    if (acc == 0x12345678u) sink[0] = acc;
  }
}

__global__ void __launch_bounds__(512) ldg_kernel(const uint4* src, size_t vec_per_cta, unsigned* sink) {
  const uint4* s = src + (size_t)blockIdx.x * vec_per_cta;
  unsigned acc = 0;
  for (size_t i = threadIdx.x; i + 7 * 512 < vec_per_cta; i += 8 * 512) {
    uint4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(s + i + u * 512));
#pragma unroll
    for (int u = 0; u < 8; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

int main() {
  const int sms = 148;
  const size_t per_cta = 8u << 20;  // 8 MiB per CTA -> 1.18 GB in total (>> L2)
  uint8_t* src;
  unsigned* sink;
  cudaMalloc(&src, per_cta * sms);
  cudaMalloc(&sink, 4);
  cudaMemset(src, 1, per_cta * sms);
  cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](const char* name, auto launch) {
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 3; ++i) launch();
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    printf("%-44s %8.1f GB/s  (%s)\n", name, 3.0 * per_cta * sms / (ms * 1e-3) / 1e9, cudaGetErrorString(err));
  };
  char name[128];
  for (int mode = 0; mode < 2; ++mode)
    for (int nprod : {1, 2, 4})
      for (int chunk : {4096, 16384, 32768})
        for (int stages : {2, 4, 8}) {
          size_t smem = 1024 + (size_t)nprod * stages * chunk;
          if (smem > 220 * 1024) continue;
          snprintf(name, sizeof name, "%s nprod=%d chunk=%5d stages=%d", mode == 0 ? "bulk  " : "ldgsts", nprod, chunk, stages);
          run(name, [&] { ring_kernel<<<sms, 32 * (nprod + 1), smem>>>(src, per_cta, chunk, stages, nprod, mode, sink); });
        }
  run("ldg.128 x8 unroll, 512 threads", [&] { ldg_kernel<<<sms, 512>>>((const uint4*)src, per_cta / 16, sink); });
  run("ldg.128 x8 unroll, 512 threads, 2 CTA/SM", [&] { ldg_kernel<<<2 * sms, 512>>>((const uint4*)src, per_cta / 32, sink); });
  return 0;
}
