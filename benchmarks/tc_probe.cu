// Mechanics probe for the tcgen05 int4 path: A operand written to TMEM with tcgen05.st.16x128b (the thread <-> (lane,
// column) map of an mma.sync fragment, i.e. of the strip-major packed word layout), B operand in shared memory in the
// canonical no-swizzle K-major layout (optionally with the two 8-row groups ALIASED, SBO = 0), tcgen05.mma M=128 N=16
// K=16 with A from TMEM, accumulator read back with tcgen05.ld.32x32b.  Compares with a host matmul.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/tc_probe benchmarks/tc_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int KT = 64;  // k per test (4 MMAs of K = 16)

__global__ void __launch_bounds__(128) probe(const __half* __restrict__ A /*[128][KT]*/, const __half* __restrict__ B /*[16][KT]*/,
                                             float* __restrict__ D /*[128][16]*/, int alias_rows) {
  __shared__ __align__(128) uint8_t bs[16 * KT * 2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  // ---- B -> shared, canonical K-major no-swizzle: 8 rows x 16 B core matrices ----
  const uint32_t LBO = alias_rows ? 128 : 256, SBO = alias_rows ? 0 : 128;
  for (int i = threadIdx.x; i < 16 * KT; i += 128) {
    const int n = i / KT, k = i % KT;
    if (alias_rows && n >= 8) continue;
    *reinterpret_cast<__half*>(bs + (k / 8) * LBO + (n / 8) * SBO + (n % 8) * 16 + (k % 8) * 2) = B[n * KT + k];
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_s)), "r"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tmem_base_s;
  const uint32_t a_col = 0, d_col = 32;  // A: 32 columns (64 halves of K), D: 16 columns

  // ---- A -> TMEM.  Warp w owns lanes 32w .. 32w+31 = strips 2w, 2w+1; thread (g, c), register 2r + h of a
  //      16x128b.x8 store -> lane 16 strip + g + 8 h, column 4 r + c (r = 0..7): column p holds k = 2p, 2p + 1 ----
  for (int sh = 0; sh < 2; ++sh) {
    const int strip = 2 * warp + sh;
    uint32_t r[16];
#pragma unroll
    for (int rr = 0; rr < 8; ++rr)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int row = 16 * strip + g + 8 * h, col = 4 * rr + c;
        const __half2 v = __halves2half2(A[row * KT + 2 * col], A[row * KT + 2 * col + 1]);
        r[2 * rr + h] = *reinterpret_cast<const uint32_t*>(&v);
      }
    const uint32_t taddr = tb + ((uint32_t)(16 * strip) << 16) + a_col;
    asm volatile(
        "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes of B -> visible to the MMA's async proxy
    // instruction descriptor: D f32, A/B f16, K-major both, N = 16, M = 128
    const uint32_t idesc = (1u << 4) | (0u << 7) | (0u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
    for (int ks = 0; ks < KT / 16; ++ks) {
      const uint32_t start = s32(bs) + ks * 2 * LBO;
      const uint64_t bdesc = (uint64_t)((start >> 4) & 0x3FFF) | ((uint64_t)((LBO >> 4) & 0x3FFF) << 16) |
                             ((uint64_t)((SBO >> 4) & 0x3FFF) << 32) | (1ull << 46);
      const uint32_t acc = ks > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tb + d_col),
          "r"(tb + a_col + 8 * ks), "l"(bdesc), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
  }
  {
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0,1,0,p;}"
                   : "=r"(ok) : "r"(s32(&bar)) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // ---- D: thread (warp w, lane l) reads lane 32 w + l, 16 columns ----
  uint32_t d[16];
  const uint32_t daddr = tb + ((uint32_t)(32 * warp) << 16) + d_col;
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]), "=r"(d[9]),
        "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
      : "r"(daddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 16; ++j) D[(32 * warp + lane) * 16 + j] = __uint_as_float(d[j]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(64));
}

// ---- kind::i8: A u8 [128][64] in TMEM (4 k per 32-bit column), B s8 [16][64] in shared memory (8 rows x 16 B core
//      matrices = 16 k per row), D s32.  Two MMAs of K = 32. ----
__global__ void __launch_bounds__(128) probe_i8(const uint8_t* __restrict__ A, const int8_t* __restrict__ B, int* __restrict__ D,
                                                int alias_rows) {
  __shared__ __align__(128) uint8_t bs[16 * KT];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t LBO = alias_rows ? 128 : 256, SBO = alias_rows ? 0 : 128;
  for (int i = threadIdx.x; i < 16 * KT; i += 128) {
    const int n = i / KT, k = i % KT;
    if (alias_rows && n >= 8) continue;
    bs[(k / 16) * LBO + (n / 8) * SBO + (n % 8) * 16 + (k % 16)] = (uint8_t)B[n * KT + k];
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_s)), "r"(64));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tmem_base_s;
  const uint32_t a_col = 0, d_col = 32;  // A: 16 columns (64 bytes of K), D: 16 columns
  for (int sh = 0; sh < 2; ++sh) {
    const int strip = 2 * warp + sh;
    uint32_t r[8];
#pragma unroll
    for (int rr = 0; rr < 4; ++rr)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int row = 16 * strip + g + 8 * h, col = 4 * rr + c;
        r[2 * rr + h] = *reinterpret_cast<const uint32_t*>(A + row * KT + 4 * col);
      }
    const uint32_t taddr = tb + ((uint32_t)(16 * strip) << 16) + a_col;
    asm volatile("tcgen05.st.sync.aligned.16x128b.x4.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
  }
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    // D s32 (2), A unsigned 8-bit (0), B signed 8-bit (1), N = 16, M = 128
    const uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
    for (int ks = 0; ks < KT / 32; ++ks) {
      const uint32_t start = s32(bs) + ks * 2 * LBO;
      const uint64_t bdesc = (uint64_t)((start >> 4) & 0x3FFF) | ((uint64_t)((LBO >> 4) & 0x3FFF) << 16) |
                             ((uint64_t)((SBO >> 4) & 0x3FFF) << 32) | (1ull << 46);
      const uint32_t acc = ks > 0;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tb + d_col),
          "r"(tb + a_col + 8 * ks), "l"(bdesc), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
  }
  {
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0,1,0,p;}"
                   : "=r"(ok) : "r"(s32(&bar)) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t d[16];
  const uint32_t daddr = tb + ((uint32_t)(32 * warp) << 16) + d_col;
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]), "=r"(d[9]),
        "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
      : "r"(daddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 16; ++j) D[(32 * warp + lane) * 16 + j] = (int)d[j];
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(64));
}

// ---- rates: (a) back-to-back tcgen05.mma M128 N{16,32,64} K16 with A in TMEM, (b) tcgen05.st.16x128b.x8 from 16 warps ----
__global__ void __launch_bounds__(512) rate_kernel(int mode, int n_mma, int iters, long long* out) {
  __shared__ __align__(128) uint8_t bs[8192];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(8) uint64_t bar;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 2048; i += 512) reinterpret_cast<uint32_t*>(bs)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tb = tmem_base_s;
  long long t0 = clock64();
  if (mode == 0 || mode == 3 || mode == 4 || mode == 5) {
    if (threadIdx.x == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const uint32_t idesc = (mode == 3 ? ((2u << 4) | (1u << 10)) : (1u << 4)) | (((uint32_t)n_mma >> 3) << 17) | ((128u >> 4) << 24);
      const uint64_t bdesc = (uint64_t)((s32(bs) >> 4) & 0x3FFF) | ((uint64_t)(128 >> 4) << 16) | (0ull << 32) | (1ull << 46);
      // the issuing thread must not be the bottleneck: loop-invariant operands, 16 MMAs per trip
      const uint32_t d0 = tb + 256, d1 = tb + 320;
      if (mode == 4 || mode == 5) {
        __shared__ __align__(8) uint64_t dummy[4];
        for (int j = 0; j < 4; ++j) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&dummy[j])), "r"(1 << 19));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const uint32_t idesc8 = (2u << 4) | (1u << 10) | (((uint32_t)n_mma >> 3) << 17) | ((128u >> 4) << 24);
        for (int i = 0; i < iters; i += 16) {
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const uint64_t bd = mode == 5 ? bdesc + (uint64_t)(((i + u) & 31) * 16) : bdesc;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(mode == 4 ? d0 + 16 * (u >> 2) : d0),
                "r"(tb + 8 * u), "l"(bd), "r"(idesc8), "r"(mode == 4 ? ((u & 3) ? 1 : 0) : 1)
                : "memory");
            if (mode == 4 && (u & 3) == 3)
              asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&dummy[u >> 2])) : "memory");
          }
        }
      } else if (mode == 3) {
        for (int i = 0; i < iters; i += 16) {
#pragma unroll
          for (int u = 0; u < 16; ++u)
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"((u & 1) ? d1 : d0),
                "r"(tb + 8 * u), "l"(bdesc), "r"(idesc), "r"(1)
                : "memory");
        }
      } else {
        for (int i = 0; i < iters; i += 16) {
#pragma unroll
          for (int u = 0; u < 16; ++u)
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"((u & 1) ? d1 : d0),
                "r"(tb + 8 * u), "l"(bdesc), "r"(idesc), "r"(1)
                : "memory");
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
      uint32_t ok = 0;
      while (!ok)
        asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0,1,0,p;}"
                     : "=r"(ok) : "r"(s32(&bar)) : "memory");
    }
  } else {
    uint32_t r[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) r[j] = threadIdx.x + j;
    const uint32_t taddr = tb + ((uint32_t)(32 * (warp & 3)) << 16) + 32 * (warp >> 2);
    for (int i = 0; i < iters; ++i) {
      asm volatile(
          "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(
              taddr + ((i & 1) << 20)),
          "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
          "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
          : "memory");
      if (mode == 2) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512));
}

int main() {
  {
    long long* dout;
    cudaMalloc(&dout, 8);
    const int iters = 4096;
    for (int kind : {0, 3}) for (int n : {16, 32, 64, 128}) {
      long long c = 1LL << 60, cmax = 0;
      cudaError_t e = cudaSuccess;
      for (int rep = 0; rep < 12; ++rep) {
        rate_kernel<<<148, 512>>>(kind, n, iters, dout);
        long long cc = 0;
        e = cudaDeviceSynchronize();
        cudaMemcpy(&cc, dout, 8, cudaMemcpyDeviceToHost);
        if (cc < c) c = cc;
        if (cc > cmax) cmax = cc;
      }
      printf("[min..max over 12 runs: %.1f .. %.1f] ", (double)c / iters, (double)cmax / iters);
      printf("tcgen05.mma kind::%s M128 N%-3d (A in TMEM): %.1f clk per MMA  (%s)\n", kind ? "i8  K32" : "f16 K16", n, (double)c / iters, cudaGetErrorString(e));
    }
    for (int nn : {8, 16}) for (int mode : {4, 5}) {
      long long c = 1LL << 60;
      cudaError_t e = cudaSuccess;
      for (int rep = 0; rep < 5; ++rep) {
        rate_kernel<<<148, 512>>>(mode, nn, iters, dout);
        long long cc = 0;
        e = cudaDeviceSynchronize();
        cudaMemcpy(&cc, dout, 8, cudaMemcpyDeviceToHost);
        if (cc < c) c = cc;
      }
      printf("kind::i8 N%d, %s: %.1f clk per MMA (%s)\n", nn, mode == 4 ? "new accumulator + commit every 4 MMAs" : "one accumulator, descriptor changes every MMA",
             (double)c / iters, cudaGetErrorString(e));
    }
    for (int mode : {1, 2}) {
      rate_kernel<<<148, 512>>>(mode, 16, iters, dout);
      long long c = 0;
      cudaError_t e = cudaDeviceSynchronize();
      cudaMemcpy(&c, dout, 8, cudaMemcpyDeviceToHost);
      printf("tcgen05.st.16x128b.x8 from 16 warps%s: %.1f clk per store per warp, %.1f B/clk/SM  (%s)\n",
             mode == 2 ? " + wait::st each" : "", (double)c / iters, 16.0 * 2048 * iters / c, cudaGetErrorString(e));
    }
  }

  std::vector<__half> hA(128 * KT), hB(16 * KT);
  srand(1);
  for (auto& v : hA) v = __float2half((float)(rand() % 31 - 15));
  for (auto& v : hB) v = __float2half((float)(rand() % 15 - 7) * 0.25f);
  __half *dA, *dB;
  float* dD;
  cudaMalloc(&dA, hA.size() * 2);
  cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dD, 128 * 16 * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  for (int alias = 0; alias < 2; ++alias) {
    cudaMemset(dD, 0xff, 128 * 16 * 4);
    probe<<<1, 128>>>(dA, dB, dD, alias);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> hD(128 * 16);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 16; ++n) {
        double ref = 0;
        const int nb = alias ? n % 8 : n;
        for (int k = 0; k < KT; ++k) ref += (double)__half2float(hA[m * KT + k]) * __half2float(hB[nb * KT + k]);
        const double err = fabs(ref - hD[m * 16 + n]);
        if (err > 1e-3) ++bad;
        if (err > maxerr) maxerr = err;
      }
    printf("alias_rows=%d: %s, max |err| = %g, mismatches = %d / 2048;  D[0][0..3] = %g %g %g %g, D[17][0..1] = %g %g\n", alias,
           cudaGetErrorString(e), maxerr, bad, hD[0], hD[1], hD[2], hD[3], hD[17 * 16], hD[17 * 16 + 1]);
  }
  {
    std::vector<uint8_t> hA8(128 * KT);
    std::vector<int8_t> hB8(16 * KT);
    srand(2);
    for (auto& v : hA8) v = rand() % 16;
    for (auto& v : hB8) v = (int8_t)(rand() % 256 - 128);
    uint8_t* dA8;
    int8_t* dB8;
    int* dD8;
    cudaMalloc(&dA8, hA8.size());
    cudaMalloc(&dB8, hB8.size());
    cudaMalloc(&dD8, 128 * 16 * 4);
    cudaMemcpy(dA8, hA8.data(), hA8.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(dB8, hB8.data(), hB8.size(), cudaMemcpyHostToDevice);
    for (int alias = 0; alias < 2; ++alias) {
      cudaMemset(dD8, 0xff, 128 * 16 * 4);
      probe_i8<<<1, 128>>>(dA8, dB8, dD8, alias);
      cudaError_t e = cudaDeviceSynchronize();
      std::vector<int> hD(128 * 16);
      cudaMemcpy(hD.data(), dD8, hD.size() * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 16; ++n) {
          int ref = 0;
          const int nb = alias ? n % 8 : n;
          for (int k = 0; k < KT; ++k) ref += (int)hA8[m * KT + k] * (int)hB8[nb * KT + k];
          if (ref != hD[m * 16 + n]) ++bad;
        }
      printf("kind::i8 alias_rows=%d: %s, mismatches = %d / 2048; D[0][0..2] = %d %d %d\n", alias, cudaGetErrorString(e), bad,
             hD[0], hD[1], hD[2]);
    }
  }
  return 0;
}
