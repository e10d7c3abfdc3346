// int8 weight-only linear, decode path (M <= 16): strip-parallel TMA-bulk ring + mma.sync, the int8 sibling of
// gemv_int4_fused.cu (same roles, same PDL chaining, same RMSNorm prologue / residual / SiLU*mul epilogues).
//
//   y[m][n] = bf16( bf16( sum_k x[m][k] * q[n][k] ) * scale[n] )          (plain_layout.py:250-266: mm rounds, then * scale)
//
// Weights are stored BIASED (u = q + 128, 0..255) in the strip-major int8 layout (include/ao_b200.h), so a byte pair
// becomes two fp16 values 1024 + u with ONE byte permute (0x64 in the high byte: fp16 1024.0 has ulp 1) — 2 ALU ops per 4
// weights, no xor, no int->float conversion.  The bias is removed on the accumulator, per 64 k:
//     sum x (u - 128) = sum x (1024 + u) - 1152 * sum x       (the accumulator of a segment starts at -1152 * sum x).
// Activations are staged once per CTA as fp16 after a per-row power-of-two scaling (exact for every bf16 value within
// 2^-27 of the row's maximum), as in the int4 kernel.  Per byte of weight this kernel issues 0.5 ALU op and 1/256 MMA
// (int4: 1.25 and 1/128), so unlike the int4 kernel it is bound by the copy engine, not by the ALU / HMMA pipes.
//
// Decomposition: CTA b owns a contiguous range of 16-row strips (one contiguous byte range of the weight); the producer
// warp issues ONE 16 KB bulk copy per stage (16 k-tiles = 1024 k of a strip); 16 consumer warps take one 64-k tile each;
// the epilogue warp sums the 16 partials of a finished strip, applies scale and the fused epilogue.
#include <cuda_fp16.h>

#include <cstdlib>
#include <type_traits>

#include "common.cuh"

namespace ao {

constexpr int k8Consumers = 16;
constexpr int k8Threads = (k8Consumers + 2) * 32;  // + TMA producer warp + epilogue warp
constexpr int k8RedBufs = 2;
constexpr int k8StageTiles = 16;                   // k-tiles (64 k x 16 rows = 1 KB) per stage
constexpr int k8StageBytes = k8StageTiles * 1024;

enum { k8EpiPlain = 0, k8EpiResidual = 1, k8EpiSiluMul = 2 };

struct Gemv8Params {
  const uint8_t* packed;          // strip-major int8, biased: [Np/16][Kp/64][2][32][16 B]
  const __nv_bfloat16* scale;     // [N_out] (rows in the packed order)
  const __nv_bfloat16* x;         // [M][ldx]
  const __nv_bfloat16* norm_w;    // [Kx] or null
  const __nv_bfloat16* residual;  // [M][ld_res] or null
  __nv_bfloat16* y;
  int M, stages, epilogue;
  int KT;                         // Kp / 64
  int64_t Kx, ldx, N_out, ldy, ld_res;
  int n_strips, strips_base, strips_rem;
  float eps;
  uint32_t x_row_bytes;           // 2 * Kp + 16
  uint32_t off_x, off_C, off_red, off_misc;
};

__device__ __forceinline__ float warp_sum8(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max8(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ void mma8_f16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// bytes (b0, b1) / (b2, b3) of a word -> fp16x2 (1024 + b0, 1024 + b1) / (1024 + b2, 1024 + b3)
__device__ __forceinline__ uint32_t u8_f16_lo(uint32_t w) { return __byte_perm(w, 0x64646464u, 0x4140); }
__device__ __forceinline__ uint32_t u8_f16_hi(uint32_t w) { return __byte_perm(w, 0x64646464u, 0x4342); }

// NH > 0: M = 1 and every consumer thread owns <= NH chunks of the activation row (register-held staging); NH = 0: the
// general staging.  One instantiation per case — the staging code runs once per launch with a cold instruction cache, so
// code that a launch does not need is not compiled into its kernel (profiles/r01c_imma_code_size_ab.txt).
__device__ __forceinline__ uint32_t mul_bf16x2_8(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

template <int MT, int NH>
__global__ void __launch_bounds__(k8Threads, (MT == 1 ? 2 : 1)) int8_gemv_fused_kernel(const Gemv8Params p) {
  extern __shared__ __align__(128) uint8_t smem8[];
  uint8_t* ring = smem8;
  uint8_t* x_s = smem8 + p.off_x;                                // [M][x_row_bytes]: bf16 while staging, then scaled fp16
  float* C_s = reinterpret_cast<float*>(smem8 + p.off_C);        // [Kp/64][8*MT]: -1152 * (sum of the staged x of the segment)
  float* red = reinterpret_cast<float*>(smem8 + p.off_red);      // [2][16 warps][MT][32][4]
  float* part = reinterpret_cast<float*>(smem8 + p.off_misc);    // [16 rows][16 warps][2 = {sum sq, amax}]
  float* row_scale = part + 16 * k8Consumers * 2;                // [16]: 2^e (* rsqrt(mean sq + eps))
  float* row_dn = row_scale + 16;                                // [16]: 2^-e
  uint64_t* full = reinterpret_cast<uint64_t*>(row_dn + 16);
  uint64_t* empty = full + p.stages;
  uint64_t* red_full = empty + p.stages;
  uint64_t* red_free = red_full + k8RedBufs;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const int b = blockIdx.x;
  const int s0 = b * p.strips_base + min(b, p.strips_rem), s1 = s0 + p.strips_base + (b < p.strips_rem ? 1 : 0);
  const int KT = p.KT;
  const int n_chunks = (KT + k8StageTiles - 1) / k8StageTiles;

  grid_launch_dependents();

  if (warp == k8Consumers) {
    // ===================== producer =====================
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], k8Consumers);
      }
      for (int s = 0; s < k8RedBufs; ++s) {
        mbar_init(&red_full[s], k8Consumers);
        mbar_init(&red_free[s], 1);
      }
      fence_barrier_init();
    }
    __syncwarp();
    named_bar_arrive(2, k8Threads);
    if (lane == 0) {
      int slot = 0;
      uint32_t par = 1;
      for (int strip = s0; strip < s1; ++strip) {
        const uint8_t* wrow = p.packed + (int64_t)strip * KT * 1024;
        for (int ch = 0; ch < n_chunks; ++ch) {
          const int kt0 = ch * k8StageTiles;
          const int cnt = min(k8StageTiles, KT - kt0);
          mbar_wait(&empty[slot], par);
          mbar_arrive_expect_tx(&full[slot], (uint32_t)(cnt * 1024));
          bulk_g2s(ring + (size_t)slot * k8StageBytes, wrow + (int64_t)kt0 * 1024, (uint32_t)(cnt * 1024), &full[slot]);
          if (++slot == stages) {
            slot = 0;
            par ^= 1u;
          }
        }
      }
    }
    __syncwarp();
    grid_dependency_wait();  // every kernel of a PDL chain must wait before it exits (transitive ordering)
    return;
  }

  if (warp == k8Consumers + 1) {
    // ===================== epilogue warp =====================
    const int g = lane >> 2, c = lane & 3;
    grid_dependency_wait();
    named_bar_sync(2, k8Threads);
    for (int strip = s0; strip < s1; ++strip) {
      const int li = strip - s0;
      const int rbuf = li % k8RedBufs;
      const uint32_t use = (uint32_t)(li / k8RedBufs);
      const float* rb = red + (size_t)rbuf * (k8Consumers * MT * 128);
      // the strip's two scales, fetched before the wait
      const int64_t n0 = 16 * (int64_t)strip + g;
      const float sc0 = n0 < p.N_out ? __bfloat162float(p.scale[n0]) : 0.f;
      const float sc1 = n0 + 8 < p.N_out ? __bfloat162float(p.scale[n0 + 8]) : 0.f;
      mbar_wait(&red_full[rbuf], use & 1u);
      float4 sum[MT];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        // warp order, packed fp32x2 adds: the same sums with half the instructions on the launch's tail
        uint64_t slo = 0ull, shi = 0ull;
#pragma unroll
        for (int w = 0; w < k8Consumers; ++w) {
          const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(rb + ((w * MT + mt) * 32 + lane) * 4);
          asm("add.rn.f32x2 %0, %0, %1;" : "+l"(slo) : "l"(v.x));
          asm("add.rn.f32x2 %0, %0, %1;" : "+l"(shi) : "l"(v.y));
        }
        sum[mt] = make_float4(__uint_as_float((uint32_t)slo), __uint_as_float((uint32_t)(slo >> 32)),
                              __uint_as_float((uint32_t)shi), __uint_as_float((uint32_t)(shi >> 32)));
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&red_free[rbuf]);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        const float vals[2][2] = {{sum[mt].x, sum[mt].z}, {sum[mt].y, sum[mt].w}};  // [col j][row half]
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int m = 8 * mt + 2 * c + j;
          if (m >= p.M) continue;
          const float rsc = row_scale[m];
          // mm result rounded to bf16, THEN the scale multiply rounds again (plain_layout.py:256-262)
          const float v0 = bf16r(bf16r(vals[j][0] * rsc) * sc0), v1 = bf16r(bf16r(vals[j][1] * rsc) * sc1);
          if (p.epilogue == k8EpiSiluMul) {
            const int64_t n = 8 * (int64_t)strip + g;  // rows g = gate (w1), rows g+8 = up (w3)
            if (2 * n < p.N_out) {
              const float sg = bf16r(v0 / (1.f + __expf(-v0)));
              p.y[m * p.ldy + n] = __float2bfloat16_rn(sg * v1);
            }
          } else if (p.epilogue == k8EpiResidual) {
            if (n0 < p.N_out) p.y[m * p.ldy + n0] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0]) + v0);
            if (n0 + 8 < p.N_out)
              p.y[m * p.ldy + n0 + 8] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0 + 8]) + v1);
          } else {
            if (n0 < p.N_out) p.y[m * p.ldy + n0] = __float2bfloat16_rn(v0);
            if (n0 + 8 < p.N_out) p.y[m * p.ldy + n0 + 8] = __float2bfloat16_rn(v1);
          }
        }
      }
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, c = lane & 3;
  const int tid = threadIdx.x;  // 0 .. 511
  constexpr int kCT = k8Consumers * 32;
  constexpr int kHold = NH > 0 ? NH : 1;  // chunks a thread keeps in registers between the two staging passes
  const int chunks_per_row = KT * 8;  // 16-byte chunks of 8 elements over Kp
  const int n_iter = (chunks_per_row + kCT - 1) / kCT;
  // Fast path (M = 1, Kp <= 16384): every thread owns <= 4 chunks, all of their loads are issued at once, the normalised
  // bf16 values stay in registers between the two passes, every warp reduces the row statistics itself, and the norm
  // weights (independent of the previous kernel, a DRAM miss) are fetched BEFORE the dependency wait.
  constexpr bool hold = NH > 0;  // host: p.M == 1 && n_iter <= NH
  uint4 xr[kHold], wn4[kHold];
#pragma unroll
  for (int i = 0; i < kHold; ++i) xr[i] = wn4[i] = make_uint4(0u, 0u, 0u, 0u);
  if (hold && p.norm_w != nullptr) {
#pragma unroll
    for (int i = 0; i < kHold; ++i) {
      const int ci = i * kCT + tid;
      if ((int64_t)ci * 8 < p.Kx) wn4[i] = __ldg(reinterpret_cast<const uint4*>(p.norm_w + (int64_t)ci * 8));
    }
  }
  grid_dependency_wait();

  // ---- stage the activations: x (* norm_w) -> per-row power-of-two scaling -> fp16; per-64-k correction sums ----
  if constexpr (hold) {
    float ss = 0.f, amax = 0.f;
#pragma unroll
    for (int i = 0; i < kHold; ++i) {
      const int ci = i * kCT + tid;
      if ((int64_t)ci * 8 < p.Kx) xr[i] = *reinterpret_cast<const uint4*>(p.x + (int64_t)ci * 8);
    }
#pragma unroll
    for (int i = 0; i < kHold; ++i) {
      float f[8];
      if (p.norm_w != nullptr) {
        unpack8(xr[i], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
        // bf16(x * w): the fp32 product of two bf16 is exact, so one packed bf16x2 multiply IS the reference's
        // compute-in-fp32-then-round (and leaves the values packed for the second pass)
        xr[i] = make_uint4(mul_bf16x2_8(xr[i].x, wn4[i].x), mul_bf16x2_8(xr[i].y, wn4[i].y), mul_bf16x2_8(xr[i].z, wn4[i].z),
                           mul_bf16x2_8(xr[i].w, wn4[i].w));
      }
      unpack8(xr[i], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(f[e]));
    }
    ss = warp_sum8(ss);
    amax = warp_max8(amax);
    if (lane == 0) {
      part[warp * 2] = ss;
      part[warp * 2 + 1] = amax;
    }
    named_bar_sync(1, kCT);
    float t = 0.f;
    amax = 0.f;
    if (lane < k8Consumers) {
      t = part[lane * 2];
      amax = part[lane * 2 + 1];
    }
    t = warp_sum8(t);
    amax = warp_max8(amax);
    int e = 0;
    if (amax > 0.f) e = (int)((__float_as_uint(amax) >> 23) & 0xFF) - 127 - 13;
    e = max(-100, min(100, e));
    const float dn = __uint_as_float((uint32_t)(127 - e) << 23);
    if (warp == 0 && lane == 0) {
      const float up = __uint_as_float((uint32_t)(e + 127) << 23);
      row_scale[0] = (p.norm_w != nullptr) ? up * rsqrtf(t / (float)p.Kx + p.eps) : up;  // read by the epilogue warp
    }
#pragma unroll
    for (int it = 0; it < kHold; ++it) {
      if (it < n_iter) {
        const int ci = it * kCT + tid;
        float xs = 0.f;
        if (ci < chunks_per_row) {
          float f[8];
          unpack8(xr[it], f);
          const __half2 h0 = __floats2half2_rn(f[0] * dn, f[1] * dn), h1 = __floats2half2_rn(f[2] * dn, f[3] * dn);
          const __half2 h2 = __floats2half2_rn(f[4] * dn, f[5] * dn), h3 = __floats2half2_rn(f[6] * dn, f[7] * dn);
          *reinterpret_cast<uint4*>(x_s + (size_t)ci * 16) =
              make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                         *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
          const float2 a0 = __half22float2(h0), a1 = __half22float2(h1), a2 = __half22float2(h2), a3 = __half22float2(h3);
          xs = ((a0.x + a0.y) + (a1.x + a1.y)) + ((a2.x + a2.y) + (a3.x + a3.y));
        }
#pragma unroll
        for (int o = 4; o; o >>= 1) xs += __shfl_xor_sync(0xffffffffu, xs, o);
        if (ci < chunks_per_row && (ci & 7) == 0) C_s[(ci >> 3) * 8 * MT] = -1152.f * xs;
      }
    }
  } else {
    // rows in batches of four: the four activation loads of a thread are in flight together (one exposed L2 round trip
    // per batch instead of one per row)
    for (int m0 = 0; m0 < p.M; m0 += 4) {
      float ss4[4] = {0.f, 0.f, 0.f, 0.f}, am4[4] = {0.f, 0.f, 0.f, 0.f};
      for (int ci = tid; ci < chunks_per_row; ci += kCT) {
        const int64_t k0 = (int64_t)ci * 8;
        uint4 v4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v4[j] = make_uint4(0, 0, 0, 0);
          if (m0 + j < p.M && k0 < p.Kx) v4[j] = *reinterpret_cast<const uint4*>(p.x + (int64_t)(m0 + j) * p.ldx + k0);
        }
        float wn[8];
        const bool normed = p.norm_w != nullptr && k0 < p.Kx;
        if (normed) unpack8(__ldg(reinterpret_cast<const uint4*>(p.norm_w + k0)), wn);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (m0 + j >= p.M) break;
          uint4 v = v4[j];
          float f[8];
          unpack8(v, f);
          if (normed) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              ss4[j] = fmaf(f[e], f[e], ss4[j]);
              f[e] = bf16r(f[e] * wn[e]);
            }
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) o[e] = float_to_bf16_bits(f[2 * e]) | (float_to_bf16_bits(f[2 * e + 1]) << 16);
            v = make_uint4(o[0], o[1], o[2], o[3]);
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) am4[j] = fmaxf(am4[j], fabsf(f[e]));
          *reinterpret_cast<uint4*>(x_s + (size_t)(m0 + j) * p.x_row_bytes + (size_t)ci * 16) = v;
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float ss = warp_sum8(ss4[j]), amax = warp_max8(am4[j]);
        if (lane == 0 && m0 + j < p.M) {
          part[((m0 + j) * k8Consumers + warp) * 2] = ss;
          part[((m0 + j) * k8Consumers + warp) * 2 + 1] = amax;
        }
      }
    }
    named_bar_sync(1, kCT);
    if (warp < p.M) {
      float t = 0.f, amax = 0.f;
      if (lane < k8Consumers) {
        t = part[(warp * k8Consumers + lane) * 2];
        amax = part[(warp * k8Consumers + lane) * 2 + 1];
      }
      t = warp_sum8(t);
      amax = warp_max8(amax);
      if (lane == 0) {
        // 2^e brings the row's largest magnitude to [2^13, 2^14)
        int e = 0;
        if (amax > 0.f) e = (int)((__float_as_uint(amax) >> 23) & 0xFF) - 127 - 13;
        e = max(-100, min(100, e));
        const float up = __uint_as_float((uint32_t)(e + 127) << 23), dn = __uint_as_float((uint32_t)(127 - e) << 23);
        row_dn[warp] = dn;
        row_scale[warp] = (p.norm_w != nullptr) ? up * rsqrtf(t / (float)p.Kx + p.eps) : up;
      }
    }
    named_bar_sync(1, kCT);
    for (int m = 0; m < p.M; ++m) {
      const float dn = row_dn[m];
      for (int base = 0; base < chunks_per_row; base += kCT) {
        const int ci = base + tid;
        float xs = 0.f;
        if (ci < chunks_per_row) {
          uint4* ptr = reinterpret_cast<uint4*>(x_s + (size_t)m * p.x_row_bytes + (size_t)ci * 16);
          float f[8];
          unpack8(*ptr, f);
          const __half2 h0 = __floats2half2_rn(f[0] * dn, f[1] * dn), h1 = __floats2half2_rn(f[2] * dn, f[3] * dn);
          const __half2 h2 = __floats2half2_rn(f[4] * dn, f[5] * dn), h3 = __floats2half2_rn(f[6] * dn, f[7] * dn);
          *ptr = make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                            *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
          const float2 a0 = __half22float2(h0), a1 = __half22float2(h1), a2 = __half22float2(h2), a3 = __half22float2(h3);
          xs = ((a0.x + a0.y) + (a1.x + a1.y)) + ((a2.x + a2.y) + (a3.x + a3.y));
        }
        // a 64-k segment = 8 consecutive chunks = 8 consecutive lanes (kCT % 8 == 0, so segments never straddle `base`)
#pragma unroll
        for (int o = 4; o; o >>= 1) xs += __shfl_xor_sync(0xffffffffu, xs, o);
        if (ci < chunks_per_row && (ci & 7) == 0) C_s[(ci >> 3) * 8 * MT + m] = -1152.f * xs;
      }
    }
  }
  // one barrier closes the staging (activations, correction sums, row scales visible to every consumer and to the epilogue
  // warp) and meets the producer, which initialised the mbarriers long ago
  named_bar_sync(2, k8Threads);

  // ---- main loop: warp w owns k-tile w of every stage ----
  constexpr uint32_t CS = 8 * MT * 4;  // bytes of one segment's record
  const uint32_t ring_u = smem_u32(ring);
  uint32_t w_off = (uint32_t)warp * 1024u + (uint32_t)lane * 16u;
  uint32_t C_u0 = smem_u32(C_s) + (uint32_t)warp * CS + (uint32_t)c * 8u;
  uint32_t x_u0[MT];
  bool x_on[MT];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    x_on[mt] = g + 8 * mt < p.M;
    x_u0[mt] = smem_u32(x_s) + (uint32_t)min(g + 8 * mt, p.M - 1) * p.x_row_bytes + (uint32_t)c * 16u + (uint32_t)warp * 128u;
  }
  asm volatile("" : "+r"(w_off), "+r"(C_u0));
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) asm volatile("" : "+r"(x_u0[mt]));
  int slot = 0;
  uint32_t par = 0;
  uint32_t stage_u = ring_u;

  for (int strip = s0; strip < s1; ++strip) {
    float yacc[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int e = 0; e < 4; ++e) yacc[mt][e] = 0.f;

    uint32_t xo = 0, C_u = C_u0;
    for (int ch = 0; ch < n_chunks; ++ch) {
      mbar_wait(&full[slot], par);
      if (warp < KT - ch * k8StageTiles) {
        const uint4 wa = lds128(stage_u + w_off), wb = lds128(stage_u + w_off + 512u);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          uint4 xa = make_uint4(0, 0, 0, 0), xb = make_uint4(0, 0, 0, 0);
          if (x_on[mt]) {
            xa = lds128(x_u0[mt] + xo);
            xb = lds128(x_u0[mt] + xo + 64u);
          }
          const float2 nc = lds64f(C_u + mt * 32);
          float a0[4] = {nc.x, nc.y, nc.x, nc.y}, a1[4] = {0.f, 0.f, 0.f, 0.f};
          mma8_f16_16816(a0, u8_f16_lo(wa.x), u8_f16_lo(wa.y), u8_f16_hi(wa.x), u8_f16_hi(wa.y), xa.x, xa.y);
          mma8_f16_16816(a1, u8_f16_lo(wa.z), u8_f16_lo(wa.w), u8_f16_hi(wa.z), u8_f16_hi(wa.w), xa.z, xa.w);
          mma8_f16_16816(a0, u8_f16_lo(wb.x), u8_f16_lo(wb.y), u8_f16_hi(wb.x), u8_f16_hi(wb.y), xb.x, xb.y);
          mma8_f16_16816(a1, u8_f16_lo(wb.z), u8_f16_lo(wb.w), u8_f16_hi(wb.z), u8_f16_hi(wb.w), xb.z, xb.w);
#pragma unroll
          for (int e = 0; e < 4; ++e) yacc[mt][e] += a0[e] + a1[e];
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      xo += k8StageTiles * 128u;
      C_u += k8StageTiles * CS;
      stage_u += k8StageBytes;
      if (++slot == stages) {
        slot = 0;
        par ^= 1u;
        stage_u = ring_u;
      }
    }

    const int li = strip - s0;
    const int rbuf = li % k8RedBufs;
    const uint32_t use = (uint32_t)(li / k8RedBufs);
    float* rb = red + (size_t)rbuf * (k8Consumers * MT * 128);
    if (use > 0) mbar_wait(&red_free[rbuf], (use - 1) & 1u);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
      *reinterpret_cast<float4*>(rb + ((warp * MT + mt) * 32 + lane) * 4) =
          make_float4(yacc[mt][0], yacc[mt][1], yacc[mt][2], yacc[mt][3]);
    __syncwarp();
    if (lane == 0) mbar_arrive(&red_full[rbuf]);
  }
}

// ------------------------------------------------------------------------------------------------ host side
struct Plan8 {
  int stages;
  size_t smem;
  uint32_t off_x, off_C, off_red, off_misc, x_row_bytes;
};

static bool plan8(int M, int MT, int64_t Kp, int max_stages, Plan8& pl) {
  pl.x_row_bytes = (uint32_t)(2 * Kp + 16);
  const size_t x_bytes = ((size_t)M * pl.x_row_bytes + 127) / 128 * 128;
  const size_t C_bytes = ((size_t)(Kp / 64) * 8 * MT * 4 + 127) / 128 * 128;
  const size_t red_bytes = (size_t)k8RedBufs * k8Consumers * MT * 128 * 4;
  const size_t misc = (16 * k8Consumers * 2 + 32) * 4 + (2 * 16 + 2 * k8RedBufs) * 8 + 64;
  const size_t fixed = x_bytes + C_bytes + red_bytes + misc;
  const size_t two_per_sm = 113 * 1024, one_per_sm = 226 * 1024;
  int st = 0;
  if (MT == 1 && fixed + 4 * (size_t)k8StageBytes <= two_per_sm) {
    st = (int)((two_per_sm - fixed) / k8StageBytes);
  } else if (fixed + 3 * (size_t)k8StageBytes <= one_per_sm) {
    st = (int)((one_per_sm - fixed) / k8StageBytes);
  } else {
    return false;
  }
  if (st > max_stages) st = max_stages;
  if (st > 16) st = 16;
  pl.stages = st;
  size_t off = (size_t)st * k8StageBytes;
  pl.off_x = (uint32_t)off; off += x_bytes;
  pl.off_C = (uint32_t)off; off += C_bytes;
  pl.off_red = (uint32_t)off; off += red_bytes;
  pl.off_misc = (uint32_t)off; off += misc;
  pl.smem = off;
  return true;
}

static int sm_count8() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n = v > 0 ? v : kNumSMs;
  }
  return n;
}

template <int MT, int NH>
static int launch8(Gemv8Params& p, const Plan8& pl, bool pdl, cudaStream_t s) {
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(int8_gemv_fused_kernel<MT, NH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  const int max_grid = sm_count8() * (pl.smem <= 113 * 1024 && MT == 1 ? 2 : 1);
  // balanced waves: every CTA gets the same number of strips (+-1 at most on the last few), e.g. 1792 strips on 296 slots
  // -> 256 CTAs x 7 instead of 16 CTAs with a 7th strip while 280 idle
  int grid = p.n_strips < max_grid ? p.n_strips : max_grid;
  {
    static const bool balance = []() { const char* e = std::getenv("AO_B200_BALANCE"); return e ? std::atoi(e) != 0 : false; }();
    const int per = (p.n_strips + grid - 1) / grid;
    if (balance) grid = (p.n_strips + per - 1) / per;
  }
  p.strips_base = p.n_strips / grid;
  p.strips_rem = p.n_strips % grid;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(k8Threads);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, int8_gemv_fused_kernel<MT, NH>, p) != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

// returns AO_B200_ERR_WORKSPACE when the activations do not fit the staging (caller falls back)
int int8_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Np,
                    int64_t Kp, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps, const void* residual,
                    int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s) {
  Gemv8Params p;
  p.packed = static_cast<const uint8_t*>(packed);
  p.scale = static_cast<const __nv_bfloat16*>(scale);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.norm_w = static_cast<const __nv_bfloat16*>(norm_w);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.epilogue = epilogue;
  p.KT = (int)(Kp / 64);
  p.Kx = Kx; p.ldx = ldx; p.N_out = N_out; p.ldy = ldy; p.ld_res = ld_res;
  p.n_strips = (int)((N_out + 15) / 16);  // padding strips beyond N_out are never read
  (void)Np;
  p.eps = eps;
  const int MT = M <= 8 ? 1 : 2;
  Plan8 pl;
  if (!plan8((int)M, MT, Kp, max_stages, pl)) return AO_B200_ERR_WORKSPACE;
  p.stages = pl.stages;
  p.x_row_bytes = pl.x_row_bytes;
  p.off_x = pl.off_x; p.off_C = pl.off_C; p.off_red = pl.off_red; p.off_misc = pl.off_misc;
  if (MT == 2) return launch8<2, 0>(p, pl, pdl, s);
  // register-held staging: M = 1 and <= 4 chunks of 8 activations per consumer thread (Kp <= 16384)
  const int n_iter = (p.KT * 8 + k8Consumers * 32 - 1) / (k8Consumers * 32);
  if (M == 1 && n_iter <= 1) return launch8<1, 1>(p, pl, pdl, s);
  if (M == 1 && n_iter <= 2) return launch8<1, 2>(p, pl, pdl, s);
  if (M == 1 && n_iter <= 4) return launch8<1, 4>(p, pl, pdl, s);
  return launch8<1, 0>(p, pl, pdl, s);
}

}  // namespace ao
