// int4 weight-only linear, decode path for M <= 2 (every group size): the integer-tensor-core sibling of gemv_int4_fused.cu.
//
// Same skeleton (strip-parallel, one 16 KB bulk copy + one scale/zero copy per stage into a shared-memory ring, 16
// consumer warps, producer warp, epilogue warp, PDL, RMSNorm prologue, residual / SiLU*mul epilogues), but the inner
// product runs on mma.sync.m16n8k32.s32.u8.s8:
//   * nibbles -> bytes is 1 shift + 2 AND per packed word (the fp16 path needs 5 ops per word), and one MMA covers 32 k
//     (fp16: 16) at the same issue cost (measured 8.2 vs 8.4 clk per MMA per SM sub-partition, profiles/r01_mma_sync_rate.txt):
//     about 40 % fewer issue slots per byte of weight than gemv_int4_fused.cu, whose cap is the ALU + HMMA pipes;
//   * the activation row is staged ONCE per CTA as 23-bit fixed point (per-row power-of-two scale, exact for every bf16
//     value within 2^-15 of the row's maximum) split into three balanced int8 limbs, and the limbs ride in three of the
//     eight MMA columns (M = 2: six), so the three partial products cost nothing extra;
//   * int32 accumulation over a segment (min(G, 128) k of one group) is exact; the accumulators start at the integer image
//     of 1.5 * 2^23, so int -> float is one subtraction; the affine dequant is applied per segment on the accumulator in fp32,
//         y[col] += s_g * acc[col] + (z_g - 8 s_g) * S[col],   S[col] = sum of the limb over the segment,
//     and the limbs are recombined (65536 a + 256 b + c) once per strip in the epilogue warp.  A warp's share of a stage
//     is 128 k: one segment at G >= 128 (G = 256: two warps share a group's scale / zero), two at G = 64, four at G = 32.
// Arithmetic: fp32-dequant semantics, like gemv_int4_fused.cu; the products are exact integers.
#include <cstdlib>

#include "common.cuh"

namespace ao {

constexpr int kIConsumers = 16;
constexpr int kIThreads = (kIConsumers + 2) * 32;
constexpr int kIRedBufs = 2;
constexpr int kIStageTiles = 32;                       // 64-k tiles per stage: 2 per consumer warp = one 128-k group
constexpr int kIStageW = kIStageTiles * 512;           // 16 KB of packed weights
template <int G>
struct IGeom {
  static constexpr int kSegK = G < 128 ? G : 128;        // k covered by one accumulator fix-up
  static constexpr int kNSeg = 128 / kSegK;              // segments in a warp's 128 k of a stage
  static constexpr int kSegLanes = kSegK / 8;            // 8-element chunks (= staging lanes) per segment
  static constexpr int kStageSZ = (kIStageTiles * 64 / G) * 64;  // groups of a stage x 16 rows x (scale | zero << 16)
  static constexpr int kStageBytes = kIStageW + kStageSZ;
};
constexpr uint32_t kIMagicI = 0x4B400000u;             // bits of 1.5 * 2^23: int adds land in the mantissa (ulp 1)
constexpr float kIMagicF = 12582912.0f;

struct GemvIParams {
  const uint8_t* packed;          // strip-major [Np/16][Kp/64][512 B]
  const uint32_t* sz;             // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* x;
  const __nv_bfloat16* norm_w;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
  int M, stages, epilogue, hold;
  int KT;                         // Kp / 64 (Kp % max(G, 128) == 0)
  int64_t Kx, ldx, N_out, ldy, ld_res;
  int n_strips, strips_base, strips_rem;  // CTA b owns strips [b base + min(b, rem), ...)
  float eps;
  uint32_t limb_row_bytes;        // Kp + 32 (row stride = 32 mod 128: the limb rows of one load sit in different banks)
  uint32_t off_x, off_S, off_red, off_misc;
};

__device__ __forceinline__ float warp_sum_i(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max_i(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint32_t absmax_bits(uint32_t m, float v) { return max(m, __float_as_uint(v) & 0x7fffffffu); }
__device__ __forceinline__ uint32_t warp_max_bits(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// D(16x8, s32) += A(16x32, u8, row) * B(32x8, s8, col)
__device__ __forceinline__ void imma_16832(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                           uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t mul_bf16x2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint2 lds64u(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}

// NH > 0: M = 1 and every thread owns <= NH 8-element chunks of the activation row (register-held staging); NH = 0: the
// general two-pass staging.  One instantiation per case: the staging code runs once per launch, instruction-cache cold,
// so what is not needed is not compiled in.
template <int G, int NH>
__global__ void __launch_bounds__(kIThreads, 2) int4_gemv_imma_kernel(const GemvIParams p) {
  using GE = IGeom<G>;
  constexpr int kIStageBytes = GE::kStageBytes;
  extern __shared__ __align__(128) uint8_t smem_i[];
  uint8_t* ring = smem_i;
  uint8_t* x_s = smem_i + p.off_x;                             // [3 M limb rows][limb_row_bytes] int8, k permuted per 8
  float* S_s = reinterpret_cast<float*>(smem_i + p.off_S);     // [Kp/kSegK][8 cols]: limb sums of the segment
  float* red = reinterpret_cast<float*>(smem_i + p.off_red);   // [2][16 warps][32][4]
  float* part = reinterpret_cast<float*>(smem_i + p.off_misc); // [2 rows][16 warps][2 = {sum sq, amax}]
  float* row_scale = part + 2 * kIConsumers * 2;               // [2]: 2^-e (* rsqrt(mean sq + eps))
  float* row_up = row_scale + 2;                               // (spare)
  uint64_t* full = reinterpret_cast<uint64_t*>(row_up + 2);
  uint64_t* empty = full + p.stages;
  uint64_t* red_full = empty + p.stages;
  uint64_t* red_free = red_full + kIRedBufs;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const int b = blockIdx.x;
  const int s0 = b * p.strips_base + min(b, p.strips_rem), s1 = s0 + p.strips_base + (b < p.strips_rem ? 1 : 0);
  const int KT = p.KT;

  grid_launch_dependents();

  if (warp == kIConsumers) {
    // ===================== producer =====================
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], kIConsumers);
      }
      for (int s = 0; s < kIRedBufs; ++s) {
        mbar_init(&red_full[s], kIConsumers);
        mbar_init(&red_free[s], 1);
      }
      fence_barrier_init();
    }
    __syncwarp();
    named_bar_arrive(2, kIThreads);
    if (lane == 0) {
      int slot = 0;
      uint32_t par = 1;
      const int n_groups = KT * 64 / G;
      for (int strip = s0; strip < s1; ++strip) {
        const uint8_t* wrow = p.packed + (int64_t)strip * KT * 512;
        const uint32_t* srow = p.sz + (int64_t)strip * n_groups * 16;
        for (int kt0 = 0; kt0 < KT; kt0 += kIStageTiles) {
          const int cnt = min(kIStageTiles, KT - kt0);
          uint8_t* dst = ring + (size_t)slot * kIStageBytes;
          mbar_wait(&empty[slot], par);
          const uint32_t sz_bytes = (uint32_t)(cnt * 64 / G) * 64u;
          mbar_arrive_expect_tx(&full[slot], (uint32_t)(cnt * 512) + sz_bytes);
          bulk_g2s(dst + kIStageW, srow + (int64_t)(kt0 * 64 / G) * 16, sz_bytes, &full[slot]);
          bulk_g2s(dst, wrow + (int64_t)kt0 * 512, (uint32_t)(cnt * 512), &full[slot]);
          if (++slot == stages) {
            slot = 0;
            par ^= 1u;
          }
        }
      }
    }
    __syncwarp();
    grid_dependency_wait();
    return;
  }

  if (warp == kIConsumers + 1) {
    // ===================== epilogue warp =====================
    const int g = lane >> 2, c = lane & 3;
    grid_dependency_wait();
    named_bar_sync(2, kIThreads);
    for (int strip = s0; strip < s1; ++strip) {
      const int li = strip - s0;
      const int rbuf = li % kIRedBufs;
      const uint32_t use = (uint32_t)(li / kIRedBufs);
      const float* rb = red + (size_t)rbuf * (kIConsumers * 128);
      mbar_wait(&red_full[rbuf], use & 1u);
      // 16 warps' partial sums, added in warp order with packed fp32x2 adds (half the instructions of the scalar chain)
      uint64_t vlo = 0ull, vhi = 0ull;
#pragma unroll
      for (int w = 0; w < kIConsumers; ++w) {
        const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(rb + (w * 32 + lane) * 4);
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vlo) : "l"(t.x));
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vhi) : "l"(t.y));
      }
      const float4 v = make_float4(__uint_as_float((uint32_t)vlo), __uint_as_float((uint32_t)(vlo >> 32)),
                                   __uint_as_float((uint32_t)vhi), __uint_as_float((uint32_t)(vhi >> 32)));
      __syncwarp();
      if (lane == 0) mbar_arrive(&red_free[rbuf]);
      // limb recombination: columns (3 m, 3 m + 1, 3 m + 2) = limbs (2^16, 2^8, 2^0) of activation row m; lane (g, c)
      // holds columns 2 c, 2 c + 1 of weight rows g (x, y) and g + 8 (z, w)
      float4 t;
      t.x = __shfl_down_sync(0xffffffffu, v.x, 1);
      t.y = __shfl_down_sync(0xffffffffu, v.y, 1);
      t.z = __shfl_down_sync(0xffffffffu, v.z, 1);
      t.w = __shfl_down_sync(0xffffffffu, v.w, 1);
      const int m = c;
      float r0, r1;  // lanes c = 0, 1: rows g, g + 8 of activation row c (other lanes: don't care)
      if (m == 0) {
        r0 = fmaf(65536.f, v.x, fmaf(256.f, v.y, t.x));
        r1 = fmaf(65536.f, v.z, fmaf(256.f, v.w, t.z));
      } else {
        r0 = fmaf(65536.f, v.y, fmaf(256.f, t.x, t.y));
        r1 = fmaf(65536.f, v.w, fmaf(256.f, t.z, t.w));
      }
      if (m >= p.M || m > 1) continue;
      const float rsc = row_scale[m];
      const float v0 = r0 * rsc, v1 = r1 * rsc;
      if (p.epilogue == 2) {
        const int64_t n = 8 * (int64_t)strip + g;  // rows g = gate (w1), rows g + 8 = up (w3)  (model.py:486)
        if (2 * n < p.N_out) {
          const float gate = bf16r(v0), up = bf16r(v1);
          const float sg = bf16r(gate / (1.f + __expf(-gate)));
          p.y[m * p.ldy + n] = __float2bfloat16_rn(sg * up);
        }
      } else {
        const int64_t n0 = 16 * (int64_t)strip + g;
        if (p.epilogue == 1) {
          if (n0 < p.N_out) p.y[m * p.ldy + n0] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0]) + bf16r(v0));
          if (n0 + 8 < p.N_out)
            p.y[m * p.ldy + n0 + 8] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0 + 8]) + bf16r(v1));
        } else {
          if (n0 < p.N_out) p.y[m * p.ldy + n0] = __float2bfloat16_rn(v0);
          if (n0 + 8 < p.N_out) p.y[m * p.ldy + n0 + 8] = __float2bfloat16_rn(v1);
        }
      }
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, c = lane & 3;
  const int tid = threadIdx.x;  // 0 .. 511
  constexpr int kCT = kIConsumers * 32;
  constexpr int kHold = NH > 0 ? NH : 1;  // chunks a thread keeps in registers between the two staging passes
  const int chunks_per_row = KT * 8;  // 8-element chunks over Kp
  const int n_iter = (chunks_per_row + kCT - 1) / kCT;
  // Fast path (M = 1, Kp <= 16384): every thread owns <= 4 chunks; all of their loads are issued at once and the
  // normalised bf16 values stay in registers between the two passes.  The norm weights do not depend on the previous
  // kernel, so they are fetched BEFORE the dependency wait (they are a DRAM miss: read once per token).
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

  // ---- stage the activations.  Pass 1: sum of squares and max magnitude of x (* norm_w).  Pass 2: fixed point,
  //      limbs, per-group limb sums (general path: x re-read — it is L1 / L2 resident and nothing writes it while this
  //      kernel runs). ----
  auto load_norm = [&](int m, int64_t k0, float (&f)[8], float& ss) {
    unpack8(*reinterpret_cast<const uint4*>(p.x + (int64_t)m * p.ldx + k0), f);
    if (p.norm_w != nullptr) {
      float wn[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(p.norm_w + k0)), wn);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        ss = fmaf(f[e], f[e], ss);
        f[e] = bf16r(f[e] * wn[e]);
      }
    }
  };
  {
    if constexpr (hold) {
      float ss = 0.f;
      uint32_t amax = 0u;  // bits of the largest |x| (a NaN has the largest bits: it propagates)
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
          xr[i] = make_uint4(mul_bf16x2(xr[i].x, wn4[i].x), mul_bf16x2(xr[i].y, wn4[i].y), mul_bf16x2(xr[i].z, wn4[i].z),
                             mul_bf16x2(xr[i].w, wn4[i].w));
        }
        unpack8(xr[i], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) amax = absmax_bits(amax, f[e]);
      }
      ss = warp_sum_i(ss);
      amax = warp_max_bits(amax);
      if (lane == 0) {
        part[warp * 2] = ss;
        part[warp * 2 + 1] = __uint_as_float(amax);
      }
    } else {
      for (int m = 0; m < p.M; ++m) {
        float ss = 0.f;
        uint32_t amax = 0u;
        for (int ci = tid; ci < chunks_per_row; ci += kCT) {
          const int64_t k0 = (int64_t)ci * 8;
          if (k0 >= p.Kx) continue;
          float f[8];
          load_norm(m, k0, f, ss);
#pragma unroll
          for (int e = 0; e < 8; ++e) amax = absmax_bits(amax, f[e]);
        }
        ss = warp_sum_i(ss);
        amax = warp_max_bits(amax);
        if (lane == 0) {
          part[(m * kIConsumers + warp) * 2] = ss;
          part[(m * kIConsumers + warp) * 2 + 1] = __uint_as_float(amax);
        }
      }
    }
    named_bar_sync(1, kCT);
    // every warp reduces the 16 partials of every row itself (no second hand-over through shared memory)
    float up_m[2] = {1.f, 1.f};
    constexpr int kRows = hold ? 1 : 2;  // the register-held path is M = 1
#pragma unroll
    for (int m = 0; m < kRows; ++m) {
      float t = 0.f;
      uint32_t amax_b = 0u;
      if (m < p.M && lane < kIConsumers) {
        t = part[(m * kIConsumers + lane) * 2];
        amax_b = __float_as_uint(part[(m * kIConsumers + lane) * 2 + 1]);
      }
      t = warp_sum_i(t);
      amax_b = warp_max_bits(amax_b);
      const bool bad = amax_b >= 0x7f800000u;  // an Inf / NaN activation: the reference's bf16 matmul propagates it, so do we
      const float amax = bad ? 0.f : __uint_as_float(amax_b);
      // 2^e brings the row's largest magnitude into [2^21, 2^22): |round(x 2^e)| <= 2^22 fits three balanced int8 limbs
      int e = 0;
      if (amax > 0.f) e = 21 - ((int)((__float_as_uint(amax) >> 23) & 0xFF) - 127);
      e = max(-106, min(100, e));  // -106 covers the largest finite bf16; above 100 (amax < 2^-79) the values just get smaller
      up_m[m] = __uint_as_float((uint32_t)(e + 127) << 23);
      if (warp == m && lane == 0 && m < p.M) {
        const float dn = __uint_as_float((uint32_t)(127 - e) << 23);
        const float rs = (p.norm_w != nullptr) ? dn * rsqrtf(t / (float)p.Kx + p.eps) : dn;
        row_scale[m] = bad ? __uint_as_float(0x7fc00000u) : rs;  // read by the epilogue warp (after barrier 2)
      }
    }
    // one chunk (8 activations f[] of row m at chunk index ci) -> limbs in shared memory + group sums.  Executed by whole
    // warps in uniform control flow (the group sums are shuffles); `valid` = the chunk exists.
    auto emit = [&](int m, int ci, bool valid, const float (&f)[8], float up) {
      uint8_t* rows = x_s + (size_t)(3 * m) * p.limb_row_bytes;
      int sl0 = 0, sl1 = 0, sl2 = 0;
      if (valid) {
        // Balanced limbs without per-digit carries: X + 0x808080 = sum (a_i + 128) 256^i, so its bytes xor 0x80 ARE the
        // two's-complement limbs a_i.  x * 2^e + 1.5 * 2^23 puts X + 2^22 into the float's mantissa (|X| <= 2^22 - 2^14), one
        // integer add turns that into X + 0x808080, byte permutes gather limb i of the 8 elements in the
        // order of the weight bytes after the nibble split: (k0, k0+4, k0+1, k0+5 | k0+2, k0+6, k0+3, k0+7).
        uint32_t v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          // no clamp needed: |f| <= amax < 2^(22 - e) - 2^(14 - e) (amax is a bf16), so |X| <= 2^22 - 2^14
          v[e] = __float_as_uint(fmaf(f[e], up, kIMagicF)) + 0x00408080u;  // low 24 bits = X + 0x808080
        }
        uint32_t lw[3][2];
#pragma unroll
        for (int l = 0; l < 3; ++l) {
          // limb l lives in byte (2 - l) of every v: selectors pick that byte of the two sources
          const uint32_t sel = l == 0 ? 0x0062u : (l == 1 ? 0x0051u : 0x0040u);
          const uint32_t p04 = __byte_perm(v[0], v[4], sel), p15 = __byte_perm(v[1], v[5], sel);   // (k0, k0+4), (k0+1, k0+5)
          const uint32_t p26 = __byte_perm(v[2], v[6], sel), p37 = __byte_perm(v[3], v[7], sel);
          lw[l][0] = __byte_perm(p04, p15, 0x5410) ^ 0x80808080u;
          lw[l][1] = __byte_perm(p26, p37, 0x5410) ^ 0x80808080u;
        }
        sl0 = __dp4a((int)lw[0][0], 0x01010101, __dp4a((int)lw[0][1], 0x01010101, 0));
        sl1 = __dp4a((int)lw[1][0], 0x01010101, __dp4a((int)lw[1][1], 0x01010101, 0));
        sl2 = __dp4a((int)lw[2][0], 0x01010101, __dp4a((int)lw[2][1], 0x01010101, 0));
        *reinterpret_cast<uint2*>(rows + (size_t)ci * 8) = make_uint2(lw[0][0], lw[0][1]);
        *reinterpret_cast<uint2*>(rows + p.limb_row_bytes + (size_t)ci * 8) = make_uint2(lw[1][0], lw[1][1]);
        *reinterpret_cast<uint2*>(rows + 2 * p.limb_row_bytes + (size_t)ci * 8) = make_uint2(lw[2][0], lw[2][1]);
      }
      // a segment = kSegLanes consecutive chunks = consecutive lanes (kCT % 16 == 0)
#pragma unroll
      for (int o = GE::kSegLanes / 2; o; o >>= 1) {
        sl0 += __shfl_xor_sync(0xffffffffu, sl0, o);
        sl1 += __shfl_xor_sync(0xffffffffu, sl1, o);
        sl2 += __shfl_xor_sync(0xffffffffu, sl2, o);
      }
      if (valid && (ci % GE::kSegLanes) == 0) {
        float* S = S_s + (ci / GE::kSegLanes) * 8 + 3 * m;
        S[0] = (float)sl0; S[1] = (float)sl1; S[2] = (float)sl2;
      }
    };
    if constexpr (hold) {
#pragma unroll
      for (int it = 0; it < kHold; ++it) {
        if (it < n_iter) {
          const int ci = it * kCT + tid;
          float f[8];
          unpack8(xr[it], f);  // zero beyond Kx
          emit(0, ci, ci < chunks_per_row, f, up_m[0]);
        }
      }
    } else {
      for (int m = 0; m < p.M; ++m) {
        for (int it = 0; it < n_iter; ++it) {
          const int ci = it * kCT + tid;
          const int64_t k0 = (int64_t)ci * 8;
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = 0.f;
          if (ci < chunks_per_row && k0 < p.Kx) {
            float dummy = 0.f;
            load_norm(m, k0, f, dummy);
          }
          emit(m, ci, ci < chunks_per_row, f, m == 0 ? up_m[0] : up_m[1]);
        }
      }
    }
  }
  // one barrier closes the staging (limbs, group sums, row scale visible to every consumer and to the epilogue warp) and
  // meets the producer, which initialised the mbarriers long ago
  named_bar_sync(2, kIThreads);

  // ---- main loop: warp w owns k-tiles 2 w, 2 w + 1 (128 k) of every stage ----
  const uint32_t ring_u = smem_u32(ring);
  uint32_t w_off = (uint32_t)warp * 1024u + (uint32_t)lane * 16u;
  uint32_t sz_off = (uint32_t)kIStageW + (uint32_t)(G >= 128 ? warp * 128 / G : warp * GE::kNSeg) * 64u + (uint32_t)g * 4u;
  uint32_t S_u0 = smem_u32(S_s) + (uint32_t)(warp * GE::kNSeg) * 32u + (uint32_t)c * 8u;
  const bool x_on = g < 3 * p.M;  // MMA column g = (activation row g / 3, limb g % 3)
  uint32_t x_u0 = smem_u32(x_s) + (uint32_t)min(g, 3 * p.M - 1) * p.limb_row_bytes + (uint32_t)warp * 128u + (uint32_t)c * 8u;
  asm volatile("" : "+r"(w_off), "+r"(sz_off), "+r"(S_u0), "+r"(x_u0));
  int slot = 0;
  uint32_t par = 0;
  uint32_t stage_u = ring_u;

  for (int strip = s0; strip < s1; ++strip) {
    float yacc[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t xo = 0, S_u = S_u0;
    for (int kt0 = 0; kt0 < KT; kt0 += kIStageTiles) {
      mbar_wait(&full[slot], par);
      if (2 * warp < KT - kt0) {
        const uint4 wa = lds128(stage_u + w_off), wb = lds128(stage_u + w_off + 512u);
        uint2 xb[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) xb[h] = x_on ? lds64u(x_u0 + xo + 32u * h) : make_uint2(0u, 0u);
        constexpr uint32_t kM = 0x0f0f0f0fu;
        // half h: words (row g, row g + 8); low nibbles = k (k0, k0+4, k0+1, k0+5), high nibbles = (k0+2, k0+6, k0+3, k0+7)
        const uint32_t wl[4][2] = {{wa.x, wa.y}, {wa.z, wa.w}, {wb.x, wb.y}, {wb.z, wb.w}};
        if constexpr (GE::kNSeg == 1) {
          const uint32_t p0 = lds32(stage_u + sz_off), p1 = lds32(stage_u + sz_off + 32u);
          const float2 S = lds64f(S_u);
          int a0[4] = {(int)kIMagicI, (int)kIMagicI, (int)kIMagicI, (int)kIMagicI}, a1[4] = {0, 0, 0, 0};
          imma_16832(a0, wl[0][0] & kM, wl[0][1] & kM, (wl[0][0] >> 4) & kM, (wl[0][1] >> 4) & kM, xb[0].x, xb[0].y);
          imma_16832(a1, wl[1][0] & kM, wl[1][1] & kM, (wl[1][0] >> 4) & kM, (wl[1][1] >> 4) & kM, xb[1].x, xb[1].y);
          imma_16832(a0, wl[2][0] & kM, wl[2][1] & kM, (wl[2][0] >> 4) & kM, (wl[2][1] >> 4) & kM, xb[2].x, xb[2].y);
          imma_16832(a1, wl[3][0] & kM, wl[3][1] & kM, (wl[3][0] >> 4) & kM, (wl[3][1] >> 4) & kM, xb[3].x, xb[3].y);
          const float sc0 = __uint_as_float(p0 << 16), sc1 = __uint_as_float(p1 << 16);
          const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0 & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1 & 0xFFFF0000u));
          const float f0 = __int_as_float(a0[0] + a1[0]) - kIMagicF, f1 = __int_as_float(a0[1] + a1[1]) - kIMagicF;
          const float f2 = __int_as_float(a0[2] + a1[2]) - kIMagicF, f3 = __int_as_float(a0[3] + a1[3]) - kIMagicF;
          yacc[0] = fmaf(sc0, f0, fmaf(zz0, S.x, yacc[0]));
          yacc[1] = fmaf(sc0, f1, fmaf(zz0, S.y, yacc[1]));
          yacc[2] = fmaf(sc1, f2, fmaf(zz1, S.x, yacc[2]));
          yacc[3] = fmaf(sc1, f3, fmaf(zz1, S.y, yacc[3]));
        } else {
          constexpr int kHS = 4 / GE::kNSeg;  // 32-k halves per segment
#pragma unroll
          for (int sg = 0; sg < GE::kNSeg; ++sg) {
            const uint32_t p0 = lds32(stage_u + sz_off + 64u * sg), p1 = lds32(stage_u + sz_off + 64u * sg + 32u);
            const float2 S = lds64f(S_u + 32u * sg);
            int a[4] = {(int)kIMagicI, (int)kIMagicI, (int)kIMagicI, (int)kIMagicI};
#pragma unroll
            for (int hh = 0; hh < kHS; ++hh) {
              const int h = sg * kHS + hh;
              imma_16832(a, wl[h][0] & kM, wl[h][1] & kM, (wl[h][0] >> 4) & kM, (wl[h][1] >> 4) & kM, xb[h].x, xb[h].y);
            }
            const float sc0 = __uint_as_float(p0 << 16), sc1 = __uint_as_float(p1 << 16);
            const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0 & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1 & 0xFFFF0000u));
            yacc[0] = fmaf(sc0, __int_as_float(a[0]) - kIMagicF, fmaf(zz0, S.x, yacc[0]));
            yacc[1] = fmaf(sc0, __int_as_float(a[1]) - kIMagicF, fmaf(zz0, S.y, yacc[1]));
            yacc[2] = fmaf(sc1, __int_as_float(a[2]) - kIMagicF, fmaf(zz1, S.x, yacc[2]));
            yacc[3] = fmaf(sc1, __int_as_float(a[3]) - kIMagicF, fmaf(zz1, S.y, yacc[3]));
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      xo += kIStageTiles * 64u;
      S_u += (kIStageTiles * 64 / GE::kSegK) * 32u;
      stage_u += kIStageBytes;
      if (++slot == stages) {
        slot = 0;
        par ^= 1u;
        stage_u = ring_u;
      }
    }

    const int li = strip - s0;
    const int rbuf = li % kIRedBufs;
    const uint32_t use = (uint32_t)(li / kIRedBufs);
    float* rb = red + (size_t)rbuf * (kIConsumers * 128);
    if (use > 0) mbar_wait(&red_free[rbuf], (use - 1) & 1u);
    *reinterpret_cast<float4*>(rb + (warp * 32 + lane) * 4) = make_float4(yacc[0], yacc[1], yacc[2], yacc[3]);
    __syncwarp();
    if (lane == 0) mbar_arrive(&red_full[rbuf]);
  }
}

// ------------------------------------------------------------------------------------------------ host side
struct PlanI {
  int stages;
  size_t smem;
  uint32_t off_x, off_S, off_red, off_misc, limb_row_bytes;
};

static bool plan_imma(int M, int64_t Kp, int G, int max_stages, PlanI& pl) {
  const int seg_k = G < 128 ? G : 128;
  const size_t kIStageBytes = (size_t)kIStageW + (size_t)(kIStageTiles * 64 / G) * 64;
  pl.limb_row_bytes = (uint32_t)(Kp + 32);
  const size_t x_bytes = ((size_t)3 * M * pl.limb_row_bytes + 127) / 128 * 128;
  const size_t S_bytes = ((size_t)(Kp / seg_k) * 32 + 127) / 128 * 128;
  const size_t red_bytes = (size_t)kIRedBufs * kIConsumers * 128 * 4;
  const size_t misc = (2 * kIConsumers * 2 + 4) * 4 + (2 * 16 + 2 * kIRedBufs) * 8 + 64;
  const size_t fixed = x_bytes + S_bytes + red_bytes + misc;
  const size_t two_per_sm = 113 * 1024, one_per_sm = 226 * 1024;
  int st = 0;
  if (fixed + 4 * kIStageBytes <= two_per_sm) {
    st = (int)((two_per_sm - fixed) / kIStageBytes);
  } else if (fixed + 3 * kIStageBytes <= one_per_sm) {
    st = (int)((one_per_sm - fixed) / kIStageBytes);
  } else {
    return false;
  }
  if (st > max_stages) st = max_stages;
  if (st > 16) st = 16;
  pl.stages = st;
  size_t off = (size_t)st * kIStageBytes;
  pl.off_x = (uint32_t)off; off += x_bytes;
  pl.off_S = (uint32_t)off; off += S_bytes;
  pl.off_red = (uint32_t)off; off += red_bytes;
  pl.off_misc = (uint32_t)off; off += misc;
  pl.smem = off;
  return true;
}

// M <= 2, G in {32, 64, 128, 256}.  Returns AO_B200_ERR_UNSUPPORTED / AO_B200_ERR_WORKSPACE when the shape is not for this kernel.
int int4_gemv_imma(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                   int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps,
                   const void* residual, int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s) {
  if (M > 2 || (G != 32 && G != 64 && G != 128 && G != 256) || Kp % (G > 128 ? G : 128) != 0) return AO_B200_ERR_UNSUPPORTED;
  (void)Np;
  GemvIParams p;
  p.packed = static_cast<const uint8_t*>(packed);
  p.sz = static_cast<const uint32_t*>(sz);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.norm_w = static_cast<const __nv_bfloat16*>(norm_w);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.epilogue = epilogue;
  {
    static const int hold = []() { const char* e = std::getenv("AO_B200_IMMA_HOLD"); return e ? std::atoi(e) : 1; }();
    p.hold = hold;
  }
  p.KT = (int)(Kp / 64);
  p.Kx = Kx; p.ldx = ldx; p.N_out = N_out; p.ldy = ldy; p.ld_res = ld_res;
  p.n_strips = (int)((N_out + 15) / 16);
  p.eps = eps;
  PlanI pl;
  if (!plan_imma((int)M, Kp, G, max_stages, pl)) return AO_B200_ERR_WORKSPACE;
  p.stages = pl.stages;
  p.limb_row_bytes = pl.limb_row_bytes;
  p.off_x = pl.off_x; p.off_S = pl.off_S; p.off_red = pl.off_red; p.off_misc = pl.off_misc;
  using KernelFn = void (*)(const GemvIParams);
  static const KernelFn kernels[4][4] = {
      {int4_gemv_imma_kernel<32, 0>, int4_gemv_imma_kernel<32, 1>, int4_gemv_imma_kernel<32, 2>, int4_gemv_imma_kernel<32, 4>},
      {int4_gemv_imma_kernel<64, 0>, int4_gemv_imma_kernel<64, 1>, int4_gemv_imma_kernel<64, 2>, int4_gemv_imma_kernel<64, 4>},
      {int4_gemv_imma_kernel<128, 0>, int4_gemv_imma_kernel<128, 1>, int4_gemv_imma_kernel<128, 2>, int4_gemv_imma_kernel<128, 4>},
      {int4_gemv_imma_kernel<256, 0>, int4_gemv_imma_kernel<256, 1>, int4_gemv_imma_kernel<256, 2>, int4_gemv_imma_kernel<256, 4>}};
  const int gi = G == 32 ? 0 : (G == 64 ? 1 : (G == 128 ? 2 : 3));
  // register-held staging: M = 1 and <= 4 chunks of 8 activations per consumer thread (Kp <= 16384)
  const int n_iter = (p.KT * 8 + kIConsumers * 32 - 1) / (kIConsumers * 32);
  const int hi = (p.hold != 0 && M == 1 && n_iter <= 4) ? (n_iter <= 1 ? 1 : (n_iter <= 2 ? 2 : 3)) : 0;
  const KernelFn kern = kernels[gi][hi];
  static bool attr_done[4][4] = {};
  if (!attr_done[gi][hi]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done[gi][hi] = true;
  }
  static int sms = 0;
  if (sms == 0) {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    sms = v > 0 ? v : kNumSMs;
  }
  // AO_B200_IMMA_CTAS: CTAs of ONE launch per SM.  2 (default) = more warps per launch; 1 leaves the second slot to the NEXT
  // kernel of the PDL chain, whose producer then prefetches its weights while this launch is still running.
  static const int ctas = []() { const char* e = std::getenv("AO_B200_IMMA_CTAS"); const int v = e ? std::atoi(e) : 2; return v < 1 ? 1 : (v > 2 ? 2 : v); }();
  static const int one_below = []() { const char* e = std::getenv("AO_B200_IMMA_CTAS1_BELOW"); return e ? std::atoi(e) : 0; }();
  const int max_grid = sms * (pl.smem <= 113 * 1024 ? (p.n_strips <= one_below ? 1 : ctas) : 1);
  int grid = p.n_strips < max_grid ? p.n_strips : max_grid;
  {
    // balanced waves (off by default: slower in the decode step): every CTA gets the same number of strips
    static const bool balance = []() { const char* e = std::getenv("AO_B200_BALANCE"); return e ? std::atoi(e) != 0 : false; }();
    const int per = (p.n_strips + grid - 1) / grid;
    if (balance) grid = (p.n_strips + per - 1) / per;
  }
  // (Splitting strips along K between CTAs — 256-k units, parked partial sums + tickets — was tried and removed: wherever
  // it applied the launch got 3 - 4 us slower, more than the balance won; profiles/r01c_imma_splitk_ab.txt.)
  p.strips_base = p.n_strips / grid;
  p.strips_rem = p.n_strips % grid;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kIThreads);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, kern, p) != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

}  // namespace ao
