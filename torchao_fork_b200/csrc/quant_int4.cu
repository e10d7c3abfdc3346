// int4 group-wise quantize / pack / unpack / dequant kernels (bit-exact with the reference's
// eager bf16 arithmetic; see oracle/ao_oracle.py for the op-by-op restatement).
#include "common.cuh"

namespace ao {

// qparams for one (row, group) from its min / max.  Every reference op boundary
// (quant_primitives.py:1294-1300) is a bf16 rounding.
struct QParams {
  float scale, zero, min_val;
};

__device__ __forceinline__ QParams tinygemm_qparams(float mn, float mx, float eps, int div_mode) {
  QParams p;
  const float d = bf16r(mx - mn);
  float s = (div_mode == AO_B200_DIV_RECIP) ? bf16r(d * (1.0f / 15.0f)) : bf16r(__fdiv_rn(d, 15.0f));
  s = bf16r(fmaxf(s, eps));
  const float t8 = bf16r(s * 8.0f);
  p.scale = s;
  p.zero = bf16r(mn + t8);
  p.min_val = bf16r(p.zero - t8);  // quant_primitives.py:569 recomputes it from the rounded zero
  return p;
}

__device__ __forceinline__ uint32_t tinygemm_quant1(float w, float min_val, float scale) {
  const float a = bf16r(w - min_val);
  const float b = bf16r(__fdiv_rn(a, scale));
  const float r = fminf(fmaxf(rintf(b), 0.0f), 15.0f);
  return static_cast<uint32_t>(r);
}

// ---- fast exact path of the fused kernel -------------------------------------------------------------------------
// The per-element work of _quantize_affine_tinygemm is  q = clamp(rint(bf16(bf16(w - min_val) / scale)), 0, 15).  Done
// naively it is ~22 instructions per weight, four of them on the quarter-rate conversion pipe (two bf16 roundings,
// rint, float->int) plus an IEEE division: the kernel ran at 13-20 % of the HBM roofline, instruction-bound.  Here:
//   * min / max and the subtraction stay in packed bf16x2 (HMNMX2 / HADD2.BF16 round exactly like the reference's
//     fp32-compute-then-round, and take one instruction per TWO weights);
//   * the division is a * (1/s) + one Newton step with a reciprocal computed once per group.  Both operands carry 8
//     significant bits, so the true quotient is either exactly representable in fp32 (then the residual is exact and
//     the step returns it exactly) or at least 2^-18 (relative) away from any bf16 rounding boundary, while the step's
//     error is < 2^-22: bf16(quotient) is the reference's value in every case;
//   * rint is an add of 1.5 * 2^23 (round-to-nearest-even in the FMA pipe), the nibbles are assembled with FMAs
//     (q * 16^slot, exact below 2^16) and converted to integer once per four weights.
struct QFast {
  float s, rs;        // scale and its reciprocal
  uint32_t minv2;     // (min_val, min_val) as bf16x2
};
__device__ __forceinline__ uint32_t bf2_min(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("min.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t bf2_max(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t bf2_sub(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ float fast_quot(float a, const QFast& p) {
  const float x0 = a * p.rs;
  const float r = fmaf(-p.s, x0, a);
  return fmaf(r, p.rs, x0);
}
// 8 weights (4 bf16x2 registers: element pairs (0,1) (2,3) (4,5) (6,7)) -> one packed word, nibble slots
// [k0,k0+2,k0+4,k0+6,k0+1,k0+3,k0+5,k0+7]: even elements fill the low 16 bits, odd elements the high 16 bits
__device__ __forceinline__ uint32_t quant_word_fast(const uint4& raw, const QFast& p) {
  const uint32_t w2[4] = {raw.x, raw.y, raw.z, raw.w};
  constexpr float kMagic = 12582912.0f;  // 1.5 * 2^23
  constexpr uint32_t k15 = 0x41704170u;  // (15.0, 15.0) bf16x2
  float lo = 0.f, hi = 0.f, wgt = 1.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t a2 = bf2_sub(w2[j], p.minv2);                                  // bf16(w - min_val)
    const float q0 = fast_quot(__uint_as_float(a2 << 16), p), q1 = fast_quot(__uint_as_float(a2 & 0xFFFF0000u), p);
    uint32_t t2;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(t2) : "f"(q1), "f"(q0));            // bf16(a / scale)
    t2 = bf2_max(bf2_min(t2, k15), 0u);                                           // clamp commutes with rint at 0 and 15
    const float r0 = (__uint_as_float(t2 << 16) + kMagic) - kMagic;               // rint, ties to even
    const float r1 = (__uint_as_float(t2 & 0xFFFF0000u) + kMagic) - kMagic;
    lo = fmaf(r0, wgt, lo);
    hi = fmaf(r1, wgt, hi);
    wgt *= 16.f;
  }
  return static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
}

// ---------------------------------------------------------------------------------------
// Fused quantize + pack into the super-tile layout (include/ao_b200.h).  One warp owns a unit = 16 rows x UK columns, UK = max(G, 64):
// it loads the unit's tiles in the strip layout's own lane order (so the packed words fall
// out of registers already in place), reduces min/max over the 4 lanes that share a row,
// and writes 512-byte tiles with one coalesced 16-byte store per lane.
// ---------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) int4_quantize_pack_kernel(
    const __nv_bfloat16* __restrict__ w, int64_t N, int64_t K, int64_t ldw, int64_t Np, int64_t Kp,
    float eps, int div_mode, uint4* __restrict__ packed, uint32_t* __restrict__ sz) {
  constexpr int UK = G > 64 ? G : 64;  // columns per unit
  constexpr int T = UK / 64;           // tiles per unit
  constexpr int NG = UK / G;           // groups per unit (2 for G=32, else 1)
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  const int64_t units_k = Kp / UK;
  const int64_t num_units = (Np / 16) * units_k;
  const int64_t warp_global = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t warp_stride = (int64_t)gridDim.x * (blockDim.x >> 5);

  for (int64_t unit = warp_global; unit < num_units; unit += warp_stride) {
    const int64_t strip = unit / units_k;
    const int64_t uk = unit % units_k;
    uint4 raw[T][4];
#pragma unroll
    for (int t = 0; t < T; ++t) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t row = strip * 16 + g + 8 * (j & 1);
        const int64_t k0 = uk * UK + t * 64 + 32 * (j >> 1) + 8 * c;
        raw[t][j] = (row < N && k0 < K) ? ldg128(w + row * ldw + k0) : make_uint4(0, 0, 0, 0);
      }
    }
    // min / max per (row slot r = j&1, group gi), in packed bf16x2 (exact: comparisons only)
    uint32_t mn2[2][NG], mx2[2][NG];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int gi = 0; gi < NG; ++gi) {
        mn2[r][gi] = 0x7F807F80u;  // (+inf, +inf)
        mx2[r][gi] = 0xFF80FF80u;  // (-inf, -inf)
      }
#pragma unroll
    for (int t = 0; t < T; ++t)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gi = (NG == 2) ? (j >> 1) : 0;
        const uint32_t v[4] = {raw[t][j].x, raw[t][j].y, raw[t][j].z, raw[t][j].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          mn2[j & 1][gi] = bf2_min(mn2[j & 1][gi], v[i]);
          mx2[j & 1][gi] = bf2_max(mx2[j & 1][gi], v[i]);
        }
      }
    QParams qp[2][NG];
    QFast qf[2][NG];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int gi = 0; gi < NG; ++gi) {
        uint32_t a2 = mn2[r][gi], b2 = mx2[r][gi];
        a2 = bf2_min(a2, __shfl_xor_sync(0xffffffffu, a2, 1));
        a2 = bf2_min(a2, __shfl_xor_sync(0xffffffffu, a2, 2));
        b2 = bf2_max(b2, __shfl_xor_sync(0xffffffffu, b2, 1));
        b2 = bf2_max(b2, __shfl_xor_sync(0xffffffffu, b2, 2));
        const float a = fminf(__uint_as_float(a2 << 16), __uint_as_float(a2 & 0xFFFF0000u));
        const float b = fmaxf(__uint_as_float(b2 << 16), __uint_as_float(b2 & 0xFFFF0000u));
        qp[r][gi] = tinygemm_qparams(a, b, eps, div_mode);
        qf[r][gi].s = qp[r][gi].scale;
        qf[r][gi].rs = __frcp_rn(qp[r][gi].scale);
        const uint32_t mb = float_to_bf16_bits(qp[r][gi].min_val);
        qf[r][gi].minv2 = mb | (mb << 16);
      }
    // quantize + pack
#pragma unroll
    for (int t = 0; t < T; ++t) {
      uint32_t words[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) words[j] = quant_word_fast(raw[t][j], qf[j & 1][(NG == 2) ? (j >> 1) : 0]);
      const int64_t kt = uk * T + t;
      packed[tile_slot(strip, kt, Kp / 64) * 32 + lane] = make_uint4(words[0], words[1], words[2], words[3]);
    }
    if (c == 0) {
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int gi = 0; gi < NG; ++gi) {
          const int64_t row = strip * 16 + g + 8 * r;
          const int64_t grp = uk * NG + gi;
          sz[sz_index(grp, row, Kp / G)] = float_to_bf16_bits(qp[r][gi].scale) | (float_to_bf16_bits(qp[r][gi].zero) << 16);
        }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Stand-alone steps (API parity with choose_qparams / quantize; one warp per (row, group)).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) int4_choose_qparams_kernel(
    const __nv_bfloat16* __restrict__ w, int64_t N, int64_t K, int64_t ldw, int G, float eps, int div_mode,
    __nv_bfloat16* __restrict__ scale, __nv_bfloat16* __restrict__ zero) {
  const int lane = threadIdx.x & 31;
  const int64_t groups = K / G;
  const int64_t total = N * groups;
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t u = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); u < total; u += stride) {
    const int64_t row = u / groups, grp = u % groups;
    const __nv_bfloat16* p = w + row * ldw + grp * G;
    float mn = INFINITY, mx = -INFINITY;
    for (int i = lane; i < G; i += 32) {
      const float v = __bfloat162float(p[i]);
      mn = fminf(mn, v);
      mx = fmaxf(mx, v);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (lane == 0) {
      const QParams q = tinygemm_qparams(mn, mx, eps, div_mode);
      scale[u] = __float2bfloat16_rn(q.scale);
      zero[u] = __float2bfloat16_rn(q.zero);
    }
  }
}

__global__ void __launch_bounds__(256) int4_quantize_kernel(
    const __nv_bfloat16* __restrict__ w, int64_t N, int64_t K, int64_t ldw, int G,
    const __nv_bfloat16* __restrict__ scale, const __nv_bfloat16* __restrict__ zero, int32_t* __restrict__ q) {
  const int64_t total = N * K;
  const int64_t groups = K / G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / K, k = i % K;
    const int64_t u = row * groups + k / G;
    const float s = __bfloat162float(scale[u]);
    const float z = __bfloat162float(zero[u]);
    const float min_val = bf16r(z - bf16r(s * 8.0f));
    q[i] = static_cast<int32_t>(tinygemm_quant1(__bfloat162float(w[row * ldw + k]), min_val, s));
  }
}

// ---------------------------------------------------------------------------------------
// strip layout <-> plain
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) int4_pack_kernel(const int32_t* __restrict__ q, int64_t N, int64_t K,
                                                        int64_t Np, int64_t Kp, int pad_value,
                                                        uint4* __restrict__ packed) {
  const int64_t total = (Np / 16) * (Kp / 64) * 32;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int lane = i & 31;
    int64_t strip, kt;
    tile_of_slot(i >> 5, Kp / 64, strip, kt);
    const int g = lane >> 2, c = lane & 3;
    uint32_t words[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t row = strip * 16 + g + 8 * (j & 1);
      const int64_t k0 = kt * 64 + 32 * (j >> 1) + 8 * c;
      uint32_t word = 0;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int v = (row < N && k0 + e < K) ? q[row * K + k0 + e] : pad_value;
        word |= (static_cast<uint32_t>(v) & 0xFu) << (4 * nib_slot(e));
      }
      words[j] = word;
    }
    packed[i] = make_uint4(words[0], words[1], words[2], words[3]);
  }
}

template <bool kDequant>
__global__ void __launch_bounds__(256) int4_unpack_kernel(const uint4* __restrict__ packed,
                                                          const uint32_t* __restrict__ sz, int64_t Np, int64_t Kp,
                                                          int G, int32_t* __restrict__ q,
                                                          __nv_bfloat16* __restrict__ out) {
  const int64_t total = (Np / 16) * (Kp / 64) * 32;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int lane = i & 31;
    int64_t strip, kt;
    tile_of_slot(i >> 5, Kp / 64, strip, kt);
    const int g = lane >> 2, c = lane & 3;
    const uint4 v = packed[i];
    const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t row = strip * 16 + g + 8 * (j & 1);
      const int64_t k0 = kt * 64 + 32 * (j >> 1) + 8 * c;
      if constexpr (kDequant) {
        const uint32_t p = sz[sz_index(k0 / G, row, Kp / G)];
        const uint32_t s2 = __byte_perm(p, p, 0x1010), z2 = __byte_perm(p, p, 0x3232);
        uint32_t d[4];
        dequant_word(words[j], s2, z2, d);
        *reinterpret_cast<uint4*>(out + row * Kp + k0) = make_uint4(d[0], d[1], d[2], d[3]);
      } else {
        int32_t e[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) e[t] = (words[j] >> (4 * nib_slot(t))) & 0xF;
        int4* dst = reinterpret_cast<int4*>(q + row * Kp + k0);
        dst[0] = make_int4(e[0], e[1], e[2], e[3]);
        dst[1] = make_int4(e[4], e[5], e[6], e[7]);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// tinygemm layout (aten._convert_weight_to_int4pack) <-> plain.  One thread per packed word.
// word (nTile, kOuter, t, j): row = 8 nTile + t/4; kb = (kOuter*ikt + 2j)*16 + 2 (t%4);
// values v0..v7 at k = kb, kb+1, kb+8, kb+9, kb+16, kb+17, kb+24, kb+25
// (tensor_core_tiled_layout.cu:136-159), nibble slots v0,v2,v4,v6,v1,v3,v5,v7 (:201-208).
// ---------------------------------------------------------------------------------------
template <int kMode>  // 0 = unpack to int32, 1 = dequant to bf16, 2 = pack from int32
__global__ void __launch_bounds__(256) tinygemm_layout_kernel(int32_t* __restrict__ packed,
                                                              const uint32_t* __restrict__ sz, int64_t N, int64_t K,
                                                              int G, int ikt, int32_t* __restrict__ q,
                                                              __nv_bfloat16* __restrict__ out) {
  const int half = ikt / 2;
  const int64_t k_outer = K / (ikt * 16);
  const int64_t total = (N / 8) * k_outer * 32 * half;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = i % half;
    const int t = (i / half) & 31;
    const int64_t tile = i / (half * 32);
    const int64_t ko = tile % k_outer, nt = tile / k_outer;
    const int64_t row = nt * 8 + (t >> 2);
    const int64_t kb = (ko * ikt + 2 * j) * 16 + 2 * (t & 3);
    const int64_t ks[4] = {kb, kb + 8, kb + 16, kb + 24};
    if constexpr (kMode == 2) {
      uint32_t word = 0;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int2 v = *reinterpret_cast<const int2*>(q + row * K + ks[p]);
        word |= (static_cast<uint32_t>(v.x) & 0xFu) << (4 * nib_slot(2 * p));
        word |= (static_cast<uint32_t>(v.y) & 0xFu) << (4 * nib_slot(2 * p + 1));
      }
      packed[i] = static_cast<int32_t>(word);
    } else {
      const uint32_t word = static_cast<uint32_t>(packed[i]);
      if constexpr (kMode == 1) {
        const uint32_t sp = sz[(kb / G) * N + row];
        const uint32_t s2 = __byte_perm(sp, sp, 0x1010), z2 = __byte_perm(sp, sp, 0x3232);
        uint32_t d[4];
        dequant_word(word, s2, z2, d);
#pragma unroll
        for (int p = 0; p < 4; ++p) *reinterpret_cast<uint32_t*>(out + row * K + ks[p]) = d[p];
      } else {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          int2 v;
          v.x = (word >> (4 * nib_slot(2 * p))) & 0xF;
          v.y = (word >> (4 * nib_slot(2 * p + 1))) & 0xF;
          *reinterpret_cast<int2*>(q + row * K + ks[p]) = v;
        }
      }
    }
  }
}

static inline int grid_for(int64_t work_items, int per_block) {
  int64_t b = (work_items + per_block - 1) / per_block;
  const int64_t cap = (int64_t)kNumSMs * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace ao

using namespace ao;

static bool valid_group(int g) { return g == 32 || g == 64 || g == 128 || g == 256; }

extern "C" int ao_b200_int4_quantize_pack(const void* w, int64_t N, int64_t K, int64_t ldw, int64_t Np,
                                          int64_t Kp, int group_size, float eps, int div_mode, void* packed,
                                          void* scale_and_zero, void* stream) {
  AO_CHECK_ARG(w && packed && scale_and_zero);
  AO_CHECK_ARG(N > 0 && K > 0 && Np >= N && Kp >= K);
  if (!valid_group(group_size)) return AO_B200_ERR_UNSUPPORTED;
  const int64_t uk = group_size > 64 ? group_size : 64;
  AO_CHECK_ARG(K % group_size == 0 && Kp % uk == 0 && Np % 16 == 0);
  AO_CHECK_ARG(ldw % 8 == 0 && ao_aligned(w, 16) && ao_aligned(packed, 16) && ao_aligned(scale_and_zero, 4));
  AO_CHECK_ARG(div_mode == AO_B200_DIV_IEEE || div_mode == AO_B200_DIV_RECIP);
  const int64_t units = (Np / 16) * (Kp / uk);
  const int grid = grid_for(units, 8);
  auto s = static_cast<cudaStream_t>(stream);
  auto wp = static_cast<const __nv_bfloat16*>(w);
  auto pk = static_cast<uint4*>(packed);
  auto sz = static_cast<uint32_t*>(scale_and_zero);
  switch (group_size) {
    case 32: int4_quantize_pack_kernel<32><<<grid, 256, 0, s>>>(wp, N, K, ldw, Np, Kp, eps, div_mode, pk, sz); break;
    case 64: int4_quantize_pack_kernel<64><<<grid, 256, 0, s>>>(wp, N, K, ldw, Np, Kp, eps, div_mode, pk, sz); break;
    case 128: int4_quantize_pack_kernel<128><<<grid, 256, 0, s>>>(wp, N, K, ldw, Np, Kp, eps, div_mode, pk, sz); break;
    default: int4_quantize_pack_kernel<256><<<grid, 256, 0, s>>>(wp, N, K, ldw, Np, Kp, eps, div_mode, pk, sz); break;
  }
  return ao_launch_status();
}

extern "C" int ao_b200_int4_choose_qparams(const void* w, int64_t N, int64_t K, int64_t ldw, int group_size,
                                           float eps, int div_mode, void* scale, void* zero, void* stream) {
  AO_CHECK_ARG(w && scale && zero && N > 0 && K > 0 && group_size > 0 && K % group_size == 0 && ldw >= K);
  AO_CHECK_ARG(div_mode == AO_B200_DIV_IEEE || div_mode == AO_B200_DIV_RECIP);
  const int64_t units = N * (K / group_size);
  int4_choose_qparams_kernel<<<grid_for(units, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(w), N, K, ldw, group_size, eps, div_mode,
      static_cast<__nv_bfloat16*>(scale), static_cast<__nv_bfloat16*>(zero));
  return ao_launch_status();
}

extern "C" int ao_b200_int4_quantize(const void* w, int64_t N, int64_t K, int64_t ldw, int group_size,
                                     const void* scale, const void* zero, int32_t* q, void* stream) {
  AO_CHECK_ARG(w && scale && zero && q && N > 0 && K > 0 && group_size > 0 && K % group_size == 0 && ldw >= K);
  int4_quantize_kernel<<<grid_for(N * K, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(w), N, K, ldw, group_size, static_cast<const __nv_bfloat16*>(scale),
      static_cast<const __nv_bfloat16*>(zero), q);
  return ao_launch_status();
}

extern "C" int ao_b200_int4_pack(const int32_t* q, int64_t N, int64_t K, int64_t Np, int64_t Kp, int pad_value,
                                 void* packed, void* stream) {
  AO_CHECK_ARG(q && packed && N > 0 && K > 0 && Np >= N && Kp >= K && Np % 16 == 0 && Kp % 64 == 0);
  AO_CHECK_ARG(ao_aligned(packed, 16));
  const int64_t total = (Np / 16) * (Kp / 64) * 32;
  int4_pack_kernel<<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      q, N, K, Np, Kp, pad_value, static_cast<uint4*>(packed));
  return ao_launch_status();
}

extern "C" int ao_b200_int4_unpack(const void* packed, int64_t Np, int64_t Kp, int32_t* q, void* stream) {
  AO_CHECK_ARG(packed && q && Np > 0 && Kp > 0 && Np % 16 == 0 && Kp % 64 == 0);
  AO_CHECK_ARG(ao_aligned(packed, 16) && ao_aligned(q, 16));
  const int64_t total = (Np / 16) * (Kp / 64) * 32;
  int4_unpack_kernel<false><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(packed), nullptr, Np, Kp, 0, q, nullptr);
  return ao_launch_status();
}

extern "C" int ao_b200_int4_dequant(const void* packed, const void* scale_and_zero, int64_t Np, int64_t Kp,
                                    int group_size, void* out, void* stream) {
  AO_CHECK_ARG(packed && scale_and_zero && out && Np > 0 && Kp > 0 && Np % 16 == 0 && Kp % 64 == 0);
  if (!valid_group(group_size)) return AO_B200_ERR_UNSUPPORTED;
  AO_CHECK_ARG(Kp % group_size == 0 && ao_aligned(packed, 16) && ao_aligned(out, 16));
  const int64_t total = (Np / 16) * (Kp / 64) * 32;
  int4_unpack_kernel<true><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(packed), static_cast<const uint32_t*>(scale_and_zero), Np, Kp, group_size,
      nullptr, static_cast<__nv_bfloat16*>(out));
  return ao_launch_status();
}

static int tinygemm_args_ok(int64_t N, int64_t K, int ikt) {
  if (!(ikt == 2 || ikt == 4 || ikt == 8)) return AO_B200_ERR_UNSUPPORTED;
  if (N <= 0 || K <= 0 || N % 8 != 0 || K % (ikt * 16) != 0) return AO_B200_ERR_INVALID_ARG;
  return AO_B200_OK;
}

extern "C" int ao_b200_tinygemm_unpack(const int32_t* packed, int64_t N, int64_t K, int inner_k_tiles, int32_t* q,
                                       void* stream) {
  AO_CHECK_ARG(packed && q);
  if (int e = tinygemm_args_ok(N, K, inner_k_tiles)) return e;
  const int64_t total = N * K / 8;
  tinygemm_layout_kernel<0><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      const_cast<int32_t*>(packed), nullptr, N, K, 0, inner_k_tiles, q, nullptr);
  return ao_launch_status();
}

extern "C" int ao_b200_tinygemm_dequant(const int32_t* packed, const void* scale_and_zero, int64_t N, int64_t K,
                                        int group_size, int inner_k_tiles, void* out, void* stream) {
  AO_CHECK_ARG(packed && scale_and_zero && out);
  if (!valid_group(group_size)) return AO_B200_ERR_UNSUPPORTED;
  if (int e = tinygemm_args_ok(N, K, inner_k_tiles)) return e;
  AO_CHECK_ARG(K % group_size == 0);
  const int64_t total = N * K / 8;
  tinygemm_layout_kernel<1><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      const_cast<int32_t*>(packed), static_cast<const uint32_t*>(scale_and_zero), N, K, group_size, inner_k_tiles,
      nullptr, static_cast<__nv_bfloat16*>(out));
  return ao_launch_status();
}

extern "C" int ao_b200_tinygemm_pack(const int32_t* q, int64_t N, int64_t K, int inner_k_tiles, int32_t* packed,
                                     void* stream) {
  AO_CHECK_ARG(packed && q);
  if (int e = tinygemm_args_ok(N, K, inner_k_tiles)) return e;
  const int64_t total = N * K / 8;
  tinygemm_layout_kernel<2><<<grid_for(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      packed, nullptr, N, K, 0, inner_k_tiles, const_cast<int32_t*>(q), nullptr);
  return ao_launch_status();
}
