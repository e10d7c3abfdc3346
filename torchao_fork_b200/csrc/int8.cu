// int8 weight-only: symmetric per-row quantize (bit-exact with the reference's eager bf16 ops), the strip-major
// pack / unpack, and the fallback decode GEMV  y = bf16( bf16(x @ q^T) * scale ).
#include "common.cuh"

namespace ao {

// ---------------------------------------------------------------------------------------
// quantize: one CTA per row.  choose_qparams_affine SYMMETRIC (quant_primitives.py:1537-1565):
//   amax = max(-min(mn,0), max(mx,0)); scale = clamp(bf16(amax / 127.5), eps)
// quantize_affine (:453-456): q = clamp(round(bf16(w * bf16(1/scale))), -128, 127)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) int8_quantize_kernel(const __nv_bfloat16* __restrict__ w, int64_t N, int64_t K,
                                                            int64_t ldw, float eps, int div_mode,
                                                            int8_t* __restrict__ q, __nv_bfloat16* __restrict__ scale) {
  __shared__ float s_mn[8], s_mx[8];
  __shared__ float s_inv;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t row = blockIdx.x; row < N; row += gridDim.x) {
    const __nv_bfloat16* p = w + row * ldw;
    float mn = INFINITY, mx = -INFINITY;
    const bool vec = (K % 8 == 0) && (ldw % 8 == 0) && ((reinterpret_cast<uintptr_t>(w) & 15) == 0);
    if (vec) {
      // packed bf16x2 min / max (comparisons only: exact), one instruction per two weights
      uint32_t mn2 = 0x7F807F80u, mx2 = 0xFF80FF80u;
      for (int64_t i = threadIdx.x; i < K / 8; i += blockDim.x) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + i);
        const uint32_t r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          asm("min.bf16x2 %0, %0, %1;" : "+r"(mn2) : "r"(r[e]));
          asm("max.bf16x2 %0, %0, %1;" : "+r"(mx2) : "r"(r[e]));
        }
      }
      mn = fminf(__uint_as_float(mn2 << 16), __uint_as_float(mn2 & 0xFFFF0000u));
      mx = fmaxf(__uint_as_float(mx2 << 16), __uint_as_float(mx2 & 0xFFFF0000u));
    } else {
      for (int64_t i = threadIdx.x; i < K; i += blockDim.x) {
        const float v = __bfloat162float(p[i]);
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (lane == 0) {
      s_mn[warp] = mn;
      s_mx[warp] = mx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 1; i < 8; ++i) {
        mn = fminf(mn, s_mn[i]);
        mx = fmaxf(mx, s_mx[i]);
      }
      const float amax = fmaxf(-fminf(mn, 0.f), fmaxf(mx, 0.f));
      float s = (div_mode == AO_B200_DIV_RECIP) ? bf16r(amax * (1.0f / 127.5f)) : bf16r(__fdiv_rn(amax, 127.5f));
      s = bf16r(fmaxf(s, eps));
      scale[row] = __float2bfloat16_rn(s);
      s_inv = bf16r(__fdiv_rn(1.0f, s));
    }
    __syncthreads();
    const float inv = s_inv;
    int8_t* qo = q + row * K;
    if (vec) {
      // bf16(w * inv) is ONE packed multiply (the fp32 product of two bf16 values is exact, so the reference's
      // compute-in-fp32-then-round is a single rounding = mul.rn.bf16x2); clamp in bf16x2 (commutes with rint at the
      // integer bounds); rint = add 1.5 * 2^23, whose low mantissa byte is then the two's-complement int8
      const uint32_t ib = float_to_bf16_bits(inv), inv2 = ib | (ib << 16);
      constexpr float kMagic = 12582912.0f;
      for (int64_t i = threadIdx.x; i < K / 8; i += blockDim.x) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + i);
        const uint32_t r[4] = {v.x, v.y, v.z, v.w};
        uint32_t u[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          uint32_t t2;
          asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(t2) : "r"(r[e]), "r"(inv2));
          asm("min.bf16x2 %0, %0, %1;" : "+r"(t2) : "r"(0x42FE42FEu));  // 127.0
          asm("max.bf16x2 %0, %0, %1;" : "+r"(t2) : "r"(0xC300C300u));  // -128.0
          u[2 * e] = __float_as_uint(__uint_as_float(t2 << 16) + kMagic);
          u[2 * e + 1] = __float_as_uint(__uint_as_float(t2 & 0xFFFF0000u) + kMagic);
        }
        const uint32_t o0 = __byte_perm(__byte_perm(u[0], u[1], 0x0040), __byte_perm(u[2], u[3], 0x0040), 0x5410);
        const uint32_t o1 = __byte_perm(__byte_perm(u[4], u[5], 0x0040), __byte_perm(u[6], u[7], 0x0040), 0x5410);
        *reinterpret_cast<uint2*>(qo + i * 8) = make_uint2(o0, o1);
      }
    } else {
      for (int64_t i = threadIdx.x; i < K; i += blockDim.x) {
        const float r = fminf(fmaxf(rintf(bf16r(__bfloat162float(p[i]) * inv)), -128.f), 127.f);
        qo[i] = static_cast<int8_t>(static_cast<int>(r));
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------
// strip-major int8 layout (include/ao_b200.h): uint8 [Np/16][Kp/64][2 halves][32 lanes][16 B], BIASED (u = q + 128).
// Lane l = (g = l >> 2, c = l & 3) of half h owns 4 words; word j holds row 16 s + g + 8 (j & 1),
// k0 = 64 kt + 32 h + 8 c + 4 (j >> 1), bytes k0 .. k0 + 3.  One thread per 16-byte chunk.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) int8_pack_kernel(const int8_t* __restrict__ q, int64_t N, int64_t K, int64_t Np,
                                                        int64_t Kp, uint4* __restrict__ packed) {
  const int64_t KT = Kp / 64, total = (Np / 16) * KT * 64;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int lane = (int)(idx & 31), h = (int)((idx >> 5) & 1);
    const int64_t tile = idx >> 6, s = tile / KT, kt = tile - s * KT;
    const int g = lane >> 2, c = lane & 3;
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t row = 16 * s + g + 8 * (j & 1), k0 = 64 * kt + 32 * h + 8 * c + 4 * (j >> 1);
      uint32_t v = 0x80808080u;  // q = 0
      if (row < N) {
        if (k0 + 4 <= K && (K & 3) == 0) {
          v = *reinterpret_cast<const uint32_t*>(q + row * K + k0) ^ 0x80808080u;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (k0 + e < K) v = (v & ~(0xFFu << (8 * e))) | ((((uint32_t)(uint8_t)q[row * K + k0 + e]) ^ 0x80u) << (8 * e));
        }
      }
      w[j] = v;
    }
    packed[idx] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void __launch_bounds__(256) int8_unpack_kernel(const uint4* __restrict__ packed, int64_t Np, int64_t Kp, int64_t N,
                                                          int64_t K, int8_t* __restrict__ q) {
  const int64_t KT = Kp / 64, total = (Np / 16) * KT * 64;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int lane = (int)(idx & 31), h = (int)((idx >> 5) & 1);
    const int64_t tile = idx >> 6, s = tile / KT, kt = tile - s * KT;
    const int g = lane >> 2, c = lane & 3;
    const uint4 v = packed[idx];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t row = 16 * s + g + 8 * (j & 1), k0 = 64 * kt + 32 * h + 8 * c + 4 * (j >> 1);
      if (row >= N) continue;
      const uint32_t t = w[j] ^ 0x80808080u;
      if (k0 + 4 <= K && (K & 3) == 0) {
        *reinterpret_cast<uint32_t*>(q + row * K + k0) = t;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (k0 + e < K) q[row * K + k0 + e] = (int8_t)((t >> (8 * e)) & 0xFFu);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Fallback GEMV (M <= 16) for shapes whose activations do not fit the fused kernel's shared-memory staging
// (gemv_int8_fused.cu).  A CTA owns a strip; warps split the k-tiles; a lane's 16-byte chunk is rows g / g + 8 of a
// 32-k half, converted exactly to bf16 (byte-insert into the fp32 1.5 * 2^23 magic, subtract, pack pairs).
// ---------------------------------------------------------------------------------------
constexpr int kI8Warps = 8;

template <int MT>
__global__ void __launch_bounds__(kI8Warps * 32) int8_gemv_kernel(const uint8_t* __restrict__ packed,
                                                                  const __nv_bfloat16* __restrict__ scale,
                                                                  const __nv_bfloat16* __restrict__ x,
                                                                  __nv_bfloat16* __restrict__ y, int M, int64_t N_out,
                                                                  int64_t Kx, int KT, int64_t ldx, int64_t ldy) {
  __shared__ float red[kI8Warps][MT][32][4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, c = lane & 3;
  const int64_t strips = (N_out + 15) / 16;
  const int n_half = 2 * KT;  // 32-k halves of a strip, 512 bytes each

  for (int64_t strip = blockIdx.x; strip < strips; strip += gridDim.x) {
    float acc[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[mt][i] = 0.f;
    const uint8_t* wp = packed + strip * (int64_t)KT * 1024 + lane * 16;

    constexpr int U = 4;
    for (int h0 = warp * U; h0 < n_half; h0 += kI8Warps * U) {
      uint4 a[U];
      uint4 xv[U][MT];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int hh = h0 + u;
        const bool ok = hh < n_half;
        a[u] = ok ? ldg128(wp + (int64_t)hh * 512) : make_uint4(0x80808080u, 0x80808080u, 0x80808080u, 0x80808080u);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          const int m = g + 8 * mt;
          const int64_t k0 = (int64_t)hh * 32 + 8 * c;
          xv[u][mt] = (ok && m < M && k0 < Kx) ? __ldg(reinterpret_cast<const uint4*>(x + m * ldx + k0)) : make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        uint32_t g01, g23, h01, h23, g45, g67, h45, h67;
        u8x4_to_bf16x2x2(a[u].x, g01, g23);  // row g,     k 8c .. 8c+3
        u8x4_to_bf16x2x2(a[u].y, h01, h23);  // row g + 8
        u8x4_to_bf16x2x2(a[u].z, g45, g67);  // row g,     k 8c+4 .. 8c+7
        u8x4_to_bf16x2x2(a[u].w, h45, h67);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_bf16_16816(acc[mt], g01, h01, g23, h23, xv[u][mt].x, xv[u][mt].y);
          mma_bf16_16816(acc[mt], g45, h45, g67, h67, xv[u][mt].z, xv[u][mt].w);
        }
      }
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
      *reinterpret_cast<float4*>(&red[warp][mt][lane][0]) = make_float4(acc[mt][0], acc[mt][1], acc[mt][2], acc[mt][3]);
    __syncthreads();
    if (warp < MT) {
      const int mt = warp;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < kI8Warps; ++w) {
        const float4 v = *reinterpret_cast<const float4*>(&red[w][mt][lane][0]);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      const int64_t r0 = strip * 16 + g, r1 = r0 + 8;
      const bool v0 = r0 < N_out, v1 = r1 < N_out;
      const int m0 = 2 * c + 8 * mt;
      const float sc0 = v0 ? __bfloat162float(scale[r0]) : 0.f;
      const float sc1 = v1 ? __bfloat162float(scale[r1]) : 0.f;
      // plain_layout.py:256-262: mm rounds to bf16, THEN the scale multiply rounds again
      if (m0 < M) {
        if (v0) y[m0 * ldy + r0] = __float2bfloat16_rn(bf16r(s.x) * sc0);
        if (v1) y[m0 * ldy + r1] = __float2bfloat16_rn(bf16r(s.z) * sc1);
      }
      if (m0 + 1 < M) {
        if (v0) y[(m0 + 1) * ldy + r0] = __float2bfloat16_rn(bf16r(s.y) * sc0);
        if (v1) y[(m0 + 1) * ldy + r1] = __float2bfloat16_rn(bf16r(s.w) * sc1);
      }
    }
    __syncthreads();
  }
}

int int8_gemv(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Kp,
              void* y, int64_t N_out, int64_t ldy, cudaStream_t s) {
  const int64_t strips = (N_out + 15) / 16;
  const int grid = static_cast<int>(strips < (int64_t)kNumSMs * 8 ? strips : (int64_t)kNumSMs * 8);
  auto xp = static_cast<const __nv_bfloat16*>(x);
  auto yp = static_cast<__nv_bfloat16*>(y);
  auto sp = static_cast<const __nv_bfloat16*>(scale);
  auto wp = static_cast<const uint8_t*>(packed);
  if (M <= 8)
    int8_gemv_kernel<1><<<grid, kI8Warps * 32, 0, s>>>(wp, sp, xp, yp, (int)M, N_out, Kx, (int)(Kp / 64), ldx, ldy);
  else
    int8_gemv_kernel<2><<<grid, kI8Warps * 32, 0, s>>>(wp, sp, xp, yp, (int)M, N_out, Kx, (int)(Kp / 64), ldx, ldy);
  return ao_launch_status();
}

}  // namespace ao

extern "C" int ao_b200_int8_pack(const int8_t* q, int64_t N, int64_t K, int64_t Np, int64_t Kp, void* packed, void* stream) {
  AO_CHECK_ARG(q && packed && N > 0 && K > 0 && Np >= N && Kp >= K && Np % 16 == 0 && Kp % 64 == 0);
  AO_CHECK_ARG(ao_aligned(packed, 16) && ((K & 3) != 0 || ao_aligned(q, 4)));
  const int64_t total = (Np / 16) * (Kp / 64) * 64;
  const int64_t blocks = (total + 255) / 256;
  ao::int8_pack_kernel<<<(unsigned)(blocks < 148 * 32 ? blocks : 148 * 32), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      q, N, K, Np, Kp, static_cast<uint4*>(packed));
  return ao_launch_status();
}

extern "C" int ao_b200_int8_unpack(const void* packed, int64_t Np, int64_t Kp, int64_t N, int64_t K, int8_t* q, void* stream) {
  AO_CHECK_ARG(q && packed && N > 0 && K > 0 && Np >= N && Kp >= K && Np % 16 == 0 && Kp % 64 == 0);
  AO_CHECK_ARG(ao_aligned(packed, 16) && ((K & 3) != 0 || ao_aligned(q, 4)));
  const int64_t total = (Np / 16) * (Kp / 64) * 64;
  const int64_t blocks = (total + 255) / 256;
  ao::int8_unpack_kernel<<<(unsigned)(blocks < 148 * 32 ? blocks : 148 * 32), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint4*>(packed), Np, Kp, N, K, q);
  return ao_launch_status();
}

extern "C" int ao_b200_int8_quantize(const void* w, int64_t N, int64_t K, int64_t ldw, float eps, int div_mode,
                                     int8_t* q, void* scale, void* stream) {
  AO_CHECK_ARG(w && q && scale && N > 0 && K > 0 && ldw >= K);
  AO_CHECK_ARG(div_mode == AO_B200_DIV_IEEE || div_mode == AO_B200_DIV_RECIP);
  AO_CHECK_ARG(K % 8 != 0 || ao_aligned(q, 8));
  const int grid = static_cast<int>(N < (int64_t)ao::kNumSMs * 8 ? N : (int64_t)ao::kNumSMs * 8);
  ao::int8_quantize_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(w), N, K, ldw, eps, div_mode, q, static_cast<__nv_bfloat16*>(scale));
  return ao_launch_status();
}
