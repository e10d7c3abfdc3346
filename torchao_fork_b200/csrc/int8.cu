// int8 weight-only: symmetric per-row quantize (bit-exact with the reference's eager bf16 ops)
// and the decode GEMV  y = bf16( bf16(x @ q^T) * scale ).
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
      for (int64_t i = threadIdx.x; i < K / 8; i += blockDim.x) {
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(p) + i), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          mn = fminf(mn, f[e]);
          mx = fmaxf(mx, f[e]);
        }
      }
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
      for (int64_t i = threadIdx.x; i < K / 8; i += blockDim.x) {
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(p) + i), f);
        uint32_t o[2] = {0, 0};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float r = fminf(fmaxf(rintf(bf16r(f[e] * inv)), -128.f), 127.f);
          o[e >> 2] |= (static_cast<uint32_t>(static_cast<int>(r)) & 0xFFu) << (8 * (e & 3));
        }
        *reinterpret_cast<uint2*>(qo + i * 8) = make_uint2(o[0], o[1]);
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
// GEMV (M <= 16).  q is plain row-major int8 [N][K] (PlainAQTTensorImpl.int_data, kept
// bit-identical to the reference's tensor).  A CTA owns 16 rows; warps split K; lane (g,c)
// reads 16 bytes of row g and of row g+8 per 64-wide k step and converts int8 -> bf16 exactly
// (xor 0x80, byte-insert into the fp32 1.5*2^23 magic, subtract, pack pairs to bf16x2).
// ---------------------------------------------------------------------------------------
constexpr int kI8Warps = 8;

// 4 int8 (one register) -> 2 bf16x2 registers: (b0,b1), (b2,b3)
__device__ __forceinline__ void i8x4_to_bf16x2x2(uint32_t v, uint32_t& o01, uint32_t& o23) {
  const uint32_t u = v ^ 0x80808080u;  // b + 128 in 0..255
  constexpr uint32_t kMagic = 0x4B400000u;         // 1.5 * 2^23
  constexpr float kBias = 12582912.0f + 128.0f;
  const float f0 = __uint_as_float(__byte_perm(u, kMagic, 0x7650)) - kBias;
  const float f1 = __uint_as_float(__byte_perm(u, kMagic, 0x7651)) - kBias;
  const float f2 = __uint_as_float(__byte_perm(u, kMagic, 0x7652)) - kBias;
  const float f3 = __uint_as_float(__byte_perm(u, kMagic, 0x7653)) - kBias;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o01) : "f"(f1), "f"(f0));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o23) : "f"(f3), "f"(f2));
}

template <int MT>
__global__ void __launch_bounds__(kI8Warps * 32) int8_gemv_kernel(const int8_t* __restrict__ q,
                                                                  const __nv_bfloat16* __restrict__ scale,
                                                                  const __nv_bfloat16* __restrict__ x,
                                                                  __nv_bfloat16* __restrict__ y, int M, int64_t N,
                                                                  int64_t K, int64_t ldx, int64_t ldy) {
  __shared__ float red[kI8Warps][MT][32][4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, c = lane & 3;
  const int64_t strips = (N + 15) / 16;
  const int64_t k_steps = (K + 63) / 64;  // K % 16 == 0 checked by the host; tail chunks predicated

  for (int64_t strip = blockIdx.x; strip < strips; strip += gridDim.x) {
    float acc[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[mt][i] = 0.f;
    const int64_t r0 = strip * 16 + g, r1 = r0 + 8;
    const bool v0 = r0 < N, v1 = r1 < N;
    const int8_t* p0 = q + r0 * K + 16 * c;
    const int8_t* p1 = q + r1 * K + 16 * c;

    constexpr int U = 4;
    for (int64_t ks0 = warp * U; ks0 < k_steps; ks0 += kI8Warps * U) {
      uint4 a0[U], a1[U];
      uint4 xv[U][MT][2];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t ks = ks0 + u;
        const bool ok = ks < k_steps && ks * 64 + 16 * c < K;
        a0[u] = (ok && v0) ? ldg128(p0 + ks * 64) : make_uint4(0, 0, 0, 0);
        a1[u] = (ok && v1) ? ldg128(p1 + ks * 64) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          const int m = g + 8 * mt;
          const __nv_bfloat16* xp = x + m * ldx + ks * 64 + 16 * c;
          const bool xo = ok && m < M;
          xv[u][mt][0] = xo ? __ldg(reinterpret_cast<const uint4*>(xp)) : make_uint4(0, 0, 0, 0);
          xv[u][mt][1] = xo ? __ldg(reinterpret_cast<const uint4*>(xp) + 1) : make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t w0[4] = {a0[u].x, a0[u].y, a0[u].z, a0[u].w};
        const uint32_t w1[4] = {a1[u].x, a1[u].y, a1[u].z, a1[u].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {  // 4 k's per register -> one MMA (k-slots {2c,2c+1},{2c+8,2c+9})
          uint32_t lo01, lo23, hi01, hi23;
          i8x4_to_bf16x2x2(w0[i], lo01, lo23);
          i8x4_to_bf16x2x2(w1[i], hi01, hi23);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            const uint4 xa = xv[u][mt][i >> 1];
            const uint32_t b0 = (i & 1) ? xa.z : xa.x, b1 = (i & 1) ? xa.w : xa.y;
            mma_bf16_16816(acc[mt], lo01, hi01, lo23, hi23, b0, b1);
          }
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

int int8_gemv(const void* x, int64_t M, int64_t K, int64_t ldx, const int8_t* q, const void* scale, int64_t N,
              void* y, int64_t ldy, cudaStream_t s) {
  const int64_t strips = (N + 15) / 16;
  const int grid = static_cast<int>(strips < (int64_t)kNumSMs * 8 ? strips : (int64_t)kNumSMs * 8);
  auto xp = static_cast<const __nv_bfloat16*>(x);
  auto yp = static_cast<__nv_bfloat16*>(y);
  auto sp = static_cast<const __nv_bfloat16*>(scale);
  if (M <= 8)
    int8_gemv_kernel<1><<<grid, kI8Warps * 32, 0, s>>>(q, sp, xp, yp, (int)M, N, K, ldx, ldy);
  else
    int8_gemv_kernel<2><<<grid, kI8Warps * 32, 0, s>>>(q, sp, xp, yp, (int)M, N, K, ldx, ldy);
  return ao_launch_status();
}

}  // namespace ao

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
