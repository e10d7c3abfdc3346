// int4 weight-only linear, decode path (M <= 16): HBM-bound split-K GEMV on the strip layout.
//
//   y[m][n] = bf16( sum_k fp32( x[m][k] * W'[n][k] ) ),  W' = bf16_rn((q-8)*s + z)
//
// Weights are the A operand of mma.sync.m16n8k16 (16 rows x 16 k), the (<= 8 per tile)
// activation rows are the B operand, so one 16-byte load per lane = 32 weights = 4 MMAs and the
// only CUDA-core work per weight is the exact lop3/bf16-fma dequant.  The MMA's k index is a
// free permutation as long as A and B agree: lane (g,c) pairs weight word (row, k0..k0+7) with
// the 16 activation bytes x[m][k0..k0+7], both in natural order.
#include "common.cuh"

namespace ao {

constexpr int kGemvWarps = 8;

template <int G, int MT>
__global__ void __launch_bounds__(kGemvWarps * 32) int4_gemv_strip_kernel(
    const uint4* __restrict__ packed, const uint32_t* __restrict__ sz, const __nv_bfloat16* __restrict__ x,
    __nv_bfloat16* __restrict__ y, int M, int64_t Kx, int64_t ldx, int64_t Np, int64_t Kp, int64_t N_out,
    int64_t ldy) {
  __shared__ float red[kGemvWarps][MT][32][4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, c = lane & 3;
  const int64_t k_tiles = Kp / 64;
  const int64_t strips = Np / 16;

  for (int64_t strip = blockIdx.x; strip < strips; strip += gridDim.x) {
    float acc[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[mt][i] = 0.f;

    const uint4* wp = packed + lane;
    const int64_t row_lo = strip * 16 + g;

    constexpr int U = 4;  // tiles in flight per warp
    for (int64_t kt0 = warp * U; kt0 < k_tiles; kt0 += kGemvWarps * U) {
      uint4 wv[U];
      uint32_t szv[U][2][2];  // [tile][k half][row slot]
      uint4 xv[U][MT][2];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t kt = kt0 + u;
        const bool ok = kt < k_tiles;
        wv[u] = ok ? ldg128(wp + tile_slot(strip, kt, k_tiles) * 32) : make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int64_t grp = (kt * 64 + 32 * h) / G;
          const bool need = ok && (h == 0 || G == 32);
          szv[u][h][0] = need ? __ldg(sz + sz_index(grp, row_lo, Kp / G)) : 0u;
          szv[u][h][1] = need ? __ldg(sz + sz_index(grp, row_lo + 8, Kp / G)) : 0u;
        }
        if (G != 32) {
          szv[u][1][0] = szv[u][0][0];
          szv[u][1][1] = szv[u][0][1];
        }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int m = g + 8 * mt;
            const int64_t k0 = kt * 64 + 32 * h + 8 * c;
            xv[u][mt][h] = (ok && m < M && k0 < Kx) ? __ldg(reinterpret_cast<const uint4*>(x + m * ldx + k0))
                                                     : make_uint4(0, 0, 0, 0);
          }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t words[4] = {wv[u].x, wv[u].y, wv[u].z, wv[u].w};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t lo[4], hi[4];
          const uint32_t p0 = szv[u][h][0], p1 = szv[u][h][1];
          dequant_word(words[2 * h], __byte_perm(p0, p0, 0x1010), __byte_perm(p0, p0, 0x3232), lo);
          dequant_word(words[2 * h + 1], __byte_perm(p1, p1, 0x1010), __byte_perm(p1, p1, 0x3232), hi);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            const uint4 xx = xv[u][mt][h];
            mma_bf16_16816(acc[mt], lo[0], hi[0], lo[1], hi[1], xx.x, xx.y);
            mma_bf16_16816(acc[mt], lo[2], hi[2], lo[3], hi[3], xx.z, xx.w);
          }
        }
      }
    }
    // cross-warp split-K reduction
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
      *reinterpret_cast<float4*>(&red[warp][mt][lane][0]) = make_float4(acc[mt][0], acc[mt][1], acc[mt][2], acc[mt][3]);
    __syncthreads();
    if (warp < MT) {
      const int mt = warp;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < kGemvWarps; ++w) {
        const float4 v = *reinterpret_cast<const float4*>(&red[w][mt][lane][0]);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      // C fragment: (row g, cols 2c,2c+1), (row g+8, cols 2c,2c+1); cols are batch rows m
      const int m0 = 2 * c + 8 * mt;
      const int64_t n0 = strip * 16 + g;
      if (m0 < M) {
        if (n0 < N_out) y[m0 * ldy + n0] = __float2bfloat16_rn(s.x);
        if (n0 + 8 < N_out) y[m0 * ldy + n0 + 8] = __float2bfloat16_rn(s.z);
      }
      if (m0 + 1 < M) {
        if (n0 < N_out) y[(m0 + 1) * ldy + n0] = __float2bfloat16_rn(s.y);
        if (n0 + 8 < N_out) y[(m0 + 1) * ldy + n0 + 8] = __float2bfloat16_rn(s.w);
      }
    }
    __syncthreads();
  }
}

template <int G>
static int launch_gemv(int mt, const uint4* packed, const uint32_t* sz, const __nv_bfloat16* x, __nv_bfloat16* y,
                       int M, int64_t Kx, int64_t ldx, int64_t Np, int64_t Kp, int64_t N_out, int64_t ldy,
                       cudaStream_t s) {
  const int64_t strips = Np / 16;
  const int grid = static_cast<int>(strips < (int64_t)kNumSMs * 8 ? strips : (int64_t)kNumSMs * 8);
  if (mt == 1)
    int4_gemv_strip_kernel<G, 1><<<grid, kGemvWarps * 32, 0, s>>>(packed, sz, x, y, M, Kx, ldx, Np, Kp, N_out, ldy);
  else
    int4_gemv_strip_kernel<G, 2><<<grid, kGemvWarps * 32, 0, s>>>(packed, sz, x, y, M, Kx, ldx, Np, Kp, N_out, ldy);
  return ao_launch_status();
}

int int4_gemv(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
              int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, cudaStream_t s) {
  auto pk = static_cast<const uint4*>(packed);
  auto szp = static_cast<const uint32_t*>(sz);
  auto xp = static_cast<const __nv_bfloat16*>(x);
  auto yp = static_cast<__nv_bfloat16*>(y);
  const int mt = M <= 8 ? 1 : 2;
  switch (G) {
    case 32: return launch_gemv<32>(mt, pk, szp, xp, yp, (int)M, Kx, ldx, Np, Kp, N_out, ldy, s);
    case 64: return launch_gemv<64>(mt, pk, szp, xp, yp, (int)M, Kx, ldx, Np, Kp, N_out, ldy, s);
    case 128: return launch_gemv<128>(mt, pk, szp, xp, yp, (int)M, Kx, ldx, Np, Kp, N_out, ldy, s);
    case 256: return launch_gemv<256>(mt, pk, szp, xp, yp, (int)M, Kx, ldx, Np, Kp, N_out, ldy, s);
  }
  return AO_B200_ERR_UNSUPPORTED;
}

}  // namespace ao
