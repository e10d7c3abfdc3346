// int4 weight-only linear, decode path (M <= 16): persistent, TMA-bulk-staged, stream-K GEMV.
//
//   y[m][n] = bf16( sum_k fp32( x[m][k] * W'[n][k] ) ),  W' = bf16_rn((q-8)*s + z)
//
// Structure (one CTA per SM):
//   * the packed weight is a flat sequence of T = (Np/128)*(Kp/64) super-tiles of 4 KB; CTA b owns the
//     contiguous range [b*T/B, (b+1)*T/B) — every SM streams the same number of bytes (+-4 KB);
//   * a producer warp feeds a shared-memory ring with cp.async.bulk (one 16 KB weight copy + the
//     512-byte scale/zero blocks per stage, completion on an mbarrier), independent of the consumers'
//     progress, so the bytes in flight per SM are bounded by the ring (>= 128 KB), not by registers;
//   * kConsumers consumer warps take super-tiles round-robin: LDS.128 -> exact lop3/bf16-fma dequant ->
//     mma.sync.m16n8k16 with the weights as the A operand (16 rows x 16 k) and the <= 8 activation rows
//     as B; the same activation fragment serves the 8 sub-tiles of a super-tile;
//   * a CTA's range is cut into segments by 128-row super-strip; at the end of a segment the warps
//     reduce through shared memory (deterministic order); a segment covering the whole K extent is
//     written as bf16, otherwise it becomes a partial record in the workspace and the LAST CTA to
//     arrive for that super-strip (atomic counter) sums the records in CTA order and writes y;
//   * PDL: weights never depend on the previous kernel, so the producer starts streaming before
//     griddepcontrol.wait; consumers wait before touching x / y / workspace.
#include "common.cuh"

namespace ao {

constexpr int kConsumers = 16;       // consumer warps
constexpr int kStageSupers = 4;      // super-tiles per pipeline stage
constexpr int kSuperBytes = 4096;
constexpr int kRedFloats = kConsumers * 128 * 4;

struct GemvStreamParams {
  const uint4* packed;
  const uint32_t* sz;
  const __nv_bfloat16* x;
  __nv_bfloat16* y;
  float* part;    // [B][2][128][MP]
  int* counters;  // [Np/128], zero on entry, zero on exit
  int M, MP;
  int64_t Kx, ldx, Np, N_out, ldy, KT, T;
  int stages;
};

template <int G>
__host__ __device__ constexpr int sz_bytes_per_super() { return G == 32 ? 1024 : 512; }

__device__ __forceinline__ int64_t cta_start(int64_t b, int64_t T, int64_t B) { return (b * T) / B; }

template <int G, int MT>
__global__ void __launch_bounds__((kConsumers + 1) * 32, 1) int4_gemv_stream_kernel(const GemvStreamParams p) {
  constexpr int SZB = sz_bytes_per_super<G>();
  constexpr int kStageBytes = kStageSupers * (kSuperBytes + SZB);
  extern __shared__ __align__(128) uint8_t smem[];
  const int stages = p.stages;
  uint8_t* ring = smem;
  float* red = reinterpret_cast<float*>(smem + (size_t)stages * kStageBytes);
  uint64_t* full = reinterpret_cast<uint64_t*>(red + kRedFloats);
  uint64_t* empty = full + stages;
  int* s_flag = reinterpret_cast<int*>(empty + stages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t B = gridDim.x, b = blockIdx.x;
  const int64_t t0 = cta_start(b, p.T, B), t1 = cta_start(b + 1, p.T, B);
  const int64_t n_super = t1 - t0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kStageSupers);
    }
    fence_barrier_init();
  }
  __syncthreads();
  grid_launch_dependents();

  if (warp == kConsumers) {
    // ===================== producer (one warp; lane 0 owns the barriers, lanes share the copies) ============
    // 32-bit incremental (S, kt) bookkeeping: no 64-bit divides on the critical issue path.
    const int n_sup = static_cast<int>(n_super);
    const int KT = static_cast<int>(p.KT);
    int S = static_cast<int>(t0 / p.KT);
    int kt = static_cast<int>(t0 - (int64_t)S * p.KT);
    int slot = 0;
    uint32_t par = 1;  // parity to wait on for "slot is free": first lap passes immediately
    const uint4* wsrc = p.packed + t0 * (kSuperBytes / 16);
    for (int done = 0; done < n_sup; done += kStageSupers) {
      const int cnt = min(kStageSupers, n_sup - done);
      uint8_t* wdst = ring + (size_t)slot * kStageBytes;
      uint8_t* sdst = wdst + kStageSupers * kSuperBytes;
      if (lane == 0) {
        mbar_wait(&empty[slot], par);
        mbar_arrive_expect_tx(&full[slot], static_cast<uint32_t>(cnt * (kSuperBytes + SZB)));
        bulk_g2s(wdst, wsrc, static_cast<uint32_t>(cnt * kSuperBytes), &full[slot]);
      }
      __syncwarp();
      // scale/zero blocks: lane j < cnt * (SZB/512) copies one 512-byte block
      constexpr int kBlk = SZB / 512;
      if (lane < cnt * kBlk) {
        const int j = lane / kBlk, h = lane % kBlk;
        int kj = kt + j, Sj = S;
        if (kj >= KT) {
          kj -= KT;
          Sj += 1;
        }
        const int grp = (kj * 64) / G + h;
        bulk_g2s(sdst + j * SZB + h * 512, p.sz + (int64_t)grp * p.Np + 128 * (int64_t)Sj, 512, &full[slot]);
      }
      wsrc += cnt * (kSuperBytes / 16);
      kt += cnt;
      if (kt >= KT) {
        kt -= KT;
        S += 1;
      }
      if (++slot == stages) {
        slot = 0;
        par ^= 1u;
      }
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, c = lane & 3;
  const int tid = threadIdx.x;  // 0 .. kConsumers*32-1
  grid_dependency_wait();       // x, y and the workspace belong to the dependency chain

  const int KT = static_cast<int>(p.KT);
  const int n_sup = static_cast<int>(n_super);
  const int64_t S_first = t0 / p.KT, S_last = (t1 - 1) / p.KT;
  for (int64_t S = S_first; S <= S_last; ++S) {
    const int64_t seg_lo = max(t0, S * p.KT), seg_hi = min(t1, (S + 1) * p.KT);
    float acc[8][MT][4];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[r][mt][e] = 0.f;

    const int first = static_cast<int>(seg_lo - t0), last = static_cast<int>(seg_hi - t0);
    int i = first + ((warp - (first % kConsumers) + kConsumers) % kConsumers);
    int kt = static_cast<int>(t0 + i - S * p.KT);
    int stage = i / kStageSupers;
    int slot = stage % stages;
    uint32_t par = static_cast<uint32_t>(stage / stages) & 1u;
    for (; i < last; i += kConsumers, kt += kConsumers) {
      // activation fragments for this k-tile (shared by the 8 sub-tiles)
      uint4 xv[MT][2];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int m = g + 8 * mt;
          const int k0 = kt * 64 + 32 * h + 8 * c;
          xv[mt][h] = (m < p.M && k0 < p.Kx) ? __ldg(reinterpret_cast<const uint4*>(p.x + m * p.ldx + k0))
                                              : make_uint4(0, 0, 0, 0);
        }
      mbar_wait(&full[slot], par);
      const uint8_t* wsrc = ring + (size_t)slot * kStageBytes + (i % kStageSupers) * kSuperBytes;
      const uint32_t* ssrc = reinterpret_cast<const uint32_t*>(ring + (size_t)slot * kStageBytes +
                                                               kStageSupers * kSuperBytes + (i % kStageSupers) * SZB);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const uint4 wv = *reinterpret_cast<const uint4*>(wsrc + r * 512 + lane * 16);
        const uint32_t words[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint32_t* sh = ssrc + ((G == 32) ? h * 128 : 0);
          const uint32_t p0 = sh[16 * r + g], p1 = sh[16 * r + g + 8];
          uint32_t lo[4], hi[4];
          dequant_word(words[2 * h], __byte_perm(p0, p0, 0x1010), __byte_perm(p0, p0, 0x3232), lo);
          dequant_word(words[2 * h + 1], __byte_perm(p1, p1, 0x1010), __byte_perm(p1, p1, 0x3232), hi);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            const uint4 xx = xv[mt][h];
            mma_bf16_16816(acc[r][mt], lo[0], hi[0], lo[1], hi[1], xx.x, xx.y);
            mma_bf16_16816(acc[r][mt], lo[2], hi[2], lo[3], hi[3], xx.z, xx.w);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      // advance kConsumers super-tiles = kConsumers / kStageSupers stages
      slot += kConsumers / kStageSupers;
      while (slot >= stages) {
        slot -= stages;
        par ^= 1u;
      }
    }
    (void)n_sup;

    // ---- segment flush: cross-warp reduction, 4 activation rows (MMA columns) per pass ----
    const bool whole = (seg_lo == S * p.KT) && (seg_hi == (S + 1) * p.KT);
    const int rec_slot = (S == S_first) ? 0 : 1;
    float* rec = p.part + ((b * 2 + rec_slot) * 128) * p.MP;
    const int passes = (p.M + 3) >> 2;
    for (int ps = 0; ps < passes; ++ps) {
      const int mt = ps >> 1;
      if ((c >> 1) == (ps & 1)) {
        float* dst = red + warp * 512 + 2 * (c & 1);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          float a0, a1, a2, a3;
          if (MT == 2 && mt == 1) {
            a0 = acc[r][MT - 1][0]; a1 = acc[r][MT - 1][1]; a2 = acc[r][MT - 1][2]; a3 = acc[r][MT - 1][3];
          } else {
            a0 = acc[r][0][0]; a1 = acc[r][0][1]; a2 = acc[r][0][2]; a3 = acc[r][0][3];
          }
          *reinterpret_cast<float2*>(dst + (16 * r + g) * 4) = make_float2(a0, a1);
          *reinterpret_cast<float2*>(dst + (16 * r + g + 8) * 4) = make_float2(a2, a3);
        }
      }
      named_bar_sync(1, kConsumers * 32);
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < kConsumers; ++w) v += red[w * 512 + tid];
      const int row = tid >> 2, m = 4 * ps + (tid & 3);
      const int64_t n = 128 * S + row;
      if (whole) {
        if (m < p.M && n < p.N_out) p.y[m * p.ldy + n] = __float2bfloat16_rn(v);
      } else {
        rec[row * p.MP + m] = v;
      }
      named_bar_sync(1, kConsumers * 32);
    }
    if (!whole) {
      // publish the record; the last CTA to arrive for super-strip S reduces all of them in CTA order
      __threadfence();
      named_bar_sync(1, kConsumers * 32);
      const int64_t b_first = ((S * p.KT + 1) * B + p.T - 1) / p.T - 1;
      const int64_t b_last = (((S + 1) * p.KT) * B + p.T - 1) / p.T - 1;
      if (tid == 0) {
        const int old = atomicAdd(&p.counters[S], 1);
        *s_flag = (old == static_cast<int>(b_last - b_first)) ? 1 : 0;
      }
      named_bar_sync(1, kConsumers * 32);
      if (*s_flag) {
        __threadfence();
        const int row = tid >> 2;
        const int64_t n = 128 * S + row;
        for (int ps = 0; ps < passes; ++ps) {
          const int m = 4 * ps + (tid & 3);
          float v = 0.f;
          for (int64_t bb = b_first; bb <= b_last; ++bb) {
            const int sl = (S == cta_start(bb, p.T, B) / p.KT) ? 0 : 1;
            v += __ldcg(p.part + ((bb * 2 + sl) * 128 + row) * p.MP + m);
          }
          if (m < p.M && n < p.N_out) p.y[m * p.ldy + n] = __float2bfloat16_rn(v);
        }
        if (tid == 0) p.counters[S] = 0;
      }
      named_bar_sync(1, kConsumers * 32);
    }
  }
}

static int g_sm_count = 0;
static int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0, n = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    g_sm_count = n > 0 ? n : kNumSMs;
  }
  return g_sm_count;
}

int64_t int4_gemv_stream_workspace_bytes(int64_t Np) {
  const int64_t counters = ((Np / 128) * 4 + 255) / 256 * 256;
  const int64_t parts = (int64_t)sm_count() * 2 * 128 * 16 * 4;
  return counters + parts;
}

template <int G, int MT>
static int launch_stream(const GemvStreamParams& p, int grid, bool pdl, cudaStream_t s) {
  constexpr int SZB = sz_bytes_per_super<G>();
  constexpr int kStageBytes = kStageSupers * (kSuperBytes + SZB);
  const size_t smem = (size_t)p.stages * kStageBytes + kRedFloats * 4 + p.stages * 16 + 16;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(int4_gemv_stream_kernel<G, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             227 * 1024) != cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3((kConsumers + 1) * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, int4_gemv_stream_kernel<G, MT>, p) != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

int int4_gemv_stream(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz,
                     int64_t Np, int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, void* workspace,
                     int64_t workspace_bytes, int stages, bool pdl, cudaStream_t s) {
  if (workspace_bytes < int4_gemv_stream_workspace_bytes(Np) || !workspace) return AO_B200_ERR_WORKSPACE;
  GemvStreamParams p;
  p.packed = static_cast<const uint4*>(packed);
  p.sz = static_cast<const uint32_t*>(sz);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.counters = static_cast<int*>(workspace);
  p.part = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + ((Np / 128) * 4 + 255) / 256 * 256);
  p.M = static_cast<int>(M);
  p.MP = 4 * static_cast<int>((M + 3) / 4);
  p.Kx = Kx; p.ldx = ldx; p.Np = Np; p.N_out = N_out; p.ldy = ldy;
  p.KT = Kp / 64;
  p.T = (Np / 128) * p.KT;
  p.stages = stages;
  const int grid = static_cast<int>(p.T < sm_count() ? p.T : sm_count());
  const int mt = M <= 8 ? 1 : 2;
#define AO_LAUNCH(GG) \
  return mt == 1 ? launch_stream<GG, 1>(p, grid, pdl, s) : launch_stream<GG, 2>(p, grid, pdl, s)
  switch (G) {
    case 32: AO_LAUNCH(32);
    case 64: AO_LAUNCH(64);
    case 128: AO_LAUNCH(128);
    case 256: AO_LAUNCH(256);
  }
#undef AO_LAUNCH
  return AO_B200_ERR_UNSUPPORTED;
}

}  // namespace ao
