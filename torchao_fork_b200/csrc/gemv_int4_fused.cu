// int4 weight-only linear, decode path (M <= 16) — v3: strip-parallel, TMA-bulk ring, deferred-affine dequant,
// fused RMSNorm prologue and residual / SiLU*mul epilogues, PDL-chained.
//
//   y[m][n] = sum_g [ s[g][n] * sum_{k in g} x[m][k] * (q[n][k] - 8)  +  z[g][n] * sum_{k in g} x[m][k] ]
//
// i.e. the affine dequant (q-8)*s+z is applied ONCE PER GROUP in fp32 to the tensor-core partial sum instead of
// once per weight in bf16 (that is the arithmetic of the reference's CPU op _weight_int4pack_mm_for_cpu, which
// dequantises in fp32 without rounding W' to bf16 — SURVEY.md §8 a15; the CUDA op rounds W' to bf16 first, a10;
// the two differ by the bf16 rounding of each W', well inside the stated GEMM tolerance).  What is left per packed
// word is 3 shifts + 4 lop3 (nibble -> bf16 value 128+q with the 0x4300 magic) + one mma.sync: half the issue
// slots of the exact per-weight bf16 fma, which is what made v2 ALU-bound at ~54 % of HBM peak.
//     sum x*(q-8) = sum x*(128+q) - 136 * X_g,   X_g = sum_{k in g} x[k]  (computed once per CTA from the staged x)
//
// Decomposition: a work item is one 16-row strip over the whole K; CTA b owns strips [b*S/B, (b+1)*S/B).  No
// cross-CTA reduction, no workspace, deterministic.  The layout is strip-major, so the CTA's share is ONE contiguous
// byte range of the weight (and of the scale/zero tensor): the producer issues one 16 KB cp.async.bulk plus one
// <= 4 KB scale/zero copy per stage into a shared-memory ring (many small bulk copies measured ~40 ns each and
// lost DRAM page locality: 20 % of HBM peak); 8 consumer warps split every 32-k-tile stage 4 k-tiles each, and one (rotating) warp sums the 8 partials of
// a finished strip and runs the epilogue while the others go on.  The ring is sized so two CTAs of consecutive
// kernels fit on one SM: with programmatic dependent launch the next GEMV's weights are already in shared memory
// when the previous kernel retires (weights never depend on the previous kernel; x, y do -> griddepcontrol.wait).
#include <cuda_fp16.h>

#include <cstdlib>
#include <type_traits>

#include "common.cuh"

#ifndef AO_CHAINS
#define AO_CHAINS 2
#endif
#ifndef AO_STAGE_TILES
#define AO_STAGE_TILES 32
#endif
#ifndef AO_EXP
#define AO_EXP 0  // benchmarks/gemv_trace.cu ablations only
#endif

namespace ao {

constexpr int kFConsumers = 16;
constexpr int kFThreads = (kFConsumers + 2) * 32;  // + TMA producer warp + epilogue warp
constexpr int kRedBufs = 2;                       // partial-sum buffers between the consumers and the epilogue warp
constexpr int kStageTiles = AO_STAGE_TILES;  // k-tiles (of 64 k) per stage: 16 KB of weights per 32
constexpr int kTilesPerWarp = kStageTiles / kFConsumers;  // 2 k-tiles = 128 k per warp per stage

enum { kEpiPlain = 0, kEpiResidual = 1, kEpiSiluMul = 2 };

// AO_GEMV_TRACE (benchmarks/gemv_trace.cu only): per-warp, per-stage SM clock stamps of one CTA
#ifdef AO_GEMV_TRACE
constexpr int kTraceIts = 256, kTraceKinds = 6;
__device__ long long ao_trace[kTraceKinds][kFConsumers + 2][kTraceIts];
__device__ int ao_trace_cta = 0;
#define AO_TRACE(kind, it)                                                                  \
  do {                                                                                      \
    if ((int)blockIdx.x == ao_trace_cta && (threadIdx.x & 31) == 0 && (it) < kTraceIts)     \
      ao_trace[kind][threadIdx.x >> 5][it] = clock64();                                     \
  } while (0)
#else
#define AO_TRACE(kind, it) do { } while (0)
#endif

struct GemvFusedParams {
  const uint8_t* packed;      // strip-major [Np/16][Kp/64][512 B]
  const uint32_t* sz;         // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* x;     // [M][ldx]
  const __nv_bfloat16* norm_w;  // [Kx] or null: x_eff = bf16(x * norm_w), result scaled by rsqrt(mean(x^2)+eps)
  const __nv_bfloat16* residual;  // [M][ld_res] or null (kEpiResidual)
  __nv_bfloat16* y;
  int M, stages, epilogue, copy_split;
  int KT;                     // Kp / 64
  int64_t Kx, ldx, Np, N_out, ldy, ld_res;
  int n_strips;               // Np / 16
  int strips_base, strips_rem;  // n_strips / grid, n_strips % grid
  float eps;
  uint32_t x_row_bytes;       // 2 * Kp + 16
  uint32_t off_x, off_X, off_red, off_misc;  // shared-memory carve-up (bytes)
};

template <int G>
__host__ __device__ constexpr int stage_sz_bytes() { return (kStageTiles * 64 / G) * 64; }
template <int G>
__host__ __device__ constexpr int stage_bytes() { return kStageTiles * 512 + stage_sz_bytes<G>(); }

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// D(16x8, f32) += A(16x16, f16, row) * B(16x8, f16, col)
__device__ __forceinline__ void mma_f16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                              uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// (w & mask) | 0x64006400: a nibble pair OR-ed into the mantissa of fp16 1024.0 (ulp 1, 10 mantissa bits):
// mask 0x000f000f -> 1024 + q, mask 0x00f000f0 -> 1024 + 16 q.  Two of the four nibble alignments of a word need no
// shift at all and the other two share one (w >> 8): 5 ALU-pipe ops per 8 weights (bf16's 7-bit mantissa needs 7).
__device__ __forceinline__ uint32_t nib_f16_lo(uint32_t w) {
  uint32_t h;
  asm("lop3.b32 %0, %1, 0x000f000f, 0x64006400, 0xea;" : "=r"(h) : "r"(w));
  return h;
}
__device__ __forceinline__ uint32_t nib_f16_hi(uint32_t w) {
  uint32_t h;
  asm("lop3.b32 %0, %1, 0x00f000f0, 0x64006400, 0xea;" : "=r"(h) : "r"(w));
  return h;
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

template <int G, int MT>
__global__ void __launch_bounds__(kFThreads, (MT == 1 ? 2 : 1)) int4_gemv_fused_kernel(const GemvFusedParams p) {
  constexpr int kStageB = stage_bytes<G>();
  constexpr int SEG = G < 128 ? G : 128;  // the correction sums are kept per min(G, 128) k: a warp owns 128 k of a stage
  constexpr int U32 = SEG / 32;           // 32-k halves per accumulation segment
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* x_s = smem + p.off_x;                                // [M][x_row_bytes]: bf16 while staging, then scaled fp16
  float* X_s = reinterpret_cast<float*>(smem + p.off_X);        // [Kp/SEG][2 = {X, C}][8*MT]
  float* red = reinterpret_cast<float*>(smem + p.off_red);      // [2][16 warps][MT][32][4]
  float* part = reinterpret_cast<float*>(smem + p.off_misc);    // [16 rows][16 warps][2 = {sum sq, amax}]
  float* row_scale = part + 16 * kFConsumers * 2;               // [16]: 2^e (* rsqrt(mean sq + eps))
  float* row_dn = row_scale + 16;                               // [16]: 2^-e
  uint64_t* full = reinterpret_cast<uint64_t*>(row_dn + 16);
  uint64_t* empty = full + p.stages;
  uint64_t* red_full = empty + p.stages;                        // [2]
  uint64_t* red_free = red_full + kRedBufs;                            // [2]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  AO_TRACE(5, 0);
  const int stages = p.stages;
  const int b = blockIdx.x;
  // strips [s0, s1) of this CTA: base = n_strips / grid each, the first `rem` CTAs one more (host-computed: no division)
  const int s0 = b * p.strips_base + min(b, p.strips_rem), s1 = s0 + p.strips_base + (b < p.strips_rem ? 1 : 0);
  const int KT = p.KT;
  const int n_chunks = (KT + kStageTiles - 1) / kStageTiles;

  // The producer warp initialises the mbarriers and starts streaming at once; everybody else meets it on named
  // barrier 2 only when they first need an mbarrier (after the activation staging), so the first stages are in
  // flight while the kernel is still waiting for its predecessor (PDL) and staging x.
  grid_launch_dependents();
  AO_TRACE(5, 1);

  if (warp == kFConsumers) {
    // ===================== producer: lane 0 owns the barriers; `copy_split` lanes share the weight copy ===============
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], kFConsumers);
      }
      for (int s = 0; s < kRedBufs; ++s) {
        mbar_init(&red_full[s], kFConsumers);
        mbar_init(&red_free[s], 1);
      }
      fence_barrier_init();
    }
    __syncwarp();
    named_bar_arrive(2, kFThreads);
    {
      int slot = 0;
      uint32_t par = 1;
      const int n_groups = KT * 64 / G;
      const int split = p.copy_split;
      for (int strip = s0; strip < s1; ++strip) {
        const uint8_t* wrow = p.packed + (int64_t)strip * KT * 512;
        const uint32_t* srow = p.sz + (int64_t)strip * n_groups * 16;
        for (int ch = 0; ch < n_chunks; ++ch) {
          const int kt0 = ch * kStageTiles;
          const int cnt = min(kStageTiles, KT - kt0);
          const int n_grp = cnt * 64 / G;
          uint8_t* wdst = ring + (size_t)slot * kStageB;
          if (lane == 0) {
            AO_TRACE(0, (strip - s0) * n_chunks + ch);
            mbar_wait(&empty[slot], par);
            AO_TRACE(1, (strip - s0) * n_chunks + ch);
            mbar_arrive_expect_tx(&full[slot], (uint32_t)(cnt * 512 + n_grp * 64));
            bulk_g2s(wdst + kStageTiles * 512, srow + (int64_t)(kt0 * 64 / G) * 16, (uint32_t)(n_grp * 64), &full[slot]);
          }
          __syncwarp();
          if (lane < split) {
            const int per = (cnt + split - 1) / split;           // k-tiles per piece
            const int t0 = lane * per, t1 = min(cnt, t0 + per);
            if (t1 > t0)
              bulk_g2s(wdst + t0 * 512, wrow + (int64_t)(kt0 + t0) * 512, (uint32_t)((t1 - t0) * 512), &full[slot]);
          }
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

  if (warp == kFConsumers + 1) {
    // ===================== epilogue warp: sums the 16 partials of every finished strip, applies the row scale and
    //                       the residual / SiLU*mul epilogue, stores =====================
    const int g = lane >> 2, c = lane & 3;
    grid_dependency_wait();
    named_bar_sync(2, kFThreads);  // mbarriers initialised
    for (int strip = s0; strip < s1; ++strip) {
      const int li = strip - s0;
      const int rbuf = li % kRedBufs;
      const uint32_t use = (uint32_t)(li / kRedBufs);
      const float* rb = red + (size_t)rbuf * (kFConsumers * MT * 128);
      mbar_wait(&red_full[rbuf], use & 1u);
      AO_TRACE(4, li);
      float4 sum[MT];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        // warp order, packed fp32x2 adds: the same sums with half the instructions on the launch's tail
        uint64_t slo = 0ull, shi = 0ull;
#pragma unroll
        for (int w = 0; w < kFConsumers; ++w) {
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
        const float vals[2][2] = {{sum[mt].x, sum[mt].z}, {sum[mt].y, sum[mt].w}};  // [col j][row half]: rows g, g+8
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int m = 8 * mt + 2 * c + j;
          if (m >= p.M) continue;
          const float rsc = row_scale[m];
          const float v0 = vals[j][0] * rsc, v1 = vals[j][1] * rsc;
          if (p.epilogue == kEpiSiluMul) {
            // rows g = gate (w1), rows g+8 = up (w3) of output index 8*strip + g  (model.py:486)
            const int64_t n = 8 * (int64_t)strip + g;
            if (2 * n < p.N_out) {
              const float gate = bf16r(v0), up = bf16r(v1);
              const float sg = bf16r(gate / (1.f + __expf(-gate)));
              p.y[m * p.ldy + n] = __float2bfloat16_rn(sg * up);
            }
          } else {
            const int64_t n0 = 16 * (int64_t)strip + g;
            if (p.epilogue == kEpiResidual) {
              if (n0 < p.N_out)
                p.y[m * p.ldy + n0] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0]) + bf16r(v0));
              if (n0 + 8 < p.N_out)
                p.y[m * p.ldy + n0 + 8] =
                    __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n0 + 8]) + bf16r(v1));
            } else {
              if (n0 < p.N_out) p.y[m * p.ldy + n0] = __float2bfloat16_rn(v0);
              if (n0 + 8 < p.N_out) p.y[m * p.ldy + n0 + 8] = __float2bfloat16_rn(v1);
            }
          }
        }
      }
      AO_TRACE(5, li);
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, c = lane & 3;
  const int tid = threadIdx.x;  // 0 .. 511
  constexpr int kCT = kFConsumers * 32;
  grid_dependency_wait();
  AO_TRACE(5, 2);

  // ---- stage the activations: x (* norm_w) -> per-row power-of-two scaling -> fp16, the entries that meet a
  //      "1024 + 16 q" weight pre-divided by 16; then the per-segment correction sums ----
  const int chunks_per_row = KT * 8;  // 16-byte chunks of 8 elements over Kp
  constexpr int CPS = SEG / 8;  // chunks per segment: 4, 8 or 16
  {
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
        const float ss = warp_sum_f(ss4[j]), amax = warp_max_f(am4[j]);
        if (lane == 0 && m0 + j < p.M) {
          part[((m0 + j) * kFConsumers + warp) * 2] = ss;
          part[((m0 + j) * kFConsumers + warp) * 2 + 1] = amax;
        }
      }
    }
    named_bar_sync(1, kCT);
    AO_TRACE(5, 3);
    if (warp < p.M) {
      // warp w finishes row w: 16 (sum sq, amax) partials -> 2^e and the epilogue scale
      float t = 0.f, amax = 0.f;
      if (lane < kFConsumers) {
        t = part[(warp * kFConsumers + lane) * 2];
        amax = part[(warp * kFConsumers + lane) * 2 + 1];
      }
      t = warp_sum_f(t);
      amax = warp_max_f(amax);
      if (lane == 0) {
        // 2^e: brings the row's largest magnitude to [2^13, 2^14): every bf16 value within 2^-27 of it is exact in fp16
        int e = 0;
        if (amax > 0.f) e = (int)((__float_as_uint(amax) >> 23) & 0xFF) - 127 - 13;
        e = max(-100, min(100, e));
        const float up = __uint_as_float((uint32_t)(e + 127) << 23), dn = __uint_as_float((uint32_t)(127 - e) << 23);
        row_dn[warp] = dn;
        row_scale[warp] = (p.norm_w != nullptr) ? up * rsqrtf(t / (float)p.Kx + p.eps) : up;
      }
    }
    named_bar_sync(1, kCT);
    AO_TRACE(5, 4);
    // rescale to fp16 and, in the same pass, the per-segment correction sums: a segment is SEG / 8 consecutive chunks
    // = SEG / 8 consecutive lanes.  X[seg][m] = sum of the scaled activations, C[seg][m] = 1024 * (sum of the STORED
    // values) + 8 * X, stored negated:  sum_k x_k (q_k - 8) = (MMA result) - C  ->  the accumulator starts at -C.
    // (Columns m >= M of X_s stay unwritten: they only feed MMA columns that are never stored.)
    for (int m = 0; m < p.M; ++m) {
      const float dn = row_dn[m], dn16 = dn * 0.0625f;
      for (int base = 0; base < chunks_per_row; base += kCT) {
        const int ci = base + tid;
        float xs = 0.f, xt = 0.f;
        if (ci < chunks_per_row) {
          uint4* ptr = reinterpret_cast<uint4*>(x_s + (size_t)m * p.x_row_bytes + (size_t)ci * 16);
          float f[8];
          unpack8(*ptr, f);
          // element pairs (0,1),(4,5) meet "1024 + q" weights, pairs (2,3),(6,7) meet "1024 + 16 q" weights
          const __half2 h0 = __floats2half2_rn(f[0] * dn, f[1] * dn), h1 = __floats2half2_rn(f[2] * dn16, f[3] * dn16);
          const __half2 h2 = __floats2half2_rn(f[4] * dn, f[5] * dn), h3 = __floats2half2_rn(f[6] * dn16, f[7] * dn16);
          *ptr = make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                            *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
          const float2 a0 = __half22float2(h0), a1 = __half22float2(h1), a2 = __half22float2(h2), a3 = __half22float2(h3);
          const float plain = (a0.x + a0.y) + (a2.x + a2.y), div16 = (a1.x + a1.y) + (a3.x + a3.y);
          xs = plain + div16;
          xt = fmaf(16.f, div16, plain);
        }
#pragma unroll
        for (int o = CPS / 2; o; o >>= 1) {
          xs += __shfl_xor_sync(0xffffffffu, xs, o);
          xt += __shfl_xor_sync(0xffffffffu, xt, o);
        }
        if (ci < chunks_per_row && (ci % CPS) == 0) {
          const int sg = ci / CPS;
          X_s[(sg * 2 + 0) * 8 * MT + m] = xt;
          X_s[(sg * 2 + 1) * 8 * MT + m] = -fmaf(1024.f, xs, 8.f * xt);
        }
      }
    }
    named_bar_sync(1, kCT);
    AO_TRACE(5, 5);
    AO_TRACE(5, 6);
  }
  named_bar_sync(2, kFThreads);  // the producer has initialised the mbarriers (long ago)

  // ---- main loop.  Everything a warp touches in a stage sits at a FIXED offset from the stage base (its two
  //      k-tiles, its scale/zero group(s)) or from a per-chunk base (its 128 activations, its correction sums): all
  //      addresses are 32-bit shared-window addresses + immediates, so the loop body is loads, 5 ALU ops per packed
  //      word, one mma.sync per word and the per-segment fix-up — nothing else. ----
  constexpr uint32_t XS = 2 * 8 * MT * 4;                          // bytes of one segment's {X, -C} record
  constexpr uint32_t kXChunk = (kStageTiles * 64 / SEG) * XS;      // correction records per stage
  const int tl0 = warp * kTilesPerWarp;                            // this warp's first k-tile inside a stage
  const uint32_t ring_u = smem_u32(ring);
  const uint32_t w_off = (uint32_t)tl0 * 512u + (uint32_t)lane * 16u;
  const uint32_t sz_off = kStageTiles * 512u + (uint32_t)(((tl0 * 64) / G) * 16 + g) * 4u;
  const uint32_t X_u0 = smem_u32(X_s) + (uint32_t)((tl0 * 64) / SEG) * XS + (uint32_t)c * 8u;
  // activation fragments: only the lanes whose MMA column is a real row (g + 8 mt < M) load; the others keep zeros
  // (their columns are never stored) — for M = 1 that is one shared-memory wavefront per load instead of four
  uint32_t x_u0[MT];
  bool x_on[MT];
  uint4 xx[MT];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    x_on[mt] = g + 8 * mt < p.M;
    x_u0[mt] = smem_u32(x_s) + (uint32_t)min(g + 8 * mt, p.M - 1) * p.x_row_bytes + (uint32_t)c * 16u + (uint32_t)tl0 * 128u;
    xx[mt] = make_uint4(0, 0, 0, 0);
  }
  // keep the per-lane offsets in registers (opaque to the compiler, which otherwise re-derives them from %tid in
  // every stage: ~12 extra ALU-pipe instructions per stage on the pipe that already carries the dequant)
  uint32_t w_off_r = w_off, sz_off_r = sz_off, X_u0_r = X_u0;
  asm volatile("" : "+r"(w_off_r), "+r"(sz_off_r), "+r"(X_u0_r));
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) asm volatile("" : "+r"(x_u0[mt]));
  const int n_full = KT / kStageTiles;  // chunks [0, n_full) are complete stages
  int slot = 0;
  uint32_t par = 0;
  uint32_t stage_u = ring_u;

  // one stage of one warp: two k-tiles = four 32-k halves; FULL = every tile of the stage is valid
  auto consume = [&](auto full_tag, const int cnt, const uint32_t xo, const uint32_t X_u, float (&yacc)[MT][4]) {
    constexpr bool FULL = decltype(full_tag)::value;
    uint4 wv[kTilesPerWarp];
#pragma unroll
    for (int t = 0; t < kTilesPerWarp; ++t) wv[t] = lds128(stage_u + w_off_r + t * 512u);
    // kChains independent accumulators per segment: back-to-back mma.sync on ONE accumulator serialise on the HMMA
    // latency (~50 clk under load: removing the MMAs altogether took the kernel from 82 to 56 us)
    constexpr int kChains = (2 * U32 < AO_CHAINS) ? 2 * U32 : AO_CHAINS;
    float acc[kChains][MT][4];
#pragma unroll
    for (int t = 0; t < kTilesPerWarp; ++t) {
      if (!FULL && tl0 + t >= cnt) continue;
      const uint32_t words[4] = {wv[t].x, wv[t].y, wv[t].z, wv[t].w};
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int q32 = 2 * t + h;       // 32-k half inside the warp's share of the stage
        const int si = q32 / U32;        // segment inside the warp's share
        const int j0 = (2 * (q32 % U32)) % kChains, j1 = (2 * (q32 % U32) + 1) % kChains;
        const uint32_t wl = words[2 * h], wh = words[2 * h + 1];  // rows g and g + 8
        const uint32_t wl8 = wl >> 8, wh8 = wh >> 8;
#if AO_EXP == 2
        const uint32_t a0l = wl, a1l = wl, a2l = wl8, a3l = wl8, a0h = wh, a1h = wh, a2h = wh8, a3h = wh8;
#else
        const uint32_t a0l = nib_f16_lo(wl), a1l = nib_f16_hi(wl), a2l = nib_f16_lo(wl8), a3l = nib_f16_hi(wl8);
        const uint32_t a0h = nib_f16_lo(wh), a1h = nib_f16_hi(wh), a2h = nib_f16_lo(wh8), a3h = nib_f16_hi(wh8);
#endif
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          if (x_on[mt]) xx[mt] = lds128(x_u0[mt] + xo + q32 * 64);
          if (2 * (q32 % U32) < kChains) {
            if (j0 == 0) {  // chain 0 starts at -C, so the chains sum to  sum_k x_k (q_k - 8)
              const float2 nc = lds64f(X_u + si * XS + 8 * MT * 4 + mt * 32);
              acc[0][mt][0] = nc.x; acc[0][mt][1] = nc.y; acc[0][mt][2] = nc.x; acc[0][mt][3] = nc.y;
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) acc[j0][mt][e] = 0.f;
            }
          }
          if (2 * (q32 % U32) + 1 < kChains) {
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[j1][mt][e] = 0.f;
          }
#if AO_EXP == 1
          asm volatile("" ::"r"(a0l), "r"(a0h), "r"(a1l), "r"(a1h), "r"(a2l), "r"(a2h), "r"(a3l), "r"(a3h), "r"(xx[mt].x),
                       "r"(xx[mt].y), "r"(xx[mt].z), "r"(xx[mt].w));
#else
          mma_f16_16816(acc[j0][mt], a0l, a0h, a1l, a1h, xx[mt].x, xx[mt].y);
          mma_f16_16816(acc[j1][mt], a2l, a2h, a3l, a3h, xx[mt].z, xx[mt].w);
#endif
        }
        if (q32 % U32 == U32 - 1) {
          // segment complete: y += s * acc + z * X   (fp32)
          const uint32_t ga = stage_u + sz_off_r + (uint32_t)((si * SEG) / G) * 64u;
          const uint32_t p0 = lds32(ga), p1 = lds32(ga + 32u);
          const float sc0 = __uint_as_float(p0 << 16), sc1 = __uint_as_float(p1 << 16);
          const float z0 = __uint_as_float(p0 & 0xFFFF0000u), z1 = __uint_as_float(p1 & 0xFFFF0000u);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            float a[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              a[e] = acc[0][mt][e];
#pragma unroll
              for (int j = 1; j < kChains; ++j) a[e] += acc[j][mt][e];
            }
            const float2 X = lds64f(X_u + si * XS + mt * 32);
            yacc[mt][0] = fmaf(sc0, a[0], fmaf(z0, X.x, yacc[mt][0]));
            yacc[mt][1] = fmaf(sc0, a[1], fmaf(z0, X.y, yacc[mt][1]));
            yacc[mt][2] = fmaf(sc1, a[2], fmaf(z1, X.x, yacc[mt][2]));
            yacc[mt][3] = fmaf(sc1, a[3], fmaf(z1, X.y, yacc[mt][3]));
          }
        }
      }
    }
  };

  for (int strip = s0; strip < s1; ++strip) {
    float yacc[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int e = 0; e < 4; ++e) yacc[mt][e] = 0.f;

    uint32_t xo = 0, X_u = X_u0_r;
    for (int ch = 0; ch < n_chunks; ++ch) {
      AO_TRACE(0, (strip - s0) * n_chunks + ch);
#if AO_EXP != 3 && AO_EXP != 5
      mbar_wait(&full[slot], par);
#endif
      AO_TRACE(1, (strip - s0) * n_chunks + ch);
      if (ch < n_full) {
        consume(std::true_type{}, kStageTiles, xo, X_u, yacc);
      } else if (tl0 < KT - ch * kStageTiles) {
        consume(std::false_type{}, KT - ch * kStageTiles, xo, X_u, yacc);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      AO_TRACE(2, (strip - s0) * n_chunks + ch);
      xo += kStageTiles * 128u;
      X_u += kXChunk;
      stage_u += kStageB;
      if (++slot == stages) {
        slot = 0;
        par ^= 1u;
        stage_u = ring_u;
      }
    }

    // ---- strip done: the 16 partials go to shared memory and the warp goes straight on to the next strip; the
    //      epilogue warp sums them.  (A consumer doubling as the reducer falls behind by the reduction, the next
    //      reducer has to wait for it, and the lag chains: strip time became compute + reduce, 2700 clk.) ----
#if AO_EXP == 4 || AO_EXP == 5
    if (yacc[0][0] == 123.456f) p.y[strip] = __float2bfloat16_rn(yacc[0][1] + yacc[0][2] + yacc[0][3]);
    continue;
#endif
    const int li = strip - s0;
    const int rbuf = li % kRedBufs;
    const uint32_t use = (uint32_t)(li / kRedBufs);  // n-th use of this buffer
    float* rb = red + (size_t)rbuf * (kFConsumers * MT * 128);
    if (use > 0) mbar_wait(&red_free[rbuf], (use - 1) & 1u);  // the epilogue of kRedBufs strips ago has read it
    AO_TRACE(3, li);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
      *reinterpret_cast<float4*>(rb + ((warp * MT + mt) * 32 + lane) * 4) =
          make_float4(yacc[mt][0], yacc[mt][1], yacc[mt][2], yacc[mt][3]);
    __syncwarp();
    if (lane == 0) mbar_arrive(&red_full[rbuf]);
  }
}

// ------------------------------------------------------------------------------------------------ host side
static int fused_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n = v > 0 ? v : kNumSMs;
  }
  return n;
}

struct FusedPlan {
  int stages;
  size_t smem;
  uint32_t off_x, off_X, off_red, off_misc, x_row_bytes;
};

// shared-memory plan; returns false when the activations do not fit (caller falls back to the strip kernel)
template <int G>
static bool plan_fused(int M, int MT, int64_t Kp, int max_stages, FusedPlan& pl) {
  const size_t stage_b = stage_bytes<G>();
  pl.x_row_bytes = (uint32_t)(2 * Kp + 16);
  const size_t x_bytes = ((size_t)M * pl.x_row_bytes + 127) / 128 * 128;
  const size_t X_bytes = ((size_t)(Kp / (G < 128 ? G : 128)) * 2 * 8 * MT * 4 + 127) / 128 * 128;
  const size_t red_bytes = (size_t)kRedBufs * kFConsumers * MT * 128 * 4;
  const size_t misc = (16 * kFConsumers * 2 + 32) * 4 + (2 * 16 + 2 * kRedBufs) * 8 + 64;
  const size_t fixed = x_bytes + X_bytes + red_bytes + misc;
  static const size_t two_per_sm = []() {  // AO_B200_SMEM_KB: shared-memory budget per CTA (default: two CTAs per SM)
    const char* e = std::getenv("AO_B200_SMEM_KB");
    return (size_t)(e ? std::atoi(e) : 113) * 1024;
  }();
  const size_t one_per_sm = 226 * 1024;
  int st = 0;
  static const size_t min2 = []() {  // fewest ring stages for which two co-resident CTAs still beat one with a deep ring
    const char* e = std::getenv("AO_B200_MIN_STAGES2");
    return (size_t)(e ? std::atoi(e) : 4);  // 3 measured slower on w2 (K = 14336): 394 vs 350 us per 32 layers
  }();
  if (fixed + min2 * stage_b <= two_per_sm) {
    st = (int)((two_per_sm - fixed) / stage_b);
  } else if (fixed + 3 * stage_b <= one_per_sm) {
    st = (int)((one_per_sm - fixed) / stage_b);
  } else {
    return false;
  }
  if (st > max_stages) st = max_stages;
  if (st > 16) st = 16;
  pl.stages = st;
  size_t off = (size_t)st * stage_b;
  pl.off_x = (uint32_t)off; off += x_bytes;
  pl.off_X = (uint32_t)off; off += X_bytes;
  pl.off_red = (uint32_t)off; off += red_bytes;
  pl.off_misc = (uint32_t)off; off += misc;
  pl.smem = off;
  return true;
}

template <int G, int MT>
static int launch_fused(GemvFusedParams& p, const FusedPlan& pl, bool pdl, cudaStream_t s) {
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(int4_gemv_fused_kernel<G, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) !=
        cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  // AO_B200_CTAS: co-resident CTAs per SM of ONE launch (the 113 KB plan leaves room for two)
  static const int ctas_per_sm = []() { const char* e = std::getenv("AO_B200_CTAS"); int v = e ? std::atoi(e) : 2; return v < 1 ? 1 : v; }();
  const int max_grid = fused_sm_count() * (pl.smem <= 113 * 1024 && MT == 1 ? ctas_per_sm : 1);
  const int grid = p.n_strips < max_grid ? p.n_strips : max_grid;
  p.strips_base = p.n_strips / grid;
  p.strips_rem = p.n_strips % grid;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, int4_gemv_fused_kernel<G, MT>, p) != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

// returns AO_B200_ERR_WORKSPACE when the shape does not fit this kernel (caller may fall back)
int int4_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                    int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps,
                    const void* residual, int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s) {
  GemvFusedParams p;
  p.packed = static_cast<const uint8_t*>(packed);
  p.sz = static_cast<const uint32_t*>(sz);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.norm_w = static_cast<const __nv_bfloat16*>(norm_w);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.epilogue = epilogue;
  p.KT = (int)(Kp / 64);
  p.Kx = Kx; p.ldx = ldx; p.Np = Np; p.N_out = N_out; p.ldy = ldy; p.ld_res = ld_res;
  p.n_strips = (int)(Np / 16);
  p.eps = eps;
  const int MT = M <= 8 ? 1 : 2;
  FusedPlan pl;
  bool ok = false;
  switch (G) {
    case 32: ok = plan_fused<32>((int)M, MT, Kp, max_stages, pl); break;
    case 64: ok = plan_fused<64>((int)M, MT, Kp, max_stages, pl); break;
    case 128: ok = plan_fused<128>((int)M, MT, Kp, max_stages, pl); break;
    case 256: ok = plan_fused<256>((int)M, MT, Kp, max_stages, pl); break;
    default: return AO_B200_ERR_UNSUPPORTED;
  }
  if (!ok) return AO_B200_ERR_WORKSPACE;
  p.stages = pl.stages;
  {
    static const int cs = []() { const char* e = std::getenv("AO_B200_COPY_SPLIT"); int v = e ? std::atoi(e) : 1; return v < 1 ? 1 : (v > 32 ? 32 : v); }();
    p.copy_split = cs;
  }
  p.x_row_bytes = pl.x_row_bytes;
  p.off_x = pl.off_x; p.off_X = pl.off_X; p.off_red = pl.off_red; p.off_misc = pl.off_misc;
#define AO_LF(GG) return MT == 1 ? launch_fused<GG, 1>(p, pl, pdl, s) : launch_fused<GG, 2>(p, pl, pdl, s)
  switch (G) {
    case 32: AO_LF(32);
    case 64: AO_LF(64);
    case 128: AO_LF(128);
    case 256: AO_LF(256);
  }
#undef AO_LF
  return AO_B200_ERR_UNSUPPORTED;
}

}  // namespace ao
