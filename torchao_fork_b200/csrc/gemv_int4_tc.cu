// int4 weight-only linear, decode path (M <= 5) on the 5th-generation tensor core: tcgen05.mma kind::i8 with the
// A operand (weights) written to TENSOR MEMORY by the dequant warps and the accumulator read back per quantization
// group.
//
//   y[m][n] = sum_g [ s[g][n] * sum_{k in g} x[m][k] (q[n][k] - 8)  +  z[g][n] * sum_{k in g} x[m][k] ]
//
// Integer formulation: every activation row is turned into 24-bit fixed point relative to its largest magnitude,
// x = 2^e * xi, and xi is split into three balanced base-256 digits (int8 "limbs").  The limbs of the M rows are the
// N <= 16 columns of the MMA, the unsigned nibbles q are its A operand: D[n][3m + l] = sum_k q[n][k] * limb_l(xi[m][k])
// is EXACT in int32, and  sum_k x (q - 8) = 2^e ( D2 * 65536 + D1 * 256 + D0 - 8 * sum_k xi ).  What is left per packed
// word (8 weights) on the CUDA cores is one shift and two lop3 (nibbles -> bytes), a quarter of a shared-memory load and
// a quarter of a tensor-memory store: ~4 issue slots against ~14 for the mma.sync kernel (gemv_int4_fused.cu), whose
// ALU pipe is what caps it at ~70 % of HBM peak.  (Measured, benchmarks/tc_probe.cu: kind::i8 M128 N<=64 K32 with A in
// TMEM retires one MMA per 45 clocks = 2 KB of int4 per 45 clocks per SM = 13 TB/s; 16x128b tensor-memory stores run at
// 480 B/clk/SM.)
//
// The packed layout is the strip-major one unchanged: the 16-byte lane chunk of a (16-row strip, 64-k tile) holds
// rows g and g + 8, which is exactly the (lane, column) pattern of a tcgen05.st.16x128b, so a 128-row block is 8 strips
// stored to TMEM lanes 16 s .. 16 s + 15.  The k order inside a 32-k block that this produces is baked into the staging
// of the activations (the B operand in shared memory, canonical no-swizzle K-major layout).
//
// Roles: 16 dequant warps (quadrant = warp % 4 owns two strips, warp / 4 picks the k-tiles), 4 epilogue warps
// (thread = output row: tcgen05.ld per group, int -> float, scale/zero fix-up, final store), 1 TMA producer warp,
// 1 MMA-issuing warp (also owns the TMEM allocation).
#include <cuda_fp16.h>

#include <cstdlib>

#include "common.cuh"

namespace ao {

constexpr int kTcDequantWarps = 16, kTcEpiWarps = 4;
constexpr int kTcProducerWarp = kTcDequantWarps + kTcEpiWarps, kTcMmaWarp = kTcProducerWarp + 1;
constexpr int kTcThreads = (kTcMmaWarp + 1) * 32;  // 704
constexpr int kTcStageTiles = 8;                   // 64-k tiles per stage: 512 k, 8 strips x 4 KB = 32 KB of weights
constexpr int kTcStageK = kTcStageTiles * 64;
constexpr int kTcABufs = 3;                        // TMEM A ring: 128 columns (512 k of int8) per buffer
constexpr int kTcDBufs = 8;                        // TMEM accumulator ring: N columns per buffer
constexpr int kTcLimbs = 3;

#ifdef AO_TC_TRACE
constexpr int kTcTraceIts = 128;
__device__ long long ao_tc_trace[6][kTcMmaWarp + 1][kTcTraceIts];
#define TC_TRACE(kind, it)                                                              \
  do {                                                                                  \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (it) < kTcTraceIts)               \
      ao_tc_trace[kind][threadIdx.x >> 5][it] = clock64();                              \
  } while (0)
#else
#define TC_TRACE(kind, it) do { } while (0)
#endif

struct GemvTcParams {
  const uint8_t* packed;      // strip-major [Np/16][Kp/64][512 B]
  const uint32_t* sz;         // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* x;     // [M][ldx]
  const __nv_bfloat16* norm_w;
  const __nv_bfloat16* residual;
  __nv_bfloat16* y;
  int M, N_mma, stages, epilogue;
  int KT;                     // Kp / 64
  int64_t Kx, ldx, N_out, ldy, ld_res;
  int n_blocks;               // Np / 128 (row blocks of 8 strips)
  int n_chunks;               // stages (512 k) per row block
  int units_base, units_rem;  // work units (row block, stage) per CTA: base, and one more for the first `rem` CTAs
  int win;                    // activation window of a CTA: min(n_chunks, units_base + 1) stages
  float eps;
  uint32_t* tickets;          // [n_blocks][4 quadrants], zero between launches
  float* partials;            // [grid][2][5][128]
  uint32_t off_x, off_sx, off_misc, off_tmp;  // shared-memory carve-up (bytes)
};

// first work unit of CTA c / CTA that owns unit v (units are dealt out contiguously, row-block major)
__host__ __device__ __forceinline__ int tc_first_unit(int c, int base, int rem) { return c * base + (c < rem ? c : rem); }
__host__ __device__ __forceinline__ int tc_cta_of_unit(int v, int base, int rem) {
  const int big = rem * (base + 1);
  return v < big ? v / (base + 1) : rem + (v - big) / base;
}

template <int G>
__host__ __device__ constexpr int tc_stage_sz_bytes() { return 8 * (kTcStageK / G) * 64; }
template <int G>
__host__ __device__ constexpr int tc_stage_bytes() { return 8 * kTcStageTiles * 512; }  // weights only: scale/zero by LDG

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D(tmem, s32) (+)= A(tmem, u8 [128][32]) * B(smem, s8 [N][32])
__device__ __forceinline__ void tc_mma_i8(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_st_16x128b_x4(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.16x128b.x4.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tc_ld_32x32b_x8(uint32_t taddr, uint32_t (&d)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tc_ld_32x32b_x16(uint32_t taddr, uint32_t (&d)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]), "=r"(d[9]),
        "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
      : "r"(taddr));
}

// one lane of a converged warp (the compiler keeps the surrounding code on the uniform datapath, unlike `lane == 0`)
__device__ __forceinline__ bool tc_elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\tselp.u32 %0, 1, 0, px;\n\t}" : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ float tc_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float tc_warp_max(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// position of logical k (inside a 32-k block) in the B operand: chunk r' (16 k of the MMA's 32) and byte inside it.
// The A side: word nibble i holds k0 + {0,2,4,6,1,3,5,7}[i]; (w & 0x0f0f0f0f) keeps nibbles 0,2,4,6 = k0 + {0,4,1,5},
// ((w >> 4) & 0x0f0f0f0f) keeps nibbles 1,3,5,7 = k0 + {2,6,3,7}; lane c's two registers land in columns c and 4 + c.
__host__ __device__ __forceinline__ constexpr int tc_b_offset(int k32) {
  const int c = k32 >> 3, o = k32 & 7;
  const int rp = (o >> 1) & 1, b = ((o & 1) << 1) | (o >> 2);
  return rp * 1000 + 4 * c + b;  // rp * (chunk stride, substituted by the caller) + byte
}

template <int G, int NM /* MMA N: 8 (M <= 2) or 16 (M <= 5) */>
__global__ void __launch_bounds__(kTcThreads, 1) int4_gemv_tc_kernel(const GemvTcParams p) {
  constexpr int kStageB = tc_stage_bytes<G>();
  constexpr int GPS = kTcStageK / G;        // groups per stage (2 .. 16)
  constexpr int MPG = G / 32;               // MMAs (K = 32) per group
  constexpr uint32_t CH = NM * 16;          // bytes of one 16-k chunk of the B operand (NM rows x 16 B)
  constexpr uint32_t kXStage = 16 * 2 * CH; // B operand bytes per stage (16 blocks of 32 k)
  constexpr int kCT = kTcDequantWarps * 32; // staging threads
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* xs = smem + p.off_x;                                  // B operand: [win][16][2][NM][16] int8
  float* sx = reinterpret_cast<float*>(smem + p.off_sx);         // [win][GPS][8]: sum of the fixed-point activations
  float* part = reinterpret_cast<float*>(smem + p.off_misc);     // [8 rows][16 warps][2]
  float* row_scale = part + 8 * kTcDequantWarps * 2;             // [8]: 2^-sh (* rsqrt(mean sq + eps))
  float* row_dn = row_scale + 8;                                 // [8]: 2^sh
  uint64_t* full = reinterpret_cast<uint64_t*>(row_dn + 8);
  uint64_t* empty = full + p.stages;
  uint64_t* a_full = empty + p.stages;
  uint64_t* a_empty = a_full + kTcABufs;
  uint64_t* d_full = a_empty + kTcABufs;
  uint64_t* d_empty = d_full + kTcDBufs;
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(d_empty + kTcDBufs);
  __nv_bfloat16* xtmp = reinterpret_cast<__nv_bfloat16*>(smem + p.off_tmp);  // [M][win * 512] staging copy (bf16)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const int b = blockIdx.x;
  const int KT = p.KT, n_chunks = p.n_chunks;
  const int n_groups = KT * 64 / G;
  // work units [u0, u1): unit u = (row block u / n_chunks, stage u % n_chunks); the activation window of the CTA starts
  // at stage w0 and covers `win_len` consecutive stages (cyclically): local stage index lc = (ch - w0) mod n_chunks
  const int u0 = tc_first_unit(b, p.units_base, p.units_rem), u1 = tc_first_unit(b + 1, p.units_base, p.units_rem);
  const int rb_first = u0 / n_chunks, w0 = u0 - rb_first * n_chunks;
  const int win_len = min(n_chunks, u1 - u0);
  grid_launch_dependents();
  TC_TRACE(5, 0);

  if (warp == kTcMmaWarp) {
    // ---- barriers + tensor-memory allocation, then the other warps are let in (named barrier 2) ----
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], kTcDequantWarps);
      }
      for (int s = 0; s < kTcABufs; ++s) {
        mbar_init(&a_full[s], kTcDequantWarps);
        mbar_init(&a_empty[s], 1);
      }
      for (int s = 0; s < kTcDBufs; ++s) {
        mbar_init(&d_full[s], 1);
        mbar_init(&d_empty[s], kTcEpiWarps);
      }
      fence_barrier_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    tc_fence_before();
  }
  named_bar_sync(2, kTcThreads);  // barriers initialised, TMEM base published
  tc_fence_after();
  const uint32_t tb = *tmem_base_s;
  TC_TRACE(5, 1);
  const uint32_t a_col0 = tb, d_col0 = tb + kTcABufs * 128;

  if (warp == kTcProducerWarp) {
    // ===================== TMA producer: per stage 8 strips x 4 KB, lanes 0..7 issue one strip each =====================
    int rb = rb_first, ch = w0;
    for (int it = 0; it < u1 - u0; ++it) {
      const int slot = it % stages;
      const uint32_t par = ((uint32_t)(it / stages) & 1u) ^ 1u;
      const int cnt = min(kTcStageTiles, KT - ch * kTcStageTiles);
      uint8_t* dst = ring + (size_t)slot * kStageB;
      TC_TRACE(0, it);
      if (lane == 0) {
        mbar_wait(&empty[slot], par);
        mbar_arrive_expect_tx(&full[slot], (uint32_t)(8 * cnt * 512));
      }
      __syncwarp();
      if (lane < 8) {
        const int64_t strip = (int64_t)rb * 8 + lane;
        bulk_g2s(dst + lane * (kTcStageTiles * 512), p.packed + (strip * KT + (int64_t)ch * kTcStageTiles) * 512,
                 (uint32_t)(cnt * 512), &full[slot]);
      }
      TC_TRACE(1, it);
      if (++ch == n_chunks) { ch = 0; ++rb; }
    }
    __syncwarp();
    grid_dependency_wait();
  } else if (warp == kTcMmaWarp) {
    // ===================== MMA issuer: the whole warp runs the loop (uniform control flow keeps the operands in
    //                       uniform registers), one elected lane issues =====================
    named_bar_sync(3, kCT + 32 + kTcEpiWarps * 32);  // activations staged (B operand, correction sums)
    constexpr uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | (((uint32_t)NM >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint64_t desc_hi = ((uint64_t)((CH >> 4) & 0x3FFF) << 16) | ((uint64_t)((NM > 8 ? 128u : 0u) >> 4) << 32) | (1ull << 46);
    const uint32_t xs_u = smem_u32(xs);
    uint32_t dcount = 0;
    int ch = w0, lc = 0;
    for (int it = 0; it < u1 - u0; ++it) {
      const int abuf = it % kTcABufs;
      const int cnt = min(kTcStageTiles, KT - ch * kTcStageTiles);
      TC_TRACE(0, it);
      mbar_wait(&a_full[abuf], (uint32_t)(it / kTcABufs) & 1u);
      tc_fence_after();
      TC_TRACE(1, it);
      const uint32_t a_base = a_col0 + abuf * 128;
      const uint32_t x_base = xs_u + (uint32_t)lc * kXStage;
      const int n_grp = cnt * 2 / MPG;
      for (int gj = 0; gj < n_grp; ++gj) {
        const uint32_t dbuf = dcount % kTcDBufs;
        mbar_wait(&d_empty[dbuf], ((dcount / kTcDBufs) & 1u) ^ 1u);
        tc_fence_after();
        if (tc_elect_one()) {
#pragma unroll
          for (int jj = 0; jj < MPG; ++jj) {
            const int j = gj * MPG + jj;  // 32-k block inside the stage
            const uint64_t bdesc = desc_hi | (uint64_t)(((x_base + (uint32_t)j * 2 * CH) >> 4) & 0x3FFF);
            tc_mma_i8(d_col0 + dbuf * NM, a_base + 8 * j, bdesc, idesc, jj ? 1u : 0u);
          }
          tc_commit(&d_full[dbuf]);
        }
        __syncwarp();
        ++dcount;
      }
      if (tc_elect_one()) tc_commit(&a_empty[abuf]);
      __syncwarp();
      TC_TRACE(2, it);
      if (++ch == n_chunks) ch = 0;
      if (++lc == n_chunks) lc = 0;
    }
    __syncwarp();
    grid_dependency_wait();
  } else if (warp >= kTcDequantWarps) {
    // ===================== epilogue warps: thread = output row of the 128-row block =====================
    const int q = warp - kTcDequantWarps;
    const int row_in_block = 32 * q + lane;
    grid_dependency_wait();
    named_bar_sync(3, kCT + 32 + kTcEpiWarps * 32);
    // scale/zero of this thread's row: straight from global memory (6 % of the traffic, 64-byte runs), fetched one
    // stage ahead, so the weight ring is released by the dequant warps alone
    const uint32_t* sz_row = p.sz + (2 * q + (lane >> 4)) * (int64_t)n_groups * 16 + (lane & 15);
    auto load_sz = [&](int rb, int ch, uint32_t (&w)[GPS]) {
      const int cnt = min(kTcStageTiles, KT - ch * kTcStageTiles);
      const int ngrp = cnt * 64 / G;
      const uint32_t* src = sz_row + ((int64_t)rb * 8 * n_groups + (int64_t)ch * GPS) * 16;
#pragma unroll
      for (int gl = 0; gl < GPS; ++gl) w[gl] = gl < ngrp ? __ldg(src + gl * 16) : 0u;
    };
    // final scale + epilogue + store of the M outputs of row n
    auto finalize = [&](int rb, const float (&acc)[5]) {
      const int64_t n = (int64_t)rb * 128 + row_in_block;
#pragma unroll
      for (int m = 0; m < 5; ++m) {
        if (3 * m + 2 >= NM || m >= p.M) break;
        const float v = acc[m] * row_scale[m];
        if (p.epilogue == 2) {
          // rows i < 8 of a strip = gate (w1), rows i + 8 = up (w3) of output 8 strip + i  (model.py:486)
          const float other = __shfl_xor_sync(0xffffffffu, v, 8);
          const int64_t o = 8 * (n >> 4) + (n & 7);
          if ((n & 8) == 0 && 2 * o < p.N_out) {
            const float gate = bf16r(v), up = bf16r(other);
            const float sg = bf16r(gate / (1.f + __expf(-gate)));
            p.y[m * p.ldy + o] = __float2bfloat16_rn(sg * up);
          }
        } else if (n < p.N_out) {
          if (p.epilogue == 1)
            p.y[m * p.ldy + n] = __float2bfloat16_rn(__bfloat162float(p.residual[m * p.ld_res + n]) + bf16r(v));
          else
            p.y[m * p.ldy + n] = __float2bfloat16_rn(v);
        }
      }
    };
    uint32_t dcount = 0;
    uint32_t sz_next[GPS];
    int rb = rb_first, ch = w0, lc = 0;
    if (u0 < u1) load_sz(rb, ch, sz_next);
    float yacc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    int seg_ch0 = ch;  // first stage of the current row-block segment
    for (int it = 0; it < u1 - u0; ++it) {
      const int cnt = min(kTcStageTiles, KT - ch * kTcStageTiles);
      const int ngrp = cnt * 64 / G;
      TC_TRACE(0, it);
      uint32_t szw[GPS];
#pragma unroll
      for (int gl = 0; gl < GPS; ++gl) szw[gl] = sz_next[gl];
      const bool seg_end = (ch == n_chunks - 1) || (it == u1 - u0 - 1);
      if (it + 1 < u1 - u0) load_sz(ch + 1 < n_chunks ? rb : rb + 1, ch + 1 < n_chunks ? ch + 1 : 0, sz_next);
      TC_TRACE(1, it);
#pragma unroll
      for (int gl = 0; gl < GPS; ++gl) {
        if (gl >= ngrp) break;
        const uint32_t dbuf = dcount % kTcDBufs;
        mbar_wait(&d_full[dbuf], (dcount / kTcDBufs) & 1u);
        tc_fence_after();
        uint32_t d[NM];
        if constexpr (NM == 8) tc_ld_32x32b_x8(d_col0 + dbuf * NM + ((uint32_t)(32 * q) << 16), d);
        else tc_ld_32x32b_x16(d_col0 + dbuf * NM + ((uint32_t)(32 * q) << 16), d);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&d_empty[dbuf]);
        ++dcount;
        const float s = __uint_as_float(szw[gl] << 16), z = __uint_as_float(szw[gl] & 0xFFFF0000u);
        const float t = fmaf(-8.f, s, z);
        const float* sxg = sx + (size_t)(lc * GPS + gl) * 8;
#pragma unroll
        for (int m = 0; m < 5; ++m) {
          if (3 * m + 2 >= NM || m >= p.M) break;
          const float val = fmaf(fmaf((float)(int)d[3 * m + 2], 256.f, (float)(int)d[3 * m + 1]), 256.f, (float)(int)d[3 * m]);
          yacc[m] = fmaf(s, val, fmaf(t, sxg[m], yacc[m]));
        }
      }
      TC_TRACE(2, it);
      if (seg_end) {
        if (seg_ch0 == 0 && ch == n_chunks - 1) {
          finalize(rb, yacc);  // the whole K of this row block was ours
        } else {
          // partial sums of a row block shared with neighbouring CTAs: park them, take a ticket; the last CTA to arrive
          // adds the contributions in CTA order (deterministic) and finishes the rows.  No CTA ever waits for another.
          const int slot = (rb == rb_first) ? 0 : 1;
          float* mine = p.partials + ((size_t)(b * 2 + slot) * 5) * 128 + row_in_block;
#pragma unroll
          for (int m = 0; m < 5; ++m) {
            if (3 * m + 2 >= NM || m >= p.M) break;
            __stcg(mine + m * 128, yacc[m]);
          }
          __threadfence();
          __syncwarp();
          const int c_first = tc_cta_of_unit(rb * n_chunks, p.units_base, p.units_rem);
          const int c_last = tc_cta_of_unit(rb * n_chunks + n_chunks - 1, p.units_base, p.units_rem);
          uint32_t ticket = 0;
          if (lane == 0) ticket = atomicAdd(&p.tickets[rb * 4 + q], 1u);
          ticket = __shfl_sync(0xffffffffu, ticket, 0);
          if (ticket == (uint32_t)(c_last - c_first)) {
            __threadfence();
            float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            for (int c = c_first; c <= c_last; ++c) {
              const int cs = (tc_first_unit(c, p.units_base, p.units_rem) / n_chunks == rb) ? 0 : 1;
              const float* src = p.partials + ((size_t)(c * 2 + cs) * 5) * 128 + row_in_block;
#pragma unroll
              for (int m = 0; m < 5; ++m) {
                if (3 * m + 2 >= NM || m >= p.M) break;
                acc[m] += __ldcg(src + m * 128);
              }
            }
            if (lane == 0) p.tickets[rb * 4 + q] = 0;  // ready for the next launch
            finalize(rb, acc);
          }
        }
#pragma unroll
        for (int m = 0; m < 5; ++m) yacc[m] = 0.f;
        seg_ch0 = 0;
      }
      if (++ch == n_chunks) { ch = 0; ++rb; }
      if (++lc == n_chunks) lc = 0;
    }
  } else {
    // ===================== dequant warps (activation staging first) =====================
    const int tid = threadIdx.x;  // 0 .. 511
    grid_dependency_wait();
    {
      // local stage index of global stage c in this CTA's window, or -1
      auto local_stage = [&](int c) {
        int l = c - w0;
        if (l < 0) l += n_chunks;
        return l < win_len ? l : -1;
      };
      const int chunks_per_row = KT * 8;  // 16-byte chunks of 8 elements over Kp
      const int64_t tmp_row = (int64_t)win_len * kTcStageK;
      // pass 1: the whole row (sum of squares, max magnitude); the window's part is kept as bf16
      for (int m = 0; m < p.M; ++m) {
        float ss = 0.f, amax = 0.f;
        for (int ci = tid; ci < chunks_per_row; ci += kCT) {
          const int64_t k0 = (int64_t)ci * 8;
          uint4 v = make_uint4(0, 0, 0, 0);
          if (k0 < p.Kx) v = *reinterpret_cast<const uint4*>(p.x + (int64_t)m * p.ldx + k0);
          float f[8];
          unpack8(v, f);
          if (p.norm_w != nullptr && k0 < p.Kx) {
            float wn[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(p.norm_w + k0)), wn);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              ss = fmaf(f[e], f[e], ss);
              f[e] = bf16r(f[e] * wn[e]);
            }
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) o[e] = float_to_bf16_bits(f[2 * e]) | (float_to_bf16_bits(f[2 * e + 1]) << 16);
            v = make_uint4(o[0], o[1], o[2], o[3]);
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(f[e]));
          const int l = local_stage(ci >> 6);
          if (l >= 0) *reinterpret_cast<uint4*>(xtmp + (int64_t)m * tmp_row + (int64_t)l * kTcStageK + (ci & 63) * 8) = v;
        }
        ss = tc_warp_sum(ss);
        amax = tc_warp_max(amax);
        if (lane == 0) {
          part[(m * kTcDequantWarps + warp) * 2] = ss;
          part[(m * kTcDequantWarps + warp) * 2 + 1] = amax;
        }
      }
      named_bar_sync(1, kCT);
      if (warp < p.M) {
        float t = 0.f, amax = 0.f;
        if (lane < kTcDequantWarps) {
          t = part[(warp * kTcDequantWarps + lane) * 2];
          amax = part[(warp * kTcDequantWarps + lane) * 2 + 1];
        }
        t = tc_warp_sum(t);
        amax = tc_warp_max(amax);
        if (lane == 0) {
          // 2^sh brings the row's largest magnitude into [2^21, 2^22): 23-bit fixed point, three balanced int8 limbs
          int sh = 0;
          if (amax > 0.f) sh = 21 - ((int)((__float_as_uint(amax) >> 23) & 0xFF) - 127);
          sh = max(-100, min(100, sh));
          const float up = __uint_as_float((uint32_t)(sh + 127) << 23), dn = __uint_as_float((uint32_t)(127 - sh) << 23);
          row_dn[warp] = up;
          row_scale[warp] = (p.norm_w != nullptr) ? dn * rsqrtf(t / (float)p.Kx + p.eps) : dn;
        }
      }
      named_bar_sync(1, kCT);
      // pass 2 (window only): fixed point, limbs into the B operand layout, per-group sums
      constexpr int CPG = G / 8;  // chunks per group: 4 .. 32 consecutive lanes
      const int n_win_chunks = win_len * 64;
      for (int m = 0; m < p.M; ++m) {
        const float up = row_dn[m];
        for (int base = 0; base < n_win_chunks; base += kCT) {
          const int wi = base + tid;  // chunk index inside the window: local stage wi / 64, 8-element chunk wi % 64
          int sum = 0;
          if (wi < n_win_chunks) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4*>(xtmp + (int64_t)m * tmp_row + (int64_t)wi * 8), f);
            int xi[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              xi[e] = __float2int_rn(f[e] * up);
              sum += xi[e];
            }
            // byte b of chunk r' holds element {0,4,1,5}[b] (r' = 0) / {2,6,3,7}[b] (r' = 1)  — see tc_b_offset
            const int blk = wi >> 2, c = wi & 3;  // 32-k block inside the window
            uint8_t* dst = xs + (size_t)blk * 2 * CH + 4 * c;
#pragma unroll
            for (int l = 0; l < kTcLimbs; ++l) {
              uint32_t w0_ = 0, w1_ = 0;
              constexpr int o0[4] = {0, 4, 1, 5}, o1[4] = {2, 6, 3, 7};
#pragma unroll
              for (int bb = 0; bb < 4; ++bb) {
                w0_ |= ((uint32_t)xi[o0[bb]] & 0xFFu) << (8 * bb);
                w1_ |= ((uint32_t)xi[o1[bb]] & 0xFFu) << (8 * bb);
              }
              const int n = 3 * m + l;
              const uint32_t roff = (uint32_t)((n >> 3) * 128 + (n & 7) * 16);
              *reinterpret_cast<uint32_t*>(dst + roff) = w0_;
              *reinterpret_cast<uint32_t*>(dst + CH + roff) = w1_;
              // next balanced digit: remove the signed low byte, shift
#pragma unroll
              for (int e = 0; e < 8; ++e) xi[e] = (xi[e] - (int)(int8_t)(xi[e] & 0xFF)) >> 8;
            }
          }
#pragma unroll
          for (int o = CPG / 2; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
          if (wi < n_win_chunks && (wi % CPG) == 0) sx[(size_t)(wi / CPG) * 8 + m] = (float)sum;
        }
      }
      TC_TRACE(5, 2);
      fence_proxy_async();  // the B operand is read by the tensor core through the async proxy
      named_bar_arrive(3, kCT + 32 + kTcEpiWarps * 32);
    }
    // ---- main loop: nibbles -> bytes -> tensor memory ----
    const int q = warp & 3, jslot = warp >> 2;
    const uint32_t ring_u = smem_u32(ring);
    int ch = w0;
    for (int it = 0; it < u1 - u0; ++it) {
      const int slot = it % stages, abuf = it % kTcABufs;
      const int cnt = min(kTcStageTiles, KT - ch * kTcStageTiles);
      TC_TRACE(0, it);
      mbar_wait(&full[slot], (uint32_t)(it / stages) & 1u);
      TC_TRACE(3, it);
      mbar_wait(&a_empty[abuf], ((uint32_t)(it / kTcABufs) & 1u) ^ 1u);
      tc_fence_after();
      TC_TRACE(1, it);
      const uint32_t stage_u = ring_u + slot * kStageB;
#pragma unroll
      for (int sh = 0; sh < 2; ++sh) {
#pragma unroll
        for (int tt = 0; tt < 2; ++tt) {
          const int t = 2 * jslot + tt;
          if (t < cnt) {
            const uint4 w = lds128(stage_u + (uint32_t)(((2 * q + sh) * kTcStageTiles + t) * 512 + lane * 16));
            uint32_t r[8];
            r[0] = w.x & 0x0F0F0F0Fu; r[1] = w.y & 0x0F0F0F0Fu;
            r[2] = (w.x >> 4) & 0x0F0F0F0Fu; r[3] = (w.y >> 4) & 0x0F0F0F0Fu;
            r[4] = w.z & 0x0F0F0F0Fu; r[5] = w.w & 0x0F0F0F0Fu;
            r[6] = (w.z >> 4) & 0x0F0F0F0Fu; r[7] = (w.w >> 4) & 0x0F0F0F0Fu;
            tc_st_16x128b_x4(a_col0 + abuf * 128 + t * 16 + ((uint32_t)(32 * q + 16 * sh) << 16), r);
          }
        }
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&empty[slot]);
        mbar_arrive(&a_full[abuf]);
      }
      TC_TRACE(2, it);
      if (++ch == n_chunks) ch = 0;
    }
  }

  // ---- teardown: nobody touches tensor memory any more, the allocating warp frees it ----
  tc_fence_before();
  __syncthreads();
  if (warp == kTcMmaWarp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512));
}

// ------------------------------------------------------------------------------------------------ host side
static int tc_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    n = v > 0 ? v : kNumSMs;
  }
  return n;
}

// workspace: [grid][2][5][128] fp32 partial sums (fixed size, FIRST: launches of different shapes share one
// workspace, and only the tickets must stay zero between launches) + [n_blocks][4] tickets
// (the ticket area has a FIXED size, so that whatever follows it in a shared workspace — the prefill GEMM's staging
// buffer — never overlaps the tickets of a launch with more row blocks)
constexpr size_t kTcTicketBytes = 1u << 20;  // 65536 row blocks = 8.4 M output rows
static size_t tc_ticket_bytes(int64_t Np) { (void)Np; return kTcTicketBytes; }
static size_t tc_partial_bytes() { return (size_t)tc_sm_count() * 2 * 5 * 128 * 4; }
int64_t int4_gemv_tc_workspace_bytes(int64_t Np) {
  if (Np % 128 != 0 || (size_t)(Np / 128) * 16 > kTcTicketBytes) return 0;
  return (int64_t)(tc_ticket_bytes(Np) + tc_partial_bytes());
}

template <int G, int NM>
static int launch_tc(GemvTcParams& p, bool pdl, cudaStream_t s) {
  constexpr int GPS = kTcStageK / G;
  const int units = p.n_blocks * p.n_chunks;
  const int grid = units < tc_sm_count() ? units : tc_sm_count();
  p.units_base = units / grid;
  p.units_rem = units % grid;
  p.win = p.n_chunks < p.units_base + 1 ? p.n_chunks : p.units_base + 1;
  const size_t stage_b = tc_stage_bytes<G>();
  const size_t x_bytes = (size_t)p.win * 16 * 2 * NM * 16;
  const size_t sx_bytes = ((size_t)p.win * GPS * 32 + 127) / 128 * 128;
  const size_t tmp_bytes = ((size_t)p.M * p.win * kTcStageK * 2 + 127) / 128 * 128;
  const size_t misc_fixed = (8 * kTcDequantWarps * 2 + 16) * 4 + (2 * kTcABufs + 2 * kTcDBufs) * 8 + 64;
  const size_t budget = 227 * 1024;
  const size_t fixed = x_bytes + sx_bytes + tmp_bytes + misc_fixed + 2 * 16 * 8 + 128;
  if (fixed + 3 * stage_b > budget) return AO_B200_ERR_WORKSPACE;
  int st = (int)((budget - fixed) / stage_b);
  if (st > 16) st = 16;
  p.stages = st;
  size_t off = (size_t)st * stage_b;
  p.off_x = (uint32_t)off; off += x_bytes;
  p.off_sx = (uint32_t)off; off += sx_bytes;
  p.off_tmp = (uint32_t)off; off += tmp_bytes;
  p.off_misc = (uint32_t)off; off += misc_fixed + 2 * (size_t)st * 8 + 128;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(int4_gemv_tc_kernel<G, NM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kTcThreads);
  cfg.dynamicSmemBytes = off;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  if (cudaLaunchKernelEx(&cfg, int4_gemv_tc_kernel<G, NM>, p) != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

// returns AO_B200_ERR_WORKSPACE / AO_B200_ERR_UNSUPPORTED when the shape does not fit this kernel (caller falls back)
int int4_gemv_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                 int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps, const void* residual,
                 int64_t ld_res, int epilogue, void* workspace, int64_t workspace_bytes, bool pdl, cudaStream_t s) {
  if (M > 5 || Np % 128 != 0) return AO_B200_ERR_UNSUPPORTED;
  if (int4_gemv_tc_workspace_bytes(Np) == 0) return AO_B200_ERR_UNSUPPORTED;
  if (workspace == nullptr || workspace_bytes < int4_gemv_tc_workspace_bytes(Np)) return AO_B200_ERR_WORKSPACE;
  GemvTcParams p;
  p.packed = static_cast<const uint8_t*>(packed);
  p.sz = static_cast<const uint32_t*>(sz);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.norm_w = static_cast<const __nv_bfloat16*>(norm_w);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.epilogue = epilogue;
  p.KT = (int)(Kp / 64);
  p.Kx = Kx; p.ldx = ldx; p.N_out = N_out; p.ldy = ldy; p.ld_res = ld_res;
  p.n_blocks = (int)(Np / 128);
  p.n_chunks = (p.KT + kTcStageTiles - 1) / kTcStageTiles;
  p.eps = eps;
  p.partials = static_cast<float*>(workspace);
  p.tickets = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(workspace) + tc_partial_bytes());
  static const bool force16 = std::getenv("AO_B200_TC_N16") != nullptr;
  const bool n8 = M <= 2 && !force16;
  p.N_mma = n8 ? 8 : 16;
#define AO_LT(GG) return n8 ? launch_tc<GG, 8>(p, pdl, s) : launch_tc<GG, 16>(p, pdl, s)
  switch (G) {
    case 32: AO_LT(32);
    case 64: AO_LT(64);
    case 128: AO_LT(128);
    case 256: AO_LT(256);
  }
#undef AO_LT
  return AO_B200_ERR_UNSUPPORTED;
}

}  // namespace ao
