// int4 weight-only linear, prefill path (M > 16): fused dequant + GEMM on tcgen05 / tensor memory.
//
//   y[m][n] = sum_k x[m][k] * W'[n][k],   W'[n][k] = bf16_rn((q[n][k] - 8) * scale[g][n] + zero[g][n])
//
// i.e. exactly the reference's tinygemm math (aten._weight_int4pack_mm, the op behind
// torchao/dtypes/uintx/tensor_core_tiled_layout.py:_linear_bf16_act_uint4_weight_impl): every weight is dequantised
// to bf16 with ONE rounding, products accumulate in fp32.  The tensor-core mapping is transposed with respect to the
// usual GEMM so that the dequantised weights never touch shared memory:
//
//   MMA M (128 TMEM lanes)  = 128 weight rows (8 strips of the strip-major packed layout)
//   MMA K                   = 16 k per instruction, 4 instructions per 64-k tile
//   MMA N (<= 256)          = NT activation rows (a tile of the prompt)
//   A operand               = W' in TENSOR MEMORY, written by the dequant warps with tcgen05.st.16x128b: the 16-byte
//                             lane chunk of the packed layout holds rows g / g + 8 of a strip, which is the
//                             (lane, column) pattern of that store shape
//   B operand               = activations in shared memory (canonical no-swizzle K-major core matrices), brought in by
//                             TMA bulk copies from a pre-permuted copy of x (see x_permute_kernel)
//   D                       = fp32 [128 weight rows][NT] in tensor memory, accumulated over the WHOLE K, read back once.
//
// The k order inside a 32-k block that the 16x128b store produces (column p = 4 i + c holds k = 8 c + 2 i, 2 i + 1) is
// baked into the pre-permuted activations, so neither side shuffles anything in the main loop.
//
// Roles (704 threads): warps 0-3 epilogue (warp = TMEM lane quadrant), 4-19 dequant (quadrant = warp % 4, one strip
// each, even / odd stages by warp set), 20 TMA producer, 21 MMA issue (+ TMEM allocation).
//
// Thread-block cluster (CL = 4, 2 or 1 CTAs along the weight rows): the CTAs of a cluster work on the same prompt
// tile, so each one fetches 1 / CL of the activation tile and MULTICASTS it into all of them; without this the
// activation re-reads (32 KB per 64-k stage per CTA against 4 KB of weights) are bound by L2 bandwidth at about a
// third of the tensor rate.  A ring slot is free when the MMAs of ALL cluster CTAs on it have retired (multicast
// tcgen05.commit onto every CTA's `empty` barrier).
#include <cuda_bf16.h>

#include <cstdlib>

#include "common.cuh"

namespace ao {

#ifndef AO_GM_DQ_SETS
#define AO_GM_DQ_SETS 2   // 3 sets (960 threads) measured: no gain, the dequant is bound by ALU issue (profiles/r02_gemm_trace.txt)
#endif
constexpr int kGmDqSets = AO_GM_DQ_SETS;            // sets of 8 dequant warps (one strip each); set s takes the 64-k tiles s, s + sets, ...
constexpr int kGmEpiWarps = 4, kGmDqWarps = 8 * kGmDqSets;
constexpr int kGmProducerWarp = kGmEpiWarps + kGmDqWarps, kGmMmaWarp = kGmProducerWarp + 1;
constexpr int kGmThreads = (kGmMmaWarp + 1) * 32;  // 704
constexpr int kGmWSlots = 3, kGmWSlotBytes = 8 * 4 * 512;  // weight ring: 16 KB slices of the 128-row block (256 k of int4, 128 k of int8)
constexpr int kGmMaxStages = 8;                            // activation ring slots (64 k each)
constexpr int kGmHeader = 1024;                            // mbarriers + TMEM base address

#ifdef AO_GM_TRACE
constexpr int kGmTraceIts = 256;
__device__ long long ao_gm_trace[4][kGmMmaWarp + 1][kGmTraceIts];
#define GM_TRACE(kind, it)                                                                               \
  do {                                                                                                   \
    if (blockIdx.x == 0 && blockIdx.y == 0 && (threadIdx.x & 31) == 0 && (it) < kGmTraceIts)             \
      ao_gm_trace[kind][threadIdx.x >> 5][it] = clock64();                                               \
  } while (0)
#else
#define GM_TRACE(kind, it) do { } while (0)
#endif

struct GemmTcParams {
  const uint8_t* packed;     // strip-major [Np/16][Kp/64][512 B]
  const uint32_t* sz;        // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* scale8;  // int8 (G == 0): per-row scale [N_out], applied to the bf16-rounded accumulator
  const __nv_bfloat16* xt;   // permuted activations [Mpad/NT][Kp/64][8 chunks][NT][8]
  __nv_bfloat16* y;          // [M][ldy]
  int M, KT, stages, tmem_cols, cluster, vec_store;
  int kt_per_split;          // split-K (grid.z): this many 64-k tiles per split, a multiple of the weight-ring slice
  float* part;               // split-K: fp32 partial sums [splits][Mpad][Np] (null: one split, bf16 output written directly)
  int64_t Mpad, N_out, ldy, Np;
};

// ---- activations -> chunk-major, k-permuted copy ------------------------------------------------------------------
// xt[m / NT][kt][q][m % NT][c * 2 + e] = x[m][64 kt + 32 half + 8 c + 2 i + e],  q = 4 half + i; rows >= M and columns
// >= Kx are zero.  The (NT rows x 64 k) tile of one pipeline stage is one contiguous NT * 128 bytes, already in the
// core-matrix order of the MMA's shared-memory operand.  One thread = one (row, 64-k tile): 128 bytes in, 8 x 16
// bytes out (coalesced over rows).
__global__ void __launch_bounds__(256) x_permute_kernel(const __nv_bfloat16* __restrict__ x, int64_t M, int64_t Kx,
                                                        int64_t ldx, int64_t Mpad, int KT, int NT,
                                                        __nv_bfloat16* __restrict__ xt) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Mpad * KT) return;
  const int64_t kt = idx / Mpad, m = idx - kt * Mpad;
  uint32_t in[32];
#pragma unroll
  for (int v = 0; v < 8; ++v) {
    uint4 t = make_uint4(0u, 0u, 0u, 0u);
    const int64_t k = kt * 64 + v * 8;
    if (m < M && k < Kx) t = *reinterpret_cast<const uint4*>(x + m * ldx + k);  // Kx % 8 == 0: all or nothing
    in[4 * v] = t.x, in[4 * v + 1] = t.y, in[4 * v + 2] = t.z, in[4 * v + 3] = t.w;
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int half = q >> 2, i = q & 3;
    const uint4 o = make_uint4(in[16 * half + i], in[16 * half + 4 + i], in[16 * half + 8 + i], in[16 * half + 12 + i]);
    *reinterpret_cast<uint4*>(xt + ((((m / NT) * KT + kt) * 8 + q) * NT + m % NT) * 8) = o;
  }
}

// ---- tcgen05 helpers ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void gm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void gm_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D(tmem, f32 [128][N]) (+)= A(tmem, bf16 [128][16]) * B(smem, bf16 [N][16])
__device__ __forceinline__ void gm_mma_bf16(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void gm_st_16x128b_x8(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void gm_ld_32x32b_x32(uint32_t taddr, uint32_t (&d)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
      "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]), "=r"(d[9]),
        "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15]), "=r"(d[16]), "=r"(d[17]), "=r"(d[18]),
        "=r"(d[19]), "=r"(d[20]), "=r"(d[21]), "=r"(d[22]), "=r"(d[23]), "=r"(d[24]), "=r"(d[25]), "=r"(d[26]), "=r"(d[27]),
        "=r"(d[28]), "=r"(d[29]), "=r"(d[30]), "=r"(d[31])
      : "r"(taddr));
}
__device__ __forceinline__ void gm_commit_multicast(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void gm_bulk_g2s_multicast(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                                      uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void gm_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t gm_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ bool gm_elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\tselp.u32 %0, 1, 0, px;\n\t}" : "=r"(pred));
  return pred != 0;
}

// ---- the GEMM -----------------------------------------------------------------------------------------------------
// grid (Np / 128, Mpad / NT), clusters of CL CTAs along x.  Shared memory: 1 KB of mbarriers, the WEIGHT ring
// (kGmWSlots x [8 strips x 4 tiles x 512 B]: a strip's consecutive k-tiles are contiguous in the packed layout, so a
// 256-k slice of the row block is 8 copies of 2 KB), the ACTIVATION ring (`stages` x [8 chunks x NT x 16 B], one 64-k
// tile per slot, each CTA of the cluster multicasting its 1 / CL share as ONE copy).  A bulk copy costs its issuing
// warp ~100 clocks, so copies per 64 k are what bounds the producer: 3 here (first version: 10, producer-bound at
// 1200 clocks per tile against 512 of MMA).  Tensor memory: columns [0, NT) accumulator, NT + 32 slot ..: A tiles.
// G == 0: int8 weights (strip-major biased int8, 1 KB per 16 x 64 tile, per-row scale in the epilogue) — same pipeline,
// the dequant warps convert bytes to bf16 exactly instead of applying the int4 affine.
template <int G, int NT>
__global__ void __launch_bounds__(kGmThreads, 1) int4_gemm_tc_kernel(const GemmTcParams p) {
  extern __shared__ __align__(1024) uint8_t gm_smem[];
  constexpr bool kI8 = (G == 0);
  constexpr int kTileB = kI8 ? 1024 : 512;                    // bytes of one (strip, 64-k) tile
  constexpr int kSlotTiles = kGmWSlotBytes / (8 * kTileB);    // k-tiles per weight-ring slot: 4 (int4) / 2 (int8)
  constexpr uint32_t kXStage = NT * 128;
  constexpr uint32_t kChunk = NT * 16;  // bytes between consecutive 8-k chunks of the B tile (descriptor LBO)
  const int stages = p.stages, KTa = p.KT;
  // split-K: this CTA covers k-tiles [kt_lo, kt_lo + KT) of the row block (the whole K when grid.z == 1)
  const int kt_lo = blockIdx.z * p.kt_per_split;
  const int KT = min(p.kt_per_split, KTa - kt_lo);
  uint64_t* wfull = reinterpret_cast<uint64_t*>(gm_smem);
  uint64_t* wempty = wfull + kGmWSlots;
  uint64_t* xfull = wempty + kGmWSlots;
  uint64_t* empty = xfull + kGmMaxStages;
  uint64_t* a_full = empty + kGmMaxStages;
  uint64_t* d_full = a_full + kGmMaxStages;
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(d_full + 1);
  uint8_t* wring = gm_smem + kGmHeader;
  uint8_t* xring = wring + kGmWSlots * kGmWSlotBytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nb = blockIdx.x, mt = blockIdx.y;
  const int CL = p.cluster;

  if (warp == kGmMmaWarp) {
    if (lane == 0) {
      for (int s = 0; s < kGmWSlots; ++s) {
        mbar_init(&wfull[s], 1);
        mbar_init(&wempty[s], 8 * kSlotTiles);  // one arrival per (dequant warp = strip, tile)
      }
      for (int s = 0; s < stages; ++s) {
        mbar_init(&xfull[s], 1);
        mbar_init(&empty[s], CL);
        mbar_init(&a_full[s], 8);
      }
      mbar_init(d_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_s)), "r"(p.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    gm_fence_before();
  }
  __syncthreads();
  gm_cluster_sync();  // every CTA's barriers exist before a peer multicasts into them
  gm_fence_after();
  const uint32_t crank = CL > 1 ? gm_cluster_rank() : 0u;
  const uint16_t cmask = (uint16_t)((1u << CL) - 1);
  const uint32_t tb = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
  const uint32_t wring_u = smem_u32(wring), xring_u = smem_u32(xring);

  if (warp == kGmProducerWarp) {
    // ---- TMA producer, activations: one multicast copy (this CTA's share of the stage) per 64-k tile ----------------
    const uint32_t share = kXStage / CL;
    const uint8_t* xsrc = reinterpret_cast<const uint8_t*>(p.xt) + ((int64_t)mt * KTa + kt_lo) * kXStage + crank * share;
    for (int it = 0; it < KT; ++it) {
      GM_TRACE(0, it);
      const int slot = it % stages;
      if (it >= stages) mbar_wait(&empty[slot], ((it / stages) - 1) & 1);
      GM_TRACE(1, it);
      if (gm_elect_one()) {
        mbar_arrive_expect_tx(&xfull[slot], kXStage);
        gm_bulk_g2s_multicast(xring + slot * kXStage + crank * share, xsrc + (int64_t)it * kXStage, share, &xfull[slot], cmask);
      }
      __syncwarp();
    }
  } else if (warp == kGmMmaWarp) {
    // ---- MMA issue: 4 x (M128, N = NT, K16) per 64-k tile, whole warp in uniform control flow ---------------------
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)NT >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint64_t desc_hi = ((uint64_t)((kChunk >> 4) & 0x3FFF) << 16) | ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
    for (int it = 0; it < KT; ++it) {
      const int slot = it % stages;
      const uint32_t par = (it / stages) & 1;
      GM_TRACE(0, it);
      mbar_wait(&a_full[slot], par);
      GM_TRACE(1, it);
      mbar_wait(&xfull[slot], par);
      gm_fence_after();
      GM_TRACE(2, it);
      if (gm_elect_one()) {
        const uint32_t bx = xring_u + slot * kXStage;
        const uint32_t a0 = tb + NT + slot * 32;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t bdesc = desc_hi | (uint64_t)(((bx + (uint32_t)ks * 2 * kChunk) >> 4) & 0x3FFF);
          gm_mma_bf16(tb, a0 + 8 * ks, bdesc, idesc, (it | ks) ? 1u : 0u);
        }
        gm_commit_multicast(&empty[slot], cmask);
        if (it == KT - 1) gm_commit(d_full);
      }
      __syncwarp();
      GM_TRACE(3, it);
    }
  } else if (warp >= kGmEpiWarps) {
    // ---- dequant warps: packed words -> bf16 W' (reference arithmetic, one rounding) -> tensor memory -------------
    const int dq = (warp - kGmEpiWarps) & 7, set = (warp - kGmEpiWarps) >> 3, qd = warp & 3, strip = 2 * qd + (dq >> 2);
    const int g = lane >> 2;
    constexpr int GG = kI8 ? 64 : G;  // (int8: no per-group parameters; GG only keeps the int4 index math well-formed)
    constexpr int kGroupsPerTile = GG >= 64 ? 1 : 64 / GG;
    const int n_groups = KTa * 64 / GG;
    const uint32_t* szp = kI8 ? nullptr : p.sz + ((int64_t)(nb * 8 + strip) * n_groups) * 16 + g;
    const uint32_t taddr0 = tb + ((uint32_t)(32 * qd + 16 * (dq >> 2)) << 16) + NT;
    uint32_t szn[kGroupsPerTile][2];
    if constexpr (!kI8) {
      const int64_t grp = GG >= 64 ? (int64_t)(kt_lo + set) * 64 / GG : (int64_t)(kt_lo + set) * kGroupsPerTile;
#pragma unroll
      for (int h = 0; h < kGroupsPerTile; ++h) {
        szn[h][0] = __ldg(szp + (grp + h) * 16);
        szn[h][1] = __ldg(szp + (grp + h) * 16 + 8);
      }
    }
    for (int it = set; it < KT; it += kGmDqSets) {
      const int slot = it % stages;
      uint32_t s2[kGroupsPerTile][2], z2[kGroupsPerTile][2];
      if constexpr (!kI8) {
#pragma unroll
        for (int h = 0; h < kGroupsPerTile; ++h)
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            s2[h][r] = __byte_perm(szn[h][r], 0, 0x1010);
            z2[h][r] = __byte_perm(szn[h][r], 0, 0x3232);
          }
        if (it + kGmDqSets < KT) {
          const int64_t grp = GG >= 64 ? (int64_t)(kt_lo + it + kGmDqSets) * 64 / GG : (int64_t)(kt_lo + it + kGmDqSets) * kGroupsPerTile;
#pragma unroll
          for (int h = 0; h < kGroupsPerTile; ++h) {
            szn[h][0] = __ldg(szp + (grp + h) * 16);
            szn[h][1] = __ldg(szp + (grp + h) * 16 + 8);
          }
        }
      }
      const int wc = it / kSlotTiles, ws = wc % kGmWSlots;
      GM_TRACE(0, it);
      mbar_wait(&wfull[ws], (wc / kGmWSlots) & 1);
      GM_TRACE(1, it);
      const uint32_t wtile = wring_u + ws * kGmWSlotBytes + strip * (kSlotTiles * kTileB) + (it % kSlotTiles) * kTileB + lane * 16;
      uint32_t r[16];
      if constexpr (kI8) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {  // lane chunk of a half: words (row g, k 0-3), (g+8, k 0-3), (g, k 4-7), (g+8, k 4-7)
          const uint4 w4 = lds128(wtile + half * 512);
          const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int hrow = j & 1, i0 = 2 * (j >> 1);
            u8x4_to_bf16x2x2(w[j], r[2 * (4 * half + i0) + hrow], r[2 * (4 * half + i0 + 1) + hrow]);
          }
        }
      } else {
        const uint4 w4 = lds128(wtile);
        const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // word j: row g + 8 (j & 1), k half j >> 1
          const int half = j >> 1, hrow = j & 1;
          const int h = kGroupsPerTile > 1 ? half * (kGroupsPerTile / 2) : 0;
          uint32_t o[4];
          dequant_word(w[j], s2[h][hrow], z2[h][hrow], o);
#pragma unroll
          for (int i = 0; i < 4; ++i) r[2 * (4 * half + i) + hrow] = o[i];
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&wempty[ws]);  // the words are in registers
      if (it >= stages) mbar_wait(&empty[slot], ((it / stages) - 1) & 1);  // the MMAs that read this A slot retired
      gm_fence_after();
      GM_TRACE(2, it);
      gm_st_16x128b_x8(taddr0 + slot * 32, r);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      GM_TRACE(3, it);
      gm_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_full[slot]);
    }
  } else {
    if (warp == 0) {
      // ---- TMA producer, weights (this warp has nothing else to do until the accumulator is final): slice wc = k-tiles
      // 4 wc .. 4 wc + 3 of the row block, 8 copies of 2 KB, into ring slot wc % kGmWSlots as soon as the dequant
      // warps have released it.  Kept off the activation producer: a bulk copy costs its issuing warp ~130 clocks.
      const uint8_t* wsrc = p.packed + tile_slot((int64_t)nb * 8 + (lane & 7), kt_lo, KTa) * kTileB;
      for (int wc = 0; wc * kSlotTiles < KT; ++wc) {
        const int ws = wc % kGmWSlots, t0 = wc * kSlotTiles;
        if (wc >= kGmWSlots) mbar_wait(&wempty[ws], ((wc / kGmWSlots) - 1) & 1);
        const uint32_t tiles = (uint32_t)(KT - t0 < kSlotTiles ? KT - t0 : kSlotTiles);
        if (lane == 0) mbar_arrive_expect_tx(&wfull[ws], 8 * kTileB * tiles);
        __syncwarp();
        if (lane < 8)
          bulk_g2s(wring + ws * kGmWSlotBytes + lane * (kSlotTiles * kTileB), wsrc + (int64_t)t0 * kTileB, kTileB * tiles, &wfull[ws]);
        __syncwarp();
      }
    }
    // ---- epilogue: thread = weight row (TMEM lane); fp32 -> bf16, transposed through the (now idle) rings so that
    // the global stores are 16-byte pieces of 256-byte output rows -----------------------------------------------
    GM_TRACE(0, 0);
    mbar_wait(d_full, 0);
    gm_fence_after();
    GM_TRACE(1, 0);
    const int nl = 32 * warp + lane;
    if (p.part != nullptr) {
      // split-K: this split's fp32 partial sums straight from tensor memory (thread = weight row: a warp's 32 rows are one
      // 128-byte line per activation row); split_reduce_kernel sums the splits and applies the output rounding
      float* dst = p.part + ((int64_t)blockIdx.z * p.Mpad + (int64_t)mt * NT) * p.Np + (int64_t)nb * 128 + nl;
#pragma unroll 1
      for (int cc = 0; cc < NT / 32; ++cc) {
        uint32_t d[32];
        gm_ld_32x32b_x32(tb + ((uint32_t)(32 * warp) << 16) + 32 * cc, d);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if ((int64_t)mt * NT + 32 * cc < p.M) {
          float* q = dst + (int64_t)(32 * cc) * p.Np;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            *q = __uint_as_float(d[j]);
            q += p.Np;
          }
        }
      }
      gm_fence_before();
    } else {
    constexpr uint32_t kPitch = 128 * 2 + 16;
    const uint32_t tile_u = wring_u;
    float sc8 = 1.f;
    if constexpr (kI8) {
      const int64_t n = (int64_t)nb * 128 + nl;
      sc8 = n < p.N_out ? __bfloat162float(p.scale8[n]) : 0.f;
    }
#pragma unroll 1
    for (int cc = 0; cc < NT / 32; ++cc) {
      uint32_t d[32];
      gm_ld_32x32b_x32(tb + ((uint32_t)(32 * warp) << 16) + 32 * cc, d);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        // int8: mm result rounded to bf16, THEN multiplied by the row scale and rounded again (plain_layout.py:256-262)
        const unsigned short b = kI8 ? __bfloat16_as_ushort(__float2bfloat16_rn(bf16r(__uint_as_float(d[j])) * sc8))
                                     : __bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(d[j])));
        asm volatile("st.shared.u16 [%0], %1;" ::"r"(tile_u + (uint32_t)(32 * cc + j) * kPitch + nl * 2), "h"(b) : "memory");
      }
    }
    gm_fence_before();
    named_bar_sync(1, kGmEpiWarps * 32);
    const int t = threadIdx.x, col = t & 15;
    const int64_t n0 = (int64_t)nb * 128 + col * 8;
#pragma unroll 4
    for (int r0 = 0; r0 < NT; r0 += 8) {
      const int row = r0 + (t >> 4);
      const int64_t m = (int64_t)mt * NT + row;
      if (m >= p.M || n0 >= p.N_out) continue;
      const uint4 v = lds128(tile_u + (uint32_t)row * kPitch + col * 16);
      __nv_bfloat16* dst = p.y + m * p.ldy + n0;
      if (p.vec_store && n0 + 8 <= p.N_out) {
        *reinterpret_cast<uint4*>(dst) = v;
      } else {
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (n0 + e < p.N_out) dst[e] = __ushort_as_bfloat16((unsigned short)(vv[e >> 1] >> (16 * (e & 1))));
      }
    }
    GM_TRACE(2, 0);
    }
  }
  __syncthreads();
  gm_cluster_sync();  // peers may still be arriving on this CTA's barriers
  if (warp == kGmMmaWarp) {
    gm_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(p.tmem_cols));
  }
}

// ---- split-K epilogue: y = bf16( sum over splits ) ; int8: y = bf16( bf16(sum) * scale[n] ) ------------------------------
__global__ void __launch_bounds__(256) split_reduce_kernel(const float* __restrict__ part, int splits, int64_t Mpad, int64_t Np,
                                                           int64_t M, int64_t N_out, const __nv_bfloat16* __restrict__ scale8,
                                                           __nv_bfloat16* __restrict__ y, int64_t ldy) {
  const int64_t n4 = (N_out + 3) / 4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M * n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / n4, n = (i - m * n4) * 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int sp = 0; sp < splits; ++sp) {  // fixed order: deterministic
      const float4 v = *reinterpret_cast<const float4*>(part + ((int64_t)sp * Mpad + m) * Np + n);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    const float r[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (n + e >= N_out) break;
      const float v = scale8 ? bf16r(r[e]) * __bfloat162float(scale8[n + e]) : r[e];
      y[m * ldy + n + e] = __float2bfloat16_rn(v);
    }
  }
}

// ---- host ---------------------------------------------------------------------------------------------------------
static int gm_tile_rows(int64_t M) { return M <= 64 ? 64 : (M <= 128 ? 128 : 256); }

// split-K plan: when (row blocks x prompt tiles) leaves most of the 148 SMs idle (small M and / or small N), the K range
// is cut into `splits` pieces of `kt_per` 64-k tiles (a multiple of 4, >= 16), each piece its own CTA (grid.z).  Pure
// function of the shape, so the workspace size is too.
static void gm_split_plan(int64_t M, int64_t Np, int64_t Kp, int& splits, int& kt_per) {
  const int NT = gm_tile_rows(M);
  const int64_t ctas = (Np / 128) * ((M + NT - 1) / NT);
  const int KT = (int)(Kp / 64);
  splits = 1;
  kt_per = KT;
  static const int enable = []() { const char* e = std::getenv("AO_B200_GEMM_SPLITK"); return e ? std::atoi(e) : 1; }();
  // measured (profiles/r01b_gemm_splitk.jsonl): pays for long K and few prompt rows (4096 x 14336, M = 32: 84 -> 45 us);
  // at K = 4096 the extra reduce launch costs more than the idle SMs, at M = 512 the partial sums are 25 MB
  if (!enable || ctas >= 112 || KT < 128 || M > 128) return;
  int want = (int)((148 + ctas - 1) / ctas);           // enough CTAs for one wave
  if (want > 8) want = 8;
  if (want > KT / 16) want = KT / 16;                   // at least 1024 k per split
  if (want < 2) return;
  kt_per = ((KT + want - 1) / want + 3) / 4 * 4;
  splits = (KT + kt_per - 1) / kt_per;
  if (splits < 2) { splits = 1; kt_per = KT; }
}

// bytes of the workspace the GEMM needs for M rows: the permuted-activation staging copy + (split-K) the fp32 partial
// sums (0: shape not handled by this kernel)
int64_t int4_gemm_tc_workspace_bytes(int64_t M, int64_t Np, int64_t Kp) {
  if (Np % 128 != 0 || Kp % 64 != 0 || M <= 0) return 0;
  const int NT = gm_tile_rows(M);
  const int64_t Mpad = (M + NT - 1) / NT * NT;
  int splits, kt_per;
  gm_split_plan(M, Np, Kp, splits, kt_per);
  return Mpad * Kp * 2 + (splits > 1 ? (int64_t)splits * Mpad * Np * 4 : 0);
}

template <int G, int NT>
static int launch_gemm_tc(GemmTcParams& p, int64_t Np, int splits, cudaStream_t s) {
  constexpr size_t kStageBytes = NT * 128;
  constexpr size_t kFixed = kGmHeader + kGmWSlots * kGmWSlotBytes;
  static int max_smem = -1;
  if (max_smem < 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  int stages = (int)(((size_t)max_smem - kFixed) / kStageBytes);
  const int max_by_tmem = (512 - NT) / 32;
  if (stages > max_by_tmem) stages = max_by_tmem;
  if (stages > kGmMaxStages) stages = kGmMaxStages;
  if (stages > p.kt_per_split) stages = p.kt_per_split;
  if (stages < 1) return AO_B200_ERR_UNSUPPORTED;
  p.stages = stages;
  int cols = 32;
  while (cols < NT + 32 * stages) cols <<= 1;
  p.tmem_cols = cols;
  size_t smem = kFixed + stages * kStageBytes;
  const size_t epi = kGmHeader + (size_t)NT * (128 * 2 + 16);  // the epilogue's transposed output tile overlays the rings
  if (smem < epi) smem = epi;
  auto kern = int4_gemm_tc_kernel<G, NT>;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess) return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  const int64_t n_blocks = Np / 128;
  static const int max_cl = [] {
    const char* e = std::getenv("AO_B200_GEMM_CLUSTER");
    return e ? std::atoi(e) : 4;
  }();
  p.cluster = (n_blocks % 4 == 0 && max_cl >= 4) ? 4 : ((n_blocks % 2 == 0 && max_cl >= 2) ? 2 : 1);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)n_blocks, (unsigned)(p.Mpad / NT), (unsigned)splits);
  cfg.blockDim = dim3(kGmThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.cluster;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, p) == cudaSuccess ? AO_B200_OK : AO_B200_ERR_CUDA;
}

// x [M][ldx] bf16 (first Kx columns valid), y [M][ldy] bf16; workspace: >= int4_gemm_tc_workspace_bytes(M, Np, Kp).
// G == 0: int8 weights (`sz` is then the bf16 per-row scale)
static int gemm_tc_any(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                       int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes,
                       cudaStream_t s) {
  const int64_t need = int4_gemm_tc_workspace_bytes(M, Np, Kp);
  if (need == 0 || Np / 128 > 2147483647 || M > ((int64_t)1 << 24)) return AO_B200_ERR_UNSUPPORTED;
  if (workspace == nullptr || workspace_bytes < need || !ao_aligned(workspace, 16)) return AO_B200_ERR_WORKSPACE;
  const int NT = gm_tile_rows(M);
  GemmTcParams p{};
  p.packed = static_cast<const uint8_t*>(packed);
  p.sz = G ? static_cast<const uint32_t*>(sz) : nullptr;
  p.scale8 = G ? nullptr : static_cast<const __nv_bfloat16*>(sz);
  p.xt = static_cast<const __nv_bfloat16*>(workspace);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.KT = (int)(Kp / 64);
  p.Mpad = (M + NT - 1) / NT * NT;
  p.N_out = N_out;
  p.ldy = ldy;
  p.Np = Np;
  p.vec_store = (ldy % 8 == 0 && ao_aligned(y, 16)) ? 1 : 0;
  if (p.Mpad / NT > 65535) return AO_B200_ERR_UNSUPPORTED;
  int splits = 1;
  gm_split_plan(M, Np, Kp, splits, p.kt_per_split);
  p.part = splits > 1 ? reinterpret_cast<float*>(static_cast<char*>(workspace) + p.Mpad * Kp * 2) : nullptr;
  const int64_t items = p.Mpad * p.KT;
  x_permute_kernel<<<(unsigned)((items + 255) / 256), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), M, Kx, ldx, p.Mpad, p.KT, NT,
                                                                   static_cast<__nv_bfloat16*>(workspace));
  if (cudaGetLastError() != cudaSuccess) return AO_B200_ERR_CUDA;
  int rc = AO_B200_ERR_UNSUPPORTED;
#define AO_GM(GG)                                                              \
  switch (NT) {                                                                \
    case 64: rc = launch_gemm_tc<GG, 64>(p, Np, splits, s); break;             \
    case 128: rc = launch_gemm_tc<GG, 128>(p, Np, splits, s); break;           \
    default: rc = launch_gemm_tc<GG, 256>(p, Np, splits, s); break;            \
  }                                                                            \
  break;
  switch (G) {
    case 0: AO_GM(0)
    case 32: AO_GM(32)
    case 64: AO_GM(64)
    case 128: AO_GM(128)
    case 256: AO_GM(256)
  }
#undef AO_GM
  if (rc != AO_B200_OK || splits == 1) return rc;
  const int64_t red_items = M * ((N_out + 3) / 4);
  const int64_t blocks = (red_items + 255) / 256;
  split_reduce_kernel<<<(unsigned)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, s>>>(p.part, splits, p.Mpad, Np, M, N_out, p.scale8,
                                                                                  p.y, ldy);
  return ao_launch_status();
}

int int4_gemm_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np, int64_t Kp,
                 int G, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, cudaStream_t s) {
  if (!(G == 32 || G == 64 || G == 128 || G == 256)) return AO_B200_ERR_UNSUPPORTED;
  return gemm_tc_any(x, M, Kx, ldx, packed, sz, Np, Kp, G, y, N_out, ldy, workspace, workspace_bytes, s);
}

// int8 weight-only prefill GEMM: packed = strip-major biased int8 (Np % 128 == 0), scale bf16 [N_out]
int int8_gemm_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Np,
                 int64_t Kp, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, cudaStream_t s) {
  return gemm_tc_any(x, M, Kx, ldx, packed, scale, Np, Kp, 0, y, N_out, ldy, workspace, workspace_bytes, s);
}

}  // namespace ao
