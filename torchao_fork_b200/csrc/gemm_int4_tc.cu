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

constexpr int kGmEpiWarps = 4, kGmDqWarps = 16;  // dequant warps: two sets of 8 (one strip each), alternating stages
constexpr int kGmProducerWarp = kGmEpiWarps + kGmDqWarps, kGmMmaWarp = kGmProducerWarp + 1;
constexpr int kGmThreads = (kGmMmaWarp + 1) * 32;  // 704
constexpr int kGmWeightStage = 8 * 512;            // 8 strips x one 64-k tile

struct GemmTcParams {
  const uint8_t* packed;     // strip-major [Np/16][Kp/64][512 B]
  const uint32_t* sz;        // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* xt;   // permuted activations [Kp/8 chunks][Mpad][8]
  __nv_bfloat16* y;          // [M][ldy]
  int M, KT, stages, tmem_cols, cluster;
  int64_t Mpad, N_out, ldy;
};

// ---- activations -> chunk-major, k-permuted copy ------------------------------------------------------------------
// xt[(kt * 8 + q) * Mpad + m][c * 2 + e] = x[m][64 kt + 32 half + 8 c + 2 i + e],  q = 4 half + i, rows >= M and
// columns >= Kx are zero.  One thread = one (row, 64-k tile): 128 bytes in, 8 x 16 bytes out (coalesced over rows).
__global__ void __launch_bounds__(256) x_permute_kernel(const __nv_bfloat16* __restrict__ x, int64_t M, int64_t Kx,
                                                        int64_t ldx, int64_t Mpad, int KT, __nv_bfloat16* __restrict__ xt) {
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
    *reinterpret_cast<uint4*>(xt + ((kt * 8 + q) * Mpad + m) * 8) = o;
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
// grid (Np / 128, Mpad / NT).  Shared memory: `stages` x [8 x 512 B packed weights | 8 chunks x NT x 16 B activations],
// then the mbarriers.  Tensor memory: columns [0, NT) accumulator, NT + 32 slot .. + 32: the A tile of ring slot `slot`.
template <int G, int NT>
__global__ void __launch_bounds__(kGmThreads, 1) int4_gemm_tc_kernel(const GemmTcParams p) {
  extern __shared__ __align__(1024) uint8_t gm_smem[];
  constexpr uint32_t kStageBytes = kGmWeightStage + NT * 128;
  constexpr uint32_t kChunk = NT * 16;  // bytes between consecutive 8-k chunks of the B tile (descriptor LBO)
  const int stages = p.stages, KT = p.KT;
  uint64_t* full = reinterpret_cast<uint64_t*>(gm_smem + (size_t)stages * kStageBytes);
  uint64_t* empty = full + stages;
  uint64_t* a_full = empty + stages;
  uint64_t* d_full = a_full + stages;
  uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(d_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nb = blockIdx.x, mt = blockIdx.y;

  if (warp == kGmMmaWarp) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], p.cluster);
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
  const int CL = p.cluster;
  const uint32_t crank = CL > 1 ? gm_cluster_rank() : 0u;
  const uint16_t cmask = (uint16_t)((1u << CL) - 1);
  const uint32_t tb = *reinterpret_cast<volatile uint32_t*>(tmem_base_s);
  const uint32_t smem0 = smem_u32(gm_smem);

  if (warp == kGmProducerWarp) {
    // ---- TMA producer: lanes 0..7 one weight strip each; lanes 8.. one activation chunk each, this CTA's share
    // (chunks q = rank, rank + CL, ...) multicast to the whole cluster ----------------------------------------------
    const uint8_t* wsrc = p.packed + tile_slot((int64_t)nb * 8 + (lane & 7), 0, KT) * 512;
    const int xq = (int)crank + (lane - 8) * CL;  // chunk this lane fetches (lanes 8 .. 8 + 8 / CL - 1)
    const __nv_bfloat16* xsrc = p.xt + ((int64_t)(xq & 7) * p.Mpad + (int64_t)mt * NT) * 8;
    for (int it = 0; it < KT; ++it) {
      const int slot = it % stages;
      if (it >= stages) mbar_wait(&empty[slot], ((it / stages) - 1) & 1);
      uint8_t* st = gm_smem + (size_t)slot * kStageBytes;
      if (lane == 0) mbar_arrive_expect_tx(&full[slot], kStageBytes);
      __syncwarp();
      if (lane < 8) bulk_g2s(st + lane * 512, wsrc + (int64_t)it * 512, 512, &full[slot]);
      else if (xq < 8)
        gm_bulk_g2s_multicast(st + kGmWeightStage + xq * kChunk, xsrc + (int64_t)it * 8 * p.Mpad * 8, kChunk, &full[slot], cmask);
    }
  } else if (warp == kGmMmaWarp) {
    // ---- MMA issue: 4 x (M128, N = NT, K16) per 64-k tile, whole warp in uniform control flow ---------------------
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)NT >> 3) << 17) | ((128u >> 4) << 24);
    constexpr uint64_t desc_hi = ((uint64_t)((kChunk >> 4) & 0x3FFF) << 16) | ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
    for (int it = 0; it < KT; ++it) {
      const int slot = it % stages;
      const uint32_t par = (it / stages) & 1;
      mbar_wait(&a_full[slot], par);
      mbar_wait(&full[slot], par);
      gm_fence_after();
      if (gm_elect_one()) {
        const uint32_t bx = smem0 + slot * kStageBytes + kGmWeightStage;
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
    }
  } else if (warp >= kGmEpiWarps) {
    // ---- dequant warps: packed words -> bf16 W' (reference arithmetic, one rounding) -> tensor memory -------------
    const int dq = (warp - kGmEpiWarps) & 7, set = (warp - kGmEpiWarps) >> 3, qd = warp & 3, strip = 2 * qd + (dq >> 2);
    const int g = lane >> 2;
    constexpr int kGroupsPerTile = G >= 64 ? 1 : 64 / G;
    const int n_groups = KT * 64 / G;
    const uint32_t* szp = p.sz + ((int64_t)(nb * 8 + strip) * n_groups) * 16 + g;
    const uint32_t taddr0 = tb + ((uint32_t)(32 * qd + 16 * (dq >> 2)) << 16) + NT;
    uint32_t szn[kGroupsPerTile][2];
    {
      const int64_t grp = G >= 64 ? (int64_t)set * 64 / G : (int64_t)set * kGroupsPerTile;
#pragma unroll
      for (int h = 0; h < kGroupsPerTile; ++h) {
        szn[h][0] = __ldg(szp + (grp + h) * 16);
        szn[h][1] = __ldg(szp + (grp + h) * 16 + 8);
      }
    }
    for (int it = set; it < KT; it += 2) {
      const int slot = it % stages;
      uint32_t s2[kGroupsPerTile][2], z2[kGroupsPerTile][2];
#pragma unroll
      for (int h = 0; h < kGroupsPerTile; ++h)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          s2[h][r] = __byte_perm(szn[h][r], 0, 0x1010);
          z2[h][r] = __byte_perm(szn[h][r], 0, 0x3232);
        }
      if (it + 2 < KT) {
        const int64_t grp = G >= 64 ? (int64_t)(it + 2) * 64 / G : (int64_t)(it + 2) * kGroupsPerTile;
#pragma unroll
        for (int h = 0; h < kGroupsPerTile; ++h) {
          szn[h][0] = __ldg(szp + (grp + h) * 16);
          szn[h][1] = __ldg(szp + (grp + h) * 16 + 8);
        }
      }
      mbar_wait(&full[slot], (it / stages) & 1);
      const uint4 w4 = lds128(smem0 + slot * kStageBytes + strip * 512 + lane * 16);
      const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
      uint32_t r[16];
#pragma unroll
      for (int j = 0; j < 4; ++j) {  // word j: row g + 8 (j & 1), k half j >> 1
        const int half = j >> 1, hrow = j & 1;
        const int h = kGroupsPerTile > 1 ? half * (kGroupsPerTile / 2) : 0;
        uint32_t o[4];
        dequant_word(w[j], s2[h][hrow], z2[h][hrow], o);
#pragma unroll
        for (int i = 0; i < 4; ++i) r[2 * (4 * half + i) + hrow] = o[i];
      }
      gm_st_16x128b_x8(taddr0 + slot * 32, r);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      gm_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_full[slot]);
    }
  } else {
    // ---- epilogue: thread = weight row (TMEM lane), 32 prompt rows per tcgen05.ld ---------------------------------
    mbar_wait(d_full, 0);
    gm_fence_after();
    const int64_t n = (int64_t)nb * 128 + 32 * warp + lane;
#pragma unroll 1
    for (int cc = 0; cc < NT / 32; ++cc) {
      uint32_t d[32];
      gm_ld_32x32b_x32(tb + ((uint32_t)(32 * warp) << 16) + 32 * cc, d);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      const int64_t m0 = (int64_t)mt * NT + 32 * cc;
      if (n < p.N_out) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (m0 + j < p.M) p.y[(m0 + j) * p.ldy + n] = __float2bfloat16_rn(__uint_as_float(d[j]));
      }
    }
    gm_fence_before();
  }
  __syncthreads();
  gm_cluster_sync();  // peers may still be arriving on this CTA's barriers
  if (warp == kGmMmaWarp) {
    gm_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(p.tmem_cols));
  }
}

// ---- host ---------------------------------------------------------------------------------------------------------
static int gm_tile_rows(int64_t M) { return M <= 64 ? 64 : (M <= 128 ? 128 : 256); }

// bytes of the permuted-activation staging buffer the GEMM needs for M rows (0: shape not handled by this kernel)
int64_t int4_gemm_tc_workspace_bytes(int64_t M, int64_t Np, int64_t Kp) {
  if (Np % 128 != 0 || Kp % 64 != 0 || M <= 0) return 0;
  const int NT = gm_tile_rows(M);
  const int64_t Mpad = (M + NT - 1) / NT * NT;
  return Mpad * Kp * 2;
}

template <int G, int NT>
static int launch_gemm_tc(GemmTcParams& p, int64_t Np, cudaStream_t s) {
  constexpr size_t kStageBytes = kGmWeightStage + NT * 128;
  static int max_smem = -1;
  if (max_smem < 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  int stages = (int)(((size_t)max_smem - 1024) / kStageBytes);
  const int max_by_tmem = (512 - NT) / 32;
  if (stages > max_by_tmem) stages = max_by_tmem;
  if (stages > 8) stages = 8;
  if (stages > p.KT) stages = p.KT;
  if (stages < 1) return AO_B200_ERR_UNSUPPORTED;
  p.stages = stages;
  int cols = 32;
  while (cols < NT + 32 * stages) cols <<= 1;
  p.tmem_cols = cols;
  const size_t smem = stages * kStageBytes + (3 * stages + 1) * 8 + 16;
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
  cfg.gridDim = dim3((unsigned)n_blocks, (unsigned)(p.Mpad / NT));
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

// x [M][ldx] bf16 (first Kx columns valid), y [M][ldy] bf16; workspace: >= int4_gemm_tc_workspace_bytes(M, Np, Kp)
int int4_gemm_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np, int64_t Kp,
                 int G, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, cudaStream_t s) {
  const int64_t need = int4_gemm_tc_workspace_bytes(M, Np, Kp);
  if (need == 0 || Np / 128 > 2147483647 || M > ((int64_t)1 << 24)) return AO_B200_ERR_UNSUPPORTED;
  if (workspace == nullptr || workspace_bytes < need || !ao_aligned(workspace, 16)) return AO_B200_ERR_WORKSPACE;
  const int NT = gm_tile_rows(M);
  GemmTcParams p{};
  p.packed = static_cast<const uint8_t*>(packed);
  p.sz = static_cast<const uint32_t*>(sz);
  p.xt = static_cast<const __nv_bfloat16*>(workspace);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.M = (int)M;
  p.KT = (int)(Kp / 64);
  p.Mpad = (M + NT - 1) / NT * NT;
  p.N_out = N_out;
  p.ldy = ldy;
  if (p.Mpad / NT > 65535) return AO_B200_ERR_UNSUPPORTED;
  const int64_t items = p.Mpad * p.KT;
  x_permute_kernel<<<(unsigned)((items + 255) / 256), 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), M, Kx, ldx, p.Mpad, p.KT,
                                                                   static_cast<__nv_bfloat16*>(workspace));
  if (cudaGetLastError() != cudaSuccess) return AO_B200_ERR_CUDA;
#define AO_GM(GG)                                                        \
  switch (NT) {                                                          \
    case 64: return launch_gemm_tc<GG, 64>(p, Np, s);                    \
    case 128: return launch_gemm_tc<GG, 128>(p, Np, s);                  \
    default: return launch_gemm_tc<GG, 256>(p, Np, s);                   \
  }
  switch (G) {
    case 32: AO_GM(32)
    case 64: AO_GM(64)
    case 128: AO_GM(128)
    case 256: AO_GM(256)
  }
#undef AO_GM
  return AO_B200_ERR_UNSUPPORTED;
}

}  // namespace ao
