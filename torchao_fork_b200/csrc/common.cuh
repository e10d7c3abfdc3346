// Shared device helpers for libao_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ao_b200.h"

#define AO_CHECK_ARG(cond) \
  do {                     \
    if (!(cond)) return AO_B200_ERR_INVALID_ARG; \
  } while (0)

static inline int ao_launch_status() {
  return cudaGetLastError() == cudaSuccess ? AO_B200_OK : AO_B200_ERR_CUDA;
}

static inline bool ao_aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

namespace ao {

constexpr int kNumSMs = 148;

// fp32 -> bf16 (nearest even) -> fp32: the rounding every torch eager op on a bf16 tensor applies.
__device__ __forceinline__ float bf16r(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float bf16_bits_to_float(uint32_t b) { return __uint_as_float(b << 16); }
__device__ __forceinline__ uint32_t float_to_bf16_bits(float x) {
  return static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(x)));
}

// 8 bf16 (one uint4) -> 8 floats
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t r[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(r[i] << 16);
    f[2 * i + 1] = __uint_as_float(r[i] & 0xFFFF0000u);
  }
}

__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t fma_bf16x2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// One packed word (8 nibbles, slot order k0,k2,k4,k6,k1,k3,k5,k7) -> 4 bf16x2 registers
// out[i] = ( W'[k0+2i], W'[k0+2i+1] ),  W' = bf16_rn((q-8)*s + z)   [single rounding].
// Same arithmetic as the reference's convert_i4x8_to_bf16x2x4 + __hfma2
// (tensor_core_tiled_layout.cu:41-117,194): (q | 0x4300) is the bf16 128+q, one exact
// bf16 fma with (1, -136) gives q-8, a second bf16 fma applies scale and zero.
__device__ __forceinline__ void dequant_word(uint32_t w, uint32_t s2, uint32_t z2, uint32_t (&out)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint32_t h;
    const uint32_t t = w >> (4 * i);
    asm("lop3.b32 %0, %1, 0x000f000f, 0x43004300, 0xea;" : "=r"(h) : "r"(t));  // (t & m) | magic
    h = fma_bf16x2(h, 0x3F803F80u, 0xC308C308u);                               // -> q - 8
    out[i] = fma_bf16x2(h, s2, z2);
  }
}

// 4 biased int8 (u = q + 128, one register) -> 2 bf16x2 registers (q0,q1), (q2,q3), exact: byte-insert into the fp32
// 1.5 * 2^23 magic, subtract, pack pairs
__device__ __forceinline__ void u8x4_to_bf16x2x2(uint32_t u, uint32_t& o01, uint32_t& o23) {
  constexpr uint32_t kMagic = 0x4B400000u;         // 1.5 * 2^23
  constexpr float kBias = 12582912.0f + 128.0f;
  const float f0 = __uint_as_float(__byte_perm(u, kMagic, 0x7650)) - kBias;
  const float f1 = __uint_as_float(__byte_perm(u, kMagic, 0x7651)) - kBias;
  const float f2 = __uint_as_float(__byte_perm(u, kMagic, 0x7652)) - kBias;
  const float f3 = __uint_as_float(__byte_perm(u, kMagic, 0x7653)) - kBias;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o01) : "f"(f1), "f"(f0));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o23) : "f"(f3), "f"(f2));
}

// D(16x8, f32) += A(16x16, bf16, row) * B(16x8, bf16, col)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                               uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- mbarrier / bulk-async-copy (TMA unit, 1-D) / programmatic-dependent-launch PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// shared-memory loads by 32-bit shared-window address (constant offsets fold into the instruction's immediate)
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ float2 lds64f(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// global -> shared bulk copy (bytes % 16 == 0, both addresses 16-byte aligned); completes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// PDL: wait until the grids this launch programmatically depends on have completed (no-op otherwise)
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// PDL: let the next kernel in the stream start launching (its pre-wait prologue overlaps our tail)
__device__ __forceinline__ void grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// strip-major layout addressing: 16-row strip s16, k-tile kt -> index of the 512-byte sub-tile.  All of a strip's
// sub-tiles are contiguous (K * 8 bytes), so a CTA that owns a range of strips streams one contiguous byte range.
__host__ __device__ __forceinline__ int64_t tile_slot(int64_t s16, int64_t kt, int64_t k_tiles) {
  return s16 * k_tiles + kt;
}
__host__ __device__ __forceinline__ void tile_of_slot(int64_t slot, int64_t k_tiles, int64_t& s16, int64_t& kt) {
  s16 = slot / k_tiles;
  kt = slot - s16 * k_tiles;
}
// scale_and_zero, strip-major: uint32 (scale | zero << 16) [Np/16][n_groups][16 rows]
__host__ __device__ __forceinline__ int64_t sz_index(int64_t grp, int64_t row, int64_t n_groups) {
  return ((row >> 4) * n_groups + grp) * 16 + (row & 15);
}

// nibble slot of value i (0..7) inside a word
__host__ __device__ __forceinline__ constexpr int nib_slot(int i) { return (i & 1) * 4 + (i >> 1); }

}  // namespace ao
