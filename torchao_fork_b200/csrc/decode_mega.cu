// One decode step of the gpt-fast Llama (torchao/_models/llama/model.py:342-486, generate.py:111-117) as ONE persistent,
// cooperative launch: every SM runs one CTA for the whole token, and the int4 weight stream never drains between
// projections.
//
// Why: with one launch per projection (round 1: 161 launches per token) each launch pays a fixed 3.5 - 5 us (dependency
// resolution, cold instruction cache, activation staging, pipeline fill, drain) around 1.3 - 9.6 us of streaming; the step
// ran at 0.43 of the HBM roofline.  Here a CTA is
//   * 1 producer warp that walks the CTA's share of EVERY weight of the model in order (bulk-copy ring in shared memory:
//     16 KB of nibbles + the groups' scale / zero per stage) and never waits for anything but a free ring slot: while the
//     consumers exchange activations between two projections, the ring fills with the next projection's weights;
//   * 16 consumer warps: integer tensor-core inner product of gemv_int4_imma.cu (nibbles -> bytes, mma.sync m16n8k32
//     u8 x s8, activations as three balanced int8 limbs of 23-bit fixed point, exact int32 group sums, affine dequant on
//     the accumulator in fp32) + the attention of the token + the staging of the next projection's activations;
//   * 1 epilogue warp: cross-warp sums, limb recombination, SiLU*mul, publishing the outputs.
//
// Hand-over between projections (every output element depends on the whole input vector: an all-to-all over the CTAs):
// activations travel through L2 as 8-byte PACKETS {2 bf16 | 32-bit tag} written with single 8-byte stores, tag = global
// phase number.  A packet is either old or complete, so no fence is needed anywhere; a per-phase arrival counter (one
// relaxed atomic per CTA) tells the pollers when a sweep is worth issuing, the tags make the sweep itself safe.  The
// residual stream h never leaves the SM: every CTA keeps a bf16 copy in shared memory and adds the published deltas
// (wo / w2 outputs) itself — bit-identical on every CTA, and the same code path sums the partial deltas of `world`
// tensor-parallel ranks (packets pushed into the peers' exchange buffers over NVLink, rank order, fp32).
//
// Arithmetic (parity with the reference, tests/test_mega.py): RMSNorm exactly as model.py:498-503,
// bf16(bf16(x * rsqrt(mean(x^2) + eps)) * w); GEMV with fp32-dequant semantics like gemv_int4_imma.cu; residual adds
// h = bf16(h + bf16(delta)); SiLU*mul bf16(bf16(silu(bf16 g)) * bf16 u); RoPE in fp32 on bf16 cos / sin, rounded to bf16;
// attention scores / softmax / PV in fp32, output rounded to bf16.  Non-finite activations propagate as NaN.
//
// Every wait in the kernel is bounded (globaltimer): on a time-out the CTA raises a sticky status word, stops waiting
// and runs its remaining control flow without blocking, so a bug or a dead peer ends the launch instead of hanging the GPU.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace ao {
namespace mega {

constexpr int kC = 16;                       // consumer warps
constexpr int kThreads = (kC + 2) * 32;      // + producer + epilogue
constexpr int kCT = kC * 32;
constexpr int kStageTiles = 32;              // 64-k tiles per stage: 2 per consumer warp
constexpr int kStageW = kStageTiles * 512;   // 16 KB of packed weights
constexpr int kRedBufs = 2;
constexpr uint32_t kMagicI = 0x4B400000u;    // bits of 1.5 * 2^23
constexpr float kMagicF = 12582912.0f;
constexpr int kMaxWorld = 8;
constexpr int kHG = 4;                       // q heads per attention pass
constexpr int kQPad = 136;                   // q row stride in floats (two 64-dim halves, the second shifted by 4 floats)
constexpr int kPartStride = 132;             // attention partial: 128 acc + m + l (+ pad)

enum { PH_GEMV = 0, PH_ATTN = 1 };
enum { IN_H = 0, IN_PK = 1 };
enum { OUT_PK = 0, OUT_DELTA = 1, OUT_SILU = 2, OUT_PLAIN = 3 };
constexpr uint32_t kNone = 0xffffffffu;

struct alignas(16) Phase {
  const uint8_t* packed;          // strip-major int4 [Np/16][Kp/64][512 B]
  const uint32_t* sz;             // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* norm_w;    // IN_H: RMSNorm weight [Kx]
  __nv_bfloat16* out_plain;       // OUT_PLAIN: bf16 [N_out]
  __nv_bfloat16* k_cache;         // PH_ATTN: [Hkv][S][128]
  __nv_bfloat16* v_cache;
  uint32_t in_off;                // packet offset in the exchange region: IN_PK input, IN_H deltas [world][Kx / 2] (kNone: no delta), ATTN qkv
  uint32_t out_off;               // packet offset of the output (OUT_DELTA: + rank * N_out / 2 in every peer's region)
  uint32_t aux_off;               // ATTN: split partials
  int type, in_kind, out_kind;
  int KT, n_strips, N_out, Kx, rot;
};

struct Params {
  const Phase* phases;
  int n_phases, n_cta;
  uint64_t* xchg[kMaxWorld];      // exchange regions: own at [rank], the peers' as mapped here
  int world, rank;
  const __nv_bfloat16* emb;
  const int64_t* tok;
  int32_t* pos;
  const __nv_bfloat16* freqs;     // [block_size][64][2] (cos, sin)
  uint64_t* counter;              // arrivals, monotonic over the life of the state buffer
  uint32_t* step;                 // launches so far
  uint32_t* status;               // sticky error bits
  int dim, H, Hkv, S, attn_bpc, attn_max_split;
  float eps, attn_scale;
  int stages;
  uint32_t off_x, off_S, off_red, off_h, off_misc;
  unsigned long long timeout_ns;
  unsigned long long* trace;      // optional: [n_cta][n_phases][8] clock64 stamps (benchmarks/mega_trace.py)
};

// ------------------------------------------------------------------------------------------------ small helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ uint32_t warp_max_u(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ void imma_16832(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                           uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint2 lds64u(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// packets: single 8-byte accesses that bypass the (incoherent) L1.  `.sys` scope: the exchange regions are also written
// by NVLink peers; on local memory it costs the same as `.gpu`.
__device__ __forceinline__ uint64_t ld_pk(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ ulonglong2 ld_pk2(const uint64_t* p) {
  ulonglong2 v;
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_pk(uint64_t* p, uint32_t data, uint32_t tag) {
  const uint64_t v = (uint64_t)data | ((uint64_t)tag << 32);
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_ctr(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_ctr(uint64_t* p) {
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], 1;" ::"l"(p) : "memory");
}

__device__ __forceinline__ void stamp(unsigned long long* trace, int n_phases, int i, int k) {
  if (trace != nullptr) trace[((size_t)blockIdx.x * n_phases + i) * 8 + k] = (unsigned long long)clock64();
}

struct Ctx {
  volatile uint32_t* abort_s;   // shared: this CTA stopped waiting
  uint32_t* status;             // global sticky status
  unsigned long long timeout_ns;
};
__device__ __noinline__ void raise_abort(const Ctx& c, uint32_t code) {
  *c.abort_s = 1u;
  atomicOr(c.status, code);
}
// bounded mbarrier wait
__device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity, const Ctx& c, uint32_t code) {
  const unsigned long long t0 = globaltimer();
  uint32_t n = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (*c.abort_s) return;
    if ((++n & 15u) == 0 && globaltimer() - t0 > c.timeout_ns) {
      raise_abort(c, code);
      return;
    }
  }
}
__device__ __forceinline__ void mbar_wait_b(uint64_t* bar, uint32_t parity, const Ctx& c, uint32_t code) {
  if (mbar_try_wait(bar, parity)) return;
  mbar_wait_slow(bar, parity, c, code);
}
// bounded wait for the arrival counter (one lane per CTA)
__device__ __noinline__ void wait_counter(const uint64_t* ctr, uint64_t target, const Ctx& c) {
  if (ld_ctr(ctr) >= target) return;
  const unsigned long long t0 = globaltimer();
  uint32_t n = 0;
  for (;;) {
    if (ld_ctr(ctr) >= target) return;
    if (*c.abort_s) return;
    __nanosleep(20);
    if ((++n & 63u) == 0) {
      if (globaltimer() - t0 > c.timeout_ns) {
        raise_abort(c, 2u);
        return;
      }
      if (*reinterpret_cast<volatile uint32_t*>(c.status) != 0u) {  // somebody else gave up: do not wait for them
        *c.abort_s = 1u;
        return;
      }
    }
  }
}
// bounded wait for one packet
__device__ __noinline__ uint32_t poll_pk_slow(const uint64_t* p, uint32_t tag, const Ctx& c) {
  const unsigned long long t0 = globaltimer();
  uint32_t n = 0;
  for (;;) {
    const uint64_t v = ld_pk(p);
    if ((uint32_t)(v >> 32) == tag) return (uint32_t)v;
    if (*c.abort_s) return 0u;
    if ((++n & 63u) == 0 && globaltimer() - t0 > c.timeout_ns) {
      raise_abort(c, 4u);
      return 0u;
    }
  }
}
__device__ __forceinline__ uint32_t poll_pk(const uint64_t* p, uint32_t tag, const Ctx& c) {
  const uint64_t v = ld_pk(p);
  if ((uint32_t)(v >> 32) == tag) return (uint32_t)v;
  return poll_pk_slow(p, tag, c);
}
// 4 consecutive packets (8 bf16 = one staging chunk); p is 32-byte aligned
__device__ __forceinline__ uint4 poll_chunk(const uint64_t* p, uint32_t tag, const Ctx& c) {
  const ulonglong2 a = ld_pk2(p), b = ld_pk2(p + 2);
  uint4 r = make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)b.x, (uint32_t)b.y);
  const bool ok = ((uint32_t)(a.x >> 32) == tag) & ((uint32_t)(a.y >> 32) == tag) & ((uint32_t)(b.x >> 32) == tag) &
                  ((uint32_t)(b.y >> 32) == tag);
  if (!ok) {
    r.x = poll_pk_slow(p, tag, c);
    r.y = poll_pk_slow(p + 1, tag, c);
    r.z = poll_pk_slow(p + 2, tag, c);
    r.w = poll_pk_slow(p + 3, tag, c);
  }
  return r;
}
__device__ __forceinline__ uint4 load_chunk(const uint64_t* p) {  // packets known to be present
  const ulonglong2 a = ld_pk2(p), b = ld_pk2(p + 2);
  return make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)b.x, (uint32_t)b.y);
}

__device__ __forceinline__ void strip_range(int b, int n_cta, int n_strips, int rot, int& s0, int& s1) {
  int bb = b + rot;
  if (bb >= n_cta) bb -= n_cta;
  const int base = n_strips / n_cta, rem = n_strips - base * n_cta;
  s0 = bb * base + min(bb, rem);
  s1 = s0 + base + (bb < rem ? 1 : 0);
}

template <int G>
struct Geom {
  static constexpr int kSegK = G < 128 ? G : 128;
  static constexpr int kNSeg = 128 / kSegK;
  static constexpr int kSegLanes = kSegK / 8;
  static constexpr int kStageSZ = (kStageTiles * 64 / G) * 64;
  static constexpr int kStageBytes = kStageW + kStageSZ;
};

// ------------------------------------------------------------------------------------------------ attention phase
// Attention.forward for the new token (model.py:405-440): one (kv head, sequence split) per CTA, the Gq q heads of the
// group share every K / V row.  16-position blocks: two lanes per position (64 dims each) for q.K, lane-parallel
// dims for P.V; the 16 warps' (m, l, acc) are merged through shared memory, splits through {fp32 | tag} packets.
// Its own function (not inlined): the walk wants the whole register file, the caller's state is parked once per phase.
__device__ __noinline__ void attention_phase(const Params& p, const Phase* ph, uint8_t* x_s, uint64_t* xq, int pos,
                                             uint32_t tag_in, uint32_t tag_out, const Ctx& ctx) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x;

  const int H = p.H, Hkv = p.Hkv, S = p.S;
  const int Gq = H / Hkv;
  const int len = min(pos + 1, S);
  const int nblk = (len + 15) >> 4;
  int nsplit = min(p.attn_max_split, (nblk + p.attn_bpc - 1) / p.attn_bpc);
  const int bps = (nblk + nsplit - 1) / nsplit;
  nsplit = (nblk + bps - 1) / bps;
  const bool active = b < Hkv * nsplit;
  if (active) {
    const int kvh = b % Hkv, split = b / Hkv;
    const bool owner = split == (nblk - 1) / bps && pos < S;
    float* q_s = reinterpret_cast<float*>(x_s);                 // [Gq][kQPad]
    float* part_s = q_s + ((Gq * kQPad + 31) & ~31);            // [16 warps][kHG][kPartStride]
    __nv_bfloat16* kc = ph->k_cache + (int64_t)kvh * S * 128;
    __nv_bfloat16* vc = ph->v_cache + (int64_t)kvh * S * 128;
    const uint64_t* qkv = xq + ph->in_off;
    for (int idx = tid; idx < (Gq + 2) * 64; idx += kCT) {
      const int which = idx >> 6, pr = idx & 63;
      const int elem = which < Gq ? (kvh * Gq + which) * 128 : (which == Gq ? (H + kvh) * 128 : (H + Hkv + kvh) * 128);
      const uint32_t raw = poll_pk(qkv + (elem >> 1) + pr, tag_in, ctx);
      uint32_t o = raw;
      float r0 = 0.f, r1 = 0.f;
      if (which <= Gq) {
        const uint32_t cs = __ldg(reinterpret_cast<const uint32_t*>(p.freqs) + (int64_t)min(pos, S - 1) * 64 + pr);
        const float x0 = __uint_as_float(raw << 16), x1 = __uint_as_float(raw & 0xFFFF0000u);
        const float cc = __uint_as_float(cs << 16), ss = __uint_as_float(cs & 0xFFFF0000u);
        r0 = bf16r(x0 * cc - x1 * ss);
        r1 = bf16r(x1 * cc + x0 * ss);
        o = float_to_bf16_bits(r0) | (float_to_bf16_bits(r1) << 16);
      }
      if (which < Gq) {
        const int d = 2 * pr + (pr >= 32 ? 4 : 0);
        *reinterpret_cast<float2*>(q_s + which * kQPad + d) = make_float2(r0 * p.attn_scale, r1 * p.attn_scale);
      } else if (owner) {
        __nv_bfloat16* dst = (which == Gq ? kc : vc) + (int64_t)pos * 128 + 2 * pr;
        *reinterpret_cast<uint32_t*>(dst) = o;
      }
    }
    named_bar_sync(1, kCT);  // q in shared memory; the new k / v row is visible to this CTA
    const int blk_end = min((split + 1) * bps, nblk);
    const int pj_idx = lane >> 1, half = lane & 1;
    for (int hp = 0; hp < Gq; hp += kHG) {
      float m[kHG], l[kHG], acc[kHG][4];
#pragma unroll
      for (int h = 0; h < kHG; ++h) {
        m[h] = -INFINITY;
        l[h] = 0.f;
        acc[h][0] = acc[h][1] = acc[h][2] = acc[h][3] = 0.f;
      }
      for (int blk = split * bps + warp; blk < blk_end; blk += kC) {
        const int j = blk * 16 + pj_idx;
        const bool valid = j < len;
        uint4 kk[8];
        {
          const uint4* krow = reinterpret_cast<const uint4*>(kc + (int64_t)(valid ? j : blk * 16) * 128 + half * 64);
#pragma unroll
          for (int q = 0; q < 8; ++q) kk[q] = krow[q];
        }
        float sc[kHG];
#pragma unroll
        for (int h = 0; h < kHG; ++h) sc[h] = 0.f;
        // (asm volatile shared loads: q is block-invariant, and the compiler would otherwise hoist all 64 loads = 256
        // registers out of the walk)
        const uint32_t qbase = smem_u32(q_s) + (uint32_t)half * 272u;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float k0 = __uint_as_float(kk[q].x << 16), k1 = __uint_as_float(kk[q].x & 0xFFFF0000u);
          const float k2 = __uint_as_float(kk[q].y << 16), k3 = __uint_as_float(kk[q].y & 0xFFFF0000u);
          const float k4 = __uint_as_float(kk[q].z << 16), k5 = __uint_as_float(kk[q].z & 0xFFFF0000u);
          const float k6 = __uint_as_float(kk[q].w << 16), k7 = __uint_as_float(kk[q].w & 0xFFFF0000u);
#pragma unroll
          for (int h = 0; h < kHG; ++h) {
            const uint32_t qr = qbase + (uint32_t)(min(hp + h, Gq - 1) * kQPad + 8 * q) * 4u;
            const uint4 qa = lds128(qr), qb = lds128(qr + 16u);
            sc[h] += __uint_as_float(qa.x) * k0 + __uint_as_float(qa.y) * k1 + __uint_as_float(qa.z) * k2 +
                     __uint_as_float(qa.w) * k3 + __uint_as_float(qb.x) * k4 + __uint_as_float(qb.y) * k5 +
                     __uint_as_float(qb.z) * k6 + __uint_as_float(qb.w) * k7;
          }
        }
#pragma unroll
        for (int h = 0; h < kHG; ++h) {
          const float s = sc[h] + __shfl_xor_sync(0xffffffffu, sc[h], 1);
          sc[h] = valid ? s : -INFINITY;
        }
        float pj[kHG];
#pragma unroll
        for (int h = 0; h < kHG; ++h) {
          float mb = sc[h];
#pragma unroll
          for (int o = 16; o > 1; o >>= 1) mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
          const float m_new = fmaxf(m[h], mb);  // finite: the block holds at least one position
          const float corr = __expf(m[h] - m_new);
          pj[h] = valid ? __expf(sc[h] - m_new) : 0.f;
          float ps = pj[h];
#pragma unroll
          for (int o = 16; o > 1; o >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, o);
          l[h] = l[h] * corr + ps;
          acc[h][0] *= corr; acc[h][1] *= corr; acc[h][2] *= corr; acc[h][3] *= corr;
          m[h] = m_new;
        }
        const int cnt = min(16, len - blk * 16);
        uint2 vr[16];
#pragma unroll
        for (int jj = 0; jj < 16; ++jj)
          vr[jj] = jj < cnt ? *reinterpret_cast<const uint2*>(vc + (int64_t)(blk * 16 + jj) * 128 + lane * 4)
                            : make_uint2(0u, 0u);
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
          const float v0 = __uint_as_float(vr[jj].x << 16), v1 = __uint_as_float(vr[jj].x & 0xFFFF0000u);
          const float v2 = __uint_as_float(vr[jj].y << 16), v3 = __uint_as_float(vr[jj].y & 0xFFFF0000u);
#pragma unroll
          for (int h = 0; h < kHG; ++h) {
            const float pw = __shfl_sync(0xffffffffu, pj[h], 2 * jj);
            acc[h][0] += pw * v0; acc[h][1] += pw * v1; acc[h][2] += pw * v2; acc[h][3] += pw * v3;
          }
        }
      }
#pragma unroll
      for (int h = 0; h < kHG; ++h) {
        float* pp = part_s + (warp * kHG + h) * kPartStride;
        *reinterpret_cast<float4*>(pp + lane * 4) = make_float4(acc[h][0], acc[h][1], acc[h][2], acc[h][3]);
        if (lane == 0) {
          pp[128] = m[h];
          pp[129] = l[h];
        }
      }
      named_bar_sync(1, kCT);
      if (warp < kHG && hp + warp < Gq) {
        // merge the 16 warps' partials of head hp + warp
        const int head = kvh * Gq + hp + warp;
        float ms = -INFINITY, ls = 0.f;
        if (lane < kC) {
          ms = part_s[(lane * kHG + warp) * kPartStride + 128];
          ls = part_s[(lane * kHG + warp) * kPartStride + 129];
        }
        float mm = ms;
#pragma unroll
        for (int o = 16; o; o >>= 1) mm = fmaxf(mm, __shfl_xor_sync(0xffffffffu, mm, o));
        const float wgt = ls > 0.f ? __expf(ms - mm) : 0.f;
        float ll = warp_sum(ls * wgt);
        float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int s = 0; s < kC; ++s) {
          const float ws = __shfl_sync(0xffffffffu, wgt, s);
          const float4 po = *reinterpret_cast<const float4*>(part_s + (s * kHG + warp) * kPartStride + lane * 4);
          a[0] += po.x * ws; a[1] += po.y * ws; a[2] += po.z * ws; a[3] += po.w * ws;
        }
        if (nsplit > 1) {
          // partial packets {fp32 | tag}: [head][split][kPartStride]
          uint64_t* pp = xq + ph->aux_off + ((size_t)head * p.attn_max_split) * kPartStride;
          if (split > 0) {
            uint64_t* mine = pp + (size_t)split * kPartStride;
#pragma unroll
            for (int e = 0; e < 4; ++e) st_pk(mine + lane * 4 + e, __float_as_uint(a[e]), tag_out);
            if (lane == 0) st_pk(mine + 128, __float_as_uint(mm), tag_out);
            if (lane == 1) st_pk(mine + 129, __float_as_uint(ll), tag_out);
          } else {
            for (int s = 1; s < nsplit; ++s) {
              const uint64_t* theirs = pp + (size_t)s * kPartStride;
              const float m2 = __uint_as_float(poll_pk(theirs + 128, tag_out, ctx));
              const float l2 = __uint_as_float(poll_pk(theirs + 129, tag_out, ctx));
              float a2[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) a2[e] = __uint_as_float(poll_pk(theirs + lane * 4 + e, tag_out, ctx));
              const float mn = fmaxf(mm, m2);
              const float w1 = __expf(mm - mn), w2 = l2 > 0.f ? __expf(m2 - mn) : 0.f;
              ll = ll * w1 + l2 * w2;
#pragma unroll
              for (int e = 0; e < 4; ++e) a[e] = a[e] * w1 + a2[e] * w2;
              mm = mn;
            }
          }
        }
        if (split == 0) {
          const float inv = 1.f / ll;
          const uint32_t o0 = float_to_bf16_bits(a[0] * inv) | (float_to_bf16_bits(a[1] * inv) << 16);
          const uint32_t o1 = float_to_bf16_bits(a[2] * inv) | (float_to_bf16_bits(a[3] * inv) << 16);
          uint64_t* dst = xq + ph->out_off + ((head * 128 + lane * 4) >> 1);
          st_pk(dst, o0, tag_out);
          st_pk(dst + 1, o1, tag_out);
        }
      }
      named_bar_sync(1, kCT);  // part_s is reused by the next pass / the next phase's staging
    }
  }
}

// ------------------------------------------------------------------------------------------------ the kernel
template <int G>
__global__ void __launch_bounds__(kThreads, 1) decode_mega_kernel(const Params p) {
  using GE = Geom<G>;
  constexpr int kStageBytes = GE::kStageBytes;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* x_s = smem + p.off_x;                                  // limbs [3][Kp + 32] int8 / attention scratch
  float* S_s = reinterpret_cast<float*>(smem + p.off_S);          // [Kp / kSegK][8]: limb sums per segment
  float* red = reinterpret_cast<float*>(smem + p.off_red);        // [2][16 warps][32][4]
  uint4* h_s = reinterpret_cast<uint4*>(smem + p.off_h);          // residual stream, bf16 [dim]
  float* part = reinterpret_cast<float*>(smem + p.off_misc);      // [16 warps][2]: {sum sq, max |.| bits}
  float* row_scale = part + 2 * kC;                               // [2] (GEMV-phase parity)
  volatile uint32_t* abort_s = reinterpret_cast<volatile uint32_t*>(row_scale + 2);
  uint64_t* full = reinterpret_cast<uint64_t*>(row_scale + 4);
  uint64_t* empty = full + p.stages;
  uint64_t* red_full = empty + p.stages;
  uint64_t* red_free = red_full + kRedBufs;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x, n_cta = p.n_cta;
  const int stages = p.stages;
  const int n_phases = p.n_phases;
  const Phase* __restrict__ phases = p.phases;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kC);
    }
    for (int s = 0; s < kRedBufs; ++s) {
      mbar_init(&red_full[s], kC);
      mbar_init(&red_free[s], 1);
    }
    *abort_s = 0u;
    fence_barrier_init();
  }
  __syncthreads();
  Ctx ctx{abort_s, p.status, p.timeout_ns};

  const uint32_t step = *reinterpret_cast<volatile uint32_t*>(p.step);
  const int pos = *reinterpret_cast<volatile int32_t*>(p.pos);
  const uint32_t tag0 = step * (uint32_t)n_phases;                           // phase i publishes with tag tag0 + i + 1
  const uint64_t cnt0 = (uint64_t)step * (uint64_t)n_phases * (uint64_t)n_cta;  // counter before this launch
  uint64_t* xq = p.xchg[p.rank];

  if (warp == kC) {
    // ===================== producer: the CTA's share of every weight, in phase order =====================
    if (lane == 0) {
      int slot = 0;
      uint32_t par = 1;
      for (int i = 0; i < n_phases; ++i) {
        const Phase* ph = phases + i;
        if (ph->type != PH_GEMV) continue;
        const int KT = ph->KT;
        int s0, s1;
        strip_range(b, n_cta, ph->n_strips, ph->rot, s0, s1);
        const int n_groups = KT * 64 / G;
        const uint8_t* packed = ph->packed;
        const uint32_t* sz = ph->sz;
        for (int strip = s0; strip < s1; ++strip) {
          const uint8_t* wrow = packed + (int64_t)strip * KT * 512;
          const uint32_t* srow = sz + (int64_t)strip * n_groups * 16;
          for (int kt0 = 0; kt0 < KT; kt0 += kStageTiles) {
            const int cnt = min(kStageTiles, KT - kt0);
            uint8_t* dst = ring + (size_t)slot * kStageBytes;
            mbar_wait_b(&empty[slot], par, ctx, 8u);
            const uint32_t sz_bytes = (uint32_t)(cnt * 64 / G) * 64u;
            mbar_arrive_expect_tx(&full[slot], (uint32_t)(cnt * 512) + sz_bytes);
            bulk_g2s(dst + kStageW, srow + (int64_t)(kt0 * 64 / G) * 16, sz_bytes, &full[slot]);
            bulk_g2s(dst, wrow + (int64_t)kt0 * 512, (uint32_t)(cnt * 512), &full[slot]);
            if (++slot == stages) {
              slot = 0;
              par ^= 1u;
            }
          }
        }
        stamp(p.trace, n_phases, i, 5);
      }
    }
    return;
  }

  if (warp == kC + 1) {
    // ===================== epilogue warp =====================
    const int g = lane >> 2, c = lane & 3;
    uint32_t li = 0;   // strips handed over so far (all phases)
    int gphase = 0;    // GEMV phases so far
    for (int i = 0; i < n_phases; ++i) {
      const Phase* ph = phases + i;
      if (ph->type != PH_GEMV) continue;
      int s0, s1;
      strip_range(b, n_cta, ph->n_strips, ph->rot, s0, s1);
      const int out_kind = ph->out_kind;
      const int N_out = ph->N_out;
      const uint32_t tag = tag0 + (uint32_t)i + 1u;
      const uint32_t out_off = ph->out_off;
      __nv_bfloat16* out_plain = ph->out_plain;
      for (int strip = s0; strip < s1; ++strip, ++li) {
        const int rbuf = (int)(li & 1u);
        const uint32_t use = li >> 1;
        const float* rb = red + (size_t)rbuf * (kC * 128);
        mbar_wait_b(&red_full[rbuf], use & 1u, ctx, 16u);
        uint64_t vlo = 0ull, vhi = 0ull;
#pragma unroll
        for (int w = 0; w < kC; ++w) {
          const ulonglong2 t = *reinterpret_cast<const ulonglong2*>(rb + (w * 32 + lane) * 4);
          asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vlo) : "l"(t.x));
          asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vhi) : "l"(t.y));
        }
        const float4 v = make_float4(__uint_as_float((uint32_t)vlo), __uint_as_float((uint32_t)(vlo >> 32)),
                                     __uint_as_float((uint32_t)vhi), __uint_as_float((uint32_t)(vhi >> 32)));
        __syncwarp();
        if (lane == 0) mbar_arrive(&red_free[rbuf]);
        // limb recombination (M = 1): lane (g, c = 0) holds limbs 0, 1 of rows g (x, y) and g + 8 (z, w); limb 2 sits in
        // lane (g, c = 1).x / .z
        const float t2x = __shfl_down_sync(0xffffffffu, v.x, 1), t2z = __shfl_down_sync(0xffffffffu, v.z, 1);
        const float rsc = row_scale[gphase & 1];
        const float v0 = fmaf(65536.f, v.x, fmaf(256.f, v.y, t2x)) * rsc;   // row g      (valid on lanes c == 0)
        const float v1 = fmaf(65536.f, v.z, fmaf(256.f, v.w, t2z)) * rsc;   // row g + 8
        if (out_kind == OUT_SILU) {
          // rows g = gate (w1), rows g + 8 = up (w3) of output n = 8 strip + g   (model.py:486)
          const float gate = bf16r(v0), up = bf16r(v1);
          const float sg = bf16r(gate / (1.f + __expf(-gate)));
          const uint32_t o = float_to_bf16_bits(sg * up);
          const uint32_t on = __shfl_down_sync(0xffffffffu, o, 4);  // row g + 1
          const int n = 8 * strip + g;
          if (c == 0 && (g & 1) == 0 && n < N_out) st_pk(xq + out_off + (n >> 1), o | (on << 16), tag);
        } else if (out_kind == OUT_PLAIN) {
          const int n0 = 16 * strip + g;
          if (c == 0) {
            if (n0 < N_out) out_plain[n0] = __float2bfloat16_rn(v0);
            if (n0 + 8 < N_out) out_plain[n0 + 8] = __float2bfloat16_rn(v1);
          }
        } else {
          const uint32_t o0 = float_to_bf16_bits(v0), o1 = float_to_bf16_bits(v1);
          const uint32_t o0n = __shfl_down_sync(0xffffffffu, o0, 4), o1n = __shfl_down_sync(0xffffffffu, o1, 4);
          const int n0 = 16 * strip + g;
          if (c == 0 && (g & 1) == 0) {
            const uint32_t d0 = o0 | (o0n << 16), d1 = o1 | (o1n << 16);
            if (out_kind == OUT_PK) {
              if (n0 < N_out) st_pk(xq + out_off + (n0 >> 1), d0, tag);
              if (n0 + 8 < N_out) st_pk(xq + out_off + ((n0 + 8) >> 1), d1, tag);
            } else {  // OUT_DELTA: this rank's slot in every rank's region (own first, then staggered over the peers)
              const uint32_t off = out_off + (uint32_t)p.rank * (uint32_t)(N_out >> 1);
              for (int d = 0; d < p.world; ++d) {
                int r = p.rank + d;
                if (r >= p.world) r -= p.world;
                uint64_t* dst = p.xchg[r] + off;
                if (n0 < N_out) st_pk(dst + (n0 >> 1), d0, tag);
                if (n0 + 8 < N_out) st_pk(dst + ((n0 + 8) >> 1), d1, tag);
              }
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        red_ctr(p.counter);
        stamp(p.trace, n_phases, i, 4);
      }
      ++gphase;
    }
    if (b == 0 && lane == 0) {  // every CTA read both at its start, long before CTA 0 can get here
      *reinterpret_cast<volatile int32_t*>(p.pos) = pos + 1;
      *reinterpret_cast<volatile uint32_t*>(p.step) = step + 1u;
    }
    return;
  }

  // ===================== consumers =====================
  const int g = lane >> 2, c = lane & 3;
  {
    // residual stream := embedding row of the fed token (model.py:358)
    const int64_t tok = *reinterpret_cast<const volatile int64_t*>(p.tok);
    const uint4* erow = reinterpret_cast<const uint4*>(p.emb + tok * (int64_t)p.dim);
    for (int ci = tid; ci < p.dim / 8; ci += kCT) h_s[ci] = __ldg(erow + ci);
  }
  const uint32_t ring_u = smem_u32(ring);
  int slot = 0;
  uint32_t par = 0;
  uint32_t stage_u = ring_u;
  uint32_t li = 0;
  int gphase = 0;

  for (int i = 0; i < n_phases; ++i) {
    const Phase* ph = phases + i;
    const uint32_t tag_in = tag0 + (uint32_t)i;  // tag of phase i - 1
    const uint32_t tag_out = tag_in + 1u;
    // RMSNorm weights do not depend on anybody: request them before waiting (they are a DRAM miss, once per token)
    const bool is_gemv = ph->type == PH_GEMV;
    const int in_kind = ph->in_kind;
    const int Kx = ph->Kx;
    const __nv_bfloat16* norm_w = ph->norm_w;
    uint4 wn_first = make_uint4(0u, 0u, 0u, 0u);
    if (is_gemv && in_kind == IN_H && tid * 8 < Kx) wn_first = __ldg(reinterpret_cast<const uint4*>(norm_w) + tid);
    if (tid == 0) stamp(p.trace, n_phases, i, 0);
    if (i > 0) {
      if (tid == 0) wait_counter(p.counter, cnt0 + (uint64_t)i * (uint64_t)n_cta, ctx);
    }
    named_bar_sync(1, kCT);  // B1: phase i - 1 is published; every consumer warp of this CTA has left its main loop

    if (!is_gemv) {
      // ================= attention of the new token (model.py:405-440), one kv head x sequence split per CTA =================
      if (tid == 0) stamp(p.trace, n_phases, i, 1);
      attention_phase(p, ph, x_s, xq, pos, tag_in, tag_out, ctx);
      if (tid == 0) {
        red_ctr(p.counter);
        stamp(p.trace, n_phases, i, 4);
      }
      continue;
    }

    // ================= GEMV phase: stage the activation row as int8 limbs =================
    const int KT = ph->KT;
    const int chunks_p = KT * 8;                 // 8-element chunks over Kp
    const int n_chunks = Kx >> 3;
    const uint32_t limb_row = (uint32_t)KT * 64u + 32u;
    const uint64_t* in_pk = ph->in_off == kNone ? nullptr : xq + ph->in_off;
    if (tid == 0) stamp(p.trace, n_phases, i, 1);
    float ss = 0.f;
    uint32_t amax_u = 0u;  // max |.| as bits: an Inf / NaN input shows up as >= 0x7f800000
    if (in_kind == IN_H) {
      // h += sum over ranks of the published deltas (rank order, fp32), then statistics of h and of h * w
      for (int ci = tid; ci < n_chunks; ci += kCT) {
        uint4 hv = h_s[ci];
        if (in_pk != nullptr) {
          float hf[8], acc[8];
          unpack8(hv, hf);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = 0.f;
          for (int r0 = 0; r0 < p.world; r0 += 4) {
            uint4 d[4];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
              if (r0 + rr < p.world) d[rr] = poll_chunk(in_pk + (size_t)(r0 + rr) * (Kx >> 1) + ci * 4, tag_in, ctx);
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
              if (r0 + rr < p.world) {
                float df[8];
                unpack8(d[rr], df);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] += df[e];
              }
          }
          uint32_t o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e)
            o[e] = float_to_bf16_bits(hf[2 * e] + acc[2 * e]) | (float_to_bf16_bits(hf[2 * e + 1] + acc[2 * e + 1]) << 16);
          hv = make_uint4(o[0], o[1], o[2], o[3]);
          h_s[ci] = hv;
        }
        const uint4 wn = ci == tid ? wn_first : __ldg(reinterpret_cast<const uint4*>(norm_w) + ci);
        float f[8], wf[8];
        unpack8(hv, f);
        unpack8(wn, wf);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          ss = fmaf(f[e], f[e], ss);
          amax_u = max(amax_u, __float_as_uint(f[e] * wf[e]) & 0x7fffffffu);
        }
      }
    } else {
      for (int ci = tid; ci < n_chunks; ci += kCT) {
        const uint4 xv = poll_chunk(in_pk + ci * 4, tag_in, ctx);
        float f[8];
        unpack8(xv, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) amax_u = max(amax_u, __float_as_uint(f[e]) & 0x7fffffffu);
      }
    }
    ss = warp_sum(ss);
    amax_u = warp_max_u(amax_u);
    if (lane == 0) {
      part[warp * 2] = ss;
      part[warp * 2 + 1] = __uint_as_float(amax_u);
    }
    named_bar_sync(1, kCT);  // B2
    if (tid == 0) stamp(p.trace, n_phases, i, 6);
    float up, rnorm = 1.f;
    {
      float t = 0.f;
      uint32_t am = 0u;
      if (lane < kC) {
        t = part[lane * 2];
        am = __float_as_uint(part[lane * 2 + 1]);
      }
      t = warp_sum(t);
      am = warp_max_u(am);
      float bound = __uint_as_float(am);
      bool bad = am >= 0x7f800000u;
      if (in_kind == IN_H) {
        rnorm = rsqrtf(t / (float)Kx + p.eps);
        // |bf16(bf16(x r) w)| <= |x w| r (1 + 2^-8)^2: a bound within 2x of the true maximum costs at most one of the 23 bits
        bound = bound * rnorm * 1.0079f;
        bad = bad || !(bound < 3.0e38f);
      }
      // 2^e brings the bound into [2^21, 2^22): |round(x 2^e)| <= 2^22 - 2^14 for every bf16 x below the bound
      int e = 0;
      if (bound > 0.f) e = 21 - ((int)((__float_as_uint(bound) >> 23) & 0xFF) - 127);
      e = max(-106, min(100, e));
      up = __uint_as_float((uint32_t)(e + 127) << 23);
      if (warp == 0 && lane == 0)
        row_scale[gphase & 1] = bad ? __uint_as_float(0x7fc00000u) : __uint_as_float((uint32_t)(127 - e) << 23);
    }
    {
      uint8_t* rows = x_s;
      const int n_iter = (chunks_p + kCT - 1) / kCT;
      for (int it = 0; it < n_iter; ++it) {
        const int ci = it * kCT + tid;
        const bool valid = ci < chunks_p;
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = 0.f;
        if (ci < n_chunks) {
          if (in_kind == IN_H) {
            float wf[8];
            unpack8(h_s[ci], f);
            unpack8(__ldg(reinterpret_cast<const uint4*>(norm_w) + ci), wf);
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = bf16r(bf16r(f[e] * rnorm) * wf[e]);   // RMSNorm.forward, model.py:498-503
          } else {
            unpack8(load_chunk(in_pk + ci * 4), f);
          }
        }
        int sl0 = 0, sl1 = 0, sl2 = 0;
        if (valid) {
          // balanced limbs without carries: X + 0x808080 = sum (a_i + 128) 256^i (see gemv_int4_imma.cu)
          uint32_t v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = __float_as_uint(fmaf(f[e], up, kMagicF)) + 0x00408080u;
          uint32_t lw[3][2];
#pragma unroll
          for (int l = 0; l < 3; ++l) {
            const uint32_t sel = l == 0 ? 0x0062u : (l == 1 ? 0x0051u : 0x0040u);
            const uint32_t p04 = __byte_perm(v[0], v[4], sel), p15 = __byte_perm(v[1], v[5], sel);
            const uint32_t p26 = __byte_perm(v[2], v[6], sel), p37 = __byte_perm(v[3], v[7], sel);
            lw[l][0] = __byte_perm(p04, p15, 0x5410) ^ 0x80808080u;
            lw[l][1] = __byte_perm(p26, p37, 0x5410) ^ 0x80808080u;
          }
          sl0 = __dp4a((int)lw[0][0], 0x01010101, __dp4a((int)lw[0][1], 0x01010101, 0));
          sl1 = __dp4a((int)lw[1][0], 0x01010101, __dp4a((int)lw[1][1], 0x01010101, 0));
          sl2 = __dp4a((int)lw[2][0], 0x01010101, __dp4a((int)lw[2][1], 0x01010101, 0));
          *reinterpret_cast<uint2*>(rows + (size_t)ci * 8) = make_uint2(lw[0][0], lw[0][1]);
          *reinterpret_cast<uint2*>(rows + limb_row + (size_t)ci * 8) = make_uint2(lw[1][0], lw[1][1]);
          *reinterpret_cast<uint2*>(rows + 2 * limb_row + (size_t)ci * 8) = make_uint2(lw[2][0], lw[2][1]);
        }
#pragma unroll
        for (int o = GE::kSegLanes / 2; o; o >>= 1) {
          sl0 += __shfl_xor_sync(0xffffffffu, sl0, o);
          sl1 += __shfl_xor_sync(0xffffffffu, sl1, o);
          sl2 += __shfl_xor_sync(0xffffffffu, sl2, o);
        }
        if (valid && (ci % GE::kSegLanes) == 0) {
          float* Sp = S_s + (ci / GE::kSegLanes) * 8;
          Sp[0] = (float)sl0; Sp[1] = (float)sl1; Sp[2] = (float)sl2; Sp[3] = 0.f;
        }
      }
    }
    named_bar_sync(1, kCT);  // B3: limbs, segment sums and the row scale are in place
    if (tid == 0) stamp(p.trace, n_phases, i, 2);

    // ================= main loop: warp w owns k-tiles 2 w, 2 w + 1 (128 k) of every stage =================
    int s0, s1;
    strip_range(b, n_cta, ph->n_strips, ph->rot, s0, s1);
    const uint32_t w_off = (uint32_t)warp * 1024u + (uint32_t)lane * 16u;
    const uint32_t sz_off = (uint32_t)kStageW + (uint32_t)(G >= 128 ? warp * 128 / G : warp * GE::kNSeg) * 64u + (uint32_t)g * 4u;
    const uint32_t S_u0 = smem_u32(S_s) + (uint32_t)(warp * GE::kNSeg) * 32u + (uint32_t)c * 8u;
    const bool x_on = g < 3;  // MMA column g = limb g of the activation row
    const uint32_t x_u0 = smem_u32(x_s) + (uint32_t)min(g, 2) * limb_row + (uint32_t)warp * 128u + (uint32_t)c * 8u;
    for (int strip = s0; strip < s1; ++strip, ++li) {
      float yacc[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t xo = 0, S_u = S_u0;
      for (int kt0 = 0; kt0 < KT; kt0 += kStageTiles) {
        mbar_wait_b(&full[slot], par, ctx, 32u);
        if (2 * warp < KT - kt0) {
          const uint4 wa = lds128(stage_u + w_off), wb = lds128(stage_u + w_off + 512u);
          uint2 xb[4];
#pragma unroll
          for (int h = 0; h < 4; ++h) xb[h] = x_on ? lds64u(x_u0 + xo + 32u * h) : make_uint2(0u, 0u);
          constexpr uint32_t kM = 0x0f0f0f0fu;
          const uint32_t wl[4][2] = {{wa.x, wa.y}, {wa.z, wa.w}, {wb.x, wb.y}, {wb.z, wb.w}};
          if constexpr (GE::kNSeg == 1) {
            const uint32_t p0 = lds32(stage_u + sz_off), p1 = lds32(stage_u + sz_off + 32u);
            const float2 S = lds64f(S_u);
            int a0[4] = {(int)kMagicI, (int)kMagicI, (int)kMagicI, (int)kMagicI}, a1[4] = {0, 0, 0, 0};
            imma_16832(a0, wl[0][0] & kM, wl[0][1] & kM, (wl[0][0] >> 4) & kM, (wl[0][1] >> 4) & kM, xb[0].x, xb[0].y);
            imma_16832(a1, wl[1][0] & kM, wl[1][1] & kM, (wl[1][0] >> 4) & kM, (wl[1][1] >> 4) & kM, xb[1].x, xb[1].y);
            imma_16832(a0, wl[2][0] & kM, wl[2][1] & kM, (wl[2][0] >> 4) & kM, (wl[2][1] >> 4) & kM, xb[2].x, xb[2].y);
            imma_16832(a1, wl[3][0] & kM, wl[3][1] & kM, (wl[3][0] >> 4) & kM, (wl[3][1] >> 4) & kM, xb[3].x, xb[3].y);
            const float sc0 = __uint_as_float(p0 << 16), sc1 = __uint_as_float(p1 << 16);
            const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0 & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1 & 0xFFFF0000u));
            const float f0 = __int_as_float(a0[0] + a1[0]) - kMagicF, f1 = __int_as_float(a0[1] + a1[1]) - kMagicF;
            const float f2 = __int_as_float(a0[2] + a1[2]) - kMagicF, f3 = __int_as_float(a0[3] + a1[3]) - kMagicF;
            yacc[0] = fmaf(sc0, f0, fmaf(zz0, S.x, yacc[0]));
            yacc[1] = fmaf(sc0, f1, fmaf(zz0, S.y, yacc[1]));
            yacc[2] = fmaf(sc1, f2, fmaf(zz1, S.x, yacc[2]));
            yacc[3] = fmaf(sc1, f3, fmaf(zz1, S.y, yacc[3]));
          } else {
            constexpr int kHS = 4 / GE::kNSeg;
#pragma unroll
            for (int sg = 0; sg < GE::kNSeg; ++sg) {
              const uint32_t p0 = lds32(stage_u + sz_off + 64u * sg), p1 = lds32(stage_u + sz_off + 64u * sg + 32u);
              const float2 S = lds64f(S_u + 32u * sg);
              int a[4] = {(int)kMagicI, (int)kMagicI, (int)kMagicI, (int)kMagicI};
#pragma unroll
              for (int hh = 0; hh < kHS; ++hh) {
                const int h = sg * kHS + hh;
                imma_16832(a, wl[h][0] & kM, wl[h][1] & kM, (wl[h][0] >> 4) & kM, (wl[h][1] >> 4) & kM, xb[h].x, xb[h].y);
              }
              const float sc0 = __uint_as_float(p0 << 16), sc1 = __uint_as_float(p1 << 16);
              const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0 & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1 & 0xFFFF0000u));
              yacc[0] = fmaf(sc0, __int_as_float(a[0]) - kMagicF, fmaf(zz0, S.x, yacc[0]));
              yacc[1] = fmaf(sc0, __int_as_float(a[1]) - kMagicF, fmaf(zz0, S.y, yacc[1]));
              yacc[2] = fmaf(sc1, __int_as_float(a[2]) - kMagicF, fmaf(zz1, S.x, yacc[2]));
              yacc[3] = fmaf(sc1, __int_as_float(a[3]) - kMagicF, fmaf(zz1, S.y, yacc[3]));
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        xo += kStageTiles * 64u;
        S_u += (kStageTiles * 64 / GE::kSegK) * 32u;
        stage_u += kStageBytes;
        if (++slot == stages) {
          slot = 0;
          par ^= 1u;
          stage_u = ring_u;
        }
      }
      const int rbuf = (int)(li & 1u);
      const uint32_t use = li >> 1;
      float* rb = red + (size_t)rbuf * (kC * 128);
      if (use > 0) mbar_wait_b(&red_free[rbuf], (use - 1) & 1u, ctx, 64u);
      *reinterpret_cast<float4*>(rb + (warp * 32 + lane) * 4) = make_float4(yacc[0], yacc[1], yacc[2], yacc[3]);
      __syncwarp();
      if (lane == 0) mbar_arrive(&red_full[rbuf]);
    }
    if (tid == 0) stamp(p.trace, n_phases, i, 3);
    ++gphase;
  }
}

// ------------------------------------------------------------------------------------------------ host side
struct SmemPlan {
  int stages;
  size_t smem;
  uint32_t off_x, off_S, off_red, off_h, off_misc;
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static bool plan_smem(int G, int64_t Kmax, int dim, int Gq, SmemPlan& pl) {
  const int seg_k = G < 128 ? G : 128;
  const size_t stage_bytes = (size_t)kStageW + (size_t)(kStageTiles * 64 / G) * 64;
  const size_t limbs = (size_t)3 * (size_t)(Kmax + 32);
  const size_t attn = ((size_t)((Gq * kQPad + 31) & ~31) + (size_t)kC * kHG * kPartStride) * 4;
  const size_t x_bytes = align_up(limbs > attn ? limbs : attn, 128);
  const size_t S_bytes = align_up((size_t)(Kmax / seg_k) * 32, 128);
  const size_t red_bytes = (size_t)kRedBufs * kC * 128 * 4;
  const size_t h_bytes = align_up((size_t)dim * 2, 128);
  const size_t misc = align_up((2 * kC + 4) * 4 + (2 * 16 + 2 * kRedBufs) * 8, 128);
  const size_t fixed = x_bytes + S_bytes + red_bytes + h_bytes + misc;
  const size_t cap = 227 * 1024;
  if (fixed + 3 * stage_bytes > cap) return false;
  int st = (int)((cap - fixed) / stage_bytes);
  if (st > 16) st = 16;
  pl.stages = st;
  size_t off = (size_t)st * stage_bytes;
  pl.off_x = (uint32_t)off; off += x_bytes;
  pl.off_S = (uint32_t)off; off += S_bytes;
  pl.off_red = (uint32_t)off; off += red_bytes;
  pl.off_h = (uint32_t)off; off += h_bytes;
  pl.off_misc = (uint32_t)off; off += misc;
  pl.smem = off;
  return true;
}

struct Offsets {
  uint32_t qkv, att, d1, mid, d2, aux, total;  // in packets
};

static Offsets xchg_layout(const ao_b200_mega_model& m) {
  Offsets o;
  const uint32_t qkv_n = (uint32_t)(m.n_head + 2 * m.n_kv_head) * 128u;
  auto al = [](uint32_t v) { return (v + 15u) & ~15u; };  // 128-byte granules
  uint32_t off = 0;
  o.qkv = off; off += al(qkv_n / 2);
  o.att = off; off += al((uint32_t)m.n_head * 128u / 2);
  o.d1 = off; off += al((uint32_t)m.world * (uint32_t)m.dim / 2);
  o.mid = off; off += al((uint32_t)m.inter / 2);
  o.d2 = off; off += al((uint32_t)m.world * (uint32_t)m.dim / 2);
  o.aux = off; off += al((uint32_t)m.n_head * (uint32_t)(m.n_cta / (m.n_kv_head > 0 ? m.n_kv_head : 1) + 1) * kPartStride);
  o.total = off;
  return o;
}

static int check_model(const ao_b200_mega_model* m) {
  AO_CHECK_ARG(m && m->layers && m->n_layer > 0 && m->n_cta > 0);
  AO_CHECK_ARG(m->head_dim == 128 && m->n_kv_head > 0 && m->n_head % m->n_kv_head == 0 && m->n_head / m->n_kv_head <= 8);
  AO_CHECK_ARG(m->dim % 128 == 0 && m->inter % 128 == 0 && (m->n_head * 128) % 128 == 0 && m->vocab > 0);
  AO_CHECK_ARG(m->world >= 1 && m->world <= kMaxWorld && m->rank >= 0 && m->rank < m->world);
  AO_CHECK_ARG(m->emb && m->freqs && m->tok && m->pos && m->logits && m->state && m->out_packed && m->out_sz && m->out_norm);
  AO_CHECK_ARG(m->max_seq > 0 && m->n_kv_head <= m->n_cta);
  for (int r = 0; r < m->world; ++r) AO_CHECK_ARG(m->xchg[r] != nullptr);
  if (m->group_size != 32 && m->group_size != 64 && m->group_size != 128 && m->group_size != 256) return AO_B200_ERR_UNSUPPORTED;
  if (m->group_size == 256 && (m->dim % 256 != 0 || m->inter % 256 != 0 || (m->n_head * 128) % 256 != 0)) return AO_B200_ERR_UNSUPPORTED;
  return AO_B200_OK;
}

}  // namespace mega
}  // namespace ao

using namespace ao::mega;

extern "C" int64_t ao_b200_decode_mega_plan_bytes(const ao_b200_mega_model* m) {
  if (check_model(m) != AO_B200_OK) return 0;
  return (int64_t)sizeof(Phase) * (5 * (int64_t)m->n_layer + 1);
}

extern "C" int64_t ao_b200_decode_mega_xchg_bytes(const ao_b200_mega_model* m) {
  if (!m || m->n_kv_head <= 0 || m->n_cta <= 0 || m->world < 1) return 0;
  return (int64_t)xchg_layout(*m).total * 8;
}

extern "C" int64_t ao_b200_decode_mega_state_bytes(void) { return 64; }

extern "C" int ao_b200_decode_mega_setup(const ao_b200_mega_model* m, void* plan_dev, int64_t plan_bytes, void* stream) {
  const int rc = check_model(m);
  if (rc != AO_B200_OK) return rc;
  AO_CHECK_ARG(plan_dev && plan_bytes >= ao_b200_decode_mega_plan_bytes(m));
  const Offsets o = xchg_layout(*m);
  const int n_phases = 5 * m->n_layer + 1;
  std::vector<Phase> ph((size_t)n_phases);
  memset(ph.data(), 0, sizeof(Phase) * ph.size());
  int rot = 0;
  auto gemv = [&](Phase& q, const void* packed, const void* sz, int64_t N, int64_t K) {
    q.type = PH_GEMV;
    q.packed = static_cast<const uint8_t*>(packed);
    q.sz = static_cast<const uint32_t*>(sz);
    q.KT = (int)(K / 64);
    q.Kx = (int)K;
    q.N_out = (int)N;
    q.n_strips = (int)((N + 15) / 16);
    q.rot = rot;
    rot = (rot + q.n_strips % m->n_cta) % m->n_cta;  // the CTAs that got an extra strip rotate from phase to phase
  };
  const int64_t qkv_n = (int64_t)(m->n_head + 2 * m->n_kv_head) * 128, att_n = (int64_t)m->n_head * 128;
  for (int l = 0; l < m->n_layer; ++l) {
    const ao_b200_mega_layer& L = m->layers[l];
    AO_CHECK_ARG(L.wqkv_packed && L.wqkv_sz && L.wo_packed && L.wo_sz && L.w13_packed && L.w13_sz && L.w2_packed && L.w2_sz);
    AO_CHECK_ARG(L.attn_norm && L.ffn_norm && L.k_cache && L.v_cache);
    Phase* q = &ph[(size_t)5 * l];
    gemv(q[0], L.wqkv_packed, L.wqkv_sz, qkv_n, m->dim);
    q[0].in_kind = IN_H; q[0].in_off = l == 0 ? kNone : o.d2; q[0].norm_w = static_cast<const __nv_bfloat16*>(L.attn_norm);
    q[0].out_kind = OUT_PK; q[0].out_off = o.qkv;
    q[1].type = PH_ATTN; q[1].in_off = o.qkv; q[1].out_off = o.att; q[1].aux_off = o.aux;
    q[1].k_cache = static_cast<__nv_bfloat16*>(L.k_cache); q[1].v_cache = static_cast<__nv_bfloat16*>(L.v_cache);
    gemv(q[2], L.wo_packed, L.wo_sz, m->dim, att_n);
    q[2].in_kind = IN_PK; q[2].in_off = o.att; q[2].out_kind = OUT_DELTA; q[2].out_off = o.d1;
    gemv(q[3], L.w13_packed, L.w13_sz, 2 * (int64_t)m->inter, m->dim);
    q[3].N_out = m->inter;  // outputs after SiLU*mul (n_strips stays 2 I / 16)
    q[3].in_kind = IN_H; q[3].in_off = o.d1; q[3].norm_w = static_cast<const __nv_bfloat16*>(L.ffn_norm);
    q[3].out_kind = OUT_SILU; q[3].out_off = o.mid;
    gemv(q[4], L.w2_packed, L.w2_sz, m->dim, m->inter);
    q[4].in_kind = IN_PK; q[4].in_off = o.mid; q[4].out_kind = OUT_DELTA; q[4].out_off = o.d2;
  }
  Phase& f = ph[(size_t)5 * m->n_layer];
  gemv(f, m->out_packed, m->out_sz, m->vocab, m->dim);
  f.in_kind = IN_H; f.in_off = o.d2; f.norm_w = static_cast<const __nv_bfloat16*>(m->out_norm);
  f.out_kind = OUT_PLAIN; f.out_plain = static_cast<__nv_bfloat16*>(m->logits);
  auto s = static_cast<cudaStream_t>(stream);
  if (cudaMemcpyAsync(plan_dev, ph.data(), sizeof(Phase) * ph.size(), cudaMemcpyHostToDevice, s) != cudaSuccess) return AO_B200_ERR_CUDA;
  if (cudaStreamSynchronize(s) != cudaSuccess) return AO_B200_ERR_CUDA;  // ph is a host temporary (set-up time only)
  return AO_B200_OK;
}

extern "C" int ao_b200_decode_mega_step(const ao_b200_mega_model* m, const void* plan_dev, void* stream) {
  const int rc = check_model(m);
  if (rc != AO_B200_OK) return rc;
  AO_CHECK_ARG(plan_dev);
  const int Gq = m->n_head / m->n_kv_head;
  const int64_t att_n = (int64_t)m->n_head * 128;
  int64_t Kmax = m->dim > m->inter ? m->dim : m->inter;
  if (att_n > Kmax) Kmax = att_n;
  SmemPlan pl;
  if (!plan_smem(m->group_size, Kmax, m->dim, Gq, pl)) return AO_B200_ERR_WORKSPACE;
  Params p;
  memset(&p, 0, sizeof p);
  p.phases = static_cast<const Phase*>(plan_dev);
  p.n_phases = 5 * m->n_layer + 1;
  p.n_cta = m->n_cta;
  for (int r = 0; r < m->world; ++r) p.xchg[r] = static_cast<uint64_t*>(m->xchg[r]);
  p.world = m->world;
  p.rank = m->rank;
  p.emb = static_cast<const __nv_bfloat16*>(m->emb);
  p.tok = m->tok;
  p.pos = m->pos;
  p.freqs = static_cast<const __nv_bfloat16*>(m->freqs);
  uint8_t* st = static_cast<uint8_t*>(m->state);
  p.counter = reinterpret_cast<uint64_t*>(st);
  p.step = reinterpret_cast<uint32_t*>(st + 8);
  p.status = reinterpret_cast<uint32_t*>(st + 12);
  p.dim = m->dim; p.H = m->n_head; p.Hkv = m->n_kv_head; p.S = m->max_seq;
  p.attn_bpc = m->attn_blocks_per_cta > 0 ? m->attn_blocks_per_cta : 32;
  p.attn_max_split = m->n_cta / m->n_kv_head;
  p.eps = m->eps;
  p.attn_scale = 1.0f / sqrtf((float)m->head_dim);
  p.stages = pl.stages;
  p.off_x = pl.off_x; p.off_S = pl.off_S; p.off_red = pl.off_red; p.off_h = pl.off_h; p.off_misc = pl.off_misc;
  p.timeout_ns = (unsigned long long)(m->timeout_ms > 0 ? m->timeout_ms : 2000) * 1000000ull;
  p.trace = static_cast<unsigned long long*>(m->trace);
  using KernelFn = void (*)(const Params);
  KernelFn kern = nullptr;
  int gi = 0;
  switch (m->group_size) {
    case 32: kern = decode_mega_kernel<32>; gi = 0; break;
    case 64: kern = decode_mega_kernel<64>; gi = 1; break;
    case 128: kern = decode_mega_kernel<128>; gi = 2; break;
    default: kern = decode_mega_kernel<256>; gi = 3; break;
  }
  static const bool debug = std::getenv("AO_B200_DEBUG") != nullptr;
  static bool attr_done[4] = {};
  if (!attr_done[gi]) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      if (debug) fprintf(stderr, "ao_b200_decode_mega_step: cudaFuncSetAttribute: %s\n", cudaGetErrorString(e));
      cudaGetLastError();
      return AO_B200_ERR_CUDA;
    }
    attr_done[gi] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)m->n_cta);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = pl.smem;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident or the launch fails: the CTAs wait on one another
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p);
  if (le != cudaSuccess) {
    if (debug) fprintf(stderr, "ao_b200_decode_mega_step: launch (grid %d, smem %zu, stages %d): %s\n", m->n_cta, pl.smem, pl.stages, cudaGetErrorString(le));
    cudaGetLastError();
    return AO_B200_ERR_CUDA;
  }
  return ao_launch_status();
}
