// One decode step of the gpt-fast Llama (torchao/_models/llama/model.py:342-486, generate.py:111-117) as ONE persistent,
// cooperative launch: every SM runs one CTA for the whole token, and the weight stream never drains between projections.
//
// Why: with one launch per projection (round 1: 161 launches per token) each launch pays a fixed 3.5 - 5 us (dependency
// resolution, cold instruction cache, activation staging, pipeline fill, drain) around 1.3 - 9.6 us of streaming; the step
// ran at 0.43 of the HBM roofline.  Here a CTA is
//   * 1 producer warp that walks the CTA's share of EVERY byte the token streams, in order — the int4 weights of every
//     projection (16 KB of nibbles + the groups' scale / zero per stage) and, for the CTAs that own a slice of the
//     attention, the K / V cache rows (32 positions per stage) — through ONE bulk-copy ring in shared memory.  It never
//     waits for anything but a free ring slot: while the consumers hand activations over between two projections the
//     ring fills with the next projection's weights, and the K / V rows of a layer are on chip before its q exists;
//   * 16 consumer warps: integer tensor-core inner product of gemv_int4_imma.cu (nibbles -> bytes, mma.sync m16n8k32
//     u8 x s8, activations as three balanced int8 limbs of 23-bit fixed point, exact int32 group sums, affine dequant on
//     the accumulator in fp32), the attention of the token out of the ring, the staging of the next projection's input;
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
//
// Code organisation: the three roles and the consumer's three jobs (input staging, main loop, attention) are separate
// NON-inlined functions — each gets the whole register file for its own loop instead of sharing one allocation across a
// 900-line kernel (the first version kept ring indices in local memory inside the main loop).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"

// build-time A/B switches (benchmarks/ab_build.sh)
#ifndef AO_MEGA_POLL_SLEEP
#define AO_MEGA_POLL_SLEEP 0    // ns between two polls of a packet that has not arrived
#endif
#ifndef AO_MEGA_INLINE_PHASES
#define AO_MEGA_INLINE_PHASES 1  // staging / main loop / attention inlined into the consumers' phase loop (no ABI calls on the hand-over path)
#endif
#if AO_MEGA_INLINE_PHASES
#define AO_PHASE_FN __device__ __forceinline__
#else
#define AO_PHASE_FN __device__ __noinline__
#endif
#ifndef AO_MEGA_BATCH_SWEEP
#define AO_MEGA_BATCH_SWEEP 0   // a sweep has all its packet loads in flight before it looks at the first tag
#endif

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
constexpr int kAttnPos = 32;                 // cache positions per ring stage
constexpr int kKvPitch = 136;                // elements per row of the private K / V cache: 272 B, so that the 8 rows of an
                                             // ldmatrix land in 8 different 16-byte bank groups (a 256 B pitch: 8-way conflicts)
constexpr int kKvTileBytes = kAttnPos * kKvPitch * 2;   // 8704 B of K rows, then 8704 B of V rows
constexpr int kAttnStageBytes = 2 * kKvTileBytes;       // 17408
constexpr int kAttnRound = 8;                // tiles walked concurrently by a CTA (one warp each)
constexpr int kPartStride = 132;             // attention partial: 128 acc + m + l (+ pad)
constexpr int kHdrBytes = 1024;              // shared-memory header in front of the ring: wait context + a copy of Params

enum { PH_GEMV = 0, PH_ATTN = 1 };
enum { IN_H = 0, IN_PK = 1 };
enum { OUT_PK = 0, OUT_DELTA = 1, OUT_SILU = 2, OUT_PLAIN = 3 };
constexpr uint32_t kNone = 0xffffffffu;

// Per-layer pointers (global memory, filled by ao_b200_decode_mega_setup, copied to shared memory at kernel start) and
// per-kind constants (kernel parameters = constant bank): a phase's description costs no global load on the critical path
// (the first version read a 96-byte Phase record per phase from global memory: a DRAM miss per phase per SM, because the
// weight stream evicts it from L2 between tokens).
struct alignas(16) LayerPtrs {      // 112 bytes
  const uint8_t* w[4];              // wqkv, wo, w13, w2: strip-major int4 [Np/16][Kp/64][512 B]
  const uint32_t* sz[4];            // strip-major [Np/16][Kp/G][16] (scale | zero << 16)
  const __nv_bfloat16* norm[2];     // attention_norm, ffn_norm
  __nv_bfloat16* kc;                // [Hkv][S][128]: the model's KVCache (appended to, never read here)
  __nv_bfloat16* vc;
  __nv_bfloat16* kp;                // [Hkv][S_pad][136]: the kernel's own copy with padded rows (read through the ring)
  __nv_bfloat16* vp;
};
enum { K_WQKV = 0, K_ATTN = 1, K_WO = 2, K_W13 = 3, K_W2 = 4, K_OUT = 5, kKinds = 6 };
struct Kind {
  uint32_t in_off;                // packet offset in the exchange region: IN_PK input, IN_H deltas [world (+ 1: their sum)][Kx / 2], ATTN qkv
  uint32_t out_off;               // packet offset of the output (OUT_DELTA: + rank * N_out / 2 in every peer's region)
  uint32_t aux_off;               // ATTN: split partials
  int type, in_kind, out_kind;
  int KT, n_strips, N_out, Kx, rot_inc;
  int gate;                       // arrivals the sweep of this phase does not wait for (their packets are polled directly)
};
// a phase as the roles see it (registers)
struct Phase {
  const uint8_t* packed;
  const uint32_t* sz;
  const __nv_bfloat16* norm_w;
  __nv_bfloat16* out_plain;
  __nv_bfloat16* k_cache;
  __nv_bfloat16* v_cache;
  __nv_bfloat16* k_pad;
  __nv_bfloat16* v_pad;
  uint32_t in_off, out_off, aux_off;
  int type, in_kind, out_kind;
  int KT, n_strips, N_out, Kx, rot_inc, gate;
};

struct Params {
  const LayerPtrs* layers;        // global [n_layer]
  Kind kinds[kKinds];
  const uint8_t* out_w;           // the output projection
  const uint32_t* out_sz;
  const __nv_bfloat16* out_norm;
  __nv_bfloat16* logits;
  int n_layer, n_phases, n_cta;
  uint64_t* xchg[kMaxWorld];      // exchange regions: own at [rank], the peers' as mapped here
  int world, rank;
  const __nv_bfloat16* emb;
  const int64_t* tok;
  int32_t* pos;
  const __nv_bfloat16* freqs;     // [block_size][64][2] (cos, sin)
  uint64_t* counter;              // arrivals, monotonic over the life of the state buffer
  uint32_t* step;                 // launches so far
  uint32_t* status;               // sticky error bits
  int dim, H, Hkv, S, S_pad, attn_spc, attn_max_split;
  float eps, attn_scale;
  int stages;
  uint32_t off_x, off_S, off_red, off_h, off_misc, off_lay;
  unsigned long long timeout_ns;
  unsigned long long* trace;      // optional: [n_cta][n_phases][8] clock64 stamps (benchmarks/mega_trace.py)
  int tp_reduce;                  // world > 1: sum the ranks' partial outputs once per rank (reduce_slice) instead of in every CTA
  int dbg;                        // profiling switches (AO_B200_MEGA_DEBUG): 1 = main loops skip the math (stream only), 2 = count slow polls, 4 = no evict-first hint on the weight stream
};

// wait context at the start of dynamic shared memory: reachable from every (non-inlined) function without arguments
struct SCtx {
  uint32_t abort;                 // this CTA stopped waiting
  uint32_t pad;
  uint32_t* status;
  unsigned long long timeout_ns;
};
__device__ __forceinline__ SCtx* sctx() {
  extern __shared__ __align__(128) uint8_t smem_mega[];
  return reinterpret_cast<SCtx*>(smem_mega);
}
// The kernel parameters, copied to shared memory at kernel start: the non-inlined role functions read them with LDS
// instead of generic loads from the parameter window (a reference to a __grid_constant__ parameter: every p.x was a
// long-scoreboard stall on the critical path between two projections).
static_assert(sizeof(Params) + 64 <= kHdrBytes, "Params must fit the shared-memory header");
__device__ __forceinline__ const Params& sparams() {
  extern __shared__ __align__(128) uint8_t smem_mega[];
  return *reinterpret_cast<const Params*>(smem_mega + 64);
}

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
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// mbarriers by 32-bit shared-window address
__device__ __forceinline__ bool bar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s_u(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar) : "memory");
}
// The weight stream (4 GB per token through a 126 MB L2) is marked evict-first: what the kernel re-reads every layer or
// every token — its own code, the exchange packets, the norm weights, the K / V tiles — then stays resident in L2.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
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

// ------------------------------------------------------------------------------------------------ bounded waits
__device__ __noinline__ void raise_abort(uint32_t code) {
  SCtx* c = sctx();
  *reinterpret_cast<volatile uint32_t*>(&c->abort) = 1u;
  atomicOr(c->status, code);
}
__device__ __forceinline__ bool aborted() { return *reinterpret_cast<volatile uint32_t*>(&sctx()->abort) != 0u; }

__device__ __noinline__ void bar_wait_slow(uint32_t bar, uint32_t parity, uint32_t code) {
  const unsigned long long t0 = globaltimer();
  const unsigned long long lim = sctx()->timeout_ns;
  uint32_t n = 0;
  while (!bar_try(bar, parity)) {
    if (aborted()) return;
    if ((++n & 15u) == 0 && globaltimer() - t0 > lim) {
      raise_abort(code);
      return;
    }
  }
}
__device__ __forceinline__ void bar_wait(uint32_t bar, uint32_t parity, uint32_t code) {
  if (bar_try(bar, parity)) return;
  bar_wait_slow(bar, parity, code);
}
// the arrival counter (one lane per CTA polls; no back-off: 148 pollers of one line are no load for the L2)
__device__ __noinline__ void wait_counter(const uint64_t* ctr, uint64_t target) {
  if (ld_ctr(ctr) >= target) return;
  const unsigned long long t0 = globaltimer();
  const unsigned long long lim = sctx()->timeout_ns;
  uint32_t n = 0;
  for (;;) {
    if (ld_ctr(ctr) >= target) return;
    if ((++n & 63u) == 0) {
      if (aborted()) return;
      if (globaltimer() - t0 > lim) {
        raise_abort(2u);
        return;
      }
      if (*reinterpret_cast<volatile uint32_t*>(sctx()->status) != 0u) {  // somebody else gave up: do not wait for them
        *reinterpret_cast<volatile uint32_t*>(&sctx()->abort) = 1u;
        return;
      }
    }
  }
}
// one packet
__device__ __noinline__ uint32_t poll_pk_slow(const uint64_t* p, uint32_t tag) {
  if (sparams().dbg & 2) atomicAdd(sctx()->status + 1, 1u);  // statistics: how often a poll ran ahead of its packet (state word 4)
  const unsigned long long t0 = globaltimer();
  const unsigned long long lim = sctx()->timeout_ns;
  uint32_t n = 0;
  for (;;) {
    const uint64_t v = ld_pk(p);
    if ((uint32_t)(v >> 32) == tag) return (uint32_t)v;
#if AO_MEGA_POLL_SLEEP > 0
    __nanosleep(AO_MEGA_POLL_SLEEP);
#endif
    if ((++n & 63u) == 0) {
      if (aborted()) return 0u;
      if (globaltimer() - t0 > lim) {
        raise_abort(4u);
        return 0u;
      }
    }
  }
}
__device__ __forceinline__ uint32_t poll_pk(const uint64_t* p, uint32_t tag) {
  const uint64_t v = ld_pk(p);
  if ((uint32_t)(v >> 32) == tag) return (uint32_t)v;
  return poll_pk_slow(p, tag);
}
// 4 consecutive packets (8 bf16 = one staging chunk); p is 32-byte aligned.  The slow path is the normal way in for the
// last arrivals of a phase: the arrival counter only gates the sweep ("almost everybody has published"), the tags decide.
__device__ __noinline__ uint4 poll_chunk_slow(const uint64_t* p, uint32_t tag) {
  if (sparams().dbg & 2) atomicAdd(sctx()->status + 1, 1u);  // statistics (AO_B200_MEGA_DEBUG=2): sweeps that ran ahead
  const unsigned long long t0 = globaltimer();
  const unsigned long long lim = sctx()->timeout_ns;
  uint32_t n = 0;
  for (;;) {
    const ulonglong2 a = ld_pk2(p), b = ld_pk2(p + 2);
    if (((uint32_t)(a.x >> 32) == tag) & ((uint32_t)(a.y >> 32) == tag) & ((uint32_t)(b.x >> 32) == tag) &
        ((uint32_t)(b.y >> 32) == tag))
      return make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)b.x, (uint32_t)b.y);
#if AO_MEGA_POLL_SLEEP > 0
    __nanosleep(AO_MEGA_POLL_SLEEP);
#endif
    if ((++n & 63u) == 0) {
      if (aborted()) break;
      if (globaltimer() - t0 > lim) {
        raise_abort(4u);
        break;
      }
    }
  }
  return make_uint4(0u, 0u, 0u, 0u);
}
__device__ __forceinline__ uint4 poll_chunk(const uint64_t* p, uint32_t tag) {
  const ulonglong2 a = ld_pk2(p), b = ld_pk2(p + 2);
  const bool ok = ((uint32_t)(a.x >> 32) == tag) & ((uint32_t)(a.y >> 32) == tag) & ((uint32_t)(b.x >> 32) == tag) &
                  ((uint32_t)(b.y >> 32) == tag);
  if (ok) return make_uint4((uint32_t)a.x, (uint32_t)a.y, (uint32_t)b.x, (uint32_t)b.y);
  return poll_chunk_slow(p, tag);
}
// the same in two halves, so that a sweep can have all its loads in flight before it looks at the first tag
struct ChunkLd {
  ulonglong2 a, b;
};
__device__ __forceinline__ ChunkLd chunk_load(const uint64_t* p) {
  ChunkLd c;
#if !AO_MEGA_BATCH_SWEEP
  c.a = c.b = make_ulonglong2(0ull, 0ull);
  return c;
#endif
  c.a = ld_pk2(p);
  c.b = ld_pk2(p + 2);
  return c;
}
__device__ __forceinline__ uint4 chunk_take(const ChunkLd& c, const uint64_t* p, uint32_t tag) {
#if !AO_MEGA_BATCH_SWEEP
  return poll_chunk(p, tag);
#endif
  const bool ok = ((uint32_t)(c.a.x >> 32) == tag) & ((uint32_t)(c.a.y >> 32) == tag) & ((uint32_t)(c.b.x >> 32) == tag) &
                  ((uint32_t)(c.b.y >> 32) == tag);
  if (ok) return make_uint4((uint32_t)c.a.x, (uint32_t)c.a.y, (uint32_t)c.b.x, (uint32_t)c.b.y);
  return poll_chunk_slow(p, tag);
}
// Staging chunk of (round c, lane, warp): the chunks a warp stages are exactly the ones its main loop reads (warp w owns
// k in [2048 s + 128 w, + 128) of every stage s: chunks 256 s + 16 w + 0..15), lanes 0-15 take stage 2 c, lanes 16-31
// stage 2 c + 1 — so the limbs need no CTA barrier before the main loop, only a __syncwarp.
__device__ __forceinline__ int stage_chunk(int c, int lane, int warp) { return 256 * (2 * c + (lane >> 4)) + 16 * warp + (lane & 15); }

__device__ __forceinline__ void strip_range(int b, int n_cta, int n_strips, int rot, int& s0, int& s1) {
  int bb = b + rot;
  if (bb >= n_cta) bb -= n_cta;
  const int base = n_strips / n_cta, rem = n_strips - base * n_cta;
  s0 = bb * base + min(bb, rem);
  s1 = s0 + base + (bb < rem ? 1 : 0);
}

// which slice of a layer's attention this CTA owns: (kv head, range of 32-position tiles).  The producer (which streams
// the K / V rows) and the consumers (which eat them) evaluate the same function.
struct AttnPart {
  int active, kvh, split, nsplit, t0, t1, n_t, len;
};
__device__ __forceinline__ AttnPart attn_partition(int pos, int S, int Hkv, int spc, int max_split, int b) {
  AttnPart a;
  a.len = min(pos + 1, S);                              // positions 0 .. pos (the new token included)
  a.n_t = (a.len + kAttnPos - 1) / kAttnPos;
  int ns = min(max_split, (a.n_t + spc - 1) / spc);
  const int tps = (a.n_t + ns - 1) / ns;                // tiles per split
  ns = (a.n_t + tps - 1) / tps;
  a.nsplit = ns;
  a.active = b < Hkv * ns;
  a.kvh = b % Hkv;
  a.split = b / Hkv;
  a.t0 = a.split * tps;
  a.t1 = min(a.t0 + tps, a.n_t);
  return a;
}

template <int G>
struct Geom {
  static constexpr int kSegK = G < 128 ? G : 128;
  static constexpr int kNSeg = 128 / kSegK;
  static constexpr int kSegLanes = kSegK / 8;
  static constexpr int kStageSZ = (kStageTiles * 64 / G) * 64;
  static constexpr int kStageBytes = kStageW + kStageSZ;                                         // a weight stage
  static constexpr int kSlotBytes = kStageBytes > kAttnStageBytes ? kStageBytes : kAttnStageBytes;  // a ring slot
};

// shared-memory addresses of the barriers, derived from the first one
struct Bars {
  uint32_t full, empty, red_full, red_free;
};
__device__ __forceinline__ Bars make_bars(uint32_t first, int stages) {
  Bars b;
  b.full = first;
  b.empty = first + 8u * stages;
  b.red_full = b.empty + 8u * stages;
  b.red_free = b.red_full + 8u * kRedBufs;
  return b;
}

// phase i -> its description (constant bank + the layer's pointers in shared memory)
__device__ __forceinline__ Phase get_phase(const Params& p, int i) {
  extern __shared__ __align__(128) uint8_t smem_mega[];
  const LayerPtrs* lay_s = reinterpret_cast<const LayerPtrs*>(smem_mega + p.off_lay);
  const int layer = i / 5;
  const int kind = layer == p.n_layer ? (int)K_OUT : i - 5 * layer;
  const Kind& k = p.kinds[kind];
  Phase ph;
  ph.in_off = (i == 0) ? kNone : k.in_off;   // the first projection of the token has no delta to add
  ph.out_off = k.out_off; ph.aux_off = k.aux_off;
  ph.type = k.type; ph.in_kind = k.in_kind; ph.out_kind = k.out_kind;
  ph.KT = k.KT; ph.n_strips = k.n_strips; ph.N_out = k.N_out; ph.Kx = k.Kx; ph.rot_inc = k.rot_inc; ph.gate = k.gate;
  ph.packed = nullptr; ph.sz = nullptr; ph.norm_w = nullptr; ph.out_plain = nullptr; ph.k_cache = nullptr; ph.v_cache = nullptr;
  ph.k_pad = nullptr; ph.v_pad = nullptr;
  if (kind == K_OUT) {
    ph.packed = p.out_w; ph.sz = p.out_sz; ph.norm_w = p.out_norm; ph.out_plain = p.logits;
  } else if (kind == K_ATTN) {
    ph.k_cache = lay_s[layer].kc; ph.v_cache = lay_s[layer].vc; ph.k_pad = lay_s[layer].kp; ph.v_pad = lay_s[layer].vp;
  } else {
    const int wi = kind == K_WQKV ? 0 : kind - 1;
    ph.packed = lay_s[layer].w[wi]; ph.sz = lay_s[layer].sz[wi];
    if (kind == K_WQKV) ph.norm_w = lay_s[layer].norm[0];
    if (kind == K_W13) ph.norm_w = lay_s[layer].norm[1];
  }
  return ph;
}
__device__ __forceinline__ int next_rot(int rot, int inc, int n_cta) {
  rot += inc;
  return rot >= n_cta ? rot - n_cta : rot;
}

// ------------------------------------------------------------------------------------------------ producer warp
// The CTA's share of every byte of the token, in phase order, through the ring.
template <int G>
__device__ __noinline__ void producer_loop(int pos, uint32_t ring_u, uint32_t bar0) {
  constexpr int kStageBytes = Geom<G>::kSlotBytes;
  const Params& p = sparams();
  const int b = blockIdx.x, n_cta = p.n_cta, stages = p.stages, n_phases = p.n_phases;
  const Bars bars = make_bars(bar0, stages);
  int slot = 0, rot = 0;
  uint32_t par = 1;
  const uint64_t pol = l2_evict_first_policy();
  const bool stream_hint = (p.dbg & 4) == 0;  // AO_B200_MEGA_DEBUG=4: weights with the default L2 policy (A/B)
  for (int i = 0; i < n_phases; ++i) {
    const Phase phv = get_phase(p, i);
    const Phase* ph = &phv;
    if (ph->type == PH_GEMV) {
      const int KT = ph->KT;
      int s0, s1;
      strip_range(b, n_cta, ph->n_strips, rot, s0, s1);
      rot = next_rot(rot, ph->rot_inc, n_cta);
      const int n_groups = KT * 64 / G;
      const uint8_t* wrow = ph->packed + (int64_t)s0 * KT * 512;
      const uint32_t* srow = ph->sz + (int64_t)s0 * n_groups * 16;
      for (int strip = s0; strip < s1; ++strip) {
        for (int kt0 = 0; kt0 < KT; kt0 += kStageTiles) {
          const int cnt = min(kStageTiles, KT - kt0);
          const uint32_t dst = ring_u + (uint32_t)slot * kStageBytes;
          bar_wait(bars.empty + 8u * slot, par, 8u);
          const uint32_t sz_bytes = (uint32_t)(cnt * 64 / G) * 64u;
          bar_expect_tx(bars.full + 8u * slot, (uint32_t)(cnt * 512) + sz_bytes);
          if (stream_hint) {
            bulk_g2s_hint(dst, wrow + (int64_t)kt0 * 512, (uint32_t)(cnt * 512), bars.full + 8u * slot, pol);
            bulk_g2s_hint(dst + kStageW, srow + (int64_t)(kt0 * 64 / G) * 16, sz_bytes, bars.full + 8u * slot, pol);
          } else {
            bulk_g2s_u(dst, wrow + (int64_t)kt0 * 512, (uint32_t)(cnt * 512), bars.full + 8u * slot);
            bulk_g2s_u(dst + kStageW, srow + (int64_t)(kt0 * 64 / G) * 16, sz_bytes, bars.full + 8u * slot);
          }
          if (++slot == stages) {
            slot = 0;
            par ^= 1u;
          }
        }
        wrow += (int64_t)KT * 512;
        srow += (int64_t)n_groups * 16;
      }
    } else {
      // K / V rows of this CTA's attention slice out of the padded private cache: rows [32 t, 32 t + 32) of the kv head,
      // K into the first 8704 B of the slot, V into the second.  Rows written by EARLIER launches only; the new token's
      // row is patched in by the consumers (attention_phase).
      const AttnPart a = attn_partition(pos, p.S, p.Hkv, p.attn_spc, p.attn_max_split, b);
      if (a.active) {
        const __nv_bfloat16* kc = ph->k_pad + (int64_t)a.kvh * p.S_pad * kKvPitch;
        const __nv_bfloat16* vc = ph->v_pad + (int64_t)a.kvh * p.S_pad * kKvPitch;
        for (int t = a.t0; t < a.t1; ++t) {
          const int64_t e0 = (int64_t)t * kAttnPos * kKvPitch;
          const uint32_t dst = ring_u + (uint32_t)slot * kStageBytes;
          bar_wait(bars.empty + 8u * slot, par, 8u);
          bar_expect_tx(bars.full + 8u * slot, (uint32_t)kAttnStageBytes);
          bulk_g2s_u(dst, kc + e0, (uint32_t)kKvTileBytes, bars.full + 8u * slot);
          bulk_g2s_u(dst + (uint32_t)kKvTileBytes, vc + e0, (uint32_t)kKvTileBytes, bars.full + 8u * slot);
          if (++slot == stages) {
            slot = 0;
            par ^= 1u;
          }
        }
      }
    }
    stamp(p.trace, n_phases, i, 5);
  }
}

// ------------------------------------------------------------------------------------------------ epilogue warp
// Per strip: sum the 16 consumer warps' partials, recombine the limbs, apply the row scale, publish.
__device__ __noinline__ void epilogue_loop(int pos, uint32_t step, uint32_t red_u, uint32_t bar0, const float* row_scale) {
  const Params& p = sparams();
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  const int b = blockIdx.x, n_cta = p.n_cta, n_phases = p.n_phases;
  const Bars bars = make_bars(bar0, p.stages);
  const uint32_t tag0 = step * (uint32_t)n_phases;
  uint64_t* xq = p.xchg[p.rank];
  uint32_t li = 0;   // strips handed over so far (all phases)
  int gphase = 0;    // GEMV phases so far
  int rot = 0;
  for (int i = 0; i < n_phases; ++i) {
    const Phase phv = get_phase(p, i);
    const Phase* ph = &phv;
    if (ph->type != PH_GEMV) continue;
    int s0, s1;
    strip_range(b, n_cta, ph->n_strips, rot, s0, s1);
    rot = next_rot(rot, ph->rot_inc, n_cta);
    const int out_kind = ph->out_kind;
    const int N_out = ph->N_out;
    const uint32_t tag = tag0 + (uint32_t)i + 1u;
    const uint32_t out_off = ph->out_off;
    __nv_bfloat16* out_plain = ph->out_plain;
    for (int strip = s0; strip < s1; ++strip, ++li) {
      const uint32_t rbuf = li & 1u;
      const uint32_t use = li >> 1;
      const uint32_t rb = red_u + rbuf * (kC * 128 * 4) + (uint32_t)lane * 16u;
      bar_wait(bars.red_full + 8u * rbuf, use & 1u, 16u);
      uint64_t vlo = 0ull, vhi = 0ull;
#pragma unroll
      for (int w = 0; w < kC; ++w) {
        const uint4 t = lds128(rb + (uint32_t)w * 512u);
        const uint64_t tx = (uint64_t)t.x | ((uint64_t)t.y << 32), ty = (uint64_t)t.z | ((uint64_t)t.w << 32);
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vlo) : "l"(tx));
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(vhi) : "l"(ty));
      }
      const float4 v = make_float4(__uint_as_float((uint32_t)vlo), __uint_as_float((uint32_t)(vlo >> 32)),
                                   __uint_as_float((uint32_t)vhi), __uint_as_float((uint32_t)(vhi >> 32)));
      __syncwarp();
      if (lane == 0) bar_arrive(bars.red_free + 8u * rbuf);
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
        const uint32_t d0 = o0 | (o0n << 16), d1 = o1 | (o1n << 16);  // valid on lanes c == 0, g even (lanes 0, 8, 16, 24)
        if (out_kind == OUT_PK || p.world == 1) {
          const int n0 = 16 * strip + g;
          if (c == 0 && (g & 1) == 0) {
            if (n0 < N_out) st_pk(xq + out_off + (n0 >> 1), d0, tag);
            if (n0 + 8 < N_out) st_pk(xq + out_off + ((n0 + 8) >> 1), d1, tag);
          }
        } else {
          // OUT_DELTA, world > 1: this rank's slot in every rank's region.  The 4 packet pairs x 8 ranks are spread over
          // the 32 lanes (one destination rank per lane group) instead of 16 stores in a row from 4 lanes: the stores to
          // the peers leave in one issue slot each, and the arrival count is not held back by them.
          const int src = (lane & 3) * 8, rd = lane >> 2;  // source lane (row pair g = 2 (lane & 3)), destination rank offset
          const uint32_t e0 = __shfl_sync(0xffffffffu, d0, src), e1 = __shfl_sync(0xffffffffu, d1, src);
          const int n0 = 16 * strip + 2 * (lane & 3);
          if (rd < p.world) {
            int r = p.rank + rd;  // own region first, then staggered over the peers
            if (r >= p.world) r -= p.world;
            uint64_t* dst = p.xchg[r] + out_off + (uint32_t)p.rank * (uint32_t)(N_out >> 1);
            if (n0 < N_out) st_pk(dst + (n0 >> 1), e0, tag);
            if (n0 + 8 < N_out) st_pk(dst + ((n0 + 8) >> 1), e1, tag);
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
}

// ------------------------------------------------------------------------------------------------ consumers: input staging
// This code runs once per projection between two streaming loops, on 4 warps per scheduler: its cost is its instruction
// count (and its footprint in the instruction cache), not memory traffic — the first version's 2300 - 3000-instruction
// staging functions took 3 us per projection.  Hence: one instantiation per (kind, chunks per thread), straight-line,
// packed bf16x2 arithmetic wherever it is exact, no division, no generic fall-back.
__device__ __forceinline__ uint32_t bf2_mul(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t bf2_add(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t bf2_absmax_nan(uint32_t m, uint32_t v) {  // max(m, |v|) per half, NaN wins
  uint32_t d;
  const uint32_t av = v & 0x7fff7fffu;
  asm("max.NaN.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(m), "r"(av));
  return d;
}
__device__ __forceinline__ uint32_t cvt_bf2(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// one chunk (8 activations as 4 bf16x2 words, k order) -> three int8 limb rows in shared memory + the segment's limb
// sums.  Executed by whole warps in uniform control flow (the segment sums are shuffles); `valid` = the chunk exists.
// Balanced limbs without carries: X + 0x808080 = sum (a_i + 128) 256^i, so its bytes xor 0x80 ARE the two's-complement
// limbs a_i; x * 2^e + 1.5 * 2^23 puts X + 2^22 into the float's mantissa (see gemv_int4_imma.cu).
template <int G>
__device__ __forceinline__ void emit_limbs(uint32_t x_u, uint32_t limb_row, uint32_t S_u, int ci, bool valid, const uint4 y,
                                           float up) {
  using GE = Geom<G>;
  int s01 = 0, s2 = 0;  // limb sums of the chunk: (limb 1 << 16) + limb 0 (each |.| <= 1024), limb 2
  if (valid) {
    const uint32_t w[4] = {y.x, y.y, y.z, y.w};
    uint32_t v[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      v[2 * e] = __float_as_uint(fmaf(__uint_as_float(w[e] << 16), up, kMagicF)) + 0x00408080u;
      v[2 * e + 1] = __float_as_uint(fmaf(__uint_as_float(w[e] & 0xFFFF0000u), up, kMagicF)) + 0x00408080u;
    }
    uint32_t lw[3][2];
#pragma unroll
    for (int l = 0; l < 3; ++l) {
      // limb l lives in byte (2 - l) of every v; gathered in the order of the weight bytes after the nibble split:
      // (k0, k0+4, k0+1, k0+5 | k0+2, k0+6, k0+3, k0+7)
      const uint32_t sel = l == 0 ? 0x0062u : (l == 1 ? 0x0051u : 0x0040u);
      const uint32_t p04 = __byte_perm(v[0], v[4], sel), p15 = __byte_perm(v[1], v[5], sel);
      const uint32_t p26 = __byte_perm(v[2], v[6], sel), p37 = __byte_perm(v[3], v[7], sel);
      lw[l][0] = __byte_perm(p04, p15, 0x5410) ^ 0x80808080u;
      lw[l][1] = __byte_perm(p26, p37, 0x5410) ^ 0x80808080u;
    }
    const int sl0 = __dp4a((int)lw[0][0], 0x01010101, __dp4a((int)lw[0][1], 0x01010101, 0));
    const int sl1 = __dp4a((int)lw[1][0], 0x01010101, __dp4a((int)lw[1][1], 0x01010101, 0));
    s2 = __dp4a((int)lw[2][0], 0x01010101, __dp4a((int)lw[2][1], 0x01010101, 0));
    s01 = sl1 * 65536 + sl0;
    const uint32_t a = x_u + (uint32_t)ci * 8u;
    asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(lw[0][0]), "r"(lw[0][1]) : "memory");
    asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a + limb_row), "r"(lw[1][0]), "r"(lw[1][1]) : "memory");
    asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a + 2u * limb_row), "r"(lw[2][0]), "r"(lw[2][1]) : "memory");
  }
  // a segment = kSegLanes consecutive chunks = consecutive lanes; two sums travel in one register (|sum| <= 16384 each)
#pragma unroll
  for (int o = GE::kSegLanes / 2; o; o >>= 1) {
    s01 += __shfl_xor_sync(0xffffffffu, s01, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (valid && (ci & (GE::kSegLanes - 1)) == 0) {
    const int sl0 = (int)(short)(s01 & 0xffff);
    const int sl1 = (s01 - sl0) >> 16;
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(S_u + (uint32_t)(ci / GE::kSegLanes) * 32u), "f"((float)sl0),
                 "f"((float)sl1), "f"((float)s2), "f"(0.f) : "memory");
  }
}

// second level of the statistics reduction: every warp reduces the 16 warps' partials itself; returns the power of two
// `up` that brings the bound into [2^21, 2^22) and (warp 0) publishes the row scale for the epilogue warp.
__device__ __forceinline__ float finish_stats(const float* part, int lane, int warp, bool with_norm, float inv_k, float eps,
                                              float* row_scale_slot, float& rnorm) {
  float t = 0.f;
  uint32_t am = 0u;
  if (lane < kC) {
    t = part[lane * 2];
    am = __float_as_uint(part[lane * 2 + 1]);
  }
#pragma unroll
  for (int o = 8; o; o >>= 1) {
    t += __shfl_xor_sync(0xffffffffu, t, o);
    am = max(am, __shfl_xor_sync(0xffffffffu, am, o));
  }
  t = __shfl_sync(0xffffffffu, t, 0);
  am = __shfl_sync(0xffffffffu, am, 0);
  float bound = __uint_as_float(am);
  bool bad = am >= 0x7f800000u;  // an Inf / NaN activation
  rnorm = 1.f;
  if (with_norm) {
    rnorm = rsqrtf(fmaf(t, inv_k, eps));
    // |bf16(bf16(x r) w)| <= |bf16(x w)| r (1 + 2^-8)^3: a bound within 2x of the true maximum costs at most one of the 23 bits
    bound = bound * rnorm * 1.0158f;
    bad = bad || !(bound < 3.0e38f);
  }
  // 2^e brings the bound into [2^21, 2^22): |round(x 2^e)| <= 2^22 - 2^14 for every bf16 x below the bound
  int e = 0;
  if (bound > 0.f) e = 21 - ((int)((__float_as_uint(bound) >> 23) & 0xFF) - 127);
  e = max(-106, min(100, e));
  // every warp publishes the (identical) row scale itself: the epilogue warp reads it after ANY warp's first hand-over
  // of the phase, and only a warp's own stores are ordered before its own mbarrier arrive
  if (lane == 0) *row_scale_slot = bad ? __uint_as_float(0x7fc00000u) : __uint_as_float((uint32_t)(127 - e) << 23);
  return __uint_as_float((uint32_t)(e + 127) << 23);
}

// world > 1: every rank sums the partial outputs of all ranks ONCE, its CTAs taking a slice of the vector each, and
// publishes the bf16 sums in a local slot behind the ranks' slots; the sweep then reads one slot, as at world == 1.  (The
// first version had every CTA read all `world` slots: 8 x 32 KB x 148 CTAs out of L2 per hand-over at world 8 — 6 us.)
// One thread per (chunk, rank); butterfly sum over the ranks (every lane ends with the same bits; every rank computes
// the same tree over the same packets, so the ranks stay bit-identical); bf16(sum) is what all_reduce(bf16) gives.
__device__ __noinline__ void reduce_slice(const uint64_t* in_pk, int n_chunks, uint32_t tag_in) {
  const Params& p = sparams();
  const int world = p.world, tid = threadIdx.x;
  const int sh = 31 - __clz(world);
  const int cpc = (n_chunks + p.n_cta - 1) / p.n_cta;
  const int r = tid & (world - 1);
  // (chunk, rank) pairs of this CTA's slice, kCT >> sh chunks per pass (one pass on a full grid: 4 - 7 chunks per CTA);
  // the trip count is warp-uniform (the butterfly needs whole warps)
  for (int jb = (tid & ~31) >> sh; jb < cpc; jb += kCT >> sh) {
    const int j = jb + ((tid & 31) >> sh);
    const int ci = (int)blockIdx.x * cpc + j;
    const bool on = j < cpc && ci < n_chunks;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = 0.f;
    if (on) unpack8(poll_chunk(in_pk + (size_t)r * (size_t)(n_chunks * 4) + ci * 4, tag_in), f);
    for (int o = 1; o < world; o <<= 1) {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] += __shfl_xor_sync(0xffffffffu, f[e], o);
    }
    if (on && r == 0) {
      uint64_t* dst = const_cast<uint64_t*>(in_pk) + (size_t)world * (size_t)(n_chunks * 4) + ci * 4;
      const uint64_t t = (uint64_t)tag_in << 32;
      const uint64_t a = cvt_bf2(f[0], f[1]) | t, b = cvt_bf2(f[2], f[3]) | t, c = cvt_bf2(f[4], f[5]) | t, d = cvt_bf2(f[6], f[7]) | t;
      asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1,%2};" ::"l"(dst), "l"(a), "l"(b) : "memory");
      asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1,%2};" ::"l"(dst + 2), "l"(c), "l"(d) : "memory");
    }
  }
}

// IN_H: h += the published delta (world > 1: the sum slot, see reduce_slice; h stays in shared memory; one packed bf16x2
// add == bf16(h + d) exactly), then RMSNorm.forward (model.py:498-503): bf16(bf16(h * rsqrt(mean(h^2) + eps)) * w).
// The norm weights are requested BEFORE the wait (a DRAM miss, once per token).  NCH = chunks of 8 per thread (K <= 4096 NCH).
template <int G, int NCH>
AO_PHASE_FN void stage_h(int n_chunks, uint32_t in_off, const __nv_bfloat16* norm_w_p, float inv_k,
                                     int i, uint32_t tag_in, uint64_t cnt_target, float* row_scale_slot, uint32_t x_u,
                                     uint32_t S_u, uint4* h_s, float* part) {
  const Params& p = sparams();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t limb_row = (uint32_t)n_chunks * 8u + 32u;
  const uint64_t* in_pk = in_off == kNone ? nullptr : p.xchg[p.rank] + in_off;
  const uint4* norm_w = reinterpret_cast<const uint4*>(norm_w_p);
  uint4 wn[NCH], hv[NCH];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int ci = stage_chunk(c, lane, warp);
    wn[c] = ci < n_chunks ? __ldg(norm_w + ci) : make_uint4(0u, 0u, 0u, 0u);
    hv[c] = make_uint4(0u, 0u, 0u, 0u);
  }
  if (tid == 0) stamp(p.trace, p.n_phases, i, 0);
  int n_slot = p.world;
  if (in_pk != nullptr && p.tp_reduce) {  // the first warps out of their main loop sum this CTA's slice of the partials
    reduce_slice(in_pk, n_chunks, tag_in);
    in_pk += (size_t)p.world * (size_t)(n_chunks * 4);  // the sum slot
    n_slot = 1;
  }
  if (tid == 0 && i > 0) wait_counter(p.counter, cnt_target);
  named_bar_sync(1, kCT);  // B1: phase i - 1 is published; every consumer warp of this CTA has left its main loop
  if (tid == 0) stamp(p.trace, p.n_phases, i, 1);
  float ss = 0.f;
  uint32_t am2 = 0u;  // packed max |bf16(h w)|
  ChunkLd l0[NCH], l1[NCH];  // slot 0 (world 1; or the sum slot) and slot 1 (world 2 without reduce_slice): loads in flight first
  if (in_pk != nullptr) {
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int ci = stage_chunk(c, lane, warp);
      if (ci < n_chunks) {
        l0[c] = chunk_load(in_pk + ci * 4);
        if (n_slot > 1) l1[c] = chunk_load(in_pk + (size_t)(n_chunks * 4) + ci * 4);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int ci = stage_chunk(c, lane, warp);
    if (ci < n_chunks) {
      uint4 h = h_s[ci];
      if (in_pk != nullptr) {
        const uint4 d = chunk_take(l0[c], in_pk + ci * 4, tag_in);
        if (n_slot == 1) {
          h = make_uint4(bf2_add(h.x, d.x), bf2_add(h.y, d.y), bf2_add(h.z, d.z), bf2_add(h.w, d.w));
        } else {  // world 2: summing both slots here is cheaper than the extra hand-over of reduce_slice (measured)
          float acc[8], hf[8], df[8];
          unpack8(d, acc);
          unpack8(chunk_take(l1[c], in_pk + (size_t)(n_chunks * 4) + ci * 4, tag_in), df);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] += df[e];
          for (int r = 2; r < n_slot; ++r) {  // (AO_B200_MEGA_TP_REDUCE=0 at world > 2)
            unpack8(poll_chunk(in_pk + (size_t)r * (n_chunks * 4) + ci * 4, tag_in), df);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += df[e];
          }
          unpack8(h, hf);
          h = make_uint4(cvt_bf2(hf[0] + acc[0], hf[1] + acc[1]), cvt_bf2(hf[2] + acc[2], hf[3] + acc[3]),
                         cvt_bf2(hf[4] + acc[4], hf[5] + acc[5]), cvt_bf2(hf[6] + acc[6], hf[7] + acc[7]));
        }
        h_s[ci] = h;
      }
      hv[c] = h;
      float f[8];
      unpack8(h, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
      am2 = bf2_absmax_nan(am2, bf2_mul(h.x, wn[c].x));
      am2 = bf2_absmax_nan(am2, bf2_mul(h.y, wn[c].y));
      am2 = bf2_absmax_nan(am2, bf2_mul(h.z, wn[c].z));
      am2 = bf2_absmax_nan(am2, bf2_mul(h.w, wn[c].w));
    }
  }
  if (tid == 0) stamp(p.trace, p.n_phases, i, 7);
  uint32_t am = max(am2 << 16, am2 & 0xffff0000u);  // as fp32 bits (a NaN has the largest bits)
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
    am = max(am, __shfl_xor_sync(0xffffffffu, am, o));
  }
  if (lane == 0) {
    part[warp * 2] = ss;
    part[warp * 2 + 1] = __uint_as_float(am);
  }
  named_bar_sync(1, kCT);  // B2
  if (tid == 0) stamp(p.trace, p.n_phases, i, 6);
  float rnorm;
  const float up = finish_stats(part, lane, warp, true, inv_k, p.eps, row_scale_slot, rnorm);
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int ci = stage_chunk(c, lane, warp);
    if (NCH == 1 || 512 * c < n_chunks) {  // warp-uniform: stage 2 c exists
      float f[8];
      unpack8(hv[c], f);
      uint4 y;
      y.x = bf2_mul(cvt_bf2(f[0] * rnorm, f[1] * rnorm), wn[c].x);
      y.y = bf2_mul(cvt_bf2(f[2] * rnorm, f[3] * rnorm), wn[c].y);
      y.z = bf2_mul(cvt_bf2(f[4] * rnorm, f[5] * rnorm), wn[c].z);
      y.w = bf2_mul(cvt_bf2(f[6] * rnorm, f[7] * rnorm), wn[c].w);
      emit_limbs<G>(x_u, limb_row, S_u, ci, ci < n_chunks, y, up);
    }
  }
  __syncwarp();  // the warp's own limbs and segment sums are in place (the row scale was published before B2's readers left)
  if (tid == 0) stamp(p.trace, p.n_phases, i, 2);
}

// IN_PK: the packets of the previous phase as they are.  The thread keeps its chunks in registers between the two passes.
template <int G, int NCH>
AO_PHASE_FN void stage_x(int n_chunks, uint32_t in_off, int i, uint32_t tag_in, uint64_t cnt_target,
                                     float* row_scale_slot, uint32_t x_u, uint32_t S_u, float* part) {
  const Params& p = sparams();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t limb_row = (uint32_t)n_chunks * 8u + 32u;
  const uint64_t* in_pk = p.xchg[p.rank] + in_off;
  if (tid == 0) {
    stamp(p.trace, p.n_phases, i, 0);
    wait_counter(p.counter, cnt_target);
  }
  named_bar_sync(1, kCT);  // B1
  if (tid == 0) stamp(p.trace, p.n_phases, i, 1);
  uint4 xv[NCH];
  uint32_t am2 = 0u;
  {
    // every load of the sweep is in flight before the first tag is looked at: one L2 round trip, not NCH of them
    ChunkLd ld[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int ci = stage_chunk(c, lane, warp);
      if (ci < n_chunks) ld[c] = chunk_load(in_pk + ci * 4);
    }
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int ci = stage_chunk(c, lane, warp);
      xv[c] = make_uint4(0u, 0u, 0u, 0u);
      if (ci < n_chunks) {
        xv[c] = chunk_take(ld[c], in_pk + ci * 4, tag_in);
        am2 = bf2_absmax_nan(bf2_absmax_nan(bf2_absmax_nan(bf2_absmax_nan(am2, xv[c].x), xv[c].y), xv[c].z), xv[c].w);
      }
    }
  }
  if (tid == 0) stamp(p.trace, p.n_phases, i, 7);
  uint32_t am = max(am2 << 16, am2 & 0xffff0000u);
#pragma unroll
  for (int o = 16; o; o >>= 1) am = max(am, __shfl_xor_sync(0xffffffffu, am, o));
  if (lane == 0) {
    part[warp * 2] = 0.f;
    part[warp * 2 + 1] = __uint_as_float(am);
  }
  named_bar_sync(1, kCT);  // B2
  if (tid == 0) stamp(p.trace, p.n_phases, i, 6);
  float rnorm;
  const float up = finish_stats(part, lane, warp, false, 0.f, 0.f, row_scale_slot, rnorm);
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int ci = stage_chunk(c, lane, warp);
    if (NCH == 1 || 512 * c < n_chunks) emit_limbs<G>(x_u, limb_row, S_u, ci, ci < n_chunks, xv[c], up);
  }
  __syncwarp();
  if (tid == 0) stamp(p.trace, p.n_phases, i, 2);
}

// ------------------------------------------------------------------------------------------------ consumers: main loop
// one stage of one strip for one warp (its 128 k x 16 rows): yacc += sum over the segments of  s * acc + (z - 8 s) * S
template <int NS>
__device__ __forceinline__ void stage_math(const uint4 wa, const uint4 wb, const uint32_t (&p0)[NS], const uint32_t (&p1)[NS],
                                           const uint2 (&xb)[4], const float2 (&S)[NS], float (&yacc)[4]) {
  constexpr uint32_t kM = 0x0f0f0f0fu;
  // half h: words (row g, row g + 8); low nibbles = k (k0, k0+4, k0+1, k0+5), high nibbles = (k0+2, k0+6, k0+3, k0+7)
  const uint32_t wl[4][2] = {{wa.x, wa.y}, {wa.z, wa.w}, {wb.x, wb.y}, {wb.z, wb.w}};
  if constexpr (NS == 1) {
    int a0[4] = {(int)kMagicI, (int)kMagicI, (int)kMagicI, (int)kMagicI}, a1[4] = {0, 0, 0, 0};
    imma_16832(a0, wl[0][0] & kM, wl[0][1] & kM, (wl[0][0] >> 4) & kM, (wl[0][1] >> 4) & kM, xb[0].x, xb[0].y);
    imma_16832(a1, wl[1][0] & kM, wl[1][1] & kM, (wl[1][0] >> 4) & kM, (wl[1][1] >> 4) & kM, xb[1].x, xb[1].y);
    imma_16832(a0, wl[2][0] & kM, wl[2][1] & kM, (wl[2][0] >> 4) & kM, (wl[2][1] >> 4) & kM, xb[2].x, xb[2].y);
    imma_16832(a1, wl[3][0] & kM, wl[3][1] & kM, (wl[3][0] >> 4) & kM, (wl[3][1] >> 4) & kM, xb[3].x, xb[3].y);
    const float sc0 = __uint_as_float(p0[0] << 16), sc1 = __uint_as_float(p1[0] << 16);
    const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0[0] & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1[0] & 0xFFFF0000u));
    const float f0 = __int_as_float(a0[0] + a1[0]) - kMagicF, f1 = __int_as_float(a0[1] + a1[1]) - kMagicF;
    const float f2 = __int_as_float(a0[2] + a1[2]) - kMagicF, f3 = __int_as_float(a0[3] + a1[3]) - kMagicF;
    yacc[0] = fmaf(sc0, f0, fmaf(zz0, S[0].x, yacc[0]));
    yacc[1] = fmaf(sc0, f1, fmaf(zz0, S[0].y, yacc[1]));
    yacc[2] = fmaf(sc1, f2, fmaf(zz1, S[0].x, yacc[2]));
    yacc[3] = fmaf(sc1, f3, fmaf(zz1, S[0].y, yacc[3]));
  } else {
    constexpr int kHS = 4 / NS;  // 32-k halves per segment
#pragma unroll
    for (int sg = 0; sg < NS; ++sg) {
      int a[4] = {(int)kMagicI, (int)kMagicI, (int)kMagicI, (int)kMagicI};
#pragma unroll
      for (int hh = 0; hh < kHS; ++hh) {
        const int h = sg * kHS + hh;
        imma_16832(a, wl[h][0] & kM, wl[h][1] & kM, (wl[h][0] >> 4) & kM, (wl[h][1] >> 4) & kM, xb[h].x, xb[h].y);
      }
      const float sc0 = __uint_as_float(p0[sg] << 16), sc1 = __uint_as_float(p1[sg] << 16);
      const float zz0 = fmaf(-8.f, sc0, __uint_as_float(p0[sg] & 0xFFFF0000u)), zz1 = fmaf(-8.f, sc1, __uint_as_float(p1[sg] & 0xFFFF0000u));
      yacc[0] = fmaf(sc0, __int_as_float(a[0]) - kMagicF, fmaf(zz0, S[sg].x, yacc[0]));
      yacc[1] = fmaf(sc0, __int_as_float(a[1]) - kMagicF, fmaf(zz0, S[sg].y, yacc[1]));
      yacc[2] = fmaf(sc1, __int_as_float(a[2]) - kMagicF, fmaf(zz1, S[sg].x, yacc[2]));
      yacc[3] = fmaf(sc1, __int_as_float(a[3]) - kMagicF, fmaf(zz1, S[sg].y, yacc[3]));
    }
  }
}

// Warp w owns k-tiles 2 w, 2 w + 1 (128 k) of every stage.  XREG (KT <= 64: K <= 4096, at most two stages per strip): the
// warp's slice of the activation limbs and their segment sums are the same for every strip, so they are read from shared
// memory ONCE per phase and stay in registers.  A stage's slot is released as soon as its words are in registers (the
// loads are unconditional — a warp past the end of a short stage reads stale bytes it never uses).
// Returns the ring state (slot | parity << 8).
template <int G, bool XREG>
AO_PHASE_FN uint32_t gemv_main(uint32_t ring_u, uint32_t bar0, int stages, int KT, int n_my, uint32_t x_u,
                                           uint32_t limb_row, uint32_t S_u, uint32_t red_u, uint32_t state, uint32_t li,
                                           int dbg) {
  using GE = Geom<G>;
  constexpr int kStageBytes = GE::kSlotBytes;
  constexpr int NS = GE::kNSeg;
  constexpr uint32_t kSOff = (kStageTiles * 64 / GE::kSegK) * 32u;  // S_s bytes per stage
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t full0 = bar0, empty0 = bar0 + 8u * stages, red_full0 = empty0 + 8u * stages, red_free0 = red_full0 + 8u * kRedBufs;
  int slot = (int)(state & 0xffu);
  uint32_t par = state >> 8;
  const uint32_t w_off = (uint32_t)warp * 1024u + (uint32_t)lane * 16u;
  const uint32_t sz_off = (uint32_t)kStageW + (uint32_t)(G >= 128 ? warp * 128 / G : warp * NS) * 64u + (uint32_t)g * 4u;
  const uint32_t S_u0 = S_u + (uint32_t)(warp * NS) * 32u + (uint32_t)c * 8u;
  const uint32_t x_u0 = x_u + (uint32_t)min(g, 2) * limb_row + (uint32_t)warp * 128u + (uint32_t)c * 8u;
  const bool x_on = g < 3;  // MMA column g = limb g of the activation row
  const bool skip = (dbg & 1) != 0;
  const uint32_t red_w = red_u + (uint32_t)(warp * 32 + lane) * 16u;

  uint2 xr[2][4];
  float2 Sr[2][NS];
  if constexpr (XREG) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      const bool on = 2 * warp < KT - kStageTiles * s;
#pragma unroll
      for (int h = 0; h < 4; ++h) xr[s][h] = (on && x_on) ? lds64u(x_u0 + 2048u * s + 32u * h) : make_uint2(0u, 0u);
#pragma unroll
      for (int sg = 0; sg < NS; ++sg) Sr[s][sg] = on ? lds64f(S_u0 + kSOff * s + 32u * sg) : make_float2(0.f, 0.f);
    }
  }
  const int n_st = XREG ? 1 : (KT + kStageTiles - 1) / kStageTiles;  // XREG: the (<= 2) stages are unrolled below
  for (int n = 0; n < n_my; ++n) {
    float yacc[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t xo = 0, So = 0;
    int kt0 = 0;
    for (int st = 0; st < n_st; ++st) {
#pragma unroll
      for (int s = 0; s < (XREG ? 2 : 1); ++s) {
        if (!XREG || kStageTiles * s < KT) {
          const uint32_t stage_u = ring_u + (uint32_t)slot * kStageBytes;
          bar_wait(full0 + 8u * slot, par, 32u);
          const uint4 wa = lds128(stage_u + w_off), wb = lds128(stage_u + w_off + 512u);
          uint32_t p0[NS], p1[NS];
#pragma unroll
          for (int sg = 0; sg < NS; ++sg) {
            p0[sg] = lds32(stage_u + sz_off + 64u * sg);
            p1[sg] = lds32(stage_u + sz_off + 64u * sg + 32u);
          }
          uint2 xb[4];
          float2 S[NS];
          if constexpr (XREG) {
#pragma unroll
            for (int h = 0; h < 4; ++h) xb[h] = xr[s][h];
#pragma unroll
            for (int sg = 0; sg < NS; ++sg) S[sg] = Sr[s][sg];
          } else {
#pragma unroll
            for (int h = 0; h < 4; ++h) xb[h] = x_on ? lds64u(x_u0 + xo + 32u * h) : make_uint2(0u, 0u);
#pragma unroll
            for (int sg = 0; sg < NS; ++sg) S[sg] = lds64f(S_u0 + So + 32u * sg);
          }
          __syncwarp();
          if (lane == 0) bar_arrive(empty0 + 8u * slot);
          if (++slot == stages) {
            slot = 0;
            par ^= 1u;
          }
          const bool on = 2 * warp < KT - (XREG ? kStageTiles * s : kt0);
          if (on && !skip) stage_math<NS>(wa, wb, p0, p1, xb, S, yacc);
        }
      }
      xo += kStageTiles * 64u;
      So += kSOff;
      kt0 += kStageTiles;
    }
    // hand the strip's partial sums to the epilogue warp
    const uint32_t rbuf = li & 1u, use = li >> 1;
    if (use > 0) bar_wait(red_free0 + 8u * rbuf, (use - 1) & 1u, 64u);
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(red_w + rbuf * (kC * 128 * 4)), "f"(yacc[0]), "f"(yacc[1]),
                 "f"(yacc[2]), "f"(yacc[3]) : "memory");
    __syncwarp();
    if (lane == 0) bar_arrive(red_full0 + 8u * rbuf);
    ++li;
  }
  return (uint32_t)slot | (par << 8);
}

// ------------------------------------------------------------------------------------------------ consumers: attention
// Attention.forward for the new token (model.py:405-440) on the tensor cores.  One (kv head, range of 32-position tiles)
// per CTA; the GQ q heads of the group are the M rows of the MMAs (padded to 16), so every K / V row is read once for
// all of them.  The K / V rows arrive through the weight ring out of the kernel's padded cache (the producer requested
// them long before q existed); ONE WARP walks one tile:
//     S^T = K Q^T     16 x mma.m16n8k16 (bf16, fp32 accumulate): positions are the M rows (K fragments by ldmatrix.x4 out of
//                     272-byte rows: no bank conflicts), the GQ <= 8 heads the N columns — no accumulator is wasted
//     online softmax  on the accumulator fragments (a lane holds 4 positions x 2 heads: 3 shuffles per reduction)
//     P^T -> B frags  movmatrix (register transpose of the 8x8 blocks), P rounded to bf16 like the reference's softmax output
//     O^T += V^T P^T  16 x mma.m16n8k16: dims are the M rows, V^T fragments by ldmatrix.x4.trans
// about 150 instructions per tile against 4000 for the first version (fp32 FMAs + shuffle reductions: 1 us per 32
// positions of every layer's critical path).  Up to 8 tiles are walked concurrently; the warps' (m, l, O) are merged through
// shared memory, the splits of a long sequence through {fp32 | tag} packets gathered by split 0.
// The new token's k (roped) and v come out of the qkv packets: the CTA that owns the last tile appends them to both
// caches (for the next launches) and patches them into its copy of the tile.
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D(16x8, fp32) += A(16x16, bf16, row) * B(16x8, bf16, col)
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// transpose of an 8x8 bf16 matrix held one row-pair per lane (lane (g, c): row g, columns 2c, 2c + 1)
__device__ __forceinline__ uint32_t movmatrix_t(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

template <int GQ>
AO_PHASE_FN uint32_t attention_phase(__nv_bfloat16* k_cache, __nv_bfloat16* v_cache, __nv_bfloat16* k_pad,
                                                 __nv_bfloat16* v_pad, uint32_t in_off, uint32_t out_off, uint32_t aux_off, int i,
                                                 int pos, uint32_t tag_in, uint64_t cnt_target, uint32_t ring_u,
                                                 uint32_t slot_bytes, uint32_t bar0, uint8_t* x_s, uint32_t state) {
  const Params& p = sparams();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x, n_phases = p.n_phases;
  const int H = p.H, Hkv = p.Hkv, S = p.S;
  const int stages = p.stages;
  const Bars bars = make_bars(bar0, stages);
  const AttnPart a = attn_partition(pos, S, Hkv, p.attn_spc, p.attn_max_split, b);
  uint64_t* xq = p.xchg[p.rank];
  const uint32_t tag_out = tag_in + 1u;
  // scratch in the limb region
  uint32_t* q_s = reinterpret_cast<uint32_t*>(x_s);                // [GQ][68] bf16x2: roped q, rows of 272 B
  uint32_t* kv_new = q_s + 8 * (kKvPitch / 2);                     // [2][64] bf16x2: roped k, v of the new token
  float* part_s = reinterpret_cast<float*>(kv_new + 128);          // [16 warps][GQ][kPartStride]
  constexpr int kItems = (GQ + 2) * 64;                            // (q heads, k, v) x 64 pairs
  constexpr int kIt = (kItems + kCT - 1) / kCT;
  // cos / sin of the position: independent of everybody, requested before the wait
  uint32_t cs[kIt];
#pragma unroll
  for (int it = 0; it < kIt; ++it) {
    const int idx = it * kCT + tid;
    cs[it] = 0u;
    if (a.active && idx < (GQ + 1) * 64)
      cs[it] = __ldg(reinterpret_cast<const uint32_t*>(p.freqs) + (int64_t)min(pos, S - 1) * 64 + (idx & 63));
  }
  if (tid == 0) {
    stamp(p.trace, n_phases, i, 0);
    wait_counter(p.counter, cnt_target);
  }
  named_bar_sync(1, kCT);  // B1: qkv is published; every consumer warp has left the wqkv main loop
  if (tid == 0) stamp(p.trace, n_phases, i, 1);
  if (!a.active) {
    if (tid == 0) {
      red_ctr(p.counter);
      stamp(p.trace, n_phases, i, 4);
    }
    return state;
  }
  const bool owner = a.split == a.nsplit - 1 && pos < S;
  {
    const uint64_t* qkv = xq + in_off;
#pragma unroll
    for (int it = 0; it < kIt; ++it) {
      const int idx = it * kCT + tid;
      if (idx < kItems) {
        const int which = idx >> 6, pr = idx & 63;
        const int elem = which < GQ ? (a.kvh * GQ + which) * 128 : (which == GQ ? (H + a.kvh) * 128 : (H + Hkv + a.kvh) * 128);
        const uint32_t raw = poll_pk(qkv + (elem >> 1) + pr, tag_in);
        uint32_t o = raw;
        if (which <= GQ) {  // RoPE, model.py:547-559: fp32 math on bf16 cos / sin, rounded to bf16
          const float x0 = __uint_as_float(raw << 16), x1 = __uint_as_float(raw & 0xFFFF0000u);
          const float cc = __uint_as_float(cs[it] << 16), sn = __uint_as_float(cs[it] & 0xFFFF0000u);
          o = cvt_bf2(x0 * cc - x1 * sn, x1 * cc + x0 * sn);
        }
        if (which < GQ) {
          q_s[which * (kKvPitch / 2) + pr] = o;
        } else {
          kv_new[(which - GQ) * 64 + pr] = o;
          if (owner) {  // KVCache.update (model.py:175-192) + the kernel's padded copy
            const bool is_k = which == GQ;
            *reinterpret_cast<uint32_t*>((is_k ? k_cache : v_cache) + ((int64_t)a.kvh * S + pos) * 128 + 2 * pr) = o;
            *reinterpret_cast<uint32_t*>((is_k ? k_pad : v_pad) + ((int64_t)a.kvh * p.S_pad + pos) * kKvPitch + 2 * pr) = o;
          }
        }
      }
    }
  }
  named_bar_sync(1, kCT);  // q and the new k / v are in shared memory
  if (tid == 0) stamp(p.trace, n_phases, i, 2);

  int slot = (int)(state & 0xffu);
  uint32_t par = state >> 8;
  const int g = lane >> 2, c = lane & 3;
  const float scale = p.attn_scale;
  // this lane's slice of the state: heads 2 c, 2 c + 1 (columns of the accumulators); rows = dims 16 dm + g, + 8
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  float Ot[8][4];
#pragma unroll
  for (int dm = 0; dm < 8; ++dm) Ot[dm][0] = Ot[dm][1] = Ot[dm][2] = Ot[dm][3] = 0.f;
  const int round_max = min(kAttnRound, stages - 1);
  // ldmatrix row addresses of this lane: matrix mi = lane >> 3, row r8 = lane & 7
  const int mi = lane >> 3, r8 = lane & 7;
  // K (A operand, row = position): matrices (rows 0-7 | 8-15) x (dims 0-7 | 8-15) in the order a0..a3
  const uint32_t k_lane = (uint32_t)(8 * (mi & 1) + r8) * (kKvPitch * 2) + (uint32_t)(mi >> 1) * 16u;   // + tile + 16 mt rows + 32 ks bytes
  // V^T (A operand, row = dim): stored blocks (positions 0-7 | 8-15 -> a0/a1 vs a2/a3) x (dims 0-7 | 8-15)
  const uint32_t v_lane = (uint32_t)kKvTileBytes + (uint32_t)(8 * (mi >> 1) + r8) * (kKvPitch * 2) + (uint32_t)(mi & 1) * 16u;
  const uint32_t q_lane = smem_u32(q_s) + (uint32_t)g * (kKvPitch * 2) + (uint32_t)c * 4u;  // Q^T (B operand): head g, dims 2 c, + 1

  for (int tb = a.t0; tb < a.t1; tb += round_max) {
    const int R = min(round_max, a.t1 - tb);
    {
      int sl = slot;
      uint32_t pr = par;
      for (int r = 0; r < R; ++r) {  // every warp waits for every tile of the round (the patch below needs a CTA barrier)
        bar_wait(bars.full + 8u * sl, pr, 32u);
        if (++sl == stages) {
          sl = 0;
          pr ^= 1u;
        }
      }
    }
    if (owner && a.n_t - 1 < tb + R) {  // CTA-uniform: the round holds the tile of the new position (always the last one)
      int sl = slot + (a.n_t - 1 - tb);
      if (sl >= stages) sl -= stages;
      const uint32_t stage_u = ring_u + (uint32_t)sl * slot_bytes;
      const int row = pos - (a.n_t - 1) * kAttnPos;
      if (tid < 128)
        sts32(stage_u + (uint32_t)(tid >> 6) * kKvTileBytes + (uint32_t)row * (kKvPitch * 2) + (uint32_t)(tid & 63) * 4u, kv_new[tid]);
      fence_proxy_async();  // generic writes into a buffer the bulk-copy engine will overwrite later
      named_bar_sync(1, kCT);
    }
    if (tid == 0 && tb == a.t0) stamp(p.trace, n_phases, i, 6);  // the first round's tiles are in shared memory
    if (warp < R) {
      int sl = slot + warp;
      if (sl >= stages) sl -= stages;
      const uint32_t stage_u = ring_u + (uint32_t)sl * slot_bytes;
      const int jt = (tb + warp) * kAttnPos;
      // ---- S^T = K Q^T: 2 blocks of 16 positions x 8 k-steps of 16 dims
      float st[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) st[mt][0] = st[mt][1] = st[mt][2] = st[mt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t qb0 = g < GQ ? lds32(q_lane + 32u * ks) : 0u, qb1 = g < GQ ? lds32(q_lane + 32u * ks + 16u) : 0u;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          uint32_t ka[4];
          ldsm_x4(ka, stage_u + k_lane + (uint32_t)(16 * mt) * (kKvPitch * 2) + 32u * ks);
          mma_bf16(st[mt], ka, qb0, qb1);
        }
      }
      // ---- online softmax: this lane holds positions g, g + 8, g + 16, g + 24 of the tile for heads 2 c, 2 c + 1
      float corr[2];
      uint32_t pk[2][2];  // P^T blocks (8 positions x 8 heads) as bf16x2 (heads 2c | 2c + 1): [16-position block][first | second 8]
      {
        float mt2[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const bool ok = jt + 16 * mt + 8 * hf + g < a.len;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              st[mt][2 * hf + e] = ok ? st[mt][2 * hf + e] * scale : -INFINITY;
              mt2[e] = fmaxf(mt2[e], st[mt][2 * hf + e]);
            }
          }
        }
        float m_use[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          mt2[e] = fmaxf(mt2[e], __shfl_xor_sync(0xffffffffu, mt2[e], 4));
          mt2[e] = fmaxf(mt2[e], __shfl_xor_sync(0xffffffffu, mt2[e], 8));
          mt2[e] = fmaxf(mt2[e], __shfl_xor_sync(0xffffffffu, mt2[e], 16));
          const float m_new = fmaxf(m_run[e], mt2[e]);
          m_use[e] = m_new == -INFINITY ? 0.f : m_new;
          corr[e] = __expf(m_run[e] - m_use[e]);
          m_run[e] = m_new;
        }
        float ls[2] = {0.f, 0.f};
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            // exp(-inf) = 0 past the sequence; rounded to bf16 like softmax(...).to(bf16) @ V, and the row sum uses the
            // ROUNDED probabilities so that the weights and their normaliser agree
            const uint32_t w = cvt_bf2(__expf(st[mt][2 * hf] - m_use[0]), __expf(st[mt][2 * hf + 1] - m_use[1]));
            pk[mt][hf] = w;
            ls[0] += __uint_as_float(w << 16);
            ls[1] += __uint_as_float(w & 0xFFFF0000u);
          }
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          ls[e] += __shfl_xor_sync(0xffffffffu, ls[e], 4);
          ls[e] += __shfl_xor_sync(0xffffffffu, ls[e], 8);
          ls[e] += __shfl_xor_sync(0xffffffffu, ls[e], 16);
          l_run[e] = l_run[e] * corr[e] + ls[e];
        }
      }
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        Ot[dm][0] *= corr[0]; Ot[dm][2] *= corr[0];
        Ot[dm][1] *= corr[1]; Ot[dm][3] *= corr[1];
      }
      // ---- O^T += V^T P^T: 8 blocks of 16 dims x 2 k-steps of 16 positions
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        const uint32_t pb0 = movmatrix_t(pk[ks][0]), pb1 = movmatrix_t(pk[ks][1]);  // lane (g, c): head g, positions 2c, 2c + 1
#pragma unroll
        for (int dm = 0; dm < 8; ++dm) {
          uint32_t va[4];
          ldsm_x4_t(va, stage_u + v_lane + (uint32_t)(16 * ks) * (kKvPitch * 2) + 32u * dm);
          mma_bf16(Ot[dm], va, pb0, pb1);
        }
      }
    }
    __syncwarp();
    for (int r = 0; r < R; ++r) {  // every warp releases every slot of the round
      if (lane == 0) bar_arrive(bars.empty + 8u * slot);
      if (++slot == stages) {
        slot = 0;
        par ^= 1u;
      }
    }
  }
  if (tid == 0) stamp(p.trace, n_phases, i, 3);
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int h = 2 * c + e;
    if (h < GQ) {
      float* pp = part_s + (warp * GQ + h) * kPartStride;
#pragma unroll
      for (int dm = 0; dm < 8; ++dm) {
        pp[16 * dm + g] = Ot[dm][e];
        pp[16 * dm + g + 8] = Ot[dm][2 + e];
      }
      if (g == 0) {
        pp[128] = m_run[e];
        pp[129] = l_run[e];
      }
    }
  }
  named_bar_sync(1, kCT);
  if (warp < GQ) {
    // merge the 16 warps' partials of head kvh GQ + warp
    const int head = a.kvh * GQ + warp;
    float ms = -INFINITY, ls = 0.f;
    if (lane < kC) {
      ms = part_s[(lane * GQ + warp) * kPartStride + 128];
      ls = part_s[(lane * GQ + warp) * kPartStride + 129];
    }
    float mm = ms;
#pragma unroll
    for (int o = 16; o; o >>= 1) mm = fmaxf(mm, __shfl_xor_sync(0xffffffffu, mm, o));
    const float wgt = ls > 0.f ? __expf(ms - mm) : 0.f;
    float ll = warp_sum(ls * wgt);
    float o4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < kC; ++s) {
      const float ws = __shfl_sync(0xffffffffu, wgt, s);
      const float4 po = *reinterpret_cast<const float4*>(part_s + (s * GQ + warp) * kPartStride + lane * 4);
      o4[0] += po.x * ws; o4[1] += po.y * ws; o4[2] += po.z * ws; o4[3] += po.w * ws;
    }
    if (a.nsplit > 1) {
      // split partials as packets {fp32 | tag}: [head][split][kPartStride]
      uint64_t* pp = xq + aux_off + ((size_t)head * p.attn_max_split) * kPartStride;
      if (a.split > 0) {
        uint64_t* mine = pp + (size_t)a.split * kPartStride;
#pragma unroll
        for (int e = 0; e < 4; ++e) st_pk(mine + lane * 4 + e, __float_as_uint(o4[e]), tag_out);
        if (lane == 0) st_pk(mine + 128, __float_as_uint(mm), tag_out);
        if (lane == 1) st_pk(mine + 129, __float_as_uint(ll), tag_out);
      } else {
        for (int s = 1; s < a.nsplit; ++s) {
          const uint64_t* theirs = pp + (size_t)s * kPartStride;
          const float m2 = __uint_as_float(poll_pk(theirs + 128, tag_out));
          const float l2 = __uint_as_float(poll_pk(theirs + 129, tag_out));
          float a2[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) a2[e] = __uint_as_float(poll_pk(theirs + lane * 4 + e, tag_out));
          const float mn = fmaxf(mm, m2);
          const float w1 = __expf(mm - mn), w2 = l2 > 0.f ? __expf(m2 - mn) : 0.f;
          ll = ll * w1 + l2 * w2;
#pragma unroll
          for (int e = 0; e < 4; ++e) o4[e] = o4[e] * w1 + a2[e] * w2;
          mm = mn;
        }
      }
    }
    if (a.split == 0) {
      const float inv = 1.f / ll;
      uint64_t* dst = xq + out_off + ((head * 128 + lane * 4) >> 1);
      st_pk(dst, cvt_bf2(o4[0] * inv, o4[1] * inv), tag_out);
      st_pk(dst + 1, cvt_bf2(o4[2] * inv, o4[3] * inv), tag_out);
    }
  }
  if (tid == 0) {
    red_ctr(p.counter);
    stamp(p.trace, n_phases, i, 4);
  }
  return (uint32_t)slot | (par << 8);
}

// ------------------------------------------------------------------------------------------------ the kernel
// ------------------------------------------------------------------------------------------------ consumers: one phase
struct ConsumerCtx {
  uint32_t tag0;
  uint64_t cnt0;
  int pos;
  uint32_t ring_u, bar0, red_u, x_u, S_u;
  uint8_t* x_s;
  uint4* h_s;
  float* part;
  float* row_scale;
  uint32_t state;  // ring slot | parity << 8
  uint32_t li;     // strips handed to the epilogue warp so far
  int gphase, rot;
};

// phase i = kind KIND of layer l (the shape-specialised loop: no division, no select on the kind)
template <int KIND>
__device__ __forceinline__ Phase get_phase_k(const Params& p, int l, int i) {
  extern __shared__ __align__(128) uint8_t smem_mega[];
  const LayerPtrs* lay_s = reinterpret_cast<const LayerPtrs*>(smem_mega + p.off_lay);
  const Kind& k = p.kinds[KIND];
  Phase ph;
  ph.in_off = (i == 0) ? kNone : k.in_off;
  ph.out_off = k.out_off; ph.aux_off = k.aux_off;
  ph.type = k.type; ph.in_kind = k.in_kind; ph.out_kind = k.out_kind;
  ph.KT = k.KT; ph.n_strips = k.n_strips; ph.N_out = k.N_out; ph.Kx = k.Kx; ph.rot_inc = k.rot_inc; ph.gate = k.gate;
  ph.packed = nullptr; ph.sz = nullptr; ph.norm_w = nullptr; ph.out_plain = nullptr; ph.k_cache = nullptr; ph.v_cache = nullptr;
  ph.k_pad = nullptr; ph.v_pad = nullptr;
  if constexpr (KIND == K_OUT) {
    ph.norm_w = p.out_norm;
  } else if constexpr (KIND == K_ATTN) {
    ph.k_cache = lay_s[l].kc; ph.v_cache = lay_s[l].vc; ph.k_pad = lay_s[l].kp; ph.v_pad = lay_s[l].vp;
  } else if constexpr (KIND == K_WQKV) {
    ph.norm_w = lay_s[l].norm[0];
  } else if constexpr (KIND == K_W13) {
    ph.norm_w = lay_s[l].norm[1];
  }
  return ph;  // (the consumers need neither the weight pointers nor the output pointer: producer / epilogue warp)
}

template <int G, int GQ>
__device__ __forceinline__ void consumer_attn(const Phase& ph, int i, ConsumerCtx& cx) {
  const Params& p = sparams();
  const uint32_t tag_in = cx.tag0 + (uint32_t)i;
  const uint64_t cnt_target = cx.cnt0 + (uint64_t)i * (uint64_t)p.n_cta - (uint64_t)ph.gate;
  cx.state = attention_phase<GQ>(ph.k_cache, ph.v_cache, ph.k_pad, ph.v_pad, ph.in_off, ph.out_off, ph.aux_off, i, cx.pos, tag_in,
                                 cnt_target, cx.ring_u, Geom<G>::kSlotBytes, cx.bar0, cx.x_s, cx.state);
}

// INK / NCH: the phase's input kind and staging width when known at compile time (-1 / 0: decided here)
template <int G, int INK, int NCH>
__device__ __forceinline__ void consumer_gemv(const Phase& ph, int i, ConsumerCtx& cx) {
  const Params& p = sparams();
  const int n_cta = p.n_cta;
  const uint32_t tag_in = cx.tag0 + (uint32_t)i;  // tag of phase i - 1
  // arrivals to wait for before sweeping: all but `gate` — the stragglers' packets are polled directly, which takes one
  // L2 round trip off the critical path (measured: profiles/)
  const uint64_t cnt_target = cx.cnt0 + (uint64_t)i * (uint64_t)n_cta - (uint64_t)ph.gate;
  const int KT = ph.KT;
  const int n_chunks = KT * 8;  // Kp == Kx in this kernel (check_model)
  const int nch = NCH > 0 ? NCH : ((KT + kStageTiles - 1) / kStageTiles + 1) / 2;  // stages of the strip, two per staging round
  float* rs = cx.row_scale + (cx.gphase & 1);
  float* part_g = cx.part + (cx.gphase & 1) * 2 * kC;
  bool in_h;
  if constexpr (INK < 0) in_h = ph.in_kind == IN_H; else in_h = INK == IN_H;
  if (in_h) {
    if constexpr (INK < 0 || INK == IN_H) {
      const float inv_k = 1.0f / (float)ph.Kx;
      if constexpr (NCH == 0) {
        if (nch <= 1) stage_h<G, 1>(n_chunks, ph.in_off, ph.norm_w, inv_k, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, cx.h_s, part_g);
        else stage_h<G, 2>(n_chunks, ph.in_off, ph.norm_w, inv_k, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, cx.h_s, part_g);
      } else {
        stage_h<G, (NCH <= 1 ? 1 : 2)>(n_chunks, ph.in_off, ph.norm_w, inv_k, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, cx.h_s, part_g);
      }
    }
  } else {
    if constexpr (INK < 0 || INK == IN_PK) {
      if constexpr (NCH == 0) {
        if (nch <= 1) stage_x<G, 1>(n_chunks, ph.in_off, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, part_g);
        else if (nch <= 2) stage_x<G, 2>(n_chunks, ph.in_off, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, part_g);
        else if (nch <= 4) stage_x<G, 4>(n_chunks, ph.in_off, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, part_g);
        else stage_x<G, 8>(n_chunks, ph.in_off, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, part_g);
      } else {
        stage_x<G, NCH>(n_chunks, ph.in_off, i, tag_in, cnt_target, rs, cx.x_u, cx.S_u, part_g);
      }
    }
  }
  int s0, s1;
  strip_range((int)blockIdx.x, n_cta, ph.n_strips, cx.rot, s0, s1);
  cx.rot = next_rot(cx.rot, ph.rot_inc, n_cta);
  const uint32_t limb_row = (uint32_t)KT * 64u + 32u;
  bool xreg;
  if constexpr (NCH == 0) xreg = KT <= 2 * kStageTiles; else xreg = NCH == 1;
  if (xreg) {
    if constexpr (NCH <= 1)
      cx.state = gemv_main<G, true>(cx.ring_u, cx.bar0, p.stages, KT, s1 - s0, cx.x_u, limb_row, cx.S_u, cx.red_u, cx.state, cx.li, p.dbg);
  } else {
    if constexpr (NCH != 1)
      cx.state = gemv_main<G, false>(cx.ring_u, cx.bar0, p.stages, KT, s1 - s0, cx.x_u, limb_row, cx.S_u, cx.red_u, cx.state, cx.li, p.dbg);
  }
  cx.li += (uint32_t)(s1 - s0);
  if (threadIdx.x == 0) stamp(p.trace, p.n_phases, i, 3);
  ++cx.gphase;
}

// GQ_T == 0: any shape.  Otherwise the model's shape class is a template argument: q heads per kv head, and the chunks of
// 8 activations per thread when staging dim (NH_T), the attention output (NWO_T) and the intermediate vector (NW2_T).
template <int G, int GQ_T, int NH_T, int NWO_T, int NW2_T>
__global__ void __launch_bounds__(kThreads, 1) decode_mega_kernel(const __grid_constant__ Params p_arg) {
  using GE = Geom<G>;
  constexpr int kStageBytes = GE::kSlotBytes;
  extern __shared__ __align__(128) uint8_t smem_mega[];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&p_arg);
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem_mega + 64);
    for (int k = threadIdx.x; k < (int)(sizeof(Params) / 4); k += kThreads) dst[k] = src[k];
  }
  const Params& p = sparams();  // valid after the __syncthreads below; until then only p_arg is used
  uint8_t* ring = smem_mega + kHdrBytes;
  uint8_t* x_s = smem_mega + p_arg.off_x;                              // limbs [3][Kp + 32] int8 / attention scratch
  float* S_s = reinterpret_cast<float*>(smem_mega + p_arg.off_S);      // [Kp / kSegK][8]: limb sums per segment
  uint4* h_s = reinterpret_cast<uint4*>(smem_mega + p_arg.off_h);      // residual stream, bf16 [dim]
  float* part = reinterpret_cast<float*>(smem_mega + p_arg.off_misc);  // [16 warps][2]: {sum sq, max |.| bits}
  float* row_scale = part + 4 * kC;                                // [2] (GEMV-phase parity); part is double-buffered too
  uint64_t* bar_p = reinterpret_cast<uint64_t*>(row_scale + 4);    // full[stages], empty[stages], red_full[2], red_free[2]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x, n_cta = p_arg.n_cta;
  const int stages = p_arg.stages, n_phases = p_arg.n_phases;
  {
    const uint4* src = reinterpret_cast<const uint4*>(p_arg.layers);
    uint4* dst = reinterpret_cast<uint4*>(smem_mega + p_arg.off_lay);
    for (int k = tid; k < p_arg.n_layer * (int)(sizeof(LayerPtrs) / 16); k += kThreads) dst[k] = __ldg(src + k);
  }
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&bar_p[s], 1);
      mbar_init(&bar_p[stages + s], kC);
    }
    for (int s = 0; s < kRedBufs; ++s) {
      mbar_init(&bar_p[2 * stages + s], kC);
      mbar_init(&bar_p[2 * stages + kRedBufs + s], 1);
    }
    SCtx* c = sctx();
    c->abort = 0u;
    c->status = p_arg.status;
    c->timeout_ns = p_arg.timeout_ns;
    fence_barrier_init();
  }
  __syncthreads();

  const uint32_t step = *reinterpret_cast<volatile uint32_t*>(p_arg.step);
  const int pos = *reinterpret_cast<volatile int32_t*>(p_arg.pos);
  const uint32_t ring_u = smem_u32(ring), bar0 = smem_u32(bar_p), red_u = smem_u32(smem_mega + p_arg.off_red);

  if (warp == kC) {
    if (lane == 0) producer_loop<G>(pos, ring_u, bar0);
    return;
  }
  if (warp == kC + 1) {
    epilogue_loop(pos, step, red_u, bar0, row_scale);
    return;
  }

  // ===================== consumers =====================
  {
    // residual stream := embedding row of the fed token (model.py:358)
    const int64_t tok = *reinterpret_cast<const volatile int64_t*>(p.tok);
    const uint4* erow = reinterpret_cast<const uint4*>(p.emb + tok * (int64_t)p.dim);
    for (int ci = tid; ci < p.dim / 8; ci += kCT) h_s[ci] = __ldg(erow + ci);
  }
  ConsumerCtx cx;
  cx.tag0 = step * (uint32_t)n_phases;                              // phase i publishes with tag tag0 + i + 1
  cx.cnt0 = (uint64_t)step * (uint64_t)n_phases * (uint64_t)n_cta;  // counter before this launch
  cx.pos = pos;
  cx.ring_u = ring_u; cx.bar0 = bar0; cx.red_u = red_u;
  cx.x_s = x_s; cx.x_u = smem_u32(x_s); cx.S_u = smem_u32(S_s);
  cx.h_s = h_s; cx.part = part; cx.row_scale = row_scale;
  cx.state = 0u; cx.li = 0u; cx.gphase = 0; cx.rot = 0;
  if constexpr (GQ_T == 0) {
    // any shape: one dispatcher per phase (every phase switch jumps through it: instruction-cache misses on the critical path)
    const int GQ = p.H / p.Hkv;
    for (int i = 0; i < n_phases; ++i) {
      const Phase ph = get_phase(p, i);
      if (ph.type != PH_GEMV) {
        switch (GQ) {
          case 1: consumer_attn<G, 1>(ph, i, cx); break;
          case 2: consumer_attn<G, 2>(ph, i, cx); break;
          case 4: consumer_attn<G, 4>(ph, i, cx); break;
          default: consumer_attn<G, 8>(ph, i, cx); break;
        }
      } else {
        consumer_gemv<G, -1, 0>(ph, i, cx);
      }
    }
  } else {
    // a known shape: the layer is straight-line code in execution order (see `variants` on the host side: measured, not used)
    const int n_layer = p.n_layer;
    for (int l = 0; l < n_layer; ++l) {
      const int i = 5 * l;
      consumer_gemv<G, IN_H, NH_T>(get_phase_k<K_WQKV>(p, l, i), i, cx);
      consumer_attn<G, GQ_T>(get_phase_k<K_ATTN>(p, l, i + 1), i + 1, cx);
      consumer_gemv<G, IN_PK, NWO_T>(get_phase_k<K_WO>(p, l, i + 2), i + 2, cx);
      consumer_gemv<G, IN_H, NH_T>(get_phase_k<K_W13>(p, l, i + 3), i + 3, cx);
      consumer_gemv<G, IN_PK, NW2_T>(get_phase_k<K_W2>(p, l, i + 4), i + 4, cx);
    }
    consumer_gemv<G, IN_H, NH_T>(get_phase_k<K_OUT>(p, n_layer, 5 * n_layer), 5 * n_layer, cx);
  }
}

// ------------------------------------------------------------------------------------------------ host side
struct SmemPlan {
  int stages;
  size_t smem;
  uint32_t off_x, off_S, off_red, off_h, off_misc, off_lay;
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static bool plan_smem(int G, int64_t Kmax, int dim, int Gq, int n_layer, SmemPlan& pl) {
  const int seg_k = G < 128 ? G : 128;
  size_t stage_bytes = (size_t)kStageW + (size_t)(kStageTiles * 64 / G) * 64;
  if (stage_bytes < (size_t)kAttnStageBytes) stage_bytes = kAttnStageBytes;  // a ring slot also holds a K / V tile
  const size_t limbs = (size_t)3 * (size_t)(Kmax + 32);
  const size_t attn = (size_t)8 * kKvPitch * 2 + 512 + (size_t)kC * Gq * kPartStride * 4;  // q, new k / v, warp partials
  const size_t x_bytes = align_up(limbs > attn ? limbs : attn, 128);
  const size_t S_bytes = align_up((size_t)(Kmax / seg_k) * 32, 128);
  const size_t red_bytes = (size_t)kRedBufs * kC * 128 * 4;
  const size_t h_bytes = align_up((size_t)dim * 2, 128);
  const size_t misc = align_up((4 * kC + 4) * 4 + (2 * 16 + 2 * kRedBufs) * 8, 128);
  const size_t lay = align_up((size_t)n_layer * sizeof(LayerPtrs), 128);
  const size_t fixed = x_bytes + S_bytes + red_bytes + h_bytes + misc + lay;
  const size_t cap = 227 * 1024 - kHdrBytes;
  if (fixed + 3 * stage_bytes > cap) return false;
  int st = (int)((cap - fixed) / stage_bytes);
  if (st > 16) st = 16;
  pl.stages = st;
  size_t off = kHdrBytes + (size_t)st * stage_bytes;
  pl.off_x = (uint32_t)off; off += x_bytes;
  pl.off_S = (uint32_t)off; off += S_bytes;
  pl.off_red = (uint32_t)off; off += red_bytes;
  pl.off_h = (uint32_t)off; off += h_bytes;
  pl.off_misc = (uint32_t)off; off += misc;
  pl.off_lay = (uint32_t)off; off += lay;
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
  const uint32_t slots = m.world > 1 ? (uint32_t)m.world + 1u : 1u;  // the ranks' partial sums + their local sum
  uint32_t off = 0;
  o.qkv = off; off += al(qkv_n / 2);
  o.att = off; off += al((uint32_t)m.n_head * 128u / 2);
  o.d1 = off; off += al(slots * (uint32_t)m.dim / 2);
  o.mid = off; off += al((uint32_t)m.inter / 2);
  o.d2 = off; off += al(slots * (uint32_t)m.dim / 2);
  o.aux = off; off += al((uint32_t)m.n_head * (uint32_t)(m.n_cta / (m.n_kv_head > 0 ? m.n_kv_head : 1) + 1) * kPartStride);
  o.total = off;
  return o;
}

static int check_model(const ao_b200_mega_model* m) {
  AO_CHECK_ARG(m && m->layers && m->n_layer > 0 && m->n_cta > 0);
  AO_CHECK_ARG(m->head_dim == 128 && m->n_kv_head > 0 && m->n_head % m->n_kv_head == 0);
  const int gq = m->n_head / m->n_kv_head;
  AO_CHECK_ARG(gq == 1 || gq == 2 || gq == 4 || gq == 8);
  AO_CHECK_ARG(m->dim % 128 == 0 && m->inter % 128 == 0 && m->vocab > 0);
  if (m->dim > 8192 || m->inter > 32768 || m->n_head * 128 > 32768) return AO_B200_ERR_UNSUPPORTED;  // staging: <= 2 / 8 chunks of 8 per thread
  AO_CHECK_ARG(m->world >= 1 && m->world <= kMaxWorld && (m->world & (m->world - 1)) == 0 && m->rank >= 0 && m->rank < m->world);
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
  return (int64_t)sizeof(LayerPtrs) * (int64_t)m->n_layer;
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
  std::vector<LayerPtrs> lp((size_t)m->n_layer);
  for (int l = 0; l < m->n_layer; ++l) {
    const ao_b200_mega_layer& L = m->layers[l];
    AO_CHECK_ARG(L.wqkv_packed && L.wqkv_sz && L.wo_packed && L.wo_sz && L.w13_packed && L.w13_sz && L.w2_packed && L.w2_sz);
    AO_CHECK_ARG(L.attn_norm && L.ffn_norm && L.k_cache && L.v_cache && L.k_pad && L.v_pad);
    LayerPtrs& q = lp[(size_t)l];
    q.w[0] = static_cast<const uint8_t*>(L.wqkv_packed); q.sz[0] = static_cast<const uint32_t*>(L.wqkv_sz);
    q.w[1] = static_cast<const uint8_t*>(L.wo_packed); q.sz[1] = static_cast<const uint32_t*>(L.wo_sz);
    q.w[2] = static_cast<const uint8_t*>(L.w13_packed); q.sz[2] = static_cast<const uint32_t*>(L.w13_sz);
    q.w[3] = static_cast<const uint8_t*>(L.w2_packed); q.sz[3] = static_cast<const uint32_t*>(L.w2_sz);
    q.norm[0] = static_cast<const __nv_bfloat16*>(L.attn_norm); q.norm[1] = static_cast<const __nv_bfloat16*>(L.ffn_norm);
    q.kc = static_cast<__nv_bfloat16*>(L.k_cache); q.vc = static_cast<__nv_bfloat16*>(L.v_cache);
    q.kp = static_cast<__nv_bfloat16*>(L.k_pad); q.vp = static_cast<__nv_bfloat16*>(L.v_pad);
  }
  auto s = static_cast<cudaStream_t>(stream);
  if (cudaMemcpyAsync(plan_dev, lp.data(), sizeof(LayerPtrs) * lp.size(), cudaMemcpyHostToDevice, s) != cudaSuccess) return AO_B200_ERR_CUDA;
  if (cudaStreamSynchronize(s) != cudaSuccess) return AO_B200_ERR_CUDA;  // lp is a host temporary (set-up time only)
  return AO_B200_OK;
}

extern "C" int ao_b200_decode_mega_step(const ao_b200_mega_model* m, const void* plan_dev, void* stream) {
  const int rc = check_model(m);
  if (rc != AO_B200_OK) return rc;
  AO_CHECK_ARG(plan_dev);
  const int Gq = m->n_head / m->n_kv_head;
  const int64_t qkv_n = (int64_t)(m->n_head + 2 * m->n_kv_head) * 128, att_n = (int64_t)m->n_head * 128;
  int64_t Kmax = m->dim > m->inter ? m->dim : m->inter;
  if (att_n > Kmax) Kmax = att_n;
  SmemPlan pl;
  if (!plan_smem(m->group_size, Kmax, m->dim, Gq, m->n_layer, pl)) return AO_B200_ERR_WORKSPACE;
  Params p;
  memset(&p, 0, sizeof p);
  p.layers = static_cast<const LayerPtrs*>(plan_dev);
  p.n_layer = m->n_layer;
  p.n_phases = 5 * m->n_layer + 1;
  p.n_cta = m->n_cta;
  {
    const Offsets o = xchg_layout(*m);
    auto gemv = [&](Kind& q, int64_t N, int64_t K, int in_kind, uint32_t in_off, int out_kind, uint32_t out_off) {
      q.type = PH_GEMV;
      q.KT = (int)(K / 64);
      q.Kx = (int)K;
      q.N_out = (int)N;
      q.n_strips = (int)((N + 15) / 16);
      q.rot_inc = q.n_strips % m->n_cta;  // the CTAs that got an extra strip rotate from phase to phase
      q.in_kind = in_kind; q.in_off = in_off; q.out_kind = out_kind; q.out_off = out_off;
    };
    gemv(p.kinds[K_WQKV], qkv_n, m->dim, IN_H, o.d2, OUT_PK, o.qkv);
    p.kinds[K_ATTN].type = PH_ATTN; p.kinds[K_ATTN].in_off = o.qkv; p.kinds[K_ATTN].out_off = o.att; p.kinds[K_ATTN].aux_off = o.aux;
    gemv(p.kinds[K_WO], m->dim, att_n, IN_PK, o.att, OUT_DELTA, o.d1);
    gemv(p.kinds[K_W13], 2 * (int64_t)m->inter, m->dim, IN_H, o.d1, OUT_SILU, o.mid);
    p.kinds[K_W13].N_out = m->inter;  // outputs after SiLU*mul (n_strips stays 2 I / 16)
    gemv(p.kinds[K_W2], m->dim, m->inter, IN_PK, o.mid, OUT_DELTA, o.d2);
    gemv(p.kinds[K_OUT], m->vocab, m->dim, IN_H, o.d2, OUT_PLAIN, 0);
  }
  p.out_w = static_cast<const uint8_t*>(m->out_packed);
  p.out_sz = static_cast<const uint32_t*>(m->out_sz);
  p.out_norm = static_cast<const __nv_bfloat16*>(m->out_norm);
  p.logits = static_cast<__nv_bfloat16*>(m->logits);
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
  p.status = reinterpret_cast<uint32_t*>(st + 12);   // [status][slow polls]
  p.dim = m->dim; p.H = m->n_head; p.Hkv = m->n_kv_head; p.S = m->max_seq;
  p.S_pad = (m->max_seq + kAttnPos - 1) / kAttnPos * kAttnPos;
  p.attn_spc = m->attn_blocks_per_cta > 0 ? m->attn_blocks_per_cta : kAttnRound;
  p.attn_max_split = m->n_cta / m->n_kv_head;
  p.eps = m->eps;
  p.attn_scale = 1.0f / sqrtf((float)m->head_dim);
  p.stages = pl.stages;
  p.off_x = pl.off_x; p.off_S = pl.off_S; p.off_red = pl.off_red; p.off_h = pl.off_h; p.off_misc = pl.off_misc; p.off_lay = pl.off_lay;
  p.timeout_ns = (unsigned long long)(m->timeout_ms > 0 ? m->timeout_ms : 2000) * 1000000ull;
  p.trace = static_cast<unsigned long long*>(m->trace);
  {
    static const int dbg = []() { const char* e = std::getenv("AO_B200_MEGA_DEBUG"); return e ? std::atoi(e) : 0; }();
    p.dbg = dbg;
    static const int tpr = []() { const char* e = std::getenv("AO_B200_MEGA_TP_REDUCE"); return e ? std::atoi(e) : -1; }();
    p.tp_reduce = m->world > 1 && (tpr >= 0 ? tpr != 0 : m->world >= 4);
    static const int gate_env = []() { const char* e = std::getenv("AO_B200_MEGA_GATE"); return e ? std::atoi(e) : 32; }();
    const int gate = gate_env < 0 ? 0 : (gate_env > m->n_cta - 1 ? m->n_cta - 1 : gate_env);
    for (int k = 0; k < kKinds; ++k) p.kinds[k].gate = gate;
  }
  using KernelFn = void (*)(const Params);
  struct Variant {
    int G;
    KernelFn fn;
  };
  // One kernel per group size, any shape.  (Instantiations per shape class — GQ_T / NH_T / NWO_T / NW2_T > 0: the layer as
  // straight-line code in execution order, no dispatcher between the phases — were measured on Llama-3-8B / 70B: the
  // consumers reach the next phase's wait 0.5 - 1 us earlier and wait that much longer, because the hand-over is bound by
  // the slowest CTA's epilogue, not by the consumers' instruction path: 828 vs 833 tok/s.  Not instantiated.)
  static const Variant variants[] = {
      {32, decode_mega_kernel<32, 0, 0, 0, 0>},   {64, decode_mega_kernel<64, 0, 0, 0, 0>},
      {128, decode_mega_kernel<128, 0, 0, 0, 0>}, {256, decode_mega_kernel<256, 0, 0, 0, 0>},
  };
  constexpr int kVariants = (int)(sizeof(variants) / sizeof(variants[0]));
  KernelFn kern = nullptr;
  int gi = 0;
  for (int v = 0; v < kVariants && kern == nullptr; ++v) {
    if (variants[v].G == m->group_size) {
      kern = variants[v].fn;
      gi = v;
    }
  }
  if (kern == nullptr) return AO_B200_ERR_UNSUPPORTED;
  static const bool debug = std::getenv("AO_B200_DEBUG") != nullptr;
  static bool attr_done[kVariants] = {};
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
  // The CTAs wait on one another, so all of them must be resident.  A cooperative launch guarantees it (or fails); a
  // plain launch of <= (SMs x occupancy) CTAs gets there too as soon as whatever else runs on the GPU drains, and every
  // wait in the kernel is bounded — AO_B200_MEGA_COOP=0 selects it (a cooperative kernel node costs the serving loop
  // extra launch latency per token: profiles/).  The grid is checked against the occupancy either way.
  static const bool coop = []() { const char* e = std::getenv("AO_B200_MEGA_COOP"); return e ? std::atoi(e) != 0 : true; }();
  static int max_ctas[kVariants] = {};
  if (max_ctas[gi] == 0) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, pl.smem) != cudaSuccess) per_sm = 0;
    max_ctas[gi] = sms * per_sm;
  }
  if (m->n_cta > max_ctas[gi]) {
    if (debug) fprintf(stderr, "ao_b200_decode_mega_step: %d CTAs cannot be co-resident (limit %d)\n", m->n_cta, max_ctas[gi]);
    return AO_B200_ERR_UNSUPPORTED;
  }
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = coop ? 1 : 0;
  const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p);
  if (le != cudaSuccess) {
    if (debug) fprintf(stderr, "ao_b200_decode_mega_step: launch (grid %d, smem %zu, stages %d): %s\n", m->n_cta, pl.smem, pl.stages, cudaGetErrorString(le));
    cudaGetLastError();
    return AO_B200_ERR_CUDA;
  }
  return ao_launch_status();
}
