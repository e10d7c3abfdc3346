// Non-GEMM remainder of one Llama decode step: fused residual-add + RMSNorm,
// RoPE + KV-cache append, GQA decode attention, SiLU*mul.  All read the token position from DEVICE
// memory so a captured CUDA graph can be replayed for every token.
#include <mutex>

#include "common.cuh"

namespace ao {

// Scratch of ao_b200_decode_attention_static (bottom of this file): lives in the library image, every GPU that loads
// the library holds it until exit.  In HOST code this name is only a stand-in for the symbol — the device address
// comes from cudaGetSymbolAddress (static_scratch_ptr below), never from `&g_attn_static_scratch`.
static __device__ __align__(256) unsigned char g_attn_static_scratch[AO_B200_STATIC_SCRATCH_BYTES];

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------------------------------
// h = bf16(x + residual) (optional); y = bf16( bf16(h * rsqrt(mean(h^2) + eps)) * w )
// (model.py:489-503: norm in fp32, cast back, then multiply by the bf16 weight).  One CTA per row.
// ---------------------------------------------------------------------------------------
template <int kThreads>
__global__ void __launch_bounds__(kThreads) rmsnorm_kernel(const __nv_bfloat16* __restrict__ x,
                                                           const __nv_bfloat16* __restrict__ residual,
                                                           const __nv_bfloat16* __restrict__ weight,
                                                           __nv_bfloat16* __restrict__ h_out,
                                                           __nv_bfloat16* __restrict__ y, int64_t dim, float eps) {
  extern __shared__ float s_row[];  // dim floats
  __shared__ float s_part[kThreads / 32];
  const int64_t row = blockIdx.x;
  const __nv_bfloat16* xr = x + row * dim;
  const __nv_bfloat16* rr = residual ? residual + row * dim : nullptr;
  float ss = 0.f;
  for (int64_t i = threadIdx.x * 8; i < dim; i += kThreads * 8) {
    float f[8];
    unpack8(*reinterpret_cast<const uint4*>(xr + i), f);
    if (rr) {
      float r[8];
      unpack8(*reinterpret_cast<const uint4*>(rr + i), r);
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = bf16r(f[e] + r[e]);
      if (h_out) {
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) o[e] = float_to_bf16_bits(f[2 * e]) | (float_to_bf16_bits(f[2 * e + 1]) << 16);
        *reinterpret_cast<uint4*>(h_out + row * dim + i) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      s_row[i + e] = f[e];
      ss += f[e] * f[e];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < kThreads / 32; ++i) tot += s_part[i];
  const float rs = rsqrtf(tot / static_cast<float>(dim) + eps);
  for (int64_t i = threadIdx.x * 8; i < dim; i += kThreads * 8) {
    float wv[8];
    unpack8(*reinterpret_cast<const uint4*>(weight + i), wv);
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a = bf16r(s_row[i + 2 * e] * rs) * wv[2 * e];
      const float b = bf16r(s_row[i + 2 * e + 1] * rs) * wv[2 * e + 1];
      o[e] = float_to_bf16_bits(a) | (float_to_bf16_bits(b) << 16);
    }
    *reinterpret_cast<uint4*>(y + row * dim + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ---------------------------------------------------------------------------------------
// RoPE (model.py:547-559, fp32 math on bf16 cos/sin) on q and k + KV-cache append at *input_pos
// (KVCache.update, model.py:175-192).  qkv: [B][(H + 2 Hkv) * hd]; caches [B][Hkv][S][hd].
// One thread per (b, head, pair).
// ---------------------------------------------------------------------------------------
__global__ void rope_kv_append_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ freqs,
                                      const int32_t* __restrict__ input_pos, __nv_bfloat16* __restrict__ q_out,
                                      __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache, int B,
                                      int H, int Hkv, int hd, int S) {
  const int pos = *input_pos;
  const int half = hd / 2;
  const int heads = H + 2 * Hkv;
  const int64_t total = (int64_t)B * heads * half;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int p = i % half;
    const int head = (i / half) % heads;
    const int b = i / ((int64_t)half * heads);
    const __nv_bfloat162 v2 = *reinterpret_cast<const __nv_bfloat162*>(qkv + ((int64_t)b * heads + head) * hd + 2 * p);
    if (head < H + Hkv) {
      const __nv_bfloat162 cs = *reinterpret_cast<const __nv_bfloat162*>(freqs + ((int64_t)pos * half + p) * 2);
      const float x0 = __bfloat162float(v2.x), x1 = __bfloat162float(v2.y);
      const float c = __bfloat162float(cs.x), s = __bfloat162float(cs.y);
      __nv_bfloat162 o;
      o.x = __float2bfloat16_rn(x0 * c - x1 * s);
      o.y = __float2bfloat16_rn(x1 * c + x0 * s);
      if (head < H)
        *reinterpret_cast<__nv_bfloat162*>(q_out + ((int64_t)b * H + head) * hd + 2 * p) = o;
      else
        *reinterpret_cast<__nv_bfloat162*>(k_cache + (((int64_t)b * Hkv + (head - H)) * S + pos) * hd + 2 * p) = o;
    } else {
      *reinterpret_cast<__nv_bfloat162*>(v_cache + (((int64_t)b * Hkv + (head - H - Hkv)) * S + pos) * hd + 2 * p) = v2;
    }
  }
}

// ---------------------------------------------------------------------------------------
// GQA decode attention for one new token: out[b][h][:] = softmax(q.K^T * scale)[0..pos] . V
// One CTA per (b, kv head, split); each warp owns one q head of the group and walks the key
// positions with an online softmax; hd == 128 (4 dims per lane), lanes reduce each dot product.
// Splits over the sequence are merged by a tiny second kernel when n_split > 1.
// ---------------------------------------------------------------------------------------
template <int kGroup>  // q heads per kv head (<= 8)
__global__ void __launch_bounds__(kGroup * 32) decode_attn_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k_cache,
    const __nv_bfloat16* __restrict__ v_cache, const int32_t* __restrict__ input_pos, __nv_bfloat16* __restrict__ out,
    float* __restrict__ part_o, float* __restrict__ part_ml, int H, int Hkv, int S, float scale, int n_split) {
  constexpr int hd = 128;
  const int pos = *input_pos;
  const int kvh = blockIdx.x, b = blockIdx.y, split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = kvh * kGroup + warp;
  const int len = pos + 1;
  const int chunk = (len + n_split - 1) / n_split;
  const int j0 = split * chunk, j1 = min(len, j0 + chunk);

  float qv[4];
  {
    const uint2 raw = *reinterpret_cast<const uint2*>(q + ((int64_t)b * H + h) * hd + lane * 4);
    qv[0] = __uint_as_float(raw.x << 16) * scale;
    qv[1] = __uint_as_float(raw.x & 0xFFFF0000u) * scale;
    qv[2] = __uint_as_float(raw.y << 16) * scale;
    qv[3] = __uint_as_float(raw.y & 0xFFFF0000u) * scale;
  }
  const __nv_bfloat16* kb = k_cache + ((int64_t)b * Hkv + kvh) * S * hd + lane * 4;
  const __nv_bfloat16* vb = v_cache + ((int64_t)b * Hkv + kvh) * S * hd + lane * 4;
  float m = -INFINITY, l = 0.f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int j = j0; j < j1; ++j) {
    const uint2 kr = *reinterpret_cast<const uint2*>(kb + (int64_t)j * hd);
    float s = qv[0] * __uint_as_float(kr.x << 16) + qv[1] * __uint_as_float(kr.x & 0xFFFF0000u) +
              qv[2] * __uint_as_float(kr.y << 16) + qv[3] * __uint_as_float(kr.y & 0xFFFF0000u);
    s = warp_sum(s);
    const float m_new = fmaxf(m, s);
    const float corr = __expf(m - m_new);
    const float p = __expf(s - m_new);
    const uint2 vr = *reinterpret_cast<const uint2*>(vb + (int64_t)j * hd);
    l = l * corr + p;
    acc[0] = acc[0] * corr + p * __uint_as_float(vr.x << 16);
    acc[1] = acc[1] * corr + p * __uint_as_float(vr.x & 0xFFFF0000u);
    acc[2] = acc[2] * corr + p * __uint_as_float(vr.y << 16);
    acc[3] = acc[3] * corr + p * __uint_as_float(vr.y & 0xFFFF0000u);
    m = m_new;
  }
  if (n_split == 1) {
    const float inv = 1.f / l;
    uint2 o;
    o.x = float_to_bf16_bits(acc[0] * inv) | (float_to_bf16_bits(acc[1] * inv) << 16);
    o.y = float_to_bf16_bits(acc[2] * inv) | (float_to_bf16_bits(acc[3] * inv) << 16);
    *reinterpret_cast<uint2*>(out + ((int64_t)b * H + h) * hd + lane * 4) = o;
  } else {
    const int64_t slot = ((int64_t)b * H + h) * n_split + split;
    *reinterpret_cast<float4*>(part_o + slot * hd + lane * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    if (lane == 0) {
      part_ml[slot * 2] = m;
      part_ml[slot * 2 + 1] = l;
    }
  }
}

__global__ void decode_attn_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                         __nv_bfloat16* __restrict__ out, int n_split) {
  constexpr int hd = 128;
  const int64_t bh = blockIdx.x;
  const int d = threadIdx.x;  // 128 threads
  float m = -INFINITY;
  for (int s = 0; s < n_split; ++s) m = fmaxf(m, part_ml[(bh * n_split + s) * 2]);
  float l = 0.f, acc = 0.f;
  for (int s = 0; s < n_split; ++s) {
    const float ms = part_ml[(bh * n_split + s) * 2], ls = part_ml[(bh * n_split + s) * 2 + 1];
    const float w = (ls > 0.f) ? __expf(ms - m) : 0.f;
    l += ls * w;
    acc += part_o[(bh * n_split + s) * hd + d] * w;
  }
  out[bh * hd + d] = __float2bfloat16_rn(acc / l);
}

// ---------------------------------------------------------------------------------------
// RoPE + KV append + GQA decode attention + split merge in ONE launch (PDL-aware): what Attention.forward does for
// one new token (model.py:405-440).  Grid (Hkv, B, n_split), one warp per q head of the group.
//   * every CTA ropes its own q heads from qkv (fp32 math on bf16 cos/sin, rounded to bf16 like the reference);
//   * the CTA whose split contains position `pos` ropes the new k, appends k and v to the caches, then walks it;
//   * splits leave (m, l, acc) partials in the workspace; the LAST CTA of a (b, kv head) to finish (atomic ticket, no
//     spinning) merges them and writes the bf16 output, and resets the ticket for the next launch.
// ---------------------------------------------------------------------------------------
template <int kGroup>
__global__ void __launch_bounds__(kGroup * 32) decode_attn_fused_kernel(
    const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ freqs,
    const int32_t* __restrict__ input_pos, __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
    __nv_bfloat16* __restrict__ out, float* __restrict__ part_o, float* __restrict__ part_ml,
    unsigned int* __restrict__ tickets, int H, int Hkv, int S, float scale, int n_split) {
  constexpr int hd = 128;
  __shared__ int s_last;
  grid_launch_dependents();
  const int kvh = blockIdx.x, b = blockIdx.y, split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = kvh * kGroup + warp;
  const int heads = H + 2 * Hkv;
  grid_dependency_wait();
  const int pos = *input_pos;
  const int len = pos + 1;
  const int chunk = (len + n_split - 1) / n_split;
  const int j0 = split * chunk, j1 = min(len, j0 + chunk);

  // cos/sin of this lane's two pairs (dims 4 lane .. 4 lane + 3)
  const uint2 cs = *reinterpret_cast<const uint2*>(freqs + ((int64_t)pos * (hd / 2) + 2 * lane) * 2);
  const float c0 = __uint_as_float(cs.x << 16), s0 = __uint_as_float(cs.x & 0xFFFF0000u);
  const float c1 = __uint_as_float(cs.y << 16), s1 = __uint_as_float(cs.y & 0xFFFF0000u);
  auto rope4 = [&](uint2 raw) {
    const float x0 = __uint_as_float(raw.x << 16), x1 = __uint_as_float(raw.x & 0xFFFF0000u);
    const float x2 = __uint_as_float(raw.y << 16), x3 = __uint_as_float(raw.y & 0xFFFF0000u);
    uint2 o;
    o.x = float_to_bf16_bits(x0 * c0 - x1 * s0) | (float_to_bf16_bits(x1 * c0 + x0 * s0) << 16);
    o.y = float_to_bf16_bits(x2 * c1 - x3 * s1) | (float_to_bf16_bits(x3 * c1 + x2 * s1) << 16);
    return o;
  };
  __nv_bfloat16* kb = k_cache + ((int64_t)b * Hkv + kvh) * S * hd + lane * 4;
  __nv_bfloat16* vb = v_cache + ((int64_t)b * Hkv + kvh) * S * hd + lane * 4;
  if (j0 <= pos && pos < j1) {  // block-uniform: this CTA owns the new position
    if (warp == 0) {
      const uint2 kraw = *reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + H + kvh) * hd + lane * 4);
      const uint2 vraw = *reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + H + Hkv + kvh) * hd + lane * 4);
      *reinterpret_cast<uint2*>(kb + (int64_t)pos * hd) = rope4(kraw);
      *reinterpret_cast<uint2*>(vb + (int64_t)pos * hd) = vraw;
    }
    __syncthreads();
  }
  float qv[4];
  {
    const uint2 r = rope4(*reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + h) * hd + lane * 4));
    qv[0] = __uint_as_float(r.x << 16) * scale;
    qv[1] = __uint_as_float(r.x & 0xFFFF0000u) * scale;
    qv[2] = __uint_as_float(r.y << 16) * scale;
    qv[3] = __uint_as_float(r.y & 0xFFFF0000u) * scale;
  }
  // The walk is latency-bound (a few KB per CTA), so it is organised for memory-level parallelism: a block of 32 cache
  // positions per trip, lane j owns position jb + j for q.K (its whole 256-byte K row: 16 independent 16-byte loads, no
  // shuffles), the V rows of the block (one coalesced 8-byte load per lane per position) are requested BEFORE the scores
  // are reduced, and P.V broadcasts the probabilities lane by lane.  One DRAM round trip per 32 positions instead of
  // one per 4.
  __shared__ __align__(16) float q_s[kGroup][hd];
  *reinterpret_cast<float4*>(&q_s[warp][lane * 4]) = make_float4(qv[0], qv[1], qv[2], qv[3]);
  __syncwarp();
  float m = -INFINITY, l = 0.f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int jb = j0; jb < j1; jb += 32) {
    const int cnt = min(32, j1 - jb);
    const bool valid = lane < cnt;
    uint4 kk[16];
    {
      const uint4* krow = reinterpret_cast<const uint4*>(kb - lane * 4 + (int64_t)(jb + (valid ? lane : 0)) * hd);
#pragma unroll
      for (int i = 0; i < 16; ++i) kk[i] = krow[i];
    }
    uint2 vr[32];
#pragma unroll
    for (int jj = 0; jj < 32; ++jj)
      vr[jj] = jj < cnt ? *reinterpret_cast<const uint2*>(vb + (int64_t)(jb + jj) * hd) : make_uint2(0u, 0u);
    float sc = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float4 qa = *reinterpret_cast<const float4*>(&q_s[warp][8 * i]);
      const float4 qb = *reinterpret_cast<const float4*>(&q_s[warp][8 * i + 4]);
      sc += qa.x * __uint_as_float(kk[i].x << 16) + qa.y * __uint_as_float(kk[i].x & 0xFFFF0000u) +
            qa.z * __uint_as_float(kk[i].y << 16) + qa.w * __uint_as_float(kk[i].y & 0xFFFF0000u) +
            qb.x * __uint_as_float(kk[i].z << 16) + qb.y * __uint_as_float(kk[i].z & 0xFFFF0000u) +
            qb.z * __uint_as_float(kk[i].w << 16) + qb.w * __uint_as_float(kk[i].w & 0xFFFF0000u);
    }
    if (!valid) sc = -INFINITY;
    float mb = sc;
#pragma unroll
    for (int o = 16; o; o >>= 1) mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
    const float m_new = fmaxf(m, mb);  // finite: the block holds at least one position
    const float corr = __expf(m - m_new);
    const float pj = valid ? __expf(sc - m_new) : 0.f;
    float ps = pj;
#pragma unroll
    for (int o = 16; o; o >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, o);
    l = l * corr + ps;
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[e] *= corr;
#pragma unroll
    for (int jj = 0; jj < 32; ++jj) {
      const float pw = __shfl_sync(0xffffffffu, pj, jj);  // 0 beyond cnt
      acc[0] += pw * __uint_as_float(vr[jj].x << 16);
      acc[1] += pw * __uint_as_float(vr[jj].x & 0xFFFF0000u);
      acc[2] += pw * __uint_as_float(vr[jj].y << 16);
      acc[3] += pw * __uint_as_float(vr[jj].y & 0xFFFF0000u);
    }
    m = m_new;
  }
  if (n_split == 1) {
    const float inv = 1.f / l;
    uint2 o;
    o.x = float_to_bf16_bits(acc[0] * inv) | (float_to_bf16_bits(acc[1] * inv) << 16);
    o.y = float_to_bf16_bits(acc[2] * inv) | (float_to_bf16_bits(acc[3] * inv) << 16);
    *reinterpret_cast<uint2*>(out + ((int64_t)b * H + h) * hd + lane * 4) = o;
    return;
  }
  {
    const int64_t slot = ((int64_t)b * H + h) * n_split + split;
    *reinterpret_cast<float4*>(part_o + slot * hd + lane * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    if (lane == 0) {
      part_ml[slot * 2] = m;
      part_ml[slot * 2 + 1] = l;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(&tickets[b * Hkv + kvh], 1u);
    s_last = (t == (unsigned int)n_split - 1);
    if (s_last) tickets[b * Hkv + kvh] = 0;  // every split has arrived: ready for the next launch
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // merge the n_split partials of this warp's q head
  const int64_t bh = (int64_t)b * H + h;
  float ll = 0.f, a[4] = {0.f, 0.f, 0.f, 0.f};
  if (n_split <= 32) {
    // lane s owns split s's (m, l): one coalesced load, the maximum and the normaliser are warp reductions, the weights
    // exp(m_s - max) are computed once in parallel; the accumulation loop is then independent loads + broadcasts
    float ms = -INFINITY, ls = 0.f;
    if (lane < n_split) {
      const float2 ml = __ldcg(reinterpret_cast<const float2*>(part_ml + (bh * n_split + lane) * 2));
      ms = ml.x;
      ls = ml.y;
    }
    float mm = ms;
#pragma unroll
    for (int o = 16; o; o >>= 1) mm = fmaxf(mm, __shfl_xor_sync(0xffffffffu, mm, o));
    const float w = (ls > 0.f) ? __expf(ms - mm) : 0.f;
    ll = ls * w;
#pragma unroll
    for (int o = 16; o; o >>= 1) ll += __shfl_xor_sync(0xffffffffu, ll, o);
    const float4* pp = reinterpret_cast<const float4*>(part_o + bh * n_split * hd + lane * 4);
#pragma unroll 8
    for (int s = 0; s < n_split; ++s) {
      const float4 po = __ldcg(pp + s * (hd / 4));
      const float ws = __shfl_sync(0xffffffffu, w, s);
      a[0] += po.x * ws; a[1] += po.y * ws; a[2] += po.z * ws; a[3] += po.w * ws;
    }
  } else {
    float mm = -INFINITY;
    for (int s = 0; s < n_split; ++s) mm = fmaxf(mm, __ldcg(part_ml + (bh * n_split + s) * 2));
    for (int s = 0; s < n_split; ++s) {
      const float ms = __ldcg(part_ml + (bh * n_split + s) * 2), ls = __ldcg(part_ml + (bh * n_split + s) * 2 + 1);
      const float w = (ls > 0.f) ? __expf(ms - mm) : 0.f;
      const float4 po = __ldcg(reinterpret_cast<const float4*>(part_o + (bh * n_split + s) * hd + lane * 4));
      ll += ls * w;
      a[0] += po.x * w; a[1] += po.y * w; a[2] += po.z * w; a[3] += po.w * w;
    }
  }
  const float inv = 1.f / ll;
  uint2 o;
  o.x = float_to_bf16_bits(a[0] * inv) | (float_to_bf16_bits(a[1] * inv) << 16);
  o.y = float_to_bf16_bits(a[2] * inv) | (float_to_bf16_bits(a[3] * inv) << 16);
  *reinterpret_cast<uint2*>(out + bh * hd + lane * 4) = o;
}

// out = bf16( bf16(silu(gate)) * up ), gate_up = [B][2I] (rows of w1 then rows of w3) — model.py:486
__global__ void silu_mul_kernel(const __nv_bfloat16* __restrict__ gate_up, __nv_bfloat16* __restrict__ out, int64_t B,
                                int64_t I) {
  const int64_t total = B * I / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = (i * 8) / I, c = (i * 8) % I;
    float g[8], u[8];
    unpack8(*reinterpret_cast<const uint4*>(gate_up + b * 2 * I + c), g);
    unpack8(*reinterpret_cast<const uint4*>(gate_up + b * 2 * I + I + c), u);
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a0 = bf16r(g[2 * e] / (1.f + __expf(-g[2 * e]))) * u[2 * e];
      const float a1 = bf16r(g[2 * e + 1] / (1.f + __expf(-g[2 * e + 1]))) * u[2 * e + 1];
      o[e] = float_to_bf16_bits(a0) | (float_to_bf16_bits(a1) << 16);
    }
    *reinterpret_cast<uint4*>(out + b * I + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace ao

using namespace ao;

extern "C" int ao_b200_rmsnorm(const void* x, const void* residual, const void* weight, void* h_out, void* y,
                               int64_t rows, int64_t dim, float eps, void* stream) {
  AO_CHECK_ARG(x && weight && y && rows > 0 && dim > 0 && dim % 8 == 0 && dim * 4 <= 200 * 1024);
  AO_CHECK_ARG(ao_aligned(x, 16) && ao_aligned(weight, 16) && ao_aligned(y, 16));
  AO_CHECK_ARG((!residual || ao_aligned(residual, 16)) && (!h_out || ao_aligned(h_out, 16)));
  const size_t smem = dim * sizeof(float);
  static bool attr_set = false;
  if (!attr_set && smem > 48 * 1024) {
    cudaFuncSetAttribute(rmsnorm_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  rmsnorm_kernel<512><<<static_cast<unsigned>(rows), 512, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(residual),
      static_cast<const __nv_bfloat16*>(weight), static_cast<__nv_bfloat16*>(h_out), static_cast<__nv_bfloat16*>(y),
      dim, eps);
  return ao_launch_status();
}

extern "C" int ao_b200_rope_kv_append(const void* qkv, const void* freqs_cis, const int32_t* input_pos, void* q_out,
                                      void* k_cache, void* v_cache, int B, int H, int Hkv, int head_dim, int S,
                                      void* stream) {
  AO_CHECK_ARG(qkv && freqs_cis && input_pos && q_out && k_cache && v_cache);
  AO_CHECK_ARG(B > 0 && H > 0 && Hkv > 0 && head_dim > 0 && head_dim % 2 == 0 && S > 0);
  const int64_t total = (int64_t)B * (H + 2 * Hkv) * (head_dim / 2);
  const int grid = static_cast<int>((total + 255) / 256);
  rope_kv_append_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(qkv), static_cast<const __nv_bfloat16*>(freqs_cis), input_pos,
      static_cast<__nv_bfloat16*>(q_out), static_cast<__nv_bfloat16*>(k_cache), static_cast<__nv_bfloat16*>(v_cache),
      B, H, Hkv, head_dim, S);
  return ao_launch_status();
}

extern "C" int64_t ao_b200_decode_attention_workspace_bytes(int B, int H, int n_split) {
  return n_split <= 1 ? 0 : (int64_t)B * H * n_split * (128 + 2) * sizeof(float);
}

extern "C" int ao_b200_decode_attention(const void* q, const void* k_cache, const void* v_cache,
                                        const int32_t* input_pos, void* out, int B, int H, int Hkv, int head_dim,
                                        int S, float scale, int n_split, void* workspace, int64_t workspace_bytes,
                                        void* stream) {
  AO_CHECK_ARG(q && k_cache && v_cache && input_pos && out && B > 0 && H > 0 && Hkv > 0 && S > 0 && n_split >= 1);
  if (head_dim != 128 || H % Hkv != 0) return AO_B200_ERR_UNSUPPORTED;
  const int group = H / Hkv;
  if (workspace_bytes < ao_b200_decode_attention_workspace_bytes(B, H, n_split)) return AO_B200_ERR_WORKSPACE;
  float* part_o = static_cast<float*>(workspace);
  float* part_ml = part_o ? part_o + (int64_t)B * H * n_split * 128 : nullptr;
  auto s = static_cast<cudaStream_t>(stream);
  dim3 grid(Hkv, B, n_split);
  auto qp = static_cast<const __nv_bfloat16*>(q);
  auto kp = static_cast<const __nv_bfloat16*>(k_cache);
  auto vp = static_cast<const __nv_bfloat16*>(v_cache);
  auto op = static_cast<__nv_bfloat16*>(out);
#define AO_ATTN(G) \
  decode_attn_kernel<G><<<grid, G * 32, 0, s>>>(qp, kp, vp, input_pos, op, part_o, part_ml, H, Hkv, S, scale, n_split)
  switch (group) {
    case 1: AO_ATTN(1); break;
    case 2: AO_ATTN(2); break;
    case 4: AO_ATTN(4); break;
    case 8: AO_ATTN(8); break;
    default: return AO_B200_ERR_UNSUPPORTED;
  }
#undef AO_ATTN
  if (n_split > 1)
    decode_attn_merge_kernel<<<B * H, 128, 0, s>>>(part_o, part_ml, op, n_split);
  return ao_launch_status();
}

extern "C" int64_t ao_b200_decode_attention_fused_workspace_bytes(int B, int H, int Hkv, int n_split) {
  return (int64_t)B * H * n_split * (128 + 2) * sizeof(float) + (int64_t)B * Hkv * sizeof(unsigned int);
}

extern "C" int ao_b200_decode_attention_fused(const void* qkv, const void* freqs_cis, const int32_t* input_pos,
                                              void* k_cache, void* v_cache, void* out, int B, int H, int Hkv,
                                              int head_dim, int S, float scale, int n_split, void* workspace,
                                              int64_t workspace_bytes, void* stream) {
  AO_CHECK_ARG(qkv && freqs_cis && input_pos && k_cache && v_cache && out && workspace);
  AO_CHECK_ARG(B > 0 && H > 0 && Hkv > 0 && S > 0 && n_split >= 1 && n_split <= 65535);
  if (head_dim != 128 || H % Hkv != 0) return AO_B200_ERR_UNSUPPORTED;
  if (workspace_bytes < ao_b200_decode_attention_fused_workspace_bytes(B, H, Hkv, n_split)) return AO_B200_ERR_WORKSPACE;
  const int group = H / Hkv;
  float* part_o = static_cast<float*>(workspace);
  float* part_ml = part_o + (int64_t)B * H * n_split * 128;
  unsigned int* tickets = reinterpret_cast<unsigned int*>(part_ml + (int64_t)B * H * n_split * 2);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(Hkv, B, n_split);
  cfg.blockDim = dim3(group * 32);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  auto qp = static_cast<const __nv_bfloat16*>(qkv);
  auto fp = static_cast<const __nv_bfloat16*>(freqs_cis);
  auto kp = static_cast<__nv_bfloat16*>(k_cache);
  auto vp = static_cast<__nv_bfloat16*>(v_cache);
  auto op = static_cast<__nv_bfloat16*>(out);
  cudaError_t e;
#define AO_ATTN(G)                                                                                                  \
  e = cudaLaunchKernelEx(&cfg, decode_attn_fused_kernel<G>, qp, fp, input_pos, kp, vp, op, part_o, part_ml, tickets, H, \
                         Hkv, S, scale, n_split)
  switch (group) {
    case 1: AO_ATTN(1); break;
    case 2: AO_ATTN(2); break;
    case 4: AO_ATTN(4); break;
    case 8: AO_ATTN(8); break;
    default: return AO_B200_ERR_UNSUPPORTED;
  }
#undef AO_ATTN
  if (e != cudaSuccess) return AO_B200_ERR_CUDA;
  return ao_launch_status();
}

extern "C" int ao_b200_silu_mul(const void* gate_up, void* out, int64_t B, int64_t I, void* stream) {
  AO_CHECK_ARG(gate_up && out && B > 0 && I > 0 && I % 8 == 0 && ao_aligned(gate_up, 16) && ao_aligned(out, 16));
  const int64_t total = B * I / 8;
  const int grid = static_cast<int>((total + 255) / 256);
  silu_mul_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(gate_up),
                                                                      static_cast<__nv_bfloat16*>(out), B, I);
  return ao_launch_status();
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant of ao_b200_decode_attention: the split partials live in g_attn_static_scratch.
// ---------------------------------------------------------------------------------------
// Device address of the static scratch on the CURRENT device: resolved once per device (cudaGetSymbolAddress with that
// device current), then cached.  No allocation anywhere on this path.
static int static_scratch_ptr(void** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static void* ptrs[kMaxDevices] = {};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return AO_B200_ERR_CUDA;
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, ao::g_attn_static_scratch) != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return AO_B200_ERR_CUDA;
    }
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return AO_B200_OK;
}

extern "C" int ao_b200_decode_attention_static(const void* q, const void* k_cache, const void* v_cache,
                                               const int32_t* input_pos, void* out, int B, int H, int Hkv, int head_dim,
                                               int S, float scale, int n_split, void* stream) {
  AO_CHECK_ARG(q && k_cache && v_cache && input_pos && out && B > 0 && H > 0 && Hkv > 0 && S > 0 && n_split >= 1);
  // the partials of every (batch row, head, split) must fit the fixed array: reject, never truncate
  if (ao_b200_decode_attention_workspace_bytes(B, H, n_split) > (int64_t)AO_B200_STATIC_SCRATCH_BYTES)
    return AO_B200_ERR_STATIC_SCRATCH;
  void* scratch = nullptr;
  const int rc = static_scratch_ptr(&scratch);
  if (rc != AO_B200_OK) return rc;
  // every partial slot that the merge reads is written by the split kernel of the same call: nothing has to start at zero
  return ao_b200_decode_attention(q, k_cache, v_cache, input_pos, out, B, H, Hkv, head_dim, S, scale, n_split, scratch,
                                  (int64_t)AO_B200_STATIC_SCRATCH_BYTES, stream);
}
