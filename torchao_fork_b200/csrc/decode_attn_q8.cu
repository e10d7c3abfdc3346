// Attention.forward for one new token over an int8 KV cache: the decode-side of
// AffineQuantizedKVCache (torchao/_models/llama/model.py:198-240).  One launch does
//   RoPE on q and the new k (model.py:547-559),
//   update(): q_k, k_scale = _quantize_activation_per_token_absmax(k) (torchao/quantization/utils.py:152-182) stored at
//             *input_pos — likewise v —
//   k_out = k_cache * k_cache_scale with k_out[pos] = k (the new token's row is used UNQUANTIZED, model.py:222-225),
//   GQA softmax(q k_out^T * scale) v_out over positions 0..pos, split n_split ways, partials merged by the last CTA.
// Arithmetic of the quantizer, as torch evaluates it on CUDA for bf16 rows (every op rounds to bf16):
//   scale = clamp(bf16(amax * (1 / 127)), min = 1e-5)      [division by a host scalar = multiplication by its fp32 reciprocal]
//   q     = clamp(rint(bf16(x * bf16(1 / scale))), -127, 127)
//   k_out = bf16(q * scale)
// The cache is half the bytes of the bf16 cache: the walk streams 128 + 2 bytes per position per kv head and tensor
// instead of 256.  Same grid / walk organisation as decode_attn_fused_kernel (decode_ops.cu).
#include "common.cuh"

namespace ao {

__device__ __forceinline__ float bf16_to_f(__nv_bfloat16 v) { return __bfloat162float(v); }

// 4 int8 (one 32-bit word, little endian) times a row scale, each product rounded to bf16
__device__ __forceinline__ void deq4(uint32_t w, float s, float (&o)[4]) {
  o[0] = bf16r((float)(int8_t)(w & 0xff) * s);
  o[1] = bf16r((float)(int8_t)((w >> 8) & 0xff) * s);
  o[2] = bf16r((float)(int8_t)((w >> 16) & 0xff) * s);
  o[3] = bf16r((float)(int8_t)(w >> 24) * s);
}

// One row of 128 bf16 values held 4 per lane -> int8 x 4 (packed) and the row's bf16 scale (all lanes).
__device__ __forceinline__ uint32_t quantize_row_absmax(const float (&x)[4], float& scale_out) {
  float a = fmaxf(fmaxf(fabsf(x[0]), fabsf(x[1])), fmaxf(fabsf(x[2]), fabsf(x[3])));
#pragma unroll
  for (int o = 16; o; o >>= 1) a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, o));
  const float inv127 = 1.0f / 127.0f;
  float scale = bf16r(a * inv127);
  scale = bf16r(fmaxf(scale, 1e-5f));
  const float recip = bf16r(1.0f / scale);
  uint32_t w = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float q = rintf(bf16r(x[e] * recip));
    q = fminf(fmaxf(q, -127.f), 127.f);
    w |= ((uint32_t)(int)q & 0xffu) << (8 * e);
  }
  scale_out = scale;
  return w;
}

template <int kGroup>
__global__ void __launch_bounds__(kGroup * 32) decode_attn_fused_q8_kernel(
    const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ freqs,
    const int32_t* __restrict__ input_pos, int8_t* __restrict__ k_cache, int8_t* __restrict__ v_cache,
    __nv_bfloat16* __restrict__ k_scale, __nv_bfloat16* __restrict__ v_scale, __nv_bfloat16* __restrict__ out,
    float* __restrict__ part_o, float* __restrict__ part_ml, unsigned int* __restrict__ tickets, int H, int Hkv, int S,
    float scale, int n_split) {
  constexpr int hd = 128;
  __shared__ int s_last;
  __shared__ __align__(16) float q_s[kGroup][hd];
  __shared__ __align__(16) float kv_new[2][hd];  // the new token's k (roped) and v as the reference uses them: unquantized
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

  const uint2 cs = *reinterpret_cast<const uint2*>(freqs + ((int64_t)pos * (hd / 2) + 2 * lane) * 2);
  const float c0 = __uint_as_float(cs.x << 16), s0 = __uint_as_float(cs.x & 0xFFFF0000u);
  const float c1 = __uint_as_float(cs.y << 16), s1 = __uint_as_float(cs.y & 0xFFFF0000u);
  auto rope4 = [&](uint2 raw, float (&o)[4]) {
    const float x0 = __uint_as_float(raw.x << 16), x1 = __uint_as_float(raw.x & 0xFFFF0000u);
    const float x2 = __uint_as_float(raw.y << 16), x3 = __uint_as_float(raw.y & 0xFFFF0000u);
    o[0] = bf16r(x0 * c0 - x1 * s0);
    o[1] = bf16r(x1 * c0 + x0 * s0);
    o[2] = bf16r(x2 * c1 - x3 * s1);
    o[3] = bf16r(x3 * c1 + x2 * s1);
  };
  const int64_t row0 = ((int64_t)b * Hkv + kvh) * S;
  const int8_t* kb = k_cache + row0 * hd;
  const int8_t* vb = v_cache + row0 * hd;
  const __nv_bfloat16* ksb = k_scale + row0;
  const __nv_bfloat16* vsb = v_scale + row0;
  const bool owns_new = j0 <= pos && pos < j1;  // block-uniform
  if (owns_new) {
    if (warp == 0) {
      const uint2 kraw = *reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + H + kvh) * hd + lane * 4);
      const uint2 vraw = *reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + H + Hkv + kvh) * hd + lane * 4);
      float kr[4], vr4[4];
      rope4(kraw, kr);
      vr4[0] = __uint_as_float(vraw.x << 16); vr4[1] = __uint_as_float(vraw.x & 0xFFFF0000u);
      vr4[2] = __uint_as_float(vraw.y << 16); vr4[3] = __uint_as_float(vraw.y & 0xFFFF0000u);
      float ks, vs;
      const uint32_t kq = quantize_row_absmax(kr, ks);
      const uint32_t vq = quantize_row_absmax(vr4, vs);
      *reinterpret_cast<uint32_t*>(k_cache + (row0 + pos) * hd + lane * 4) = kq;
      *reinterpret_cast<uint32_t*>(v_cache + (row0 + pos) * hd + lane * 4) = vq;
      if (lane == 0) {
        k_scale[row0 + pos] = __float2bfloat16_rn(ks);
        v_scale[row0 + pos] = __float2bfloat16_rn(vs);
      }
      *reinterpret_cast<float4*>(&kv_new[0][lane * 4]) = make_float4(kr[0], kr[1], kr[2], kr[3]);
      *reinterpret_cast<float4*>(&kv_new[1][lane * 4]) = make_float4(vr4[0], vr4[1], vr4[2], vr4[3]);
    }
    __syncthreads();
  }
  {
    float qr[4];
    rope4(*reinterpret_cast<const uint2*>(qkv + ((int64_t)b * heads + h) * hd + lane * 4), qr);
    *reinterpret_cast<float4*>(&q_s[warp][lane * 4]) = make_float4(qr[0] * scale, qr[1] * scale, qr[2] * scale, qr[3] * scale);
  }
  __syncwarp();
  float m = -INFINITY, l = 0.f, acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int jb = j0; jb < j1; jb += 32) {
    const int cnt = min(32, j1 - jb);
    const bool valid = lane < cnt;
    const int myj = jb + (valid ? lane : 0);
    // lane j owns position jb + j for q.K: its 128-byte K row (8 independent 16-byte loads) and both row scales
    uint4 kk[8];
    {
      const uint4* krow = reinterpret_cast<const uint4*>(kb + (int64_t)myj * hd);
#pragma unroll
      for (int i = 0; i < 8; ++i) kk[i] = krow[i];
    }
    const float ks = bf16_to_f(ksb[myj]);
    const float vs = bf16_to_f(vsb[myj]);
    uint32_t vr[32];  // this lane's 4 dims of the V rows of the block
#pragma unroll
    for (int jj = 0; jj < 32; ++jj)
      vr[jj] = jj < cnt ? *reinterpret_cast<const uint32_t*>(vb + (int64_t)(jb + jj) * hd + lane * 4) : 0u;
    float sc = 0.f;
    if (owns_new && myj == pos && valid) {
#pragma unroll 8
      for (int i = 0; i < hd; i += 4) {
        const float4 qa = *reinterpret_cast<const float4*>(&q_s[warp][i]);
        const float4 ka = *reinterpret_cast<const float4*>(&kv_new[0][i]);
        sc += qa.x * ka.x + qa.y * ka.y + qa.z * ka.z + qa.w * ka.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t w[4] = {kk[i].x, kk[i].y, kk[i].z, kk[i].w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          float kf[4];
          deq4(w[t], ks, kf);
          const float4 qa = *reinterpret_cast<const float4*>(&q_s[warp][16 * i + 4 * t]);
          sc += qa.x * kf[0] + qa.y * kf[1] + qa.z * kf[2] + qa.w * kf[3];
        }
      }
    }
    if (!valid) sc = -INFINITY;
    float mb = sc;
#pragma unroll
    for (int o = 16; o; o >>= 1) mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, o));
    const float m_new = fmaxf(m, mb);
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
      const float vsj = __shfl_sync(0xffffffffu, vs, jj);
      float vf[4];
      if (owns_new && jb + jj == pos) {  // warp-uniform
        const float4 va = *reinterpret_cast<const float4*>(&kv_new[1][lane * 4]);
        vf[0] = va.x; vf[1] = va.y; vf[2] = va.z; vf[3] = va.w;
      } else {
        deq4(vr[jj], vsj, vf);
      }
      acc[0] += pw * vf[0];
      acc[1] += pw * vf[1];
      acc[2] += pw * vf[2];
      acc[3] += pw * vf[3];
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
  // merge the n_split partials of this warp's q head (a split that held no position has l == 0)
  const int64_t bh = (int64_t)b * H + h;
  float mm = -INFINITY;
  for (int s = lane; s < n_split; s += 32) mm = fmaxf(mm, __ldcg(part_ml + (bh * n_split + s) * 2));
#pragma unroll
  for (int o = 16; o; o >>= 1) mm = fmaxf(mm, __shfl_xor_sync(0xffffffffu, mm, o));
  float ll = 0.f, a[4] = {0.f, 0.f, 0.f, 0.f};
  for (int s = 0; s < n_split; ++s) {
    const float ms = __ldcg(part_ml + (bh * n_split + s) * 2), ls = __ldcg(part_ml + (bh * n_split + s) * 2 + 1);
    const float w = (ls > 0.f) ? __expf(ms - mm) : 0.f;
    const float4 po = __ldcg(reinterpret_cast<const float4*>(part_o + (bh * n_split + s) * hd + lane * 4));
    ll += ls * w;
    a[0] += po.x * w; a[1] += po.y * w; a[2] += po.z * w; a[3] += po.w * w;
  }
  const float inv = 1.f / ll;
  uint2 o;
  o.x = float_to_bf16_bits(a[0] * inv) | (float_to_bf16_bits(a[1] * inv) << 16);
  o.y = float_to_bf16_bits(a[2] * inv) | (float_to_bf16_bits(a[3] * inv) << 16);
  *reinterpret_cast<uint2*>(out + bh * hd + lane * 4) = o;
}

}  // namespace ao

using namespace ao;

extern "C" int ao_b200_decode_attention_fused_q8(const void* qkv, const void* freqs_cis, const int32_t* input_pos,
                                                 void* k_cache, void* v_cache, void* k_scale, void* v_scale, void* out,
                                                 int B, int H, int Hkv, int head_dim, int S, float scale, int n_split,
                                                 void* workspace, int64_t workspace_bytes, void* stream) {
  AO_CHECK_ARG(qkv && freqs_cis && input_pos && k_cache && v_cache && k_scale && v_scale && out && workspace);
  AO_CHECK_ARG(B > 0 && H > 0 && Hkv > 0 && S > 0 && n_split >= 1 && n_split <= 65535);
  AO_CHECK_ARG(ao_aligned(k_cache, 16) && ao_aligned(v_cache, 16) && ao_aligned(qkv, 16));
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
  auto kp = static_cast<int8_t*>(k_cache);
  auto vp = static_cast<int8_t*>(v_cache);
  auto ksp = static_cast<__nv_bfloat16*>(k_scale);
  auto vsp = static_cast<__nv_bfloat16*>(v_scale);
  auto op = static_cast<__nv_bfloat16*>(out);
  cudaError_t e;
#define AO_ATTN(G)                                                                                                       \
  e = cudaLaunchKernelEx(&cfg, decode_attn_fused_q8_kernel<G>, qp, fp, input_pos, kp, vp, ksp, vsp, op, part_o, part_ml, \
                         tickets, H, Hkv, S, scale, n_split)
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
