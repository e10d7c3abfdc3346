// HQQ (half-quadratic quantization) qparam search for Int4WeightOnlyConfig(use_hqq=True) — SURVEY.md section 8 row N1.
//
// Reference: torchao/quantization/quant_primitives.py:1998-2111 (_choose_qparams_and_quantize_affine_hqq) with
// optimize_weights_proximal_legacy (:1915-1973) and _convert_to_affinequantized_format (:1981-1995), called from
// AffineQuantizedTensor.from_hp_to_intx (affine_quantized_tensor.py:252-288).  The reference runs ~15 eager tensor ops
// per iteration over an fp32 copy of the weight; here an iteration is ONE pass over the bf16 weight (a warp per
// group, everything else in registers), and the early-stop test (a mean over the WHOLE tensor, so the iteration count
// is global) reads per-block partial sums of the error.  Arithmetic: fp32 with explicit single roundings where the
// reference has op boundaries (W * scale, + zero, ...), i.e. the reference's CPU path; its CUDA path runs the same
// iteration in fp16, which this does not imitate.  The result feeds the same packed format as the tinygemm recipe.
#include "common.cuh"

namespace ao {

constexpr int kHqqWarps = 8;

__device__ __forceinline__ float warp_sum_h(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// lane l of the group's warp owns elements l + 32 j, j < G / 32 (coalesced 64-byte rows of bf16)
template <int G>
__device__ __forceinline__ void load_group(const __nv_bfloat16* w, int64_t group, int lane, float (&v)[G / 32]) {
  const __nv_bfloat16* p = w + group * G + lane;
#pragma unroll
  for (int j = 0; j < G / 32; ++j) v[j] = __bfloat162float(p[32 * j]);
}

// scale = (15 / (max - min)).clamp(max = 2e4), zero = round(-min * scale)          (:2052-2060)
template <int G>
__global__ void __launch_bounds__(kHqqWarps * 32) hqq_init_kernel(const __nv_bfloat16* __restrict__ w, int64_t n_groups,
                                                                  float* __restrict__ scale, float* __restrict__ zero) {
  const int lane = threadIdx.x & 31;
  const int64_t group = (int64_t)blockIdx.x * kHqqWarps + (threadIdx.x >> 5);
  if (group >= n_groups) return;
  float v[G / 32];
  load_group<G>(w, group, lane, v);
  float mn = v[0], mx = v[0];
#pragma unroll
  for (int j = 1; j < G / 32; ++j) {
    mn = fminf(mn, v[j]);
    mx = fmaxf(mx, v[j]);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 0) {
    const float s = fminf(__fdiv_rn(15.f, __fsub_rn(mx, mn)), 2e4f);  // constant group: 15 / 0 = inf -> 2e4
    scale[group] = s;
    zero[group] = rintf(__fmul_rn(-mn, s));
  }
}

// One proximal iteration (:1958-1962):
//   W_q = round(W * scale + zero).clamp(0, 15);  W_r = (W_q - zero) / scale
//   W_e = sign(d) * relu(|d| - (1 / beta) * |d| ^ (p - 1)),  d = W - W_r
//   zero <- mean(W_q - (W - W_e) * scale);  err_partial[block] = sum |W - W_r|
template <int G>
__global__ void __launch_bounds__(kHqqWarps * 32) hqq_step_kernel(const __nv_bfloat16* __restrict__ w, int64_t n_groups,
                                                                  const float* __restrict__ scale, float* __restrict__ zero,
                                                                  float inv_beta, float lp_m1, float* __restrict__ err_partial,
                                                                  const int* __restrict__ stop) {
  __shared__ float s_err[kHqqWarps];
  if (stop != nullptr && *reinterpret_cast<const volatile int*>(stop) != 0) return;  // ao_b200_hqq_solve: the early stop has fired
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t group = (int64_t)blockIdx.x * kHqqWarps + warp;
  float err = 0.f;
  if (group < n_groups) {
    float v[G / 32];
    load_group<G>(w, group, lane, v);
    const float s = scale[group], z = zero[group];
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < G / 32; ++j) {
      const float wq = fminf(fmaxf(rintf(__fadd_rn(__fmul_rn(v[j], s), z)), 0.f), 15.f);
      const float wr = __fdiv_rn(__fsub_rn(wq, z), s);
      const float d = __fsub_rn(v[j], wr), ad = fabsf(d);
      float we = 0.f;  // d = 0: pow = inf, relu(-inf) = 0, sign = 0
      if (ad > 0.f) we = copysignf(fmaxf(__fsub_rn(ad, __fmul_rn(inv_beta, powf(ad, lp_m1))), 0.f), d);
      acc += __fsub_rn(wq, __fmul_rn(__fsub_rn(v[j], we), s));
      err += ad;
    }
    acc = warp_sum_h(acc);
    if (lane == 0) zero[group] = __fdiv_rn(acc, (float)G);
  }
  err = warp_sum_h(err);
  if (lane == 0) s_err[warp] = err;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kHqqWarps; ++i) t += s_err[i];  // fixed order: the early-stop decision is reproducible
    err_partial[blockIdx.x] = t;
  }
}

// W_q = round(W * scale + zero).clamp(0, 15) (:1971); scale_ao = 1 / scale (:2078), zero_ao = (8 - zero) * scale_ao
// (:1981-1995), both cast to bf16 (:2101-2102)
template <int G>
__global__ void __launch_bounds__(kHqqWarps * 32) hqq_finalize_kernel(const __nv_bfloat16* __restrict__ w, int64_t n_groups,
                                                                      const float* __restrict__ scale,
                                                                      const float* __restrict__ zero, int32_t* __restrict__ q,
                                                                      __nv_bfloat16* __restrict__ scale_out,
                                                                      __nv_bfloat16* __restrict__ zero_out) {
  const int lane = threadIdx.x & 31;
  const int64_t group = (int64_t)blockIdx.x * kHqqWarps + (threadIdx.x >> 5);
  if (group >= n_groups) return;
  float v[G / 32];
  load_group<G>(w, group, lane, v);
  const float s = scale[group], z = zero[group];
  int32_t* qp = q + group * G + lane;
#pragma unroll
  for (int j = 0; j < G / 32; ++j) qp[32 * j] = (int32_t)fminf(fmaxf(rintf(__fadd_rn(__fmul_rn(v[j], s), z)), 0.f), 15.f);
  if (lane == 0) {
    const float si = __fdiv_rn(1.f, s);
    scale_out[group] = __float2bfloat16_rn(si);
    zero_out[group] = __float2bfloat16_rn(__fmul_rn(__fsub_rn(8.f, z), si));
  }
}

}  // namespace ao

using namespace ao;

#define AO_HQQ_DISPATCH(KERNEL, ...)                                                 \
  switch (group_size) {                                                              \
    case 32: KERNEL<32><<<grid, kHqqWarps * 32, 0, s>>>(__VA_ARGS__); break;         \
    case 64: KERNEL<64><<<grid, kHqqWarps * 32, 0, s>>>(__VA_ARGS__); break;         \
    case 128: KERNEL<128><<<grid, kHqqWarps * 32, 0, s>>>(__VA_ARGS__); break;       \
    case 256: KERNEL<256><<<grid, kHqqWarps * 32, 0, s>>>(__VA_ARGS__); break;       \
    default: return AO_B200_ERR_UNSUPPORTED;                                         \
  }

extern "C" int64_t ao_b200_hqq_blocks(int64_t n_groups) { return (n_groups + kHqqWarps - 1) / kHqqWarps; }

extern "C" int ao_b200_hqq_init(const void* w, int64_t n_groups, int group_size, float* scale, float* zero, void* stream) {
  AO_CHECK_ARG(w && scale && zero && n_groups > 0 && n_groups <= ((int64_t)1 << 31) * kHqqWarps);
  auto s = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)ao_b200_hqq_blocks(n_groups);
  AO_HQQ_DISPATCH(hqq_init_kernel, static_cast<const __nv_bfloat16*>(w), n_groups, scale, zero);
  return ao_launch_status();
}

extern "C" int ao_b200_hqq_step(const void* w, int64_t n_groups, int group_size, const float* scale, float* zero,
                                double beta, double lp_norm, float* err_partial, void* stream) {
  AO_CHECK_ARG(w && scale && zero && err_partial && n_groups > 0 && beta > 0.0);
  auto s = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)ao_b200_hqq_blocks(n_groups);
  const float inv_beta = (float)(1.0 / beta), lp_m1 = (float)(lp_norm - 1.0);
  AO_HQQ_DISPATCH(hqq_step_kernel, static_cast<const __nv_bfloat16*>(w), n_groups, scale, zero, inv_beta, lp_m1, err_partial, nullptr);
  return ao_launch_status();
}

// The reference's global early stop (:1964-1972) decided on the device: current = mean |W - W_r| of the iteration that
// just ran (per-block partial sums added in a fixed order, in double); the first iteration that does not improve on the
// best so far sets `stop`, which makes the remaining step launches return at once (its zero is kept, as in the reference).
// state: [double best][int stop][int iterations run]
__global__ void __launch_bounds__(256) hqq_judge_kernel(const float* __restrict__ err_partial, int64_t n_blocks, double inv_numel,
                                                        double* __restrict__ state) {
  __shared__ double s_sum[256];
  int* flags = reinterpret_cast<int*>(state + 1);
  if (*reinterpret_cast<volatile int*>(flags) != 0) return;
  double t = 0.0;
  for (int64_t i = threadIdx.x; i < n_blocks; i += 256) t += (double)err_partial[i];
  s_sum[threadIdx.x] = t;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if ((int)threadIdx.x < o) s_sum[threadIdx.x] += s_sum[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double current = s_sum[0] * inv_numel;
    flags[1] += 1;
    if (current < state[0]) state[0] = current;
    else flags[0] = 1;
  }
}

__global__ void hqq_state_init_kernel(double* state) {
  state[0] = 1e4;
  reinterpret_cast<int*>(state + 1)[0] = 0;
  reinterpret_cast<int*>(state + 1)[1] = 0;
}

extern "C" int ao_b200_hqq_solve(const void* w, int64_t n_groups, int group_size, const float* scale, float* zero,
                                 double beta, double kappa, double lp_norm, int iters, float* err_partial, void* state,
                                 void* stream) {
  AO_CHECK_ARG(w && scale && zero && err_partial && state && n_groups > 0 && beta > 0.0 && kappa > 0.0 && iters >= 1);
  AO_CHECK_ARG(ao_aligned(state, 8));
  auto s = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)ao_b200_hqq_blocks(n_groups);
  hqq_state_init_kernel<<<1, 1, 0, s>>>(static_cast<double*>(state));  // best = 1e4 (:1956), stop = iterations = 0
  const double inv_numel = 1.0 / ((double)n_groups * (double)group_size);
  const float lp_m1 = (float)(lp_norm - 1.0);
  const int* stop = reinterpret_cast<const int*>(static_cast<const double*>(state) + 1);
  for (int it = 0; it < iters; ++it) {
    const float inv_beta = (float)(1.0 / beta);
    AO_HQQ_DISPATCH(hqq_step_kernel, static_cast<const __nv_bfloat16*>(w), n_groups, scale, zero, inv_beta, lp_m1, err_partial, stop);
    hqq_judge_kernel<<<1, 256, 0, s>>>(err_partial, (int64_t)grid, inv_numel, static_cast<double*>(state));
    beta *= kappa;
  }
  return ao_launch_status();
}

extern "C" int ao_b200_hqq_finalize(const void* w, int64_t n_groups, int group_size, const float* scale, const float* zero,
                                    int32_t* q, void* scale_out, void* zero_out, void* stream) {
  AO_CHECK_ARG(w && scale && zero && q && scale_out && zero_out && n_groups > 0);
  auto s = static_cast<cudaStream_t>(stream);
  const unsigned grid = (unsigned)ao_b200_hqq_blocks(n_groups);
  AO_HQQ_DISPATCH(hqq_finalize_kernel, static_cast<const __nv_bfloat16*>(w), n_groups, scale, zero, q,
                  static_cast<__nv_bfloat16*>(scale_out), static_cast<__nv_bfloat16*>(zero_out));
  return ao_launch_status();
}
