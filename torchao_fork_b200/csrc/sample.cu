// Token sampling for the decode step: temperature, top-k, softmax, exponential-noise argmax in ONE launch.
// Mirrors torchao/_models/llama/generate.py:79-100 (logits_to_probs + multinomial_sample_one_no_sync):
//     lg = logits.float() / max(T, 1e-5);  v = topk(lg, k);  lg[lg < v[-1]] = -inf
//     probs = softmax(lg);  next = argmax(probs / q),  q ~ Exp(1)  (caller supplies q: torch's generator stays the RNG)
// One 8-CTA cluster per batch row.  Logits are bf16, so the k-th largest value is found EXACTLY by a two-level histogram over the
// 16-bit order-preserving key (high byte, then low byte inside the selected bin): two passes over the row instead of
// a 128K-element sort.  Ties at the k-th value are all kept, like the reference's `lg < pivot` mask.
#include <cooperative_groups.h>

#include <cfloat>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace ao {
namespace {

constexpr int kSampThreads = 1024;

// order-preserving 16-bit key of a bf16 bit pattern (ascending key = ascending value; -0 < +0)
__device__ __forceinline__ uint32_t bf16_key(uint32_t b) { return (b & 0x8000u) ? (~b & 0xFFFFu) : (b | 0x8000u); }
__device__ __forceinline__ float key_to_float(uint32_t key) {
  const uint32_t b = (key & 0x8000u) ? (key & 0x7FFFu) : (~key & 0xFFFFu);
  return __uint_as_float(b << 16);
}

// warp-aggregated histogram increment: lanes with equal `bin` elect one leader that adds the group's size
__device__ __forceinline__ void hist_add(uint32_t* hist, uint32_t bin, bool valid) {
  const uint32_t act = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  const uint32_t peers = __match_any_sync(act, bin);
  if ((peers & ((1u << (threadIdx.x & 31)) - 1)) == 0) atomicAdd(&hist[bin], (uint32_t)__popc(peers));
}

// one warp: find the bin (scanning from the top) in which the descending cumulative count reaches k;
// *out_k = rank of the wanted element inside that bin (1-based, from the bin's top)
__device__ void select_bin(const uint32_t* hist, uint32_t k, uint32_t* out_bin, uint32_t* out_k) {
  const int lane = threadIdx.x & 31;  // lane l owns bins [8l, 8l + 8)
  uint32_t c[8], tot = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    c[i] = hist[8 * lane + i];
    tot += c[i];
  }
  uint32_t above = 0;  // elements in the bins of higher lanes
  for (int l = 31; l > 0; --l) {
    const uint32_t t = __shfl_sync(0xffffffffu, tot, l);
    if (l > lane) above += t;
  }
  if (above < k && above + tot >= k) {
    uint32_t kk = k - above;
#pragma unroll
    for (int i = 7; i >= 0; --i) {
      if (c[i] >= kk) {
        *out_bin = 8 * lane + i;
        *out_k = kk;
        break;
      }
      kk -= c[i];
    }
  }
}

template <typename T, typename Op>
__device__ T block_reduce(T v, Op op, T* scratch /* [32] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  T r = scratch[lane < (kSampThreads / 32) ? lane : 0];
#pragma unroll
  for (int o = 16; o; o >>= 1) r = op(r, __shfl_xor_sync(0xffffffffu, r, o));
  __syncthreads();
  return r;
}

// A cluster of kSampCluster CTAs per row: each CTA keeps its slice of the row in shared memory as 16-bit keys (one
// global read of the logits in total), partial histograms / sums / argmax candidates are exchanged through
// distributed shared memory (every CTA reads the kSampCluster partials and merges them redundantly).
constexpr int kSampCluster = 8;

__global__ void __launch_bounds__(kSampThreads) sample_kernel(const __nv_bfloat16* __restrict__ logits, int64_t ldl,
                                                              int V, int slice, float inv_temp_div, int top_k,
                                                              const float* __restrict__ noise, int64_t ldq,
                                                              int64_t* __restrict__ next_tok) {
  extern __shared__ __align__(16) unsigned short keys[];  // [slice]
  __shared__ uint32_t hist1[256], hist2[256], merged[256];
  __shared__ uint32_t sel_bin, sel_k;
  __shared__ uint32_t part_max;
  __shared__ float part_sum;
  __shared__ unsigned long long part_best;
  __shared__ float red_f[32];
  __shared__ unsigned long long red_u[32];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int row = blockIdx.y;
  const int tid = threadIdx.x;
  const float T = inv_temp_div;  // divisor max(temperature, 1e-5); <= 0 means greedy
  const int i0 = rank * slice, n = max(0, min(slice, V - i0));
  const bool use_topk = (T > 0.f) && top_k > 0 && top_k < V;
  if (tid < 256) {
    hist1[tid] = 0;
    hist2[tid] = 0;
  }
  grid_dependency_wait();
  const unsigned short* lg = reinterpret_cast<const unsigned short*>(logits + (int64_t)row * ldl) + i0;
  __syncthreads();

  // ---- pass 1 (the only global read of the logits): keys -> shared, high-byte histogram, max key ----
  uint32_t kmax = 0;
  for (int base = 0; base < n; base += kSampThreads) {
    const int i = base + tid;
    const bool valid = i < n;
    const uint32_t key = valid ? bf16_key(__ldg(lg + i)) : 0u;
    if (valid) keys[i] = (unsigned short)key;
    kmax = max(kmax, key);
    if (use_topk) hist_add(hist1, key >> 8, valid);
  }
  kmax = block_reduce<uint32_t>(kmax, [](uint32_t a, uint32_t b) { return a > b ? a : b; }, reinterpret_cast<uint32_t*>(red_f));
  if (tid == 0) part_max = kmax;
  cluster.sync();
#pragma unroll
  for (int r = 0; r < kSampCluster; ++r) kmax = max(kmax, *cluster.map_shared_rank(&part_max, r));

  if (T <= 0.f) {
    // greedy: smallest index holding the maximum
    int best = 0x7fffffff;
    for (int i = tid; i < n; i += kSampThreads)
      if (keys[i] == kmax) best = min(best, i0 + i);
    best = block_reduce<int>(best, [](int a, int b) { return a < b ? a : b; }, reinterpret_cast<int*>(red_f));
    if (tid == 0) part_best = (unsigned long long)(uint32_t)best;
    cluster.sync();
    if (rank == 0 && tid == 0) {
      unsigned long long b = part_best;
      for (int r = 1; r < kSampCluster; ++r) b = min(b, *cluster.map_shared_rank(&part_best, r));
      next_tok[row] = (int64_t)b;
    }
    cluster.sync();
    return;
  }

  uint32_t thr_key = 0;  // keep everything
  if (use_topk) {
    if (tid < 256) {
      uint32_t c = 0;
#pragma unroll
      for (int r = 0; r < kSampCluster; ++r) c += cluster.map_shared_rank(hist1, r)[tid];
      merged[tid] = c;
    }
    __syncthreads();
    if (tid < 32) select_bin(merged, (uint32_t)top_k, &sel_bin, &sel_k);
    __syncthreads();
    const uint32_t hb = sel_bin, k2 = sel_k;
    // ---- pass 2: low-byte histogram inside the selected bin ----
    for (int base = 0; base < n; base += kSampThreads) {
      const int i = base + tid;
      const uint32_t key = i < n ? keys[i] : 0u;
      hist_add(hist2, key & 0xFFu, i < n && (key >> 8) == hb);
    }
    cluster.sync();
    if (tid < 256) {
      uint32_t c = 0;
#pragma unroll
      for (int r = 0; r < kSampCluster; ++r) c += cluster.map_shared_rank(hist2, r)[tid];
      merged[tid] = c;
    }
    __syncthreads();
    if (tid < 32) select_bin(merged, k2, &sel_bin, &sel_k);
    __syncthreads();
    thr_key = (hb << 8) | sel_bin;
  }
  // ---- pass 3: softmax denominator over the kept entries ----
  const float mx = key_to_float(kmax) / T;
  float sum = 0.f;
  for (int i = tid; i < n; i += kSampThreads) {
    const uint32_t key = keys[i];
    if (key >= thr_key) sum += expf(key_to_float(key) / T - mx);
  }
  sum = block_reduce<float>(sum, [](float a, float b) { return a + b; }, red_f);
  if (tid == 0) part_sum = sum;
  cluster.sync();
  sum = 0.f;
#pragma unroll
  for (int r = 0; r < kSampCluster; ++r) sum += *cluster.map_shared_rank(&part_sum, r);
  // ---- pass 4: argmax of probs / q (first index on ties) ----
  const float* q = noise + (int64_t)row * ldq + i0;
  float bv = -1.f;
  int bi = 0x7fffffff;
  for (int i = tid; i < n; i += kSampThreads) {
    const uint32_t key = keys[i];
    if (key >= thr_key) {
      const float pr = expf(key_to_float(key) / T - mx) / sum;
      const float sc = pr / __ldg(q + i);
      if (sc > bv) {
        bv = sc;
        bi = i0 + i;
      }
    }
  }
  // pack (score, ~index) so that a max-reduce picks the largest score and, on ties, the smallest index
  unsigned long long pk = ((unsigned long long)__float_as_uint(fmaxf(bv, 0.f)) << 32) | (uint32_t)(0x7fffffff - bi);
  if (bv < 0.f) pk = 0ull;
  pk = block_reduce<unsigned long long>(pk, [](unsigned long long a, unsigned long long b) { return a > b ? a : b; }, red_u);
  if (tid == 0) part_best = pk;
  cluster.sync();
  if (rank == 0 && tid == 0) {
    unsigned long long b = part_best;
    for (int r = 1; r < kSampCluster; ++r) b = max(b, *cluster.map_shared_rank(&part_best, r));
    next_tok[row] = 0x7fffffff - (int)(b & 0xffffffffull);
  }
  cluster.sync();  // nobody leaves while its shared memory may still be read
}

}  // namespace
}  // namespace ao

extern "C" int ao_b200_sample(const void* logits, int64_t ldl, int64_t B, int64_t V, float temperature, int top_k,
                              const void* noise, int64_t ldq, void* next_tok, void* stream) {
  AO_CHECK_ARG(logits != nullptr && next_tok != nullptr && B >= 1 && B <= 65535 && V >= 1 && ldl >= V);
  AO_CHECK_ARG(temperature == 0.f || (noise != nullptr && ldq >= V));
  const int slice = (int)((V + ao::kSampCluster - 1) / ao::kSampCluster);
  const size_t smem = (size_t)slice * 2;
  if (smem > 200 * 1024) return AO_B200_ERR_UNSUPPORTED;  // V > ~800K
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(ao::sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess)
      return AO_B200_ERR_CUDA;
    attr_done = true;
  }
  const float div = temperature == 0.f ? 0.f : (temperature > 1e-5f ? temperature : 1e-5f);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(ao::kSampCluster, (unsigned)B);
  cfg.blockDim = dim3(ao::kSampThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = ao::kSampCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, ao::sample_kernel, static_cast<const __nv_bfloat16*>(logits), ldl, (int)V, slice, div, top_k,
                         static_cast<const float*>(noise), ldq, static_cast<int64_t*>(next_tok)) != cudaSuccess)
    return AO_B200_ERR_CUDA;
  return ao_launch_status();
}
