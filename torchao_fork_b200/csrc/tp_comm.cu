// Tensor-parallel exchange for the decode step: ONE kernel per row-parallel projection does the all-reduce of the
// partial sums over NVLink peer memory AND the residual add,  h += sum_r partial_r,  instead of an NCCL all-reduce
// (~17 us for these 8-16 KB messages inside a CUDA graph, measured) followed by an add kernel.
//
// Every rank owns one device buffer (cudaMalloc + CUDA IPC handle, opened by all peers):
//     [header: epoch, done-ticket, status][data: 2 slots x world x max_n / 2 x {2 bf16, tag}]
// The payload travels in 8-byte packets {two bf16, 32-bit epoch tag} written with single 8-byte stores, so a packet
// is either absent (old tag) or complete: no fences, no separate flag, one NVLink one-way latency.  (First version:
// 16-byte data stores + __threadfence_system + release flag + acquire poll = 10.3 us per call; this one: see
// profiles/.)  A call with epoch e (kept in the rank's own header, so a replayed CUDA graph keeps counting) uses slot
// e & 1:
//   1. every thread pushes its 4 packets (8 bf16 of `partial`) into data[slot][my rank] of every peer,
//   2. polls its own data[slot][r] packets until their tag reads e (bounded spin, status word set on time-out),
//   3. sums the `world` values in RANK ORDER in fp32 (own from registers) and adds the residual: every rank computes
//      bit-identical h, so the replicated residual stream cannot drift apart.
// Two slots suffice: a rank can only be one call ahead of the slowest peer (it needs that peer's packets of the
// current call to finish it).  All ranks must issue the same sequence of calls (they replay the same graph).
#include <cuda_bf16.h>

#include <cstring>

#include "common.cuh"

namespace ao {

constexpr int kTpMaxWorld = 8, kTpMaxCtas = 32, kTpThreads = 512;
constexpr int kTpHeaderBytes = 256;
constexpr long long kTpSpinClocks = 4000000000ll;  // ~2 s at 1.9 GHz

struct TpPeers {
  uint8_t* buf[kTpMaxWorld];
};

__device__ __forceinline__ void st_packet(void* p, uint32_t data, uint32_t tag) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(data), "r"(tag) : "memory");
}
__device__ __forceinline__ uint2 ld_packet(const void* p) {
  uint2 r;
  asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}

__global__ void __launch_bounds__(kTpThreads, 1)
tp_allreduce_add_kernel(const __nv_bfloat16* __restrict__ partial, __nv_bfloat16* __restrict__ h, int64_t n, TpPeers peers,
                        int rank, int world, int64_t max_n) {
  grid_dependency_wait();  // `partial` comes from the previous kernel of the stream (PDL)
  uint8_t* mine = peers.buf[rank];
  uint32_t* hdr = reinterpret_cast<uint32_t*>(mine);
  const uint32_t e = *reinterpret_cast<volatile uint32_t*>(hdr) + 1;  // every CTA reads it before the last one bumps it
  const int slot = e & 1;
  const int64_t i = ((int64_t)blockIdx.x * kTpThreads + threadIdx.x) * 8;  // 8 bf16 per thread = 4 packets
  // packet j of this thread (elements i + 2 j, i + 2 j + 1) from rank r sits at
  //   data_off + (r * max_n / 2 + pk0 + 32 j) * 8,  pk0 = (global warp * 4) * 32 + lane:
  // one warp-wide store instruction covers 256 contiguous bytes = two full 128-byte lines on the link (lane-major
  // packets would make it 32 partial sectors)
  // a rank's region is sized for whole warps (256 elements = 128 packets each): a partly active last warp still writes
  // its packets at warp * 128 + lane + 32 j, which must stay inside the region (max_n only has to be a multiple of 8)
  const size_t region_pk = (size_t)((max_n + 255) / 256) * 128;
  const size_t data_off = kTpHeaderBytes + ((size_t)slot * world) * region_pk * 8;
  const size_t pk0 = (size_t)((blockIdx.x * kTpThreads + threadIdx.x) >> 5) * 128 + (threadIdx.x & 31);
  if (i < n) {
    const uint4 v = *reinterpret_cast<const uint4*>(partial + i);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    for (int d = 1; d < world; ++d) {  // staggered so that the ranks do not all hit the same peer first
      const int p = (rank + d) % world;
      uint8_t* dst = peers.buf[p] + data_off + ((size_t)rank * region_pk + pk0) * 8;
#pragma unroll
      for (int j = 0; j < 4; ++j) st_packet(dst + 256 * j, w[j], e);
    }
    const uint4 hv = *reinterpret_cast<const uint4*>(h + i);  // in flight while the packets cross the link
    // poll: ALL packets of ALL peers are requested before any tag is looked at (one L2 round trip per sweep instead of
    // one per packet: 6.4 -> see profiles/ us at world 2, and no growth with the world size)
    uint32_t u[kTpMaxWorld - 1][4];
    const long long t0 = clock64();
    for (;;) {
      bool all = true;
#pragma unroll
      for (int d = 1; d < kTpMaxWorld; ++d) {
        if (d < world) {
          const int r = (rank + d) % world;
          const uint8_t* src = mine + data_off + ((size_t)r * region_pk + pk0) * 8;
          uint2 pk[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) pk[j] = ld_packet(src + 256 * j);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            all = all && (pk[j].y == e);
            u[d - 1][j] = pk[j].x;
          }
        }
      }
      if (all) break;
      if (clock64() - t0 > kTpSpinClocks) {  // a peer never arrived: flag the error, do not hang the GPU
        atomicExch(hdr + 2, 1u);
        break;
      }
    }
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int r = 0; r < world; ++r) {  // rank order, whatever the polling order was
      const int d = (r - rank + world) % world;
      uint32_t q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q[j] = w[j];
#pragma unroll
      for (int dd = 1; dd < kTpMaxWorld; ++dd)
        if (dd == d) {
#pragma unroll
          for (int j = 0; j < 4; ++j) q[j] = u[dd - 1][j];
        }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[2 * j] += bf16_bits_to_float(q[j] & 0xffffu);
        acc[2 * j + 1] += bf16_bits_to_float(q[j] >> 16);
      }
    }
    float hf[8];
    unpack8(hv, hf);
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      o[j] = float_to_bf16_bits(hf[2 * j] + acc[2 * j]) | (float_to_bf16_bits(hf[2 * j + 1] + acc[2 * j + 1]) << 16);
    *reinterpret_cast<uint4*>(h + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  // the last CTA to get here publishes the new epoch (all CTAs have read the old one by now)
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(hdr + 1, 1u);
    if (done == gridDim.x - 1) {
      hdr[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(hdr) = e;
    }
  }
}

}  // namespace ao

extern "C" int64_t ao_b200_tp_buffer_bytes(int world, int64_t max_elems) {
  if (world < 1 || world > ao::kTpMaxWorld || max_elems <= 0 || max_elems % 8 != 0) return 0;
  if (max_elems > (int64_t)ao::kTpMaxCtas * ao::kTpThreads * 8) return 0;
  return ao::kTpHeaderBytes + (int64_t)2 * world * ((max_elems + 255) / 256 * 128) * 8;
}

extern "C" int ao_b200_tp_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_out) {
  AO_CHECK_ARG(bytes > 0 && dev_ptr && ipc_handle_out);
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess) return AO_B200_ERR_CUDA;
  if (cudaMemset(p, 0, (size_t)bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return AO_B200_ERR_CUDA;
  cudaIpcMemHandle_t hnd;
  if (cudaIpcGetMemHandle(&hnd, p) != cudaSuccess) {
    cudaFree(p);
    return AO_B200_ERR_CUDA;
  }
  memcpy(ipc_handle_out, &hnd, sizeof hnd);
  *dev_ptr = p;
  return AO_B200_OK;
}

extern "C" int ao_b200_tp_open(const void* ipc_handle, void** peer_ptr) {
  AO_CHECK_ARG(ipc_handle && peer_ptr);
  cudaIpcMemHandle_t hnd;
  memcpy(&hnd, ipc_handle, sizeof hnd);
  void* p = nullptr;
  if (cudaIpcOpenMemHandle(&p, hnd, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
    cudaGetLastError();
    return AO_B200_ERR_CUDA;
  }
  *peer_ptr = p;
  return AO_B200_OK;
}

extern "C" int ao_b200_tp_close(void* peer_ptr) {
  return cudaIpcCloseMemHandle(peer_ptr) == cudaSuccess ? AO_B200_OK : AO_B200_ERR_CUDA;
}

extern "C" int ao_b200_tp_free(void* dev_ptr) { return cudaFree(dev_ptr) == cudaSuccess ? AO_B200_OK : AO_B200_ERR_CUDA; }

extern "C" int ao_b200_tp_status(const void* own_buf, int* status_out) {
  AO_CHECK_ARG(own_buf && status_out);
  uint32_t v = 0;
  if (cudaMemcpy(&v, static_cast<const uint8_t*>(own_buf) + 8, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return AO_B200_ERR_CUDA;
  *status_out = (int)v;
  return AO_B200_OK;
}

extern "C" int ao_b200_tp_allreduce_add(const void* partial, void* h, int64_t n, void* const* bufs, int rank, int world,
                                        int64_t max_elems, void* stream) {
  AO_CHECK_ARG(partial && h && bufs && n > 0 && n % 8 == 0 && n <= max_elems);
  AO_CHECK_ARG(world >= 1 && world <= ao::kTpMaxWorld && rank >= 0 && rank < world);
  AO_CHECK_ARG(ao_aligned(partial, 16) && ao_aligned(h, 16));
  if (ao_b200_tp_buffer_bytes(world, max_elems) == 0) return AO_B200_ERR_UNSUPPORTED;
  ao::TpPeers peers{};
  for (int r = 0; r < world; ++r) {
    AO_CHECK_ARG(bufs[r] != nullptr);
    peers.buf[r] = static_cast<uint8_t*>(bufs[r]);
  }
  const int64_t per_cta = (int64_t)ao::kTpThreads * 8;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)((n + per_cta - 1) / per_cta));
  cfg.blockDim = dim3(ao::kTpThreads);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, ao::tp_allreduce_add_kernel, static_cast<const __nv_bfloat16*>(partial),
                            static_cast<__nv_bfloat16*>(h), n, peers, rank, world, max_elems) == cudaSuccess
             ? AO_B200_OK
             : AO_B200_ERR_CUDA;
}
