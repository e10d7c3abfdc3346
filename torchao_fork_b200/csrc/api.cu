// C-ABI entry points that dispatch between kernels + misc.
#include "common.cuh"

namespace ao {
int int4_gemv(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
              int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, cudaStream_t s);
int int4_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                    int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps,
                    const void* residual, int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s);
int int4_gemv_imma(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                   int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps,
                   const void* residual, int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s);
int int4_gemv_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np,
                 int64_t Kp, int G, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps, const void* residual,
                 int64_t ld_res, int epilogue, void* workspace, int64_t workspace_bytes, bool pdl, cudaStream_t s);
int64_t int4_gemv_tc_workspace_bytes(int64_t Np);
int int4_gemm_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* sz, int64_t Np, int64_t Kp,
                 int G, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, cudaStream_t s);
int64_t int4_gemm_tc_workspace_bytes(int64_t M, int64_t Np, int64_t Kp);
int int8_gemv(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Kp,
              void* y, int64_t N_out, int64_t ldy, cudaStream_t s);
int int8_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Np,
                    int64_t Kp, void* y, int64_t N_out, int64_t ldy, const void* norm_w, float eps, const void* residual,
                    int64_t ld_res, int epilogue, int max_stages, bool pdl, cudaStream_t s);
int int8_gemm_tc(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Np,
                 int64_t Kp, void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, cudaStream_t s);
}  // namespace ao

extern "C" const char* ao_b200_error_string(int code) {
  switch (code) {
    case AO_B200_OK: return "ok";
    case AO_B200_ERR_INVALID_ARG: return "invalid argument (null/misaligned pointer or bad shape)";
    case AO_B200_ERR_UNSUPPORTED: return "unsupported configuration";
    case AO_B200_ERR_CUDA: return "CUDA launch failed";
    case AO_B200_ERR_WORKSPACE: return "workspace too small";
    case AO_B200_ERR_STATIC_SCRATCH: return "problem too large for the fixed static scratch (AO_B200_STATIC_SCRATCH_BYTES) of the allocation-free entry point";
  }
  return "unknown error";
}

extern "C" int ao_b200_abi_version(void) { return 5; }

#include <cstdlib>
namespace {
struct Tunables {
  int gemv_version, stages, pdl, gemv_tc, gemm_tc, gemv_imma;
  Tunables() {
    // tcgen05 prefill GEMM for M > 16 (0: loop the decode GEMV over 16-row chunks instead — kept for A/B testing)
    const char* gm = std::getenv("AO_B200_GEMM_TC");
    gemm_tc = gm ? std::atoi(gm) : 1;
    const char* v = std::getenv("AO_B200_GEMV");
    gemv_version = v ? std::atoi(v) : 3;
    const char* st = std::getenv("AO_B200_STAGES");
    stages = st ? std::atoi(st) : 8;
    if (stages < 2) stages = 2;
    if (stages > 16) stages = 16;
    const char* pd = std::getenv("AO_B200_PDL");
    pdl = pd ? std::atoi(pd) : 1;
    // tcgen05 / tensor-memory decode GEMV: 0 never, 1 wherever it applies (before the integer-MMA kernel), 2 (default)
    // for large weights that the integer-MMA kernel does not take (2 < M <= 5).  Against the fp16 mma.sync kernel its
    // steady-state rate is higher (5.2 vs 4.6 TB/s) at ~3 us more per launch; the integer-MMA kernel beats it at every
    // size (57344 x 8192, M = 1: 44.0 vs 51.7 us, profiles/r01c_gemv_70b_imma_vs_tc.jsonl)
    const char* tc = std::getenv("AO_B200_GEMV_TC");
    gemv_tc = tc ? std::atoi(tc) : 2;
    // integer-MMA decode GEMV (M <= 2; gemv_int4_imma.cu): 1 (default) wherever it applies, 0 never
    const char* im = std::getenv("AO_B200_GEMV_IMMA");
    gemv_imma = im ? std::atoi(im) : 1;
  }
};
const Tunables& tunables() {
  static Tunables t;
  return t;
}
}  // namespace

extern "C" int64_t ao_b200_int4_mm_workspace_bytes(int64_t M, int64_t Np, int64_t Kp) {
  // [0, gemv): tickets + partial sums of the tensor-memory GEMV's split-K (row blocks shared by neighbouring CTAs; the
  // tickets must stay zero between launches, so nothing else may ever be written there); [gemv, ...): the prefill
  // GEMM's permuted copy of the activations (M > 16).  The mma.sync kernels reduce inside the CTA and need none.
  const int64_t gemv = ao::int4_gemv_tc_workspace_bytes(Np);
  return gemv + (M > 16 ? ao::int4_gemm_tc_workspace_bytes(M, Np, Kp) : 0);
}

extern "C" int ao_b200_int4_mm(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed,
                               const void* scale_and_zero, int64_t Np, int64_t Kp, int group_size, void* y,
                               int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, void* stream) {
  AO_CHECK_ARG(x && packed && scale_and_zero && y);
  AO_CHECK_ARG(M > 0 && Np > 0 && Kp > 0 && Np % 16 == 0 && Kp % 64 == 0 && N_out > 0 && N_out <= Np);
  AO_CHECK_ARG(Kx > 0 && Kx <= Kp && Kx % 8 == 0 && ldx % 8 == 0 && ldx >= Kx && ldy >= N_out);
  AO_CHECK_ARG(ao_aligned(x, 16) && ao_aligned(packed, 16) && ao_aligned(scale_and_zero, 64) && ao_aligned(y, 2));
  if (!(group_size == 32 || group_size == 64 || group_size == 128 || group_size == 256)) return AO_B200_ERR_UNSUPPORTED;
  AO_CHECK_ARG(Kp % group_size == 0);
  auto s = static_cast<cudaStream_t>(stream);
  auto xb = static_cast<const __nv_bfloat16*>(x);
  auto yb = static_cast<__nv_bfloat16*>(y);
  if (M > 16 && tunables().gemm_tc) {
    // prefill: tensor-memory dequant-GEMM; its staging buffer sits after the GEMV region of the workspace
    const int64_t off = ao::int4_gemv_tc_workspace_bytes(Np);
    const int e = ao::int4_gemm_tc(x, M, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, y, N_out, ldy,
                                   workspace ? static_cast<char*>(workspace) + off : nullptr, workspace_bytes - off, s);
    if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
  }
  for (int64_t m0 = 0; m0 < M; m0 += 16) {
    const int64_t mc = (M - m0) < 16 ? (M - m0) : 16;
    int e = AO_B200_ERR_WORKSPACE;
    const bool tc_large = tunables().gemv_tc == 2 && Np * Kp >= (int64_t)192 << 20;
    for (int pass = 0; pass < 2; ++pass) {
      // pass 0: tensor-memory kernel when forced; integer-MMA kernel; pass 1: tensor-memory kernel for large weights
      if (((pass == 0 && tunables().gemv_tc == 1) || (pass == 1 && tc_large)) && mc <= 5) {
        e = ao::int4_gemv_tc(xb + m0 * ldx, mc, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, yb + m0 * ldy, N_out,
                             ldy, nullptr, 0.f, nullptr, 0, 0, workspace, workspace_bytes, tunables().pdl != 0, s);
        if (e == AO_B200_OK) break;
        if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
      }
      if (pass == 0 && tunables().gemv_imma && mc <= 2) {
        e = ao::int4_gemv_imma(xb + m0 * ldx, mc, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, yb + m0 * ldy, N_out,
                               ldy, nullptr, 0.f, nullptr, 0, 0, tunables().stages, tunables().pdl != 0, s);
        if (e == AO_B200_OK) break;
        if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
      }
    }
    if (e == AO_B200_OK) continue;
    if (tunables().gemv_version == 3) {
      e = ao::int4_gemv_fused(xb + m0 * ldx, mc, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, yb + m0 * ldy,
                              N_out, ldy, nullptr, 0.f, nullptr, 0, 0, tunables().stages, tunables().pdl != 0, s);
      if (e == AO_B200_OK) continue;
      if (e != AO_B200_ERR_WORKSPACE) return e;
    }
    e = ao::int4_gemv(xb + m0 * ldx, mc, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, yb + m0 * ldy, N_out,
                      ldy, s);
    if (e) return e;
  }
  return AO_B200_OK;
}

extern "C" int ao_b200_int4_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed,
                                       const void* scale_and_zero, int64_t Np, int64_t Kp, int group_size, void* y,
                                       int64_t N_out, int64_t ldy, const void* norm_weight, float norm_eps,
                                       const void* residual, int64_t ld_res, int epilogue, void* workspace,
                                       int64_t workspace_bytes, void* stream) {
  AO_CHECK_ARG(x && packed && scale_and_zero && y);
  AO_CHECK_ARG(M > 0 && M <= 16 && Np > 0 && Kp > 0 && Np % 16 == 0 && Kp % 64 == 0 && N_out > 0 && N_out <= Np);
  AO_CHECK_ARG(Kx > 0 && Kx <= Kp && Kx % 8 == 0 && ldx % 8 == 0 && ldx >= Kx);
  AO_CHECK_ARG(ao_aligned(x, 16) && ao_aligned(packed, 16) && ao_aligned(scale_and_zero, 64));
  AO_CHECK_ARG(epilogue >= 0 && epilogue <= 2);
  AO_CHECK_ARG(epilogue != 1 || residual != nullptr);
  AO_CHECK_ARG(epilogue != 2 || (N_out % 16 == 0 && ldy >= N_out / 2));
  AO_CHECK_ARG(epilogue == 2 || ldy >= N_out);
  AO_CHECK_ARG(!norm_weight || ao_aligned(norm_weight, 16));
  if (!(group_size == 32 || group_size == 64 || group_size == 128 || group_size == 256)) return AO_B200_ERR_UNSUPPORTED;
  AO_CHECK_ARG(Kp % group_size == 0);
  const bool tc_large = tunables().gemv_tc == 2 && Np * Kp >= (int64_t)192 << 20;
  for (int pass = 0; pass < 2; ++pass) {
    if ((pass == 0 && tunables().gemv_tc == 1) || (pass == 1 && tc_large)) {
      // tensor-memory kernel (M <= 5, Np % 128 == 0, activations + >= 3 ring stages must fit); else the mma.sync kernels
      const int e = ao::int4_gemv_tc(x, M, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, y, N_out, ldy, norm_weight,
                                     norm_eps, residual, ld_res, epilogue, workspace, workspace_bytes, tunables().pdl != 0,
                                     static_cast<cudaStream_t>(stream));
      if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
    }
    if (pass == 0 && tunables().gemv_imma && M <= 2) {
      const int e = ao::int4_gemv_imma(x, M, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, y, N_out, ldy, norm_weight,
                                       norm_eps, residual, ld_res, epilogue, tunables().stages, tunables().pdl != 0,
                                       static_cast<cudaStream_t>(stream));
      if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
    }
  }
  return ao::int4_gemv_fused(x, M, Kx, ldx, packed, scale_and_zero, Np, Kp, group_size, y, N_out, ldy, norm_weight,
                             norm_eps, residual, ld_res, epilogue, tunables().stages, tunables().pdl != 0,
                             static_cast<cudaStream_t>(stream));
}

extern "C" int64_t ao_b200_int8_mm_workspace_bytes(int64_t M, int64_t Np, int64_t Kp) {
  // the GEMM's permuted copy of the activations (+ split-K partial sums).  (Routing 2 < M <= 16 batches that do not fit the
  // GEMV staging through the GEMM was measured SLOWER than the plain strip GEMV: 40 vs 18-28 us on 4096 x 14336.)
  return M > 16 ? ao::int4_gemm_tc_workspace_bytes(M, Np, Kp) : 0;
}

static int int8_args_ok(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale, int64_t Np,
                        int64_t Kp, const void* y, int64_t N_out) {
  AO_CHECK_ARG(x && packed && scale && y && M > 0);
  AO_CHECK_ARG(Np > 0 && Kp > 0 && Np % 16 == 0 && Kp % 64 == 0 && N_out > 0 && N_out <= Np);
  AO_CHECK_ARG(Kx > 0 && Kx <= Kp && Kx % 8 == 0 && ldx % 8 == 0 && ldx >= Kx);
  AO_CHECK_ARG(ao_aligned(x, 16) && ao_aligned(packed, 16) && ao_aligned(scale, 2));
  return AO_B200_OK;
}

extern "C" int ao_b200_int8_mm(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed, const void* scale,
                               int64_t Np, int64_t Kp, void* y, int64_t N_out, int64_t ldy, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  if (int e = int8_args_ok(x, M, Kx, ldx, packed, scale, Np, Kp, y, N_out)) return e;
  AO_CHECK_ARG(ldy >= N_out);
  auto s = static_cast<cudaStream_t>(stream);
  auto xb = static_cast<const __nv_bfloat16*>(x);
  auto yb = static_cast<__nv_bfloat16*>(y);
  if (M > 16 && tunables().gemm_tc) {
    const int e = ao::int8_gemm_tc(x, M, Kx, ldx, packed, scale, Np, Kp, y, N_out, ldy, workspace, workspace_bytes, s);
    if (e != AO_B200_ERR_WORKSPACE && e != AO_B200_ERR_UNSUPPORTED) return e;
  }
  for (int64_t m0 = 0; m0 < M; m0 += 16) {
    const int64_t mc = (M - m0) < 16 ? (M - m0) : 16;
    int e = ao::int8_gemv_fused(xb + m0 * ldx, mc, Kx, ldx, packed, scale, Np, Kp, yb + m0 * ldy, N_out, ldy, nullptr, 0.f,
                                nullptr, 0, 0, tunables().stages, tunables().pdl != 0, s);
    if (e == AO_B200_ERR_WORKSPACE) e = ao::int8_gemv(xb + m0 * ldx, mc, Kx, ldx, packed, scale, Kp, yb + m0 * ldy, N_out, ldy, s);
    if (e) return e;
  }
  return AO_B200_OK;
}

extern "C" int ao_b200_int8_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx, const void* packed,
                                       const void* scale, int64_t Np, int64_t Kp, void* y, int64_t N_out, int64_t ldy,
                                       const void* norm_weight, float norm_eps, const void* residual, int64_t ld_res,
                                       int epilogue, void* stream) {
  if (int e = int8_args_ok(x, M, Kx, ldx, packed, scale, Np, Kp, y, N_out)) return e;
  AO_CHECK_ARG(M <= 16 && epilogue >= 0 && epilogue <= 2);
  AO_CHECK_ARG(epilogue != 1 || residual != nullptr);
  AO_CHECK_ARG(epilogue != 2 || (N_out % 16 == 0 && ldy >= N_out / 2));
  AO_CHECK_ARG(epilogue == 2 || ldy >= N_out);
  AO_CHECK_ARG(!norm_weight || ao_aligned(norm_weight, 16));
  return ao::int8_gemv_fused(x, M, Kx, ldx, packed, scale, Np, Kp, y, N_out, ldy, norm_weight, norm_eps, residual, ld_res,
                             epilogue, tunables().stages, tunables().pdl != 0, static_cast<cudaStream_t>(stream));
}


// ---- serving-loop plumbing (no reference counterpart) ----------------------------------------------------------------
// Stream-ordered wait on a 32-bit word in pinned host memory: everything enqueued on `stream` after this call starts only
// once *flag >= value.  Lets a host loop enqueue the NEXT decode step (its launch latency) while it is still reading the
// current step's token: the step's own input copy runs behind the wait, after the host has stored the token and bumped the
// flag.  Driver entry point resolved through the runtime (no link dependency on libcuda).
#include <cuda.h>
extern "C" int ao_b200_stream_wait_value32(void* stream, void* flag, uint32_t value) {
  AO_CHECK_ARG(flag != nullptr && ao_aligned(flag, 4));
  using Fn = CUresult (*)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
  static Fn fn = []() -> Fn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<Fn>(p);
  }();
  if (fn == nullptr) return AO_B200_ERR_UNSUPPORTED;
  void* dptr = nullptr;
  if (cudaHostGetDevicePointer(&dptr, flag, 0) != cudaSuccess) {
    cudaGetLastError();
    return AO_B200_ERR_INVALID_ARG;  // not pinned / mapped host memory
  }
  if (fn(static_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(dptr), value, CU_STREAM_WAIT_VALUE_GEQ) != CUDA_SUCCESS)
    return AO_B200_ERR_CUDA;
  return AO_B200_OK;
}
