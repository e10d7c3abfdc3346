/*
 * ao_b200.h — C ABI of libao_b200.so: B200 (sm_100a) kernels for torchao's
 * weight-only quantized linear path.
 *
 * Boundary rules (mirror SURVEY.md §8(b)):
 *   - plain pointers + sizes + cudaStream_t (passed as void*); no torch / ATen types;
 *   - every pointer is a DEVICE pointer on the current CUDA device unless noted;
 *   - the library never allocates, frees, synchronises or retains pointers: all
 *     outputs and workspaces are caller-provided, every launch goes to `stream`,
 *     so every entry point is CUDA-graph-capturable and re-entrant;
 *   - return value: 0 = ok, otherwise an AO_B200_ERR_* code (never throws/aborts).
 *     ao_b200_error_string() turns a code into text for the host wrapper's
 *     RuntimeError (the reference raises c10::Error from TORCH_CHECK,
 *     torchao/csrc/cuda/tensor_core_tiled_layout/tensor_core_tiled_layout.cu:236-254).
 *
 * bf16 tensors are passed as `const void*` to 2-byte elements, row-major.
 *
 * HBM layouts
 *   strip-major int4 weight (this library's format), uint32 [Np/16][Kp/64][32][4]:
 *       strip s (16 rows) x k-tile kt (64 k) = one 512-byte sub-tile; ALL sub-tiles of a strip are contiguous
 *       (Kp * 8 bytes), so a CTA that owns a range of strips streams one contiguous byte range with large TMA bulk
 *       copies.  Inside a sub-tile lane l = (g = l>>2, c = l&3) owns 16 bytes; word j of the lane:
 *       row = 16 s + g + 8 (j&1), k0 = 64 kt + 32 (j>>1) + 8 c, and holds k0..k0+7 in nibble slots
 *       [k0,k0+2,k0+4,k0+6,k0+1,k0+3,k0+5,k0+7] (slot i = bits 4i..4i+3) so that "(w >> 4i) & 0x000f000f" yields
 *       the pair (k0+2i, k0+2i+1).  A lane's 4 words are exactly the A-operand rows (g, g+8) of two m16n8k16 MMAs
 *       per k-half, and each word is one 16-byte chunk of a K-major UMMA shared-memory operand row.
 *       Np % 16 == 0 (the host pads to 128), Kp % 64 == 0.
 *   scale_and_zero, strip-major: bf16 [Np/16][Kp/G][16][2] = (scale, zero) — the numbers of the reference's
 *       pack_tinygemm_scales_and_zeros (torchao/quantization/utils.py:395-409, [Kp/G][Np][2]) with the 16 rows of a
 *       strip made contiguous, so a strip's qparams for a K range are one contiguous copy.
 *   "tinygemm" layout: int32 [N/8][K/(ikt*16)][32][ikt/2] + scale_and_zero bf16 [K/G][N][2], the formats of
 *       aten._convert_weight_to_int4pack / pack_tinygemm_scales_and_zeros (tensor_core_tiled_layout.cu:128-215);
 *       only the ao_b200_tinygemm_* functions take them.
 */
#ifndef AO_B200_H_
#define AO_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define AO_B200_API __attribute__((visibility("default")))
#else
#define AO_B200_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

enum {
  AO_B200_OK = 0,
  AO_B200_ERR_INVALID_ARG = 1,   /* null pointer, bad shape, misaligned pointer */
  AO_B200_ERR_UNSUPPORTED = 2,   /* group size / M / arch not supported        */
  AO_B200_ERR_CUDA = 3,          /* a CUDA launch failed (cudaGetLastError)    */
  AO_B200_ERR_WORKSPACE = 4,     /* caller-provided workspace too small        */
  AO_B200_ERR_STATIC_SCRATCH = 5 /* ao_b200_decode_attention_static: needs more than AO_B200_STATIC_SCRATCH_BYTES */
};

/* how `tensor / python_scalar` is evaluated when restating the reference's eager ops */
enum {
  AO_B200_DIV_IEEE = 0,  /* fp32 IEEE division      (torch CPU eager)               */
  AO_B200_DIV_RECIP = 1  /* x * (1.0f / c) in fp32  (torch CUDA eager; the default) */
};

AO_B200_API const char* ao_b200_error_string(int code);
AO_B200_API int ao_b200_abi_version(void);

/* ---------------------------------------------------------------- int4: quantize ---- */

/* Fused replacement for TensorCoreTiledLayout.pre_process (zero pad) +
 * _choose_qparams_affine_tinygemm (quant_primitives.py:1238-1306) +
 * _quantize_affine_tinygemm (quant_primitives.py:462-573) +
 * TensorCoreTiledAQTTensorImpl.from_plain (tensor_core_tiled_layout.py:262-307):
 * one pass over W.  w: bf16 [N][K] with row stride ldw (elements, % 8 == 0).
 * Rows >= N / columns >= K of the padded [Np][Kp] problem read as zero.
 * Outputs: packed (strip-major, Np x Kp), scale_and_zero bf16 strip-major [Np/16][Kp/G][16][2].
 * G in {32,64,128,256}; K % G == 0; Kp % max(G,64) == 0; Np % 16 == 0. */
AO_B200_API int ao_b200_int4_quantize_pack(const void* w, int64_t N, int64_t K, int64_t ldw,
                               int64_t Np, int64_t Kp, int group_size, float eps, int div_mode,
                               void* packed, void* scale_and_zero, void* stream);

/* _choose_qparams_affine_tinygemm alone: scale, zero bf16 [N][K/G]. */
AO_B200_API int ao_b200_int4_choose_qparams(const void* w, int64_t N, int64_t K, int64_t ldw, int group_size,
                                float eps, int div_mode, void* scale, void* zero, void* stream);

/* _quantize_affine_tinygemm alone: q int32 [N][K] in 0..15 from given bf16 scale/zero [N][K/G]. */
AO_B200_API int ao_b200_int4_quantize(const void* w, int64_t N, int64_t K, int64_t ldw, int group_size,
                          const void* scale, const void* zero, int32_t* q, void* stream);

/* ---------------------------------------------------------------- int4: layouts ----- */

/* from_plain: q int32 [N][K] (values 0..15; rows>=N / cols>=K pad with `pad_value`) -> strip-major layout. */
AO_B200_API int ao_b200_int4_pack(const int32_t* q, int64_t N, int64_t K, int64_t Np, int64_t Kp,
                      int pad_value, void* packed, void* stream);
/* strip-major layout -> int32 [Np][Kp]. */
AO_B200_API int ao_b200_int4_unpack(const void* packed, int64_t Np, int64_t Kp, int32_t* q, void* stream);
/* strip-major layout + strip-major scale_and_zero -> bf16 [Np][Kp], W' = bf16_rn((q-8)*s + z) (single rounding,
 * the dequant of aten._weight_int4pack_mm; tensor_core_tiled_layout.cu:41-117,194). */
AO_B200_API int ao_b200_int4_dequant(const void* packed, const void* scale_and_zero, int64_t Np, int64_t Kp,
                         int group_size, void* out, void* stream);

/* torchao::unpack_tensor_core_tiled_layout (torchao/ops.py:255-296, .cu:320-368):
 * tinygemm layout -> int32 [N][K]. inner_k_tiles in {2,4,8}. */
AO_B200_API int ao_b200_tinygemm_unpack(const int32_t* packed, int64_t N, int64_t K, int inner_k_tiles,
                            int32_t* q, void* stream);
/* torchao::dequantize_tensor_core_tiled_layout (ops.py:299-377, .cu:223-318). */
AO_B200_API int ao_b200_tinygemm_dequant(const int32_t* packed, const void* scale_and_zero, int64_t N, int64_t K,
                             int group_size, int inner_k_tiles, void* out, void* stream);
/* inverse of the above unpack == aten._convert_weight_to_int4pack on already-paired nibbles:
 * q int32 [N][K] -> tinygemm layout (for interchange with reference-packed checkpoints). */
AO_B200_API int ao_b200_tinygemm_pack(const int32_t* q, int64_t N, int64_t K, int inner_k_tiles,
                          int32_t* packed, void* stream);

/* ---------------------------------------------------------------- int4: linear ------ */

/* Replacement for aten._weight_int4pack_mm(x, packed, G, scale_and_zero)
 * (call site tensor_core_tiled_layout.py:104):
 *   y[m][n] = bf16( sum_k fp32( x[m][k] * W'[n][k] ) ),  n < N_out.
 * x: bf16 [M][Kx] row stride ldx (Kx <= Kp, Kx % 8 == 0; k >= Kx reads as zero — no
 * host-side F.pad of the activation); y: bf16 [M][N_out] row stride ldy.
 * M <= 16 -> HBM-bound split-K tensor-core GEMV; larger M -> tcgen05 GEMM.
 * workspace: ao_b200_int4_mm_workspace_bytes(...) bytes, zero-initialised once by the
 * caller (the kernels leave it zeroed on exit); may be NULL when that returns 0. */
AO_B200_API int64_t ao_b200_int4_mm_workspace_bytes(int64_t M, int64_t Np, int64_t Kp);
AO_B200_API int ao_b200_int4_mm(const void* x, int64_t M, int64_t Kx, int64_t ldx,
                    const void* packed, const void* scale_and_zero, int64_t Np, int64_t Kp,
                    int group_size, void* y, int64_t N_out, int64_t ldy,
                    void* workspace, int64_t workspace_bytes, void* stream);

/* Decode-path GEMV (M <= 16) with the neighbouring element-wise ops of the Llama block fused in
 * (reference: torchao/_models/llama/model.py:405-408,486,489-503):
 *   norm_weight != NULL : the activation is RMSNorm(x) * norm_weight (RMSNorm.forward), computed in the prologue
 *                         (x * norm_weight rounded to bf16; rsqrt(mean(x^2)+eps) applied to the fp32 result);
 *   epilogue 0 (plain)  : y[m][n] = bf16(acc)
 *   epilogue 1 (residual): y[m][n] = bf16(residual[m][n] + bf16(acc))          (h = x + attention(...), y may alias residual)
 *   epilogue 2 (silu*mul): the packed weight holds w1 and w3 interleaved 8 rows / 8 rows per 16-row strip
 *                         (ops.interleave_gate_up); y[m][i] = bf16(bf16(silu(bf16(gate_i))) * bf16(up_i)), i < N_out/2.
 * Arithmetic: the affine dequant is applied per group in fp32 to the tensor-core partial sums (the fp32 dequant of
 * the reference's _weight_int4pack_mm_for_cpu) — see csrc/gemv_int4_fused.cu, csrc/gemv_int4_tc.cu.
 * workspace: ao_b200_int4_mm_workspace_bytes(M, Np, Kp) bytes, zero-initialised once by the caller (the kernels leave
 * it zeroed); with NULL / too small a workspace the call still works but skips the tensor-memory kernel.  Returns
 * AO_B200_ERR_WORKSPACE when M * K does not fit any decode kernel's shared-memory staging (use ao_b200_int4_mm then). */
AO_B200_API int ao_b200_int4_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx,
                            const void* packed, const void* scale_and_zero, int64_t Np, int64_t Kp,
                            int group_size, void* y, int64_t N_out, int64_t ldy,
                            const void* norm_weight, float norm_eps, const void* residual, int64_t ld_res,
                            int epilogue, void* workspace, int64_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- int8 -------------- */

/* choose_qparams_affine(SYMMETRIC, int8, block (1,K), eps=fp32 eps) + quantize_affine
 * (quant_primitives.py:1537-1565, :398-459; caller quant_api.py:1223-1240).
 * w bf16 [N][K] -> q int8 [N][K] (plain, row-major: PlainAQTTensorImpl.int_data),
 * scale bf16 [N].  zero_point is identically 0. */
AO_B200_API int ao_b200_int8_quantize(const void* w, int64_t N, int64_t K, int64_t ldw, float eps, int div_mode,
                          int8_t* q, void* scale, void* stream);

/* HQQ qparam search for Int4WeightOnlyConfig(use_hqq=True): replaces _choose_qparams_and_quantize_affine_hqq
 * (quant_primitives.py:1998-2111, optimizer :1915-1973, format conversion :1981-1995; caller
 * affine_quantized_tensor.py:252-288).  w bf16, contiguous, viewed as [n_groups][group_size]; scale / zero fp32
 * [n_groups] are the solver's state (hqq convention  W ~ (q - zero) / scale).
 *   init:     scale = min(15 / (max - min), 2e4), zero = round(-min * scale)
 *   step:     one proximal iteration (zero updated in place); err_partial[ao_b200_hqq_blocks(n_groups)] receives the
 *             per-block sums of |W - W_r|: the caller divides their sum by the element count and applies the reference's
 *             global early stop (stop at the first iteration whose mean error does not improve), multiplying beta by
 *             kappa between iterations
 *   finalize: q int32 [n_groups][group_size] in 0..15, scale_out = bf16(1 / scale), zero_out = bf16((8 - zero) / scale):
 *             the tinygemm convention  w' = (q - 8) * scale_out + zero_out, ready for ao_b200_int4_pack. */
AO_B200_API int64_t ao_b200_hqq_blocks(int64_t n_groups);
AO_B200_API int ao_b200_hqq_init(const void* w, int64_t n_groups, int group_size, float* scale, float* zero, void* stream);
AO_B200_API int ao_b200_hqq_step(const void* w, int64_t n_groups, int group_size, const float* scale, float* zero,
                                 double beta, double lp_norm, float* err_partial, void* stream);
/* The whole solver loop without a host round trip (ABI v5): `iters` step launches, each followed by a one-block kernel
 * that applies the reference's global early stop on the device (quant_primitives.py:1964-1972: stop at the first
 * iteration whose mean |W - W_r| does not improve; that iteration's zero is kept) — the remaining step launches return
 * at once.  beta is multiplied by kappa between iterations.  state: 16 bytes of device memory, 8-byte aligned:
 * [double best][int32 stopped][int32 iterations run] (written here; readable after the stream has caught up). */
AO_B200_API int ao_b200_hqq_solve(const void* w, int64_t n_groups, int group_size, const float* scale, float* zero,
                                  double beta, double kappa, double lp_norm, int iters, float* err_partial, void* state,
                                  void* stream);
AO_B200_API int ao_b200_hqq_finalize(const void* w, int64_t n_groups, int group_size, const float* scale, const float* zero,
                                     int32_t* q, void* scale_out, void* zero_out, void* stream);

/* strip-major int8 layout (this library's format): uint8 [Np/16][Kp/64][2][32][16], BIASED bytes u = q + 128.
 *   strip s (16 rows) x k-tile kt (64 k) = 1024 contiguous bytes = two 32-k halves of 512 bytes; inside a half lane
 *   l = (g = l>>2, c = l&3) owns 16 bytes = 4 words; word j holds row 16 s + g + 8 (j&1), k0 = 64 kt + 32 h + 8 c + 4 (j>>1),
 *   bytes k0 .. k0+3.  As with the int4 layout all tiles of a strip are contiguous (Kp * 16 bytes), the lane chunk is the
 *   A fragment of two m16n8k16 MMAs and the (lane, column) pattern of a tcgen05.st.16x128b.  Rows >= N / columns >= K hold
 *   q = 0.  PlainAQTTensorImpl.int_data (plain_layout.py:47-229) is recovered exactly by ao_b200_int8_unpack. */
AO_B200_API int ao_b200_int8_pack(const int8_t* q, int64_t N, int64_t K, int64_t Np, int64_t Kp, void* packed, void* stream);
AO_B200_API int ao_b200_int8_unpack(const void* packed, int64_t Np, int64_t Kp, int64_t N, int64_t K, int8_t* q, void* stream);

/* Replacement for _linear_fp_act_int8_weight_impl (plain_layout.py:250-266):
 *   y = bf16( bf16( x @ q^T ) * scale ).   x bf16 [M][Kx] (ldx; k >= Kx reads as zero), packed strip-major int8 (Np x Kp),
 *   scale bf16 [N_out], y bf16 [M][N_out] (ldy).  M <= 16: TMA-ring GEMV (gemv_int8_fused.cu); larger M: tcgen05 GEMM
 *   (needs Np % 128 == 0 and a workspace of ao_b200_int8_mm_workspace_bytes; otherwise the GEMV is looped). */
AO_B200_API int64_t ao_b200_int8_mm_workspace_bytes(int64_t M, int64_t Np, int64_t Kp);
AO_B200_API int ao_b200_int8_mm(const void* x, int64_t M, int64_t Kx, int64_t ldx,
                    const void* packed, const void* scale, int64_t Np, int64_t Kp,
                    void* y, int64_t N_out, int64_t ldy, void* workspace, int64_t workspace_bytes, void* stream);

/* int8 sibling of ao_b200_int4_gemv_fused (same prologue / epilogue semantics, M <= 16).  For epilogue 2 the packed
 * weight and `scale` hold w1 / w3 interleaved 8 rows / 8 rows per strip.  Returns AO_B200_ERR_WORKSPACE when M * Kp does
 * not fit the shared-memory staging (use ao_b200_int8_mm then). */
AO_B200_API int ao_b200_int8_gemv_fused(const void* x, int64_t M, int64_t Kx, int64_t ldx,
                            const void* packed, const void* scale, int64_t Np, int64_t Kp,
                            void* y, int64_t N_out, int64_t ldy,
                            const void* norm_weight, float norm_eps, const void* residual, int64_t ld_res,
                            int epilogue, void* stream);

/* ---------------------------------------------------------------- decode-step glue ---- */
/* The non-GEMM remainder of one Llama decode step (reference:
 * torchao/_models/llama/model.py).  All of them read the token position from DEVICE memory
 * (int32 *input_pos) so that one captured CUDA graph can be replayed for every token. */

/* RMSNorm.forward (model.py:489-503) with an optional fused residual add:
 *   h = bf16(x + residual) (written to h_out when non-NULL);  y = bf16( bf16(h * rsqrt(mean(h^2)+eps)) * weight ).
 * x, residual, h_out, y: bf16 [rows][dim]; weight bf16 [dim]; dim % 8 == 0. */
AO_B200_API int ao_b200_rmsnorm(const void* x, const void* residual, const void* weight, void* h_out, void* y,
                    int64_t rows, int64_t dim, float eps, void* stream);

/* apply_rotary_emb (model.py:547-559) on q and k + KVCache.update (model.py:175-192) for ONE new token.
 * qkv bf16 [B][(H + 2 Hkv) * head_dim] (the wqkv output), freqs_cis bf16 [block_size][head_dim/2][2],
 * q_out bf16 [B][H][head_dim], caches bf16 [B][Hkv][S][head_dim]. */
AO_B200_API int ao_b200_rope_kv_append(const void* qkv, const void* freqs_cis, const int32_t* input_pos, void* q_out,
                           void* k_cache, void* v_cache, int B, int H, int Hkv, int head_dim, int S, void* stream);

/* scaled_dot_product_attention of one query token against cache positions 0..*input_pos, GQA without
 * repeat_interleave (model.py:465-470).  head_dim must be 128.  n_split > 1 splits the sequence over CTAs
 * (workspace: ao_b200_decode_attention_workspace_bytes).  out bf16 [B][H][head_dim]. */
AO_B200_API int64_t ao_b200_decode_attention_workspace_bytes(int B, int H, int n_split);
AO_B200_API int ao_b200_decode_attention(const void* q, const void* k_cache, const void* v_cache,
                             const int32_t* input_pos, void* out, int B, int H, int Hkv, int head_dim, int S,
                             float scale, int n_split, void* workspace, int64_t workspace_bytes, void* stream);

/* Allocation-free variant of ao_b200_decode_attention.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid allocation).  Same inputs, outputs and arithmetic as ao_b200_decode_attention, but the split partials
 * (ao_b200_decode_attention_workspace_bytes(B, H, n_split) bytes) live in a file-scope `static __device__` array of
 * AO_B200_STATIC_SCRATCH_BYTES (1 MiB: B * H * n_split <= 2016) inside the library image — every GPU that loads the
 * library holds that 1 MiB until exit.  The call makes no cudaMalloc / cudaFree / allocating library call; the array's
 * device address is resolved once per device with cudaGetSymbolAddress.  Larger problems are REJECTED with
 * AO_B200_ERR_STATIC_SCRATCH (nothing is launched).  All calls on a device share the one array, which keeps the previous
 * call's contents (nothing in it has to start at zero: each call writes every partial it later reads): calls that may
 * overlap on the GPU (different streams, concurrent graphs) RACE on it — serialise them, or use ao_b200_decode_attention
 * with caller-owned workspaces. */
#define AO_B200_STATIC_SCRATCH_BYTES (1 << 20)
AO_B200_API int ao_b200_decode_attention_static(const void* q, const void* k_cache, const void* v_cache,
                                    const int32_t* input_pos, void* out, int B, int H, int Hkv, int head_dim, int S,
                                    float scale, int n_split, void* stream);

/* Attention.forward for ONE new token in one launch (model.py:405-440): RoPE on q and the new k (model.py:547-559),
 * KV append at *input_pos (KVCache.update, model.py:175-192), GQA softmax(q K^T * scale) V over positions 0..pos split
 * n_split ways, partials merged by the last CTA to finish.  qkv bf16 [B][(H + 2 Hkv) * 128] (the wqkv output), caches
 * bf16 [B][Hkv][S][128], out bf16 [B][H * 128].  workspace: ao_b200_decode_attention_fused_workspace_bytes, ZEROED
 * once by the caller (the kernel leaves its tickets zeroed).  Programmatic-dependent-launch aware. */
AO_B200_API int64_t ao_b200_decode_attention_fused_workspace_bytes(int B, int H, int Hkv, int n_split);
AO_B200_API int ao_b200_decode_attention_fused(const void* qkv, const void* freqs_cis, const int32_t* input_pos,
                                               void* k_cache, void* v_cache, void* out, int B, int H, int Hkv,
                                               int head_dim, int S, float scale, int n_split, void* workspace,
                                               int64_t workspace_bytes, void* stream);

/* The same launch over an int8 KV cache — the decode side of AffineQuantizedKVCache (model.py:198-240): the new k (after
 * RoPE) and v rows are quantized per token and head as _quantize_activation_per_token_absmax does
 * (torchao/quantization/utils.py:152-182: symmetric, -127..127, eps 1e-5, bf16 scale; bit-exact with the reference's
 * eager CUDA evaluation) and stored at *input_pos; the walk reads k_out = bf16(q * scale) for the older positions and
 * the UNQUANTIZED new row (model.py:222-225, 230-232).  caches int8 [B][Hkv][S][128], scales bf16 [B][Hkv][S] (the
 * reference's [B][Hkv][S][1]).  Workspace as for ao_b200_decode_attention_fused. */
AO_B200_API int ao_b200_decode_attention_fused_q8(const void* qkv, const void* freqs_cis, const int32_t* input_pos,
                                      void* k_cache, void* v_cache, void* k_scale, void* v_scale, void* out,
                                      int B, int H, int Hkv, int head_dim, int S, float scale, int n_split,
                                      void* workspace, int64_t workspace_bytes, void* stream);

/* FeedForward gate (model.py:486): out = bf16( bf16(silu(gate)) * up ), gate_up bf16 [B][2*I] = [w1 x | w3 x]. */
AO_B200_API int ao_b200_silu_mul(const void* gate_up, void* out, int64_t B, int64_t I, void* stream);

/* sample (generate.py:79-100: logits_to_probs + multinomial_sample_one_no_sync) in one launch:
 *   lg = float(logits) / max(temperature, 1e-5); entries below the top_k-th largest -> -inf (ties kept, top_k <= 0 or
 *   >= V: keep all); probs = softmax(lg); next_tok[b] = argmax_i probs[i] / noise[b][i]  (first index on ties).
 * noise fp32 [B][ldq] holds Exp(1) samples drawn by the CALLER (torch's exponential_ on the reference side), so the
 * random stream stays the framework generator's.  temperature == 0: greedy argmax, noise may be null.
 * logits bf16 [B][ldl], next_tok int64 [B]. */
AO_B200_API int ao_b200_sample(const void* logits, int64_t ldl, int64_t B, int64_t V, float temperature, int top_k,
                               const void* noise, int64_t ldq, void* next_tok, void* stream);

/* ---- tensor-parallel exchange over NVLink peer memory (new: the reference has no int4 TP path; SURVEY.md section 8e
 * names the collective it replaces: ncclAllReduce(sum, bf16) of [B][dim] after the row-parallel wo and w2) ------------
 * One process per GPU.  Every rank allocates ONE exchange buffer with ao_b200_tp_alloc (cudaMalloc, zeroed; returns the
 * 64-byte CUDA IPC handle), hands the handle to its peers (any host channel), and opens theirs with ao_b200_tp_open.
 * bufs[r] = rank r's buffer as seen from this process (own pointer at bufs[rank]).
 *   ao_b200_tp_allreduce_add: h[i] = bf16( h[i] + sum_r partial_r[i] ), i < n, summed in rank order in fp32 — bit-identical
 *   on every rank.  One launch: pushes `partial` to all peers, flags, waits for theirs (bounded spin; on time-out the
 *   status word is set and the call returns garbage instead of hanging), sums, adds.  All ranks must issue the same
 *   sequence of calls (n may vary, max_elems and world may not).  n % 8 == 0, n <= max_elems <= 131072. */
AO_B200_API int64_t ao_b200_tp_buffer_bytes(int world, int64_t max_elems);
AO_B200_API int ao_b200_tp_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle_out /* 64 bytes */);
AO_B200_API int ao_b200_tp_open(const void* ipc_handle /* 64 bytes */, void** peer_ptr);
AO_B200_API int ao_b200_tp_close(void* peer_ptr);
AO_B200_API int ao_b200_tp_free(void* dev_ptr);
AO_B200_API int ao_b200_tp_status(const void* own_buf, int* status_out /* 0 ok, 1 a peer timed out */);
AO_B200_API int ao_b200_tp_allreduce_add(const void* partial, void* h, int64_t n, void* const* bufs, int rank, int world,
                                         int64_t max_elems, void* stream);

/* ---- one decode step as ONE persistent launch (csrc/decode_mega.cu) --------------------------------------------------
 * decode_one_token (torchao/_models/llama/generate.py:111-117) -> Transformer.forward for one token
 * (torchao/_models/llama/model.py:342-360: embedding, n_layer x TransformerBlock.forward :405-408 [RMSNorm :498-503, wqkv,
 * RoPE :547-559, KVCache.update :175-192, GQA attention :465-470, wo, residual, RMSNorm, w2(silu(w1 x) * w3 x) :486,
 * residual], final norm + output) with int4 weight-only linears (Int4WeightOnlyConfig, any supported group size), batch 1.
 * One cooperative launch of n_cta CTAs (one per SM) replaces the 161 launches per token of the per-projection path: the
 * weight stream never drains between projections.  The launch leaves bf16 logits[vocab] in `logits`, appends the
 * token's k / v to the caches at *pos, then sets *pos += 1 (device side: one captured graph serves every token).
 *
 * ao_b200_mega_model is filled by the HOST (plain struct, host memory); every pointer inside is a device pointer on the
 * current device except `layers` (host array of n_layer entries).  Weights are strip-major int4 + strip-major
 * scale_and_zero (see above); w13 = w1 / w3 rows interleaved 8 / 8 per 16-row strip (the SiLU*mul pairing of
 * ao_b200_int4_gemv_fused epilogue 2).  Caches bf16 [n_kv_head][max_seq][128]; freqs bf16 [>= max_seq][64][2].
 * Caller-owned device buffers:
 *   plan    ao_b200_decode_mega_plan_bytes(m) bytes, filled once by ao_b200_decode_mega_setup (synchronises `stream`:
 *           set-up time only, not capturable);
 *   state   ao_b200_decode_mega_state_bytes() bytes, ZEROED once: [u64 arrival counter][u32 launches][u32 status];
 *           status != 0 after a launch = a wait inside the kernel timed out (timeout_ms, default 2000): the outputs of
 *           that launch are garbage and state + xchg must be zeroed again before the next one;
 *   xchg[rank]  ao_b200_decode_mega_xchg_bytes(m) bytes, ZEROED once: the activation packets {2 bf16 | 32-bit tag};
 *           world > 1 (tensor parallel: wqkv / w13 / output column shards, wo / w2 row shards, n_head / n_kv_head / inter /
 *           vocab are the LOCAL sizes): xchg[r] = rank r's region as mapped into this process (CUDA IPC, like
 *           ao_b200_tp_alloc / ao_b200_tp_open); the wo / w2 partial sums are pushed into every peer's region from the
 *           GEMV epilogue and summed by the consumers of the next projection — at world >= 4 once per rank (its CTAs
 *           take a slice of the vector each and publish the bf16 sum in a local slot of the region), at world 2 in every
 *           CTA's sweep — no separate all-reduce launch; every rank adds in the same order, so the replicated residual
 *           stream is bit-identical on all ranks.  world must be a power of two (<= 8).  All ranks must launch the same
 *           sequence of steps.
 * ao_b200_decode_mega_step is CUDA-graph-capturable (cooperative kernel node). */
typedef struct ao_b200_mega_layer {
  const void* wqkv_packed; const void* wqkv_sz;
  const void* wo_packed; const void* wo_sz;
  const void* w13_packed; const void* w13_sz;
  const void* w2_packed; const void* w2_sz;
  const void* attn_norm; const void* ffn_norm;   /* bf16 [dim] */
  void* k_cache; void* v_cache;                  /* the model's KVCache, bf16 [n_kv_head][max_seq][128]: appended to */
  void* k_pad; void* v_pad;                      /* the kernel's own copy with padded rows, bf16 [n_kv_head][S_pad][136],
                                                    S_pad = max_seq rounded up to 32: must hold rows 0 .. *pos - 1 of the
                                                    caches when a launch starts (the launch appends row *pos to both) */
} ao_b200_mega_layer;

typedef struct ao_b200_mega_model {
  int32_t n_layer, dim, n_head, n_kv_head, head_dim, inter, vocab, group_size, max_seq;
  int32_t world, rank, n_cta, timeout_ms, attn_blocks_per_cta;  /* attn_blocks_per_cta: 32-position tiles per CTA before the sequence is split over more CTAs (0 = default 8) */
  float eps;
  int32_t reserved;
  const ao_b200_mega_layer* layers;   /* HOST array [n_layer] */
  const void* out_packed; const void* out_sz; const void* out_norm;
  const void* emb;                    /* bf16 [full vocab][dim] */
  const void* freqs;
  const int64_t* tok;                 /* the token to feed */
  int32_t* pos;                       /* its position; incremented by the launch */
  void* logits;                       /* bf16 [vocab] */
  void* state;
  void* xchg[8];
  void* trace;                        /* NULL, or u64 [n_cta][5 n_layer + 1][8]: per-phase clock64 stamps of every CTA (profiling) */
} ao_b200_mega_model;

AO_B200_API int64_t ao_b200_decode_mega_plan_bytes(const ao_b200_mega_model* m);
AO_B200_API int64_t ao_b200_decode_mega_xchg_bytes(const ao_b200_mega_model* m);
AO_B200_API int64_t ao_b200_decode_mega_state_bytes(void);
AO_B200_API int ao_b200_decode_mega_setup(const ao_b200_mega_model* m, void* plan_dev, int64_t plan_bytes, void* stream);
AO_B200_API int ao_b200_decode_mega_step(const ao_b200_mega_model* m, const void* plan_dev, void* stream);

/* Serving-loop plumbing, NO COUNTERPART IN THE REFERENCE: a stream-ordered wait until *flag >= value, `flag` a 32-bit
 * word in pinned host memory (cuStreamWaitValue32, resolved through the runtime).  The host loop of
 * _models/llama/generate.py enqueues the next decode step behind such a wait while it reads the current token, and
 * releases it by storing the token and then the flag.  AO_B200_ERR_UNSUPPORTED when the driver lacks the entry point. */
AO_B200_API int ao_b200_stream_wait_value32(void* stream, void* flag, uint32_t value);

#ifdef __cplusplus
}
#endif
#endif /* AO_B200_H_ */
