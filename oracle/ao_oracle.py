"""CPU oracle for the weight-only quantized linear path (TEST INFRASTRUCTURE ONLY).

This file is a numpy restatement of the reference algorithm.  It is the checker
for the CUDA kernels; it is never the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks every function
below bit-for-bit against ``tests/golden/*.npz``, which were produced by the
reference's own Python functions run on CPU in the build container
(``tests/golden/make_golden.py``, ``make_golden_hqq.py``, ``make_golden_kvq.py``; torch 2.11.0).
One exception, marked where it is defined: ``int4_row_quantize_zp`` / ``pack_int4``
restate a third-party dependency of the reference that is absent here (parity
with it is unpinned).

Conventions
-----------
* bf16 values are carried as ``np.float32`` arrays whose values are exactly
  bf16-representable; ``to_bits``/``from_bits`` convert to the uint16 pattern.
* The reference runs eager torch ops on bf16 tensors.  Each eager op computes
  in fp32 ("opmath") and rounds its result to bf16 (round-to-nearest-even).
  Every function here therefore rounds after every op, exactly where the
  reference has an op boundary.
* ``div_mode``: how ``tensor / python_scalar`` is evaluated.
    - ``"ieee"``  : fp32 IEEE division  (torch CPU eager; the golden vectors)
    - ``"recip"`` : multiply by the fp32 reciprocal of the scalar (torch CUDA
      eager: ATen's div_true_kernel_cuda multiplies by 1/b when b is a CPU
      scalar).  This is what the reference does when it quantizes a model that
      lives on the GPU, so it is the product kernels' default.
  ``tensor / tensor`` is IEEE division on both devices.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------
# bf16 helpers
# --------------------------------------------------------------------------
def bf16_round(x):
    """fp32 -> nearest-even bf16, returned as fp32 (what ``.to(bfloat16)`` does)."""
    x = np.ascontiguousarray(x, dtype=F32)
    u = x.view(np.uint32).astype(np.uint64)
    bias = ((u >> 16) & 1) + 0x7FFF
    r = ((u + bias) & 0xFFFF0000).astype(np.uint32)
    out = r.view(F32).copy()
    nan = np.isnan(x)
    if nan.any():
        out[nan] = np.nan
    return out.reshape(x.shape)


def to_bits(x):
    """bf16-valued fp32 array -> uint16 bit pattern."""
    x = np.ascontiguousarray(x, dtype=F32)
    return (x.view(np.uint32) >> 16).astype(np.uint16).reshape(x.shape)


def from_bits(b):
    """uint16 bf16 bit pattern -> fp32 array."""
    b = np.ascontiguousarray(b, dtype=np.uint16)
    return (b.astype(np.uint32) << 16).view(F32).reshape(b.shape)


def _div_scalar(x, c, div_mode):
    if div_mode == "ieee":
        return bf16_round(x.astype(F32) / F32(c))
    if div_mode == "recip":
        inv = F32(1.0) / F32(c)
        return bf16_round(x.astype(F32) * inv)
    raise ValueError(div_mode)


# --------------------------------------------------------------------------
# int4 (tinygemm flavour): qparams / quantize / dequantize
# --------------------------------------------------------------------------
def choose_qparams_tinygemm(w, group_size, eps=1e-6, div_mode="ieee"):
    """Reference: torchao/quantization/quant_primitives.py:1238-1306
    (_choose_qparams_affine_tinygemm, ASYMMETRIC, qmin 0, qmax 15, bf16 qparams).

    w: [N][K] bf16-valued.  Returns (scale, zero) each [N][K/G], bf16-valued.
    """
    n, k = w.shape
    assert k % group_size == 0
    wg = w.reshape(n, k // group_size, group_size).astype(F32)
    mn = wg.min(axis=-1)                              # :1287 amin
    mx = wg.max(axis=-1)                              # :1288 amax
    d = bf16_round(mx - mn)                           # :1294 (max - min)
    scale = _div_scalar(d, 15.0, div_mode)            # :1294  / float(qmax-qmin)
    scale = bf16_round(np.maximum(scale, F32(eps)))   # :1295 clamp(min=eps)
    t = bf16_round(scale * F32(8.0))                  # :1300 scale * mid_point
    zero = bf16_round(mn + t)                         # :1300 min + ...
    return scale, zero


def quantize_tinygemm(w, scale, zero, group_size):
    """Reference: quant_primitives.py:541-573 (_quantize_affine_tinygemm_no_dtype_cast).
    Returns int32 [N][K] in 0..15."""
    n, k = w.shape
    wg = w.reshape(n, k // group_size, group_size).astype(F32)
    s = scale.reshape(n, -1, 1).astype(F32)
    z = zero.reshape(n, -1, 1).astype(F32)
    t = bf16_round(s * F32(8.0))                      # :569 scale * mid_point
    min_val = bf16_round(z - t)                       # :569 zero - ...
    a = bf16_round(wg - min_val)                      # :570 input - min_val
    with np.errstate(divide="ignore", invalid="ignore"):
        b = bf16_round(a / s)                         # :570 / scale  (tensor/tensor: IEEE)
    q = np.clip(np.rint(b), 0, 15)                    # :570 round-half-even, clamp
    return q.astype(np.int32).reshape(n, k)


def pack_scales_and_zeros(scale, zero):
    """Reference: torchao/quantization/utils.py:395-409 -> [K/G][N][2]."""
    return np.ascontiguousarray(np.stack([scale, zero], axis=-1).transpose(1, 0, 2))


def unpack_scales_and_zeros(sz):
    """Reference: utils.py:412-414 -> scale, zero each [N][K/G]."""
    t = sz.transpose(1, 0, 2)
    return np.ascontiguousarray(t[..., 0]), np.ascontiguousarray(t[..., 1])


def _bf16_fma_exact(a, b, c):
    """round_bf16(a*b + c) with ONE rounding (what fma.rn.bf16 does).

    a*b is exact in float64 (a is a small int, b has 8 significant bits).  The
    float64 sum is exact unless the exponents of the two terms are > ~40 bits
    apart; those elements are redone with exact rational arithmetic.
    """
    a64 = a.astype(np.float64)
    b64 = b.astype(np.float64)
    c64 = c.astype(np.float64)
    p = a64 * b64
    s = p + c64
    out = bf16_round(_f64_to_f32_round_odd(s))
    # exactness check of the float64 addition (TwoSum residual)
    bb = s - p
    err = (p - (s - bb)) + (c64 - bb)
    bad = np.argwhere(err != 0)
    if bad.size:
        from fractions import Fraction
        for idx in map(tuple, bad):
            exact = Fraction(float(p[idx])) + Fraction(float(c64[idx]))
            out[idx] = _round_fraction_to_bf16(exact)
    return out


def _f64_to_f32_round_odd(x):
    """float64 -> float32 with round-to-odd so that a later bf16 RNE rounding
    equals a single rounding of the float64 value."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    f = x.astype(F32)
    inexact = f.astype(np.float64) != x
    u = f.view(np.uint32).copy()
    # move toward zero if we rounded away, then force the low bit to 1
    away = inexact & (np.abs(f.astype(np.float64)) > np.abs(x))
    u[away] -= 1
    u[inexact] |= 1
    return u.view(F32).reshape(x.shape)


def _round_fraction_to_bf16(fr):
    from fractions import Fraction
    if fr == 0:
        return F32(0.0)
    sign = -1 if fr < 0 else 1
    a = abs(fr)
    import math
    e = math.floor(math.log2(float(a))) if a > 0 else 0
    while Fraction(2) ** e > a:
        e -= 1
    while Fraction(2) ** (e + 1) <= a:
        e += 1
    ulp = Fraction(2) ** (e - 7)
    qn = a / ulp
    fl = qn.numerator // qn.denominator
    rem = qn - fl
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (fl & 1)):
        fl += 1
    return F32(sign * float(fl * ulp))


def dequantize_tinygemm(q, scale, zero, group_size):
    """W'[n][k] = bf16_rn((q-8)*s + z), ONE rounding.

    Reference: tensor_core_tiled_layout.cu:41-117 (q-8 exact via the 0x4300
    magic + fma.rn.bf16x2) and :194 (__hfma2(v, scale, zero)); python spelling
    test/test_ops.py:319-336.  Same arithmetic as aten._weight_int4pack_mm's
    in-register dequant."""
    n, k = q.shape
    qg = (q.reshape(n, k // group_size, group_size).astype(F32) - F32(8.0))
    s = np.broadcast_to(scale.reshape(n, -1, 1).astype(F32), qg.shape)
    z = np.broadcast_to(zero.reshape(n, -1, 1).astype(F32), qg.shape)
    return _bf16_fma_exact(qg, s, z).reshape(n, k)


def dequantize_affine_tinygemm_2round(q, scale, zero, group_size):
    """Reference python dequant (quant_primitives.py:939-1031,
    _dequantize_affine_tinygemm): (q - 8) * scale + zero with a bf16 rounding
    after the multiply AND after the add.  Used by AQT.dequantize()."""
    n, k = q.shape
    qg = q.reshape(n, k // group_size, group_size).astype(F32) - F32(8.0)
    s = scale.reshape(n, -1, 1).astype(F32)
    z = zero.reshape(n, -1, 1).astype(F32)
    return bf16_round(bf16_round(qg * s) + z).reshape(n, k)


# --------------------------------------------------------------------------
# packed layouts
# --------------------------------------------------------------------------
def nibble_pair_u8(q):
    """Reference: tensor_core_tiled_layout.py:276  (q[:, ::2] << 4 | q[:, 1::2]) -> uint8 [N][K/2]."""
    return ((q[:, ::2] << 4) | q[:, 1::2]).astype(np.uint8)


# nibble slot (bit offset / 4) of the i-th of 8 values inside one uint32 word:
# value order v0..v7 -> slots 0,4,1,5,2,6,3,7 (i.e. slot order v0,v2,v4,v6,v1,v3,v5,v7).
# Same in-word order as the reference (tensor_core_tiled_layout.cu:201-208); it is
# what makes "(w >> 4i) & 0x000f000f" yield the bf16x2 pair (v[2i], v[2i+1]).
_SLOT = np.array([0, 4, 1, 5, 2, 6, 3, 7], dtype=np.uint32)


def _pack_words(v):
    """v: [..., 8] ints in 0..15 -> uint32 words."""
    v = v.astype(np.uint32)
    w = np.zeros(v.shape[:-1], dtype=np.uint32)
    for i in range(8):
        w |= (v[..., i] & 0xF) << (4 * _SLOT[i])
    return w


def _unpack_words(w):
    w = w.astype(np.uint32)
    out = np.zeros(w.shape + (8,), dtype=np.int32)
    for i in range(8):
        out[..., i] = (w >> (4 * _SLOT[i])) & 0xF
    return out


def pack_tinygemm(q, inner_k_tiles=8):
    """Restatement of aten._convert_weight_to_int4pack's output layout, recovered
    from its inverse kernel, tensor_core_tiled_layout.cu:128-215.

    q int32 [N][K] -> int32 [N/8][K/(ikt*16)][32][ikt/2]."""
    n, k = q.shape
    ikt = inner_k_tiles
    assert n % 8 == 0 and k % (ikt * 16) == 0
    out = np.zeros((n // 8, k // (ikt * 16), 32, ikt // 2), dtype=np.uint32)
    t = np.arange(32)
    row_in_tile = t // 4
    for j in range(ikt // 2):
        for ko in range(k // (ikt * 16)):
            kbase0 = (ko * ikt + 2 * j) * 16                     # .cu:146
            ks0 = kbase0 + (t % 4) * 2                           # .cu:147
            ks = np.stack([ks0, ks0 + 1, ks0 + 8, ks0 + 9,       # .cu:148
                           ks0 + 16, ks0 + 17, ks0 + 24, ks0 + 25], axis=-1)  # .cu:151-153
            for nt in range(n // 8):
                rows = nt * 8 + row_in_tile                      # .cu:136
                vals = q[rows[:, None], ks]                      # [32][8]
                out[nt, ko, :, j] = _pack_words(vals)
    return out.view(np.int32)


def unpack_tinygemm(packed, inner_k_tiles=8):
    """torchao::unpack_tensor_core_tiled_layout semantics (tensor_core_tiled_layout.cu:197-215, :320-368)."""
    p = packed.view(np.uint32)
    ikt = inner_k_tiles
    nt_, ko_, _, half = p.shape
    assert half == ikt // 2
    n, k = nt_ * 8, ko_ * ikt * 16
    q = np.zeros((n, k), dtype=np.int32)
    t = np.arange(32)
    for j in range(half):
        for ko in range(ko_):
            kbase0 = (ko * ikt + 2 * j) * 16
            ks0 = kbase0 + (t % 4) * 2
            ks = np.stack([ks0, ks0 + 1, ks0 + 8, ks0 + 9,
                           ks0 + 16, ks0 + 17, ks0 + 24, ks0 + 25], axis=-1)
            for nt in range(nt_):
                rows = nt * 8 + t // 4
                q[rows[:, None], ks] = _unpack_words(p[nt, ko, :, j])
    return q


# ---- the B200 strip-major layout (this build's own; see DESIGN.md "Data layout in HBM") ----
# packed uint32 [N/16][K/64][32][4]: strip s (16 rows) x k-tile kt (64 k) = one 512-byte sub-tile; all sub-tiles of a
# strip are contiguous.  In a sub-tile lane l=(g=l>>2, c=l&3) owns 16 bytes; word j:
#   row = 16*s + g + 8*(j&1);  k0 = 64*kt + 32*(j>>1) + 8*c;  the word holds k0..k0+7 (nibble order of _pack_words).
def pack_b200(q):
    n, k = q.shape
    assert n % 16 == 0 and k % 64 == 0
    lane = np.arange(32)
    g, c = lane >> 2, lane & 3
    out = np.zeros((n // 16, k // 64, 32, 4), dtype=np.uint32)
    q5 = q.reshape(n // 16, 16, k // 64, 8, 8)        # [s][row][kt][chunk][8]
    for j in range(4):
        rows = g + 8 * (j & 1)
        chunks = c + 4 * (j >> 1)
        vals = q5[:, rows, :, chunks, :]               # -> [32][s][kt][8]
        out[:, :, :, j] = _pack_words(vals).transpose(1, 2, 0)
    return out.view(np.int32)


def unpack_b200(packed):
    p = packed.view(np.uint32)
    s_, kt_ = p.shape[0], p.shape[1]
    n, k = s_ * 16, kt_ * 64
    lane = np.arange(32)
    g, c = lane >> 2, lane & 3
    q5 = np.zeros((s_, 16, kt_, 8, 8), dtype=np.int32)
    for j in range(4):
        rows = g + 8 * (j & 1)
        chunks = c + 4 * (j >> 1)
        vals = _unpack_words(p[:, :, :, j])            # [s][kt][32][8]
        q5[:, rows, :, chunks, :] = vals.transpose(2, 0, 1, 3)
    return q5.reshape(n, k)


def pack_scales_and_zeros_b200(scale, zero):
    """scale, zero [N][K/G] -> strip-major [N/16][K/G][16][2]: the numbers of the reference's
    pack_tinygemm_scales_and_zeros (torchao/quantization/utils.py:395-409, [K/G][N][2]) with the 16 rows of a
    strip contiguous."""
    n, ng = scale.shape
    sz = np.stack([scale, zero], axis=-1)              # [N][NG][2]
    return np.ascontiguousarray(sz.reshape(n // 16, 16, ng, 2).transpose(0, 2, 1, 3))


def unpack_scales_and_zeros_b200(sz):
    s16, ng = sz.shape[0], sz.shape[1]
    t = sz.transpose(0, 2, 1, 3).reshape(s16 * 16, ng, 2)
    return np.ascontiguousarray(t[..., 0]), np.ascontiguousarray(t[..., 1])


# --------------------------------------------------------------------------
# linear
# --------------------------------------------------------------------------
def int4_weight_only_mm(x, q, scale, zero, group_size):
    """y[m][n] = bf16( sum_k fp32(x[m][k] * W'[n][k]) ), W' = dequantize_tinygemm.
    Semantics of aten._weight_int4pack_mm; summation order is
    unspecified there, here it is float64 (tolerance-compared, not bit-compared)."""
    wd = dequantize_tinygemm(q, scale, zero, group_size).astype(np.float64)
    return bf16_round((x.astype(np.float64) @ wd.T).astype(F32))


# --------------------------------------------------------------------------
# int8 weight-only
# --------------------------------------------------------------------------
def choose_qparams_int8_symmetric(w, eps=float(np.finfo(np.float32).eps), div_mode="ieee"):
    """Reference: quant_primitives.py:1537-1565 (SYMMETRIC, int8 -128..127, per row,
    scale dtype = input dtype bf16); caller quant_api.py:1223-1240.
    Returns scale [N] bf16-valued (zero_point is int64 zeros)."""
    wf = w.astype(F32)
    mn = wf.min(axis=-1)
    mx = wf.max(axis=-1)
    mn_neg = np.minimum(mn, F32(0))                    # :1540
    mx_pos = np.maximum(mx, F32(0))                    # :1541
    amax = np.maximum(-mn_neg, mx_pos)                 # :1549 (negation exact)
    scale = _div_scalar(amax, 127.5, div_mode)         # :1550  / (255/2)
    scale = bf16_round(np.maximum(scale, F32(eps)))    # :1565 clamp(min=eps)
    return scale


def quantize_int8_symmetric(w, scale):
    """Reference: quant_primitives.py:453-456: clamp(round(w * (1.0/scale)) + zp, -128, 127),
    zp = 0; `1.0/scale` and the product each round to bf16."""
    s = scale.reshape(-1, 1).astype(F32)
    inv = bf16_round(F32(1.0) / s)
    p = bf16_round(w.astype(F32) * inv)
    q = np.clip(np.rint(p), -128, 127)
    return q.astype(np.int8)


def int8_weight_only_mm(x, q, scale):
    """Reference: plain_layout.py:250-266:  y = bf16(bf16(x @ q^T) * scale)."""
    m = bf16_round((x.astype(np.float64) @ q.astype(np.float64).T).astype(F32))
    return bf16_round(m * scale.reshape(1, -1).astype(F32))


def dequantize_int8(q, scale):
    """Reference: quant_primitives.py:828-833 with zp=0: bf16(q) * scale."""
    return bf16_round(q.astype(F32) * scale.reshape(-1, 1).astype(F32))


# ---- the B200 strip-major int8 layout (this build's own; include/ao_b200.h) ----
# uint8 [N/16][K/64][2 halves][32 lanes][16]: biased bytes u = q + 128; word j (4 bytes) of lane l=(g=l>>2, c=l&3) in
# half h of tile (s, kt) holds row 16 s + g + 8 (j&1), k0 = 64 kt + 32 h + 8 c + 4 (j>>1), bytes k0..k0+3.
def pack_int8_b200(q):
    n, k = q.shape
    assert n % 16 == 0 and k % 64 == 0
    u = (q.astype(np.int16) + 128).astype(np.uint8)
    u6 = u.reshape(n // 16, 16, k // 64, 2, 4, 2, 4)   # [s][row][kt][h][c][j>>1][4]
    out = np.zeros((n // 16, k // 64, 2, 32, 4, 4), dtype=np.uint8)
    for lane in range(32):
        g, c = lane >> 2, lane & 3
        for j in range(4):
            out[:, :, :, lane, j, :] = u6[:, g + 8 * (j & 1), :, :, c, j >> 1, :]
    return out.reshape(n // 16, k // 64, 1024)


def unpack_int8_b200(packed):
    s_, kt_ = packed.shape[0], packed.shape[1]
    p = packed.reshape(s_, kt_, 2, 32, 4, 4)
    u6 = np.zeros((s_, 16, kt_, 2, 4, 2, 4), dtype=np.uint8)
    for lane in range(32):
        g, c = lane >> 2, lane & 3
        for j in range(4):
            u6[:, g + 8 * (j & 1), :, :, c, j >> 1, :] = p[:, :, :, lane, j, :]
    return (u6.reshape(s_ * 16, kt_ * 64).astype(np.int16) - 128).astype(np.int8)


# --------------------------------------------------------------------------
# HQQ (use_hqq=True): half-quadratic qparam search, then the same packed format
# --------------------------------------------------------------------------
def hqq_quantize(w, group_size, nbits=4, lp_norm=0.7, beta=10.0, kappa=1.01, iters=20):
    """quant_primitives.py:1998-2111 (_choose_qparams_and_quantize_affine_hqq, axis=1, optimize=True,
    raw_output=False, compute_dtype=bf16) with optimize_weights_proximal_legacy (:1915-1973) on its fp32 (CPU) path and
    _convert_to_affinequantized_format (:1981-1995).  ``w``: bf16-valued fp32 (N, K).  Returns q uint8 (N, K),
    scale / zero as bf16-valued fp32 (N, K/G) in the tinygemm convention  w' = (q - 8) * scale + zero.

    Parity status of THIS function: pinned to tests/golden/hqq_*.npz within a tolerance, not bit-for-bit — the
    reference's iteration uses torch's vectorised fp32 ``mean`` / ``pow`` whose last-ulp behaviour numpy does not
    reproduce; the scale (a pure min / max quantity) IS bit-exact, the optimised zero agrees to a bf16 ulp on almost every
    group and q is identical on > 99.9 % of the weights (tests/test_oracle_golden.py states the bounds)."""
    w = np.ascontiguousarray(w, dtype=F32)
    n, k = w.shape
    W = w.reshape(-1, group_size)
    max_v = F32(2 ** nbits - 1)
    mn = W.min(axis=1, keepdims=True)
    mx = W.max(axis=1, keepdims=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        scale = np.minimum(max_v / (mx - mn), F32(2e4)).astype(F32)  # .clamp(max=2e4): inf (constant group) -> 2e4
    zero = np.rint(-mn * scale).astype(F32)
    best = 1e4
    for _ in range(iters):
        Wq = np.clip(np.rint(W * scale + zero), 0, max_v).astype(F32)
        Wr = ((Wq - zero) / scale).astype(F32)
        d = (W - Wr).astype(F32)
        ad = np.abs(d)
        with np.errstate(divide="ignore", invalid="ignore"):
            shr = ad - F32(1.0 / beta) * np.power(ad, F32(lp_norm - 1)).astype(F32)
        We = (np.sign(d) * np.maximum(shr, F32(0))).astype(F32)  # |d| = 0: pow = inf, relu(-inf) = 0, sign = 0
        We[ad == 0] = 0
        zero = np.mean(Wq - (W - We) * scale, axis=1, keepdims=True, dtype=F32)
        beta *= kappa
        err = float(np.mean(np.abs(W - Wr), dtype=F32))
        if err < best:
            best = err
        else:
            break
    q = np.clip(np.rint(W * scale + zero), 0, max_v).astype(np.uint8).reshape(n, k)
    s_inv = (F32(1.0) / scale).astype(F32)
    zero_ao = ((F32(2 ** (nbits - 1)) - zero) * s_inv).astype(F32)
    return q, bf16_round(s_inv).reshape(n, -1), bf16_round(zero_ao).reshape(n, -1)


# --------------------------------------------------------------------------
# int8 KV cache (row N2): per-token symmetric absmax quantizer and the cache update
# --------------------------------------------------------------------------
def quantize_activation_per_token_absmax(x, div_mode="ieee"):
    """reference: torchao/quantization/utils.py:152-182 -> choose_qparams_affine SYMMETRIC (quant_primitives.py:1537-1575)
    + _quantize_affine_no_dtype_cast (:433-459) with quant_min = -127, quant_max = 127, eps = 1e-5, on bf16 rows
    (every op rounds to bf16).  x: bf16-valued fp32 [..., D] -> (q int8 [..., D], scale bf16-valued fp32 [...])."""
    x = np.ascontiguousarray(x, dtype=F32)
    min_val, max_val = x.min(axis=-1), x.max(axis=-1)
    max_val_pos = np.maximum(-np.minimum(min_val, F32(0)), np.maximum(max_val, F32(0)))
    scale = _div_scalar(max_val_pos, 127.0, div_mode)                      # max_val_pos / (float(quant_max - quant_min) / 2)
    scale = bf16_round(np.maximum(scale, F32(1e-5)))                        # torch.clamp(scale, min=eps), result in bf16
    recip = bf16_round(F32(1.0) / scale)                                    # 1.0 / scale == scale.reciprocal() * 1.0
    q = np.rint(bf16_round(x * recip[..., None]))                           # torch.round(input * (1.0 / scale)); zero_point = 0
    return np.clip(q, -127, 127).astype(np.int8), scale


def kv_cache_update(cache, cache_scale, pos, val, div_mode="ieee"):
    """reference: AffineQuantizedKVCache.update for ONE of k / v (torchao/_models/llama/model.py:215-232): writes the
    quantized rows and their scales at `pos` (in place) and returns the dequantized cache bf16(q * scale) with the rows at
    `pos` replaced by the unquantized values.  cache int8 [B][H][S][D], cache_scale bf16-valued fp32 [B][H][S][1],
    val bf16-valued fp32 [B][H][len(pos)][D]."""
    q, s = quantize_activation_per_token_absmax(val, div_mode)
    cache[:, :, pos] = q
    cache_scale[:, :, pos] = s[..., None]
    out = bf16_round(cache.astype(F32) * cache_scale)
    out[:, :, pos] = val
    return out


# --------------------------------------------------------------------------
# Int4Tensor (row N3): fbgemm-gpu-genai's int4_row_quantize_zp + pack_int4, restated
# (third-party, absent from /root/reference: fbgemm_gpu/experimental/gen_ai/quantize.py, >= 1.2.0 as required by
# torchao/quantization/quantize_/workflows/int4/int4_tensor.py:22-26,71-73) -- parity with fbgemm itself is UNPINNED.
# --------------------------------------------------------------------------
def int4_row_quantize_zp(w, group_size, div_mode="ieee"):
    """w bf16-valued fp32 [N][K] -> q int8 [N][K] in [-8, 7], scale / zero fp32 [K/G][N] (fp32 arithmetic throughout;
    the caller casts scale / zero to the weight's dtype, int4_tensor.py:94-95)."""
    n, k = w.shape
    x = np.ascontiguousarray(w, dtype=F32).reshape(-1, group_size)
    lo, hi = x.min(axis=1, keepdims=True), x.max(axis=1, keepdims=True)
    rng = np.maximum(hi - lo, F32(1e-6))
    scale = rng / F32(15) if div_mode == "ieee" else rng * (F32(1) / F32(15))
    zero = lo + scale * F32(8)
    q = (np.clip(np.rint((x - lo) / scale), 0, 15) - 8).astype(np.int8).reshape(n, k)
    return q, scale.reshape(n, -1).T.copy(), zero.reshape(n, -1).T.copy()


def pack_int4(q):
    """int8 [N][K] in [-8, 7] -> int8 [N][K/2]: even k in the low nibble (fbgemm pack_int4)."""
    return ((q[:, 0::2] & 0xF) | (q[:, 1::2] << 4)).astype(np.int8)


def sqnr_db(ref, x):
    """Reference: torchao/quantization/utils.py:51-55 compute_error."""
    ref = ref.astype(np.float64)
    x = x.astype(np.float64)
    ps = np.linalg.norm(ref)
    pn = np.linalg.norm(ref - x)
    if pn == 0:
        return float("inf")
    return float(20 * np.log10(ps / pn))
