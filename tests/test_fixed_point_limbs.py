"""The activation staging of the integer-MMA decode kernels (csrc/gemv_int4_imma.cu, csrc/decode_mega.cu `emit_limbs`),
restated in numpy and checked exhaustively over bf16: a row is scaled by the power of two that brings its largest
magnitude into [2^21, 2^22), rounded to an integer X by the float-magic add, and split into three BALANCED base-256
digits without carries:  X + 0x808080 = sum (a_i + 128) 256^i, so the bytes of (bits + 0x00408080) xor 0x80 are the
two's-complement limbs a_i in [-128, 127] with  X = a_0 + 256 a_1 + 65536 a_2.  The kernels rely on (1) that identity
for every bf16 value below the row bound, (2) X being exactly x * 2^e whenever x is within 2^-15 of the bound (bf16 has
8 significant bits), and (3) |limb sums| staying small enough for exact int32 accumulation."""
import numpy as np


def bf16_values():
    bits = np.arange(0, 1 << 16, dtype=np.uint32)
    x = (bits << 16).view(np.float32)
    return x[np.isfinite(x)]


def limbs_of(x, up):
    """emit_limbs: v = bits(fma(x, up, 1.5 * 2^23)) + 0x00408080; limb l = byte l of v, xor 0x80 (as int8)."""
    v = (np.float32(x) * np.float32(up) + np.float32(12582912.0)).astype(np.float32)      # exact: |x * up| <= 2^22
    u = v.view(np.uint32) + np.uint32(0x00408080)
    b = [((u >> (8 * l)) & 0xFF).astype(np.uint8) ^ np.uint8(0x80) for l in range(3)]
    return [bb.view(np.int8).astype(np.int64) for bb in b]


def test_limbs_reconstruct_the_rounded_fixed_point_value_for_every_bf16_below_the_bound():
    x_all = bf16_values()
    for bound_exp in (-20, -3, 0, 7, 30):                                  # row bound in [2^bound_exp, 2^(bound_exp + 1))
        bound = np.float32(1.99) * np.float32(2.0) ** bound_exp
        e = 21 - bound_exp                                                  # the kernels: e = 21 - exponent(bound)
        up = np.float32(2.0) ** e
        x = x_all[np.abs(x_all) <= bound]
        a0, a1, a2 = limbs_of(x, up)
        X = a0 + 256 * a1 + 65536 * a2
        want = np.rint(x.astype(np.float64) * float(up)).astype(np.int64)   # round half to even, like the fp32 magic add
        assert np.array_equal(X, want)
        assert np.abs(X).max() <= 2 ** 22 and all(np.abs(a).max() <= 128 for a in (a0, a1, a2))
        # exact (no rounding at all) for every value within 2^-14 of the bound's binade: 8 significant bits + 14 spare
        big = np.abs(x) >= np.float32(2.0) ** (bound_exp - 14)
        assert np.array_equal(X[big].astype(np.float64), x[big].astype(np.float64) * float(up))


def test_group_sums_fit_exact_int32_accumulation():
    # one MMA segment = 128 k: |sum of limb * nibble| <= 128 * 128 * 15 < 2^18; the accumulators start at the integer image of
    # 1.5 * 2^23 and a float conversion is one subtraction as long as the sum stays below 2^22 in magnitude
    assert 128 * 128 * 15 < 2 ** 18 < 2 ** 22
    # segment sums of the limbs themselves (the (z - 8 s) * S term): |S| <= 128 * 128 = 2^14, two of them travel in one
    # 32-bit register as (limb1 << 16) + limb0 in the staging shuffles
    s0, s1 = -(2 ** 14), 2 ** 14
    packed = s1 * 65536 + s0
    lo = ((packed & 0xFFFF) ^ 0x8000) - 0x8000
    assert lo == s0 and (packed - lo) >> 16 == s1
