"""Pin the CPU oracle (oracle/ao_oracle.py) against golden vectors produced by the
reference's own functions (tests/golden/make_golden.py).  Bit-exact."""
import glob
import os

import numpy as np
import pytest

import ao_oracle as o

from conftest import GOLDEN

INT4 = sorted(glob.glob(os.path.join(GOLDEN, "int4_*_g*.npz")))
INT8 = sorted(glob.glob(os.path.join(GOLDEN, "int8_*.npz")))


def test_golden_present():
    assert len(INT4) >= 6 and len(INT8) >= 3


@pytest.mark.parametrize("path", INT4, ids=os.path.basename)
def test_int4_qparams_quantize_bit_exact(path):
    g = np.load(path)
    gs = int(g["group_size"])
    w = o.from_bits(g["w"])
    s, z = o.choose_qparams_tinygemm(w, gs, div_mode="ieee")
    assert np.array_equal(o.to_bits(s), g["scale"])
    assert np.array_equal(o.to_bits(z), g["zero"])
    q = o.quantize_tinygemm(w, s, z, gs)
    assert np.array_equal(q.astype(np.uint8), g["q"])
    assert np.array_equal(o.to_bits(o.pack_scales_and_zeros(s, z)), g["scale_and_zero"])
    s2, z2 = o.unpack_scales_and_zeros(o.from_bits(g["scale_and_zero"]))
    assert np.array_equal(s2, s) and np.array_equal(z2, z)


@pytest.mark.parametrize("path", INT4, ids=os.path.basename)
def test_int4_python_dequant_bit_exact(path):
    g = np.load(path)
    gs = int(g["group_size"])
    dq = o.dequantize_affine_tinygemm_2round(
        g["q"].astype(np.int32), o.from_bits(g["scale"]), o.from_bits(g["zero"]), gs)
    assert np.array_equal(o.to_bits(dq), g["dequant_2round"])


@pytest.mark.parametrize("path", INT4, ids=os.path.basename)
def test_int4_fma_dequant_vs_cpu_kernel(path):
    """aten._weight_int4pack_mm_for_cpu dequantises (q-8)*s+z in fp32 and rounds the
    identity-matmul result once: our single-rounding fma dequant must agree except
    where fp32 double rounding bites (never, for these vectors, beyond 1 bf16 ulp)."""
    g = np.load(path)
    gs = int(g["group_size"])
    dq = o.dequantize_tinygemm(g["q"].astype(np.int32), o.from_bits(g["scale"]), o.from_bits(g["zero"]), gs)
    ref = o.from_bits(g["dequant_cpu_mm"])
    a, b = o.to_bits(dq).astype(np.int32), g["dequant_cpu_mm"].astype(np.int32)
    # equal up to 1 ulp, and equal nearly everywhere
    finite = np.isfinite(ref)
    assert np.abs(a - b)[finite].max() <= 1
    assert (a == b)[finite].mean() > 0.999


@pytest.mark.parametrize("path", INT4, ids=os.path.basename)
def test_int4_mm_vs_cpu_reference(path):
    g = np.load(path)
    if "edge" in path:
        pytest.skip("edge weights hold 1e4-scale rows; covered by dequant tests")
    gs = int(g["group_size"])
    y = o.int4_weight_only_mm(o.from_bits(g["x"]), g["q"].astype(np.int32),
                              o.from_bits(g["scale"]), o.from_bits(g["zero"]), gs)
    ref = o.from_bits(g["y_cpu"])
    assert o.sqnr_db(ref, y) >= 40.0


def test_int4_padded_case():
    g = np.load(os.path.join(GOLDEN, "int4_padded_321x256.npz"))
    gs = int(g["group_size"])
    w = o.from_bits(g["w"])
    n, k = [int(v) for v in g["padded_shape"]]
    wp = np.zeros((n, k), dtype=np.float32)
    wp[: w.shape[0], : w.shape[1]] = w
    s, z = o.choose_qparams_tinygemm(wp, gs)
    q = o.quantize_tinygemm(wp, s, z, gs)
    assert np.array_equal(o.to_bits(s), g["scale"])
    assert np.array_equal(o.to_bits(z), g["zero"])
    assert np.array_equal(q.astype(np.uint8), g["q"])


@pytest.mark.parametrize("path", INT8, ids=os.path.basename)
def test_int8_bit_exact(path):
    g = np.load(path)
    w = o.from_bits(g["w"])
    s = o.choose_qparams_int8_symmetric(w, div_mode="ieee")
    assert np.array_equal(o.to_bits(s), g["scale"])
    q = o.quantize_int8_symmetric(w, s)
    assert np.array_equal(q, g["q"])
    assert np.array_equal(o.to_bits(o.dequantize_int8(q, s)), g["dequant"])


@pytest.mark.parametrize("path", INT8, ids=os.path.basename)
def test_int8_mm_vs_cpu_reference(path):
    g = np.load(path)
    y = o.int8_weight_only_mm(o.from_bits(g["x"]), g["q"], o.from_bits(g["scale"]))
    ref = o.from_bits(g["y_cpu"])
    assert o.sqnr_db(ref, y) >= 40.0


@pytest.mark.parametrize("shape", [(128, 64), (256, 256), (384, 1024)])
def test_layout_round_trips(shape):
    rng = np.random.default_rng(shape[1])
    q = rng.integers(0, 16, shape).astype(np.int32)
    p = o.pack_b200(q)
    assert p.shape == (shape[0] // 16, shape[1] // 64, 32, 4)
    assert np.array_equal(o.unpack_b200(p), q)
    # the layout definition of include/ao_b200.h, spelled out element by element
    pu = p.view(np.uint32)
    for (st, kt, lane, j) in [(0, 0, 0, 0), (shape[0] // 16 - 1, shape[1] // 64 - 1, 31, 3), (5, 0, 13, 2)]:
        g, c = lane >> 2, lane & 3
        row, k0 = 16 * st + g + 8 * (j & 1), 64 * kt + 32 * (j >> 1) + 8 * c
        w = int(pu[st, kt, lane, j])
        assert [(w >> (4 * s)) & 15 for s in (0, 4, 1, 5, 2, 6, 3, 7)] == list(q[row, k0:k0 + 8])
    # strip-major qparams are a pure permutation of the reference's (K/G, N, 2) tensor
    sc = rng.standard_normal((shape[0], shape[1] // 32)).astype(np.float32)
    zr = rng.standard_normal((shape[0], shape[1] // 32)).astype(np.float32)
    szb = o.pack_scales_and_zeros_b200(sc, zr)
    assert szb.shape == (shape[0] // 16, shape[1] // 32, 16, 2)
    s2, z2 = o.unpack_scales_and_zeros_b200(szb)
    assert np.array_equal(s2, sc) and np.array_equal(z2, zr)
    ref = o.pack_scales_and_zeros(sc, zr)
    assert np.array_equal(szb[3, 1, 7], ref[1, 3 * 16 + 7])
    if shape[1] % 128 == 0:
        for ikt in (2, 4, 8):
            assert np.array_equal(o.unpack_tinygemm(o.pack_tinygemm(q, ikt), ikt), q)


def test_recip_mode_differs_rarely():
    rng = np.random.default_rng(0)
    w = o.bf16_round(rng.standard_normal((256, 4096)).astype(np.float32))
    s0, _ = o.choose_qparams_tinygemm(w, 128, div_mode="ieee")
    s1, _ = o.choose_qparams_tinygemm(w, 128, div_mode="recip")
    assert (s0 != s1).mean() < 1e-2


def test_int8_b200_layout_round_trip_and_word_contract():
    rng = np.random.default_rng(7)
    q = rng.integers(-128, 128, size=(48, 192), dtype=np.int8)
    p = o.pack_int8_b200(q)
    assert p.shape == (3, 3, 1024) and p.dtype == np.uint8
    assert np.array_equal(o.unpack_int8_b200(p), q)
    # the contract the kernels rely on: half h, lane (g, c), word j -> row g + 8 (j & 1), k0 = 32 h + 8 c + 4 (j >> 1)
    s, kt, h, g, c, j = 1, 2, 1, 5, 3, 2
    word = p[s, kt].reshape(2, 32, 4, 4)[h, 4 * g + c, j]
    row, k0 = 16 * s + g + 8 * (j & 1), 64 * kt + 32 * h + 8 * c + 4 * (j >> 1)
    assert np.array_equal(word.astype(np.int16) - 128, q[row, k0:k0 + 4].astype(np.int16))


# ------------------------------------------------------------------------------------------------
# HQQ: tolerance-pinned, see the docstring of ao_oracle.hqq_quantize
# ------------------------------------------------------------------------------------------------
HQQ = sorted(glob.glob(os.path.join(GOLDEN, "hqq_*.npz")))


def hqq_agreement(q, s, z, g):
    """(fraction of identical q, max |dq|, scale bit-exact?, fraction of groups whose zero is within one bf16 ulp or
    1/128 of the group's scale of the golden zero)"""
    gq, gs_, gz = g["q"], o.from_bits(g["scale"]), o.from_bits(g["zero"])
    dz = np.abs(z - gz)
    tol = np.maximum(np.abs(gz), np.abs(gs_)) * 2.0 ** -7
    return ((q == gq).mean(), int(np.abs(q.astype(np.int32) - gq.astype(np.int32)).max()),
            np.array_equal(o.to_bits(s), g["scale"]), (dz <= tol).mean())


def test_hqq_golden_present():
    assert len(HQQ) >= 4


@pytest.mark.parametrize("path", HQQ, ids=os.path.basename)
def test_hqq_oracle_matches_reference_within_tolerance(path):
    g = np.load(path)
    q, s, z = o.hqq_quantize(o.from_bits(g["w"]), int(g["group_size"]))
    same_q, max_dq, scale_exact, zero_ok = hqq_agreement(q, s, z, g)
    assert scale_exact                      # 1 / (15 / (max - min)) has no iteration in it
    assert same_q >= 0.999 and max_dq <= 1  # a zero one ulp off moves a weight sitting on a rounding boundary by one step
    assert zero_ok >= 0.99


def test_hqq_lowers_the_lp_error_of_heavy_tailed_groups():
    # what the solver is for: || W - dequant ||_p^p with p = 0.7 must not be worse than min / max (tinygemm) qparams
    g = np.load(os.path.join(GOLDEN, "hqq_heavy_24x512_g256.npz"))
    gs = int(g["group_size"])
    w = o.from_bits(g["w"])
    q, s, z = o.hqq_quantize(w, gs)
    s0, z0 = o.choose_qparams_tinygemm(w, gs)
    q0 = o.quantize_tinygemm(w, s0, z0, gs)
    rep = lambda a: np.repeat(a, gs, axis=1)
    e_hqq = np.abs(w - ((q.astype(np.float32) - 8) * rep(s) + rep(z))) ** 0.7
    e_mm = np.abs(w - ((q0.astype(np.float32) - 8) * rep(s0) + rep(z0))) ** 0.7
    assert e_hqq.mean() < e_mm.mean()
