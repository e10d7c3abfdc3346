"""torch.library registration of the B200 kernels under the reference's `torchao` namespace.

Mirrors torchao/ops.py: `lib = torch.library.Library("torchao", "FRAGMENT")`, `lib.define(schema)`
(ops.py:12-21), fake/meta impls via `torch.library.register_fake` (ops.py:73-77, 274-296, 328-377);
the CUDA impls, which in the reference are `TORCH_LIBRARY_IMPL(torchao, CUDA, m)` in C++
(tensor_core_tiled_layout.cu:370-374), are Python functions that pass raw device pointers and the
current CUDA stream to the C ABI of libao_b200.so (include/ao_b200.h).  There is no CPU impl.
"""
from typing import Tuple

import torch
from torch import Tensor

from . import _lib

lib = torch.library.Library("torchao", "FRAGMENT")

_SCHEMAS = {
    # kept verbatim from the reference (ops.py:16-21)
    "unpack_tensor_core_tiled_layout": "(Tensor packed_w, int inner_k_tiles) -> Tensor",
    "dequantize_tensor_core_tiled_layout": "(Tensor packed_w, Tensor scales_and_zeros, int group_size, int inner_k_tiles) -> Tensor",
    # new ops of the B200 path
    "b200_int4_quantize_pack": "(Tensor w, int group_size, float eps, bool cuda_div_semantics) -> (Tensor, Tensor)",
    "b200_int4_choose_qparams": "(Tensor w, int group_size, float eps, bool cuda_div_semantics) -> (Tensor, Tensor)",
    "b200_int4_quantize": "(Tensor w, Tensor scale, Tensor zero, int group_size) -> Tensor",
    "b200_int4_pack": "(Tensor int_data) -> Tensor",
    "b200_int4_hqq_quantize": "(Tensor w, int group_size) -> (Tensor, Tensor, Tensor)",
    "b200_int4_unpack": "(Tensor packed_w) -> Tensor",
    "b200_int4_dequantize": "(Tensor packed_w, Tensor scales_and_zeros, int group_size) -> Tensor",
    "b200_tinygemm_pack": "(Tensor int_data, int inner_k_tiles) -> Tensor",
    "b200_int4_weight_only_mm": "(Tensor x, Tensor packed_w, int group_size, Tensor scales_and_zeros, int out_features) -> Tensor",
    "b200_int8_quantize": "(Tensor w, float eps, bool cuda_div_semantics) -> (Tensor, Tensor)",
    "b200_int8_pack": "(Tensor int_data) -> Tensor",
    "b200_int8_unpack": "(Tensor packed_w, int out_features, int in_features) -> Tensor",
    "b200_int8_weight_only_mm": "(Tensor x, Tensor packed_w, Tensor scale) -> Tensor",
    "b200_int8_gemv_fused": "(Tensor x, Tensor packed_w, Tensor scale, Tensor? norm_weight, float norm_eps, "
                            "Tensor? residual, int epilogue) -> Tensor",
}


def _defined(name):
    try:
        getattr(torch.ops.torchao, name)
        return True
    except (AttributeError, RuntimeError):
        return False


for _name, _schema in _SCHEMAS.items():
    # if the reference's torchao/ops.py was imported first its two schemas already exist
    if not _defined(_name):
        lib.define(_name + _schema)


def register_custom_op(name):
    def decorator(func):
        return torch.library.register_fake(f"{name}")(func)

    return decorator


def _impl_cuda(name):
    def decorator(func):
        lib.impl(name, func, "CUDA")
        return func

    return decorator


def _stream(t: Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream


def _round_up(a, b):
    return (a + b - 1) // b * b


def padded_int4_shape(n: int, k: int, group_size: int) -> Tuple[int, int]:
    """Super-tile layout needs N % 128 == 0 and K % max(G, 64) == 0 (the reference pads to N % 8, K % 1024,
    tensor_core_tiled_layout.py:127-135; padding only ever adds whole all-zero groups / all-zero rows)."""
    return _round_up(n, 128), _round_up(k, max(group_size, 64))


def _check_cuda_bf16(t, what):
    torch._check(t.is_cuda, lambda: f"{what} must be a CUDA tensor (the B200 path has no CPU fallback)")
    torch._check(t.dtype is torch.bfloat16, lambda: f"{what} must be bfloat16, got {t.dtype}")


def _check_group(group_size):
    torch._check(group_size in (32, 64, 128, 256), lambda: "qGroupSize must be 32, 64, 128, or 256")


# ------------------------------------------------------------------ reference-schema ops
def unpack_tensor_core_tiled_layout(packed_w: Tensor, inner_k_tiles: int) -> Tensor:
    """Inverse of aten._convert_weight_to_int4pack: (N/8, K/(ikt*16), 32, ikt/2) int32 -> (N, K) int32.
    Reference wrapper: torchao/ops.py:255-272."""
    return torch.ops.torchao.unpack_tensor_core_tiled_layout.default(packed_w=packed_w, inner_k_tiles=inner_k_tiles)


def _tinygemm_shape_checks(packed_w, inner_k_tiles):
    torch._check(packed_w.dim() == 4, lambda: f"packed weight should be a 4d tensor, got {packed_w.dim()}D")
    torch._check(packed_w.dtype is torch.int32, lambda: f"weight must be INT32, got {packed_w.dtype}")
    torch._check(inner_k_tiles in (2, 4, 8), lambda: "inner_k_tiles must be 2, 4, or 8")
    torch._check(packed_w.size(2) == 32, lambda: "packed weight must have 32 at dim 2")
    torch._check(packed_w.size(3) == inner_k_tiles // 2, lambda: "packed weight must have inner_k_tiles/2 at dim 3")
    return packed_w.size(0) * 8, packed_w.size(1) * inner_k_tiles * 16


@register_custom_op("torchao::unpack_tensor_core_tiled_layout")
def _(packed_w: Tensor, inner_k_tiles: int) -> Tensor:
    n, k = _tinygemm_shape_checks(packed_w, inner_k_tiles)
    return torch.empty((n, k), dtype=torch.int32, device=packed_w.device)


@_impl_cuda("unpack_tensor_core_tiled_layout")
def _unpack_tinygemm_cuda(packed_w: Tensor, inner_k_tiles: int) -> Tensor:
    n, k = _tinygemm_shape_checks(packed_w, inner_k_tiles)
    torch._check(packed_w.is_contiguous(), lambda: "packed weight must be contiguous")
    out = torch.empty((n, k), dtype=torch.int32, device=packed_w.device)
    with torch.cuda.device(packed_w.device):
        _lib.check(_lib.lib().ao_b200_tinygemm_unpack(packed_w.data_ptr(), n, k, inner_k_tiles, out.data_ptr(),
                                                      _stream(packed_w)), "tinygemm_unpack")
    return out


def dequantize_tensor_core_tiled_layout(packed_w: Tensor, scales_and_zeros: Tensor, group_size: int,
                                        inner_k_tiles: int) -> Tensor:
    """tinygemm-packed int4 + (K/G, N, 2) bf16 scales_and_zeros -> (N, K) bf16, with the single-rounding
    bf16 fma dequant of aten._weight_int4pack_mm.  Reference wrapper: torchao/ops.py:299-325."""
    return torch.ops.torchao.dequantize_tensor_core_tiled_layout.default(packed_w, scales_and_zeros, group_size,
                                                                         inner_k_tiles)


def _sz_checks(scales_and_zeros, n, k, group_size):
    torch._check(scales_and_zeros.dtype is torch.bfloat16, lambda: "scales_and_zeros must be bfloat16")
    torch._check(scales_and_zeros.dim() == 3, lambda: f"scales_and_zeros must be 3D, got {scales_and_zeros.dim()}")
    _check_group(group_size)
    torch._check(scales_and_zeros.size(0) == k // group_size, lambda: "scales_and_zeros must have K // qGroupSize at dim 0")
    torch._check(scales_and_zeros.size(1) == n, lambda: "scales_and_zeros must have N at dim 1")
    torch._check(scales_and_zeros.size(2) == 2, lambda: "scales_and_zeros must have 2 at dim 2")


@register_custom_op("torchao::dequantize_tensor_core_tiled_layout")
def _(packed_w: Tensor, scales_and_zeros: Tensor, group_size: int, inner_k_tiles: int) -> Tensor:
    n, k = _tinygemm_shape_checks(packed_w, inner_k_tiles)
    _sz_checks(scales_and_zeros, n, k, group_size)
    return torch.empty((n, k), dtype=torch.bfloat16, device=packed_w.device)


@_impl_cuda("dequantize_tensor_core_tiled_layout")
def _dequant_tinygemm_cuda(packed_w: Tensor, scales_and_zeros: Tensor, group_size: int, inner_k_tiles: int) -> Tensor:
    n, k = _tinygemm_shape_checks(packed_w, inner_k_tiles)
    _sz_checks(scales_and_zeros, n, k, group_size)
    torch._check(packed_w.is_contiguous() and scales_and_zeros.is_contiguous(), lambda: "inputs must be contiguous")
    out = torch.empty((n, k), dtype=torch.bfloat16, device=packed_w.device)
    with torch.cuda.device(packed_w.device):
        _lib.check(_lib.lib().ao_b200_tinygemm_dequant(packed_w.data_ptr(), scales_and_zeros.data_ptr(), n, k,
                                                       group_size, inner_k_tiles, out.data_ptr(), _stream(packed_w)),
                   "tinygemm_dequant")
    return out


# ------------------------------------------------------------------ B200 int4 ops
def _w_checks(w, group_size):
    torch._check(w.dim() == 2, lambda: f"weight must be 2-D, got {w.dim()}D")
    _check_cuda_bf16(w, "weight")
    _check_group(group_size)
    torch._check(w.size(1) % group_size == 0, lambda: f"in_features {w.size(1)} not divisible by group_size {group_size}")


def _rows(w):
    """(tensor, leading dimension) for a 2-D bf16 tensor whose rows are contiguous and 16-byte aligned."""
    if w.stride(1) != 1 or w.stride(0) % 8 != 0 or w.stride(0) < w.size(1) or w.data_ptr() % 16 != 0:
        w = w.contiguous()
        if w.stride(0) % 8 != 0:
            w = torch.nn.functional.pad(w, (0, (-w.size(1)) % 8))[:, : w.size(1)]
    return w, w.stride(0)


@register_custom_op("torchao::b200_int4_quantize_pack")
def _(w: Tensor, group_size: int, eps: float, cuda_div_semantics: bool):
    _check_group(group_size)
    torch._check(w.dim() == 2, lambda: "weight must be 2-D")
    np_, kp = padded_int4_shape(w.size(0), w.size(1), group_size)
    return (torch.empty((np_ // 16, kp // 64, 32, 4), dtype=torch.int32, device=w.device),
            torch.empty((np_ // 16, kp // group_size, 16, 2), dtype=torch.bfloat16, device=w.device))


@_impl_cuda("b200_int4_quantize_pack")
def _int4_quantize_pack_cuda(w: Tensor, group_size: int, eps: float, cuda_div_semantics: bool):
    _w_checks(w, group_size)
    n, k = w.shape
    np_, kp = padded_int4_shape(n, k, group_size)
    w, ldw = _rows(w)
    packed = torch.empty((np_ // 16, kp // 64, 32, 4), dtype=torch.int32, device=w.device)
    sz = torch.empty((np_ // 16, kp // group_size, 16, 2), dtype=torch.bfloat16, device=w.device)
    with torch.cuda.device(w.device):
        _lib.check(_lib.lib().ao_b200_int4_quantize_pack(
            w.data_ptr(), n, k, ldw, np_, kp, group_size, eps,
            _lib.DIV_RECIP if cuda_div_semantics else _lib.DIV_IEEE,
            packed.data_ptr(), sz.data_ptr(), _stream(w)), "int4_quantize_pack")
    return packed, sz


@register_custom_op("torchao::b200_int4_choose_qparams")
def _(w: Tensor, group_size: int, eps: float, cuda_div_semantics: bool):
    shape = (w.size(0), w.size(1) // group_size)
    return (torch.empty(shape, dtype=torch.bfloat16, device=w.device),
            torch.empty(shape, dtype=torch.bfloat16, device=w.device))


@_impl_cuda("b200_int4_choose_qparams")
def _int4_choose_qparams_cuda(w: Tensor, group_size: int, eps: float, cuda_div_semantics: bool):
    torch._check(w.dim() == 2 and w.size(1) % group_size == 0, lambda: "bad weight shape for group_size")
    _check_cuda_bf16(w, "weight")
    w = w.contiguous()
    n, k = w.shape
    scale = torch.empty((n, k // group_size), dtype=torch.bfloat16, device=w.device)
    zero = torch.empty_like(scale)
    with torch.cuda.device(w.device):
        _lib.check(_lib.lib().ao_b200_int4_choose_qparams(
            w.data_ptr(), n, k, k, group_size, eps, _lib.DIV_RECIP if cuda_div_semantics else _lib.DIV_IEEE,
            scale.data_ptr(), zero.data_ptr(), _stream(w)), "int4_choose_qparams")
    return scale, zero


@register_custom_op("torchao::b200_int4_quantize")
def _(w: Tensor, scale: Tensor, zero: Tensor, group_size: int):
    return torch.empty(w.shape, dtype=torch.int32, device=w.device)


@_impl_cuda("b200_int4_quantize")
def _int4_quantize_cuda(w: Tensor, scale: Tensor, zero: Tensor, group_size: int):
    torch._check(w.dim() == 2 and w.size(1) % group_size == 0, lambda: "bad weight shape for group_size")
    _check_cuda_bf16(w, "weight")
    _check_cuda_bf16(scale, "scale")
    _check_cuda_bf16(zero, "zero_point")
    n, k = w.shape
    torch._check(scale.numel() == n * (k // group_size) and zero.numel() == scale.numel(), lambda: "qparam size mismatch")
    w, scale, zero = w.contiguous(), scale.contiguous(), zero.contiguous()
    q = torch.empty((n, k), dtype=torch.int32, device=w.device)
    with torch.cuda.device(w.device):
        _lib.check(_lib.lib().ao_b200_int4_quantize(w.data_ptr(), n, k, k, group_size, scale.data_ptr(),
                                                    zero.data_ptr(), q.data_ptr(), _stream(w)), "int4_quantize")
    return q


# HQQ solver constants: optimize_weights_proximal_legacy's opt_params (quant_primitives.py:1924-1930)
HQQ_LP_NORM, HQQ_BETA, HQQ_KAPPA, HQQ_ITERS = 0.7, 1e1, 1.01, 20


@register_custom_op("torchao::b200_int4_hqq_quantize")
def _(w: Tensor, group_size: int):
    shape = (w.size(0), w.size(1) // group_size)
    return (torch.empty(w.shape, dtype=torch.int32, device=w.device),
            torch.empty(shape, dtype=torch.bfloat16, device=w.device),
            torch.empty(shape, dtype=torch.bfloat16, device=w.device))


@_impl_cuda("b200_int4_hqq_quantize")
def _int4_hqq_quantize_cuda(w: Tensor, group_size: int):
    """`_choose_qparams_and_quantize_affine_hqq(w, nbits=4, group_size, axis=1, compute_dtype=bf16, raw_output=False)`
    (quant_primitives.py:1998-2111): q int32 (N, K) in 0..15, scale / zero bf16 (N, K/G) in the tinygemm convention.
    One kernel per solver iteration plus a one-block kernel that applies the reference's global early stop (a mean over
    the whole tensor) on the device: the launches after the stop return at once, the host never waits."""
    _w_checks(w, group_size)
    w = w.contiguous()
    n, k = w.shape
    n_groups = n * (k // group_size)
    L = _lib.lib()
    scale = torch.empty(n_groups, dtype=torch.float32, device=w.device)
    zero = torch.empty_like(scale)
    err = torch.empty(int(L.ao_b200_hqq_blocks(n_groups)), dtype=torch.float32, device=w.device)
    q = torch.empty((n, k), dtype=torch.int32, device=w.device)
    scale_out = torch.empty((n, k // group_size), dtype=torch.bfloat16, device=w.device)
    zero_out = torch.empty_like(scale_out)
    with torch.cuda.device(w.device):
        st = _stream(w)
        _lib.check(L.ao_b200_hqq_init(w.data_ptr(), n_groups, group_size, scale.data_ptr(), zero.data_ptr(), st), "hqq_init")
        # early_stop=True (:1968-1972) is decided on the device: no host round trip per iteration
        state = torch.empty(2, dtype=torch.float64, device=w.device)
        _lib.check(L.ao_b200_hqq_solve(w.data_ptr(), n_groups, group_size, scale.data_ptr(), zero.data_ptr(), HQQ_BETA,
                                       HQQ_KAPPA, HQQ_LP_NORM, HQQ_ITERS, err.data_ptr(), state.data_ptr(), st), "hqq_solve")
        _lib.check(L.ao_b200_hqq_finalize(w.data_ptr(), n_groups, group_size, scale.data_ptr(), zero.data_ptr(),
                                          q.data_ptr(), scale_out.data_ptr(), zero_out.data_ptr(), st), "hqq_finalize")
    return q, scale_out, zero_out


@register_custom_op("torchao::b200_int4_pack")
def _(int_data: Tensor):
    np_, kp = _round_up(int_data.size(0), 128), _round_up(int_data.size(1), 64)
    return torch.empty((np_ // 16, kp // 64, 32, 4), dtype=torch.int32, device=int_data.device)


@_impl_cuda("b200_int4_pack")
def _int4_pack_cuda(int_data: Tensor):
    torch._check(int_data.dim() == 2 and int_data.dtype is torch.int32, lambda: "int_data must be 2-D int32")
    int_data = int_data.contiguous()
    n, k = int_data.shape
    np_, kp = _round_up(n, 128), _round_up(k, 64)
    packed = torch.empty((np_ // 16, kp // 64, 32, 4), dtype=torch.int32, device=int_data.device)
    with torch.cuda.device(int_data.device):
        _lib.check(_lib.lib().ao_b200_int4_pack(int_data.data_ptr(), n, k, np_, kp, 0, packed.data_ptr(),
                                                _stream(int_data)), "int4_pack")
    return packed


def _strip_shape_checks(packed_w):
    torch._check(packed_w.dim() == 4 and tuple(packed_w.shape[2:]) == (32, 4),
                 lambda: f"strip-major weight must be (N/16, K/64, 32, 4), got {tuple(packed_w.shape)}")
    torch._check(packed_w.dtype is torch.int32, lambda: f"weight must be INT32, got {packed_w.dtype}")
    return packed_w.size(0) * 16, packed_w.size(1) * 64


def _sz_checks_b200(scales_and_zeros, n, k, group_size):
    """strip-major scale/zero: bf16 (N/16, K/G, 16, 2)"""
    torch._check(scales_and_zeros.dtype is torch.bfloat16, lambda: "scales_and_zeros must be bfloat16")
    _check_group(group_size)
    torch._check(tuple(scales_and_zeros.shape) == (n // 16, k // group_size, 16, 2),
                 lambda: f"scales_and_zeros must be (N/16, K/G, 16, 2) = {(n // 16, k // group_size, 16, 2)}, "
                         f"got {tuple(scales_and_zeros.shape)}")


def pack_scales_and_zeros_b200(scale: Tensor, zero: Tensor) -> Tensor:
    """scale, zero (N, K/G) -> strip-major bf16 (N/16, K/G, 16, 2); N % 16 == 0.  (The reference's
    pack_tinygemm_scales_and_zeros, torchao/quantization/utils.py:395-409, gives (K/G, N, 2): same numbers, the rows
    of one 16-row strip made contiguous so that a strip's qparams are one bulk copy.)"""
    n, ng = scale.shape
    sz = torch.stack([scale.to(torch.bfloat16), zero.to(torch.bfloat16)], dim=-1)  # (N, NG, 2)
    return sz.reshape(n // 16, 16, ng, 2).permute(0, 2, 1, 3).contiguous()


def unpack_scales_and_zeros_b200(sz: Tensor) -> Tuple[Tensor, Tensor]:
    """strip-major (N/16, K/G, 16, 2) -> scale, zero (N, K/G)"""
    s16, ng = sz.shape[0], sz.shape[1]
    t = sz.permute(0, 2, 1, 3).reshape(s16 * 16, ng, 2)
    return t[..., 0].contiguous(), t[..., 1].contiguous()


def sz_b200_to_tinygemm(sz: Tensor) -> Tensor:
    """strip-major (N/16, K/G, 16, 2) -> the reference's (K/G, N, 2)"""
    s16, ng = sz.shape[0], sz.shape[1]
    return sz.permute(1, 0, 2, 3).reshape(ng, s16 * 16, 2).contiguous()


def sz_tinygemm_to_b200(sz: Tensor) -> Tensor:
    """the reference's (K/G, N, 2) -> strip-major (N/16, K/G, 16, 2)"""
    ng, n = sz.shape[0], sz.shape[1]
    return sz.reshape(ng, n // 16, 16, 2).permute(1, 0, 2, 3).contiguous()


@register_custom_op("torchao::b200_int4_unpack")
def _(packed_w: Tensor):
    n, k = _strip_shape_checks(packed_w)
    return torch.empty((n, k), dtype=torch.int32, device=packed_w.device)


@_impl_cuda("b200_int4_unpack")
def _int4_unpack_cuda(packed_w: Tensor):
    n, k = _strip_shape_checks(packed_w)
    packed_w = packed_w.contiguous()
    out = torch.empty((n, k), dtype=torch.int32, device=packed_w.device)
    with torch.cuda.device(packed_w.device):
        _lib.check(_lib.lib().ao_b200_int4_unpack(packed_w.data_ptr(), n, k, out.data_ptr(), _stream(packed_w)),
                   "int4_unpack")
    return out


@register_custom_op("torchao::b200_int4_dequantize")
def _(packed_w: Tensor, scales_and_zeros: Tensor, group_size: int):
    n, k = _strip_shape_checks(packed_w)
    _sz_checks_b200(scales_and_zeros, n, k, group_size)
    return torch.empty((n, k), dtype=torch.bfloat16, device=packed_w.device)


@_impl_cuda("b200_int4_dequantize")
def _int4_dequant_cuda(packed_w: Tensor, scales_and_zeros: Tensor, group_size: int):
    n, k = _strip_shape_checks(packed_w)
    _sz_checks_b200(scales_and_zeros, n, k, group_size)
    packed_w, scales_and_zeros = packed_w.contiguous(), scales_and_zeros.contiguous()
    out = torch.empty((n, k), dtype=torch.bfloat16, device=packed_w.device)
    with torch.cuda.device(packed_w.device):
        _lib.check(_lib.lib().ao_b200_int4_dequant(packed_w.data_ptr(), scales_and_zeros.data_ptr(), n, k, group_size,
                                                   out.data_ptr(), _stream(packed_w)), "int4_dequant")
    return out


@register_custom_op("torchao::b200_tinygemm_pack")
def _(int_data: Tensor, inner_k_tiles: int):
    n, k = int_data.shape
    return torch.empty((n // 8, k // (inner_k_tiles * 16), 32, inner_k_tiles // 2), dtype=torch.int32,
                       device=int_data.device)


@_impl_cuda("b200_tinygemm_pack")
def _tinygemm_pack_cuda(int_data: Tensor, inner_k_tiles: int):
    torch._check(int_data.dim() == 2 and int_data.dtype is torch.int32, lambda: "int_data must be 2-D int32")
    n, k = int_data.shape
    torch._check(inner_k_tiles in (2, 4, 8), lambda: "inner_k_tiles must be 2, 4, or 8")
    torch._check(n % 8 == 0 and k % (inner_k_tiles * 16) == 0, lambda: "N % 8 and K % (inner_k_tiles*16) must be 0")
    int_data = int_data.contiguous()
    out = torch.empty((n // 8, k // (inner_k_tiles * 16), 32, inner_k_tiles // 2), dtype=torch.int32,
                      device=int_data.device)
    with torch.cuda.device(int_data.device):
        _lib.check(_lib.lib().ao_b200_tinygemm_pack(int_data.data_ptr(), n, k, inner_k_tiles, out.data_ptr(),
                                                    _stream(int_data)), "tinygemm_pack")
    return out


def convert_tinygemm_to_b200(packed_w: Tensor, inner_k_tiles: int) -> Tensor:
    """Re-tile a weight packed by aten._convert_weight_to_int4pack (reference checkpoints) into the super-tile layout."""
    return torch.ops.torchao.b200_int4_pack(unpack_tensor_core_tiled_layout(packed_w, inner_k_tiles))


# ------------------------------------------------------------------ linear ops
@register_custom_op("torchao::b200_int4_weight_only_mm")
def _(x: Tensor, packed_w: Tensor, group_size: int, scales_and_zeros: Tensor, out_features: int):
    torch._check(x.dim() == 2, lambda: "x must be 2-D")
    return torch.empty((x.size(0), out_features), dtype=torch.bfloat16, device=x.device)


_WORKSPACES = {}


def _workspace(device, nbytes, tag="int4"):
    """Per (device, stream, tag) scratch.  The int4 buffer carries the tensor-memory GEMV's tickets (must stay zero), so
    the int8 GEMM's staging copy lives in its own buffer."""
    if nbytes == 0:
        return None
    key = (device, torch.cuda.current_stream(device).cuda_stream, tag)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws


@_impl_cuda("b200_int4_weight_only_mm")
def _int4_mm_cuda(x: Tensor, packed_w: Tensor, group_size: int, scales_and_zeros: Tensor, out_features: int):
    """Same contract as aten._weight_int4pack_mm(x, packed, G, scale_and_zero)
    (call site torchao/dtypes/uintx/tensor_core_tiled_layout.py:104) on the super-tile layout."""
    np_, kp = _strip_shape_checks(packed_w)
    _sz_checks_b200(scales_and_zeros, np_, kp, group_size)
    _check_cuda_bf16(x, "activation")
    torch._check(x.dim() == 2 and x.size(1) <= kp, lambda: f"x must be (M, K<= {kp}), got {tuple(x.shape)}")
    torch._check(0 < out_features <= np_, lambda: "out_features out of range")
    torch._check(packed_w.is_contiguous() and scales_and_zeros.is_contiguous(), lambda: "weights must be contiguous")
    m = x.size(0)
    y = torch.empty((m, out_features), dtype=torch.bfloat16, device=x.device)
    if m == 0:
        return y
    if x.size(1) % 8 != 0:
        x = torch.nn.functional.pad(x, (0, (-x.size(1)) % 8))
    x, ldx = _rows(x)
    L = _lib.lib()
    ws_bytes = L.ao_b200_int4_mm_workspace_bytes(m, np_, kp)
    ws = _workspace(x.device, ws_bytes)
    with torch.cuda.device(x.device):
        _lib.check(L.ao_b200_int4_mm(x.data_ptr(), m, x.size(1), ldx, packed_w.data_ptr(), scales_and_zeros.data_ptr(),
                                     np_, kp, group_size, y.data_ptr(), out_features, out_features,
                                     ws.data_ptr() if ws is not None else None, ws_bytes, _stream(x)), "int4_mm")
    return y


@register_custom_op("torchao::b200_int8_quantize")
def _(w: Tensor, eps: float, cuda_div_semantics: bool):
    return (torch.empty(w.shape, dtype=torch.int8, device=w.device),
            torch.empty((w.size(0),), dtype=torch.bfloat16, device=w.device))


@_impl_cuda("b200_int8_quantize")
def _int8_quantize_cuda(w: Tensor, eps: float, cuda_div_semantics: bool):
    torch._check(w.dim() == 2, lambda: "weight must be 2-D")
    _check_cuda_bf16(w, "weight")
    w = w.contiguous()
    n, k = w.shape
    q = torch.empty((n, k), dtype=torch.int8, device=w.device)
    scale = torch.empty((n,), dtype=torch.bfloat16, device=w.device)
    with torch.cuda.device(w.device):
        _lib.check(_lib.lib().ao_b200_int8_quantize(w.data_ptr(), n, k, k, eps,
                                                    _lib.DIV_RECIP if cuda_div_semantics else _lib.DIV_IEEE,
                                                    q.data_ptr(), scale.data_ptr(), _stream(w)), "int8_quantize")
    return q, scale


def padded_int8_shape(n: int, k: int) -> Tuple[int, int]:
    """strip-major int8 layout: N -> x128 (row blocks of the prefill GEMM), K -> x64 (k-tiles)."""
    return _round_up(n, 128), _round_up(k, 64)


def _int8_packed_checks(packed_w):
    torch._check(packed_w.is_cuda and packed_w.dtype is torch.uint8 and packed_w.dim() == 3 and packed_w.size(2) == 1024
                 and packed_w.is_contiguous(),
                 lambda: "packed int8 weight must be a contiguous CUDA uint8 tensor (Np/16, Kp/64, 1024)")
    return packed_w.size(0) * 16, packed_w.size(1) * 64


@register_custom_op("torchao::b200_int8_pack")
def _(int_data: Tensor):
    np_, kp = padded_int8_shape(int_data.size(0), int_data.size(1))
    return torch.empty((np_ // 16, kp // 64, 1024), dtype=torch.uint8, device=int_data.device)


@_impl_cuda("b200_int8_pack")
def _int8_pack_cuda(int_data: Tensor):
    """PlainAQTTensorImpl.int_data int8 (N, K) -> strip-major biased int8 (include/ao_b200.h)."""
    torch._check(int_data.is_cuda and int_data.dtype is torch.int8 and int_data.dim() == 2,
                 lambda: "int_data must be a 2-D CUDA int8 tensor")
    int_data = int_data.contiguous()
    n, k = int_data.shape
    np_, kp = padded_int8_shape(n, k)
    packed = torch.empty((np_ // 16, kp // 64, 1024), dtype=torch.uint8, device=int_data.device)
    with torch.cuda.device(int_data.device):
        _lib.check(_lib.lib().ao_b200_int8_pack(int_data.data_ptr(), n, k, np_, kp, packed.data_ptr(), _stream(int_data)),
                   "int8_pack")
    return packed


@register_custom_op("torchao::b200_int8_unpack")
def _(packed_w: Tensor, out_features: int, in_features: int):
    return torch.empty((out_features, in_features), dtype=torch.int8, device=packed_w.device)


@_impl_cuda("b200_int8_unpack")
def _int8_unpack_cuda(packed_w: Tensor, out_features: int, in_features: int):
    np_, kp = _int8_packed_checks(packed_w)
    torch._check(0 < out_features <= np_ and 0 < in_features <= kp, lambda: "out/in features exceed the packed shape")
    q = torch.empty((out_features, in_features), dtype=torch.int8, device=packed_w.device)
    with torch.cuda.device(packed_w.device):
        _lib.check(_lib.lib().ao_b200_int8_unpack(packed_w.data_ptr(), np_, kp, out_features, in_features, q.data_ptr(),
                                                  _stream(packed_w)), "int8_unpack")
    return q


@register_custom_op("torchao::b200_int8_weight_only_mm")
def _(x: Tensor, packed_w: Tensor, scale: Tensor):
    return torch.empty((x.size(0), scale.numel()), dtype=torch.bfloat16, device=x.device)


@_impl_cuda("b200_int8_weight_only_mm")
def _int8_mm_cuda(x: Tensor, packed_w: Tensor, scale: Tensor):
    """y = (x @ int_data.T.to(bf16)) * scale with the reference's two bf16 roundings
    (torchao/dtypes/uintx/plain_layout.py:250-266); out_features = scale.numel()."""
    _check_cuda_bf16(x, "activation")
    _check_cuda_bf16(scale, "scale")
    np_, kp = _int8_packed_checks(packed_w)
    n = scale.numel()
    torch._check(x.dim() == 2 and x.size(1) <= kp, lambda: "x / weight K mismatch")
    torch._check(0 < n <= np_, lambda: "scale must have one entry per output row")
    scale = scale.contiguous()
    m = x.size(0)
    y = torch.empty((m, n), dtype=torch.bfloat16, device=x.device)
    if m == 0:
        return y
    if x.size(1) % 8 != 0:
        x = torch.nn.functional.pad(x, (0, (-x.size(1)) % 8))
    x, ldx = _rows(x)
    L = _lib.lib()
    ws_bytes = L.ao_b200_int8_mm_workspace_bytes(m, np_, kp)
    ws = _workspace(x.device, ws_bytes, "int8")
    with torch.cuda.device(x.device):
        _lib.check(L.ao_b200_int8_mm(x.data_ptr(), m, x.size(1), ldx, packed_w.data_ptr(), scale.data_ptr(), np_, kp,
                                     y.data_ptr(), n, n, ws.data_ptr() if ws is not None else None, ws_bytes, _stream(x)),
                   "int8_mm")
    return y


@register_custom_op("torchao::b200_int8_gemv_fused")
def _(x, packed_w, scale, norm_weight, norm_eps, residual, epilogue):
    n = scale.numel() // 2 if epilogue == 2 else scale.numel()
    return torch.empty((x.size(0), n), dtype=torch.bfloat16, device=x.device)


@_impl_cuda("b200_int8_gemv_fused")
def _int8_gemv_fused_cuda(x, packed_w, scale, norm_weight, norm_eps, residual, epilogue):
    """int8 Llama-block GEMV with RMSNorm prologue / residual or SiLU*mul epilogue (include/ao_b200.h)."""
    np_, kp = _int8_packed_checks(packed_w)
    _check_cuda_bf16(x, "activation")
    _check_cuda_bf16(scale, "scale")
    out_features = scale.numel()
    torch._check(x.dim() == 2 and 0 < x.size(0) <= 16 and x.size(1) <= kp and x.size(1) % 8 == 0,
                 lambda: f"x must be (M<=16, K<={kp}, K%8==0), got {tuple(x.shape)}")
    torch._check(0 < out_features <= np_, lambda: "out_features out of range")
    torch._check(epilogue in (0, 1, 2), lambda: "epilogue must be 0 (plain), 1 (residual) or 2 (silu*mul)")
    x, ldx = _rows(x)
    scale = scale.contiguous()
    m = x.size(0)
    n = out_features // 2 if epilogue == 2 else out_features
    y = torch.empty((m, n), dtype=torch.bfloat16, device=x.device)
    ld_res = 0
    if epilogue == 1:
        torch._check(residual is not None and tuple(residual.shape) == (m, n), lambda: "residual must be (M, N)")
        _check_cuda_bf16(residual, "residual")
        residual = residual.contiguous()
        ld_res = residual.stride(0)
    if norm_weight is not None:
        _check_cuda_bf16(norm_weight, "norm_weight")
        torch._check(norm_weight.numel() == x.size(1), lambda: "norm_weight must have K elements")
        norm_weight = norm_weight.contiguous()
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().ao_b200_int8_gemv_fused(
            x.data_ptr(), m, x.size(1), ldx, packed_w.data_ptr(), scale.data_ptr(), np_, kp, y.data_ptr(), out_features, n,
            None if norm_weight is None else norm_weight.data_ptr(), norm_eps,
            None if residual is None else residual.data_ptr(), ld_res, epilogue, _stream(x)), "int8_gemv_fused")
    return y


def interleave_gate_up_int8(q1: Tensor, s1: Tensor, q3: Tensor, s3: Tensor) -> Tuple[Tensor, Tensor]:
    """w1 (gate) / w3 (up) plain int8 (I, K) + scales (I,) -> fused (2I, K) / (2I,) whose 16-row strip t holds rows
    8t..8t+7 of w1 as rows 0..7 and rows 8t..8t+7 of w3 as rows 8..15 (the SiLU*mul epilogue's pairing)."""
    torch._check(q1.shape == q3.shape and q1.size(0) % 8 == 0, lambda: "w1 / w3 shape mismatch")
    i, k = q1.shape
    q = torch.stack([q1.view(i // 8, 8, k), q3.view(i // 8, 8, k)], dim=1).reshape(2 * i, k)
    s = torch.stack([s1.view(i // 8, 8), s3.view(i // 8, 8)], dim=1).reshape(2 * i)
    return q.contiguous(), s.contiguous()


# ------------------------------------------------------------------ fused decode GEMV (M <= 16)
EPILOGUE_PLAIN, EPILOGUE_RESIDUAL, EPILOGUE_SILU_MUL = 0, 1, 2

if not _defined("b200_int4_gemv_fused"):
    lib.define("b200_int4_gemv_fused(Tensor x, Tensor packed_w, int group_size, Tensor scales_and_zeros, "
               "int out_features, Tensor? norm_weight, float norm_eps, Tensor? residual, int epilogue) -> Tensor")


@register_custom_op("torchao::b200_int4_gemv_fused")
def _(x, packed_w, group_size, scales_and_zeros, out_features, norm_weight, norm_eps, residual, epilogue):
    n = out_features // 2 if epilogue == EPILOGUE_SILU_MUL else out_features
    return torch.empty((x.size(0), n), dtype=torch.bfloat16, device=x.device)


@_impl_cuda("b200_int4_gemv_fused")
def _int4_gemv_fused_cuda(x, packed_w, group_size, scales_and_zeros, out_features, norm_weight, norm_eps, residual,
                          epilogue):
    """Llama-block GEMV with RMSNorm prologue / residual or SiLU*mul epilogue (include/ao_b200.h)."""
    np_, kp = _strip_shape_checks(packed_w)
    _sz_checks_b200(scales_and_zeros, np_, kp, group_size)
    _check_cuda_bf16(x, "activation")
    torch._check(x.dim() == 2 and 0 < x.size(0) <= 16 and x.size(1) <= kp and x.size(1) % 8 == 0,
                 lambda: f"x must be (M<=16, K<={kp}, K%8==0), got {tuple(x.shape)}")
    torch._check(0 < out_features <= np_, lambda: "out_features out of range")
    torch._check(epilogue in (0, 1, 2), lambda: "epilogue must be 0 (plain), 1 (residual) or 2 (silu*mul)")
    x, ldx = _rows(x)
    m = x.size(0)
    n = out_features // 2 if epilogue == EPILOGUE_SILU_MUL else out_features
    y = torch.empty((m, n), dtype=torch.bfloat16, device=x.device)
    ld_res = 0
    if epilogue == EPILOGUE_RESIDUAL:
        torch._check(residual is not None and tuple(residual.shape) == (m, n), lambda: "residual must be (M, N)")
        _check_cuda_bf16(residual, "residual")
        residual = residual.contiguous()
        ld_res = residual.stride(0)
    if norm_weight is not None:
        _check_cuda_bf16(norm_weight, "norm_weight")
        torch._check(norm_weight.numel() == x.size(1), lambda: "norm_weight must have K elements")
        norm_weight = norm_weight.contiguous()
    L = _lib.lib()
    ws_bytes = L.ao_b200_int4_mm_workspace_bytes(m, np_, kp)
    ws = _workspace(x.device, ws_bytes)
    with torch.cuda.device(x.device):
        _lib.check(L.ao_b200_int4_gemv_fused(
            x.data_ptr(), m, x.size(1), ldx, packed_w.data_ptr(), scales_and_zeros.data_ptr(), np_, kp, group_size,
            y.data_ptr(), out_features, n, None if norm_weight is None else norm_weight.data_ptr(), norm_eps,
            None if residual is None else residual.data_ptr(), ld_res, epilogue,
            ws.data_ptr() if ws is not None else None, ws_bytes, _stream(x)), "int4_gemv_fused")
    return y


def interleave_gate_up(packed1: Tensor, sz1: Tensor, packed3: Tensor, sz3: Tensor) -> Tuple[Tensor, Tensor]:
    """Fuse w1 (gate) and w3 (up) for the SiLU*mul epilogue: fused 16-row strip t holds rows 8t..8t+7 of w1 as its rows
    0..7 and rows 8t..8t+7 of w3 as its rows 8..15, so one MMA lane owns gate_i and up_i of the same i.  Pure
    permutation of packed words / scale rows: each row keeps its own quantization (rows are quantized independently)."""
    torch._check(packed1.shape == packed3.shape and sz1.shape == sz3.shape, lambda: "w1 / w3 shape mismatch")
    S16, KT = packed1.shape[0], packed1.shape[1]
    NG = sz1.shape[1]
    # word j = 2h + b of a lane holds row g + 8b of the strip
    p1 = packed1.reshape(S16, KT, 32, 2, 2)
    p3 = packed3.reshape(S16, KT, 32, 2, 2)
    f = torch.stack([p1, p3], dim=-1)                       # [s][kt][lane][h][b][b']; b' = 0 gate rows, 1 up rows
    f = f.permute(0, 4, 1, 2, 3, 5).reshape(2 * S16, KT, 32, 4).contiguous()   # new strip = 2 s + b, word j' = 2h + b'
    z1 = sz1.reshape(S16, NG, 2, 8, 2)                      # [s][grp][b][g][2]
    z3 = sz3.reshape(S16, NG, 2, 8, 2)
    z = torch.stack([z1, z3], dim=3)                        # [s][grp][b][b'][g][2]
    return f, z.permute(0, 2, 1, 3, 4, 5).reshape(2 * S16, NG, 16, 2).contiguous()
