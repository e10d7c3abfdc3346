"""tinygemm helpers (reference: torchao/quantization/utils.py)."""
import torch

__all__ = ["compute_error", "pack_tinygemm_scales_and_zeros", "unpack_tinygemm_scales_and_zeros",
           "get_groupwise_affine_qparams", "groupwise_affine_quantize_tensor_from_qparams",
           "groupwise_affine_dequantize_tensor_from_qparams"]


def compute_error(x, y):
    """SQNR in dB (reference: utils.py:51-55)."""
    ps = torch.linalg.norm(x.float())
    pn = torch.linalg.norm(x.float() - y.float())
    return 20 * torch.log10(ps / pn)


def pack_tinygemm_scales_and_zeros(scales, zeros, dtype=torch.bfloat16):
    """(N, K/G) scale, zero -> (K/G, N, 2).  reference: utils.py:395-409."""
    assert scales.dtype == dtype and zeros.dtype == dtype, f"scales/zeros must be {dtype}"
    assert scales.shape == zeros.shape
    return torch.stack([scales, zeros], dim=-1).transpose(-3, -2).contiguous()


def unpack_tinygemm_scales_and_zeros(scales_and_zeros):
    """reference: utils.py:412-414"""
    assert scales_and_zeros.shape[-1] == 2
    return torch.split(scales_and_zeros.transpose(-3, -2), 1, -1)


def get_groupwise_affine_qparams(w, n_bit=4, groupsize=128, dtype=torch.bfloat16):
    """reference: utils.py:325-392 (tinygemm / FLOAT zero-point flavour)."""
    from .quant_primitives import MappingType, _choose_qparams_affine_tinygemm
    assert n_bit == 4 and dtype == torch.bfloat16
    s, z = _choose_qparams_affine_tinygemm(w, MappingType.ASYMMETRIC, (1, groupsize), torch.int32, 0, 15, 1e-6,
                                           dtype, dtype)
    return s.reshape(w.shape[0], -1), z.reshape(w.shape[0], -1)


def groupwise_affine_quantize_tensor_from_qparams(w, scales, zeros, n_bit=4, groupsize=128):
    """reference: utils.py:417-458 — returns the uint8 nibble-paired (N, K/2) tensor
    (high nibble = even k; tensor_core_tiled_layout.py:276)."""
    from .quant_primitives import _quantize_affine_tinygemm
    assert n_bit == 4
    q = _quantize_affine_tinygemm(w, (1, groupsize), scales, zeros, torch.int32, 0, 15)
    return (q[:, ::2] << 4 | q[:, 1::2]).to(torch.uint8)


def groupwise_affine_dequantize_tensor_from_qparams(w_int4x8, scales, zeros, n_bit=4, groupsize=128):
    """reference: utils.py:461-500"""
    from .quant_primitives import _dequantize_affine_tinygemm
    assert n_bit == 4
    if w_int4x8.dtype == torch.uint8:
        hi = (w_int4x8 >> 4).to(torch.int32)
        lo = (w_int4x8 & 0xF).to(torch.int32)
        w_int32 = torch.stack([hi, lo], dim=-1).reshape(w_int4x8.shape[0], -1)
    else:
        w_int32 = w_int4x8
    return _dequantize_affine_tinygemm(w_int32, (1, groupsize), scales, zeros, torch.int32, 0, 15,
                                       output_dtype=scales.dtype)


@torch.no_grad()
def _quantize_activation_per_token_absmax(t):
    """reference: torchao/quantization/utils.py:152-182 — symmetric int8 per row of the last dimension, -127..127,
    eps 1e-5, scale in t's dtype (fp32 for fp16 inputs).  Written out as the tensor expressions choose_qparams_affine
    (quant_primitives.py:1537-1575, SYMMETRIC branch) and _quantize_affine_no_dtype_cast (:433-459) evaluate, in the same
    order and dtypes, so that eager results equal the reference's bit for bit on the same device.  The decode kernel
    (csrc/decode_attn_q8.cu) does the same arithmetic for one new row."""
    quant_min, quant_max, eps = -127, 127, 1e-5
    scale_dtype = torch.float32 if t.dtype == torch.float16 else t.dtype
    min_val = torch.amin(t, dim=-1, keepdim=False)
    max_val = torch.amax(t, dim=-1, keepdim=False)
    min_val_neg = torch.min(min_val, torch.zeros_like(min_val))
    max_val_pos = torch.max(max_val, torch.zeros_like(max_val))
    max_val_pos = torch.max(-min_val_neg, max_val_pos)
    scale = max_val_pos / (float(quant_max - quant_min) / 2)
    zero_point = torch.full_like(scale, int((quant_max + quant_min + 1) / 2))
    scale = torch.clamp(scale, min=eps).to(dtype=scale_dtype)
    quant = torch.clamp(torch.round(t * (1.0 / scale.unsqueeze(-1))) + zero_point.unsqueeze(-1), quant_min, quant_max)
    return quant.to(torch.int8), scale
