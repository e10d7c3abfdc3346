"""The drop-in boundary on the REAL reference package: imports the unmodified reference (oracle/_ref), applies the edit
INTEGRATION.md describes — ctypes-load libao_b200.so, CUDA backends for the reference's own
`torchao::unpack_tensor_core_tiled_layout` / `dequantize_tensor_core_tiled_layout` schemas (torchao/ops.py:16-21), the
int4 layout's `from_plain` and linear impl routed to the strip-major kernels — and then runs the checks of the
reference's own tests (restated from /root/reference/test/test_ops.py:282-467, which does not travel to the GPU box)
against it.  Test infrastructure, run in a process of its own:

    PYTHONPATH=oracle/_ref python tests/ref_binding_worker.py <path to libao_b200.so> <out.json>
"""
import ctypes
import json
import sys

import torch


def bind(lib_path):
    import torchao  # the reference: defines the schemas and their fake impls (torchao/ops.py:16-21, 255-377)

    assert "oracle/_ref" in torchao.__file__.replace("\\", "/")
    ao = ctypes.CDLL(lib_path)
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    ao.ao_b200_error_string.restype = ctypes.c_char_p
    ao.ao_b200_tinygemm_unpack.argtypes = [vp, i64, i64, i32, vp, vp]
    ao.ao_b200_tinygemm_dequant.argtypes = [vp, vp, i64, i64, i32, i32, vp, vp]
    ao.ao_b200_int4_pack.argtypes = [vp, i64, i64, i64, i64, i32, vp, vp]
    ao.ao_b200_int4_mm_workspace_bytes.argtypes = [i64, i64, i64]
    ao.ao_b200_int4_mm_workspace_bytes.restype = i64
    ao.ao_b200_int4_mm.argtypes = [vp, i64, i64, i64, vp, vp, i64, i64, i32, vp, i64, i64, vp, i64, vp]

    def check(rc):
        if rc:
            raise RuntimeError(ao.ao_b200_error_string(rc).decode())

    def stream(t):
        return torch.cuda.current_stream(t.device).cuda_stream

    impl = torch.library.Library("torchao", "IMPL")

    def unpack_cuda(packed_w, inner_k_tiles):
        n, k = packed_w.size(0) * 8, packed_w.size(1) * inner_k_tiles * 16
        out = torch.empty((n, k), dtype=torch.int32, device=packed_w.device)
        check(ao.ao_b200_tinygemm_unpack(packed_w.data_ptr(), n, k, inner_k_tiles, out.data_ptr(), stream(packed_w)))
        return out

    def dequant_cuda(packed_w, scales_and_zeros, group_size, inner_k_tiles):
        n, k = packed_w.size(0) * 8, packed_w.size(1) * inner_k_tiles * 16
        out = torch.empty((n, k), dtype=torch.bfloat16, device=packed_w.device)
        check(ao.ao_b200_tinygemm_dequant(packed_w.data_ptr(), scales_and_zeros.data_ptr(), n, k, group_size, inner_k_tiles,
                                          out.data_ptr(), stream(packed_w)))
        return out

    impl.impl("unpack_tensor_core_tiled_layout", unpack_cuda, "CUDA")
    impl.impl("dequantize_tensor_core_tiled_layout", dequant_cuda, "CUDA")

    # ---- the int4 layout routed to the strip-major kernels
    from torchao.dtypes.affine_quantized_tensor_ops import _AQT_QLINEAR_DISPATCH_TABLE
    from torchao.dtypes.uintx import tensor_core_tiled_layout as tctl

    ws_cache = {}

    def from_plain_b200(cls, int_data, scale, zero_point, _layout):
        """TensorCoreTiledAQTTensorImpl.from_plain (tensor_core_tiled_layout.py:262-307) on the strip-major layout."""
        n, k = int_data.shape
        g = k // scale.reshape(n, -1).shape[1]
        np_, kp = (n + 127) // 128 * 128, (k + max(g, 64) - 1) // max(g, 64) * max(g, 64)
        packed = torch.empty((np_ // 16, kp // 64, 32, 4), dtype=torch.int32, device=int_data.device)
        check(ao.ao_b200_int4_pack(int_data.contiguous().data_ptr(), n, k, np_, kp, 8, packed.data_ptr(), stream(int_data)))
        s = torch.zeros((np_, kp // g), dtype=torch.bfloat16, device=int_data.device)
        z = torch.zeros_like(s)
        s[:n, : k // g] = scale.reshape(n, -1)
        z[:n, : k // g] = zero_point.reshape(n, -1)
        # strip-major [Np/16][Kp/G][16][2] (include/ao_b200.h)
        sz = torch.stack([s, z], dim=-1).reshape(np_ // 16, 16, kp // g, 2).permute(0, 2, 1, 3).contiguous()
        return cls(packed, sz, False, _layout)

    def linear_b200(input_tensor, weight_tensor, bias):
        """_linear_bf16_act_uint4_weight_impl (tensor_core_tiled_layout.py:74-114) -> ao_b200_int4_mm: no activation pad,
        no output slice."""
        packed, sz = weight_tensor.tensor_impl.packed_weight, weight_tensor.tensor_impl.scale_and_zero
        n, k = weight_tensor.shape
        g = weight_tensor.block_size[-1]
        x = input_tensor.reshape(-1, k).to(torch.bfloat16).contiguous()
        np_, kp = packed.size(0) * 16, packed.size(1) * 64
        y = torch.empty((x.size(0), n), dtype=torch.bfloat16, device=x.device)
        nbytes = int(ao.ao_b200_int4_mm_workspace_bytes(x.size(0), np_, kp))
        key = (x.device, nbytes)
        if key not in ws_cache:
            ws_cache[key] = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=x.device)
        ws = ws_cache[key]
        rc = ao.ao_b200_int4_mm(x.data_ptr(), x.size(0), k, x.stride(0), packed.data_ptr(), sz.data_ptr(), np_, kp, g,
                                y.data_ptr(), n, y.stride(0), ws.data_ptr(), ws.numel(), stream(x))
        if rc:
            raise RuntimeError(f"ao_b200_int4_mm: {ao.ao_b200_error_string(rc).decode()} (M={x.size(0)} Kx={k} ldx={x.stride(0)} "
                               f"Np={np_} Kp={kp} G={g} N_out={n} packed={tuple(packed.shape)} sz={tuple(sz.shape)} "
                               f"ptrs={x.data_ptr() % 16},{packed.data_ptr() % 16},{sz.data_ptr() % 64})")
        y = y.reshape(*input_tensor.shape[:-1], n)
        if bias is not None:
            y += bias
        return y.to(input_tensor.dtype)

    def enable():
        from torchao.dtypes.affine_quantized_tensor import AffineQuantizedTensor

        tctl.TensorCoreTiledAQTTensorImpl.from_plain = classmethod(from_plain_b200)
        # what @register_layout(TensorCoreTiledLayout) does (torchao/utils.py:636-642): the constructor table holds the
        # bound from_plain, so it is re-registered; the linear dispatch entry is replaced under the same check
        AffineQuantizedTensor._LAYOUT_CONSTRUCTOR_TABLE[tctl.TensorCoreTiledLayout] = tctl.TensorCoreTiledAQTTensorImpl.from_plain
        _AQT_QLINEAR_DISPATCH_TABLE[tctl._linear_bf16_act_uint4_weight_check] = linear_b200

    return enable, impl


def sqnr(ref, x):
    ref, x = ref.double(), x.double()
    return float(10 * torch.log10(ref.pow(2).mean() / (ref - x).pow(2).mean().clamp_min(1e-300)))


def main(lib_path, out_path):
    enable, _keep = bind(lib_path)
    import torchao
    from torch.library import opcheck
    from torchao.quantization import Int4WeightOnlyConfig, quantize_
    from torchao.quantization.utils import (
        get_groupwise_affine_qparams,
        groupwise_affine_dequantize_tensor_from_qparams,
        groupwise_affine_quantize_tensor_from_qparams,
        pack_tinygemm_scales_and_zeros,
    )

    res = {"unpack": [], "dequant": [], "opcheck": [], "linear": []}
    torch.manual_seed(0)
    shapes = [(4096, 4096), (4096, 11008), (11008, 4096), (4096, 14336), (14336, 4096)]   # test_ops.py:260-266
    utils = ["test_schema", "test_autograd_registration", "test_faketensor", "test_aot_dispatch_dynamic"]   # test_ops.py:300-307
    # test_unpack_tensor_core_tiled_layout_correctness (test_ops.py:282-293): bit-exact inverse of the aten pack
    for shape in shapes:
        for ikt in (2, 4, 8):
            t = torch.randint(0, 16, dtype=torch.int, size=shape, device="cuda")
            t8 = (t[::, ::2] << 4 | t[::, 1::2]).to(torch.uint8)
            packed = torch.ops.aten._convert_weight_to_int4pack(t8, ikt)
            un = torchao.ops.unpack_tensor_core_tiled_layout(packed, ikt)
            un8 = (un[::, ::2] << 4 | un[::, 1::2]).to(torch.uint8)
            res["unpack"].append({"shape": shape, "ikt": ikt, "equal": bool(torch.equal(t8, un8))})
    # test_unpack_tensor_core_tiled_layout_op / test_dequantize_tensor_core_tiled_layout_op (test_ops.py:297-316, 470-496)
    t = torch.randint(0, 16, dtype=torch.int, size=(4096, 4096), device="cuda")
    t8 = (t[::, ::2] << 4 | t[::, 1::2]).to(torch.uint8)
    for ikt in (2, 8):
        packed = torch.ops.aten._convert_weight_to_int4pack(t8, ikt)
        r = opcheck(torch.ops.torchao.unpack_tensor_core_tiled_layout, (packed, ikt), test_utils=utils, raise_exception=False)
        res["opcheck"].append({"op": "unpack", "ikt": ikt, "result": {k: str(v) for k, v in r.items()}})
        sz = torch.randn(4096 // 128, 4096, 2, dtype=torch.bfloat16, device="cuda")
        r = opcheck(torch.ops.torchao.dequantize_tensor_core_tiled_layout, (packed, sz, 128, ikt), test_utils=utils, raise_exception=False)
        res["opcheck"].append({"op": "dequantize", "ikt": ikt, "result": {k: str(v) for k, v in r.items()}})
    # test_dequantize_tensor_core_tiled_layout_correctness_{quant_dequant, unpack_and_dequant} (test_ops.py:339-467)
    for shape in shapes[:3]:
        for ikt in (2, 8):
            for g in (32, 64, 128, 256):
                n, k = shape
                w = torch.randn(n, k, dtype=torch.bfloat16, device="cuda")
                scales, zeros = get_groupwise_affine_qparams(w, n_bit=4, groupsize=g, dtype=torch.bfloat16)
                q = groupwise_affine_quantize_tensor_from_qparams(w, scales, zeros, n_bit=4, groupsize=g)
                packed = torch.ops.aten._convert_weight_to_int4pack(q, ikt)
                sz = pack_tinygemm_scales_and_zeros(scales, zeros)
                dq_ao = groupwise_affine_dequantize_tensor_from_qparams(q, scales, zeros, n_bit=4, groupsize=g)
                eye = torch.eye(k, device="cuda", dtype=torch.bfloat16)
                dq_id = torch.ops.aten._weight_int4pack_mm(eye, packed, g, sz).t()
                dq_op = torchao.ops.dequantize_tensor_core_tiled_layout(packed, sz, g, ikt)
                un = torchao.ops.unpack_tensor_core_tiled_layout(packed, ikt)
                un8 = (un[::, ::2] << 4 | un[::, 1::2]).to(torch.uint8)
                dq_un = groupwise_affine_dequantize_tensor_from_qparams(un8, scales, zeros, n_bit=4, groupsize=g)
                res["dequant"].append({
                    "shape": shape, "ikt": ikt, "g": g,
                    "diff_op_id": float((dq_op - dq_id).abs().max()),
                    "diff_op_ao": float((dq_op - dq_ao).abs().max()), "diff_ao_id": float((dq_id - dq_ao).abs().max()),
                    "diff_un_ao": float((dq_un - dq_ao).abs().max())})
    # the int4wo case of test/dtypes/test_affine_quantized.py / test_quant_api.py:594-660: quantize_ + linear, first through
    # the reference's own aten path, then with from_plain / the linear impl routed to libao_b200 (same reference quantize_)
    for (n, k, g, m) in [(4096, 4096, 128, 1), (4096, 4096, 128, 16), (6144, 4096, 64, 128), (4096, 14336, 128, 1), (321, 256, 32, 3)]:
        lin = torch.nn.Linear(k, n, bias=False, device="cuda", dtype=torch.bfloat16)
        w = lin.weight.detach().clone()
        x = torch.randn(m, k, dtype=torch.bfloat16, device="cuda")
        y_bf16 = lin(x)
        quantize_(lin, Int4WeightOnlyConfig(group_size=g))
        y_ref = lin(x)
        res.setdefault("_pending", []).append((n, k, g, m, w, x, y_bf16, y_ref))
    enable()
    for (n, k, g, m, w, x, y_bf16, y_ref) in res.pop("_pending"):
        lin = torch.nn.Linear(k, n, bias=False, device="cuda", dtype=torch.bfloat16)
        with torch.no_grad():
            lin.weight.copy_(w)
        quantize_(lin, Int4WeightOnlyConfig(group_size=g))
        pw = lin.weight.tensor_impl.packed_weight
        assert pw.dim() == 4 and pw.shape[0] == (n + 127) // 128 * 8 and pw.shape[2:] == (32, 4), \
            f"from_plain was not routed to the strip-major layout: {tuple(pw.shape)}"
        y = lin(x)
        res["linear"].append({"n": n, "k": k, "g": g, "m": m, "sqnr_vs_reference_kernel": sqnr(y_ref, y),
                              "sqnr_vs_bf16": sqnr(y_bf16, y), "sqnr_reference_vs_bf16": sqnr(y_bf16, y_ref)})
    json.dump(res, open(out_path, "w"))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
