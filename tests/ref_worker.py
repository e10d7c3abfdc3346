"""Runs the UNMODIFIED reference (oracle/_ref = pip install --target of /root/reference) on the GPU, in a process of its
own (the reference and this repo both own the `torchao::` op namespace).  Test infrastructure, not a test:

    PYTHONPATH=oracle/_ref python tests/ref_worker.py <mode> <in.pt> <out.pt>

modes
  quant  in: {"cases": [{"name", "w" bf16 [N][K], "kind" "int4"|"int8", "group_size", "xs": [bf16 [M][K], ...]}]}
         out: per case, what quantize_(nn.Linear, Int4WeightOnlyConfig(G)) / Int8WeightOnlyConfig() left in the module
         (torchao/quantization/quant_api.py:1042-1157, :1223-1255) and lin(x):
           int4: packed_weight (tinygemm int32), scale_and_zero, dequant = aten._weight_int4pack_mm(eye(Kp)) (the
                 reference's own dequantisation, tensor_core_tiled_layout.py:465-517), ys
           int8: int_data, scale, ys
  model  in: {"cfg": ModelArgs fields, "state": bf16 state_dict, "prompt": int64 [T], "steps": n, "max_seq": S,
              "group_size": G or None}
         out: the reference Transformer (torchao/_models/llama/model.py:243), optionally quantize_'d by the reference,
              run like generate.py:103-117: prefill logits [T][V], then n single-token steps fed with the argmax of its
              own logits: logits [n][V], tokens [n + 1].  "kv_cache_quantization": True builds the caches as
              AffineQuantizedKVCache (model.py:198-240, 302-315).
  kvq    in: {"k", "v": bf16 [B][H][T][D]}
         out: AffineQuantizedKVCache.update (model.py:215-232) called once per position t = 0..T-1 with the rows of t:
              k_cache / v_cache int8, k_cache_scale / v_cache_scale, and the k_out / v_out the LAST call returned
"""
import sys

import torch


def run_quant(inp):
    from torchao.quantization import Int4WeightOnlyConfig, Int8WeightOnlyConfig, quantize_

    out = {}
    for case in inp["cases"]:
        w = case["w"].cuda()
        n, k = w.shape
        lin = torch.nn.Linear(k, n, bias=False, device="cuda", dtype=torch.bfloat16)
        with torch.no_grad():
            lin.weight.copy_(w)
        if case["kind"] == "int4":
            g = case["group_size"]
            quantize_(lin, Int4WeightOnlyConfig(group_size=g))
            impl = lin.weight.tensor_impl
            packed, sz = impl.packed_weight, impl.scale_and_zero
            kp = sz.shape[0] * g
            eye = torch.eye(kp, dtype=torch.bfloat16, device="cuda")
            deq = torch.ops.aten._weight_int4pack_mm(eye, packed, g, sz).t().contiguous()   # [Np][Kp]
            res = {"packed_weight": packed.cpu(), "scale_and_zero": sz.cpu(), "dequant": deq.cpu()}
        else:
            quantize_(lin, Int8WeightOnlyConfig())
            impl = lin.weight.tensor_impl
            res = {"int_data": impl.int_data.cpu(), "scale": impl.scale.cpu()}
        with torch.no_grad():
            res["ys"] = [lin(x.cuda()).cpu() for x in case["xs"]]
        out[case["name"]] = res
    return out


def run_model(inp):
    from torchao._models.llama.model import ModelArgs, Transformer
    from torchao.quantization import Int4WeightOnlyConfig, quantize_

    cfg = ModelArgs(**inp["cfg"])
    with torch.device("cuda"):
        model = Transformer(cfg)
    model = model.to(torch.bfloat16).eval()
    missing = model.load_state_dict({k: v.cuda() for k, v in inp["state"].items()}, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    if inp.get("group_size"):
        quantize_(model, Int4WeightOnlyConfig(group_size=inp["group_size"]))
    with torch.device("cuda"):
        model.setup_caches(max_batch_size=1, max_seq_length=inp["max_seq"],
                           kv_cache_quantization=bool(inp.get("kv_cache_quantization")))
    prompt = inp["prompt"].cuda().view(1, -1)
    T = prompt.shape[1]
    with torch.no_grad():
        lg = model(prompt, torch.arange(T, device="cuda"))          # prefill (generate.py:103-108)
        prefill_logits = lg[0].float().cpu()
        tok = lg[0, -1].float().argmax().view(1, 1)
        tokens, step_logits = [int(tok)], []
        for i in range(inp["steps"]):                                # decode_one_token (generate.py:111-117)
            lg = model(tok, torch.tensor([T + i], device="cuda"))
            step_logits.append(lg[0, -1].float().cpu())
            tok = lg[0, -1].float().argmax().view(1, 1)
            tokens.append(int(tok))
    return {"prefill_logits": prefill_logits, "step_logits": torch.stack(step_logits), "tokens": torch.tensor(tokens)}


def run_kvq(inp):
    from torchao._models.llama.model import AffineQuantizedKVCache

    k, v = inp["k"].cuda(), inp["v"].cuda()
    B, H, T, D = k.shape
    with torch.device("cuda"):
        cache = AffineQuantizedKVCache(B, T, H, D, scale_dtype=torch.bfloat16)
    with torch.no_grad():
        for t in range(T):
            k_out, v_out = cache.update(torch.tensor([t], device="cuda"), k[:, :, t : t + 1], v[:, :, t : t + 1])
    return {"k_cache": cache.k_cache.cpu(), "v_cache": cache.v_cache.cpu(), "k_cache_scale": cache.k_cache_scale.cpu(),
            "v_cache_scale": cache.v_cache_scale.cpu(), "k_out": k_out.cpu(), "v_out": v_out.cpu()}


if __name__ == "__main__":
    mode, fin, fout = sys.argv[1:4]
    import torchao

    assert "oracle/_ref" in torchao.__file__.replace("\\", "/"), f"not the reference install: {torchao.__file__}"
    inp = torch.load(fin, weights_only=True)
    res = {"quant": run_quant, "model": run_model, "kvq": run_kvq}[mode](inp)
    res["_torchao_file"] = torchao.__file__
    torch.save(res, fout)
