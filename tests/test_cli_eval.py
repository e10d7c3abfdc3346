"""generate.py's command line (reference: torchao/_models/llama/generate.py:258-1115, 1118-1262): -q parsing, --save and
re-loading what it wrote, --kv_cache_quantization; and the perplexity loop of eval.py (the restated lm-eval rolling
log-likelihood, torchao/_models/_eval.py:54-67)."""
import math

import pytest
import torch

import torchao_fork_b200 as ao
import importlib

G = importlib.import_module("torchao_fork_b200._models.llama.generate")   # (the package re-exports the function `generate`)
from torchao_fork_b200._models.llama.eval import rolling_perplexity
from torchao_fork_b200.dtypes import AffineQuantizedTensor

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("q", ["int4wo-128", "int4wo-64-hqq", "int8wo"])
def test_cli_save_and_reload(tmp_path, monkeypatch, q):
    monkeypatch.chdir(tmp_path)
    res = G.main(random_init="tiny-llama3", quantization=q, num_samples=1, max_new_tokens=12, prompt="1,2,3,4,5", save=True,
                 temperature=0.0, top_k=None)
    assert res["tokens_per_sec"] > 0 and res["model_size_gb"] > 0
    path = tmp_path / f"tiny-llama3-{q}.pt"
    assert path.is_file()
    # what --save wrote loads straight into the quantized model (eval.py / generate.py _load_model: assign=True)
    model = G._load_model(path, "cuda", torch.bfloat16)
    assert isinstance(model.layers[0].attention.wqkv.weight, AffineQuantizedTensor)
    fresh = G.apply_quantization(G._random_init_model("tiny-llama3", "cuda", torch.bfloat16), q)
    prompt = torch.tensor([1, 2, 3, 4, 5], device="cuda")
    a = G.generate(model, prompt, 10, temperature=0.0, top_k=None)
    b = G.generate(fresh, prompt, 10, temperature=0.0, top_k=None)
    assert torch.equal(a, b)


def test_cli_rejects_techniques_outside_the_path():
    m = G._random_init_model("tiny-llama3", "cuda", torch.bfloat16)
    with pytest.raises(NotImplementedError, match="int8wo, int4wo"):
        G.apply_quantization(m, "float8dq")
    with pytest.raises(AssertionError, match="group_size"):
        G.apply_quantization(m, "int4wo-48")


def test_cli_kv_cache_quantization(tmp_path):
    out = tmp_path / "result.txt"
    res = G.main(random_init="tiny-llama3", quantization="int4wo-128", num_samples=1, max_new_tokens=10, prompt="7,8,9",
                 kv_cache_quantization=True, write_result=out)
    assert res["decode_tokens_per_sec"] > 0
    assert "kv_quant: True" in out.read_text() and "quant: int4wo-128" in out.read_text()


def test_rolling_perplexity():
    model = G._random_init_model("tiny-llama3", "cuda", torch.bfloat16)
    V = model.config.vocab_size
    tokens = torch.randint(0, V, (300,), generator=torch.Generator().manual_seed(0)).cuda()
    # one window == the direct computation
    one = rolling_perplexity(model, tokens[:100], max_length=128)
    inp = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), tokens[:99]]).view(1, -1)
    model.setup_caches(1, 128, device="cuda")
    with torch.no_grad():
        lp = torch.log_softmax(model(inp, torch.arange(100, device="cuda", dtype=torch.int32))[0].float(), -1)
    nll = -float(lp.gather(1, tokens[:100].view(-1, 1)).sum())
    assert one["n_tokens"] == 100 and one["n_windows"] == 1 and abs(one["nll_sum"] - nll) < 1e-3 * abs(nll)
    # several windows, each conditioned on the token before it
    multi = rolling_perplexity(model, tokens, max_length=128)
    assert multi["n_tokens"] == 300 and multi["n_windows"] == 3 and math.isfinite(multi["token_perplexity"])
    assert 0.2 * V < multi["token_perplexity"] < 5 * V          # random weights on random tokens: about uniform
    # the quantized model stays close to the bf16 one
    ao.quantize_(model, ao.Int4WeightOnlyConfig(group_size=32))
    q = rolling_perplexity(model, tokens, max_length=128)
    assert abs(math.log(q["token_perplexity"]) - math.log(multi["token_perplexity"])) < 0.05
