"""bench.py's reference arm on the host cores (the one arm that runs without a GPU): ONE JSON line on stdout with the
contract's keys (`impl`, metric / unit / value, `cpu_baseline`, `e2e` with zero copy bytes, config.workload)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_reference_arm_prints_one_contract_line():
    if not os.path.isdir(os.path.join(ROOT, "oracle", "_ref", "torchao")):
        pytest.skip("oracle/_ref (the reference install) is missing: run __graft_entry__.build() where /root/reference exists")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=550, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "llama3_8b_int4_decode_tok_per_s" and d["unit"] == "tok/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2
    assert d["config"]["workload"] == "llama3-8b-int4"
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
