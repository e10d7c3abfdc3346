"""The boundary proved on the real reference package ("next" item 4; INTEGRATION.md sections 1 and 3):
tests/ref_binding_worker.py imports the unmodified reference from oracle/_ref, binds libao_b200.so into it exactly as
INTEGRATION.md prescribes and runs the reference's own checks (test/test_ops.py:282-496 restated: the reference's test
tree does not travel to the GPU box) — unpack bit-exact, dequant identical to the aten identity-matmul, opcheck with all
FOUR utilities (incl. test_aot_dispatch_dynamic), quantize_ + linear with the reference's dispatch routed to the
strip-major kernels."""
import json
import os
import subprocess
import sys
import tempfile

import pytest

from torchao_fork_b200 import _lib

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


@pytest.fixture(scope="module")
def result():
    if not os.path.isdir(os.path.join(REF, "torchao")):
        pytest.skip("oracle/_ref (the reference install) is missing")
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "out.json")
        env = dict(os.environ, PYTHONPATH=REF)
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_binding_worker.py"), _lib.LIB_PATH, out],
                           capture_output=True, text=True, timeout=1200, env=env, cwd=td)
        assert r.returncode == 0, f"binding worker failed:\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"
        return json.load(open(out))


def test_unpack_is_the_exact_inverse_of_the_aten_pack(result):          # test_ops.py:282-293
    assert len(result["unpack"]) == 15 and all(c["equal"] for c in result["unpack"]), result["unpack"]


def test_dequant_equals_the_identity_matmul_through_weight_int4pack_mm(result):   # test_ops.py:339-467
    assert len(result["dequant"]) == 24
    for c in result["dequant"]:
        assert c["diff_op_id"] == 0.0, c                       # same bits as the GEMM's own dequantisation
        assert c["diff_op_ao"] == c["diff_ao_id"], c          # same distance to the python dequantisation
        assert c["diff_op_ao"] < 1e-1 and c["diff_un_ao"] == 0.0, c


def test_opcheck_with_all_four_utilities_on_the_reference_schemas(result):       # test_ops.py:297-316, 470-496
    assert len(result["opcheck"]) == 4
    for c in result["opcheck"]:
        assert set(c["result"]) == {"test_schema", "test_autograd_registration", "test_faketensor", "test_aot_dispatch_dynamic"}
        assert all(v == "SUCCESS" for v in c["result"].values()), c


def test_reference_quantize_and_linear_routed_to_the_strip_major_kernels(result):
    assert len(result["linear"]) == 5
    for c in result["linear"]:
        assert c["sqnr_vs_reference_kernel"] >= 40.0, c        # vs the reference's own aten._weight_int4pack_mm
        assert c["sqnr_vs_bf16"] >= 15.0, c                     # the reference's int4wo floor (test_integration.py:1002-1004)
