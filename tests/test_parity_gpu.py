"""GPU parity tests: libfa_b200.so (through the public API -> ctypes -> C ABI) vs the fp32 oracle,
with the reference's own criterion  max|out - ref32| <= 2 * max|out_pt - ref32| + 1e-5
(tests/test_flash_attn.py:1085-1087; 3x for kvcache :2075-2076) and, where the reference FA2 kernel
is available for sm_100 (oracle/_ref), north_star's  err <= 2 * err_of_reference_kernel."""
import pytest
import torch

import _util
from tools.gpu_probe import CASES, run_case

pytestmark = pytest.mark.gpu

DEFAULT_P = [c for c in CASES if c["pmode"] == "t"]


@pytest.mark.parametrize("spec", DEFAULT_P, ids=lambda s: s["name"])
def test_case_matrix(spec):
    r = run_case(spec)
    assert not r["nan"], r
    mult = 3 if spec["kind"] == "kvcache" else 2
    assert r["err"] <= mult * r["err_pt"] + 1e-5, r
    assert r["lse_err"] < 2e-3, r
    if "err_refk" in r:
        assert r["err"] <= 2 * r["err_refk"] + 1e-5, r
    assert r["kernel_used"] == spec["kern"], r
