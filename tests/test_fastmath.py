"""CPU: csrc/chub_fastmath.h (compiled for the host) against libm over the kernel's argument ranges."""
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fastmath_against_libm(tmp_path):
    exe = str(tmp_path / "fastmath_check")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17",
                           os.path.join(ROOT, "tests", "fastmath_check.cpp"), "-o", exe])
    out = json.loads(subprocess.check_output([exe, "1500000"]).decode())
    assert out["log_curve_max_ulp"] <= 1.5 and out["exp_max_ulp"] <= 1.5
    assert out["log_normal_max_relerr_eps"] <= 2.0
    # what matters to parity: after the reference's fp32 narrowing nothing differs
    assert out["float_mismatch_log"] == 0 and out["float_mismatch_exp"] == 0 and out["float_mismatch_normal"] == 0
    assert out["div_mismatch"] == 0
    assert out["normal_edge_mismatch"] == 0   # r2 == 1.0f and its neighbours, tiny r2: no NaN, same fp32 variate as libm
