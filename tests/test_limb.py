"""Host-side check of the limb arithmetic that hist_rows_kernel uses (shifu_b200/csrc/limb.h): worst-case and random
contribution streams against exact 128-bit sums, including every bound the no-carry accumulation relies on."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_limb_arithmetic_is_exact(tmp_path):
    exe = str(tmp_path / "limb_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(HERE, "limb_check.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 failures" in out.stdout
