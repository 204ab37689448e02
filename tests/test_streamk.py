"""The stream-K range arithmetic of the kernels (csrc/w4a16_streamk.cuh: range_begin / owner_of
without 64-bit division) against the 64-bit formulas, on the host: every grid size up to 296,
every T up to P + 2100, pseudo-random and extreme T up to 2^31 - 1 (tests/streamk_check.cpp)."""
import os
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_range_arithmetic_matches_the_64_bit_formulas():
    src = os.path.join(ROOT, "tests", "streamk_check.cpp")
    with tempfile.TemporaryDirectory() as d:
        exe = os.path.join(d, "streamk_check")
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, src], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.startswith("ok ")
