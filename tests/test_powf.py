"""prt_b200/csrc/prt_powf.h — the powf the device runs — against the host libm's powf, bit for bit: every fp32 in
[0, 1.5] for y = 2.2f and 1/2.2f (the exponents the reference uses, material.cpp:24-28 / image.cpp:16-19) plus random
(x, y) pairs; ~2e9 arguments on all host threads (oracle/powf_check.cpp)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_powf_restatement_matches_libm(tmp_path):
    exe = str(tmp_path / "powf_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", os.path.join(ROOT, "oracle", "powf_check.cpp"), "-o", exe, "-lpthread"])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and " 0 mismatches" in out.stdout, out.stdout[-2000:]
