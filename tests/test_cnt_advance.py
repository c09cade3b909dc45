"""cnt_advance (csrc/cuda/cnt_advance.h): the excitation's pitch counter after m steps of "+= 1.0" without walking
them - what lets k_pulse_warp test 256 samples at once and still place every pulse on the reference's sample
(src/vocoder/excitation.rs:72-86). The header compiles for the host; tests/native/cnt_advance_check.cc compares it
with the walk bit for bit on counters right after a pulse, arbitrary mantissas below 4096, neighbours of powers of
two, integers and denormals."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cnt_advance_equals_the_walk_bit_for_bit(tmp_path):
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("needs g++")
    exe = str(tmp_path / "cnt_advance_check")
    subprocess.check_call([cxx, "-O2", "-std=c++17", "-ffp-contract=off", "-I", os.path.join(ROOT, "jbonsai_b200", "csrc", "cuda"),
                           "-o", exe, os.path.join(ROOT, "tests", "native", "cnt_advance_check.cc")])
    out = subprocess.run([exe, "300000"], capture_output=True, text=True)
    sys.stdout.write(out.stdout)
    assert out.returncode == 0 and "2100003 cases, 0 mismatches" in out.stdout
