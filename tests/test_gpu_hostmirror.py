"""GPU: the reference's own five unit tests, run against the header-only C++ mirror over the C ABI."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_mirror_runs_reference_unit_tests(tmp_path):
    exe = str(tmp_path / "host_mirror_test")
    lib = os.path.join(ROOT, "slam-rs_b200", "csrc")
    subprocess.check_call(["g++", "-std=c++17", "-O1", os.path.join(ROOT, "tests", "host_mirror_test.cpp"), "-o", exe,
                           "-L" + lib, "-lorbx", "-Wl,-rpath," + lib])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "host mirror ok" in out.stdout, out.stdout + out.stderr
