"""The C ABI called from plain C (no Python, no torch types): tests/c_abi_smoke.c is compiled with gcc against
include/fpemu_b200.h, linked with libfpemu_b200.so and the CUDA runtime, and run on the GPU."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fp8-emulation-toolkit_b200")


def test_plain_c_program_drives_the_library(tmp_path):
    if shutil.which("gcc") is None or not os.path.exists("/usr/local/cuda/lib64/libcudart.so"):
        pytest.skip("gcc or the CUDA runtime library is missing")
    exe = str(tmp_path / "c_abi_smoke")
    subprocess.check_call(["gcc", "-std=c99", "-O1", "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include",
                           os.path.join(ROOT, "tests", "c_abi_smoke.c"), "-o", exe, "-L", PKG, "-lfpemu_b200",
                           "-L", "/usr/local/cuda/lib64", "-lcudart", "-lm", "-Wl,-rpath," + PKG,
                           "-Wl,-rpath,/usr/local/cuda/lib64"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "c_abi_smoke ok" in r.stdout
