"""CPU-only checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/fpemu_b200.h declares, parses the reference's mode grammar, and rejects bad arguments with a status
code before touching the GPU (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    from fpemu_b200 import _lib
    _lib.lib()
    return _lib


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "fpemu_b200.h")).read()
    return re.findall(r"FPEMU_API\s+[\w\s\*]+?\b(fpemu_\w+)\s*\(", src)


def test_library_exports_every_declared_symbol(L):
    names = declared_symbols()
    assert len(names) >= 9
    lib = ctypes.CDLL(L.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libfpemu_b200.so does not export %s" % n
    assert set(names) == set(L.SIGNATURES), "python binding and header disagree: %s" % (set(names) ^ set(L.SIGNATURES))


def test_header_is_plain_c(tmp_path):
    """The boundary must be bindable from C (cgo / JNI / ctypes): compile the header with gcc -std=c99."""
    import subprocess
    c = tmp_path / "t.c"
    c.write_text('#include "fpemu_b200.h"\nint main(void){return FPEMU_FMT_COUNT == 7 ? 0 : 1;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                           "-o", str(tmp_path / "t.o")])


# mode strings of the reference dispatcher, mpemu/pytquant/cuda/fpemu_kernels.cu:1058-1364
REFERENCE_MODES = {
    "E5M2_RTZ": (0, 0, 0), "E5M2_RNE": (0, 1, 0), "E5M2_STOCHASTIC": (0, 2, 0), "E5M2_RNAZ": (0, 3, 0),
    "E5M2_RNTZ": (0, 4, 0), "E5M2_RPINF": (0, 5, 0), "E5M2_RNINF": (0, 6, 0),
    "E5M2_DAZ_RNE": (0, 1, 1), "E5M2_DAZ_STOCHASTIC": (0, 2, 1), "E5M2_DAZ_RNAZ": (0, 3, 1), "E5M2_DAZ_RNTZ": (0, 4, 1),
    "FLOAT16_RNE": (4, 1, 0), "FLOAT16_STOCHASTIC": (4, 2, 0), "FLOAT16_DAZ_RNE": (4, 1, 1),
    "BFLOAT16_RNE": (3, 1, 0), "BFLOAT16_STOCHASTIC": (3, 2, 0),
    "E4M3_RNE": (1, 1, 0), "E4M3_STOCHASTIC": (1, 2, 0), "E4M3_IEEE_RNE": (5, 1, 0), "E4M3_IEEE_STOCHASTIC": (5, 2, 0),
    "E3M4_RNE": (2, 1, 0), "E3M4_STOCHASTIC": (2, 2, 0), "FP4_NEAREST": (6, 1, 0),
}


def test_parse_mode_accepts_the_reference_grammar(L):
    for mode, want in REFERENCE_MODES.items():
        assert L.parse_mode(mode) == want, mode
    # truncate for the other formats (north_star "truncate"): same pipeline, rounding add 0
    assert L.parse_mode("E4M3_RTZ") == (1, 0, 0)
    assert L.parse_mode("E3M4_RTZ") == (2, 0, 0)
    assert L.parse_mode("BFLOAT16_RTZ") == (3, 0, 0)


@pytest.mark.parametrize("bad", ["", "NONE", "E4M3", "E4M3_", "e4m3_rne", "E4M3_RNEX", "E5M2_DAZ_RPINF", "FLOAT16_RTZ",
                                 "FLOAT16_DAZ_STOCHASTIC", "INT8_RNE", "E9M9_RNE"])
def test_parse_mode_rejects_unknown_strings(L, bad):
    with pytest.raises(ValueError):
        L.parse_mode(bad)


def test_python_oracle_parser_agrees(L):
    from oracle import fpemu_oracle as O
    for mode in list(REFERENCE_MODES) + ["E4M3_RTZ", "E3M4_RTZ", "BFLOAT16_RTZ"]:
        assert O.parse_mode(mode) == L.parse_mode(mode), mode


def test_argument_errors_return_status_codes_without_a_gpu(L):
    lib = L.lib()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.addressof(buf)
    assert lib.fpemu_cast(p, p, -1, 0, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast(p, p, 4, 7, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_DTYPE
    assert lib.fpemu_cast(p, p, 4, 0, 9, 1, 0, 1.0, None, 0, 0, None) == L.ERR_MODE
    assert lib.fpemu_cast(p, p, 4, 0, 1, 9, 0, 1.0, None, 0, 0, None) == L.ERR_MODE
    assert lib.fpemu_cast(p, p, 4, 0, 4, 0, 0, 1.0, None, 0, 0, None) == L.ERR_MODE      # FLOAT16_RTZ does not exist
    assert lib.fpemu_cast(p, p, 4, 0, 1, 1, 0, 70000.0, None, 0, 0, None) == L.ERR_SCALE  # reference: exit(1)
    assert lib.fpemu_cast(None, p, 4, 0, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast(p + 2, p, 4, 0, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_ALIGN
    assert lib.fpemu_cast(p, p, 0, 0, 1, 1, 0, 1.0, None, 0, 0, None) == 0                # empty tensor: nothing launched
    assert lib.fpemu_cast_blocked(p, p, 16, 5, 0, 1, 1, 0, 0, 0, None) == L.ERR_ARG        # 16 % 5 != 0
    assert lib.fpemu_cast_blocked(p, p, 16, 0, 0, 1, 1, 0, 0, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast_blocked(p, p, 16, 8, 0, 3, 1, 0, 0, 0, None) == L.ERR_MODE       # BFLOAT16 has no block_norm
    assert lib.fpemu_cast_blocked(p, p, 16, 8, 0, 0, 1, 1, 0, 0, None) == L.ERR_MODE       # nor has E5M2_DAZ
    assert lib.fpemu_amax(p, -1, 0, p, None) == L.ERR_ARG
    assert lib.fpemu_amax(p, 4, 0, None, None) == L.ERR_ARG
    assert lib.fpemu_cast_rowwise(p, p, -1, 4, 0, 1, 1, 0, 448.0, None, 0, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast_host(p, p, 16, 0, 1, 1, 0, 1.0, 0, None, None, 16, None, None, None) == L.ERR_ARG
    assert lib.fpemu_cast_host(p, p, 16, 0, 1, 1, 0, 1.0, 0, p, p, 24, None, None, None) == L.ERR_ARG
    # doubles (fpemu_cast_f64) and the peer-memory exchange: same status codes, nothing launched
    assert lib.fpemu_cast_f64(p, p, -1, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast_f64(p, p, 4, 9, 1, 0, 1.0, None, 0, 0, None) == L.ERR_MODE
    assert lib.fpemu_cast_f64(p, p, 4, 1, 1, 0, 70000.0, None, 0, 0, None) == L.ERR_SCALE
    assert lib.fpemu_cast_f64(p + 4, p, 4, 1, 1, 0, 1.0, None, 0, 0, None) == L.ERR_ALIGN
    assert lib.fpemu_cast_f64(p, p, 0, 1, 1, 0, 1.0, None, 0, 0, None) == 0
    assert lib.fpemu_exchange_floats(291, 8) == 4 * 32 * 8 * 291
    assert lib.fpemu_exchange_floats(0, 8) == L.ERR_ARG and lib.fpemu_exchange_floats(4, 17) == L.ERR_ARG
    assert lib.fpemu_amax_multi_push(p, p, 4, p, None, 2, 0, 1, None) == L.ERR_ARG          # no peer pointers
    assert lib.fpemu_cast_multi_xchg(p, p, 4, 1, 1, 0, None, 2, 1, 448.0, None, 0, None) == L.ERR_ARG
    assert lib.fpemu_cast_multi_xchg(p, p, 4, 1, 1, 0, p, 2, 0, 448.0, None, 0, None) == L.ERR_ARG   # epoch 0 is reserved
    for code in (0, -1, -2, -3, -4, -5, -6, 1, 700):
        assert len(L.strerror(code)) > 0
    with pytest.raises(ValueError):
        L.check(L.ERR_SCALE)
    with pytest.raises(RuntimeError):
        L.check(700)


def test_product_has_no_cpu_fallback_and_never_imports_the_oracle():
    """The product path must fail loudly without the CUDA library and must not route through oracle/."""
    pkg = os.path.join(ROOT, "fp8-emulation-toolkit_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.sep + "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text, "%s mentions the oracle" % f
    import fpemu_cuda
    import torch
    with pytest.raises(RuntimeError):
        fpemu_cuda.forward(torch.ones(8), "E4M3_RNE", 8, False, 1.0, False, 1)   # CPU tensor: refused, not emulated


def test_missing_library_raises(monkeypatch, L):
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", "/nonexistent/libfpemu_b200.so")
    with pytest.raises(L.FpemuError):
        L.lib()
