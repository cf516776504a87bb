"""Load the unmodified reference builds from oracle/_ref/ (see oracle/build_ref.py).

TEST INFRASTRUCTURE ONLY (tests/, bench.py's CPU-baseline legs).  The modules are the reference's own
pybind11 torch extensions, so they expose exactly `forward(input, mode, size, inplace, scale,
block_norm, block_size)` / `backward(...)` (cpp/fpemu_impl.cpp:2013-2049, cuda/fpemu_impl.cpp:27-54).
"""
import importlib.util
import os

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def _cpu_flags():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def host_has_avx512():
    return {"avx512f", "avx512bw", "avx512dq", "avx512vl"} <= _cpu_flags()


def available(name):
    return os.path.exists(os.path.join(REF_DIR, name + ".so"))


_cache = {}


def load(name):
    """name in {fpemu_cpp_ref_scalar, fpemu_cpp_ref_avx512, fpemu_cuda_ref}.  Returns the module or None."""
    if name in _cache:
        return _cache[name]
    path = os.path.join(REF_DIR, name + ".so")
    mod = None
    if os.path.exists(path):
        if name == "fpemu_cpp_ref_avx512" and not host_has_avx512():
            mod = None
        elif name != "fpemu_cpp_ref_avx512" and not ({"avx2", "f16c"} <= _cpu_flags()):
            mod = None
        else:
            import torch  # noqa: F401  (libtorch must be loaded before the extension)
            spec = importlib.util.spec_from_file_location(name, path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
    _cache[name] = mod
    return mod


def best_cpu():
    """(module, label): the AVX-512 build when the host supports it (the reference's -march=native would
    pick it), else the scalar build, else (None, reason)."""
    m = load("fpemu_cpp_ref_avx512")
    if m is not None:
        return m, "avx512"
    m = load("fpemu_cpp_ref_scalar")
    if m is not None:
        return m, "scalar"
    return None, "oracle/_ref not built"
