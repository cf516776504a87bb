#!/usr/bin/env python
"""Build the UNMODIFIED reference fpemu sources into oracle/_ref/ (test infrastructure only).

This is checker infrastructure, never the product: nothing under the product package imports it.
It compiles the reference's own translation units *where they lie* under /root/reference (no source
is copied into this repo) with plain g++ / nvcc command lines:

  fpemu_cpp_ref_avx512.so  reference CPU op, AVX-512 code paths active     (timed CPU baseline)
  fpemu_cpp_ref_scalar.so  reference CPU op, AVX-512 paths compiled out     (faithful oracle for
                           every rounding mode: the AVX-512 path ignores non-RNE deterministic
                           modes, mpemu/pytquant/cpp/fpemu_impl.cpp:818-824)
  fpemu_cuda_ref.so        reference CUDA op compiled for sm_100a           (GPU-side oracle and
                           "kernel to beat"; mpemu/pytquant/cuda/fpemu_kernels.cu)

Flags follow /root/reference/setup.py:12-16 and :26-32 with two deliberate differences:
  * `-march=native` is replaced by the explicit ISA list it expands to on an AVX-512 Xeon, because the
    .so travels to a GPU box whose host CPU may differ; bench.py picks the variant the host supports.
  * module names get a `_ref` suffix so they never shadow the product's `fpemu_cuda`.

Outputs go ONLY to oracle/_ref/ (git-ignored, not gpurun-ignored).  When /root/reference is absent
(GPU box) this script does nothing and the prebuilt files are used.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("FPEMU_REFERENCE_ROOT", "/root/reference")
CPU_SRC = os.path.join(REF, "mpemu/pytquant/cpp/fpemu_impl.cpp")
CUDA_SRCS = [os.path.join(REF, "mpemu/pytquant/cuda/fpemu_impl.cpp"),
             os.path.join(REF, "mpemu/pytquant/cuda/fpemu_kernels.cu")]

AVX512_FLAGS = ["-mavx512f", "-mavx512bw", "-mavx512dq", "-mavx512vl", "-mavx2", "-mfma", "-mbmi", "-mbmi2"]
SCALAR_FLAGS = ["-mavx2", "-mfma", "-mbmi", "-mbmi2"]  # F16C conversions stay (the scalar path needs them)


def _torch_paths():
    import torch  # noqa: F401  (heavy import only when building)
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths() + [sysconfig.get_paths()["include"]]
    lib = ce.library_paths()
    return inc, lib, list(ce.COMMON_NVCC_FLAGS)


def _newer(dst, srcs):
    if not os.path.exists(dst):
        return False
    t = os.path.getmtime(dst)
    return all(os.path.getmtime(s) <= t for s in srcs + [os.path.abspath(__file__)])


def build_cpu(name, isa_flags, inc, lib):
    dst = os.path.join(OUT, name + ".so")
    if _newer(dst, [CPU_SRC]):
        return dst
    cmd = ["/usr/bin/g++", "-O3", "-std=c++17", "-shared", "-fPIC", "-fopenmp", "-mf16c", "-mlzcnt",
           "-Wno-deprecated-declarations", "-w",
           "-DTORCH_EXTENSION_NAME=" + name, "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    cmd += isa_flags + ["-I" + p for p in inc] + [CPU_SRC, "-o", dst]
    cmd += ["-L" + p for p in lib] + ["-Wl,-rpath," + p for p in lib]
    cmd += ["-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-fopenmp"]
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return dst


def build_cuda(name, inc, lib, nvcc_common):
    dst = os.path.join(OUT, name + ".so")
    if _newer(dst, CUDA_SRCS):
        return dst
    cmd = ["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-w",
           "-ccbin", "/usr/bin/g++",
           "-gencode", "arch=compute_100a,code=sm_100a", "--expt-relaxed-constexpr",
           "-DTORCH_EXTENSION_NAME=" + name, "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"]
    cmd += nvcc_common + ["-I" + p for p in inc] + ["-I/usr/local/cuda/include"] + CUDA_SRCS + ["-o", dst]
    cmd += ["-L" + p for p in lib] + ["-Xlinker", "-rpath," + ":".join(lib)]
    cmd += ["-lc10", "-lc10_cuda", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-ltorch_python", "-lcudart"]
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return dst


def main(argv):
    if not os.path.isfile(CPU_SRC):
        print("oracle/build_ref.py: %s not present; using prebuilt oracle/_ref/ (if any)" % REF)
        return 0
    os.makedirs(OUT, exist_ok=True)
    inc, lib, nvcc_common = _torch_paths()
    want = set(argv) or {"cpu", "cuda"}
    if "cpu" in want:
        build_cpu("fpemu_cpp_ref_scalar", SCALAR_FLAGS, inc, lib)
        build_cpu("fpemu_cpp_ref_avx512", AVX512_FLAGS, inc, lib)
    if "cuda" in want:
        build_cuda("fpemu_cuda_ref", inc, lib, nvcc_common)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
