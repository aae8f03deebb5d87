#!/usr/bin/env python3
"""TEST INFRASTRUCTURE — builds the reference's own implementation of the hot path.

Compiles, from the sources where they lie under /root/reference (never copied
into this repo), into oracle/_ref/ (git-ignored, NOT gpurun-ignored):

  ref_cpu_bf16.so / ref_cpu.so   csrc/cpu/attention.cpp + csrc/cpu/cache.cpp
                                 (AVX512 + OpenMP; with / without AVX512-BF16)
  ref_cuda.so                    csrc/attention/attention_kernels.cu +
                                 csrc/cache_kernels.cu + csrc/pos_encoding_kernels.cu
                                 for sm_100a (≈15 min)

Both are bound through the small pybind shims in this directory
(ref_cpu_shim.cpp, ref_cuda_shim.cpp).  The reference's own build system
(CMake + network-fetched CUTLASS + torch 2.3 pin) is not run.

  prefix_prefill_ref.py          vllm/attention/ops/prefix_prefill.py staged verbatim (Triton; runs on the GPU box)

  refsuite/                      the reference's tests/kernels/test_{cache,attention}.py + the Python glue they import,
                                 staged verbatim (run over this repo's extension by tests/test_gpu_reference_suite.py)

Usage: python oracle/build_ref.py [cpu] [cuda] [prefill] [suite] [--force]
"""
import os
import subprocess
import sys
import sysconfig
import time

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("B200PA_REFERENCE", "/root/reference")
CSRC = os.path.join(REF, "csrc")


def _torch_flags():
    import torch
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return inc, libdir, abi


def _run(cmd, **kw):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd, **kw)


def _fresh(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def build_cpu(force=False):
    """csrc/cpu/attention.cpp, cache.cpp unchanged; flags as cmake/cpu_extension.cmake:33-60."""
    inc, libdir, abi = _torch_flags()
    srcs = [os.path.join(CSRC, "cpu", "attention.cpp"),
            os.path.join(CSRC, "cpu", "cache.cpp"),
            os.path.join(HERE, "ref_cpu_shim.cpp")]
    os.makedirs(OUT, exist_ok=True)
    for name, extra in (("ref_cpu_bf16", ["-mavx512bf16"]), ("ref_cpu", [])):
        target = os.path.join(OUT, name + ".so")
        if not force and _fresh(target, srcs):
            print("up to date:", target)
            continue
        cmd = ["g++", "-O3", "-fopenmp", "-fPIC", "-shared", "-std=c++17",
               "-mavx512f", "-mavx512vl", "-mavx512bw", "-mavx512dq", *extra,
               "-DVLLM_CPU_EXTENSION", f"-DTORCH_EXTENSION_NAME={name}",
               f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-DTORCH_API_INCLUDE_EXTENSION_H",
               "-w", "-I", os.path.join(CSRC, "cpu"), *inc, *srcs,
               "-o", target, f"-L{libdir}", f"-Wl,-rpath,{libdir}",
               "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python",
               "-L/usr/lib/gcc/x86_64-linux-gnu/13", "-lgomp"]
        _run(cmd)


def build_cuda(force=False):
    """csrc/attention/attention_kernels.cu, csrc/cache_kernels.cu unchanged for sm_100a;
    flags as CMakeLists.txt / cmake/utils.cmake:97-110 (+ENABLE_FP8), arch replaced."""
    inc, libdir, abi = _torch_flags()
    os.makedirs(os.path.join(OUT, "build"), exist_ok=True)
    cu = [os.path.join(CSRC, "attention", "attention_kernels.cu"),
          os.path.join(CSRC, "cache_kernels.cu"),
          os.path.join(CSRC, "pos_encoding_kernels.cu")]
    shim = os.path.join(HERE, "ref_cuda_shim.cpp")
    target = os.path.join(OUT, "ref_cuda.so")
    if not force and _fresh(target, cu + [shim]):
        print("up to date:", target)
        return
    objs = []
    procs = []
    for s in cu:
        o = os.path.join(OUT, "build", os.path.basename(s) + ".o")
        objs.append(o)
        if not force and _fresh(o, [s]):  # the two big translation units take ~10 min each: reuse
            print("up to date:", o)
            continue
        cmd = ["nvcc", "-c", s, "-o", o, "-O2", "-std=c++17", "--threads", "4",
               "-gencode", "arch=compute_100a,code=sm_100a", "-DENABLE_FP8",
               "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-w",
               "-DTORCH_EXTENSION_NAME=ref_cuda", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
               "-DTORCH_API_INCLUDE_EXTENSION_H",
               "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
               "-D__CUDA_NO_BFLOAT16_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__",
               "-I", CSRC, *inc]
        print("+", " ".join(cmd), flush=True)
        procs.append(subprocess.Popen(cmd))
    so = os.path.join(OUT, "build", "shim.o")
    _run(["g++", "-c", shim, "-o", so, "-O2", "-fPIC", "-std=c++17", "-w",
          "-DTORCH_EXTENSION_NAME=ref_cuda", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
          "-DTORCH_API_INCLUDE_EXTENSION_H", *inc])
    for p in procs:
        if p.wait() != 0:
            raise SystemExit("nvcc failed")
    _run(["g++", "-shared", "-o", target, so, *objs, f"-L{libdir}",
          f"-Wl,-rpath,{libdir}", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10",
          "-lc10_cuda", "-ltorch_python", "-L/usr/local/cuda/lib64", "-lcudart"])


def stage_prefix_prefill():
    """The reference's prefix-prefill is a Triton (Python) module: it cannot be compiled here, and it needs a
    GPU to run.  Stage it VERBATIM into oracle/_ref/ (git-ignored, so it never enters the repo history, but it
    travels to the GPU box like the compiled reference objects) so that tests/test_gpu_prefill.py and
    tools/bench_prefill.py can run the reference's own kernels beside ours."""
    import shutil
    src = os.path.join(REF, "vllm", "attention", "ops", "prefix_prefill.py")
    os.makedirs(OUT, exist_ok=True)
    shutil.copyfile(src, os.path.join(OUT, "prefix_prefill_ref.py"))
    print("staged", src)


SUITE_FILES = (
    "vllm/_custom_ops.py", "vllm/utils.py", "vllm/envs.py", "vllm/logger.py", "vllm/logging/__init__.py",
    "vllm/logging/formatter.py", "vllm/attention/ops/paged_attn.py", "vllm/attention/ops/prefix_prefill.py",
    "tests/kernels/test_cache.py", "tests/kernels/test_attention.py", "tests/kernels/conftest.py",
    "tests/kernels/allclose_default.py",
)


def stage_reference_suite():
    """The reference's own kernel tests and the Python glue between them and `vllm._C`, staged VERBATIM into
    oracle/_ref/refsuite/ (git-ignored; travels to the GPU box) so that tests/test_gpu_reference_suite.py can run
    them, unmodified, over this repo's extension mounted as `vllm._C`."""
    import shutil
    for rel in SUITE_FILES:
        dst = os.path.join(OUT, "refsuite", rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF, rel), dst)
    print("staged", len(SUITE_FILES), "reference files into", os.path.join(OUT, "refsuite"))


def main():
    if not os.path.isdir(CSRC):
        print(f"{REF} not present: keeping prebuilt oracle/_ref as is")
        return 0
    args = [a for a in sys.argv[1:] if not a.startswith("-")] or ["cpu"]
    force = "--force" in sys.argv
    t0 = time.time()
    if "cpu" in args:
        build_cpu(force)
    if "cuda" in args:
        build_cuda(force)
    if "prefill" in args or "cuda" in args:
        stage_prefix_prefill()
    if "suite" in args or "cuda" in args:
        stage_reference_suite()
    print(f"done in {time.time() - t0:.0f}s")
    return 0


if __name__ == "__main__":
    sys.exit(main())
