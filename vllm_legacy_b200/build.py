"""In-tree build of the B200 paged-attention libraries (sm_100a only).

  libb200pa.so   C-ABI shared library (include/b200pa.h): nvcc, no torch, static cudart
  _C.so          thin PyTorch extension over it (vllm._C drop-in: ops / cache_ops / cuda_utils)

Both land next to this file so they travel with the repo snapshot to the GPU box.
Usage: python -m vllm_legacy_b200.build [--force] [--verbose]
"""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys
import sysconfig

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
INC = os.path.join(ROOT, "include")
OBJ = os.environ.get("B200PA_BUILD_DIR", "/tmp/b200pa_build")
LIB = os.path.join(PKG, "libb200pa.so")
EXT = os.path.join(PKG, "_C.so")

HEAD_SIZES = (64, 80, 96, 112, 128, 192, 256)
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-I", INC, "-I", CSRC]
NVCC_FLAGS += os.environ.get("B200PA_EXTRA_NVCC_FLAGS", "").split()  # developer builds (e.g. -DB200PA_PREFILL_TRACE)


def _headers():
    hs = [os.path.join(INC, "b200pa.h")]
    hs += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    return hs


def _stamp(cmd, deps):
    h = hashlib.sha256(" ".join(cmd).encode())
    for d in sorted(deps):
        with open(d, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _compile(cmd, out, deps, force, verbose):
    stamp_file = out + ".stamp"
    stamp = _stamp(cmd, deps)
    if not force and os.path.exists(out) and os.path.exists(stamp_file):
        with open(stamp_file) as f:
            if f.read() == stamp:
                return False
    if verbose:
        print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    with open(stamp_file, "w") as f:
        f.write(stamp)
    return True


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs = []
    for name in ("cache.cu", "cache_flash.cu", "rope.cu", "prefill.cu", "prefill_tc.cu", "attention.cu", "attention_simt.cu", "attention_mma.cu",
                 "attention_flash.cu", "attention_noalloc.cu"):
        src = os.path.join(CSRC, name)
        obj = os.path.join(OBJ, name + ".o")
        jobs.append((["nvcc", *NVCC_FLAGS, "-c", src, "-o", obj], obj, [src, *hdrs]))
    inst = os.path.join(CSRC, "attention_mma_inst.cu")
    for d in HEAD_SIZES:
        obj = os.path.join(OBJ, f"attention_mma_inst_d{d}.o")
        jobs.append((["nvcc", *NVCC_FLAGS, f"-DB200PA_INST_D={d}", "-c", inst, "-o", obj], obj, [inst, *hdrs]))
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        changed = list(ex.map(lambda j: _compile(j[0], j[1], j[2], force, verbose), jobs))
    objs = [j[1] for j in jobs]
    # The in-tree library remembers which objects it was linked from (a build from another B200PA_BUILD_DIR, e.g. a
    # developer build with extra flags, must not survive a later default build whose objects are all up to date).
    link_stamp = _stamp(["link", *NVCC_FLAGS], [o + ".stamp" for o in objs])
    link_stamp_file = LIB + ".stamp"
    linked = os.path.exists(LIB) and os.path.exists(link_stamp_file) and open(link_stamp_file).read() == link_stamp
    relink = any(changed) or not linked
    if relink:
        cmd = ["nvcc", "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        if verbose:
            print("+", " ".join(cmd), flush=True)
        subprocess.check_call(cmd)
        with open(link_stamp_file, "w") as f:
            f.write(link_stamp)

    # ---- torch extension ------------------------------------------------------------------------
    import torch
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths("cuda"):
        inc += ["-isystem", p]
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    src = os.path.join(CSRC, "torch_binding.cpp")
    obj = os.path.join(OBJ, "torch_binding.o")
    cmd = ["g++", "-c", src, "-o", obj, "-O2", "-fPIC", "-std=c++17", "-w", "-DTORCH_EXTENSION_NAME=_C",
           f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-DTORCH_API_INCLUDE_EXTENSION_H", "-I", INC, *inc]
    ch = _compile(cmd, obj, [src, os.path.join(INC, "b200pa.h")], force, verbose)
    if ch or relink or not os.path.exists(EXT):
        cmd = ["g++", "-shared", "-o", EXT, obj, f"-L{PKG}", "-lb200pa", "-Wl,-rpath,$ORIGIN",
               f"-L{tlib}", f"-Wl,-rpath,{tlib}", "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda",
               "-ltorch_python"]
        if verbose:
            print("+", " ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return LIB, EXT


if __name__ == "__main__":
    lib, ext = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv)
    print("built", lib, ext)
