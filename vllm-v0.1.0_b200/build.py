"""Builds lib/libpaged_b200.so (sm_100a only) with plain nvcc, in-tree.

    python vllm-v0.1.0_b200/build.py [--force] [--verbose]

Each .cu is compiled to an object in parallel, then linked with the static CUDA
runtime, so the library has no dependency on torch or on a system libcudart.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libpaged_b200.so")

SOURCES = ["pa_api.cu", "cache_movers.cu", "step_tables.cu", "probe.cu", "prefill_attention.cu", "allreduce.cu", "decode_f16.cu", "decode_bf16.cu", "decode_f32.cu"]
ABI_HEADER = os.path.join(ROOT, "include", "paged_attention_b200.h")
HEADERS = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))) + [ABI_HEADER]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libpaged_b200 cannot be built")


def _digest(paths) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _deps(src: str):
    """The source plus the project headers it includes, transitively (quoted includes only), so that an
    edit to one kernel family's header does not recompile the others."""
    seen, todo = [], [os.path.join(CSRC, src)]
    while todo:
        p = todo.pop()
        if p in seen:
            continue
        seen.append(p)
        with open(p) as f:
            for line in f:
                if line.startswith('#include "'):
                    name = line.split('"')[1]
                    cand = [os.path.join(CSRC, name), os.path.join(ROOT, "include", name)]
                    todo += [c for c in cand if os.path.exists(c)]
    return [seen[0]] + sorted(seen[1:])


def _compile(nvcc: str, src: str, verbose: bool) -> str:
    obj = os.path.join(BUILD, src.replace(".cu", ".o"))
    stamp = obj + ".sha"
    want = _digest(_deps(src))
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == want:
        return obj
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    with open(stamp, "w") as f:
        f.write(want)
    return obj


TORCH_OPS = os.path.join(LIBDIR, "pa_torch_ops.so")


def build_torch_binding(force: bool = False) -> str:
    """lib/pa_torch_ops.so: csrc/torch_binding.cpp (pybind11, torch::Tensor& signatures of the
    reference) compiled with the host C++ compiler and linked against libpaged_b200.so."""
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_binding.cpp")
    stamp = TORCH_OPS + ".sha"
    # the binding is rebuilt whenever the C ABI header or the library it links changes
    want = _digest([src, ABI_HEADER, LIB]) + torch.__version__
    if not force and os.path.exists(TORCH_OPS) and os.path.exists(stamp) and open(stamp).read() == want:
        return TORCH_OPS
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-fvisibility=hidden",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-DTORCH_EXTENSION_NAME=pa_torch_ops", "-I", os.path.join(ROOT, "include"),
           "-I", sysconfig.get_paths()["include"]]
    for inc in ce.include_paths("cuda"):
        cmd += ["-isystem", inc]
    cmd += [src, "-o", TORCH_OPS, "-L", LIBDIR, "-lpaged_b200", "-Wl,-rpath,$ORIGIN",
            "-L", tlib, "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch", "-ltorch_python", f"-Wl,-rpath,{tlib}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"torch binding failed to build:\n{r.stderr[-4000:]}")
    with open(stamp, "w") as f:
        f.write(want)
    return TORCH_OPS


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = find_nvcc()
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            os.remove(os.path.join(BUILD, f))
    with cf.ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        objs = list(ex.map(lambda s: _compile(nvcc, s, verbose), SOURCES))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + \
              ["-cudart", "static", "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    build_torch_binding(force)
    print(f"[build] {os.path.relpath(LIB, ROOT)} sha256 {_digest([LIB])[:16]} built from sources {source_digest()[:16]}",
          file=sys.stderr)
    return LIB


def source_digest() -> str:
    """Digest of every source the library is compiled from (flags included): printed by build() so a
    stale lib/*.so on the GPU box is visible in the log."""
    return _digest([os.path.join(CSRC, s) for s in SOURCES] + HEADERS + [os.path.join(CSRC, "torch_binding.cpp")])


def stale_objects():
    """Sources whose object stamp does not match (no nvcc needed): empty list = the .so is current."""
    out = []
    for src in SOURCES:
        stamp = os.path.join(BUILD, src.replace(".cu", ".o")) + ".sha"
        want = _digest(_deps(src))
        if not (os.path.exists(stamp) and open(stamp).read() == want):
            out.append(src)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
