"""Builds the REFERENCE's own CUDA operators for this path into oracle/_ref/ -- test infrastructure.

    python oracle/build_ref.py          (only where /root/reference exists, i.e. the build container)

The reference's csrc/attention.cpp + csrc/attention/attention_kernels.cu and csrc/cache.cpp +
csrc/cache_kernels.cu are compiled UNMODIFIED, from where they lie under /root/reference, with
plain nvcc (the reference's setup.py / build system is not run) against the torch headers of
this image, for compute_100/sm_100 -- the reference itself ships no Blackwell target
(setup.py:53 falls back to 70..90).  Outputs: oracle/_ref/ref_attention_ops.so and
oracle/_ref/ref_cache_ops.so, oracle/_ref/ref_pos_encoding_ops.so (pybind11 modules, git-ignored; they travel to the GPU box with
the snapshot).  No reference source is copied into this repository.

Use: tests/test_gpu_vs_reference_kernels.py compares our kernels with these on the same GPU
tensors, and tools/compare_reference_kernel.py times them side by side.  The product never
loads them.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")

MODULES = {
    "ref_attention_ops": ["csrc/attention.cpp", "csrc/attention/attention_kernels.cu"],
    "ref_cache_ops": ["csrc/cache.cpp", "csrc/cache_kernels.cu"],
    "ref_pos_encoding_ops": ["csrc/pos_encoding.cpp", "csrc/pos_encoding_kernels.cu"],
}


def _flags():
    import torch
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths("cuda"):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"]]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    common = ["-O2", "-std=c++17", "-Xcompiler", "-fPIC", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
              "-DTORCH_API_INCLUDE_EXTENSION_H", "-gencode", "arch=compute_100,code=sm_100",
              "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__", "-w"]
    link = ["-shared", "-L", libdir, "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
            "-Xlinker", f"-rpath={libdir}"]
    return inc, common, link


def build(force: bool = False) -> dict:
    if not os.path.isdir(REF):
        raise RuntimeError(f"{REF} is absent: oracle/_ref can only be built in the build container")
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(OUT, exist_ok=True)
    inc, common, link = _flags()
    built = {}
    for name, srcs in MODULES.items():
        so = os.path.join(OUT, name + ".so")
        paths = [os.path.join(REF, s) for s in srcs]
        if not force and os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(p) for p in paths):
            built[name] = so
            continue
        objs = []
        for p in paths:
            obj = os.path.join(OUT, name + "_" + os.path.basename(p).rsplit(".", 1)[0] + ".o")
            cmd = [nvcc] + common + inc + [f"-DTORCH_EXTENSION_NAME={name}", "-c", p, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"reference source {p} did not compile:\n{r.stderr[-3000:]}")
            objs.append(obj)
        r = subprocess.run([nvcc] + objs + link + ["-o", so], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link of {name} failed:\n{r.stderr[-3000:]}")
        for o in objs:
            os.remove(o)
        built[name] = so
    return built


if __name__ == "__main__":
    print(build("--force" in sys.argv))
