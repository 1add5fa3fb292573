"""Builds variants of the library that differ in a -D of prefill_attention.cu (here, on the CPU box) and
times each on the GPU box with tools/prefill_bench.py.

    python tools/prefill_variants.py build NAME=-DPA_PREFILL_POLY_MASK=0x12 ...
    python tools/prefill_variants.py run            (on the GPU box)
"""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-v0.1.0_b200")
VDIR = os.path.join(PKG, "lib", "variants")


def build(specs):
    sys.path.insert(0, PKG)
    import build as pa_build
    pa_build.build()
    os.makedirs(VDIR, exist_ok=True)
    for f in glob.glob(os.path.join(VDIR, "*.so")):
        os.remove(f)
    nvcc = pa_build.find_nvcc()
    objs = [os.path.join(pa_build.BUILD, s.replace(".cu", ".o")) for s in pa_build.SOURCES if s != "prefill_attention.cu"]
    for spec in specs:
        name, flag = spec.split("=", 1)
        obj = os.path.join(VDIR, name + ".o")
        subprocess.run([nvcc] + pa_build.NVCC_FLAGS + flag.split() + ["-c", os.path.join(pa_build.CSRC, "prefill_attention.cu"), "-o", obj], check=True)
        lib = os.path.join(VDIR, f"libpaged_b200_{name}.so")
        subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs + [obj, "-cudart", "static"], check=True)
        os.remove(obj)
        print("built", lib)


def run():
    for lib in [None] + sorted(glob.glob(os.path.join(VDIR, "*.so"))):
        env = dict(os.environ)
        if lib:
            env["PA_LIBRARY"] = lib
            env["PA_BINDING"] = "ctypes"
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "prefill_bench.py"), "--json", "/tmp/pv.json"] + sys.argv[2:],
                           env=env, capture_output=True, text=True)
        rows = []
        for line in r.stdout.splitlines():
            try:
                row = json.loads(line)
                rows.append(f"{row['shape']}={row['ours_tflops']}")
            except Exception:
                pass
        print(os.path.basename(lib) if lib else "default", " ".join(rows) or r.stderr[-400:], flush=True)


if __name__ == "__main__":
    build(sys.argv[2:]) if sys.argv[1] == "build" else run()
