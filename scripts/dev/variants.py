"""Build compile-time variants of the library for A/B timing on the GPU box (development aid; not the product):
    python scripts/dev/variants.py name1=FLAG_A,FLAG_B name2=FLAG_C ...
-> cram_b200/_build/libcram_b200_<name>.so (travels with the gpurun snapshot; select with CRAM_B200_LIB)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cram_b200 import build as b  # noqa: E402


def build(name, defs):
    out = os.path.join(b.BUILD_DIR, f"libcram_b200_{name}.so")
    flags = [f for f in b.NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    srcs = [os.path.join(b.CSRC, s) for s in b.SOURCES]
    cmd = [b._nvcc(), *flags, *[f"-D{d}" for d in defs if d], "-shared", "-o", out, *srcs, "-cudart", "static"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        print(r.stdout[-3000:])
        raise SystemExit(f"variant {name} failed")
    return out


if __name__ == "__main__":
    jobs = []
    for a in sys.argv[1:]:
        name, _, d = a.partition("=")
        jobs.append((name, d.split(",") if d else []))
    os.makedirs(b.BUILD_DIR, exist_ok=True)
    with ThreadPoolExecutor(4) as ex:
        for out in ex.map(lambda j: build(*j), jobs):
            print("built", out)
