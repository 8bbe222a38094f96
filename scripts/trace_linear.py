"""Debug timeline of the decode projection kernel: builds a -DCRAM_LINEAR_TRACE copy of the library
(cram_b200/_build/libcram_b200_lintrace.so, not the product), replays a CUDA graph of 8 launches of one shape and
prints, per launch, %globaltimer stamps relative to the first CTA's entry of the first traced launch.

    python scripts/trace_linear.py build            (CPU box: cross-compile)
    python scripts/trace_linear.py run [qkv|o|gu|down] [split]   (GPU box)
stamps: 0 entry (producer lane)  1 after CTA sync  2 after griddepcontrol.wait  3 first stage landed  4 last MMA issued
        5 accumulator in registers  6 after the reduce-scatter barrier  7 stores done
"""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "cram_b200", "_build", "libcram_b200_lintrace.so")


def build():
    from cram_b200 import build as b
    srcs = [os.path.join(b.CSRC, s) for s in b.SOURCES]
    flags = [f for f in b.NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    subprocess.run([b._nvcc(), *flags, "-DCRAM_LINEAR_TRACE", "-shared", "-o", OUT, *srcs, "-cudart", "static"], check=True)
    print("built", OUT)


def run(which, split, pdl):
    os.environ["CRAM_B200_LIB"] = OUT
    import numpy as np, torch
    from cram_b200 import ops, _lib
    lib = _lib.load()
    lib.cram_debug_linear_split.argtypes = [ctypes.c_int]
    lib.cram_debug_pdl.argtypes = [ctypes.c_int]
    lib.cram_debug_linear_split(split)
    lib.cram_debug_pdl(pdl)
    H, I, QKV = 3584, 18944, 4608
    n_rows_w, k, mode = {"qkv": (QKV, H, 0), "o": (H, H, 1), "gu": (2 * I, H, 2), "down": (H, I, 1)}[which]
    dev = "cuda"
    L = 8
    ws = [(torch.randn(n_rows_w, k, device=dev) * 0.02).bfloat16() for _ in range(L)]
    x = torch.randn(64, k, device=dev).bfloat16()
    y = torch.zeros(64, n_rows_w // 2 if mode == 2 else n_rows_w, device=dev, dtype=torch.bfloat16)

    def fn():
        for w in ws:
            ops.linear_decode(x, w, None, mode=mode, out=y)

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()   # launches 8..15 -> trace slots 8..15
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    buf = np.zeros((32, 320, 8), dtype=np.uint64)
    lib.cram_debug_linear_trace.argtypes = [ctypes.c_void_p]
    assert lib.cram_debug_linear_trace(buf.ctypes.data) == 0
    t = buf[8:16].astype(np.int64)
    n_cta = int((t[0, :, 0] > 0).sum())
    t0 = t[0, :n_cta, 0].min()
    print(f"{which} split {split} pdl {pdl}: {n_cta} CTAs; ns relative to the first entry of launch 0 (min / median / max over CTAs)")
    for l in range(L):
        row = []
        for e in range(8):
            v = t[l, :n_cta, e]
            v = v[v > 0] - t0
            row.append("      -      " if v.size == 0 else f"{v.min():6d}/{int(np.median(v)):6d}/{v.max():6d}")
        print(f"launch {l}: " + "  ".join(row))


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build()
    else:
        run(sys.argv[2] if len(sys.argv) > 2 else "qkv", int(sys.argv[3]) if len(sys.argv) > 3 else 0,
            int(sys.argv[4]) if len(sys.argv) > 4 else 1)
