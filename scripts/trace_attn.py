"""Debug timeline of the tcgen05 attention kernel: builds a -DCRAM_ATTN_TRACE=<block> copy of the library
(cram_b200/_build/libcram_b200_trace.so, not the product), runs one launch at the bench shape and prints, for the
traced CTA, per-tile clock deltas of the softmax warps (stage 0 = warp 0, stage 1 = warp 4) and the MMA issuer.

    python scripts/trace_attn.py build [block]      (CPU box: cross-compile)
    python scripts/trace_attn.py run                (GPU box)
"""
import ctypes, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "cram_b200", "_build", "libcram_b200_trace.so")

def build(block, extra=(), out=OUT):
    from cram_b200 import build as b
    srcs = [os.path.join(b.CSRC, s) for s in b.SOURCES]
    flags = [f for f in b.NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    defs = ([f"-DCRAM_ATTN_TRACE={block}"] if block >= 0 else []) + [f"-D{d}" for d in extra]
    cmd = [b._nvcc(), *flags, *defs, "-shared", "-o", out, *srcs, "-cudart", "static"]
    subprocess.run(cmd, check=True)
    print("built", out)

def run():
    os.environ["CRAM_B200_LIB"] = os.environ.get("TRACE_LIB", OUT)
    import numpy as np, torch
    from cram_b200 import ops, _lib
    B, T, H, Hkv, dh = int(os.environ.get("TRACE_B", 16)), int(os.environ.get("TRACE_T", 2048)), 32, 8, 128
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    lens = [T - 3 * i for i in range(B)]
    total = sum(lens)
    qkv = torch.randn(total, (H + 2 * Hkv) * dh, device=dev, generator=g).bfloat16()
    q, k, v = qkv[:, :H * dh], qkv[:, H * dh:(H + Hkv) * dh], qkv[:, (H + Hkv) * dh:]
    cu = torch.zeros(B + 1, dtype=torch.int32); cu[1:] = torch.tensor(lens).cumsum(0); cu = cu.to(dev)
    lc = torch.zeros(total, device=dev).bfloat16(); lc[100:900] = -0.69
    causal = os.environ.get("TRACE_CAUSAL", "1") != "0"
    for _ in range(3):
        ops.attn_prefill(q, k, v, cu, T, H, Hkv, dh, logcred=lc, head_mask=0x0FFFFFFF, kernel=2, causal=causal)
    torch.cuda.synchronize()
    lib = _lib.load()
    buf = np.zeros((24, 64, 4), dtype=np.int64)
    lib.cram_debug_read_trace.argtypes = [ctypes.c_void_p]
    assert lib.cram_debug_read_trace(buf.ctypes.data) == 0
    t0 = buf[buf > 0].min() if (buf > 0).any() else 0  # (the per-tile stamps predate the persistent kernel: usually empty)
    slots = ((0, "softmax stage0 (warp 0)"), (4, "softmax stage1 (warp 4)"), (8, "mma stage0"), (9, "mma stage1"))
    if os.environ.get("TRACE_SLOTS"):
        slots = tuple((int(x), f"slot {x}") for x in os.environ["TRACE_SLOTS"].split(","))
    for slot, name in slots:
        print(f"--- {name}: per tile [e0 rel start] d(e0->e1) d(e1->e2) d(e2->e3) | iteration period")
        prev = None
        for j in range(64):
            e = buf[slot, j]
            if e[0] == 0:
                break
            per = (e[0] - prev) if prev is not None else 0
            prev = e[0]
            print(f"  j={j:2d} start={e[0]-t0:7d}  {e[1]-e[0]:6d} {e[2]-e[1]:6d} {e[3]-e[2]:6d} | {per:6d}")

    if hasattr(lib, "cram_debug_read_evlog"):
        ev = np.zeros((8, 2048, 2), dtype=np.int64)
        lib.cram_debug_read_evlog.argtypes = [ctypes.c_void_p]
        assert lib.cram_debug_read_evlog(ev.ctypes.data) == 0
        names = {0: "P take.start", 1: "P published next", 2: "P first loads issued", 3: "P all loads issued", 4: "M s0 PV+QK issued",
                 5: "M s1 PV+QK issued", 6: "M next item read", 7: "M next first QK issued", 8: "S item read", 9: "S S(j) seen",
                 10: "S tile done", 11: "S o_final seen", 12: "S epilogue done", 13: "S epilogue: O chunk in registers",
                 14: "S half 1 done / epilogue: box free", 15: "S half 0 done"}
        role = {0: "softmax0", 1: "softmax1", 2: "mma", 3: "producer", 4: "softmax0.w3", 5: "softmax1.w3"}
        lo, hi = (int(x) for x in os.environ.get("TRACE_ITEMS", "6,9").split(","))
        rows = []
        for slot in range(6):
            for n in range(2048):
                c, tag = int(ev[slot, n, 0]), int(ev[slot, n, 1])
                if c == 0:
                    break
                rows.append((c, slot, tag >> 16, (tag >> 4) & 0xFFF, tag & 15))
        rows.sort()
        sel = [r for r in rows if lo <= r[2] <= hi]
        if sel:
            t0 = sel[0][0]
            print(f"--- event log of the traced CTA, items {lo}..{hi} (cycles from the first event shown)")
            last = {}
            for c, slot, item, tile, kind in sel:
                d = c - last.get(slot, c)
                last[slot] = c
                print(f"  {c - t0:8d}  (+{d:6d})  {role[slot]:9s} item {item:3d} tile {tile:3d}  {names.get(kind, kind)}")
        # per-item summary for softmax1: item read -> first S seen, last tile done -> o_final -> epilogue done -> next item read
        print("--- per item (softmax stage 1): tiles | read->S0 | tiles phase | last tile->o_final | epilogue | total")
        cur = {}
        for c, slot, item, tile, kind in rows:
            if slot != 1:
                continue
            d = cur.setdefault(item, {})
            if kind == 8: d["read"] = c; d["nt"] = tile
            if kind == 9 and "s0" not in d: d["s0"] = c
            if kind == 10: d["last"] = c
            if kind == 11: d["of"] = c
            if kind == 12: d["done"] = c
        for item in sorted(cur):
            d = cur[item]
            if all(k in d for k in ("read", "s0", "last", "of", "done")):
                print(f"  item {item:3d}: {d['nt']:2d} | {d['s0'] - d['read']:6d} | {d['last'] - d['s0']:7d} | {d['of'] - d['last']:6d} | {d['done'] - d['of']:6d} | {d['done'] - d['read']:7d}")

    if hasattr(lib, "cram_debug_read_trace_fine"):
        fine = np.zeros((16, 16), dtype=np.int64)
        lib.cram_debug_read_trace_fine.argtypes = [ctypes.c_void_p]
        lib.cram_debug_read_trace_fine(fine.ctypes.data)
        print("--- fine stamps of warp 0 (tiles 16..31): deltas between consecutive stamps 0..12 "
              "(0 entry | c0 data, c0 done | c1 data, c1 done | c2 .. | c3 .. | 9 decided | 10 P issued | 11 st waited | 12 arrived)")
        for r in fine:
            if r[0] == 0:
                continue
            print("  " + " ".join(f"{int(r[i + 1] - r[i]):5d}" for i in range(12)))

if __name__ == "__main__":
    if sys.argv[1] == "build":      # build <block | -1 = no trace> [DEFINE=VAL ...] [out=path]
        outs = [a[4:] for a in sys.argv[3:] if a.startswith("out=")]
        build(int(sys.argv[2]) if len(sys.argv) > 2 else 0, [a for a in sys.argv[3:] if not a.startswith("out=")],
              outs[0] if outs else OUT)
    else:
        run()
