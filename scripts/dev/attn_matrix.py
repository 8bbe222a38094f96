"""Time the tcgen05 prefill attention kernel over a shape matrix, for the product library and every variant library
cram_b200/_build/libcram_b200_<name>.so (one subprocess per library: CRAM_B200_LIB is read at load time).
    python scripts/dev/attn_matrix.py            # all libraries
    python scripts/dev/attn_matrix.py --one      # (internal) the library in CRAM_B200_LIB
CRAM_B200_NO_DEAD_TILES=1 in the environment times the same kernel with dead-tile skipping off.
"""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

SHAPES = [  # name, B, T, H, Hkv, causal
    ("c16x2048", 16, 2048, 32, 8, True),
    ("n16x2048", 16, 2048, 32, 8, False),
    ("c4x8192", 4, 8192, 32, 8, True),
    ("c1x32768", 1, 32768, 32, 8, True),
    ("c64x512", 64, 512, 32, 8, True),
    # the bench workload's credibility pattern: one zero-credibility document per prompt (keys 20..419 -> tiles 1 and 2
    # are dead on the 28 biased heads); TF = the FULL causal FLOPs / time, so skipping shows as a higher figure
    ("z16x2048", 16, 2048, 32, 8, True),
]


def one():
    import torch
    from cram_b200 import ops
    dev, dh = "cuda", 128
    out = []
    only = os.environ.get("ATTN_SHAPES")
    for name, B, T, H, Hkv, causal in SHAPES:
        if only and name not in only.split(","):
            continue
        g = torch.Generator(device=dev).manual_seed(0)
        lens = [T - 3 * i for i in range(B)]
        total = sum(lens)
        qkv = torch.randn(total, (H + 2 * Hkv) * dh, device=dev, generator=g).bfloat16()
        q, k, v = qkv[:, :H * dh], qkv[:, H * dh:(H + Hkv) * dh], qkv[:, (H + Hkv) * dh:]
        cu = torch.zeros(B + 1, dtype=torch.int32)
        cu[1:] = torch.tensor(lens).cumsum(0)
        cu = cu.to(dev)
        lc = torch.zeros(total, device=dev).bfloat16()
        lc[100:min(900, total)] = -0.69
        if name.startswith("z"):
            for b in range(B):
                lc[int(cu[b]) + 20:int(cu[b]) + 420] = float("-inf")
        flops = 4.0 * dh * H * sum((t * (t + 1) / 2 if causal else t * t) for t in lens)
        reps = 10 if T <= 8192 else 4
        run = lambda: ops.attn_prefill(q, k, v, cu, T, H, Hkv, dh, logcred=lc, head_mask=0x0FFFFFFF, causal=causal, kernel=2)
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out.append(f"{name} {ms:.3f}ms {flops / ms / 1e9:.0f}TF")
        del qkv, q, k, v
    print(" | ".join(out), flush=True)


if __name__ == "__main__":
    if "--one" in sys.argv:
        one()
    else:
        libs = [("product", None)] + [(os.path.basename(p)[len("libcram_b200_"):-3], p)
                                      for p in sorted(glob.glob(os.path.join(ROOT, "cram_b200", "_build", "libcram_b200_*.so")))]
        for name, path in libs:
            if name == "trace":
                continue
            env = dict(os.environ)
            if path:
                env["CRAM_B200_LIB"] = path
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one"], env=env, capture_output=True, text=True, timeout=600)
            print(f"{name:>14}: {r.stdout.strip() or r.stderr[-400:]}", flush=True)
