"""Time the prefill-sized glue kernels (HBM streams) with CUDA events: GB/s of algorithmic bytes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cram_b200 import ops

dev = "cuda"
rows, hidden, inter = 32852, 4096, 14336


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


gu = torch.randn(rows, 2 * inter, device=dev).bfloat16()
out = torch.empty(rows, inter, device=dev, dtype=torch.bfloat16)
ms = timed(lambda: ops.silu_mul(gu, out=out))
print(f"silu_mul  [{rows} x 2x{inter}] {ms * 1e3:8.1f} us  {rows * inter * 3 * 2 / ms / 1e6:7.1f} GB/s")
del gu, out
x = torch.randn(rows, hidden, device=dev).bfloat16()
r = torch.randn(rows, hidden, device=dev).bfloat16()
w = torch.ones(hidden, device=dev).bfloat16()
o = torch.empty_like(x)
ms = timed(lambda: ops.rmsnorm(x, w, 1e-5, residual=r, out=o))
print(f"rmsnorm+res [{rows} x {hidden}] {ms * 1e3:8.1f} us  {rows * hidden * 4 * 2 / ms / 1e6:7.1f} GB/s")
ms = timed(lambda: ops.rmsnorm(x, w, 1e-5, out=o))
print(f"rmsnorm     [{rows} x {hidden}] {ms * 1e3:8.1f} us  {rows * hidden * 2 * 2 / ms / 1e6:7.1f} GB/s")
qkv = torch.randn(rows, 6144, device=dev).bfloat16()
cos = torch.randn(4096, 128, device=dev).bfloat16()
pos = torch.randint(0, 2048, (rows,), device=dev, dtype=torch.int32)
ms = timed(lambda: ops.rope_(qkv, 40, 128, cos, cos, pos))
print(f"rope        [{rows} x 40 heads] {ms * 1e3:8.1f} us  {rows * 40 * 128 * 2 * 2 / ms / 1e6:7.1f} GB/s")
