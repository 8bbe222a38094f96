"""Time the decode-step projections (cram_linear_decode) against the library GEMM + glue they replace, inside CUDA
graphs that rotate over 28 layers' weights (so neither the Python call nor the L2 hides anything).
    python scripts/profile_linear.py [replays]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from cram_b200 import ops

REPLAYS = int(sys.argv[1]) if len(sys.argv) > 1 else 10
L = 28
dev = "cuda"
H, I, QKV = 3584, 18944, 4608
torch.manual_seed(0)


def graph_time(fn):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(REPLAYS):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / REPLAYS / L  # us per launch


def case(name, n_rows_w, k, mode, bias=False):
    ws = [(torch.randn(n_rows_w, k, device=dev) * 0.02).to(torch.bfloat16) for _ in range(L)]
    x = torch.randn(64, k, device=dev).to(torch.bfloat16)
    b = torch.randn(n_rows_w, device=dev).to(torch.bfloat16) if bias else None
    n_out = n_rows_w // 2 if mode == 2 else n_rows_w
    y = torch.zeros(64, n_out, device=dev, dtype=torch.bfloat16)

    def ours():
        for w in ws:
            ops.linear_decode(x, w, b, mode=mode, out=y)

    def lib():
        for w in ws:
            if mode == 0:
                F.linear(x, w, b)
            elif mode == 1:
                y.add_(F.linear(x, w))
            else:
                ops.silu_mul(F.linear(x, w))

    t_o, t_l = graph_time(ours), graph_time(lib)
    mb = n_rows_w * k * 2 / 1e6
    print(f"{name:28s} W {mb:7.1f} MB  ours {t_o:7.2f} us ({mb / t_o * 1e3 / 1e3:5.2f} TB/s)   library+glue {t_l:7.2f} us ({mb / t_l:5.2f} TB/s)",
          flush=True)


if len(sys.argv) > 2 and sys.argv[2] == "eager":  # for ncu: a few plain launches of each shape
    for n_rows_w, k, mode in ((QKV, H, 0), (H, H, 1), (2 * I, H, 2), (H, I, 1)):
        w = (torch.randn(n_rows_w, k, device=dev) * 0.02).to(torch.bfloat16)
        x = torch.randn(64, k, device=dev).to(torch.bfloat16)
        y = torch.zeros(64, n_rows_w // 2 if mode == 2 else n_rows_w, device=dev, dtype=torch.bfloat16)
        for _ in range(2):
            ops.linear_decode(x, w, None, mode=mode, out=y)
    torch.cuda.synchronize()
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == "chain":
    # the MLP half of a decode layer as the engine runs it: rmsnorm -> gate/up (+SwiGLU) -> down (+residual)
    from cram_b200 import _lib
    import ctypes
    lib = _lib.load()
    lib.cram_debug_pdl.argtypes = [ctypes.c_int]
    wgu = [(torch.randn(2 * I, H, device=dev) * 0.02).to(torch.bfloat16) for _ in range(L)]
    wd = [(torch.randn(H, I, device=dev) * 0.02).to(torch.bfloat16) for _ in range(L)]
    nw = torch.ones(H, device=dev, dtype=torch.bfloat16)
    x = torch.randn(64, H, device=dev).to(torch.bfloat16)

    def chain(gu_ours, down_ours):
        def fn():
            pending = None
            for l in range(L):
                h = ops.rmsnorm(x, nw, 1e-6, residual=pending)
                act = ops.linear_decode(h, wgu[l], mode=2) if gu_ours else ops.silu_mul(F.linear(h, wgu[l]))
                if down_ours:
                    ops.linear_decode(act, wd[l], mode=1, out=x)
                    pending = None
                else:
                    pending = F.linear(act, wd[l])
        return fn

    for pdl in (0, 1):
        lib.cram_debug_pdl(pdl)
        for gu_ours in (False, True):
            for down_ours in (False, True):
                x.normal_()
                print(f"pdl {pdl}  gate/up {'ours' if gu_ours else 'lib '}  down {'ours' if down_ours else 'lib '}: "
                      f"{graph_time(chain(gu_ours, down_ours)):7.2f} us per layer (MLP half)", flush=True)
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == "splits":
    from cram_b200 import _lib
    import ctypes
    lib = _lib.load()
    lib.cram_debug_linear_split.argtypes = [ctypes.c_int]
    lib.cram_debug_linear_ring.argtypes = [ctypes.c_int]
    for sp, ring in ((2, 8), (4, 4), (4, 8), (8, 4)):
        lib.cram_debug_linear_split(sp)
        lib.cram_debug_linear_ring(ring)
        print("forced split", sp, "ring", ring, flush=True)
        case("qkv  4608x3584 +bias", QKV, H, 0, True)
        case("o    3584x3584 +residual", H, H, 1)
        case("down 3584x18944 +residual", H, I, 1)
    sys.exit(0)
from cram_b200 import _lib as _l
import ctypes as _c
_lb = _l.load()
_lb.cram_debug_pdl.argtypes = [_c.c_int]
for pdl in (0, 1):
    _lb.cram_debug_pdl(pdl)
    print("programmatic dependent launch", pdl, flush=True)
    case("qkv  4608x3584 +bias", QKV, H, 0, True)
    case("o    3584x3584 +residual", H, H, 1)
    case("g/u  2x18944x3584 swiglu", 2 * I, H, 2)
    case("down 3584x18944 +residual", H, I, 1)
sys.exit(0)
case("qkv  4608x3584 +bias", QKV, H, 0, True)
case("o    3584x3584 +residual", H, H, 1)
case("g/u  2x18944x3584 swiglu", 2 * I, H, 2)
case("down 3584x18944 +residual", H, I, 1)
