"""Tiny-shape pass over the three kernels with hand-rolled mbarrier / TMEM / cluster protocols, for compute-sanitizer
(memcheck / racecheck / synccheck). One tool per gpurun call:

    compute-sanitizer --tool memcheck  python scripts/sanitize_kernels.py
    compute-sanitizer --tool racecheck python scripts/sanitize_kernels.py
    compute-sanitizer --tool synccheck python scripts/sanitize_kernels.py

Every launch is also compared with a torch fp64 reference, so a sanitizer-clean run is also a correct one.

compute-sanitizer is CLOSED on this pool (round 2: "runs under it have left GPUs needing a reset", see
profiles/r2_sanitizer_closed.txt), so the script carries its own bounds evidence: every output tensor is a view into the
middle of a larger allocation whose guard bands (1 MiB before and after, a fixed byte pattern) are verified after the
launch, and inputs sit flush against the END of their allocation's used part. A stray store shows up as a damaged guard.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from cram_b200 import ops  # noqa: E402
from tests.helpers import cu_of, make_logcred, make_qkv, ref_attention  # noqa: E402

DEV = "cuda"
GUARD = 1 << 20
PATTERN = 0xA5


def guarded(shape, dtype):
    """A tensor of `shape` carved out of the middle of a byte buffer with GUARD bytes of PATTERN on either side."""
    n = int(torch.tensor(shape).prod().item()) * torch.empty((), dtype=dtype).element_size()
    n_al = (n + 255) // 256 * 256
    raw = torch.full((GUARD + n_al + GUARD,), PATTERN, dtype=torch.uint8, device=DEV)
    view = raw[GUARD:GUARD + n].view(dtype).view(*shape)
    return raw, view, n


def check_guards(raw, n, what):
    torch.cuda.synchronize()
    lo, hi = raw[:GUARD], raw[GUARD + n:]
    assert bool((lo == PATTERN).all()) and bool((hi == PATTERN).all()), f"{what}: a guard band was written"


def prefill():
    """attn_prefill_sm100_kernel: ragged lengths incl. a 1-row sequence, GQA 4, bias on two heads, both head dims."""
    for lens, H, Hkv, dh in (([1, 129, 300], 4, 1, 128), ([260], 2, 2, 64)):
        q, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, torch.bfloat16, DEV, seed=3)
        lc = make_logcred(lens, DEV, torch.bfloat16, seed=4)
        cu = cu_of(lens, DEV)
        raw, out, n = guarded((sum(lens), H * dh), torch.bfloat16)
        got = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=0b101, kernel=2, out=out)
        check_guards(raw, n, "attn_prefill_sm100")
        want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=0b101)
        err = (got.double() - want).abs().max().item()
        assert err <= 2e-2, err
        print(f"attn_prefill_sm100 lens={lens} H={H}/{Hkv} dh={dh}: max|diff|={err:.2e}")


def linear():
    """linear_decode_sm100_kernel: a split-K cluster shape (DSMEM reduce-scatter), SwiGLU and residual epilogues."""
    g = torch.Generator(device=DEV).manual_seed(1)
    for m, n, k, mode in ((5, 256, 2048, ops.LINEAR_PLAIN), (64, 128, 4096, ops.LINEAR_RESIDUAL), (7, 128, 512, ops.LINEAR_SWIGLU)):
        x = torch.randn(m, k, device=DEV, generator=g).to(torch.bfloat16)
        rows = 2 * n if mode == ops.LINEAR_SWIGLU else n
        w = (torch.randn(rows, k, device=DEV, generator=g) * k ** -0.5).to(torch.bfloat16)
        if mode == ops.LINEAR_RESIDUAL:
            res = torch.randn(m, n, device=DEV, generator=g).to(torch.bfloat16)
            raw, got, nb = guarded((m, n), torch.bfloat16)
            got.copy_(res)
            ops.linear_decode(x, w, mode=mode, out=got)
            want = res.double() + (x.double() @ w.double().T)
        elif mode == ops.LINEAR_SWIGLU:
            raw, got, nb = guarded((m, n), torch.bfloat16)
            ops.linear_decode(x, w, mode=mode, out=got)
            gu = x.double() @ w.double().T
            want = torch.nn.functional.silu(gu[:, :n]) * gu[:, n:]
        else:
            raw, got, nb = guarded((m, n), torch.bfloat16)
            ops.linear_decode(x, w, out=got)
            want = x.double() @ w.double().T
        check_guards(raw, nb, "linear_decode")
        err = (got.double() - want).abs().max().item()
        assert err <= 6e-2, err
        print(f"linear_decode m={m} n={n} k={k} mode={mode}: max|diff|={err:.2e}")


def decode():
    """attn_decode_mma_kernel (+ combine): scattered pages, GQA 7 (Qwen2) and 4, split-KV on and off."""
    for H, Hkv, splits in ((28, 4, 0), (8, 2, 3)):
        dh, page, dtype = 128, 16, torch.bfloat16
        lens = [1, 65, 300]
        B = len(lens)
        _, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, dtype, DEV, seed=7, fused=False)
        q, _, _, _ = make_qkv(B, H, Hkv, dh, dtype, DEV, seed=8, fused=False)
        pages_per = (max(lens) + page - 1) // page
        perm = torch.randperm(B * pages_per, device=DEV).to(torch.int32).view(B, pages_per)
        kc = torch.zeros(B * pages_per, page, Hkv, dh, dtype=dtype, device=DEV)
        vc = torch.zeros_like(kc)
        pos = torch.cat([torch.arange(n, device=DEV) for n in lens])
        seq = torch.repeat_interleave(torch.arange(B, device=DEV), torch.tensor(lens, device=DEV))
        slot = perm[seq, pos // page].to(torch.int64) * page + pos % page
        ops.kv_append(k, v, slot, kc, vc, Hkv, dh)
        cu = cu_of(lens, DEV)
        lcp = make_logcred(lens, DEV, dtype, seed=9)
        lc = torch.zeros(B, max(lens), dtype=dtype, device=DEV)
        for b in range(B):
            lc[b, :lens[b]] = lcp[cu[b]:cu[b + 1]]
        raw, out, nb = guarded((B, H * dh), dtype)
        got = ops.attn_decode(q, kc, vc, perm, torch.tensor(lens, dtype=torch.int32, device=DEV), max(lens), H, Hkv, dh,
                              logcred=lc, head_mask=0b1011, n_splits=splits, out=out)
        check_guards(raw, nb, "attn_decode_mma")
        want = ref_attention(q, k, v, cu_of([1] * B, DEV), cu, H, Hkv, dh, logcred=lcp, head_mask=0b1011)
        err = (got.double() - want).abs().max().item()
        assert err <= 2e-2, err
        print(f"attn_decode_mma H={H}/{Hkv} splits={splits}: max|diff|={err:.2e}")


if __name__ == "__main__":
    which = sys.argv[1:] or ["prefill", "linear", "decode"]
    for name in which:
        globals()[name]()
    torch.cuda.synchronize()
    print("sanitize_kernels: ok")
