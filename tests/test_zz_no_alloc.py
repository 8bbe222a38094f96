"""GPU: cram_attn_decode_static (the allocation-free entry point, include/cram_b200.h) == cram_attn_decode.

Collected LAST on purpose (the name sorts after every other test file): a failure
here must not hide the parity tests. Both entry points are called through the C-ABI (ctypes, cram_b200.ops). The static
variant keeps its split-KV partials in a 32 MiB device array inside the library, shared by every call and never cleared,
so the test makes two calls in a row with different shapes (the second one smaller: it runs over the first one's stale
partials) and checks each against the existing entry point and against the fp64 reference.
"""
import pytest
import torch

from cram_b200 import _lib, ops
from tests.helpers import cu_of, make_logcred, make_qkv, ref_attention, report

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-2  # the tolerance of the bf16 decode parity tests (tests/test_decode_gpu.py)


def _case(lens, H, Hkv, dh, page, dtype, seed):
    B = len(lens)
    _, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, dtype, DEV, seed=seed, fused=False)
    q, _, _, _ = make_qkv(B, H, Hkv, dh, dtype, DEV, seed=seed + 1, fused=False)
    pages_per = (max(lens) + page - 1) // page
    g = torch.Generator(device="cpu").manual_seed(seed + 2)
    bt = torch.randperm(B * pages_per, generator=g).to(torch.int32).view(B, pages_per).to(DEV)
    kc = torch.zeros(B * pages_per, page, Hkv, dh, dtype=dtype, device=DEV)
    vc = torch.zeros_like(kc)
    pos = torch.cat([torch.arange(n, device=DEV) for n in lens])
    seq = torch.repeat_interleave(torch.arange(B, device=DEV), torch.tensor(lens, device=DEV))
    slot = bt[seq, pos // page].to(torch.int64) * page + pos % page
    ops.kv_append(k, v, slot, kc, vc, Hkv, dh)
    cu = cu_of(lens, DEV)
    lcp = make_logcred(lens, DEV, dtype, seed=seed + 3)
    lc = torch.zeros(B, max(lens), dtype=dtype, device=DEV)
    for b in range(B):
        lc[b, :lens[b]] = lcp[cu[b]:cu[b + 1]]
    seqlens = torch.tensor(lens, dtype=torch.int32, device=DEV)
    want = ref_attention(q, k, v, cu_of([1] * B, DEV), cu, H, Hkv, dh, logcred=lcp, head_mask=0b1011)
    return dict(q=q, kc=kc, vc=vc, bt=bt, seqlens=seqlens, max_len=max(lens), H=H, Hkv=Hkv, dh=dh, lc=lc, want=want)


def _both(c, splits):
    args = (c["q"], c["kc"], c["vc"], c["bt"], c["seqlens"], c["max_len"], c["H"], c["Hkv"], c["dh"])
    old = ops.attn_decode(*args, logcred=c["lc"], head_mask=0b1011, n_splits=splits)
    out = torch.full_like(old, float("nan"))  # the caller's output buffer: every element must be written
    torch.cuda.synchronize()
    held, free = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    new = ops.attn_decode_static(*args, out, logcred=c["lc"], head_mask=0b1011, n_splits=splits)
    torch.cuda.synchronize()
    alloc_delta = (torch.cuda.memory_allocated() - held, free - torch.cuda.mem_get_info()[0])
    return old, new, alloc_delta


def test_static_scratch_decode_matches_decode_two_calls_in_a_row():
    dtype = torch.bfloat16
    # first call: 6 sequences x 28 heads x 5 splits of partials; second call: fewer, shorter, 3 splits and then the
    # automatic choice -- it reads and writes the same array, over what the first call left there
    first = _case([1, 65, 300, 1025, 777, 2048], 28, 4, 128, 16, dtype, seed=201)
    second = _case([130, 1, 513], 8, 2, 128, 32, dtype, seed=211)
    for name, c, splits in (("first", first, 5), ("second", second, 3), ("second, automatic splits", second, 0),
                            ("first again, one split", first, 1)):
        old, new, alloc_delta = _both(c, splits)
        assert not torch.isnan(new).any(), name
        assert report(f"static decode ({name}) vs cram_attn_decode", new, old.double()) <= TOL
        assert torch.equal(new, old), f"{name}: same kernels, same split count -> the same bits"
        assert report(f"static decode ({name}) vs fp64", new, c["want"]) <= TOL
        if name != "first":  # the first call of a process may load the module / create the context's kernels
            assert alloc_delta == (0, 0), f"{name}: device memory moved during an allocation-free call: {alloc_delta}"


def test_static_scratch_decode_rejects_launches_beyond_the_array():
    """128 sequences x 32 heads x 64 splits x 130 floats = 136 MB of partials > the 32 MiB array: a clear error, raised
    before anything is launched."""
    H, Hkv, dh, B = 32, 8, 128, 128
    q = torch.zeros(B, H * dh, dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros(B, 16, Hkv, dh, dtype=torch.bfloat16, device=DEV)
    bt = torch.arange(B, dtype=torch.int32, device=DEV).view(B, 1)
    seqlens = torch.full((B,), 16, dtype=torch.int32, device=DEV)
    out = torch.zeros_like(q)
    with pytest.raises(_lib.CramError, match="static array holds"):
        ops.attn_decode_static(q, kc, kc, bt, seqlens, 16, H, Hkv, dh, out, n_splits=64)
    assert (out == 0).all()
    # ... and the same launch with a split count that fits goes through
    ops.attn_decode_static(q, kc, kc, bt, seqlens, 16, H, Hkv, dh, out, n_splits=2)
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
