"""GPU: decode attention over the paged cache == prefill attention with one query row per sequence."""
import pytest
import torch

from cram_b200 import ops
from tests.helpers import cu_of, make_logcred, make_qkv, ref_attention, report

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("dtype,H,Hkv,dh,page", [(torch.bfloat16, 28, 4, 128, 16), (torch.bfloat16, 32, 8, 128, 64),
                                                 (torch.float32, 4, 2, 64, 16), (torch.float16, 8, 8, 64, 32)])
@pytest.mark.parametrize("splits", [0, 1, 3])
def test_decode_matches_reference(dtype, H, Hkv, dh, page, splits):
    lens = [1, 17, 300, 1025, 64]
    prompt = [1, 10, 280, 1000, 64]  # bias covers only the prompt part of each sequence
    B = len(lens)
    total = sum(lens)
    _, k, v, _ = make_qkv(total, H, Hkv, dh, dtype, DEV, seed=41, fused=False)
    qfull, _, _, _ = make_qkv(B, H, Hkv, dh, dtype, DEV, seed=42, fused=False)
    pages_per = (max(lens) + page - 1) // page
    perm = torch.randperm(B * pages_per, device=DEV).to(torch.int32).view(B, pages_per)  # scattered pages
    kc = torch.zeros(B * pages_per, page, Hkv, dh, dtype=dtype, device=DEV)
    vc = torch.zeros_like(kc)
    cu = cu_of(lens, DEV)
    pos = torch.cat([torch.arange(n, device=DEV) for n in lens])
    seq = torch.repeat_interleave(torch.arange(B, device=DEV), torch.tensor(lens, device=DEV))
    slot = perm[seq, pos // page].to(torch.int64) * page + pos % page
    ops.kv_append(k, v, slot, kc, vc, Hkv, dh)
    lc_packed = make_logcred(lens, DEV, dtype if dtype != torch.float32 else torch.float32, seed=43)
    lc = torch.zeros(B, max(lens), dtype=lc_packed.dtype, device=DEV)
    for b in range(B):
        lc[b, :lens[b]] = lc_packed[cu[b]:cu[b + 1]]
    bl = torch.tensor(prompt, dtype=torch.int32, device=DEV)
    seqlens = torch.tensor(lens, dtype=torch.int32, device=DEV)
    mask = 0b1011 & ((1 << H) - 1)
    got = ops.attn_decode(qfull, kc, vc, perm, seqlens, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, bias_len=bl,
                          n_splits=splits)
    cuq = cu_of([1] * B, DEV)
    want = ref_attention(qfull, k, v, cuq, cu, H, Hkv, dh, logcred=lc_packed, head_mask=mask, bias_len=prompt)
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    assert not torch.isnan(got).any()
    assert report(f"decode {dtype} H={H}/{Hkv} splits={splits}", got, want) <= tol


def _paged(k, v, lens, page, Hkv, dh, dtype, seed=7):
    B = len(lens)
    pages_per = (max(lens) + page - 1) // page
    g = torch.Generator(device="cpu").manual_seed(seed)
    perm = torch.randperm(B * pages_per, generator=g).to(torch.int32).view(B, pages_per).to(DEV)
    kc = torch.zeros(B * pages_per, page, Hkv, dh, dtype=dtype, device=DEV)
    vc = torch.zeros_like(kc)
    pos = torch.cat([torch.arange(n, device=DEV) for n in lens])
    seq = torch.repeat_interleave(torch.arange(B, device=DEV), torch.tensor(lens, device=DEV))
    slot = perm[seq, pos // page].to(torch.int64) * page + pos % page
    ops.kv_append(k, v, slot, kc, vc, Hkv, dh)
    return kc, vc, perm


def test_decode_edge_cases_empty_batch_and_odd_page_size():
    H, Hkv, dh, dtype = 8, 2, 128, torch.bfloat16
    # empty batch: a no-op, not an error
    q0 = torch.empty(0, H * dh, dtype=dtype, device=DEV)
    kc0 = torch.zeros(1, 16, Hkv, dh, dtype=dtype, device=DEV)
    out = ops.attn_decode(q0, kc0, kc0, torch.zeros(0, 1, dtype=torch.int32, device=DEV),
                          torch.zeros(0, dtype=torch.int32, device=DEV), 0, H, Hkv, dh)
    assert out.shape == (0, H * dh)
    # page size that is not a power of two (division path), ragged lengths incl. 1 and a tile boundary
    lens = [1, 63, 64, 65, 500]
    total, B = sum(lens), len(lens)
    _, k, v, _ = make_qkv(total, H, Hkv, dh, dtype, DEV, seed=51, fused=False)
    q, _, _, _ = make_qkv(B, H, Hkv, dh, dtype, DEV, seed=52, fused=False)
    kc, vc, bt = _paged(k, v, lens, 24, Hkv, dh, dtype)
    cu = cu_of(lens, DEV)
    lcp = make_logcred(lens, DEV, dtype, seed=53)
    lc = torch.zeros(B, max(lens), dtype=dtype, device=DEV)
    for b in range(B):
        lc[b, :lens[b]] = lcp[cu[b]:cu[b + 1]]
    seqlens = torch.tensor(lens, dtype=torch.int32, device=DEV)
    for splits in (0, 2):
        got = ops.attn_decode(q, kc, vc, bt, seqlens, max(lens), H, Hkv, dh, logcred=lc, head_mask=0b10100101, n_splits=splits)
        want = ref_attention(q, k, v, cu_of([1] * B, DEV), cu, H, Hkv, dh, logcred=lcp, head_mask=0b10100101)
        assert not torch.isnan(got).any()
        assert report(f"decode page24 splits={splits}", got, want) <= 2e-2


def test_decode_full_size_equals_prefill_kernel_last_row():
    """BASELINE configs[3] layer shape (Qwen2-7B heads, ~2.2k cached keys, page 16): the paged decode kernel and the
    tcgen05 prefill kernel run as a 1-row chunk (cu_seqlens_k) are two independent implementations of the same row."""
    H, Hkv, dh, dtype, page = 28, 4, 128, torch.bfloat16, 16
    lens = [2048 + 128 + 3 * i for i in range(8)]
    prompt = [2048 + 3 * i for i in range(8)]
    total, B = sum(lens), len(lens)
    _, k, v, _ = make_qkv(total, H, Hkv, dh, dtype, DEV, seed=61, fused=False)
    q, _, _, _ = make_qkv(B, H, Hkv, dh, dtype, DEV, seed=62, fused=False)
    kc, vc, bt = _paged(k, v, lens, page, Hkv, dh, dtype)
    cu = cu_of(lens, DEV)
    lcp = make_logcred(lens, DEV, dtype, seed=63)
    lc = torch.zeros(B, max(lens), dtype=dtype, device=DEV)
    for b in range(B):
        lc[b, :lens[b]] = lcp[cu[b]:cu[b + 1]]
    bl = torch.tensor(prompt, dtype=torch.int32, device=DEV)
    mask = 0x0FFFFF5
    got = ops.attn_decode(q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32, device=DEV), max(lens), H, Hkv, dh, logcred=lc,
                          head_mask=mask, bias_len=bl)
    via_prefill = ops.attn_prefill(q, k, v, cu_of([1] * B, DEV), 1, H, Hkv, dh, logcred=lcp, head_mask=mask, cu_seqlens_k=cu,
                                   max_seqlen_k=max(lens), bias_len=bl, kernel=2)
    assert report("decode full-size vs tcgen05 prefill row", got, via_prefill) <= 2e-2
    # no bias selected == bias of zeros, bit for bit
    a = ops.attn_decode(q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32, device=DEV), max(lens), H, Hkv, dh)
    b = ops.attn_decode(q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32, device=DEV), max(lens), H, Hkv, dh,
                        logcred=torch.zeros_like(lc), head_mask=mask, bias_len=bl)
    assert torch.equal(a, b)
