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
