"""GPU: the HBM-bound glue kernels vs the HF eager formulas evaluated by torch on the same device."""
import pytest
import torch
import torch.nn.functional as F

from cram_b200 import ops
from oracle import cram_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_rmsnorm_and_residual(dtype):
    torch.manual_seed(0)
    x = torch.randn(37, 4096, device=DEV).to(dtype)
    r = torch.randn(37, 4096, device=DEV).to(dtype)
    w = (1 + 0.1 * torch.randn(4096, device=DEV)).to(dtype)
    want = orc.rmsnorm(x, w, 1e-5)
    got = ops.rmsnorm(x.clone(), w, 1e-5)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert (got.float() - want.float()).abs().max().item() <= tol
    if dtype != torch.float32:  # identical up to the rounding of one rsqrt ulp: the bulk must be bit-equal
        assert (got == want).float().mean().item() > 0.98
    x2 = x.clone()
    got2 = ops.rmsnorm(x2, w, 1e-5, residual=r)
    assert torch.equal(x2, x + r)  # the fused residual add is the exact model-dtype add
    assert (got2.float() - orc.rmsnorm(x + r, w, 1e-5).float()).abs().max().item() <= tol * 2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_rope_matches_hf(dtype):
    torch.manual_seed(1)
    H, Hkv, dh, N = 8, 2, 128, 300
    spec = orc.ModelSpec(1, H, Hkv, dh, 64, 64, 10, rope_theta=500000.0)
    qkv = torch.randn(N, (H + 2 * Hkv) * dh, device=DEV).to(dtype)
    pos = torch.randint(0, 4000, (N,), device=DEV, dtype=torch.int32)
    cos_t, sin_t = orc.rope_cos_sin(spec, torch.arange(4096)[None], dtype)
    cos_t, sin_t = cos_t[0].to(DEV).contiguous(), sin_t[0].to(DEV).contiguous()
    want = qkv.clone()
    for name, h0, n in (("q", 0, H), ("k", H, Hkv)):
        t = qkv[:, h0 * dh:(h0 + n) * dh].view(N, n, dh)
        c, s = cos_t[pos.long()][:, None], sin_t[pos.long()][:, None]
        want[:, h0 * dh:(h0 + n) * dh] = (t * c + orc.rotate_half(t) * s).reshape(N, n * dh)
    got = qkv.clone()
    ops.rope_(got, H + Hkv, dh, cos_t, sin_t, pos)
    assert torch.equal(got[:, (H + Hkv) * dh:], qkv[:, (H + Hkv) * dh:])  # V untouched
    assert torch.equal(got, want)  # same three roundings as HF -> bit-exact


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("page_size", [16, 5])
def test_fused_rope_kv_append_equals_the_two_kernels(dtype, page_size):
    torch.manual_seed(2)
    H, Hkv, dh, N, n_pages = 7, 1, 128, 131, 40
    qkv = torch.randn(N, (H + 2 * Hkv) * dh, device=DEV).to(dtype)
    pos = torch.randint(0, 500, (N,), device=DEV, dtype=torch.int32)
    spec = orc.ModelSpec(1, H, Hkv, dh, 64, 64, 10, rope_theta=10000.0)
    cos_t, sin_t = orc.rope_cos_sin(spec, torch.arange(512)[None], dtype)
    cos_t, sin_t = cos_t[0].to(DEV).contiguous(), sin_t[0].to(DEV).contiguous()
    slot = torch.randperm(n_pages * page_size, device=DEV)[:N].to(torch.int64)
    slot[3] = -1  # a skipped row: rotated in place, not stored
    shape = (n_pages, page_size, Hkv, dh)
    kc0, vc0 = torch.randn(shape, device=DEV).to(dtype), torch.randn(shape, device=DEV).to(dtype)
    want, kc_w, vc_w = qkv.clone(), kc0.clone(), vc0.clone()
    ops.rope_(want, H + Hkv, dh, cos_t, sin_t, pos)
    ops.kv_append(want[:, H * dh:(H + Hkv) * dh], want[:, (H + Hkv) * dh:], slot, kc_w, vc_w, Hkv, dh)
    got, kc_g, vc_g = qkv.clone(), kc0.clone(), vc0.clone()
    ops.rope_kv_append_(got, H, Hkv, dh, cos_t, sin_t, pos, slot, kc_g, vc_g)
    assert torch.equal(got, want) and torch.equal(kc_g, kc_w) and torch.equal(vc_g, vc_w)
    assert not torch.equal(kc_g, kc0)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_silu_mul(dtype):
    torch.manual_seed(2)
    gu = (torch.randn(65, 2 * 1024, device=DEV) * 3).to(dtype)
    want = F.silu(gu[:, :1024]) * gu[:, 1024:]
    got = ops.silu_mul(gu)
    if dtype == torch.float32:
        assert (got - want).abs().max().item() <= 1e-5
    else:
        assert (got == want).float().mean().item() > 0.995
        assert (got.float() - want.float()).abs().max().item() <= 0.07


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_mul_all_gate_bit_patterns(dtype):
    """Every one of the 65 536 gate values of the dtype (the kernel decides the rounding of silu from a fast
    estimate and falls back to the exact expression near ties): bit-equal to T(T(g / (1 + expf(-g))) * u)
    evaluated in fp32 by torch on the same device, for several u."""
    bits = torch.arange(65536, dtype=torch.int32, device=DEV).to(torch.int16)
    g = bits.view(dtype)
    finite = torch.isfinite(g.float())
    gf = g.float()
    s = (gf / (1.0 + torch.exp(-gf))).to(dtype)  # the expression of cram_silu_mul / HF eager, rounded once
    for uval in (1.0, -0.7421875, 3.0, 0.001953125):
        u = torch.full_like(g, uval)
        gu = torch.cat([g.view(64, 1024), u.view(64, 1024)], dim=1).contiguous()
        got = ops.silu_mul(gu).reshape(-1)
        want = (s.float() * u.float()).to(dtype)
        same = (got.view(torch.int16) == want.view(torch.int16)) | ((got == 0) & (want == 0))
        bad = finite & ~same & ~(torch.isnan(got.float()) & torch.isnan(want.float()))
        assert int(bad.sum()) == 0, (uval, g[bad][:8], got[bad][:8], want[bad][:8])


def test_splice_and_answer_ie():
    torch.manual_seed(3)
    T, H, dh, hidden, V, n = 50, 4, 64, 256, 1000, 3
    clean = torch.randn(T, H * dh, device=DEV).bfloat16()
    biased = torch.randn(T, H * dh, device=DEV).bfloat16()
    heads = torch.tensor([2, 0, 3], dtype=torch.int32, device=DEV)
    out = ops.sweep_splice_heads(clean, biased, heads, H, dh)
    for i, h in enumerate(heads.tolist()):
        want = clean.clone()
        want[:, h * dh:(h + 1) * dh] = biased[:, h * dh:(h + 1) * dh]
        assert torch.equal(out[i], want)
    for dtype in (torch.float32, torch.bfloat16):
        x = torch.randn(3, T, hidden, device=DEV).to(dtype)
        nw = (1 + 0.1 * torch.randn(hidden, device=DEV)).to(dtype)
        lm = (torch.randn(V, hidden, device=DEV) * 0.06).to(dtype)
        ans = torch.tensor([5, 999, 17], dtype=torch.int64, device=DEV)
        tail = x[:, T - n:].contiguous()
        cur0, _ = ops.sweep_answer_ie(tail[:1], nw, 1e-5, lm, ans, None)
        cur, ie = ops.sweep_answer_ie(tail, nw, 1e-5, lm, ans, cur0)
        logits = F.linear(orc.rmsnorm(x, nw, 1e-5), lm)  # [3, T, V]
        for v in range(3):
            want = orc.answer_logit_sum(logits[v:v + 1].cpu(), ans.cpu())
            tol = 1e-4 if dtype == torch.float32 else 0.05
            assert abs(cur[v].item() - float(want)) <= tol, (dtype, v, cur[v].item(), float(want))
        assert ie[0].item() == 0.0
        assert torch.allclose(ie, (cur0 - cur).to(dtype).float(), atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_mul_table_path_is_bit_identical_to_the_formula(dtype):
    """Large inputs of a 16-bit dtype go through the 65 536-entry table kernel (cram_silu_mul): every gate bit pattern,
    random up values, bit-equal to T(T(g / (1 + expf(-g))) * u) and to the small-input (formula) kernel."""
    rows, inter = 4096, 4096  # 2M 16-byte items: above the table path's threshold
    g = torch.Generator(device=DEV).manual_seed(5)
    gate_bits = torch.randint(0, 65536, (rows, inter), device=DEV, generator=g, dtype=torch.int32)
    gate_bits[:16].copy_(torch.arange(65536, device=DEV, dtype=torch.int32).view(16, 4096))  # every pattern at least once
    gate = gate_bits.to(torch.int16).view(dtype)
    up = (torch.randn(rows, inter, device=DEV, generator=g) * 2).to(dtype)
    gu = torch.cat([gate, up], dim=1).contiguous()
    got = ops.silu_mul(gu)
    gf = gate.float()
    want = ((gf / (1.0 + torch.exp(-gf))).to(dtype).float() * up.float()).to(dtype)
    finite = torch.isfinite(gf)
    same = (got.view(torch.int16) == want.view(torch.int16)) | ((got == 0) & (want == 0))
    bad = finite & ~same & ~(torch.isnan(got.float()) & torch.isnan(want.float()))
    assert int(bad.sum()) == 0, (gate[bad][:8], got[bad][:8], want[bad][:8])
    small = torch.cat([ops.silu_mul(gu[i:i + 64].contiguous()) for i in range(0, 256, 64)])  # formula kernel
    a, b = got[:256].view(torch.int16), small.view(torch.int16)
    assert bool(((a == b) | ~finite[:256]).all())
