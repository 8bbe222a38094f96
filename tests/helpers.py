"""Shared helpers of the GPU parity tests (test infrastructure)."""
import math

import numpy as np
import torch


def ref_attention(q, k, v, cu_q, cu_k, H, Hkv, dh, scale=None, logcred=None, head_mask=0, bias_len=None, causal=True):
    """fp64 statement of cram_attn_prefill's contract on packed rows (include/cram_b200.h), any device.
    The bias enters exactly as the reference's hook puts it into the additive mask
    (utils/re_weighting.py:56-59) and the softmax is HF eager's (modeling_llama.py:212-218), in fp64."""
    scale = scale if scale is not None else dh ** -0.5
    cu_q = cu_q.tolist()
    cu_k = cu_k.tolist() if cu_k is not None else cu_q
    out = torch.zeros(q.shape[0], H * dh, dtype=torch.float64, device=q.device)
    rep = H // Hkv
    for s in range(len(cu_q) - 1):
        q0, q1, k0, k1 = cu_q[s], cu_q[s + 1], cu_k[s], cu_k[s + 1]
        Tq, Tk = q1 - q0, k1 - k0
        if Tq == 0:
            continue
        qq = q[q0:q1].double().view(Tq, H, dh).transpose(0, 1)
        kk = k[k0:k1].double().view(Tk, Hkv, dh).transpose(0, 1).repeat_interleave(rep, 0)
        vv = v[k0:k1].double().view(Tk, Hkv, dh).transpose(0, 1).repeat_interleave(rep, 0)
        sc = torch.matmul(qq, kk.transpose(1, 2)) * scale
        if logcred is not None and head_mask:
            bl = Tk if bias_len is None else int(bias_len[s])
            b = logcred[k0:k1].double().clone()
            b[bl:] = 0
            sel = torch.tensor([(head_mask >> h) & 1 for h in range(H)], dtype=torch.bool, device=q.device)
            sc = sc + torch.where(sel.view(H, 1, 1), b.view(1, 1, Tk), torch.zeros((), dtype=torch.float64, device=q.device))
        if causal:
            i = torch.arange(Tq, device=q.device)[:, None]
            j = torch.arange(Tk, device=q.device)[None, :]
            sc = sc.masked_fill(j > i + (Tk - Tq), float("-inf"))
        m = sc.max(-1, keepdim=True).values
        m = torch.where(torch.isinf(m), torch.zeros_like(m), m)
        p = torch.exp(sc - m)
        den = p.sum(-1, keepdim=True)
        p = torch.where(den > 0, p / den, torch.zeros_like(p))
        out[q0:q1] = torch.matmul(p, vv).transpose(0, 1).reshape(Tq, H * dh)
    return out


def make_qkv(total, H, Hkv, dh, dtype, device, seed=0, fused=True, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    width = (H + 2 * Hkv) * dh
    buf = (torch.randn(total, width, generator=g) * scale).to(dtype).to(device)
    if fused:
        return buf[:, :H * dh], buf[:, H * dh:(H + Hkv) * dh], buf[:, (H + Hkv) * dh:], buf
    return buf[:, :H * dh].contiguous(), buf[:, H * dh:(H + Hkv) * dh].contiguous(), buf[:, (H + Hkv) * dh:].contiguous(), buf


def make_logcred(lens, device, dtype=torch.bfloat16, seed=0, with_zero_doc=True):
    """Per-key log-credibility shaped like real prompts: a few document spans with k/10 scores (one 0 -> -inf)."""
    rng = np.random.default_rng(seed)
    parts = []
    for T in lens:
        cred = np.ones(T, dtype=np.float16)
        pos = min(8, T)
        d = 0
        while pos < T - 4:
            n = int(rng.integers(max(1, T // 12), max(2, T // 5)))
            sc = rng.integers(0, 11) / 10.0
            if with_zero_doc and d == 0:
                sc = 0.0
            cred[pos:min(T - 4, pos + n)] = np.float16(np.float32(sc))
            pos += n + int(rng.integers(0, 3))
            d += 1
        parts.append(torch.log(torch.from_numpy(cred)).to(dtype))
    return torch.cat(parts).to(device)


def cu_of(lens, device):
    cu = torch.zeros(len(lens) + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(lens, dtype=torch.int32).cumsum(0)
    return cu.to(device)


def report(name, got, want):
    d = (got.double() - want.double()).abs()
    print(f"[{name}] max_abs={d.max().item():.3e} mean_abs={d.mean().item():.3e} ref_max={want.abs().max().item():.3e}")
    return d.max().item()
