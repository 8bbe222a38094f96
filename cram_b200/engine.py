"""Host-side forward of a Llama / Qwen2 decoder on packed variable-length rows.

This is the runtime that carries the credibility path on a B200: the dense projections are library
GEMMs (torch -> cuBLASLt), everything between them is libcram_b200.so -- fused RMSNorm(+residual),
rotary, SwiGLU gate, the tcgen05 credibility-biased attention, paged-KV append and the decode
attention. It replaces, for the hooked layers, HF's self_attn forward under the reference's
forward_pre_hook (utils/re_weighting.py:120-124): the hook's dense mask becomes a per-key
log-credibility vector plus a per-layer head bitmask.

Rows of all sequences are packed ([sum(T_i), hidden]); ``cu_seqlens`` delimits them.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from . import ops


@dataclass
class ModelSpec:
    num_layers: int
    num_heads: int
    num_kv_heads: int
    head_dim: int
    hidden: int
    intermediate: int
    vocab: int
    rms_eps: float = 1e-6
    rope_theta: float = 10000.0
    qkv_bias: bool = False
    tie_embeddings: bool = False

    @classmethod
    def from_hf(cls, cfg) -> "ModelSpec":
        hd = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
        rp = getattr(cfg, "rope_parameters", None)
        theta = rp.get("rope_theta") if isinstance(rp, dict) and rp.get("rope_theta") else getattr(cfg, "rope_theta", 10000.0)
        return cls(cfg.num_hidden_layers, cfg.num_attention_heads, cfg.num_key_value_heads, hd, cfg.hidden_size,
                   cfg.intermediate_size, cfg.vocab_size, cfg.rms_norm_eps, float(theta),
                   qkv_bias=(cfg.model_type == "qwen2") or bool(getattr(cfg, "attention_bias", False)),
                   tie_embeddings=bool(getattr(cfg, "tie_word_embeddings", False)))


LLAMA3_8B = ModelSpec(32, 32, 8, 128, 4096, 14336, 128256, 1e-5, 500000.0)
QWEN2_7B = ModelSpec(28, 28, 4, 128, 3584, 18944, 152064, 1e-6, 1000000.0, qkv_bias=True)
LLAMA3_70B = ModelSpec(80, 64, 8, 128, 8192, 28672, 128256, 1e-5, 500000.0)
TINY = ModelSpec(2, 4, 2, 64, 256, 512, 1000, 1e-5, 10000.0)


@dataclass
class SweepCapture:
    """Per-layer tensors of one clean forward that the IE sweep re-uses (kernel 3)."""
    x_in: List[torch.Tensor] = field(default_factory=list)   # residual stream entering the layer (after the fused add)
    qkv: List[torch.Tensor] = field(default_factory=list)    # fused projections, rotary applied
    attn: List[torch.Tensor] = field(default_factory=list)   # clean attention output


@dataclass
class SuffixTail:
    """The last n rows of each of n_seq packed sequences of length T (what the IE sweep reads, re_weighting.py:160)."""
    n: int
    rows: torch.Tensor        # int64 [n_seq * n] packed row indices
    positions: torch.Tensor   # int32 [n_seq * n] position of those rows inside their sequence
    cu_q: torch.Tensor        # int32 [n_seq + 1] = arange * n

    @classmethod
    def make(cls, n_seq: int, T: int, n: int, device) -> "SuffixTail":
        base = torch.arange(n_seq, device=device, dtype=torch.int64)[:, None] * T
        off = torch.arange(T - n, T, device=device, dtype=torch.int64)[None, :]
        return cls(n, (base + off).reshape(-1), off.expand(n_seq, n).reshape(-1).to(torch.int32).contiguous(),
                   torch.arange(n_seq + 1, device=device, dtype=torch.int32) * n)


class PagedKV:
    """Paged KV cache: per layer K and V of shape [n_pages, page_size, n_kv_heads, head_dim]."""

    def __init__(self, spec: ModelSpec, n_pages: int, page_size: int, dtype, device):
        self.page_size = page_size
        shape = (spec.num_layers, n_pages, page_size, spec.num_kv_heads, spec.head_dim)
        self.k = torch.zeros(shape, dtype=dtype, device=device)
        self.v = torch.zeros(shape, dtype=dtype, device=device)


class CramEngine:
    def __init__(self, spec: ModelSpec, weights: Dict[str, torch.Tensor], max_positions: int = 8192):
        self.spec = spec
        self.w = weights
        emb = weights["embed"]
        self.device, self.dtype = emb.device, emb.dtype
        if not emb.is_cuda:
            raise ops.CramError("CramEngine needs CUDA weights: cram_b200 has no CPU fallback")
        self._rope_len = 0
        self._ensure_rope(max_positions)
        # decode steps of <= 64 sequences run the MLP through cram_linear_decode (CRAM_B200_STREAMING_MLP=0: library GEMMs)
        self.streaming_mlp = os.environ.get("CRAM_B200_STREAMING_MLP", "1") != "0"
        # the two small projections (qkv, o_proj) as well; off by default: the library GEMM is faster at 25-33 MB
        # decode step: qkv / o_proj through cram_linear_decode instead of the library GEMM (off by default: measured slower
        # in the step, profiles/r2_decode_streaming_proj_ab.txt). "1" = both, "qkv" / "o" = one of them (A/B runs).
        _sp = os.environ.get("CRAM_B200_STREAMING_PROJ", "0")
        self.streaming_proj = _sp != "0"
        self.streaming_qkv = _sp in ("1", "qkv")
        self.streaming_o = _sp in ("1", "o")
        if os.environ.get("CRAM_B200_PDL"):  # A/B timing only: 0 switches programmatic dependent launch off
            ops._lib.load().cram_debug_pdl(int(os.environ["CRAM_B200_PDL"]))
        if os.environ.get("CRAM_B200_LINEAR_RING"):  # A/B timing only: force the 4- or 8-stage ring everywhere
            ops._lib.load().cram_debug_linear_ring(int(os.environ["CRAM_B200_LINEAR_RING"]))

    # ---- construction -------------------------------------------------------------------------------
    @staticmethod
    def fuse_state_dict(spec: ModelSpec, sd: Dict[str, torch.Tensor], device, dtype=None) -> Dict[str, torch.Tensor]:
        """HF Llama/Qwen2 state dict -> fused layout (wqkv, wo, wgu, wdown per layer)."""
        def g(name):
            t = sd[name]
            return t.to(device=device, dtype=dtype or t.dtype)
        w = {"embed": g("model.embed_tokens.weight"), "norm": g("model.norm.weight")}
        w["lm_head"] = w["embed"] if spec.tie_embeddings or "lm_head.weight" not in sd else g("lm_head.weight")
        for l in range(spec.num_layers):
            p = f"model.layers.{l}."
            w[f"in_norm.{l}"] = g(p + "input_layernorm.weight")
            w[f"post_norm.{l}"] = g(p + "post_attention_layernorm.weight")
            w[f"wqkv.{l}"] = torch.cat([g(p + "self_attn.q_proj.weight"), g(p + "self_attn.k_proj.weight"),
                                        g(p + "self_attn.v_proj.weight")], 0).contiguous()
            if spec.qkv_bias and (p + "self_attn.q_proj.bias") in sd:
                w[f"bqkv.{l}"] = torch.cat([g(p + "self_attn.q_proj.bias"), g(p + "self_attn.k_proj.bias"),
                                            g(p + "self_attn.v_proj.bias")], 0).contiguous()
            w[f"wo.{l}"] = g(p + "self_attn.o_proj.weight").contiguous()
            w[f"wgu.{l}"] = torch.cat([g(p + "mlp.gate_proj.weight"), g(p + "mlp.up_proj.weight")], 0).contiguous()
            w[f"wdown.{l}"] = g(p + "mlp.down_proj.weight").contiguous()
        return w

    @classmethod
    def from_state_dict(cls, spec: ModelSpec, sd, device="cuda", dtype=None, **kw) -> "CramEngine":
        return cls(spec, cls.fuse_state_dict(spec, sd, device, dtype), **kw)

    @classmethod
    def from_hf_model(cls, model, device="cuda", **kw) -> "CramEngine":
        spec = ModelSpec.from_hf(model.config)
        return cls.from_state_dict(spec, model.state_dict(), device=device, **kw)

    @classmethod
    def random_init(cls, spec: ModelSpec, seed: int = 0, dtype=torch.bfloat16, device="cuda", std: float = 0.02,
                    **kw) -> "CramEngine":
        """Random weights of the given architecture, allocated directly on the device (HF default init N(0, 0.02))."""
        gen = torch.Generator(device=device)
        gen.manual_seed(seed)

        def n(*shape):
            return torch.empty(*shape, dtype=dtype, device=device).normal_(0.0, std, generator=gen)

        H, Hkv, dh, d, f = spec.num_heads, spec.num_kv_heads, spec.head_dim, spec.hidden, spec.intermediate
        w = {"embed": n(spec.vocab, d), "norm": torch.ones(d, dtype=dtype, device=device)}
        w["lm_head"] = w["embed"] if spec.tie_embeddings else n(spec.vocab, d)
        for l in range(spec.num_layers):
            w[f"in_norm.{l}"] = torch.ones(d, dtype=dtype, device=device)
            w[f"post_norm.{l}"] = torch.ones(d, dtype=dtype, device=device)
            w[f"wqkv.{l}"] = n((H + 2 * Hkv) * dh, d)
            if spec.qkv_bias:
                w[f"bqkv.{l}"] = n((H + 2 * Hkv) * dh)
            w[f"wo.{l}"] = n(d, H * dh)
            w[f"wgu.{l}"] = n(2 * f, d)
            w[f"wdown.{l}"] = n(d, f)
        return cls(spec, w, **kw)

    def _ensure_rope(self, n_pos: int) -> None:
        """cos/sin tables exactly as HF builds them (modeling_llama.py:73-135): fp32 angles, cast to model dtype."""
        if n_pos <= self._rope_len:
            return
        s = self.spec
        inv_freq = 1.0 / (s.rope_theta ** (torch.arange(0, s.head_dim, 2, dtype=torch.int64).to(torch.float32) / s.head_dim))
        pos = torch.arange(n_pos, dtype=torch.float32)
        freqs = pos[:, None] * inv_freq[None, :]
        emb = torch.cat((freqs, freqs), dim=-1)
        self.cos = emb.cos().to(self.dtype).to(self.device).contiguous()
        self.sin = emb.sin().to(self.dtype).to(self.device).contiguous()
        self._rope_len = n_pos

    # ---- layer pieces -------------------------------------------------------------------------------
    def _qkv(self, l: int, x: torch.Tensor, pending: Optional[torch.Tensor], positions: torch.Tensor,
             kv: Optional["PagedKV"] = None, slot_mapping: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Input norm (+ pending residual), fused QKV projection, rotary embedding; with `kv`, the rotated K and V
        rows also go to their cache slots in the same launch as the rotation."""
        s = self.spec
        h = ops.rmsnorm(x, self.w[f"in_norm.{l}"], s.rms_eps, residual=pending)  # x += pending (in place)
        if self.streaming_qkv and kv is not None and h.shape[0] <= 64 and h.dtype in (torch.bfloat16, torch.float16):
            qkv = ops.linear_decode(h, self.w[f"wqkv.{l}"], self.w.get(f"bqkv.{l}"))
        else:
            qkv = F.linear(h, self.w[f"wqkv.{l}"], self.w.get(f"bqkv.{l}"))
        if kv is None:
            ops.rope_(qkv, s.num_heads + s.num_kv_heads, s.head_dim, self.cos, self.sin, positions)
        else:
            ops.rope_kv_append_(qkv, s.num_heads, s.num_kv_heads, s.head_dim, self.cos, self.sin, positions, slot_mapping,
                                kv.k[l], kv.v[l])
        return qkv

    def split_qkv(self, qkv: torch.Tensor):
        s = self.spec
        a, b = s.num_heads * s.head_dim, (s.num_heads + s.num_kv_heads) * s.head_dim
        return qkv[:, :a], qkv[:, a:b], qkv[:, b:]

    def layer_post(self, l: int, x: torch.Tensor, attn: torch.Tensor, streaming_mlp: bool = False) -> Optional[torch.Tensor]:
        """o_proj + residual + MLP of layer l. x is updated in place; returns the pending MLP delta, or None when
        the delta is already in x (``streaming_mlp``: the <= 64-row decode step, where gate/up + SwiGLU and
        down + residual add each run as one weight-streaming launch of cram_linear_decode)."""
        s = self.spec
        if streaming_mlp and self.streaming_o:
            ops.linear_decode(attn, self.w[f"wo.{l}"], mode=ops.LINEAR_RESIDUAL, out=x)
            h = ops.rmsnorm(x, self.w[f"post_norm.{l}"], s.rms_eps)
        else:
            o = F.linear(attn, self.w[f"wo.{l}"])
            h = ops.rmsnorm(x, self.w[f"post_norm.{l}"], s.rms_eps, residual=o)
        if streaming_mlp:
            act = ops.linear_decode(h, self.w[f"wgu.{l}"], mode=ops.LINEAR_SWIGLU)
            ops.linear_decode(act, self.w[f"wdown.{l}"], mode=ops.LINEAR_RESIDUAL, out=x)
            return None
        act = ops.silu_mul(F.linear(h, self.w[f"wgu.{l}"]))
        return F.linear(act, self.w[f"wdown.{l}"])

    # ---- prefill ------------------------------------------------------------------------------------
    def prefill(self, input_ids: torch.Tensor, cu_seqlens: torch.Tensor, max_seqlen: int,
                logcred: Optional[torch.Tensor] = None, head_masks: Optional[Sequence[int]] = None,
                capture: Optional[SweepCapture] = None, kv: Optional[PagedKV] = None,
                slot_mapping: Optional[torch.Tensor] = None, positions: Optional[torch.Tensor] = None,
                attn_kernel: int = 0) -> torch.Tensor:
        """Full forward of packed sequences. Returns the final residual stream [N, hidden] (before the
        final norm). ``logcred`` is the packed per-key log-credibility [N]; ``head_masks[l]`` the
        64-bit head bitmask of layer l (0 = untouched layer)."""
        s = self.spec
        if positions is None:
            positions = packed_positions(cu_seqlens, input_ids.numel())
        self._ensure_rope(max_seqlen)
        x = F.embedding(input_ids, self.w["embed"])
        pending = None
        for l in range(s.num_layers):
            qkv = self._qkv(l, x, pending, positions, kv, slot_mapping)
            q, k, v = self.split_qkv(qkv)
            mask = int(head_masks[l]) if (head_masks is not None and logcred is not None) else 0
            attn = ops.attn_prefill(q, k, v, cu_seqlens, max_seqlen, s.num_heads, s.num_kv_heads, s.head_dim,
                                    logcred=logcred, head_mask=mask, kernel=attn_kernel)
            if capture is not None:
                capture.x_in.append(x.clone())
                capture.qkv.append(qkv)
                capture.attn.append(attn)
            pending = self.layer_post(l, x, attn)
        x.add_(pending)
        return x

    def suffix_from_attn(self, l: int, x: torch.Tensor, attn: torch.Tensor, cu_seqlens: torch.Tensor, max_seqlen: int,
                         positions: torch.Tensor, attn_kernel: int = 0, tail: Optional["SuffixTail"] = None) -> torch.Tensor:
        """Continue a forward from layer l's attention output (sweep: variants differ from here on).
        x is the residual stream entering layer l's o_proj add; it is updated in place and returned.

        With ``tail`` (the sweep only reads the last n rows of every sequence) the LAST layer does only the work those
        rows need -- exact, no approximation: K and V of every row, but Q, the attention, o_proj, the MLP and the
        residual stream for the tail rows alone. Returns the final residual stream of the tail rows [n_seq * n, hidden]."""
        s = self.spec
        last = s.num_layers - 1
        if tail is not None and l == last:
            xt = x.index_select(0, tail.rows)
            xt.add_(self.layer_post(l, xt, attn.index_select(0, tail.rows)))
            return xt
        pending = self.layer_post(l, x, attn)
        for j in range(l + 1, s.num_layers):
            if tail is not None and j == last:
                return self._last_layer_tail(j, x, pending, cu_seqlens, max_seqlen, positions, tail, attn_kernel)
            qkv = self._qkv(j, x, pending, positions)
            q, k, v = self.split_qkv(qkv)
            a = ops.attn_prefill(q, k, v, cu_seqlens, max_seqlen, s.num_heads, s.num_kv_heads, s.head_dim,
                                 kernel=attn_kernel)
            pending = self.layer_post(j, x, a)
        x.add_(pending)
        return x

    def _last_layer_tail(self, j: int, x: torch.Tensor, pending: torch.Tensor, cu_seqlens: torch.Tensor, max_seqlen: int,
                         positions: torch.Tensor, tail: "SuffixTail", attn_kernel: int) -> torch.Tensor:
        """Layer j for the tail rows only: K / V projections of every row (the tail queries attend to all of them), the Q
        projection, attention (len_q = n < len_k = T, the kernel's chunked-prefill mode), o_proj and MLP of n rows."""
        s = self.spec
        qd = s.num_heads * s.head_dim
        h = ops.rmsnorm(x, self.w[f"in_norm.{j}"], s.rms_eps, residual=pending)  # x += pending (in place)
        w, b = self.w[f"wqkv.{j}"], self.w.get(f"bqkv.{j}")
        kv = F.linear(h, w[qd:], None if b is None else b[qd:])
        ops.rope_(kv, s.num_kv_heads, s.head_dim, self.cos, self.sin, positions)
        q = F.linear(h.index_select(0, tail.rows), w[:qd], None if b is None else b[:qd])
        ops.rope_(q, s.num_heads, s.head_dim, self.cos, self.sin, tail.positions)
        kd = s.num_kv_heads * s.head_dim
        a = ops.attn_prefill(q, kv[:, :kd], kv[:, kd:], tail.cu_q, tail.n, s.num_heads, s.num_kv_heads, s.head_dim,
                             cu_seqlens_k=cu_seqlens, max_seqlen_k=max_seqlen, kernel=attn_kernel)
        xt = x.index_select(0, tail.rows)
        xt.add_(self.layer_post(j, xt, a))
        return xt

    def logits(self, x_rows: torch.Tensor) -> torch.Tensor:
        """Final norm + lm_head on selected rows of the final residual stream."""
        h = ops.rmsnorm(x_rows.contiguous(), self.w["norm"], self.spec.rms_eps)
        return F.linear(h, self.w["lm_head"])

    # ---- decode -------------------------------------------------------------------------------------
    def decode_step(self, tokens: torch.Tensor, positions: torch.Tensor, kv: PagedKV, block_table: torch.Tensor,
                    seqlens: torch.Tensor, max_seqlen: int, logcred: Optional[torch.Tensor] = None,
                    bias_len: Optional[torch.Tensor] = None, head_masks: Optional[Sequence[int]] = None,
                    workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One token per sequence. positions = index of the new token (int32); seqlens = positions + 1.
        Returns logits [B, vocab]."""
        s = self.spec
        self._ensure_rope(max_seqlen + 1)
        ps = kv.page_size
        pos64 = positions.to(torch.int64)
        slot = block_table.gather(1, (pos64 // ps)[:, None]).squeeze(1).to(torch.int64) * ps + pos64 % ps
        x = F.embedding(tokens, self.w["embed"])
        pending = None
        streaming = (self.streaming_mlp and x.shape[0] <= 64 and x.dtype in (torch.bfloat16, torch.float16)
                     and s.hidden % 8 == 0 and s.intermediate % 8 == 0)  # what cram_linear_decode takes
        for l in range(s.num_layers):
            qkv = self._qkv(l, x, pending, positions, kv, slot)
            q, k, v = self.split_qkv(qkv)
            mask = int(head_masks[l]) if (head_masks is not None and logcred is not None) else 0
            attn = ops.attn_decode(q, kv.k[l], kv.v[l], block_table, seqlens, max_seqlen, s.num_heads, s.num_kv_heads,
                                   s.head_dim, logcred=logcred, head_mask=mask, bias_len=bias_len, workspace=workspace)
            pending = self.layer_post(l, x, attn, streaming_mlp=streaming)
        if pending is not None:
            x.add_(pending)
        return self.logits(x)

    @torch.no_grad()
    def prefill_into_cache(self, prompts: Sequence[torch.Tensor], kv: Optional[PagedKV], block_table: Optional[torch.Tensor],
                           logcreds: Optional[Sequence[torch.Tensor]] = None,
                           head_masks: Optional[Sequence[int]] = None, chunk: int = 16) -> torch.Tensor:
        """Re-weighted prefill of a batch of prompts (each a 1-D int64 CUDA tensor) into the paged cache,
        `chunk` sequences per pass (bounds the activation footprint). Returns last-row logits [B, vocab].
        ``kv`` None: no decode step follows (max_new_tokens <= 1), nothing is cached."""
        dev, ps = self.device, (kv.page_size if kv is not None else 1)
        out = []
        for b0 in range(0, len(prompts), chunk):
            part = list(prompts[b0:b0 + chunk])
            lens = [int(p.numel()) for p in part]
            cu = torch.zeros(len(part) + 1, dtype=torch.int32)
            cu[1:] = torch.tensor(lens, dtype=torch.int32).cumsum(0)
            cu_d = cu.to(dev)
            ids = torch.cat(part).to(dev)
            positions = packed_positions(cu_d, ids.numel())
            slot = None
            if kv is not None:
                seq_of_row = torch.repeat_interleave(torch.arange(b0, b0 + len(part), device=dev), torch.tensor(lens, device=dev))
                pos64 = positions.to(torch.int64)
                slot = block_table[seq_of_row, pos64 // ps].to(torch.int64) * ps + pos64 % ps
            lc = torch.cat(list(logcreds[b0:b0 + chunk])).to(dev) if logcreds is not None else None
            x = self.prefill(ids, cu_d, max(lens), logcred=lc, head_masks=head_masks, kv=kv, slot_mapping=slot,
                             positions=positions)
            out.append(self.logits(x[(cu_d[1:] - 1).to(torch.int64)]))
        return torch.cat(out)

    @torch.no_grad()
    def generate(self, prompts: Sequence[torch.Tensor], logcreds: Optional[Sequence[torch.Tensor]] = None,
                 head_masks: Optional[Sequence[int]] = None, max_new_tokens: int = 100,
                 eos_token_ids: Sequence[int] = (), bad_token_ids: Sequence[int] = (), page_size: int = 16,
                 use_graph: bool = True, eos_check_every: int = 8, repetition_penalty: float = 1.0) -> List[List[int]]:
        """Greedy generation for a batch of prompts (each a 1-D int64 CUDA tensor) with per-sample
        credibility. Mirrors model.generate(do_sample=False, max_new_tokens=...) of
        utils/re_weighting.py:126-132: stop at EOS (the EOS token is kept), single-token bad words are
        never emitted; ``repetition_penalty`` != 1 is what generate() still applies from the checkpoint's
        generation_config under do_sample=False (RepetitionPenaltyLogitsProcessor: on the fp32 copy of the logits, every
        token already in the sequence, prompt included, gets logit*p if negative else logit/p). Returns the new tokens of
        every sequence.

        The decode step is captured once in a CUDA graph (DecodeSession) and replayed; generated tokens
        stay on the device and are read back every `eos_check_every` steps to test for EOS (tokens a
        sequence produces after its EOS are discarded here, as HF pads them away)."""
        s, dev = self.spec, self.device
        B = len(prompts)
        lens = [int(p.numel()) for p in prompts]
        max_total = max(lens) + max_new_tokens
        pages_per_seq = (max_total + page_size - 1) // page_size
        kv = block_table = None
        if max_new_tokens > 1:  # a single new token comes straight from the prefill's last-row logits: no cache, no decode
            kv = PagedKV(s, B * pages_per_seq, page_size, self.dtype, dev)
            block_table = torch.arange(B * pages_per_seq, dtype=torch.int32, device=dev).view(B, pages_per_seq)
        logits = self.prefill_into_cache(prompts, kv, block_table, logcreds, head_masks)
        # decode-time bias: [B, T0max] padded with zeros, keys < bias_len[b] biased
        lc_dec = bias_len = None
        if logcreds is not None:
            lc_dec = torch.zeros(B, max(lens), dtype=logcreds[0].dtype, device=dev)
            for b, lc in enumerate(logcreds):
                lc_dec[b, :lens[b]] = lc
            bias_len = torch.tensor(lens, dtype=torch.int32, device=dev)
        ban = torch.tensor(list(bad_token_ids), dtype=torch.int64, device=dev) if bad_token_ids else None
        seen = None
        if repetition_penalty != 1.0:
            seen = torch.zeros(B, s.vocab, dtype=torch.bool, device=dev)
            for b, p in enumerate(prompts):
                seen[b, p.to(dev)] = True
        first = greedy_pick(logits, ban, seen, repetition_penalty)
        eos = set(int(e) for e in eos_token_ids)
        if max_new_tokens <= 1:  # nothing to decode: no session, no graph (its capture writes token row 1)
            return [[int(t)] for t in first.tolist()] if max_new_tokens == 1 else [[] for _ in range(B)]
        sess = DecodeSession(self, kv, block_table, lens, max_new_tokens, logcred=lc_dec, bias_len=bias_len,
                             head_masks=head_masks, bad_token_ids=bad_token_ids, use_graph=use_graph, seen=seen,
                             repetition_penalty=repetition_penalty)
        sess.start(first)
        out: List[List[int]] = [[] for _ in range(B)]
        alive = [True] * B
        done_rows = 0

        def drain(upto: int) -> None:
            nonlocal done_rows
            rows = sess.out_tokens[done_rows:upto].tolist()
            for row in rows:
                for b in range(B):
                    if alive[b]:
                        out[b].append(row[b])
                        if row[b] in eos:
                            alive[b] = False
            done_rows = upto

        produced = 1  # out_tokens[0] = the token chosen from the prefill logits
        while True:
            if produced % eos_check_every == 0 or produced == max_new_tokens or not eos:
                if eos or produced == max_new_tokens:
                    drain(produced)
            if produced == max_new_tokens or not any(alive):
                break
            sess.step()
            produced += 1
        drain(produced)
        return out


class DecodeSession:
    """Greedy decode of a fixed batch over a paged KV cache with the whole per-token step -- embedding,
    every layer (library GEMMs + libcram_b200 glue + the credibility-biased decode attention), final
    norm, lm_head, bad-word ban, argmax and the position/length bookkeeping -- captured in ONE CUDA
    graph and replayed per token: ~12 launches per layer collapse into one graph launch."""

    def __init__(self, engine: CramEngine, kv: PagedKV, block_table: torch.Tensor, prompt_lens: Sequence[int],
                 max_new_tokens: int, logcred: Optional[torch.Tensor] = None, bias_len: Optional[torch.Tensor] = None,
                 head_masks: Optional[Sequence[int]] = None, bad_token_ids: Sequence[int] = (), use_graph: bool = True,
                 seen: Optional[torch.Tensor] = None, repetition_penalty: float = 1.0):
        dev = engine.device
        self.seen, self.repetition_penalty = seen, float(repetition_penalty)  # [B, vocab] bool: ids already in the sequence
        self.engine, self.kv, self.block_table = engine, kv, block_table
        self.logcred, self.bias_len, self.head_masks = logcred, bias_len, head_masks
        self.B = len(prompt_lens)
        self.max_total = int(max(prompt_lens)) + int(max_new_tokens)
        engine._ensure_rope(self.max_total + 1)
        self.prompt_lens = torch.tensor(list(prompt_lens), dtype=torch.int32, device=dev)
        self.tokens = torch.zeros(self.B, dtype=torch.int64, device=dev)
        self.positions = torch.zeros(self.B, dtype=torch.int32, device=dev)   # index of the token being fed
        self.seqlens = torch.zeros(self.B, dtype=torch.int32, device=dev)     # keys visible to it
        self.out_tokens = torch.zeros(max_new_tokens, self.B, dtype=torch.int64, device=dev)
        self.step_idx = torch.zeros(1, dtype=torch.int64, device=dev)
        self.ban = torch.tensor(list(bad_token_ids), dtype=torch.int64, device=dev) if bad_token_ids else None
        self.use_graph = use_graph
        self.graph: Optional[torch.cuda.CUDAGraph] = None

    def _body(self) -> None:
        logits = self.engine.decode_step(self.tokens, self.positions, self.kv, self.block_table, self.seqlens,
                                         self.max_total, logcred=self.logcred, bias_len=self.bias_len,
                                         head_masks=self.head_masks)
        nxt = greedy_pick(logits, self.ban, self.seen, self.repetition_penalty)
        if self.seen is not None:
            self.seen.scatter_(1, nxt[:, None], True)
        self.out_tokens.index_copy_(0, self.step_idx, nxt[None])
        self.tokens.copy_(nxt)
        self.positions.add_(1)
        self.seqlens.add_(1)
        self.step_idx.add_(1)

    def start(self, first_tokens: torch.Tensor) -> None:
        """first_tokens: the token chosen from the prefill logits (new token 0 of every sequence)."""
        self.tokens.copy_(first_tokens)
        if self.seen is not None:
            self.seen.scatter_(1, first_tokens[:, None], True)
        self.positions.copy_(self.prompt_lens)
        self.seqlens.copy_(self.prompt_lens + 1)
        self.out_tokens[0].copy_(first_tokens)
        self.step_idx.fill_(1)
        if self.use_graph and self.graph is None:
            live = [self.tokens, self.positions, self.seqlens, self.out_tokens, self.step_idx]
            if self.seen is not None:
                live.append(self.seen)
            state = [t.clone() for t in live]
            side = torch.cuda.Stream(device=self.engine.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._body()  # warm-up: cuBLAS workspaces, kernel attributes (writes only cache slots >= prompt length)
            torch.cuda.current_stream().wait_stream(side)
            for t, c in zip(live, state):
                t.copy_(c)
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._body()
            for t, c in zip(live, state):
                t.copy_(c)  # capture does not execute, but keep the state explicit

    def step(self) -> None:
        if self.graph is not None:
            self.graph.replay()
        else:
            self._body()


def greedy_pick(logits: torch.Tensor, ban: Optional[torch.Tensor], seen: Optional[torch.Tensor],
                repetition_penalty: float) -> torch.Tensor:
    """argmax of the processed last-row logits [B, vocab], as model.generate(do_sample=False) forms it: fp32 copy,
    repetition penalty on the ids in `seen`, -inf on single-token bad words."""
    if seen is not None and repetition_penalty != 1.0:
        lf = logits.float()
        logits = torch.where(seen, torch.where(lf < 0, lf * repetition_penalty, lf / repetition_penalty), lf)
    if ban is not None:
        logits[:, ban] = float("-inf")
    nxt = logits.argmax(-1)
    return nxt


def packed_positions(cu_seqlens: torch.Tensor, n_rows: int) -> torch.Tensor:
    """int32 position of every packed row inside its sequence."""
    dev = cu_seqlens.device
    lens = (cu_seqlens[1:] - cu_seqlens[:-1]).to(torch.int64)
    starts = torch.repeat_interleave(cu_seqlens[:-1].to(torch.int64), lens, output_size=n_rows)
    return (torch.arange(n_rows, device=dev, dtype=torch.int64) - starts).to(torch.int32)
