"""CPU ORACLE for the CrAM hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain torch-CPU / numpy restatement of what yeahjack/CrAM computes on its credibility-aware
attention path. Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may
import this module; nothing under ``cram_b200/`` does (the product path fails loudly when the CUDA
extension is missing -- there is no CPU fallback).

Parity status: PINNED. ``oracle/make_golden.py`` imports the unmodified reference from
``/root/reference/utils`` in the build container, runs it next to this restatement on the same seeded
inputs, asserts equality (bit-exact for indices / fp16 credibility / ranking, max-abs 0 for fp32
logits on the tiny model) and writes the fixtures under ``tests/golden/``. The ranking functions are
additionally pinned by the reference's own shipped artefacts (``datasets/*/*/heads_scores*.json``,
``selected_heads.json``), see ``tests/test_oracle_golden.py::test_ranking_kat``.

Every function cites the reference lines it follows (paths relative to /root/reference). The model
arithmetic itself lives in HF transformers (un-pinned by the reference; 5.5.0 installed here):
``models/llama/modeling_llama.py`` -- rmsnorm :52-70, rotary :73-135, rotate_half/apply :138-176,
eager_attention_forward :199-221, attention :224-292, decoder layer :295-340.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

# utils/prompts_templates.py:5-6 -- the only template the hot path uses (data contract: the span
# finder searches for "Passage-%d: " + para + "\n" inside exactly this string).
WITH_CONTEXTS = ('Given the following information: \n{paras}\nAnswer the following question based on the '
                 'given information or your internal knowledge with one or few words without the source.\n'
                 'Question: {question}\nAnswer: {answer}')


def get_prompt(context: Sequence[str], question: str, answer: str = "") -> str:
    """utils/prompt.py:4-12, type='with_contexts'."""
    paras = "\n".join(("Passage-%d: " % i) + p for i, p in enumerate(context))
    return WITH_CONTEXTS.format(question=question, paras=paras, answer=answer)


def build_prompt(tokenizer, model_name: str, question: str, paras: Sequence[str], answer: str = "") -> Tuple[str, bool]:
    """utils/re_weighting.py:64-97. Returns (prompt, add_special_tokens)."""
    add_special_tokens = True
    if model_name.find("Llama-3") != -1:
        prompt = get_prompt(paras, question, "")
        messages = [{"role": "system", "content": "You are a helpful assistant!"},
                    {"role": "user", "content": f"{prompt}"}]
        prompt = tokenizer.apply_chat_template(messages, add_generation_prompt=True, tokenize=False)
        if answer != "":
            prompt += answer
    elif model_name.find("Qwen") != -1:
        prompt = get_prompt(paras, question, "")
        messages = [{"role": "system", "content": "You are a helpful assistant."},
                    {"role": "user", "content": f"{prompt}"}]
        prompt = tokenizer.apply_chat_template(messages, add_generation_prompt=True, tokenize=False)
        add_special_tokens = False
        if answer != "":
            prompt += answer
    else:
        prompt = get_prompt(paras, question, answer)
    return prompt, add_special_tokens


def char_spans(prompt: str, paras: Sequence[str]) -> List[Tuple[int, int]]:
    """utils/re_weighting.py:100-103: first occurrence of 'Passage-i: ' + p + '\\n'; end is inclusive."""
    out = []
    for i, p in enumerate(paras):
        para = ("Passage-%d: " % i) + p + "\n"
        s = prompt.find(para)
        out.append((s, s + len(para) - 1))
    return out


def score_to_fp16(score) -> np.float16:
    """utils/re_weighting.py:111: torch.full((1, n), score).to(float16) -- int -> int64 tensor, float ->
    float32 tensor, then one rounding to fp16."""
    return torch.full((1, 1), score).to(torch.float16).numpy()[0, 0]


def span_index(offset_starts: np.ndarray, spans: Sequence[Tuple[int, int]], scores: Sequence) -> Tuple[np.ndarray, np.ndarray]:
    """utils/re_weighting.py:99-111. Returns (doc_id int32 [T] with -1 = no document, cred fp16 [T]).

    start_tok = max{idx : s >= off[idx, 0]}, end_tok likewise for e; no early exit, so tokens whose
    offset start is 0 (BOS and other post-processor tokens) always satisfy the test. Later documents
    overwrite earlier ones where spans overlap. A passage that is not found (s = -1) leaves
    start_tok = None in the reference and raises TypeError at :111.
    """
    off = np.asarray(offset_starts, dtype=np.int64)
    T = off.shape[0]
    doc_id = np.full(T, -1, dtype=np.int32)
    cred = np.ones(T, dtype=np.float16)
    for i, (s, e) in enumerate(spans):
        ok_s = np.nonzero(s >= off)[0]
        ok_e = np.nonzero(e >= off)[0]
        if ok_s.size == 0 or ok_e.size == 0:
            raise TypeError("unsupported operand type(s) for +: 'NoneType' and 'int'")
        a, b = int(ok_s.max()), int(ok_e.max())
        if b + 1 > a:
            doc_id[a:b + 1] = i
            cred[a:b + 1] = score_to_fp16(scores[i])
        elif b + 1 < a:
            # torch.full with a negative size raises in the reference
            raise RuntimeError("negative span")
    return doc_id, cred


def log_chain(cred_fp16: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    """utils/re_weighting.py:122 (torch.log on the fp16 vector) then :56 (.to(model dtype))."""
    w = torch.from_numpy(np.ascontiguousarray(cred_fp16)).to(torch.float16)
    return torch.log(w).to(dtype)


def edit_attention_mask(mask: Optional[torch.Tensor], logcred: torch.Tensor, head_idx: Sequence[int], H: int,
                        tgt_len: int, src_len: int, dtype: torch.dtype) -> torch.Tensor:
    """utils/re_weighting.py:36-61 (dense restatement; small shapes only).

    ``mask`` is HF's additive [b,1,q,k] mask (0 / finfo.min) or None; ``logcred`` is [1, T0] in any float
    dtype (cast to ``dtype`` here as :56 does). Returns the [b,H,q,k] mask.
    """
    weight_len = logcred.shape[-1]
    if mask is None:
        dev = logcred.device  # the reference builds it on hidden_states.device (:38); callers keep logcred there
        if tgt_len == 1:
            mask = torch.zeros([1, 1, tgt_len, src_len], dtype=dtype, device=dev)
        else:
            m = torch.triu(torch.full((tgt_len, src_len), torch.finfo(dtype).min, dtype=dtype, device=dev), diagonal=1)
            mask = m[None, None]
    else:
        mask = mask.clone()
    b, hd, q, k = mask.shape
    if hd == 1:
        mask = mask.repeat(1, H, 1, 1)
    expanded = logcred.reshape(1, 1, 1, weight_len).to(dtype).expand(b, H, q, weight_len)
    allowed = mask[..., :weight_len] == 0.0
    for h in head_idx:
        sub = mask[:, h, :, :weight_len]
        sub[allowed[:, h]] = expanded[:, h][allowed[:, h]]
    return mask


def eager_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, mask: torch.Tensor, scaling: float) -> torch.Tensor:
    """transformers models/llama/modeling_llama.py:199-221. q [b,H,q,dh]; k,v [b,Hkv,k,dh]; mask [b,H|1,q,k]."""
    n_rep = q.shape[1] // k.shape[1]
    k = k.repeat_interleave(n_rep, dim=1)
    v = v.repeat_interleave(n_rep, dim=1)
    w = torch.matmul(q, k.transpose(2, 3)) * scaling
    w = w + mask
    w = torch.softmax(w, dim=-1, dtype=torch.float32).to(q.dtype)
    return torch.matmul(w, v).transpose(1, 2).contiguous()


def cram_attention_closed_form(q, k, v, logcred, head_idx, scaling):
    """README.md:21 / SURVEY 8(a4): the hook is equivalent to re-weighting the post-softmax
    probabilities by cred[k] and renormalising. fp32/fp64 reference for the kernel tests."""
    b, H, T, dh = q.shape
    n_rep = H // k.shape[1]
    kk = k.repeat_interleave(n_rep, dim=1).double()
    vv = v.repeat_interleave(n_rep, dim=1).double()
    s = torch.matmul(q.double(), kk.transpose(2, 3)) * scaling
    causal = torch.triu(torch.ones(T, kk.shape[2], dtype=torch.bool), diagonal=1 + kk.shape[2] - T)
    s = s.masked_fill(causal, float("-inf"))
    p = torch.softmax(s, dim=-1)
    cred = torch.exp(logcred.double()).reshape(1, 1, 1, -1)
    out = torch.empty_like(p)
    for h in range(H):
        if h in head_idx:
            ph = p[:, h] * cred[:, 0]
            out[:, h] = ph / ph.sum(-1, keepdim=True)
        else:
            out[:, h] = p[:, h]
    return torch.matmul(out, vv).transpose(1, 2).contiguous(), out


# ----------------------------------------------------------------------------------------------
# Llama / Qwen2 forward restated on a plain state dict (HF eager arithmetic).
# ----------------------------------------------------------------------------------------------

class ModelSpec:
    """The handful of config fields the forward needs (read from a HF config or given directly)."""

    def __init__(self, num_layers, num_heads, num_kv_heads, head_dim, hidden, intermediate, vocab,
                 rms_eps=1e-6, rope_theta=10000.0, qkv_bias=False, tie_embeddings=False):
        self.num_layers, self.num_heads, self.num_kv_heads = num_layers, num_heads, num_kv_heads
        self.head_dim, self.hidden, self.intermediate, self.vocab = head_dim, hidden, intermediate, vocab
        self.rms_eps, self.rope_theta, self.qkv_bias, self.tie_embeddings = rms_eps, rope_theta, qkv_bias, tie_embeddings

    @classmethod
    def from_hf(cls, cfg):
        hd = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
        rp = getattr(cfg, "rope_parameters", None) or {}
        theta = rp.get("rope_theta", getattr(cfg, "rope_theta", 10000.0)) if isinstance(rp, dict) else getattr(cfg, "rope_theta", 10000.0)
        return cls(cfg.num_hidden_layers, cfg.num_attention_heads, cfg.num_key_value_heads, hd, cfg.hidden_size,
                   cfg.intermediate_size, cfg.vocab_size, cfg.rms_norm_eps, theta,
                   qkv_bias=(cfg.model_type == "qwen2") or bool(getattr(cfg, "attention_bias", False)),
                   tie_embeddings=bool(getattr(cfg, "tie_word_embeddings", False)))


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """modeling_llama.py:52-70."""
    dt = x.dtype
    xf = x.to(torch.float32)
    var = xf.pow(2).mean(-1, keepdim=True)
    xf = xf * torch.rsqrt(var + eps)
    return w * xf.to(dt)


def rope_cos_sin(spec: ModelSpec, position_ids: torch.Tensor, dtype: torch.dtype):
    """modeling_llama.py:73-135 (default rope, attention_scaling 1)."""
    inv_freq = 1.0 / (spec.rope_theta ** (torch.arange(0, spec.head_dim, 2, dtype=torch.int64).to(torch.float32) / spec.head_dim))
    freqs = position_ids.to(torch.float32)[:, :, None] * inv_freq[None, None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def rotate_half(x):
    x1 = x[..., : x.shape[-1] // 2]
    x2 = x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def llama_forward(sd: Dict[str, torch.Tensor], spec: ModelSpec, input_ids: torch.Tensor,
                  hooks: Optional[Dict[int, Tuple[torch.Tensor, Sequence[int]]]] = None,
                  return_attn_out: bool = False) -> torch.Tensor:
    """Full-sequence forward [1,T] -> logits [1,T,V] with the reference hook applied on ``hooks``
    = {layer: (logcred [1,T0], head_idx)} (utils/re_weighting.py:120-124 registers exactly that)."""
    hooks = hooks or {}
    dtype = sd["model.embed_tokens.weight"].dtype
    b, T = input_ids.shape
    x = torch.nn.functional.embedding(input_ids, sd["model.embed_tokens.weight"])
    pos = torch.arange(T)[None].expand(b, T)
    cos, sin = rope_cos_sin(spec, pos, dtype)
    cos, sin = cos[:, None], sin[:, None]
    base_mask = torch.triu(torch.full((T, T), torch.finfo(dtype).min, dtype=dtype), diagonal=1)[None, None].expand(b, 1, T, T)
    H, Hkv, dh = spec.num_heads, spec.num_kv_heads, spec.head_dim
    lin = torch.nn.functional.linear
    attn_outs = []
    for l in range(spec.num_layers):
        p = f"model.layers.{l}."
        h = rmsnorm(x, sd[p + "input_layernorm.weight"], spec.rms_eps)
        q = lin(h, sd[p + "self_attn.q_proj.weight"], sd.get(p + "self_attn.q_proj.bias")).view(b, T, H, dh).transpose(1, 2)
        k = lin(h, sd[p + "self_attn.k_proj.weight"], sd.get(p + "self_attn.k_proj.bias")).view(b, T, Hkv, dh).transpose(1, 2)
        v = lin(h, sd[p + "self_attn.v_proj.weight"], sd.get(p + "self_attn.v_proj.bias")).view(b, T, Hkv, dh).transpose(1, 2)
        q = q * cos + rotate_half(q) * sin
        k = k * cos + rotate_half(k) * sin
        mask = base_mask
        if l in hooks:
            logcred, heads = hooks[l]
            mask = edit_attention_mask(base_mask, logcred, heads, H, T, T, dtype)
        a = eager_attention(q, k, v, mask, dh ** -0.5).reshape(b, T, H * dh)
        if return_attn_out:
            attn_outs.append(a)
        x = x + lin(a, sd[p + "self_attn.o_proj.weight"])
        h = rmsnorm(x, sd[p + "post_attention_layernorm.weight"], spec.rms_eps)
        g = lin(h, sd[p + "mlp.gate_proj.weight"])
        u = lin(h, sd[p + "mlp.up_proj.weight"])
        x = x + lin(torch.nn.functional.silu(g) * u, sd[p + "mlp.down_proj.weight"])
    x = rmsnorm(x, sd["model.norm.weight"], spec.rms_eps)
    head = sd["model.embed_tokens.weight"] if spec.tie_embeddings else sd["lm_head.weight"]
    logits = lin(x, head)
    if return_attn_out:
        return logits, attn_outs
    return logits


def decode_step_cached(sd: Dict[str, torch.Tensor], spec: ModelSpec, token_ids: torch.Tensor, position: int,
                       k_cache: List[Optional[torch.Tensor]], v_cache: List[Optional[torch.Tensor]],
                       hooks: Optional[Dict[int, Tuple[torch.Tensor, Sequence[int]]]] = None) -> torch.Tensor:
    """One cached decode step [b,1] -> logits [b,V] the way model.generate runs it under the hook
    (utils/re_weighting.py:126-132): q_len = 1, K/V appended to the per-layer cache (k_cache[l] is
    [b,Hkv,k,dh] or None), and the hook's tgt_len == 1 branch (:39-48) turns the all-zero decode mask
    into [b,H,1,k] with logcred on the first T0 key columns of the hooked heads (:56-59)."""
    hooks = hooks or {}
    dtype = sd["model.embed_tokens.weight"].dtype
    b = token_ids.shape[0]
    x = torch.nn.functional.embedding(token_ids.view(b, 1), sd["model.embed_tokens.weight"])
    cos, sin = rope_cos_sin(spec, torch.full((b, 1), position), dtype)
    cos, sin = cos[:, None], sin[:, None]
    H, Hkv, dh = spec.num_heads, spec.num_kv_heads, spec.head_dim
    lin = torch.nn.functional.linear
    for l in range(spec.num_layers):
        p = f"model.layers.{l}."
        h = rmsnorm(x, sd[p + "input_layernorm.weight"], spec.rms_eps)
        q = lin(h, sd[p + "self_attn.q_proj.weight"], sd.get(p + "self_attn.q_proj.bias")).view(b, 1, H, dh).transpose(1, 2)
        k = lin(h, sd[p + "self_attn.k_proj.weight"], sd.get(p + "self_attn.k_proj.bias")).view(b, 1, Hkv, dh).transpose(1, 2)
        v = lin(h, sd[p + "self_attn.v_proj.weight"], sd.get(p + "self_attn.v_proj.bias")).view(b, 1, Hkv, dh).transpose(1, 2)
        q = q * cos + rotate_half(q) * sin
        k = k * cos + rotate_half(k) * sin
        k_cache[l] = k if k_cache[l] is None else torch.cat([k_cache[l], k], dim=2)
        v_cache[l] = v if v_cache[l] is None else torch.cat([v_cache[l], v], dim=2)
        src = k_cache[l].shape[2]
        if l in hooks:
            logcred, heads = hooks[l]
            mask = edit_attention_mask(None, logcred[..., :src], heads, H, 1, src, dtype)
        else:
            mask = torch.zeros(1, 1, 1, src, dtype=dtype)
        a = eager_attention(q, k_cache[l], v_cache[l], mask, dh ** -0.5).reshape(b, 1, H * dh)
        x = x + lin(a, sd[p + "self_attn.o_proj.weight"])
        h = rmsnorm(x, sd[p + "post_attention_layernorm.weight"], spec.rms_eps)
        x = x + lin(torch.nn.functional.silu(lin(h, sd[p + "mlp.gate_proj.weight"])) * lin(h, sd[p + "mlp.up_proj.weight"]),
                    sd[p + "mlp.down_proj.weight"])
    x = rmsnorm(x, sd["model.norm.weight"], spec.rms_eps)
    head = sd["model.embed_tokens.weight"] if spec.tie_embeddings else sd["lm_head.weight"]
    return lin(x, head)[:, 0]


def answer_logit_sum(logits: torch.Tensor, answer_ids: torch.Tensor) -> torch.Tensor:
    """utils/re_weighting.py:160: logits[0, -n:][arange(n), answer_ids].sum() in model dtype."""
    n = len(answer_ids)
    return logits[0, -n:][np.arange(n), answer_ids].sum()


def cal_logits(sd, spec: ModelSpec, input_ids: torch.Tensor, logcred: torch.Tensor, answer_ids: torch.Tensor) -> List[List[float]]:
    """utils/re_weighting.py:152-175: one clean forward then one full forward per (layer, head) with
    that single head hooked; IE = (ori - cur).item(), sum and difference in model dtype."""
    ori = answer_logit_sum(llama_forward(sd, spec, input_ids), answer_ids)
    out = []
    for l in range(spec.num_layers):
        row = []
        for h in range(spec.num_heads):
            cur = answer_logit_sum(llama_forward(sd, spec, input_ids, {l: (logcred, [h])}), answer_ids)
            row.append((ori - cur).item())
        out.append(row)
    return out


def greedy_generate(sd, spec: ModelSpec, input_ids: torch.Tensor, hooks, max_new_tokens: int,
                    eos_ids: Sequence[int] = (), repetition_penalty: float = 1.0, bad_ids: Sequence[int] = ()) -> List[int]:
    """utils/re_weighting.py:126-132 (do_sample=False). Re-runs the full prefix each step: with the
    hook's :57 rule (bias only on key columns < T0 where the mask is 0) a cached decode step and a
    full recompute give the same probabilities, so this is the simplest faithful restatement.

    ``repetition_penalty`` != 1 restates what model.generate still applies from the checkpoint's generation_config
    under do_sample=False (transformers generation/logits_process.py, RepetitionPenaltyLogitsProcessor: on the fp32 copy
    of the last-row logits, every token id already in the sequence -- prompt included -- gets logit*p if negative,
    logit/p otherwise); ``bad_ids`` are single-token bad_words_ids (NoBadWordsLogitsProcessor: -inf)."""
    ids = input_ids.clone()
    new = []
    for _ in range(max_new_tokens):
        T = ids.shape[1]
        hk = {}
        for l, (lc, heads) in (hooks or {}).items():
            T0 = lc.shape[-1]
            ext = torch.zeros(1, T, dtype=lc.dtype)
            ext[:, :T0] = lc
            hk[l] = (ext, heads)
        row = llama_forward(sd, spec, ids, hk)[0, -1].to(torch.float32)
        if repetition_penalty != 1.0:
            seen = ids[0].unique()
            sc = row[seen]
            row[seen] = torch.where(sc < 0, sc * repetition_penalty, sc / repetition_penalty)
        for b in bad_ids:
            row[b] = float("-inf")
        nxt = int(row.argmax())
        new.append(nxt)
        ids = torch.cat([ids, torch.tensor([[nxt]])], dim=1)
        if nxt in eos_ids:
            break
    return new


# ----------------------------------------------------------------------------------------------
# Head ranking (utils/find_best_heads.py:113-150)
# ----------------------------------------------------------------------------------------------

def combine_all(all_scores) -> List[Tuple[float, Tuple[int, int]]]:
    """utils/find_best_heads.py:113-127: float64 mean over samples, descending order.

    The reference uses np.argsort(...)[::-1] whose order inside groups of exactly equal means is
    numpy-build dependent. Tie rule defined here and in the product: value
    descending, then flat index ascending."""
    a = np.array(all_scores)
    mean = np.mean(a, axis=0)
    flat = mean.flatten()
    order = np.lexsort((np.arange(flat.size), -flat))
    idx = np.unravel_index(order, mean.shape)
    return [(float(flat[order[i]]), (int(idx[0][i]), int(idx[1][i]))) for i in range(flat.size)]


def top_k_heads(ranked, topk: int) -> Dict[int, List[int]]:
    """utils/find_best_heads.py:135-146: first k entries -> {layer: [heads...]} in first-seen order."""
    out: Dict[int, List[int]] = {}
    for _, (l, h) in ranked[:topk]:
        out.setdefault(l, []).append(h)
    return out
