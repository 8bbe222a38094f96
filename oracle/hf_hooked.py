"""The reference's hooked HF path, restated so that it can run on the GPU box -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

/root/reference does not travel to the B200 box, so the reference classes cannot be imported there. This module
re-states the part of utils/re_weighting.py that drives a HF model -- forward_pre_hook registration (:120-124), the hook
itself (:36-61, through oracle.cram_oracle.edit_attention_mask, which oracle/make_golden.py pins against the unmodified
reference), the clean / per-head forwards and the answer-logit sum of cal_logits (:159-174) -- on a real
``transformers`` LlamaForCausalLM / Qwen2ForCausalLM (eager or sdpa attention), on whatever device the model lives on.

Used by tests/test_model_gate_gpu.py (the 8B-width bf16 gate of SURVEY 8d: our engine vs HF eager + hook on the same
GPU, same weights) and by bench.py's ``reference_on_this_gpu`` leg (what a CrAM user's own code achieves on this B200).
The HF model shares the engine's weight storage (no second copy): q/k/v and gate/up are row slices of the fused matrices.
"""
from __future__ import annotations

from functools import partial
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from oracle import cram_oracle as orc


def hf_from_engine(engine, attn: str = "eager"):
    """A HF causal LM (eager attention = the reference path) whose parameters ALIAS the engine's fused weights."""
    from transformers import LlamaConfig, LlamaForCausalLM, Qwen2Config, Qwen2ForCausalLM

    s, w = engine.spec, engine.w
    kw = dict(vocab_size=s.vocab, hidden_size=s.hidden, intermediate_size=s.intermediate, num_hidden_layers=s.num_layers,
              num_attention_heads=s.num_heads, num_key_value_heads=s.num_kv_heads, rms_norm_eps=s.rms_eps,
              rope_theta=s.rope_theta, max_position_embeddings=max(8192, engine._rope_len),
              tie_word_embeddings=False, bos_token_id=0, eos_token_id=1, pad_token_id=2)
    if s.qkv_bias:
        cfg, cls = Qwen2Config(**kw), Qwen2ForCausalLM
    else:
        cfg, cls = LlamaConfig(head_dim=s.head_dim, **kw), LlamaForCausalLM
    cfg._attn_implementation = attn
    with torch.device("meta"):
        model = cls(cfg)
    H, Hkv, dh, f = s.num_heads, s.num_kv_heads, s.head_dim, s.intermediate
    a, b = H * dh, (H + Hkv) * dh
    sd = {"model.embed_tokens.weight": w["embed"], "model.norm.weight": w["norm"], "lm_head.weight": w["lm_head"]}
    for l in range(s.num_layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = w[f"in_norm.{l}"]
        sd[p + "post_attention_layernorm.weight"] = w[f"post_norm.{l}"]
        qkv = w[f"wqkv.{l}"]
        sd[p + "self_attn.q_proj.weight"], sd[p + "self_attn.k_proj.weight"], sd[p + "self_attn.v_proj.weight"] = qkv[:a], qkv[a:b], qkv[b:]
        if f"bqkv.{l}" in w:
            bq = w[f"bqkv.{l}"]
            sd[p + "self_attn.q_proj.bias"], sd[p + "self_attn.k_proj.bias"], sd[p + "self_attn.v_proj.bias"] = bq[:a], bq[a:b], bq[b:]
        sd[p + "self_attn.o_proj.weight"] = w[f"wo.{l}"]
        sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"] = w[f"wgu.{l}"][:f], w[f"wgu.{l}"][f:]
        sd[p + "mlp.down_proj.weight"] = w[f"wdown.{l}"]
    missing, unexpected = model.load_state_dict(sd, strict=False, assign=True)
    assert not [m for m in missing if "rotary" not in m], missing
    assert not unexpected, unexpected
    # the rotary module's inv_freq is a non-persistent buffer: rebuild it on the real device
    rot = type(model.model.rotary_emb)(config=cfg, device=engine.device)
    model.model.rotary_emb = rot
    return model.eval()


def _hook(module, args, kwargs, logcred: torch.Tensor, head_idx: Sequence[int], H: int):
    """utils/re_weighting.py:36-61 as a forward_pre_hook(with_kwargs=True) on model.layers.{l}.self_attn."""
    hs = kwargs["hidden_states"]
    tgt = hs.size(1)
    mask = kwargs.get("attention_mask")
    src = mask.size(-1) if mask is not None else int(kwargs["position_ids"][0][-1]) + 1
    kwargs["attention_mask"] = orc.edit_attention_mask(mask, logcred.to(hs.device), head_idx, H, tgt, src, hs.dtype)
    return args, kwargs


class hooked:
    """Context manager: hooks on the chosen layers (:120-124), removed on exit (:136-137)."""

    def __init__(self, model, logcred: torch.Tensor, layers_to_heads: Dict[int, Sequence[int]]):
        self.model, self.logcred, self.l2h = model, logcred, layers_to_heads
        self.handles = []

    def __enter__(self):
        H = self.model.config.num_attention_heads
        for layer, heads in self.l2h.items():
            mod = self.model.get_submodule(f"model.layers.{int(layer)}.self_attn")
            self.handles.append(mod.register_forward_pre_hook(partial(_hook, logcred=self.logcred, head_idx=list(heads), H=H),
                                                              with_kwargs=True))
        return self

    def __exit__(self, *e):
        for h in self.handles:
            h.remove()


@torch.no_grad()
def hooked_logits(model, input_ids: torch.Tensor, logcred: Optional[torch.Tensor],
                  layers_to_heads: Optional[Dict[int, Sequence[int]]] = None) -> torch.Tensor:
    """logits [1, T, V] of one forward with the hooks in place (:159 / :169 call shape: input_ids + attention_mask)."""
    ids = input_ids.view(1, -1)
    am = torch.ones_like(ids)
    if logcred is None or not layers_to_heads:
        return model(input_ids=ids, attention_mask=am, return_dict=True)["logits"]
    with hooked(model, logcred.view(1, -1), layers_to_heads):
        return model(input_ids=ids, attention_mask=am, return_dict=True)["logits"]


@torch.no_grad()
def cal_logits_heads(model, input_ids: torch.Tensor, logcred: torch.Tensor, answer_ids: torch.Tensor,
                     heads: Sequence[Tuple[int, int]]) -> Tuple[float, List[float]]:
    """utils/re_weighting.py:159-174 for the listed (layer, head) pairs: one clean forward, then one FULL forward per
    pair with that single head hooked, answer-logit sum and difference in model dtype, one .item() per pair."""
    n = len(answer_ids)
    ori_logits = hooked_logits(model, input_ids, None).clone()
    ori = ori_logits[0, -n:][np.arange(n), answer_ids].sum()
    out = []
    for layer, head in heads:
        cur_logits = hooked_logits(model, input_ids, logcred, {layer: [head]}).clone()
        cur = cur_logits[0, -n:][np.arange(n), answer_ids].sum()
        out.append((ori - cur).item())
    return float(ori), out
