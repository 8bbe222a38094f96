"""The tiny random-init Llama of BASELINE.json configs[0] (2 layers, 4 heads, d=256) -- TEST INFRASTRUCTURE.

Weights come from a seeded numpy PCG64 stream (not torch's default init) so the GPU box, which has no
access to the fixtures' generating run, rebuilds bit-identical fp32 weights. 2 KV heads make the
4 query heads a GQA group pair, which exercises per-q-head bias divergence inside a group
(hard part 4). Standard deviations are larger than HF's 0.02 so that attention is
not uniform and the credibility bias visibly moves logits.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch

from oracle.cram_oracle import ModelSpec

TINY_SPEC = ModelSpec(num_layers=2, num_heads=4, num_kv_heads=2, head_dim=64, hidden=256, intermediate=512,
                      vocab=1000, rms_eps=1e-5, rope_theta=10000.0)


def random_state_dict(spec: ModelSpec, seed: int = 0, dtype=torch.float32, w_std: float = 0.06,
                      emb_std: float = 0.5) -> Dict[str, torch.Tensor]:
    g = np.random.Generator(np.random.PCG64(seed))

    def n(*shape, std):
        return torch.from_numpy((g.standard_normal(shape) * std).astype(np.float32)).to(dtype)

    H, Hkv, dh, d, f = spec.num_heads, spec.num_kv_heads, spec.head_dim, spec.hidden, spec.intermediate
    sd = {"model.embed_tokens.weight": n(spec.vocab, d, std=emb_std)}
    for l in range(spec.num_layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = 1.0 + n(d, std=0.1)
        sd[p + "self_attn.q_proj.weight"] = n(H * dh, d, std=w_std)
        sd[p + "self_attn.k_proj.weight"] = n(Hkv * dh, d, std=w_std)
        sd[p + "self_attn.v_proj.weight"] = n(Hkv * dh, d, std=w_std)
        if spec.qkv_bias:
            sd[p + "self_attn.q_proj.bias"] = n(H * dh, std=0.1)
            sd[p + "self_attn.k_proj.bias"] = n(Hkv * dh, std=0.1)
            sd[p + "self_attn.v_proj.bias"] = n(Hkv * dh, std=0.1)
        sd[p + "self_attn.o_proj.weight"] = n(d, H * dh, std=w_std)
        sd[p + "post_attention_layernorm.weight"] = 1.0 + n(d, std=0.1)
        sd[p + "mlp.gate_proj.weight"] = n(f, d, std=w_std)
        sd[p + "mlp.up_proj.weight"] = n(f, d, std=w_std)
        sd[p + "mlp.down_proj.weight"] = n(d, f, std=w_std)
    sd["model.norm.weight"] = 1.0 + n(d, std=0.1)
    sd["lm_head.weight"] = n(spec.vocab, d, std=w_std)
    return sd


def tiny_state_dict(dtype=torch.float32) -> Dict[str, torch.Tensor]:
    return random_state_dict(TINY_SPEC, seed=0, dtype=dtype)


CHAT_VOCAB = 1008  # cram_b200.synth.CHAT_VOCAB: 1000 BPE ids + the chat special tokens


def chat_spec(kind: str) -> ModelSpec:
    """The tiny model of the chat-template fixtures: TINY_SPEC with room for the special tokens; "qwen" is a Qwen2
    (QKV bias)."""
    return ModelSpec(num_layers=2, num_heads=4, num_kv_heads=2, head_dim=64, hidden=256, intermediate=512,
                     vocab=CHAT_VOCAB, rms_eps=1e-5 if kind == "llama3" else 1e-6, rope_theta=10000.0,
                     qkv_bias=(kind == "qwen"))


def chat_state_dict(kind: str, meta: dict = None, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Seeded weights of the chat fixtures. With ``meta`` (tests/golden/chat_model_meta.json[kind]) the <|eot_id|>
    rigging recorded by oracle/make_golden_chat.py:rig_eot is re-applied, in fp32 before the cast."""
    sd = random_state_dict(chat_spec(kind), seed={"llama3": 21, "qwen": 22}[kind])
    if meta is not None and "eot_src_token" in meta:
        from cram_b200 import synth
        eot = 1000 + synth.LLAMA3_SPECIALS.index("<|eot_id|>")
        sd["lm_head.weight"][eot] = sd["lm_head.weight"][meta["eot_src_token"]] * meta["eot_scale"]
    return {k: v.to(dtype) for k, v in sd.items()}


def hf_model_from(spec: ModelSpec, sd: Dict[str, torch.Tensor], attn: str = "eager"):
    """A HF LlamaForCausalLM / Qwen2ForCausalLM carrying ``sd`` (eager attention = the reference path)."""
    from transformers import LlamaConfig, LlamaForCausalLM, Qwen2Config, Qwen2ForCausalLM

    kw = dict(vocab_size=spec.vocab, hidden_size=spec.hidden, intermediate_size=spec.intermediate,
              num_hidden_layers=spec.num_layers, num_attention_heads=spec.num_heads,
              num_key_value_heads=spec.num_kv_heads, rms_norm_eps=spec.rms_eps, rope_theta=spec.rope_theta,
              max_position_embeddings=8192, tie_word_embeddings=False, bos_token_id=0, eos_token_id=1, pad_token_id=2)
    if spec.qkv_bias:
        cfg, cls = Qwen2Config(**kw), Qwen2ForCausalLM
    else:
        cfg, cls = LlamaConfig(head_dim=spec.head_dim, **kw), LlamaForCausalLM
    cfg._attn_implementation = attn
    dtype = sd["model.embed_tokens.weight"].dtype
    model = cls(cfg).to(dtype).eval()
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not [m for m in missing if "rotary" not in m], missing
    assert not unexpected, unexpected
    return model


def tiny_hf_model():
    return hf_model_from(TINY_SPEC, tiny_state_dict())
