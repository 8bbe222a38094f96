"""GPU: the Llama-3 and Qwen chat-template branches (utils/re_weighting.py:64-97,127-131,155) through the drop-in
classes, against fixtures the UNMODIFIED reference produced (oracle/make_golden_chat.py -> tests/golden/chat_model.*)."""
import json
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from cram_b200 import synth
from cram_b200.engine import CramEngine, ModelSpec
from cram_b200.re_weighting import Find_Best_Heads, Re_Weighting_Strategy
from oracle.tiny_model import chat_spec, chat_state_dict

pytestmark = pytest.mark.gpu
DEV = "cuda"
HEAD_SETS = {"none": {}, "l0h2": {0: [2]}, "mixed": {"1": [0, 3], "0": [1]}, "all": {0: [0, 1, 2, 3], 1: [0, 1, 2, 3]}}


@pytest.fixture(scope="module")
def gold(golden_dir):
    return (np.load(os.path.join(golden_dir, "chat_model.npz")),
            json.load(open(os.path.join(golden_dir, "chat_model_meta.json"))))


def _setup(m):
    t = chat_spec(m["kind"])
    spec = ModelSpec(t.num_layers, t.num_heads, t.num_kv_heads, t.head_dim, t.hidden, t.intermediate, t.vocab, t.rms_eps,
                     t.rope_theta, qkv_bias=t.qkv_bias)
    eng = CramEngine.from_state_dict(spec, chat_state_dict(m["kind"], m), device=DEV)
    tok = synth.load_tokenizer(kind=m["kind"])
    # what model.generation_config carries for this branch (the engine is built from a state dict, not a HF module)
    gc = SimpleNamespace(eos_token_id=m.get("eos_token_id", 1), repetition_penalty=m.get("repetition_penalty", 1.0))
    return eng, tok, gc


@pytest.mark.parametrize("name", ["llama3", "qwen"])
def test_prompt_ids_and_credibility_bit_exact(gold, name):
    g, meta = gold
    m = meta[name]
    eng, tok, gc = _setup(m)
    rw = Re_Weighting_Strategy(m["model_name"], tokenizer=tok, engine=eng, generation_config=gc)
    for ci, c in enumerate(m["cases"]):
        inputs, w = rw.decode_with_special_attention(question=c["question"], paras=c["paras"], scores=c["scores"],
                                                     answer=c["answer"])
        assert np.array_equal(inputs["input_ids"][0].cpu().numpy(), g[f"{name}_case{ci}_ids"]), (name, ci)
        assert np.array_equal(w[0].cpu().view(torch.int16).numpy().view(np.uint16), g[f"{name}_case{ci}_cred_bits"]), (name, ci)
    # batched front end: doc_id too
    r = rw.span_index_batch([c["question"] for c in m["cases"]], [c["paras"] for c in m["cases"]],
                            [c["scores"] for c in m["cases"]], [c["answer"] for c in m["cases"]])
    want = np.concatenate([g[f"{name}_case{ci}_doc_id"] for ci in range(len(m["cases"]))])
    assert np.array_equal(r["doc_id"].cpu().numpy(), want)
    assert int(r["err"].abs().sum()) == 0


@pytest.mark.parametrize("name", ["llama3", "qwen"])
def test_generation_text_and_tokens_equal_the_reference(gold, name):
    """run_RAG_with_attention_weighting: same (prompt, output) strings; tokens incl. the stop on <|eot_id|> (Llama-3) and
    the repetition penalty + two-entry eos list of generation_config (Qwen)."""
    g, meta = gold
    m = meta[name]
    eng, tok, gc = _setup(m)
    for hs_name, hs in HEAD_SETS.items():
        rw = Re_Weighting_Strategy(m["model_name"], layers_to_be_modified=dict(hs), tokenizer=tok, engine=eng,
                                   generation_config=gc)
        if not hs:
            rw.layers_to_be_modified = {}
        prompt, out = rw.run_RAG_batch([m["question"]], [m["paras"]], [m["scores"]], max_new_tokens=m["max_new"])[0]
        assert prompt == m[f"gen_prompt_{hs_name}"], (name, hs_name)
        assert out == m[f"gen_text_{hs_name}"], (name, hs_name, out, m[f"gen_text_{hs_name}"])
        # token level, through the engine with the class's own stop / penalty settings
        inputs, w = rw.decode_with_special_attention(question=m["question"], paras=m["paras"], scores=m["scores"])
        from cram_b200 import ops
        from cram_b200.re_weighting import _heads_dict_to_masks
        new = eng.generate([inputs["input_ids"][0]], [ops.log_cred(w[0], torch.float32)],
                           _heads_dict_to_masks(rw.layers_to_be_modified, 2), max_new_tokens=m["max_new"],
                           eos_token_ids=rw._eos_ids(), repetition_penalty=rw._repetition_penalty())[0]
        assert new == g[f"{name}_gen_{hs_name}"].tolist(), (name, hs_name, new)
    if name == "llama3":
        assert rw._eos_ids() == [tok.eos_token_id, tok.convert_tokens_to_ids("<|eot_id|>")]
        assert g["llama3_gen_none"][-1] == rw._eos_ids()[1] and len(g["llama3_gen_none"]) < m["max_new"]
    else:
        assert rw._eos_ids() == m["eos_token_id"] and rw._repetition_penalty() == m["repetition_penalty"]


@pytest.mark.parametrize("name", ["llama3", "qwen"])
def test_cal_logits_ie_matrix(gold, name):
    g, meta = gold
    m = meta[name]
    eng, tok, gc = _setup(m)
    fb = Find_Best_Heads(m["model_name"], tokenizer=tok, engine=eng, generation_config=gc)
    ie = np.array(fb.cal_logits(question=m["question"], paras=m["paras"], scores=m["sweep_scores"], wrong_answer=m["wrong"]))
    want = g[f"{name}_ie_ref"]
    print(f"[{name} IE]\n", ie, "\n", want)
    assert ie.shape == want.shape and np.abs(ie - want).max() <= 1e-3
    assert abs(float(fb.ori_prob_sum) - float(g[f"{name}_ie_ori_logit_sum"])) <= 1e-3
    assert np.argsort(-ie.flatten(), kind="stable").tolist() == np.argsort(-want.flatten(), kind="stable").tolist()


def test_max_new_tokens_one(gold):
    """generate(max_new_tokens=1) (reachable through run_RAG_batch): the token of the prefill logits, no decode session."""
    g, meta = gold
    m = meta["llama3"]
    eng, tok, gc = _setup(m)
    rw = Re_Weighting_Strategy(m["model_name"], layers_to_be_modified={0: [2]}, tokenizer=tok, engine=eng, generation_config=gc)
    inputs, w = rw.decode_with_special_attention(question=m["question"], paras=m["paras"], scores=m["scores"])
    from cram_b200 import ops
    new = eng.generate([inputs["input_ids"][0]], [ops.log_cred(w[0], torch.float32)], [0b100, 0], max_new_tokens=1)
    assert new == [[int(g["llama3_gen_l0h2"][0])]]
    assert len(rw.run_RAG_batch([m["question"]], [m["paras"]], [m["scores"]], max_new_tokens=1)) == 1
