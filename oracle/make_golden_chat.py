"""Generate tests/golden/chat_model.npz + chat_model_meta.json: the reference's CHAT-TEMPLATE branches.

Runs only in the build container (needs /root/reference, read-only). Test infrastructure. The UNMODIFIED reference
classes are driven with ``model_name`` containing "Llama-3" and "Qwen" (utils/re_weighting.py:64-97), on local tokenizers
that carry the two families' public chat templates (cram_b200/synth.py: load_tokenizer(kind="llama3" | "qwen")):

  * Llama-3 branch (:65-79): system prompt "You are a helpful assistant!", template-emitted BOS + add_special_tokens=True
    -> DOUBLE BOS, content trimmed by the template, answer appended after the generation prompt (:78-79), dual EOS
    [eos, <|eot_id|>] (:130-131). The tiny model's <|eot_id|> row is rigged (see ``rig_eot``) so that generation really
    stops on the SECOND eos id.
  * Qwen branch (:80-95): system prompt "You are a helpful assistant.", add_special_tokens=False, no BOS anywhere, so
    ``tokenizer([answer]).input_ids[0, 1:]`` (:155) drops the first REAL answer token. The model is a Qwen2 (QKV bias)
    whose generation_config carries a repetition_penalty != 1 and a two-entry eos list, as the Qwen chat checkpoints ship:
    model.generate(do_sample=False) (:132) still applies both.

For each branch: span index / fp16 credibility for a set of prompts, generated text through
run_RAG_with_attention_weighting for several head sets, and the IE matrix of cal_logits -- every one first asserted
equal to the restatement in oracle/cram_oracle.py.

    python oracle/make_golden_chat.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")

from cram_b200 import synth  # noqa: E402
from oracle import cram_oracle as orc  # noqa: E402
from oracle.make_golden import cuda_to_cpu, import_reference, make_ref_instance, span_cases  # noqa: E402
from oracle.tiny_model import chat_spec, chat_state_dict, hf_model_from  # noqa: E402

BRANCHES = {
    # name -> (model_name given to the reference, tokenizer kind)
    "llama3": ("Llama-3-tiny-Instruct", "llama3"),
    "qwen": ("Qwen-tiny-Chat", "qwen"),
}
HEAD_SETS = {"none": {}, "l0h2": {0: [2]}, "mixed": {"1": [0, 3], "0": [1]}, "all": {0: [0, 1, 2, 3], 1: [0, 1, 2, 3]}}
QWEN_REP_PENALTY = 1.3  # the Qwen chat checkpoints ship 1.05-1.1; larger here so that the tiny model's output visibly depends on it
MAX_NEW = 12  # the reference hard-codes max_new_tokens=100 (:126); patched down through generation kwargs below


def hooks_of(layers_to_be_modified, w):
    return {int(l): (torch.log(w), heads) for l, heads in layers_to_be_modified.items()}


def ref_generate_tokens(ref, rw, model, q, paras, scores, max_new):
    """The reference's own run_RAG_with_attention_weighting, with model.generate wrapped to cap max_new_tokens and to
    record the token ids it returns (the reference only returns text)."""
    seen = {}
    gen = model.generate

    def capped(**kw):
        kw["max_new_tokens"] = max_new
        out = gen(**kw)
        seen["tokens"] = out[0].tolist()
        seen["prompt_len"] = kw["input_ids"].shape[1]
        return out

    model.generate = capped
    try:
        prompt_txt, out_txt = ref.run_RAG_with_attention_weighting(question=q, paras=paras, scores=scores)
    finally:
        model.generate = gen
    return prompt_txt, out_txt, seen["tokens"][seen["prompt_len"]:]


def rig_eot(sd, spec, tokenizer, ref_ids, eot_id, step=5):
    """Make <|eot_id|> the greedy choice at generation step `step` of the un-hooked model: its lm_head row becomes
    1.25 x the row of the token the model emits there. Returns the source token id (recorded in the meta file so the
    GPU test rebuilds the same weights)."""
    toks = orc.greedy_generate(sd, spec, ref_ids, {}, step + 1)
    src = toks[step]
    assert src not in toks[:step], "pick another step: the token repeats earlier"
    sd["lm_head.weight"][eot_id] = sd["lm_head.weight"][src] * 1.25
    again = orc.greedy_generate(sd, spec, ref_ids, {}, step + 3, eos_ids=(tokenizer.eos_token_id, eot_id))
    assert again[-1] == eot_id and len(again) == step + 1, (again, eot_id)
    return int(src)


def main():
    torch.set_grad_enabled(False)
    rw = import_reference()
    out, meta = {}, {}
    for name, (model_name, kind) in BRANCHES.items():
        tok = synth.load_tokenizer(kind=kind)
        spec = chat_spec(kind)
        sd = chat_state_dict(kind)
        q, paras, scores, wrong = synth.nq_sample(31 if name == "llama3" else 32, 4, 22)
        scores = [0.0, 0.4, 0.8, 1.0]
        m = dict(model_name=model_name, kind=kind, question=q, paras=paras, scores=scores, wrong=wrong, max_new=MAX_NEW)
        with cuda_to_cpu():
            # ---- rig / configure the model -----------------------------------------------------------
            prompt, ast = orc.build_prompt(tok, model_name, q, paras, "")
            ids0 = torch.tensor(tok([prompt], add_special_tokens=ast)["input_ids"])
            if name == "llama3":
                eot = tok.convert_tokens_to_ids("<|eot_id|>")
                m["eot_src_token"] = rig_eot(sd, spec, tok, ids0, eot)
                m["eot_scale"] = 1.25
            model = hf_model_from(spec, sd)
            if name == "qwen":
                gc = model.generation_config
                gc.repetition_penalty = QWEN_REP_PENALTY
                gc.eos_token_id = [tok.convert_tokens_to_ids("<|im_end|>"), tok.convert_tokens_to_ids("<|endoftext|>")]
                gc.pad_token_id = tok.convert_tokens_to_ids("<|endoftext|>")
                gc.bos_token_id = None
                m["repetition_penalty"] = QWEN_REP_PENALTY
                m["eos_token_id"] = list(gc.eos_token_id)

            # ---- (a2/a3) prompts, ids, spans, credibility ---------------------------------------------
            ref = make_ref_instance(rw, rw.Re_Weighting_Strategy, model, tok, model_name)
            cases = []
            for ci, (cq, cparas, cscores) in enumerate(span_cases(tok, 36)):
                ans = "" if ci % 3 else wrong  # every third case carries an appended answer (:78-79 / :94-95)
                inputs, w_ref = ref.decode_with_special_attention(question=cq, paras=cparas, scores=cscores, answer=ans)
                prompt, ast = orc.build_prompt(tok, model_name, cq, cparas, ans)
                enc = tok([prompt], return_offsets_mapping=True, add_special_tokens=ast)
                assert np.array_equal(np.array(enc["input_ids"][0]), inputs["input_ids"][0].numpy())
                off = np.array(enc["offset_mapping"][0], dtype=np.int64)[:, 0]
                doc_id, cred = orc.span_index(off, orc.char_spans(prompt, cparas), cscores)
                assert np.array_equal(cred.view(np.uint16), w_ref[0].numpy().view(np.uint16)), (name, ci)
                cases.append(dict(question=cq, paras=cparas, scores=[float(s) if isinstance(s, float) else int(s) for s in cscores],
                                  answer=ans))
                out[f"{name}_case{ci}_ids"] = inputs["input_ids"][0].numpy()
                out[f"{name}_case{ci}_cred_bits"] = w_ref[0].numpy().view(np.uint16)
                out[f"{name}_case{ci}_doc_id"] = doc_id
            m["cases"] = cases
            if name == "llama3":
                assert out["llama3_case0_ids"][0] == out["llama3_case0_ids"][1] == tok.bos_token_id  # the double BOS
            else:
                assert out["qwen_case0_ids"][0] == tok.convert_tokens_to_ids("<|im_start|>")

            # ---- (a6) generation through the reference's public call ----------------------------------
            eos_ids = ([tok.eos_token_id, tok.convert_tokens_to_ids("<|eot_id|>")] if name == "llama3"
                       else list(model.generation_config.eos_token_id))
            rep = QWEN_REP_PENALTY if name == "qwen" else 1.0
            for hs_name, hs in HEAD_SETS.items():
                ref = make_ref_instance(rw, rw.Re_Weighting_Strategy, model, tok, model_name, hs)
                if not hs:
                    ref.layers_to_be_modified = {}
                prompt_txt, gen_txt, new_tokens = ref_generate_tokens(ref, rw, model, q, paras, scores, MAX_NEW)
                inputs, w = ref.decode_with_special_attention(question=q, paras=paras, scores=scores)
                want = orc.greedy_generate(sd, spec, inputs["input_ids"], hooks_of(ref.layers_to_be_modified, w), MAX_NEW,
                                           eos_ids=eos_ids, repetition_penalty=rep)
                # HF pads a finished single sequence no further: its tokens end with the eos it stopped on
                assert new_tokens == want, (name, hs_name, new_tokens, want)
                out[f"{name}_gen_{hs_name}"] = np.array(new_tokens, dtype=np.int64)
                m[f"gen_text_{hs_name}"] = gen_txt
                m[f"gen_prompt_{hs_name}"] = prompt_txt
            if name == "llama3":
                g = out["llama3_gen_none"]
                assert g[-1] == tok.convert_tokens_to_ids("<|eot_id|>") and len(g) < MAX_NEW, g  # stopped on the 2nd eos id
            if name == "qwen":
                plain = orc.greedy_generate(sd, spec, inputs["input_ids"], {}, MAX_NEW, eos_ids=eos_ids)
                assert plain != out["qwen_gen_none"].tolist(), "the fixture must depend on the repetition penalty"

            # ---- (a8) IE matrix through the reference's cal_logits -----------------------------------
            fb = make_ref_instance(rw, rw.Find_Best_Heads, model, tok, model_name)
            sw_scores = [0, 1, 1, 1]
            ie_ref = fb.cal_logits(question=q, paras=paras, scores=sw_scores, wrong_answer=wrong)
            ans_ids = tok([wrong], return_tensors="pt")["input_ids"][0, 1:]
            inputs2, w2 = fb.decode_with_special_attention(question=q, paras=paras, scores=sw_scores, answer=wrong)
            ie_orc = orc.cal_logits(sd, spec, inputs2["input_ids"], torch.log(w2), ans_ids)
            assert np.abs(np.array(ie_ref) - np.array(ie_orc)).max() <= 2e-4
            if name == "qwen":  # the quirk: no BOS, so [1:] removed a real token of the answer
                full = tok([wrong])["input_ids"][0]
                assert len(ans_ids) == len(full) - 1 and full[0] != tok.convert_tokens_to_ids("<|im_start|>")
                assert inputs2["input_ids"][0, -len(full):].tolist() == full
            out[f"{name}_ie_ref"] = np.array(ie_ref, dtype=np.float64)
            out[f"{name}_ie_input_ids"] = inputs2["input_ids"][0].numpy()
            out[f"{name}_ie_answer_ids"] = ans_ids.numpy()
            out[f"{name}_ie_cred_bits"] = w2[0].numpy().view(np.uint16)
            out[f"{name}_ie_ori_logit_sum"] = np.array(float(fb.ori_prob_sum))
            m["sweep_scores"] = sw_scores
        meta[name] = m
    with open(os.path.join(GOLD, "chat_model_meta.json"), "w") as f:
        json.dump(meta, f, ensure_ascii=False, indent=1)
    np.savez_compressed(os.path.join(GOLD, "chat_model.npz"), **out)
    print("chat golden written:", {k: len(v["cases"]) for k, v in meta.items()},
          {k: v.tolist() for k, v in out.items() if "_gen_" in k})


if __name__ == "__main__":
    main()
