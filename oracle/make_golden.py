"""Generate tests/golden/* by running the UNMODIFIED reference next to the oracle restatement.

Runs only in the build container (needs /root/reference, read-only). Test infrastructure. The
reference is imported with the four shims of SURVEY.md section 8(c): utils/ on sys.path, instances
built with object.__new__, the literal "cuda" redirected to "cpu", and the local BPE tokenizer.
Every fixture written here is first checked against the restatement in oracle/cram_oracle.py
(assertions below) -- that is what pins the oracle.

    python oracle/make_golden.py            # rewrites tests/golden/*.npz / *.json
"""
from __future__ import annotations

import json
import os
import random
import sys
from functools import partial
from unittest import mock

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")

from cram_b200 import synth  # noqa: E402
from oracle import cram_oracle as orc  # noqa: E402
from oracle.tiny_model import TINY_SPEC, tiny_hf_model, tiny_state_dict  # noqa: E402


def import_reference():
    sys.path.insert(0, os.path.join(REF, "utils"))
    import re_weighting  # noqa: F401  (the reference, unmodified)
    return re_weighting


class cuda_to_cpu:
    """Redirect the literal "cuda" (utils/re_weighting.py:98,111) to "cpu" on the GPU-less box."""

    def __enter__(self):
        from transformers import BatchEncoding
        t_to, b_to = torch.Tensor.to, BatchEncoding.to

        def fix(a):
            return tuple("cpu" if (isinstance(x, str) and x == "cuda") else x for x in a)

        self.p = [mock.patch.object(torch.Tensor, "to", lambda s, *a, **k: t_to(s, *fix(a), **k)),
                  mock.patch.object(BatchEncoding, "to", lambda s, *a, **k: b_to(s, *fix(a), **k))]
        for p in self.p:
            p.start()

    def __exit__(self, *e):
        for p in self.p:
            p.stop()


def make_ref_instance(rw, cls, model, tokenizer, model_name, layers_to_be_modified=None):
    obj = object.__new__(cls)
    obj.tokenizer, obj.model, obj.model_name, obj.bad_words_ids = tokenizer, model, model_name, []
    obj.num_hidden_layers = model.config.num_hidden_layers
    obj.model_num_hidden_layers = model.config.num_hidden_layers
    obj.model_num_attention_heads = model.config.num_attention_heads
    obj.hidden_size = model.config.hidden_size
    if not layers_to_be_modified:
        layers_to_be_modified = {i: list(range(obj.model_num_attention_heads)) for i in range(obj.num_hidden_layers)}
    obj.layers_to_be_modified = layers_to_be_modified
    return obj


def golden_log_table():
    """(i) all 65 536 fp16 bit patterns through torch.log (fp16) -> bf16 / fp16 / fp32."""
    bits = np.arange(65536, dtype=np.uint16)
    w = torch.from_numpy(bits.view(np.float16).copy())
    lg = torch.log(w)  # utils/re_weighting.py:122
    out16 = lg.view(torch.int16).numpy().view(np.uint16)
    outbf = lg.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)  # :56
    np.savez_compressed(os.path.join(GOLD, "log_chain.npz"), log_fp16_bits=out16, log_bf16_bits=outbf)
    # the survey's probed anchors
    def one(x):
        i = int(np.array([x], dtype=np.float16).view(np.uint16)[0])
        return float(lg[i]), float(lg.to(torch.bfloat16)[i])
    assert one(0.25) == (-1.38671875, -1.390625), one(0.25)
    assert one(0.5) == (-0.693359375, -0.6953125)
    assert one(0.1) == (-2.302734375, -2.296875)
    assert one(1.0) == (0.0, 0.0) and one(0.0)[0] == float("-inf")


def span_cases(tokenizer, n_cases=220):
    rng = random.Random(7)
    for c in range(n_cases):
        D = rng.choice([1, 2, 3, 4, 5, 5, 8, 12, 20])
        q, paras, scores, _ = synth.nq_sample(1000 + c, D, rng.choice([3, 8, 20, 40]))
        kind = c % 6
        if kind == 1 and D >= 2:      # duplicate paragraph text (resolves by "Passage-i: " prefix)
            paras[-1] = paras[0]
        if kind == 2:                 # integer scores (int64 -> fp16 path of torch.full)
            scores = [rng.randint(0, 10) for _ in range(D)]
        if kind == 3:                 # sweep pattern
            scores = [0] + [1] * (D - 1)
        if kind == 4:                 # awkward floats
            scores = [rng.random() * rng.choice([1, 1e-3, 1e-6, 7.0]) for _ in range(D)]
        if kind == 5 and D >= 3:      # one passage is a prefix-extension of another
            paras[1] = paras[0] + " extra"
        yield q, paras, scores


def golden_spans(rw, tokenizer):
    """(ii) span index / credibility vector: reference vs restatement, bit-exact."""
    ref = make_ref_instance(rw, rw.Re_Weighting_Strategy, tiny_hf_model(), tokenizer, "tiny-llama")
    cases = []
    with cuda_to_cpu():
        for q, paras, scores in span_cases(tokenizer):
            inputs, w_ref = ref.decode_with_special_attention(question=q, paras=paras, scores=scores, answer="")
            prompt, ast = orc.build_prompt(tokenizer, "tiny-llama", q, paras, "")
            enc = tokenizer([prompt], return_offsets_mapping=True, add_special_tokens=ast)
            off = np.array(enc["offset_mapping"][0], dtype=np.int64)
            assert np.array_equal(np.array(enc["input_ids"][0]), inputs["input_ids"][0].numpy())
            spans = orc.char_spans(prompt, paras)
            doc_id, cred = orc.span_index(off[:, 0], spans, scores)
            assert np.array_equal(cred.view(np.uint16), w_ref[0].numpy().view(np.uint16)), "oracle span_index != reference"
            cases.append(dict(offset_starts=off[:, 0], spans=np.array(spans, dtype=np.int32),
                              score_bits=np.array([orc.score_to_fp16(s) for s in scores], dtype=np.float16).view(np.uint16),
                              doc_id=doc_id, cred_bits=cred.view(np.uint16)))
    flat = {}
    for i, c in enumerate(cases):
        for k, v in c.items():
            flat[f"{i}_{k}"] = v
    flat["n_cases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(GOLD, "span_index.npz"), **flat)
    return len(cases)


def golden_tiny_model(rw, tokenizer):
    """(iii)+(iv) tiny-model hooked logits / greedy tokens / IE matrix, fp32 eager on CPU."""
    model = tiny_hf_model()
    sd = tiny_state_dict()
    spec = TINY_SPEC
    q, paras, scores, wrong = synth.nq_sample(11, 4, 24)
    scores = [0.0, 0.3, 0.7, 1.0]
    head_sets = {"none": None, "l0h1": {0: [1]}, "l1h03": {1: [0, 3]}, "all": {0: [0, 1, 2, 3], 1: [0, 1, 2, 3]},
                 "strkeys": {"1": [2], "0": [0]}}
    out = {}
    with cuda_to_cpu(), torch.no_grad():
        for name, hs in head_sets.items():
            ref = make_ref_instance(rw, rw.Re_Weighting_Strategy, model, tokenizer, "tiny-llama", hs)
            if name == "none":
                ref.layers_to_be_modified = {}
            inputs, w = ref.decode_with_special_attention(question=q, paras=paras, scores=scores)
            hooks = []
            for l, heads in ref.layers_to_be_modified.items():
                m = model.get_submodule(f"model.layers.{l}.self_attn")
                hooks.append(m.register_forward_pre_hook(
                    partial(ref.edit_attention_mask, attention_weight=torch.log(w), head_idx=heads), with_kwargs=True))
            logits_ref = model(**inputs, return_dict=True)["logits"]
            for h in hooks:
                h.remove()
            hk = {int(l): (torch.log(w), heads) for l, heads in ref.layers_to_be_modified.items()}
            logits_orc = orc.llama_forward(sd, spec, inputs["input_ids"], hk)
            d = (logits_ref - logits_orc).abs().max().item()
            assert d <= 2e-5, (name, d)
            T = logits_ref.shape[1]
            rows = np.unique(np.concatenate([np.arange(0, T, 24), np.arange(T - 12, T)]))
            out["logit_rows"] = rows
            out[f"logits_{name}"] = logits_ref[0].numpy().astype(np.float32)[rows]
            # greedy generation through the reference's own public call
            prompt_txt, gen_txt = ref.run_RAG_with_attention_weighting(question=q, paras=paras, scores=scores)
            gen = model.generate(**inputs, do_sample=False, max_new_tokens=8) if name == "none" else None
            toks = orc.greedy_generate(sd, spec, inputs["input_ids"], hk, 8, eos_ids=(tokenizer.eos_token_id,))
            out[f"greedy_{name}"] = np.array(toks, dtype=np.int64)
            if gen is not None:
                assert gen[0, inputs["input_ids"].shape[1]:].tolist()[:len(toks)] == toks
            out[f"gen_text_{name}"] = np.array(gen_txt)
        out["input_ids"] = inputs["input_ids"][0].numpy()
        out["cred_bits"] = w[0].numpy().view(np.uint16)
        # IE matrix through the reference's Find_Best_Heads.cal_logits
        fb = make_ref_instance(rw, rw.Find_Best_Heads, model, tokenizer, "tiny-llama")
        sw_scores = [0, 1, 1, 1]
        ie_ref = fb.cal_logits(question=q, paras=paras, scores=sw_scores, wrong_answer=wrong)
        ans_ids = tokenizer([wrong], return_tensors="pt")["input_ids"][0, 1:]
        inputs2, w2 = fb.decode_with_special_attention(question=q, paras=paras, scores=sw_scores, answer=wrong)
        ie_orc = orc.cal_logits(sd, spec, inputs2["input_ids"], torch.log(w2), ans_ids)
        assert np.abs(np.array(ie_ref) - np.array(ie_orc)).max() <= 2e-4, (ie_ref, ie_orc)
        out["ie_ref"] = np.array(ie_ref, dtype=np.float64)
        out["ie_input_ids"] = inputs2["input_ids"][0].numpy()
        out["ie_answer_ids"] = ans_ids.numpy()
        out["ie_cred_bits"] = w2[0].numpy().view(np.uint16)
        out["ie_ori_logit_sum"] = np.array(float(fb.ori_prob_sum))
    meta = dict(question=q, paras=paras, scores=scores, wrong=wrong, sweep_scores=sw_scores)
    with open(os.path.join(GOLD, "tiny_model_meta.json"), "w") as f:
        json.dump(meta, f, ensure_ascii=False, indent=1)
    np.savez_compressed(os.path.join(GOLD, "tiny_model.npz"), **out)


def golden_ranking():
    """(v) ranking KATs: the reference's shipped IE tensors, mean rankings and top-k selections."""
    out = {}
    topk = {"nq/llama3": 894, "nq/qwen": 789, "nq/llama2_13b_chat": 1188,
            "trivia/llama3": 506, "trivia/qwen": 933, "trivia/llama2_13b_chat": 1317}
    for ds in topk:
        base = os.path.join(REF, "datasets", ds)
        key = ds.replace("/", "_")
        hs = os.path.join(base, "heads_scores.json")
        if os.path.exists(hs):
            a = np.array(json.load(open(hs)))
            f = a.astype(np.float32)
            assert (f.astype(np.float64) == a).all()
            out[key + "_scores"] = f
        mean = json.load(open(os.path.join(base, "heads_scores_mean.json")))
        out[key + "_mean_values"] = np.array([m[0] for m in mean], dtype=np.float64)
        out[key + "_mean_index"] = np.array([m[1] for m in mean], dtype=np.int32)
        sel = json.load(open(os.path.join(base, "selected_heads.json")))
        out[key + "_selected_json"] = np.array(json.dumps(sel))
        out[key + "_topk"] = np.array(topk[ds])
    np.savez_compressed(os.path.join(GOLD, "ranking_kat.npz"), **out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_grad_enabled(False)
    tokenizer = synth.load_tokenizer()
    rw = import_reference()
    golden_log_table()
    n = golden_spans(rw, tokenizer)
    golden_tiny_model(rw, tokenizer)
    golden_ranking()
    print("golden fixtures written:", sorted(os.listdir(GOLD)), "span cases:", n)


if __name__ == "__main__":
    main()
