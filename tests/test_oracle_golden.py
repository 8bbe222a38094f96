"""CPU: the oracle restatement against the committed golden vectors (made by oracle/make_golden.py from
the UNMODIFIED reference). This is what pins the oracle; the GPU tests then compare CUDA with it."""
import json
import os

import numpy as np
import torch

from oracle import cram_oracle as orc
from oracle.tiny_model import TINY_SPEC, tiny_state_dict


def test_log_chain_table(golden_dir):
    g = np.load(os.path.join(golden_dir, "log_chain.npz"))
    bits = np.arange(65536, dtype=np.uint16)
    x = bits.view(np.float16)
    ok = ~(np.isnan(x) | (x < 0))  # the credibility domain; negative / NaN inputs only differ in NaN payload
    lg16 = orc.log_chain(x, torch.float16).view(torch.int16).numpy().view(np.uint16)
    lgbf = orc.log_chain(x, torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(lg16[ok], g["log_fp16_bits"][ok])
    assert np.array_equal(lgbf[ok], g["log_bf16_bits"][ok])
    # survey anchors
    one = lambda v: float(orc.log_chain(np.array([v], dtype=np.float16), torch.bfloat16)[0])
    assert one(0.25) == -1.390625 and one(0.5) == -0.6953125 and one(0.1) == -2.296875 and one(1.0) == 0.0
    assert one(0.0) == float("-inf")


def test_span_index_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "span_index.npz"))
    n = int(g["n_cases"])
    assert n >= 200
    for i in range(n):
        off = g[f"{i}_offset_starts"]
        spans = [tuple(s) for s in g[f"{i}_spans"]]
        scores = g[f"{i}_score_bits"].view(np.float16).astype(np.float32)
        doc_id, cred = orc.span_index(off, spans, [float(s) for s in scores])
        assert np.array_equal(doc_id, g[f"{i}_doc_id"]), i
        assert np.array_equal(cred.view(np.uint16), g[f"{i}_cred_bits"]), i


def test_span_index_edge_cases():
    off = np.array([0, 0, 5, 9, 9, 14], dtype=np.int64)  # BOS (0,0) + two tokens sharing an offset
    doc_id, cred = orc.span_index(off, [(5, 9)], [0.5])
    assert doc_id.tolist() == [-1, -1, 0, 0, 0, -1]
    # later documents overwrite earlier ones on overlap
    doc_id, _ = orc.span_index(off, [(0, 9), (9, 14)], [0.1, 0.2])
    assert doc_id.tolist() == [-1, 0, 0, 0, 1, 1]  # start = LAST token with offset <= 0, i.e. index 1
    # a passage that was not found (start -1) raises like the reference
    try:
        orc.span_index(np.array([3, 4], dtype=np.int64), [(-1, 2)], [0.3])
        assert False
    except TypeError:
        pass


def test_ranking_kat(golden_dir):
    """The reference's shipped IE tensors -> mean ranking -> top-k selection."""
    g = np.load(os.path.join(golden_dir, "ranking_kat.npz"))
    keys = sorted({k[: -len("_topk")] for k in g.files if k.endswith("_topk")})
    assert len(keys) == 6
    checked = 0
    for key in keys:
        mean_vals, mean_idx = g[key + "_mean_values"], g[key + "_mean_index"]
        if key + "_scores" in g.files:
            ranked = orc.combine_all(g[key + "_scores"].astype(np.float64))
            vals = np.array([r[0] for r in ranked])
            assert np.array_equal(vals, mean_vals), key  # float64 bit-exact
            idx = np.array([r[1] for r in ranked])
            same = (idx == mean_idx).all(axis=1)
            # positions may differ only inside groups of exactly equal means
            for pos in np.nonzero(~same)[0]:
                assert (mean_vals == mean_vals[pos]).sum() > 1, (key, pos)
            checked += 1
        shipped = [(float(v), (int(i[0]), int(i[1]))) for v, i in zip(mean_vals, mean_idx)]
        sel = orc.top_k_heads(shipped, int(g[key + "_topk"]))
        want = json.loads(str(g[key + "_selected_json"]))
        assert [str(k) for k in sel.keys()] == list(want.keys()), key  # dict insertion order too
        assert all(sel[int(k)] == v for k, v in want.items()), key
    assert checked == 5


def test_tiny_model_logits_and_ie(golden_dir):
    g = np.load(os.path.join(golden_dir, "tiny_model.npz"), allow_pickle=False)
    sd, spec = tiny_state_dict(), TINY_SPEC
    ids = torch.from_numpy(g["input_ids"])[None]
    logcred = orc.log_chain(g["cred_bits"].view(np.float16), torch.float32)[None]
    rows = g["logit_rows"]
    with torch.no_grad():
        for name, hooks in (("none", {}), ("l1h03", {1: (logcred, [0, 3])})):
            lg = orc.llama_forward(sd, spec, ids, hooks)[0].numpy()[rows]
            assert np.abs(lg - g[f"logits_{name}"]).max() <= 2e-5, name
        # one row of the IE matrix (the full matrix is checked on the GPU against the same fixture)
        ids2 = torch.from_numpy(g["ie_input_ids"])[None]
        lc2 = orc.log_chain(g["ie_cred_bits"].view(np.float16), torch.float32)[None]
        ans = torch.from_numpy(g["ie_answer_ids"])
        ori = orc.answer_logit_sum(orc.llama_forward(sd, spec, ids2), ans)
        assert abs(float(ori) - float(g["ie_ori_logit_sum"])) <= 2e-4
        cur = orc.answer_logit_sum(orc.llama_forward(sd, spec, ids2, {1: (lc2, [2])}), ans)
        assert abs(float(ori - cur) - g["ie_ref"][1, 2]) <= 2e-4


def test_cached_decode_step_equals_full_recompute():
    """The oracle's cached decode step (hook's tgt_len == 1 branch, re_weighting.py:39-48,56-59) gives
    the logits of a full hooked forward over the same prefix (greedy_generate's docstring claim)."""
    spec = orc.ModelSpec(2, 4, 2, 16, 32, 64, 50, rms_eps=1e-5, rope_theta=10000.0, qkv_bias=True)
    from oracle.tiny_model import random_state_dict
    sd = random_state_dict(spec, seed=3, dtype=torch.float32)
    g = torch.Generator().manual_seed(1)
    T0, T = 9, 13
    ids = torch.randint(0, 50, (1, T), generator=g)
    lc = torch.zeros(1, T0)
    lc[:, 2:6] = float(np.log(np.float16(0.3)))
    lc[:, 6:8] = float("-inf")
    hooks = {0: (lc, [1, 3]), 1: (lc, [0])}
    kc, vc = [None] * spec.num_layers, [None] * spec.num_layers
    with torch.no_grad():
        for t in range(T):
            # during the prompt the bias only exists for keys < T0 that already are in the cache
            step = orc.decode_step_cached(sd, spec, ids[:, t], t, kc, vc, hooks)
            if t >= T0 - 1:
                ext = torch.zeros(1, t + 1)
                ext[:, :T0] = lc[:, :t + 1][:, :T0]
                full = orc.llama_forward(sd, spec, ids[:, :t + 1], {l: (ext, h) for l, (_, h) in hooks.items()})[0, -1]
                assert torch.allclose(step[0], full, atol=2e-5, rtol=1e-5), t


def test_chat_template_branches_oracle_vs_reference_golden(golden_dir):
    """Llama-3 / Qwen branches of utils/re_weighting.py:64-97 (fixtures: oracle/make_golden_chat.py, from the unmodified
    reference): prompt construction + tokenisation + span index bit-exact, greedy tokens incl. the dual-EOS stop and the
    generation_config repetition penalty, one IE entry, and the Qwen 'first real answer token dropped' quirk."""
    from cram_b200 import synth
    from oracle.tiny_model import chat_spec, chat_state_dict
    g = np.load(os.path.join(golden_dir, "chat_model.npz"))
    meta = json.load(open(os.path.join(golden_dir, "chat_model_meta.json")))
    for name in ("llama3", "qwen"):
        m = meta[name]
        tok = synth.load_tokenizer(kind=m["kind"])
        for ci, c in enumerate(m["cases"]):
            prompt, ast = orc.build_prompt(tok, m["model_name"], c["question"], c["paras"], c["answer"])
            enc = tok([prompt], return_offsets_mapping=True, add_special_tokens=ast)
            assert np.array_equal(np.array(enc["input_ids"][0]), g[f"{name}_case{ci}_ids"]), (name, ci)
            off = np.array(enc["offset_mapping"][0], dtype=np.int64)[:, 0]
            doc_id, cred = orc.span_index(off, orc.char_spans(prompt, c["paras"]), c["scores"])
            assert np.array_equal(cred.view(np.uint16), g[f"{name}_case{ci}_cred_bits"]), (name, ci)
            assert np.array_equal(doc_id, g[f"{name}_case{ci}_doc_id"]), (name, ci)
        ids0 = g[f"{name}_case0_ids"]
        if name == "llama3":
            assert ids0[0] == ids0[1] == tok.bos_token_id  # template BOS + add_special_tokens BOS (:77 + :98)
        else:
            assert ids0[0] == tok.convert_tokens_to_ids("<|im_start|>")  # no BOS at all (:93)
        # generation: one hooked head set per branch on the CPU (all four run on the GPU against the same fixture)
        spec, sd = chat_spec(m["kind"]), chat_state_dict(m["kind"], m)
        prompt, ast = orc.build_prompt(tok, m["model_name"], m["question"], m["paras"], "")
        enc = tok([prompt], return_offsets_mapping=True, add_special_tokens=ast)
        ids = torch.tensor(enc["input_ids"])
        off = np.array(enc["offset_mapping"][0], dtype=np.int64)[:, 0]
        _, cred = orc.span_index(off, orc.char_spans(prompt, m["paras"]), m["scores"])
        lc = orc.log_chain(cred, torch.float32)[None]
        eos = ([tok.eos_token_id, tok.convert_tokens_to_ids("<|eot_id|>")] if name == "llama3" else m["eos_token_id"])
        with torch.no_grad():
            for hs_name, hooks in (("none", {}), ("all", {0: (lc, [0, 1, 2, 3]), 1: (lc, [0, 1, 2, 3])})):
                toks = orc.greedy_generate(sd, spec, ids, hooks, m["max_new"], eos_ids=eos,
                                           repetition_penalty=m.get("repetition_penalty", 1.0))
                assert toks == g[f"{name}_gen_{hs_name}"].tolist(), (name, hs_name)
            if name == "llama3":  # stopped early, on the SECOND id of the dual-EOS list (:130-131)
                assert g["llama3_gen_none"][-1] == eos[1] and len(g["llama3_gen_none"]) < m["max_new"]
            ids2 = torch.from_numpy(g[f"{name}_ie_input_ids"])[None]
            lc2 = orc.log_chain(g[f"{name}_ie_cred_bits"].view(np.float16), torch.float32)[None]
            ans = torch.from_numpy(g[f"{name}_ie_answer_ids"])
            ori = orc.answer_logit_sum(orc.llama_forward(sd, spec, ids2), ans)
            assert abs(float(ori) - float(g[f"{name}_ie_ori_logit_sum"])) <= 2e-4
            cur = orc.answer_logit_sum(orc.llama_forward(sd, spec, ids2, {0: (lc2, [3])}), ans)
            assert abs(float(ori - cur) - g[f"{name}_ie_ref"][0, 3]) <= 2e-4
        full = tok([m["wrong"]])["input_ids"][0]
        if name == "qwen":   # :155 drops a REAL token when the tokenizer adds no BOS
            assert g["qwen_ie_answer_ids"].tolist() == full[1:] and len(full) == len(g["qwen_ie_answer_ids"]) + 1
            assert full[0] != tok.convert_tokens_to_ids("<|im_start|>")
        else:                # :155 drops the BOS
            assert full[0] == tok.bos_token_id and g["llama3_ie_answer_ids"].tolist() == full[1:]
