"""Golden vectors for the head-sweep DRIVERS: what the unmodified
utils/find_best_heads.py hands to Find_Best_Heads.cal_logits and what it writes.

TEST INFRASTRUCTURE. Runs in the build container only (needs /root/reference). `fire` is stubbed and the
module's `Find_Best_Heads` name is rebound to a recorder that returns a small deterministic (layers x heads)
matrix per call, so the run captures: the (question, paras, scores, wrong_answer) of every call, the
heads_scores.json / heads_scores_pos.json files, the resume-by-length rule of casual_tracing_per_head
(utils/find_best_heads.py:30-40), and casual_tracing_combine_all + find_top_k_heads on the result.
Output: tests/golden/sweep_drivers.json.

    python oracle/make_golden_sweep.py
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
CALLS = []
L, H = 3, 4


def fake_matrix(question: str, paras, scores):
    """Deterministic stand-in for the per-head probability change: depends on the call's arguments only."""
    base = sum(ord(c) for c in question) % 17
    pos = list(scores).index(0) if 0 in scores else -1
    return [[round(((base + 3 * l + 5 * h + 7 * pos + len(paras)) % 23) / 23.0 - 0.4, 6) for h in range(H)] for l in range(L)]


class _RecorderFBH:
    def __init__(self, model_name="", **kw):
        self.model_name = model_name

    def cal_logits(self, question="", paras=[], scores=[], wrong_answer=""):
        CALLS.append({"question": question, "paras": list(paras), "scores": [float(s) for s in scores],
                      "wrong_answer": wrong_answer})
        return fake_matrix(question, paras, scores)


def synth(n=5):
    return [{"id": i, "question": f"sweep question {i}?", "wrong answer": f"wrong-{i}",
             "ori_fake": [f"fake doc {i}"],
             "reranked_dense_ctxs": [f"dense doc {i}-{k}" for k in range(6)]} for i in range(n)]


def main():
    fire = types.ModuleType("fire")
    fire.Fire = lambda *a, **k: None
    sys.modules["fire"] = fire
    sys.path.insert(0, REF)
    sys.path.insert(1, os.path.join(REF, "utils"))
    import utils.find_best_heads as fbh  # the unmodified reference drivers
    fbh.Find_Best_Heads = _RecorderFBH
    fbh.tqdm = lambda x, *a, **k: x
    data = synth()
    out = {"data": data, "L": L, "H": H, "runs": []}
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            json.dump(data, open("nq_sweep.json", "w"))
            json.dump(data[:2], open("nq_sweep_first2.json", "w"))
            # 1) fresh run over the first two samples, 2) resumed run over all five (skips idx < 2)
            for name, fp, llm in [("fresh_first2", "nq_sweep_first2.json", "Meta-Llama-3-8B-Instruct"),
                                  ("resumed_all", "nq_sweep.json", "Meta-Llama-3-8B-Instruct"),
                                  ("qwen_dense_only", "nq_sweep.json", "Qwen2-7B-Instruct"),
                                  ("llama2_13b", "nq_sweep_first2.json", "Llama-2-13b-chat-hf")]:
                CALLS.clear()
                kw = dict(filepath=fp, LLM=llm, output_dir="out/nq")
                if name == "qwen_dense_only":
                    kw["contexts_type"] = ["reranked_dense"]
                fbh.casual_tracing_per_head(**kw)
                model_dir = {"Meta": "llama3", "Qwen": "qwen", "Llam": "llama2_13b_chat"}[llm[:4]]
                rel = os.path.join("out/nq", model_dir, "heads_scores.json")
                out["runs"].append({"driver": "casual_tracing_per_head", "name": name, "kwargs": kw,
                                    "calls": [dict(c) for c in CALLS], "file": rel, "content": json.load(open(rel))})
            CALLS.clear()
            kw = dict(filepath="nq_sweep_first2.json", LLM="Meta-Llama-3-8B-Instruct", output_dir="out_pos/nq")
            fbh.casual_tracing_per_head_with_position(**kw)
            rel = "out_pos/nq/llama3/heads_scores_pos.json"
            out["runs"].append({"driver": "casual_tracing_per_head_with_position", "name": "positions", "kwargs": kw,
                                "calls": [dict(c) for c in CALLS], "file": rel, "content": json.load(open(rel))})
            ranked = fbh.casual_tracing_combine_all(input_path="out/nq/llama3")
            out["ranked"] = json.load(open("out/nq/llama3/heads_scores_mean.json"))
            assert len(ranked) == L * H
            fbh.find_top_k_heads(input_path="out/nq/llama3", topk=5)
            out["selected_top5"] = json.load(open("out/nq/llama3/selected_heads.json"))
        finally:
            os.chdir(cwd)
    path = os.path.join(ROOT, "tests", "golden", "sweep_drivers.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=1)
    print("wrote", path, [(r["name"], len(r["calls"])) for r in out["runs"]])


if __name__ == "__main__":
    main()
