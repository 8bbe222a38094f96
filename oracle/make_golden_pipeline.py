"""Golden vectors for the callers either side of the path (f2 score pipeline, f4 EM/F1).

TEST INFRASTRUCTURE. Runs in the build container only (needs /root/reference). The UNMODIFIED reference
front end -- RAG_test.py:RAG_gen and eval.py -- is imported with its missing third-party imports (fire,
langchain, faiss, colorama, openai ...) stubbed out and `utils.RAG.RAG` replaced by a recorder, so that
for every sample we capture exactly what RAG_gen hands to run_RAG (question, ordered contexts, normalised
scores), the file it writes and what eval.py computes on it. Output: tests/golden/pipeline.json.

    python oracle/make_golden_pipeline.py
"""
from __future__ import annotations

import json
import os
import random
import sys
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
CALLS = []


class _RecorderRAG:
    """Stands in for utils/RAG.py:RAG: records run_RAG arguments, answers from a fixed table."""

    def __init__(self, model_name="openai", **kw):
        self.model_name, self.kw = model_name, kw

    def run_RAG(self, question="", paras=[], scores=[], type="with_contexts"):
        CALLS.append({"question": question, "paras": list(paras), "scores": [float(s) for s in scores], "type": type})
        ans = _ANSWERS[question]
        return "PROMPT::" + question, ans

    def run_RAG_without_contexts(self, question=""):
        return "PROMPT::" + question, _ANSWERS[question]


_ANSWERS = {}


def _stub_modules():
    fire = types.ModuleType("fire")
    fire.Fire = lambda *a, **k: None
    sys.modules["fire"] = fire
    sys.path.insert(0, REF)
    sys.path.insert(1, os.path.join(REF, "utils"))  # re_weighting.py:11 imports `prompt` flat
    import utils  # noqa: F401  (the reference package; __init__ is empty)
    rag = types.ModuleType("utils.RAG")
    rag.RAG = _RecorderRAG
    sys.modules["utils.RAG"] = rag


def synth_dataset(n=12, seed=0):
    rng = random.Random(seed)
    words = ["alpha", "beta", "gamma", "delta", "The", "a", "an", "river", "Paris", "1969", "moon", "Apollo", "café", "naïve"]
    data = []
    for i in range(n):
        ref = [" ".join(rng.sample(words, rng.randint(1, 3))) for _ in range(rng.randint(1, 3))]
        q = f"question number {i} about {rng.choice(words)}?"
        kind = i % 6
        if kind == 0:
            ans = ref[0]
        elif kind == 1:
            ans = "The " + ref[0].upper() + "!!"
        elif kind == 2:
            ans = ref[0].split()[0] + " extra tokens here"
        elif kind == 3:
            ans = ""
        elif kind == 4:
            ans = "completely different"
        else:
            ans = " " + ref[-1] + ". "
        _ANSWERS[q] = ans
        data.append({
            "id": i, "question": q, "reference": ref, "wrong answer": rng.choice(words),
            "ori_fake": [f"fake passage {i}-{k} " + " ".join(rng.choices(words, k=6)) for k in range(3)],
            "ori_fake_truthful_scores": [rng.randint(0, 10) for _ in range(3)],
            "ori_fake_relevant_scores": [rng.randint(0, 10) for _ in range(3)],
            "reranked_dense_ctxs": [f"dense passage {i}-{k} " + " ".join(rng.choices(words, k=6)) for k in range(5)],
            "reranked_dense_ctxs_truthful_scores": [str(rng.randint(0, 10)) if k % 2 else rng.randint(0, 10) for k in range(5)],
            "reranked_dense_ctxs_relevant_scores": [rng.randint(0, 10) for _ in range(5)],
        })
    # degenerate score rows the normalisation special-cases (RAG_test.py:96-101): all equal, all zero
    data[3]["ori_fake_truthful_scores"] = [7, 7, 7]
    data[3]["reranked_dense_ctxs_truthful_scores"] = [7, 7, 7, 7, 7]
    data[4]["ori_fake_truthful_scores"] = [0, 0, 0]
    data[4]["reranked_dense_ctxs_truthful_scores"] = [0, 0, 0, 0, 0]
    return data


CONFIGS = [
    dict(LLM="llama3", whether_re_weighting=True, modify_all_layers=True, whether_ideal_scores=True, ideal_score=0),
    dict(LLM="llama3", whether_re_weighting=True, modify_all_layers=True, whether_ideal_scores=True, ideal_score=3, fake_num=2),
    dict(LLM="qwen", whether_re_weighting=True, modify_all_layers=True, whether_ideal_scores=False),
    dict(LLM="llama2_13b_chat", whether_re_weighting=True, modify_all_layers=True, whether_ideal_scores=False, whether_relevant=True,
         topk=3),
    dict(LLM="llama3", whether_re_weighting=True, modify_all_layers=False, heads_topk=5, whether_ideal_scores=True, ideal_score=0),
    dict(LLM="llama3", whether_re_weighting=True, modify_all_layers=True, whether_ideal_scores=True, ideal_score=0, causal_num=50),
]


def main():
    _stub_modules()
    import RAG_test  # the unmodified reference front end
    import eval as ref_eval
    data = synth_dataset()
    out = {"data": data, "answers": _ANSWERS, "runs": []}
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            with open("nq_synth_bge.json", "w", encoding="utf-8") as f:
                json.dump(data, f)
            # head files for the modify_all_layers=False run (RAG_test.py:50-58 reads dataset/<name>/<LLM>)
            os.makedirs("dataset/nq/llama3")
            ranked = [[1.0 - 0.1 * i, [i % 3, (7 * i) % 4]] for i in range(12)]
            with open("dataset/nq/llama3/heads_scores_mean.json", "w") as f:
                json.dump(ranked, f)
            for cfg in CONFIGS:
                CALLS.clear()
                before = {os.path.join(dp, fn) for dp, _, fns in os.walk(".") for fn in fns}
                RAG_test.RAG_gen(filepath="nq_synth_bge.json", **cfg)
                after = {os.path.join(dp, fn) for dp, _, fns in os.walk(".") for fn in fns}
                new_files = sorted(after - before)
                details = [p for p in new_files if p.endswith(".json") and "/details/" in p]
                acc = [p for p in new_files if p.endswith("_accuracy")]
                run = {"config": cfg, "calls": [dict(c) for c in CALLS], "details_path": details[0],
                       "accuracy_path": acc[0], "accuracy_text": open(acc[0], encoding="utf-8").read(),
                       "details": json.load(open(details[0], encoding="utf-8"))}
                if not cfg.get("modify_all_layers", False):
                    run["selected_heads"] = json.load(open("dataset/nq/llama3/selected_heads.json"))
                out["runs"].append(run)
        finally:
            os.chdir(cwd)
    # eval.py primitives on a table of strings
    preds = ["The Eiffel Tower", "eiffel tower.", "", "a an the", "Émile Zola", "Emile Zola", "new  york\tcity", "42", "forty two"]
    refs = [["Eiffel Tower"], ["the eiffel tower", "tower"], ["x"], [""], ["Emile Zola"], ["Émile Zola", "zola"], ["New York City"],
            ["42", "forty-two"], ["forty-two", "42"]]
    out["eval_table"] = [{"pred": p, "refs": r, "norm": ref_eval.normalize_text(p), "EM": bool(ref_eval.EM(p, r)),
                          "F1": float(ref_eval.F1(p, r))} for p, r in zip(preds, refs)]
    path = os.path.join(ROOT, "tests", "golden", "pipeline.json")
    with open(path, "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=1)
    print("wrote", path, "runs:", len(out["runs"]), "calls per run:", [len(r["calls"]) for r in out["runs"]])


if __name__ == "__main__":
    main()
