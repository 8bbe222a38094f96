"""The head-sweep drivers against the unmodified reference drivers' recorded behaviour
(tests/golden/sweep_drivers.json, made by oracle/make_golden_sweep.py): arguments of every cal_logits call,
the files written, the resume-by-length rule, ranking and top-k selection. CPU only, a fake sweep model."""
import json
import os

import pytest

from cram_b200 import find_best_heads as fbh

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sweep_drivers.json")


def _fake_matrix(L, H, question, paras, scores):  # same formula as oracle/make_golden_sweep.py:fake_matrix
    base = sum(ord(c) for c in question) % 17
    pos = list(scores).index(0) if 0 in scores else -1
    return [[round(((base + 3 * l + 5 * h + 7 * pos + len(paras)) % 23) / 23.0 - 0.4, 6) for h in range(H)] for l in range(L)]


class _Model:
    def __init__(self, L, H):
        self.L, self.H, self.calls = L, H, []

    def cal_logits(self, question="", paras=[], scores=[], wrong_answer=""):
        self.calls.append({"question": question, "paras": list(paras), "scores": [float(s) for s in scores],
                           "wrong_answer": wrong_answer})
        return _fake_matrix(self.L, self.H, question, paras, scores)


@pytest.fixture()
def gold():
    with open(GOLD, encoding="utf-8") as f:
        return json.load(f)


def test_drivers_make_the_reference_calls_and_files(gold, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    json.dump(gold["data"], open("nq_sweep.json", "w"))
    json.dump(gold["data"][:2], open("nq_sweep_first2.json", "w"))
    for run in gold["runs"]:  # order matters: "resumed_all" continues the file "fresh_first2" left
        model = _Model(gold["L"], gold["H"])
        driver = getattr(fbh, run["driver"])
        got = driver(model=model, **run["kwargs"])
        assert model.calls == run["calls"], run["name"]
        assert json.load(open(run["file"])) == run["content"], run["name"]
        assert got == run["content"]
    assert [c["question"] for c in gold["runs"][1]["calls"]] == [d["question"] for d in gold["data"][2:]]  # resumed
    ranked = fbh.casual_tracing_combine_all(input_path="out/nq/llama3")
    assert json.load(open("out/nq/llama3/heads_scores_mean.json")) == gold["ranked"]
    assert [[v, list(i)] for v, i in ranked] == gold["ranked"]
    fbh.find_top_k_heads(input_path="out/nq/llama3", topk=5)
    assert json.load(open("out/nq/llama3/selected_heads.json")) == gold["selected_top5"]


def test_unknown_model_name_fails_like_the_reference(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    json.dump([], open("d.json", "w"))
    with pytest.raises(UnboundLocalError):  # utils/find_best_heads.py:23-28 leaves model_name unbound
        fbh.casual_tracing_per_head(filepath="d.json", LLM="mistral", model=_Model(1, 1))
