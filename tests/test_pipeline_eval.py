"""CPU: the callers either side of the path (f2 score pipeline + result files, f4 EM/F1) against
what the UNMODIFIED reference front end did on the same synthetic dataset (tests/golden/pipeline.json, made by
oracle/make_golden_pipeline.py with a recording RAG class)."""
import json
import os

import pytest

from cram_b200 import eval as ev
from cram_b200 import pipeline as pl


@pytest.fixture(scope="module")
def gold(golden_dir):
    return json.load(open(os.path.join(golden_dir, "pipeline.json"), encoding="utf-8"))


def test_eval_primitives(gold):
    for row in gold["eval_table"]:
        assert ev.normalize_text(row["pred"]) == row["norm"]
        assert bool(ev.EM(row["pred"], row["refs"])) == row["EM"]
        assert float(ev.F1(row["pred"], row["refs"])) == row["F1"]  # same arithmetic, bit-exact


def _cfg(run):
    c = dict(run["config"])
    c.pop("whether_re_weighting")
    return c


def test_contexts_and_scores_match_reference_calls(gold):
    for run in gold["runs"]:
        c = _cfg(run)
        kw = {k: c[k] for k in ("whether_ideal_scores", "ideal_score", "whether_relevant", "topk", "fake_num") if k in c}
        for sample, call in zip(gold["data"], run["calls"]):
            ctx, sc = pl.contexts_and_scores(sample, **kw)
            assert ctx == call["paras"]
            assert sc == call["scores"]  # bit-exact floats
            assert sc == sorted(sc)


class _FakeRW:
    """Stands in for Re_Weighting_Strategy on the CPU box: answers from the golden table, records its calls."""
    num_hidden_layers, model_num_attention_heads = 3, 4

    def __init__(self, answers):
        self.answers, self.calls, self.layers_to_be_modified = answers, [], {}

    def run_RAG_batch(self, questions, paras_list, scores_list, max_new_tokens=100):
        self.calls.append((list(questions), [list(p) for p in paras_list], [list(s) for s in scores_list]))
        return [("PROMPT::" + q, self.answers[q]) for q in questions]


@pytest.mark.parametrize("batch_size", [1, 5, 16])
def test_rag_gen_writes_the_reference_files(gold, tmp_path, monkeypatch, batch_size):
    monkeypatch.chdir(tmp_path)
    json.dump(gold["data"], open("nq_synth_bge.json", "w", encoding="utf-8"))
    os.makedirs("dataset/nq/llama3")
    json.dump([[1.0 - 0.1 * i, [i % 3, (7 * i) % 4]] for i in range(12)], open("dataset/nq/llama3/heads_scores_mean.json", "w"))
    for run in gold["runs"]:
        rw = _FakeRW(gold["answers"])
        c = _cfg(run)
        path = pl.RAG_gen(rw, filepath="nq_synth_bge.json", batch_size=batch_size, **c)
        assert os.path.normpath(path) == os.path.normpath(run["details_path"])
        assert json.load(open(path, encoding="utf-8")) == run["details"]
        assert open(run["accuracy_path"], encoding="utf-8").read() == run["accuracy_text"]
        flat_q = [q for call in rw.calls for q in call[0]]
        assert flat_q == [c_["question"] for c_ in run["calls"]]
        assert [s for call in rw.calls for s in call[2]] == [c_["scores"] for c_ in run["calls"]]
        if "selected_heads" in run:
            assert {str(k): v for k, v in rw.layers_to_be_modified.items()} == run["selected_heads"]
        else:  # empty selection means every head of every layer (utils/re_weighting.py:32-33)
            assert rw.layers_to_be_modified == {l: [0, 1, 2, 3] for l in range(3)}
