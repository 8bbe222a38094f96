"""EM / F1 scoring of generated answers: the step right after the path.

Same function names, arguments, file formats and edge-case behaviour as the reference's eval.py:13-94
(normalize_text: NFD, lower-case, strip punctuation, drop articles, collapse whitespace; EM is False for an
empty answer; F1 returns 1 when both token lists are empty). Host string work, nothing for the GPU here.
Pinned by tests/golden/pipeline.json, produced by the unmodified reference (oracle/make_golden_pipeline.py).
"""
from __future__ import annotations

import json
import re
import string
import unicodedata
from collections import Counter
from typing import List

import numpy as np

_ARTICLES = re.compile(r"\b(a|an|the)\b")
_PUNCT = set(string.punctuation)


def normalize_text(s: str) -> str:
    """eval.py:13-31."""
    s = unicodedata.normalize("NFD", s).lower()
    s = "".join(ch for ch in s if ch not in _PUNCT)
    s = _ARTICLES.sub(" ", s)
    return " ".join(s.split())


def f1_score(prediction: str, ground_truth: str):
    """eval.py:34-54 (token-level F1 on the normalised strings)."""
    pred, truth = normalize_text(prediction).split(), normalize_text(ground_truth).split()
    if len(pred) == len(truth) == 0:
        print("\n#> F1 Metric: Rare edge case of len(prediction_tokens) == len(ground_truth_tokens) == 0.\n")
        return 1
    num_same = sum((Counter(pred) & Counter(truth)).values())
    if num_same == 0:
        return 0
    precision, recall = 1.0 * num_same / len(pred), 1.0 * num_same / len(truth)
    return (2 * precision * recall) / (precision + recall)


def EM(answer: str, references: List[str]) -> bool:
    """eval.py:57-65."""
    if answer == "":
        return False
    a = normalize_text(answer)
    return any(normalize_text(ref) == a for ref in references)


def F1(prediction: str, answers_list: List[str]):
    """eval.py:68-71."""
    assert type(answers_list) == list
    return max(f1_score(prediction, ans) for ans in answers_list)


def Eval(file_path: str = ""):
    """eval.py:74-94: annotate every record with Eval / f1_score in place and write the `_accuracy` file next
    to the details folder."""
    with open(file_path, "r", encoding="utf-8") as f:
        data = json.load(f)
    cnt, f1_list = 0, []
    for d in data:
        if EM(d["LLM_answer"], d["reference"]):
            d["Eval"] = "Correct"
            cnt += 1
        else:
            d["Eval"] = "Wrong"
        d["f1_score"] = F1(d["LLM_answer"], d["reference"])
        f1_list.append(d["f1_score"])
    with open(file_path, "w", encoding="utf-8") as f:
        json.dump(data, f, ensure_ascii=False, indent=4)
    with open(file_path.replace(".json", "_accuracy").replace("/details", ""), "w", encoding="utf-8") as f:
        f.write(f"Accuracy: {cnt/len(data):.4f}\n")
        f.write(f"EM: {cnt}/{len(data)}\n")
        f.write(f"F1: {np.mean(f1_list):.4f}\n")
