"""The caller in front of the path: RAG_test.py's credibility-score pipeline and result
files for the re-weighting experiments, as a BATCHED front end feeding per-sample credibility to
Re_Weighting_Strategy.run_RAG_batch (the reference loops at b = 1 through RAG.run_RAG, utils/RAG.py:131-134).

Kept from the reference (RAG_test.py): which contexts and scores a sample contributes (:64-90), the score
normalisation int 0-10 -> min-shift -> max-normalise with its two degenerate cases (:91-101), the stable
ascending sort of (score, context) pairs (:102-106), the record format (Log_gen, :15-23), the result folder /
file naming (:136-171) and the final Eval call. Not kept: retriever, reranker, OpenAI paths, prompt-score
baselines ("with_contexts_score", "with_contexts_cag") and paragraph deletion -- other experiments, off the path.
Pinned by tests/golden/pipeline.json (the unmodified RAG_gen run with a recording RAG class).
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Sequence, Tuple

import numpy as np

from .eval import Eval
from .find_best_heads import find_top_k_heads


def Log_gen(sample, prompt, result):
    """RAG_test.py:15-23."""
    return {"id": sample["id"], "prompt": prompt, "question": sample["question"], "LLM_answer": result,
            "reference": sample["reference"], "wrong answer": sample["wrong answer"]}


def contexts_and_scores(sample: dict, contexts_type: Sequence[str] = ("ori_fake", "reranked_dense"),
                        whether_ideal_scores: bool = False, ideal_score: int = 0, whether_relevant: bool = False,
                        topk: int = 4, fake_num: int = 1) -> Tuple[List[str], List[float]]:
    """RAG_test.py:64-106 for whether_re_weighting=True, whether_delete_false_paras=False: the documents of one
    sample in the order they enter the prompt (ascending credibility) and their scores in [0, 1]."""
    contexts: List[str] = []
    truthful, relevant = [], []
    if "ori_fake" in contexts_type:
        contexts.extend(sample["ori_fake"][:fake_num])
        if whether_ideal_scores:
            truthful.extend([ideal_score] * fake_num)
            relevant.extend([10] * fake_num)
        else:
            truthful.extend(sample["ori_fake_truthful_scores"][:fake_num])
            if whether_relevant:
                relevant.extend(sample["ori_fake_relevant_scores"][:fake_num])
    if "reranked_dense" in contexts_type:
        contexts.extend(sample["reranked_dense_ctxs"][:topk])
        if whether_ideal_scores:
            truthful.extend([10] * topk)
            relevant.extend([10] * topk)
        else:
            truthful.extend(sample["reranked_dense_ctxs_truthful_scores"][:topk])
            if whether_relevant:
                relevant.extend(sample["reranked_dense_ctxs_relevant_scores"][:topk])
    if whether_relevant:
        scores = np.array([int(s) for s in truthful]) * np.array([int(s) for s in relevant])
    else:
        scores = np.array([int(s) for s in truthful])
        if not whether_ideal_scores:
            scores -= np.min(scores)
    if np.max(scores) != 0:
        scores = scores / np.max(scores)
    else:
        scores = scores + 1
    combined = sorted(zip(scores, contexts), key=lambda x: x[0])  # stable: ties keep document order
    scores, contexts = zip(*combined)
    return list(contexts), [float(s) for s in scores]


def result_path(filepath: str, LLM: str, type: str = "with_contexts",
                contexts_type: Sequence[str] = ("ori_fake", "reranked_dense"), whether_ideal_scores: bool = False,
                ideal_score: int = 0, modify_all_layers: bool = False, heads_topk: int = 559, topk: int = 4,
                fake_num: int = 1, causal_num: int = 100) -> str:
    """RAG_test.py:136-171 for the re-weighting runs: where the details JSON of a run goes."""
    stem = filepath.split(".")[0]
    if not whether_ideal_scores:
        folder = f"./results_real/{stem}/{LLM}/{type}/details"
    elif causal_num == 100:
        folder = f"./results{ideal_score}/{stem}/{LLM}/{type}/details"
    else:
        folder = f"./results{ideal_score}_{causal_num}/{stem}/{LLM}/{type}/details"
    parts = ["_".join(contexts_type) + f"_top{topk}", "weighting"]
    if whether_ideal_scores:
        parts.append("ideal")
    if not modify_all_layers:
        parts.append(f"heads{heads_topk}")
    filename = "_".join(parts) + ".json"
    if fake_num != 1:
        filename = filename.replace("ori_fake", f"ori_fake_{fake_num}")
    return os.path.join(folder, filename)


def head_set_for_run(filepath: str, LLM: str, modify_all_layers: bool, heads_topk: int, causal_num: int = 100) -> Dict:
    """RAG_test.py:49-60: the top-k head selection file of the dataset/model, or {} for "all heads"."""
    if modify_all_layers:
        return dict()
    root = "dataset" if causal_num == 100 else f"Head_data{causal_num}"
    input_path = os.path.join(root, filepath.split("_")[0], LLM)
    find_top_k_heads(input_path=input_path, topk=heads_topk)
    with open(os.path.join(input_path, "selected_heads.json"), "r", encoding="utf-8") as f:
        return json.load(f)


def RAG_gen(rw, filepath: str = "nq_1000_bge.json", LLM: str = "llama3", type: str = "with_contexts",
            contexts_type: Sequence[str] = ("ori_fake", "reranked_dense"), whether_ideal_scores: bool = False,
            ideal_score: int = 0, modify_all_layers: bool = False, heads_topk: int = 559, whether_relevant: bool = False,
            topk: int = 4, fake_num: int = 1, causal_num: int = 100, batch_size: int = 16, max_new_tokens: int = 100,
            run_eval: bool = True) -> str:
    """RAG_test.py:RAG_gen with whether_re_weighting=True, batched: `batch_size` samples per call of
    rw.run_RAG_batch (per-sample credibility inside one packed prefill + one paged decode). `rw` is a
    Re_Weighting_Strategy whose layers_to_be_modified is set here from the run's head file. Returns the path of
    the details JSON (annotated by Eval, with the `_accuracy` file beside its folder, as the reference leaves it)."""
    assert type == "with_contexts", "only the re-weighting experiment is on the path"
    with open(filepath, "r", encoding="utf-8") as f:
        data = json.load(f)
    layers = head_set_for_run(filepath, LLM, modify_all_layers, heads_topk, causal_num)
    if not layers:  # utils/re_weighting.py:32-33: empty selection = every head of every layer
        layers = {l: list(range(rw.model_num_attention_heads)) for l in range(rw.num_hidden_layers)}
    rw.layers_to_be_modified = layers
    file_path = result_path(filepath, LLM, type, contexts_type, whether_ideal_scores, ideal_score, modify_all_layers,
                            heads_topk, topk, fake_num, causal_num)
    os.makedirs(os.path.dirname(file_path), exist_ok=True)
    prompt_and_answer = []
    for b0 in range(0, len(data), batch_size):
        batch = data[b0:b0 + batch_size]
        prepared = [contexts_and_scores(s, contexts_type, whether_ideal_scores, ideal_score, whether_relevant, topk, fake_num)
                    for s in batch]
        results = rw.run_RAG_batch([s["question"] for s in batch], [c for c, _ in prepared], [sc for _, sc in prepared],
                                   max_new_tokens=max_new_tokens)
        for s, (prompt, result) in zip(batch, results):
            prompt_and_answer.append(Log_gen(sample=s, prompt=prompt, result=result))
        with open(file_path, "w", encoding="utf8") as file:
            json.dump(prompt_and_answer, file, ensure_ascii=False, indent=4)
    if run_eval:
        Eval(file_path)
    return file_path
