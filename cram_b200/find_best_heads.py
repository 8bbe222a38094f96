"""Head ranking / selection and the IE sweep drivers (reference utils/find_best_heads.py), same
function names, arguments and JSON formats.

Ranking (:113-150) is host integer/float64 work; it calls the C ABI's cram_rank_heads. Tie rule: the
reference's np.argsort()[::-1] leaves the order inside groups of exactly equal means to the numpy
build; here it is defined as value descending, then flat index ascending.
The drivers (:13-110) keep the heads_scores.json format and the resume-by-length rule.
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Sequence, Tuple

import numpy as np

from . import ops


def rank_mean_scores(all_scores) -> List[Tuple[float, Tuple[int, int]]]:
    """[N, L, H] IE samples -> [(mean, (layer, head)), ...] sorted as casual_tracing_combine_all does."""
    a = np.asarray(all_scores, dtype=np.float64)
    _, L, H = a.shape
    order, mean = ops.rank_heads(a)
    return [(float(mean[i]), (int(order[i]) // H, int(order[i]) % H)) for i in range(order.size)]


def select_top_k(ranked: Sequence, topk: int) -> Dict[int, List[int]]:
    """First k ranked heads -> {layer: [heads...]} in first-seen layer order (:135-146)."""
    out: Dict[int, List[int]] = {}
    for _, (l, h) in list(ranked)[:topk]:
        out.setdefault(l, []).append(h)
    return out


def casual_tracing_combine_all(input_path: str = "datasets/nq/Llama-3"):
    with open(os.path.join(input_path, "heads_scores.json"), "r", encoding="utf-8") as f:
        data = json.load(f)
    ranked = rank_mean_scores(data)
    with open(os.path.join(input_path, "heads_scores_mean.json"), "w", encoding="utf-8") as f:
        json.dump(ranked, f, ensure_ascii=False, indent=4)
    return ranked


def find_top_k_heads(input_path: str = "datasets/triviaqa", topk: int = 10):
    with open(os.path.join(input_path, "heads_scores_mean.json"), "r", encoding="utf-8") as f:
        data = json.load(f)
    selected = select_top_k([(v, tuple(i)) for v, i in data], topk)
    with open(os.path.join(input_path, "selected_heads.json"), "w", encoding="utf-8") as f:
        json.dump(selected, f, ensure_ascii=False, indent=4)
    return selected


def _model_dir_name(LLM: str) -> str:
    if "Llama-2-13b" in LLM:
        return "llama2_13b_chat"
    if "Llama-3" in LLM:
        return "llama3"
    if "Qwen" in LLM:
        return "qwen"
    raise UnboundLocalError("cannot access local variable 'model_name'")  # the reference's failure mode


def _dist_of(model):
    """(rank, world_size, group) of a sharded Find_Best_Heads (its IESweep carries them); (0, 1, None) otherwise."""
    sw = getattr(model, "sweep", None)
    return int(getattr(sw, "rank", 0)), int(getattr(sw, "world_size", 1)), getattr(sw, "group", None)


def _load_resume_state(file_path: str, rank: int, world: int, group):
    """The reference's resume rule (:30-35) under torchrun: RANK 0 reads heads_scores.json and every rank gets ITS view
    (one broadcast), so that all ranks skip the same samples and stay in step for cal_logits' all_gather."""
    state = []
    if rank == 0 and os.path.exists(file_path):
        with open(file_path, "r", encoding="utf8") as file:
            state = json.load(file)
    if world > 1:
        import torch.distributed as dist
        box = [state]
        dist.broadcast_object_list(box, src=0, group=group)
        state = box[0]
    return state


def _write_json_rank0(obj, folder: str, file_path: str, rank: int) -> None:
    """Only rank 0 writes (every rank holds the same matrix after the all_gather); temp file + os.replace so that a
    reader -- or a restart after a kill -- never sees a truncated file."""
    if rank != 0:
        return
    os.makedirs(folder, exist_ok=True)
    tmp = file_path + ".tmp"
    with open(tmp, "w", encoding="utf8") as file:
        json.dump(obj, file, ensure_ascii=False, indent=4)
    os.replace(tmp, file_path)


def casual_tracing_per_head(filepath: str = "nq_100_bge.json", LLM: str = "Meta-Llama-3-8B-Instruct",
                            output_dir: str = "datasets/nq", contexts_type: list = ["ori_fake", "reranked_dense"],
                            topk: int = 4, model=None):
    """utils/find_best_heads.py:13-58. `model` lets a caller pass a ready Find_Best_Heads (no hub access here).
    Under torchrun (a Find_Best_Heads built with rank / world_size) every rank runs this same loop: rank 0 owns the
    file, the resume state is broadcast from it, and every rank returns the same list."""
    with open(filepath, "r", encoding="utf-8") as f:
        data = json.load(f)
    model_name = _model_dir_name(LLM)
    folder = os.path.join(output_dir, model_name)
    file_path = os.path.join(folder, "heads_scores.json")
    if model is None:
        from .re_weighting import Find_Best_Heads
        model = Find_Best_Heads(model_name=LLM)
    rank, world, group = _dist_of(model)
    all_prob_changes = _load_resume_state(file_path, rank, world, group)
    for idx, sample in enumerate(data):
        if idx < len(all_prob_changes):  # resume rule: skip what the file already holds
            continue
        contexts, scores = [], []
        if "ori_fake" in contexts_type:
            contexts.extend(sample["ori_fake"])
            scores.extend([0])
        if "reranked_dense" in contexts_type:
            contexts.extend(sample["reranked_dense_ctxs"][:topk])
            scores.extend([1, 1, 1, 1])
        prob_change = model.cal_logits(question=sample["question"], paras=contexts, scores=scores,
                                       wrong_answer=sample["wrong answer"])
        all_prob_changes.append(prob_change)
        _write_json_rank0(all_prob_changes, folder, file_path, rank)
    return all_prob_changes


def casual_tracing_per_head_with_position(filepath: str = "nq_100_bge.json", LLM: str = "Meta-Llama-3-8B-Instruct",
                                          output_dir: str = "datasets/nq",
                                          contexts_type: list = ["ori_fake", "reranked_dense"], topk: int = 4, model=None):
    """utils/find_best_heads.py:61-110: the fake document moved through the 5 positions."""
    with open(filepath, "r", encoding="utf-8") as f:
        data = json.load(f)
    model_name = _model_dir_name(LLM)
    if model is None:
        from .re_weighting import Find_Best_Heads
        model = Find_Best_Heads(model_name=LLM)
    all_prob_changes = []
    folder = os.path.join(output_dir, model_name)
    rank, _, _ = _dist_of(model)
    for sample in data:
        per_position = []
        ori_fake = sample["ori_fake"]
        ctxs = sample["reranked_dense_ctxs"][:topk]
        for position in range(5):
            contexts, scores = [], []
            if "reranked_dense" in contexts_type:
                contexts.extend(ctxs[:position])
                scores.extend([1] * position)
            if "ori_fake" in contexts_type:
                contexts.extend(ori_fake)
                scores.append(0)
            if "reranked_dense" in contexts_type:
                contexts.extend(ctxs[position:])
                scores.extend([1] * (4 - position))
            per_position.append(model.cal_logits(question=sample["question"], paras=contexts, scores=scores,
                                                 wrong_answer=sample["wrong answer"]))
        all_prob_changes.append(per_position)
        _write_json_rank0(all_prob_changes, folder, os.path.join(folder, "heads_scores_pos.json"), rank)
    return all_prob_changes
