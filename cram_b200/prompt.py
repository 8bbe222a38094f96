"""Prompt format of the credibility path (data contract, not code under optimisation).

The span finder searches the prompt for exactly ``"Passage-%d: " + para + "\\n"``
(reference utils/re_weighting.py:100-102), so the 'with_contexts' template of
utils/prompts_templates.py:5-6 and the joining rule of utils/prompt.py:4-12 are part of the path's
input format. Only the two QA prompt types the path uses are provided.
"""
from __future__ import annotations

from typing import Sequence

prompt_dict = {
    "qa": {
        "without_contexts": ("Answer the following question based on your internal knowledge with one or few "
                             "words.\nQuestion: {question}\nAnswer: "),
        "with_contexts": ("Given the following information: \n{paras}\nAnswer the following question based on the "
                          "given information or your internal knowledge with one or few words without the source.\n"
                          "Question: {question}\nAnswer: {answer}"),
    },
}


def get_prompt(context: Sequence[str] = (), question: str = "", answer: str = "", type: str = "with_contexts",
               scores: Sequence = ()) -> str:
    if type == "with_contexts":
        paras = "\n".join(("Passage-%d: " % i) + para for i, para in enumerate(context))
        return prompt_dict["qa"][type].format(question=question, paras=paras, answer=answer)
    if type == "without_contexts":
        return prompt_dict["qa"][type].format(question=question)
    raise ValueError(f"prompt type {type!r} is outside the credibility path (only 'with_contexts' / 'without_contexts')")
