"""Seeded synthetic inputs for the CrAM hot path: NQ-shaped prompts and a local tokenizer.

There is no network, no dataset and no tokenizer file on the boxes this runs on, so everything
here is generated from a seed:

* a pseudo-English word list (seed 0) from which questions and DPR-style ~100-word passages are drawn;
* a byte-level BPE tokenizer trained offline on that text (vocab 1000, ``<s>`` prepended through a
  ``TemplateProcessing`` post-processor so that the BOS token carries the (0, 0) offset the
  reference's span finder silently relies on, /root/reference/utils/re_weighting.py:104-110);
* credibility scores shaped like the reference's experiment driver produces them: integers 0..10,
  min-shifted, max-normalised and the documents sorted ascending by score
  (/root/reference/RAG_test.py:92-107).

The trained tokenizer is committed as ``cram_b200/data/tokenizer.json`` so the GPU box needs no
training step; ``python -m cram_b200.synth`` regenerates it.
"""
from __future__ import annotations

import json
import os
import random
from typing import List, Sequence, Tuple

_DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
TOKENIZER_JSON = os.path.join(_DATA_DIR, "tokenizer.json")

_ONSETS = ["b", "br", "c", "ch", "d", "dr", "f", "fl", "g", "gr", "h", "j", "k", "l", "m", "n", "p",
           "pl", "qu", "r", "s", "sh", "st", "t", "th", "tr", "v", "w", "y", "z", ""]
_NUCLEI = ["a", "e", "i", "o", "u", "ai", "ea", "ee", "ou", "io"]
_CODAS = ["", "n", "r", "s", "t", "l", "m", "nd", "st", "ck", "ng", "rt"]
_FUNCTION_WORDS = ["the", "of", "and", "in", "to", "a", "is", "was", "for", "by", "on", "with", "as",
                   "at", "from", "that", "which", "it", "its", "were", "an", "be", "this", "has"]


def word_list(n: int = 2000, seed: int = 0) -> List[str]:
    """Deterministic pseudo-English vocabulary (content words + a few function words)."""
    rng = random.Random(seed)
    words = []
    seen = set(_FUNCTION_WORDS)
    while len(words) < n:
        syl = rng.choice([1, 2, 2, 3])
        w = "".join(rng.choice(_ONSETS) + rng.choice(_NUCLEI) + rng.choice(_CODAS) for _ in range(syl))
        if len(w) >= 3 and w not in seen:
            seen.add(w)
            words.append(w)
    return words


_WORDS = None


def _words() -> List[str]:
    global _WORDS
    if _WORDS is None:
        _WORDS = word_list()
    return _WORDS


def sentence(rng: random.Random, n_words: int) -> str:
    ws = _words()
    out = []
    for i in range(n_words):
        if i % 3 == 1:
            out.append(rng.choice(_FUNCTION_WORDS))
        else:
            w = rng.choice(ws)
            if rng.random() < 0.08:
                w = w.capitalize()
            out.append(w)
    s = " ".join(out)
    return s[0].upper() + s[1:]


def passage(rng: random.Random, n_words: int = 100) -> str:
    """A DPR-style passage: a few sentences, ~n_words words, no newline inside."""
    parts = []
    left = n_words
    while left > 0:
        k = min(left, rng.randint(8, 22))
        parts.append(sentence(rng, k) + rng.choice([".", ".", ".", ",", ";"]))
        left -= k
    text = " ".join(parts)
    if rng.random() < 0.15:  # multi-byte characters exercise byte-fallback tokens sharing an offset
        text += " " + rng.choice(["café", "naïve", "Zürich", "東京", "São Paulo", "—"]) + "."
    return text


def question(rng: random.Random) -> str:
    n = rng.randint(8, 15)
    return rng.choice(["who", "what", "when", "where", "which"]) + " " + sentence(rng, n - 1).lower() + "?"


def normalised_scores(raw: Sequence[int]) -> List[float]:
    """Integers 0..10 -> min-shift -> max-normalise, as /root/reference/RAG_test.py:92-103 does."""
    lo = min(raw)
    shifted = [s - lo for s in raw]
    hi = max(shifted)
    if hi == 0:
        return [1.0 for _ in shifted]
    return [s / hi for s in shifted]


def nq_sample(seed: int, n_docs: int = 5, words_per_doc: int = 100,
              sweep_scores: bool = False) -> Tuple[str, List[str], List[float], str]:
    """One NQ-shaped sample: (question, paras, scores, wrong_answer).

    ``sweep_scores`` gives the IE-sweep pattern [0, 1, 1, ...] of
    /root/reference/utils/find_best_heads.py:43-48 (first document is the fake one).
    Otherwise scores follow the RAG_test.py normalisation and documents are sorted ascending.
    """
    rng = random.Random(0xC4A3 + seed)
    q = question(rng)
    paras = [passage(rng, words_per_doc) for _ in range(n_docs)]
    if sweep_scores:
        scores = [0] + [1] * (n_docs - 1)
    else:
        raw = [rng.randint(0, 10) for _ in range(n_docs)]
        sc = normalised_scores(raw)
        order = sorted(range(n_docs), key=lambda i: sc[i])
        paras = [paras[i] for i in order]
        scores = [sc[i] for i in order]
    wrong = " ".join(rng.choice(_words()) for _ in range(rng.randint(2, 4)))
    return q, paras, scores, wrong


def training_corpus(n: int = 400) -> List[str]:
    rng = random.Random(12345)
    docs = [passage(rng, 100) for _ in range(n)]
    docs += [question(rng) for _ in range(n // 4)]
    docs += ["Given the following information: \nPassage-%d: " % (i % 20) for i in range(40)]
    docs += ["Answer the following question based on the given information or your internal knowledge "
             "with one or few words without the source.\nQuestion: \nAnswer: "] * 8
    return docs


def train_tokenizer(path: str = TOKENIZER_JSON, vocab_size: int = 1000) -> str:
    """Train the byte-level BPE and write tokenizer.json (deterministic for a given tokenizers build)."""
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers, processors, trainers

    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tok.decoder = decoders.ByteLevel()
    trainer = trainers.BpeTrainer(vocab_size=vocab_size, special_tokens=["<s>", "</s>", "<pad>"],
                                  initial_alphabet=pre_tokenizers.ByteLevel.alphabet(), show_progress=False)
    tok.train_from_iterator(training_corpus(), trainer)
    tok.post_processor = processors.TemplateProcessing(
        single="<s> $A", pair="<s> $A $B", special_tokens=[("<s>", tok.token_to_id("<s>"))])
    os.makedirs(os.path.dirname(path), exist_ok=True)
    tok.save(path)
    return path


# The public chat templates of the two model families the reference special-cases
# (/root/reference/utils/re_weighting.py:64-95). Llama-3-Instruct: the template itself emits the BOS token, so that
# tokenising the rendered string with add_special_tokens=True (:98) yields a DOUBLE BOS; the message content is
# trimmed. Qwen (ChatML): no BOS anywhere, the reference tokenises with add_special_tokens=False (:93).
LLAMA3_CHAT_TEMPLATE = (
    "{% set loop_messages = messages %}{% for message in loop_messages %}"
    "{% set content = '<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n'+ message['content'] | trim + '<|eot_id|>' %}"
    "{% if loop.index0 == 0 %}{% set content = bos_token + content %}{% endif %}{{ content }}{% endfor %}"
    "{% if add_generation_prompt %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}{% endif %}")
QWEN_CHAT_TEMPLATE = (
    "{% for message in messages %}{% if loop.first and messages[0]['role'] != 'system' %}"
    "{{ '<|im_start|>system\nYou are a helpful assistant.<|im_end|>\n' }}{% endif %}"
    "{{'<|im_start|>' + message['role'] + '\n' + message['content'] + '<|im_end|>' + '\n'}}{% endfor %}"
    "{% if add_generation_prompt %}{{ '<|im_start|>assistant\n' }}{% endif %}")
LLAMA3_SPECIALS = ["<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
QWEN_SPECIALS = ["<|im_start|>", "<|im_end|>", "<|endoftext|>"]
CHAT_VOCAB = 1008  # 1000 BPE entries + room for the added special tokens (embedding rows of the chat test models)


def load_tokenizer(path: str = TOKENIZER_JSON, kind: str = "raw"):
    """The committed tokenizer wrapped as a HF fast tokenizer (gives real ``offset_mapping``).

    ``kind``: "raw" -- BOS ``<s>`` prepended by the post-processor, no chat template (the reference's Llama-2 / else
    branch); "llama3" -- the same BOS post-processor plus the Llama-3 header / ``<|eot_id|>`` special tokens and chat
    template (double BOS under add_special_tokens=True); "qwen" -- ChatML special tokens and template, NO BOS
    post-processor, eos ``<|im_end|>`` (so ``tokenizer([answer]).input_ids[0, 1:]`` drops a real token,
    /root/reference/utils/re_weighting.py:155). The BPE vocabulary (ids 0..999) is shared by all three.
    """
    from transformers import PreTrainedTokenizerFast

    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} missing: run `python -m cram_b200.synth` to train it")
    if kind == "raw":
        return PreTrainedTokenizerFast(tokenizer_file=path, bos_token="<s>", eos_token="</s>", pad_token="<pad>")
    from tokenizers import AddedToken, Tokenizer, processors

    tok = Tokenizer.from_file(path)
    if kind == "llama3":
        tok.add_special_tokens([AddedToken(t, special=True, normalized=False) for t in LLAMA3_SPECIALS])
        return PreTrainedTokenizerFast(tokenizer_object=tok, bos_token="<s>", eos_token="</s>", pad_token="<pad>",
                                       chat_template=LLAMA3_CHAT_TEMPLATE)
    if kind == "qwen":
        tok.add_special_tokens([AddedToken(t, special=True, normalized=False) for t in QWEN_SPECIALS])
        tok.post_processor = processors.ByteLevel(trim_offsets=False)
        return PreTrainedTokenizerFast(tokenizer_object=tok, eos_token="<|im_end|>", pad_token="<|endoftext|>",
                                       chat_template=QWEN_CHAT_TEMPLATE)
    raise ValueError(f"unknown tokenizer kind {kind!r}")


def sample_with_token_target(tokenizer, seed: int, n_docs: int, target_tokens: int, build_prompt,
                             sweep_scores: bool = False, tol: float = 0.02):
    """Adjust words-per-doc until the built prompt tokenises to ``target_tokens`` within ``tol``.

    ``build_prompt(question, paras) -> str`` is the prompt builder the caller uses (so chat templates
    count). Returns (question, paras, scores, wrong_answer, n_tokens).
    """
    wpd = max(4, int(target_tokens / (n_docs * 2.6)))
    best = None
    for _ in range(24):
        q, paras, scores, wrong = nq_sample(seed, n_docs, wpd, sweep_scores)
        n = len(tokenizer(build_prompt(q, paras))["input_ids"])
        if best is None or abs(n - target_tokens) < abs(best[4] - target_tokens):
            best = (q, paras, scores, wrong, n)
        if abs(n - target_tokens) <= tol * target_tokens:
            break
        wpd = max(4, int(round(wpd * target_tokens / max(n, 1))))
    return best


if __name__ == "__main__":
    p = train_tokenizer()
    t = load_tokenizer(p)
    q, paras, scores, wrong = nq_sample(0)
    print(p, len(json.load(open(p))["model"]["vocab"]), len(t(" ".join(paras))["input_ids"]), scores, wrong)
