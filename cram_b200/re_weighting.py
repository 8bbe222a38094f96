"""Drop-in replacements for the reference's ``Re_Weighting_Strategy`` / ``Find_Best_Heads``
(utils/re_weighting.py:22-175): same constructor arguments, attributes, method names, argument
meaning and return values, with the hooked-HF-eager internals replaced by the B200 engine.

What changes underneath
  * decode_with_special_attention: the T x D Python scan over a CUDA offset tensor (:100-111) is one
    launch of kernel (2) (span index + fp16 credibility vector, bit-exact);
  * run_RAG_with_attention_weighting: no forward_pre_hook and no dense [b,H,q,k] mask -- the engine's
    attention kernels take log(credibility) per key and a head bitmask per layer;
  * cal_logits: 1 + L*H full forwards become one cached clean forward + L batched suffix forwards.
There is no CPU path: constructing either class without a CUDA device raises.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from .engine import CramEngine, ModelSpec
from .prompt import get_prompt
from .sweep import IESweep

# kept for import compatibility with the reference module (utils/re_weighting.py:13-16); unused there too
RAG_prompt1 = """Given the following information: \n"""
RAG_prompt2 = """Answer the following question based on the given information or your internal knowledge with one or few words without the source.
Question: {question}
Answer: {answer}"""


def _heads_dict_to_masks(layers_to_be_modified: dict, num_layers: int) -> List[int]:
    masks = [0] * num_layers
    for layer, heads in layers_to_be_modified.items():
        li = int(layer)  # keys may be JSON strings (RAG_test.py:50-58)
        if not 0 <= li < num_layers:
            raise AttributeError(f"model.layers.{layer}.self_attn does not exist")  # get_submodule's error class
        masks[li] |= ops.heads_to_mask(heads)
    return masks


class Re_Weighting_Strategy:

    def __init__(self, model_name: str = "Llama-2-13b-chat-hf", layers_to_be_modified: dict = dict(), bad_words_ids=[],
                 *, tokenizer=None, model=None, engine: Optional[CramEngine] = None, device: str = "cuda",
                 generation_config=None):
        """Reference signature (utils/re_weighting.py:24). The keyword-only extras let a caller hand in an
        already-loaded tokenizer / HF model / engine (the reference always loads by name, which needs the hub).
        ``generation_config`` defaults to ``model.generation_config``: model.generate(do_sample=False) at :132 still
        applies its eos_token_id list and repetition_penalty (the Qwen chat checkpoints ship both)."""
        if not torch.cuda.is_available():
            raise ops.CramError("Re_Weighting_Strategy needs a CUDA device: cram_b200 has no CPU fallback")
        if tokenizer is None:
            from transformers import AutoTokenizer
            tokenizer = AutoTokenizer.from_pretrained(model_name)
        self.tokenizer = tokenizer
        if engine is None:
            if model is None:
                from transformers import AutoModelForCausalLM
                model = AutoModelForCausalLM.from_pretrained(model_name, torch_dtype="auto")
            engine = CramEngine.from_hf_model(model, device=device)
        self.model = model  # HF module (config / weights source); the forward runs in self.engine
        self.generation_config = generation_config if generation_config is not None else getattr(model, "generation_config", None)
        self.engine = engine
        self.model_name = model_name
        self.bad_words_ids = bad_words_ids
        self.num_hidden_layers = engine.spec.num_layers
        self.model_num_attention_heads = engine.spec.num_heads
        if not layers_to_be_modified:
            layers_to_be_modified = {i: list(range(self.model_num_attention_heads)) for i in range(self.num_hidden_layers)}
        self.layers_to_be_modified = layers_to_be_modified

    # ---- compatibility shim ---------------------------------------------------------------------------
    def edit_attention_mask(self, module, input_args: tuple, input_kwargs: dict, attention_weight, head_idx: list = []):
        """The reference's forward_pre_hook (utils/re_weighting.py:36-61), kept so existing callers that
        register it on a HF module still work. The engine never calls it: it materialises the dense
        [b,H,q,k] mask the kernels exist to avoid."""
        weight_len = attention_weight.size()[-1]
        hs = input_kwargs["hidden_states"]
        dtype, device = hs.dtype, hs.device
        H = self.model_num_attention_heads
        if input_kwargs["attention_mask"] is None:
            tgt_len = hs.size()[1]
            src_len = int(input_kwargs["position_ids"][0][-1]) + 1
            if tgt_len == 1:
                mask = torch.zeros([1, 1, tgt_len, src_len], dtype=dtype, device=device)
            else:
                m = torch.triu(torch.full((tgt_len, src_len), torch.finfo(dtype).min, dtype=dtype, device=device), diagonal=1)
                mask = m[None, None]
        else:
            mask = input_kwargs["attention_mask"].clone()
        if mask.size(1) == 1:
            mask = mask.repeat(1, H, 1, 1)
        w = attention_weight.reshape(1, 1, 1, weight_len).to(dtype=dtype, device=device)
        sel = torch.zeros(H, dtype=torch.bool, device=device)
        if len(head_idx):
            sel[torch.as_tensor(list(head_idx), device=device)] = True
        sub = mask[..., :weight_len]
        allowed = (sub == 0.0) & sel.view(1, H, 1, 1)
        mask[..., :weight_len] = torch.where(allowed, w.expand_as(sub), sub)
        input_kwargs["attention_mask"] = mask
        return input_args, input_kwargs

    # ---- prompt + span index --------------------------------------------------------------------------
    def _build_prompt(self, question: str, paras: Sequence[str], answer: str) -> Tuple[str, bool]:
        """utils/re_weighting.py:64-97."""
        add_special_tokens = True
        if self.model_name.find("Llama-3") != -1:
            prompt = get_prompt(context=paras, question=question, answer="", type="with_contexts")
            messages = [{"role": "system", "content": "You are a helpful assistant!"},
                        {"role": "user", "content": f"{prompt}"}]
            prompt = self.tokenizer.apply_chat_template(messages, add_generation_prompt=True, tokenize=False)
            if answer != "":
                prompt += answer
        elif self.model_name.find("Qwen") != -1:
            prompt = get_prompt(context=paras, question=question, answer="", type="with_contexts")
            messages = [{"role": "system", "content": "You are a helpful assistant."},
                        {"role": "user", "content": f"{prompt}"}]
            prompt = self.tokenizer.apply_chat_template(messages, add_generation_prompt=True, tokenize=False)
            add_special_tokens = False
            if answer != "":
                prompt += answer
        else:
            prompt = get_prompt(context=paras, question=question, answer=answer, type="with_contexts")
        return prompt, add_special_tokens

    def _encode_batch(self, questions, paras_list, scores_list, answers):
        """Host side of decode_with_special_attention for a batch: prompts, token ids, offsets, char spans."""
        prompts, flags = [], []
        for q, ps, a in zip(questions, paras_list, answers):
            pr, f = self._build_prompt(q, ps, a)
            prompts.append(pr)
            flags.append(f)
        ids, offs, sb, se, sc, cu, cud = [], [], [], [], [], [0], [0]
        # one tokenizer call for the whole batch (the fast tokenizer encodes the prompts in parallel); no padding, so
        # every prompt gets exactly the ids / offsets of the reference's b = 1 call (:98)
        enc_all = self.tokenizer(prompts, return_offsets_mapping=True, add_special_tokens=flags[0]) if prompts else None
        for b, (pr, f, ps, ss) in enumerate(zip(prompts, flags, paras_list, scores_list)):
            ids.append(np.asarray(enc_all["input_ids"][b], dtype=np.int64))
            offs.append(np.asarray(enc_all["offset_mapping"][b], dtype=np.int64).reshape(-1, 2)[:, 0].copy())
            cu.append(cu[-1] + len(ids[-1]))
            for i, p in enumerate(ps):
                para = ("Passage-%d: " % i) + p + "\n"
                s = pr.find(para)
                sb.append(s)
                se.append(s + len(para) - 1)
                # torch.full((1, n), score): python int -> int64, float -> float32 (utils/re_weighting.py:111)
                sc.append(np.float32(ss[i]))
            cud.append(cud[-1] + len(ps))
        return prompts, ids, offs, sb, se, sc, cu, cud

    def span_index_batch(self, questions, paras_list, scores_list, answers=None):
        """Batched kernel-(2) front end. Returns a dict with packed device tensors: input_ids, cu_seqlens,
        doc_id, cred (fp16), logcred (model dtype), lens (host list) and prompts (host strings)."""
        dev = self.engine.device
        answers = answers if answers is not None else [""] * len(questions)
        prompts, ids, offs, sb, se, sc, cu, cud = self._encode_batch(questions, paras_list, scores_list, answers)
        host = lambda a, dt: torch.from_numpy(np.asarray(a, dtype=dt))
        # one pinned staging buffer per array, async copies on the current stream
        def up(t):
            return t.pin_memory().to(dev, non_blocking=True)
        ids_d = up(host(np.concatenate(ids), np.int64))
        off_d = up(host(np.concatenate(offs), np.int64))
        cu_d, cud_d = up(host(cu, np.int32)), up(host(cud, np.int32))
        n_docs = max(len(sb), 1)
        sb_d = up(host(sb if sb else [0], np.int32))
        se_d = up(host(se if se else [0], np.int32))
        sc_d = up(host(sc if sc else [0.0], np.float32))
        doc_id, cred, logcred, err = ops.span_index(off_d, cu_d, sb_d, se_d, sc_d, cud_d, logcred_dtype=self.engine.dtype)
        return dict(prompts=prompts, input_ids=ids_d, cu_seqlens=cu_d, doc_id=doc_id, cred=cred, logcred=logcred,
                    err=err, lens=[len(x) for x in ids], n_docs=n_docs)

    def decode_with_special_attention(self, question: str = "", paras: list = [], scores: list = [], answer: str = ""):
        """utils/re_weighting.py:63-113. Returns (model_inputs, attention_weight fp16 [1, T]) on the GPU."""
        r = self.span_index_batch([question], [paras], [scores], [answer])
        self._raise_on_span_errors([int(r["err"][0].item())])
        T = r["lens"][0]
        model_inputs = {"input_ids": r["input_ids"].view(1, T),
                        "attention_mask": torch.ones(1, T, dtype=torch.int64, device=r["input_ids"].device)}
        self._last_prompt_text = r["prompts"][0]
        return model_inputs, r["cred"].view(1, T)

    # ---- generation -----------------------------------------------------------------------------------
    def _eos_ids(self) -> List[int]:
        """Stop tokens of model.generate at utils/re_weighting.py:132. Llama-3: the explicit pair of :130-131, which
        overrides the checkpoint's own list. Every other model: generation_config.eos_token_id (int or list) -- the
        tokenizer's eos is only the fallback for an engine built without a HF model."""
        if self.model_name.find("Llama-3") != -1:
            return [self.tokenizer.eos_token_id, self.tokenizer.convert_tokens_to_ids("<|eot_id|>")]
        gc_eos = getattr(self.generation_config, "eos_token_id", None)
        if gc_eos is not None:
            return [int(e) for e in (gc_eos if isinstance(gc_eos, (list, tuple)) else [gc_eos])]
        return [self.tokenizer.eos_token_id] if self.tokenizer.eos_token_id is not None else []

    def _repetition_penalty(self) -> float:
        rp = getattr(self.generation_config, "repetition_penalty", None)
        return float(rp) if rp is not None else 1.0

    @staticmethod
    def _raise_on_span_errors(errs) -> None:
        """err bits of cram_span_index (include/cram_b200.h) -> the reference's exception classes."""
        if any(e & 1 for e in errs):
            # a passage was not found verbatim in the prompt: the reference fails at :111 with this error
            raise TypeError("unsupported operand type(s) for +: 'NoneType' and 'int'")
        if any(e & 2 for e in errs):
            raise RuntimeError("Trying to create tensor with negative dimension")
        if any(e & 4 for e in errs):
            raise ops.CramError("more passages in one prompt than cram_span_index supports; the tail would be un-weighted")

    def _bad_tokens(self) -> List[int]:
        out = []
        for w in self.bad_words_ids or []:
            if len(w) != 1:
                raise ops.CramError("only single-token bad_words_ids are supported (the reference uses [[13]])")
            out.append(int(w[0]))
        return out

    @torch.no_grad()
    def run_RAG_with_attention_weighting(self, question: str = "", paras: list = [], scores: list = []):
        """utils/re_weighting.py:115-138 -> (prompt, output)."""
        return self.run_RAG_batch([question], [paras], [scores])[0]

    @torch.no_grad()
    def run_RAG_batch(self, questions, paras_list, scores_list, max_new_tokens: int = 100):
        """Extension: per-sample credibility for a batch; equals the reference looped at b=1."""
        r = self.span_index_batch(questions, paras_list, scores_list)
        self._raise_on_span_errors(r["err"].tolist())
        masks = _heads_dict_to_masks(self.layers_to_be_modified, self.num_hidden_layers)
        cu = np.cumsum([0] + r["lens"])
        prompts_ids = [r["input_ids"][cu[b]:cu[b + 1]] for b in range(len(questions))]
        logcreds = [r["logcred"][cu[b]:cu[b + 1]] for b in range(len(questions))]
        new = self.engine.generate(prompts_ids, logcreds, masks, max_new_tokens=max_new_tokens,
                                   eos_token_ids=self._eos_ids(), bad_token_ids=self._bad_tokens(),
                                   repetition_penalty=self._repetition_penalty())
        results = []
        for b in range(len(questions)):
            full = prompts_ids[b].tolist() + new[b]
            prompt = self.tokenizer.decode(full[1:len(prompts_ids[b])])
            output = self.tokenizer.decode(full[1:-1])          # :133 drops the first and the last token
            output = output[output.find(prompt) + len(prompt):]  # :134-135
            results.append((prompt, output))
        return results


class Find_Best_Heads(Re_Weighting_Strategy):

    def __init__(self, model_name: str = "Llama-2-13b-chat-hf", layers_to_be_modified: list = [], *, rank: int = 0,
                 world_size: int = 1, group=None, **kw):
        """utils/re_weighting.py:143-150 (layers_to_be_modified is accepted and dropped, as there)."""
        super().__init__(model_name=model_name, **kw)
        torch.manual_seed(42)
        torch.cuda.manual_seed_all(42)
        self.model_num_hidden_layers = self.engine.spec.num_layers
        self.model_num_attention_heads = self.engine.spec.num_heads
        self.hidden_size = self.engine.spec.hidden
        self.sweep = IESweep(self.engine, rank=rank, world_size=world_size, group=group)

    @torch.no_grad()
    def cal_logits(self, question: str = "", paras: list = [], scores: list = [], right_answer: str = "",
                   wrong_answer: str = "my name is"):
        """utils/re_weighting.py:152-175 -> list[L][H] of IE = clean - single-head-hooked answer logit sum.
        Quirks kept: answer ids drop their first token (:155) and are scored at their own positions (:160)."""
        answer = wrong_answer
        answer_ids = torch.as_tensor(self.tokenizer([answer])["input_ids"][0][1:], dtype=torch.int64)
        model_inputs, attention_weight = self.decode_with_special_attention(question=question, paras=paras, scores=scores,
                                                                            answer=answer)
        ids = model_inputs["input_ids"][0]
        logcred = ops.log_cred(attention_weight[0], self.engine.dtype)  # torch.log(fp16) then .to(dtype), :166/:56
        ie = self.sweep.cal_ie(ids, logcred, answer_ids.to(ids.device))
        self.ori_prob_sum = self.sweep.last_ori[0]
        self.ori_logits = None  # the full [1,T,V] logits are never materialised here
        return ie.cpu().tolist()
