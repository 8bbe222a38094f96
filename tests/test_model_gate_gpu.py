"""GPU: the model-level bf16 gate of SURVEY 8(d) at Llama-3-8B WIDTH -- our engine against HF LlamaForCausalLM (eager
attention) under the reference's mask-rewriting hook, same GPU, same weights (the HF module aliases the engine's storage).

The hook arithmetic is oracle/hf_hooked.py + oracle.cram_oracle.edit_attention_mask (pinned against the unmodified
reference on CPU: tests/test_hf_hooked.py, oracle/make_golden.py). Three forwards of the same 4-layer, hidden-4096,
32q/8kv, vocab-128256 model on one NQ-shaped ~1k-token prompt with a k/10 credibility vector (incl. a 0 document):

    ours_bf16   CramEngine.prefill (tcgen05 attention, in-kernel bias)        -> logits [T, V] bf16
    ref_bf16    HF eager + hook, bf16  (the path north_star names)            -> logits [T, V] bf16
    truth_fp32  HF eager + hook, fp32 weights upcast from the same bf16 values

north_star quotes "max-abs <= 2e-2 on logits, identical greedy tokens". Measured on B200 (round 2, this test): |logit|
reaches 7.25, where ONE bf16 ulp is 3.1e-2, and the REFERENCE's own bf16 forward is 2.6e-2 (mean) / 2.3e-1 (max) away from
its fp32 forward -- so no bf16 implementation, the reference included, can meet 2e-2 max-abs at this width; ours is
2.0e-2 (mean) / 1.6e-1 (max) from the fp32 truth and 2.4e-2 / 2.0e-1 from the reference's bf16 result. The gate therefore
asserts what can be defended, with the reference's own error measured beside ours, and prints the rest:
  (a) our error against the fp32 truth is no larger than the reference's own bf16 error, max and mean (x1.1 slack);
  (b) our distance from the reference's bf16 logits is no larger than the reference's own distance from the fp32 truth
      (x1.25 slack), max and mean;
  (c) greedy tokens: identical on every row whose ref_bf16 top-1/top-2 margin exceeds the largest |ours - ref_bf16| (rows
      where neither bf16 result is decisive are counted and reported); over ALL rows ours agrees with the fp32
      argmax at least as often as the reference's bf16 forward does (93 % of the rows agree between the two bf16 runs).
"""
import numpy as np
import pytest
import torch

from cram_b200 import ops, synth
from cram_b200.engine import CramEngine, ModelSpec
from cram_b200.re_weighting import Re_Weighting_Strategy
from oracle import hf_hooked

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_8b_width_bf16_logits_vs_hf_eager_hooked():
    spec = ModelSpec(4, 32, 8, 128, 4096, 14336, 128256, 1e-5, 500000.0)
    eng = CramEngine.random_init(spec, seed=0, dtype=torch.bfloat16, device=DEV)
    tok = synth.load_tokenizer()
    layers = {0: list(range(0, 32, 2)), 1: [5], 3: list(range(32))}  # half / one / all heads; layer 2 untouched
    rw = Re_Weighting_Strategy("tiny-llama", layers_to_be_modified=layers, tokenizer=tok, engine=eng)
    q, paras, scores, _, n_tok = synth.sample_with_token_target(
        tok, 5, 5, 1024, lambda qq, pp: rw._build_prompt(qq, pp, "")[0])
    scores = [0.0, 0.2, 0.5, 0.9, 1.0]
    inputs, w = rw.decode_with_special_attention(question=q, paras=paras, scores=scores)
    ids = inputs["input_ids"][0]
    T = ids.numel()
    assert 900 <= T <= 1150
    logcred = ops.log_cred(w[0], torch.bfloat16)
    cu = torch.tensor([0, T], dtype=torch.int32, device=DEV)
    masks = [ops.heads_to_mask(layers.get(l, [])) for l in range(spec.num_layers)]
    ours = eng.logits(eng.prefill(ids, cu, T, logcred=logcred, head_masks=masks, attn_kernel=2)).float()

    lc16 = torch.log(w[0])  # what the reference hands to its hook (:122): fp16 log, cast to model dtype inside (:56)
    hf = hf_hooked.hf_from_engine(eng, attn="eager")
    ref = hf_hooked.hooked_logits(hf, ids, lc16, layers)[0].float()
    del hf
    eng32 = CramEngine(spec, {k: v.float() for k, v in eng.w.items()})
    hf32 = hf_hooked.hf_from_engine(eng32, attn="eager")
    assert not torch.backends.cuda.matmul.allow_tf32
    truth = hf_hooked.hooked_logits(hf32, ids, lc16, layers)[0]
    clean = hf_hooked.hooked_logits(hf32, ids, None)[0]
    del hf32, eng32

    scale = truth.abs().max().item()
    ulp = 2.0 ** (np.floor(np.log2(scale)) - 7)  # bf16 spacing at the largest logit
    e_ours, e_ref = (ours - truth).abs().max().item(), (ref - truth).abs().max().item()
    d = (ours - ref).abs()
    print(f"[8B-width gate] T={T} |logit|max={scale:.2f} bf16 ulp there={ulp:.3e}")
    print(f"  ours-vs-fp32 max {e_ours:.3e} mean {(ours - truth).abs().mean().item():.3e}")
    print(f"  HF-eager-bf16-vs-fp32 max {e_ref:.3e} mean {(ref - truth).abs().mean().item():.3e}")
    print(f"  ours-vs-HF-eager-bf16 max {d.max().item():.3e} mean {d.mean().item():.3e}  (north_star quotes 2e-2)")
    print(f"  hook effect on the fp32 logits: max {(truth - clean).abs().max().item():.3e}")
    assert (truth - clean).abs().max().item() > 10 * e_ref  # the bias really moves the logits at this scale
    m_ours, m_ref = (ours - truth).abs().mean().item(), (ref - truth).abs().mean().item()
    assert e_ours <= 1.1 * e_ref and m_ours <= 1.1 * m_ref                    # (a)
    assert d.max().item() <= 1.25 * e_ref and d.mean().item() <= 1.25 * m_ref  # (b)
    top2 = ref.topk(2, dim=-1).values
    sure = (top2[:, 0] - top2[:, 1]) > d.max().item()
    same = ours.argmax(-1) == ref.argmax(-1)
    same32 = ours.argmax(-1) == truth.argmax(-1)
    ref32 = ref.argmax(-1) == truth.argmax(-1)
    print(f"  greedy: {int(same.sum())}/{T} rows identical to HF bf16; {int(sure.sum())} rows with a decisive margin, "
          f"{int((~same & sure).sum())} of those differ; last row identical: {bool(same[-1])}; "
          f"agreement with the fp32 argmax: ours {int(same32.sum())}, HF bf16 {int(ref32.sum())}")
    assert bool(same[sure].all())                            # (c)
    # near-ties flip under ANY change of rounding order: the reference's own bf16 argmax leaves the fp32 argmax on a few
    # per cent of the rows. Ours must agree with the fp32 truth at least as often as the reference's bf16 does.
    assert same32.float().mean().item() >= ref32.float().mean().item() - 0.01
    assert same.float().mean().item() >= 0.90
