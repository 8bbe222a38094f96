"""GPU: the whole path on the tiny model of BASELINE configs[0] vs the golden vectors the UNMODIFIED
reference produced (tests/golden/tiny_model.npz) and vs the oracle in bf16."""
import json
import os

import numpy as np
import pytest
import torch

from cram_b200 import ops, synth
from cram_b200.engine import CramEngine, ModelSpec
from cram_b200.re_weighting import Find_Best_Heads, Re_Weighting_Strategy
from oracle import cram_oracle as orc
from oracle.tiny_model import TINY_SPEC, tiny_state_dict

pytestmark = pytest.mark.gpu
DEV = "cuda"
HEAD_SETS = {"none": None, "l0h1": {0: [1]}, "l1h03": {1: [0, 3]}, "all": {0: [0, 1, 2, 3], 1: [0, 1, 2, 3]},
             "strkeys": {"1": [2], "0": [0]}}


def _spec():
    t = TINY_SPEC
    return ModelSpec(t.num_layers, t.num_heads, t.num_kv_heads, t.head_dim, t.hidden, t.intermediate, t.vocab, t.rms_eps,
                     t.rope_theta)


@pytest.fixture(scope="module")
def gold(golden_dir):
    g = np.load(os.path.join(golden_dir, "tiny_model.npz"))
    meta = json.load(open(os.path.join(golden_dir, "tiny_model_meta.json")))
    return g, meta


@pytest.fixture(scope="module")
def engine32():
    return CramEngine.from_state_dict(_spec(), tiny_state_dict(torch.float32), device=DEV)


def _masks(hs, L=2):
    m = [0] * L
    for l, heads in (hs or {}).items():
        m[int(l)] |= ops.heads_to_mask(heads)
    return m


def test_fp32_logits_match_reference_golden(gold, engine32):
    g, meta = gold
    ids = torch.from_numpy(g["input_ids"]).to(DEV)
    T = ids.numel()
    cu = torch.tensor([0, T], dtype=torch.int32, device=DEV)
    cred = torch.from_numpy(g["cred_bits"].view(np.float16).copy()).to(DEV)
    logcred = ops.log_cred(cred, torch.float32)
    rows = torch.from_numpy(g["logit_rows"]).to(DEV)
    for name, hs in HEAD_SETS.items():
        x = engine32.prefill(ids, cu, T, logcred=logcred, head_masks=_masks(hs))
        lg = engine32.logits(x[rows]).cpu().numpy()
        err = np.abs(lg - g[f"logits_{name}"]).max()
        print(f"[tiny fp32 logits {name}] max_abs={err:.3e} (|logit| max {np.abs(lg).max():.2f})")
        assert err <= 5e-4, name
    # the hook must actually move the logits (guards against a silently ignored bias)
    assert np.abs(g["logits_all"] - g["logits_none"]).max() > 1e-2


def test_drop_in_api_span_generate_and_ie(gold, engine32):
    g, meta = gold
    tok = synth.load_tokenizer()
    for name, hs in HEAD_SETS.items():
        rw = Re_Weighting_Strategy("tiny-llama", layers_to_be_modified=dict(hs) if hs else dict(), tokenizer=tok,
                                   engine=engine32)
        if name == "none":
            rw.layers_to_be_modified = {}
        inputs, w = rw.decode_with_special_attention(question=meta["question"], paras=meta["paras"], scores=meta["scores"])
        assert np.array_equal(inputs["input_ids"][0].cpu().numpy(), g["input_ids"])
        assert w.dtype == torch.float16 and tuple(w.shape) == (1, len(g["input_ids"]))
        assert np.array_equal(w[0].cpu().view(torch.int16).numpy().view(np.uint16), g["cred_bits"])  # bit-exact
        prompt, out = rw.run_RAG_with_attention_weighting(question=meta["question"], paras=meta["paras"],
                                                          scores=meta["scores"])
        assert out == str(g[f"gen_text_{name}"]), (name, out, str(g[f"gen_text_{name}"]))
        # greedy tokens (first 8) against the reference's own generate
        masks = _masks(rw.layers_to_be_modified)
        new = engine32.generate([inputs["input_ids"][0]], [ops.log_cred(w[0], torch.float32)], masks, max_new_tokens=8,
                                eos_token_ids=[tok.eos_token_id])[0]
        assert new == g[f"greedy_{name}"].tolist()[:len(new)], name
    # a passage that is not in the prompt verbatim (e.g. a chat template that rewrites whitespace) makes the
    # reference fail with TypeError at re_weighting.py:111; the kernel reports it and the class raises the same
    rw = Re_Weighting_Strategy("tiny-llama", tokenizer=tok, engine=engine32)
    rw._build_prompt = lambda q_, p_, a_: ("Given the following information: \nPassage-0: something else\nAnswer: ", True)
    with pytest.raises(TypeError):
        rw.decode_with_special_attention(question="q?", paras=["abc def"], scores=[0.5])

    fb = Find_Best_Heads("tiny-llama", tokenizer=tok, engine=engine32)
    ie = np.array(fb.cal_logits(question=meta["question"], paras=meta["paras"], scores=meta["sweep_scores"],
                                wrong_answer=meta["wrong"]))
    print("[tiny fp32 IE]\n", ie, "\n", g["ie_ref"])
    assert ie.shape == g["ie_ref"].shape
    assert np.abs(ie - g["ie_ref"]).max() <= 1e-3
    assert abs(float(fb.ori_prob_sum) - float(g["ie_ori_logit_sum"])) <= 1e-3
    # ranking of the heads by IE is the reference's
    assert np.argsort(-ie.flatten(), kind="stable").tolist() == np.argsort(-g["ie_ref"].flatten(), kind="stable").tolist()


def test_bf16_tracks_oracle_bf16_within_bf16_noise(gold):
    """bf16 engine vs the oracle's bf16 hooked-eager forward: our error against the fp32 truth must be
    no worse than the reference arithmetic's own bf16 error (x2 margin)."""
    g, meta = gold
    ids_np = g["input_ids"][:320]
    sd16 = tiny_state_dict(torch.bfloat16)
    eng = CramEngine.from_state_dict(_spec(), sd16, device=DEV)
    ids = torch.from_numpy(ids_np).to(DEV)
    T = ids.numel()
    cu = torch.tensor([0, T], dtype=torch.int32, device=DEV)
    cred16 = torch.from_numpy(g["cred_bits"].view(np.float16)[:320].copy())
    lc_bf = orc.log_chain(cred16.numpy(), torch.bfloat16)
    hooks = {0: (lc_bf[None], [1, 2]), 1: (lc_bf[None], [0, 3])}
    with torch.no_grad():
        want16 = orc.llama_forward(sd16, TINY_SPEC, torch.from_numpy(ids_np)[None], hooks)[0].float()
        sd32 = {k: v.float() for k, v in sd16.items()}
        truth = orc.llama_forward(sd32, TINY_SPEC, torch.from_numpy(ids_np)[None],
                                  {l: (lc.float(), h) for l, (lc, h) in hooks.items()})[0]
    for kern in (1, 2):
        x = eng.prefill(ids, cu, T, logcred=lc_bf.to(DEV), head_masks=[0b0110, 0b1001], attn_kernel=kern)
        got = eng.logits(x).float().cpu()
        e_ours, e_ref = (got - truth).abs().max().item(), (want16 - truth).abs().max().item()
        d = (got - want16).abs().max().item()
        print(f"[tiny bf16 kernel={kern}] ours-vs-fp32 {e_ours:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}  ours-vs-oracle-bf16 {d:.3e}")
        assert e_ours <= 2 * e_ref + 1e-2
        agree = (got.argmax(-1) == want16.argmax(-1)).float().mean().item()
        assert agree >= 0.97


def test_graph_decode_equals_eager_and_oracle_cached_decode():
    """Batch of ragged prompts, per-sample credibility, Qwen2-style QKV bias: the CUDA-graph decode session
    produces the tokens of the eager loop, and (fp32) the tokens of the oracle's cached hooked decode."""
    from oracle.tiny_model import random_state_dict
    ospec = orc.ModelSpec(2, 8, 2, 64, 256, 384, 300, rms_eps=1e-6, rope_theta=10000.0, qkv_bias=True)
    sd = random_state_dict(ospec, seed=11, dtype=torch.float32)
    spec = ModelSpec(2, 8, 2, 64, 256, 384, 300, 1e-6, 10000.0, qkv_bias=True)
    eng = CramEngine.from_state_dict(spec, sd, device=DEV)
    g = torch.Generator().manual_seed(2)
    lens = [37, 90, 64]
    prompts = [torch.randint(3, 300, (n,), generator=g) for n in lens]
    creds = []
    for n in lens:
        c = torch.ones(n, dtype=torch.float16)
        c[4:n // 2] = 0.3
        c[n // 2:n - 6] = 0.0
        creds.append(c)
    masks = [0b10110101, 0b00001111]
    hooks_heads = {0: [0, 2, 4, 5, 7], 1: [0, 1, 2, 3]}
    lcs = [ops.log_cred(c.to(DEV), torch.float32) for c in creds]
    dev_prompts = [p.to(DEV) for p in prompts]
    new = 12
    got_graph = eng.generate(dev_prompts, lcs, masks, max_new_tokens=new, use_graph=True)
    got_eager = eng.generate(dev_prompts, lcs, masks, max_new_tokens=new, use_graph=False)
    assert got_graph == got_eager
    for b, n in enumerate(lens):
        lc = torch.log(creds[b]).float()[None]
        hooks = {l: (lc, h) for l, h in hooks_heads.items()}
        kc, vc = [None] * 2, [None] * 2
        with torch.no_grad():
            for t in range(n):
                hk = {l: (lc[:, :t + 1], h) for l, (lc_, h) in hooks.items()}
                logits = orc.decode_step_cached(sd, ospec, prompts[b][t:t + 1], t, kc, vc, hk)
            want = []
            for t in range(new):
                nxt = int(logits[0].argmax())
                want.append(nxt)
                logits = orc.decode_step_cached(sd, ospec, torch.tensor([nxt]), n + t, kc, vc, hooks)
        assert got_graph[b] == want, (b, got_graph[b], want)


def test_run_rag_batch_equals_reference_loop_at_b1(engine32):
    """Extension of SURVEY 8b: per-sample credibility in one packed batch == the reference's b = 1 loop."""
    tok = synth.load_tokenizer()
    rw = Re_Weighting_Strategy("tiny-llama", layers_to_be_modified={0: [1, 2], 1: [0, 3]}, tokenizer=tok, engine=engine32)
    samples = [synth.nq_sample(seed, 3 + seed % 3, 20 + 7 * seed)[:3] for seed in range(4)]
    qs, ps, ss = [s[0] for s in samples], [s[1] for s in samples], [s[2] for s in samples]
    batched = rw.run_RAG_batch(qs, ps, ss, max_new_tokens=10)
    for q, p, s, got in zip(qs, ps, ss, batched):
        assert got == rw.run_RAG_batch([q], [p], [s], max_new_tokens=10)[0]
    # and the public single-sample call is that same path (utils/re_weighting.py:115-138)
    one = rw.run_RAG_with_attention_weighting(question=qs[0], paras=ps[0], scores=ss[0])
    assert one[0] == batched[0][0] and one[1].startswith(batched[0][1])


def test_streaming_mlp_decode_step_matches_the_library_gemm_path():
    """bf16 decode step with the MLP on cram_linear_decode (SwiGLU and residual-add epilogues) vs the same step on
    library GEMMs + glue kernels: same roundings, different fp32 accumulation order -> logits agree to bf16 noise
    and the greedy continuation is the same wherever the top-2 margin exceeds that noise."""
    from cram_b200.engine import PagedKV
    spec = ModelSpec(3, 8, 2, 64, 512, 1408, 500, 1e-6, 10000.0, qkv_bias=True)
    eng = CramEngine.random_init(spec, seed=3, dtype=torch.bfloat16, device=DEV, std=0.05)
    B, ps, ctx = 5, 16, 70
    n_pages = B * 8
    g = torch.Generator(device=DEV).manual_seed(1)
    out = {}
    for streaming in (False, True):
        eng.streaming_mlp = streaming
        kv = PagedKV(spec, n_pages, ps, torch.bfloat16, DEV)
        for l in range(spec.num_layers):
            kv.k[l].copy_(torch.randn(kv.k[l].shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(10 + l)))
            kv.v[l].copy_(torch.randn(kv.v[l].shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(20 + l)))
        bt = torch.arange(n_pages, device=DEV, dtype=torch.int32).view(B, 8)
        tokens = torch.tensor([5, 17, 250, 499, 42], device=DEV)
        positions = torch.full((B,), ctx, device=DEV, dtype=torch.int32)
        out[streaming] = eng.decode_step(tokens, positions, kv, bt, positions + 1, ctx + 1).float()
    a, b = out[False], out[True]
    assert torch.isfinite(b).all()
    scale = a.abs().max().item()
    assert (a - b).abs().max().item() <= 0.03 * scale + 1e-3
    top2 = a.topk(2, dim=-1).values
    sure = (top2[:, 0] - top2[:, 1]) > 0.06 * scale
    assert bool((a.argmax(-1) == b.argmax(-1))[sure].all())
