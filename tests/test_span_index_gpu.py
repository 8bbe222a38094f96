"""GPU: kernel (2) through the C ABI vs the golden vectors of the reference and the oracle."""
import os

import numpy as np
import pytest
import torch

from cram_b200 import ops
from oracle import cram_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(offs, spans_list, scores_list, dtype=torch.bfloat16):
    cu = np.cumsum([0] + [len(o) for o in offs]).astype(np.int32)
    cud = np.cumsum([0] + [len(s) for s in spans_list]).astype(np.int32)
    sb = np.array([s[0] for sp in spans_list for s in sp] or [0], dtype=np.int32)
    se = np.array([s[1] for sp in spans_list for s in sp] or [0], dtype=np.int32)
    sc = np.array([x for ss in scores_list for x in ss] or [0], dtype=np.float32)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    off_all = np.concatenate(offs) if offs else np.zeros(0, dtype=np.int64)
    doc_id, cred, logcred, err = ops.span_index(t(off_all.astype(np.int64)), t(cu), t(sb), t(se), t(sc), t(cud), dtype)
    return doc_id.cpu().numpy(), cred.cpu(), logcred.cpu(), err.cpu().numpy(), cu


def test_golden_cases_batched(golden_dir):
    g = np.load(os.path.join(golden_dir, "span_index.npz"))
    n = int(g["n_cases"])
    offs = [g[f"{i}_offset_starts"] for i in range(n)]
    spans = [[tuple(s) for s in g[f"{i}_spans"]] for i in range(n)]
    scores = [g[f"{i}_score_bits"].view(np.float16).astype(np.float32).tolist() for i in range(n)]
    doc_id, cred, logcred, err, cu = _run(offs, spans, scores)  # ONE launch for all 220 prompts
    assert (err == 0).all()
    for i in range(n):
        a, b = cu[i], cu[i + 1]
        assert np.array_equal(doc_id[a:b], g[f"{i}_doc_id"]), i
        cb = cred[a:b].view(torch.int16).numpy().view(np.uint16)
        assert np.array_equal(cb, g[f"{i}_cred_bits"]), i
        want = orc.log_chain(g[f"{i}_cred_bits"].view(np.float16), torch.bfloat16)
        assert torch.equal(logcred[a:b].view(torch.int16), want.view(torch.int16)), i


def test_log_chain_all_fp16_inputs(golden_dir):
    g = np.load(os.path.join(golden_dir, "log_chain.npz"))
    bits = np.arange(65536, dtype=np.uint16)
    x = bits.view(np.float16)
    ok = ~(np.isnan(x) | (x < 0))
    cred = torch.from_numpy(x.copy()).to(DEV)
    l16 = ops.log_cred(cred, torch.float16).cpu().view(torch.int16).numpy().view(np.uint16)
    lbf = ops.log_cred(cred, torch.bfloat16).cpu().view(torch.int16).numpy().view(np.uint16)
    l32 = ops.log_cred(cred, torch.float32).cpu().numpy()
    assert np.array_equal(l16[ok], g["log_fp16_bits"][ok])
    assert np.array_equal(lbf[ok], g["log_bf16_bits"][ok])
    assert np.array_equal(l32[ok], g["log_fp16_bits"].view(np.float16).astype(np.float32)[ok])
    assert np.isnan(l32[~ok & ~np.isnan(x) & (x < 0)]).all()
    # torch's own CUDA log of the same fp16 vector (the reference normally runs on a GPU). torch-CPU and
    # torch-CUDA are two different logf implementations and disagree on a few of the 65 536 inputs in the
    # last fp16 bit; the golden vectors (and this kernel) follow torch-CPU, where the oracle was pinned.
    # Every credibility value the reference's pipeline can produce (k/10, RAG_test.py:92-103) agrees in all three.
    tl = torch.log(cred).cpu().view(torch.int16).numpy().view(np.uint16)
    mism = np.nonzero(ok & (tl != l16))[0]
    print(f"[log chain] torch-CUDA vs kernel/torch-CPU: {mism.size} of {int(ok.sum())} inputs differ", [hex(int(b)) for b in mism[:8]])
    assert mism.size <= 0.01 * ok.sum()
    for kk in range(11):
        b = int(np.array([np.float32(kk / 10)], dtype=np.float32).astype(np.float16).view(np.uint16)[0])
        assert tl[b] == l16[b] == g["log_fp16_bits"][b], kk


def test_edge_cases_and_error_flags():
    # empty batch entry, overlapping documents, tokens sharing an offset, BOS at (0,0), not-found passage
    offs = [np.array([0, 0, 5, 9, 9, 14], dtype=np.int64), np.zeros(0, dtype=np.int64), np.array([0, 3, 4], dtype=np.int64),
            np.array([0, 2, 4, 6], dtype=np.int64)]
    spans = [[(0, 9), (9, 14)], [], [(-1, 2)], [(4, 1)]]
    scores = [[0.1, 0.2], [], [0.3], [0.5]]
    doc_id, cred, logcred, err, cu = _run(offs, spans, scores, torch.float16)
    assert doc_id[cu[0]:cu[1]].tolist() == [-1, 0, 0, 0, 1, 1]
    want, wcred = orc.span_index(offs[0], spans[0], scores[0])
    assert np.array_equal(cred[cu[0]:cu[1]].view(torch.int16).numpy().view(np.uint16), wcred.view(np.uint16))
    assert err.tolist()[0] == 0 and err.tolist()[1] == 0
    assert err.tolist()[2] & 1  # the reference raises TypeError for this prompt
    assert err.tolist()[3] & 2  # negative span: the reference raises in torch.full
    assert (doc_id[cu[3]:cu[4]] == -1).all()


def test_random_against_oracle():
    rng = np.random.default_rng(3)
    offs, spans, scores = [], [], []
    for c in range(64):
        T = int(rng.integers(1, 3000))
        step = rng.integers(0, 6, size=T)
        step[0] = 0
        off = np.cumsum(step).astype(np.int64)
        off[rng.random(T) < 0.02] = 0  # special tokens with (0,0) offsets anywhere
        D = int(rng.integers(0, 21))
        sp = []
        for _ in range(D):
            a = int(rng.integers(0, off.max() + 1))
            b = int(rng.integers(a, off.max() + 3))
            sp.append((a, b))
        offs.append(off)
        spans.append(sp)
        scores.append([float(np.float32(rng.integers(0, 11) / 10)) for _ in range(D)])
    doc_id, cred, logcred, err, cu = _run(offs, spans, scores)
    for i in range(64):
        wd, wc = orc.span_index(offs[i], spans[i], scores[i])
        assert np.array_equal(doc_id[cu[i]:cu[i + 1]], wd), i
        assert np.array_equal(cred[cu[i]:cu[i + 1]].view(torch.int16).numpy().view(np.uint16), wc.view(np.uint16)), i
    assert (err == 0).all()


def test_full_size_properties():
    """BASELINE cfg5 size (T = 32768, 20 docs): idempotence + monotone coverage, no oracle needed."""
    T, D = 32768, 20
    off = (np.arange(T, dtype=np.int64) * 4)
    off[0] = 0
    bounds = np.linspace(40, 4 * T - 40, D + 1).astype(np.int64)
    spans = [(int(bounds[i]), int(bounds[i + 1] - 2)) for i in range(D)]
    scores = [i / (D - 1) for i in range(D)]
    a = _run([off], [spans], [scores])
    b = _run([off], [spans], [scores])
    assert np.array_equal(a[0], b[0]) and torch.equal(a[1], b[1])
    d = a[0]
    inside = d[d >= 0]
    assert (np.diff(inside) >= 0).all() and set(inside.tolist()) == set(range(D))
    wd, wc = orc.span_index(off, spans, scores)
    assert np.array_equal(d, wd)
