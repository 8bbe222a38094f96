"""GPU: kernel (1) prefill through the C ABI -- generic and tcgen05 kernels vs the oracle and vs each other."""
import numpy as np
import pytest
import torch

from cram_b200 import ops
from oracle import cram_oracle as orc
from tests.helpers import cu_of, make_logcred, make_qkv, ref_attention, report

pytestmark = pytest.mark.gpu
DEV = "cuda"
GENERIC, SM100 = 1, 2


def _oracle_hooked_eager(q, k, v, T, H, Hkv, dh, logcred, heads):
    """The reference arithmetic itself: edit_attention_mask + HF eager attention (oracle, CPU)."""
    dt = q.dtype
    qc = q.cpu().view(1, T, H, dh).transpose(1, 2)
    kc = k.cpu().view(1, T, Hkv, dh).transpose(1, 2)
    vc = v.cpu().view(1, T, Hkv, dh).transpose(1, 2)
    base = torch.triu(torch.full((T, T), torch.finfo(dt).min, dtype=dt), diagonal=1)[None, None]
    mask = orc.edit_attention_mask(base, logcred.cpu().view(1, T), heads, H, T, T, dt)
    return orc.eager_attention(qc, kc, vc, mask, dh ** -0.5).reshape(T, H * dh)


@pytest.mark.parametrize("dtype,dh,tol", [(torch.float32, 64, 2e-5), (torch.float32, 40, 2e-5), (torch.bfloat16, 128, 2e-2)])
def test_generic_vs_oracle_hooked_eager(dtype, dh, tol):
    T, H, Hkv = 333, 4, 2
    q, k, v, _ = make_qkv(T, H, Hkv, dh, dtype, DEV, seed=1)
    lc = make_logcred([T], DEV, dtype, seed=2)
    for heads in ([], [1], [0, 3], [0, 1, 2, 3]):
        got = ops.attn_prefill(q, k, v, cu_of([T], DEV), T, H, Hkv, dh, logcred=lc, head_mask=ops.heads_to_mask(heads),
                               kernel=GENERIC)
        want = _oracle_hooked_eager(q, k, v, T, H, Hkv, dh, lc, heads)
        err = report(f"generic {dtype} dh={dh} heads={heads}", got.cpu().float(), want.float())
        assert err <= tol
        assert not torch.isnan(got).any()


SM100_CASES = [
    # lens, H, Hkv, dh
    ([128], 4, 2, 128),
    ([256], 4, 4, 64),
    ([333], 4, 2, 128),
    ([1, 127, 129, 500, 64], 8, 2, 128),
    ([1000, 77], 4, 1, 64),
    ([2048], 8, 2, 128),
    ([300, 129], 64, 8, 128),     # Llama-3-70B heads: GQA group of 8, head_mask bits >= 32
    ([515, 64], 28, 4, 128),      # Qwen2-7B heads: GQA group of 7
    ([257], 40, 40, 128),         # Llama-2-13B heads (MHA), bits >= 32
    ([8192], 4, 1, 128),          # T >> 2048: 64 key tiles per row block, 32 CTAs per head
]


@pytest.mark.parametrize("lens,H,Hkv,dh", SM100_CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_sm100_vs_reference(lens, H, Hkv, dh, dtype):
    total = sum(lens)
    q, k, v, _ = make_qkv(total, H, Hkv, dh, dtype, DEV, seed=5)
    cu = cu_of(lens, DEV)
    lc = make_logcred(lens, DEV, dtype, seed=6)
    # no head / one low head / highest + middle + lowest head (bits >= 32 when H > 32) / all heads
    for mask in (0, 0b10, (1 << (H - 1)) | (1 << (H // 2)) | 1, (1 << H) - 1):
        got, lse = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100,
                                    return_lse=True)
        want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=mask)
        err = report(f"sm100 {dtype} lens={lens} H={H}/{Hkv} dh={dh} mask={mask:b}", got, want)
        assert not torch.isnan(got).any()
        assert err <= 2e-2
        gen, lse_g = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=GENERIC,
                                      return_lse=True)
        assert report("  vs generic", got, gen) <= 2e-2
        fin = torch.isfinite(lse_g)
        assert torch.equal(fin, torch.isfinite(lse))
        assert (lse[fin] - lse_g[fin]).abs().max().item() <= 2e-2


def test_sm100_small_vs_oracle_hooked_eager():
    """Same inputs through the reference's own arithmetic (bf16 eager + dense hooked mask)."""
    T, H, Hkv, dh = 384, 4, 2, 128
    q, k, v, _ = make_qkv(T, H, Hkv, dh, torch.bfloat16, DEV, seed=9)
    lc = make_logcred([T], DEV, torch.bfloat16, seed=10)
    for heads in ([], [2], [0, 1, 2, 3]):
        got = ops.attn_prefill(q, k, v, cu_of([T], DEV), T, H, Hkv, dh, logcred=lc, head_mask=ops.heads_to_mask(heads),
                               kernel=SM100)
        want = _oracle_hooked_eager(q, k, v, T, H, Hkv, dh, lc, heads)
        assert report(f"sm100 vs oracle eager heads={heads}", got.cpu().float(), want.float()) <= 2e-2


def test_sm100_properties_full_size():
    """Llama-3-8B layer shape, T = 2048 x 4 sequences: properties that need no oracle."""
    H, Hkv, dh = 32, 8, 128
    lens = [2048, 2011, 1987, 2048]
    total = sum(lens)
    q, k, v, _ = make_qkv(total, H, Hkv, dh, torch.bfloat16, DEV, seed=11)
    cu = cu_of(lens, DEV)
    lc = make_logcred(lens, DEV, torch.bfloat16, seed=12)
    mask = 0x0F0F00FF
    clean = ops.attn_prefill(q, k, v, cu, 2048, H, Hkv, dh, kernel=SM100)
    biased = ops.attn_prefill(q, k, v, cu, 2048, H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
    cv, bv = clean.view(total, H, dh), biased.view(total, H, dh)
    for h in range(H):
        if not (mask >> h) & 1:
            assert torch.equal(cv[:, h], bv[:, h]), h  # untouched heads are bit-identical
    assert not torch.equal(cv[:, 0], bv[:, 0])
    # zero log-credibility == clean
    zero = ops.attn_prefill(q, k, v, cu, 2048, H, Hkv, dh, logcred=torch.zeros_like(lc), head_mask=mask, kernel=SM100)
    assert torch.equal(zero, clean)
    # V = 1 -> every output is exactly the probability mass = 1
    ones = torch.ones_like(v)
    o = ops.attn_prefill(q, k, ones, cu, 2048, H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
    assert (o.float() - 1).abs().max().item() <= 8e-3
    # run-to-run determinism
    again = ops.attn_prefill(q, k, v, cu, 2048, H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
    assert torch.equal(again, biased)
    # spot check against fp64 on two sequences' worth of heads
    want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=mask)
    assert report("sm100 8B-shape", biased, want) <= 2e-2


def test_chunked_prefill_and_bias_len():
    """len_q < len_k (queries are the tail of the keys) and a bias that covers only the prompt part."""
    H, Hkv, dh = 4, 2, 128
    lens_k, lens_q = [300, 513], [44, 129]
    q, _, _, _ = make_qkv(sum(lens_q), H, Hkv, dh, torch.bfloat16, DEV, seed=21)
    _, k, v, _ = make_qkv(sum(lens_k), H, Hkv, dh, torch.bfloat16, DEV, seed=22)
    lc = make_logcred(lens_k, DEV, torch.bfloat16, seed=23)
    bl = torch.tensor([250, 400], dtype=torch.int32, device=DEV)
    cuq, cuk = cu_of(lens_q, DEV), cu_of(lens_k, DEV)
    want = ref_attention(q, k, v, cuq, cuk, H, Hkv, dh, logcred=lc, head_mask=0b0110, bias_len=bl.tolist())
    for kern in (GENERIC, SM100):
        got = ops.attn_prefill(q, k, v, cuq, max(lens_q), H, Hkv, dh, logcred=lc, head_mask=0b0110, cu_seqlens_k=cuk,
                               max_seqlen_k=max(lens_k), bias_len=bl, kernel=kern)
        assert report(f"chunked kernel={kern}", got, want) <= 2e-2


def test_all_keys_zero_credibility_row_is_defined():
    """A row whose every visible key has credibility 0 (log = -inf): output defined as 0, never NaN."""
    T, H, Hkv, dh = 130, 2, 1, 64
    q, k, v, _ = make_qkv(T, H, Hkv, dh, torch.bfloat16, DEV, seed=31)
    lc = torch.full((T,), float("-inf"), dtype=torch.bfloat16, device=DEV)
    lc[100:] = 0
    for kern in (GENERIC, SM100):
        got = ops.attn_prefill(q, k, v, cu_of([T], DEV), T, H, Hkv, dh, logcred=lc, head_mask=0b01, kernel=kern)
        g = got.view(T, H, dh)
        assert not torch.isnan(got).any()
        assert (g[:100, 0] == 0).all() and (g[100:, 0] != 0).any()
        want = ref_attention(q, k, v, cu_of([T], DEV), None, H, Hkv, dh, logcred=lc, head_mask=0b01)
        assert report(f"zero-cred kernel={kern}", got, want) <= 2e-2


def test_sm100_persistent_queue_ragged_batches():
    """The persistent kernel walks a dynamic work queue: many more items than SMs, row blocks past the end of short
    sequences (never published), stages without tiles, single-row sequences, and the chunked mode (len_q < len_k) --
    every CTA crosses many item boundaries with TMEM, barriers and rings alive."""
    rng = np.random.default_rng(3)
    H, Hkv, dh = 16, 4, 128
    for trial in range(3):
        lens = [int(x) for x in rng.integers(1, 900, size=37)] + [1, 128, 129, 256, 257, 1025]
        rng.shuffle(lens)
        q, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, torch.bfloat16, DEV, seed=70 + trial)
        cu = cu_of(lens, DEV)
        lc = make_logcred(lens, DEV, torch.bfloat16, seed=80 + trial)
        mask = 0b1010011001011101
        got = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
        want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=mask)
        assert not torch.isnan(got).any()
        assert report(f"persistent ragged trial {trial} ({len(lens)} sequences)", got, want) <= 2e-2
        again = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
        assert torch.equal(got, again)  # the queue hands items to different CTAs from run to run; the bits do not move
    # chunked: queries are the tail of the keys, many sequences
    lens_k = [int(x) for x in rng.integers(300, 1500, size=24)]
    lens_q = [int(min(lk, x)) for lk, x in zip(lens_k, rng.integers(1, 400, size=24))]
    qf, _, _, _ = make_qkv(sum(lens_q), H, Hkv, dh, torch.bfloat16, DEV, seed=91)
    _, k, v, _ = make_qkv(sum(lens_k), H, Hkv, dh, torch.bfloat16, DEV, seed=92)
    lc = make_logcred(lens_k, DEV, torch.bfloat16, seed=93)
    cuq, cuk = cu_of(lens_q, DEV), cu_of(lens_k, DEV)
    got = ops.attn_prefill(qf, k, v, cuq, max(lens_q), H, Hkv, dh, logcred=lc, head_mask=0xF0F0, cu_seqlens_k=cuk,
                           max_seqlen_k=max(lens_k), kernel=SM100)
    want = ref_attention(qf, k, v, cuq, cuk, H, Hkv, dh, logcred=lc, head_mask=0xF0F0)
    assert report("persistent chunked", got, want) <= 2e-2


def test_sm100_rows_that_see_no_key_give_zero():
    """len_k < len_q under the causal alignment: the first len_q - len_k query rows see no key at all. They are work
    items without tiles (the softmax warps write zeros, LSE = -inf), next to items that do have tiles."""
    H, Hkv, dh = 4, 2, 128
    lens_q, lens_k = [300, 40, 513], [100, 40, 200]
    q, _, _, _ = make_qkv(sum(lens_q), H, Hkv, dh, torch.bfloat16, DEV, seed=95)
    _, k, v, _ = make_qkv(sum(lens_k), H, Hkv, dh, torch.bfloat16, DEV, seed=96)
    cuq, cuk = cu_of(lens_q, DEV), cu_of(lens_k, DEV)
    for kern in (GENERIC, SM100):
        got, lse = ops.attn_prefill(q, k, v, cuq, max(lens_q), H, Hkv, dh, cu_seqlens_k=cuk, max_seqlen_k=max(lens_k),
                                    kernel=kern, return_lse=True)
        want = ref_attention(q, k, v, cuq, cuk, H, Hkv, dh)
        assert not torch.isnan(got).any()
        assert report(f"no-key rows kernel={kern}", got, want) <= 2e-2
        assert (got[:200] == 0).all() and torch.isinf(lse[:, :200]).all()  # rows 0..199 of sequence 0 see nothing


def _dead_doc_logcred(lens, dead, dtype=torch.bfloat16):
    """log-credibility 0 everywhere except the listed (sequence, first key, last key + 1) spans, which are -inf."""
    parts = [torch.zeros(T, dtype=dtype) for T in lens]
    for s, a, b in dead:
        parts[s][a:b] = float("-inf")
    return torch.cat(parts).to(DEV)


def test_sm100_dead_tiles_are_skipped_exactly():
    """Dead-tile skipping (live_tiles_kernel + LiveTiles): 128-key tiles whose keys all have zero credibility are never
    visited on the biased heads. Spans that cover whole tiles, a dead prefix (row blocks that see no live key at all ->
    zeros), a sequence that is dead end to end, spans that cover no whole tile, and heads outside the mask, whose
    output must stay bit-identical to clean attention."""
    H, Hkv, dh = 8, 2, 128
    lens = [2048, 1500, 700, 130, 257]
    dead = [(0, 20, 420), (0, 1024, 1536), (1, 0, 512), (2, 100, 200), (2, 300, 390), (3, 0, 130), (4, 128, 256)]
    mask = 0b10110101
    for dtype in (torch.bfloat16, torch.float16):
        q, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, dtype, DEV, seed=101)
        cu = cu_of(lens, DEV)
        lc = _dead_doc_logcred(lens, dead, dtype)
        got = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
        want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=mask)
        assert not torch.isnan(got).any()
        assert report(f"dead tiles {dtype}", got, want) <= 2e-2
        gen = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=GENERIC)
        assert report(f"dead tiles {dtype} vs generic kernel", got, gen.double()) <= 2e-2
        clean = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, kernel=SM100)
        g, c = got.view(-1, H, dh), clean.view(-1, H, dh)
        for h in range(H):
            if not (mask >> h) & 1:
                assert torch.equal(g[:, h], c[:, h]), f"head {h} carries no bias and must not move"
        biased = [h for h in range(H) if (mask >> h) & 1]
        o1, o3 = int(cu[1]), int(cu[3])
        assert (g[o1:o1 + 512][:, biased] == 0).all()        # sequence 1: rows 0..511 see only dead keys
        assert (g[o1 + 512:o1 + 640][:, biased] != 0).any()
        assert (g[o3:o3 + 130][:, biased] == 0).all()        # sequence 3 is dead end to end
        again = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=mask, kernel=SM100)
        assert torch.equal(got, again)


def test_sm100_dead_tiles_chunked_bias_len_and_long_fallback():
    """(a) chunked prefill with bias_len: keys >= bias_len carry no bias whatever their log-credibility says, so their
    tiles are live; (b) a sequence of more than 64 tiles (8192 keys) does not fit the mask layout and runs every tile."""
    H, Hkv, dh = 4, 2, 128
    lens_k, lens_q = [1000, 700], [300, 700]
    q, _, _, _ = make_qkv(sum(lens_q), H, Hkv, dh, torch.bfloat16, DEV, seed=111)
    _, k, v, _ = make_qkv(sum(lens_k), H, Hkv, dh, torch.bfloat16, DEV, seed=112)
    lc = torch.full((sum(lens_k),), float("-inf"), dtype=torch.bfloat16, device=DEV)
    lc[1000 + 384:1000 + 600] = 0
    bl = torch.tensor([600, 650], dtype=torch.int32, device=DEV)
    cuq, cuk = cu_of(lens_q, DEV), cu_of(lens_k, DEV)
    got = ops.attn_prefill(q, k, v, cuq, max(lens_q), H, Hkv, dh, logcred=lc, head_mask=0b0111, cu_seqlens_k=cuk,
                           max_seqlen_k=max(lens_k), bias_len=bl, kernel=SM100)
    want = ref_attention(q, k, v, cuq, cuk, H, Hkv, dh, logcred=lc, head_mask=0b0111, bias_len=bl.tolist())
    assert not torch.isnan(got).any()
    assert report("dead tiles, chunked + bias_len", got, want) <= 2e-2
    assert (got.view(-1, H, dh)[300:300 + 384, :3] == 0).all()  # sequence 1, rows 0..383: every visible key is dead

    # 65 tiles: beyond the 64-bit live-tile mask of a sequence -- every tile runs (dead spans still give exact zeros)
    T, H, Hkv = 65 * 128, 2, 1
    q, k, v, _ = make_qkv(T, H, Hkv, dh, torch.bfloat16, DEV, seed=113)
    lc = _dead_doc_logcred([T], [(0, 50, 3000), (0, 5000, 5300), (0, 8190, 8320)])
    got = ops.attn_prefill(q, k, v, cu_of([T], DEV), T, H, Hkv, dh, logcred=lc, head_mask=0b01, kernel=SM100)
    want = ref_attention(q, k, v, cu_of([T], DEV), None, H, Hkv, dh, logcred=lc, head_mask=0b01)
    assert report("65 tiles: no masks, every tile runs", got, want) <= 2e-2
    T = 64 * 128  # one full word, bit 63 dead
    q, k, v, _ = make_qkv(T, H, Hkv, dh, torch.bfloat16, DEV, seed=114)
    lc = _dead_doc_logcred([T], [(0, 50, 3000), (0, 8000, 8192)])
    got = ops.attn_prefill(q, k, v, cu_of([T], DEV), T, H, Hkv, dh, logcred=lc, head_mask=0b01, kernel=SM100)
    want = ref_attention(q, k, v, cu_of([T], DEV), None, H, Hkv, dh, logcred=lc, head_mask=0b01)
    assert report("64 tiles: masks, bit 63", got, want) <= 2e-2


def test_sm100_dead_tiles_long_sequences_and_mask_layout_limits():
    """The layouts the live-tile masks do not cover run every tile and must give the same answers: sequences of more than
    8192 keys (20 000 and 9 000 here, with long dead spans and a dead prefix), and more than 1024 prompts in a batch."""
    H, Hkv, dh = 2, 1, 128
    lens = [20000, 9000]
    q, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, torch.bfloat16, DEV, seed=121)
    dead = [(0, 300, 2500), (0, 7000, 9500), (0, 8192 + 100, 16384 + 700), (1, 0, 8200)]
    lc = _dead_doc_logcred(lens, dead)
    cu = cu_of(lens, DEV)
    got = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=0b10, kernel=SM100)
    want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=0b10)
    assert not torch.isnan(got).any()
    assert report("dead spans, 157 + 71 tiles (no masks)", got, want) <= 2e-2
    clean = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, kernel=SM100)
    assert torch.equal(got.view(-1, H, dh)[:, 0], clean.view(-1, H, dh)[:, 0])  # head 0 carries no bias
    assert (got.view(-1, H, dh)[20000:20000 + 8200, 1] == 0).all()               # sequence 1: a dead prefix of 8200 keys
    del q, k, v, got, want, clean
    torch.cuda.empty_cache()

    rng = np.random.default_rng(5)
    lens = [int(x) for x in rng.integers(100, 300, size=1025)]
    q, k, v, _ = make_qkv(sum(lens), H, Hkv, dh, torch.bfloat16, DEV, seed=122)
    cu = cu_of(lens, DEV)
    lc = _dead_doc_logcred(lens, [(b, 0, 128) for b in range(0, 1025, 3)])
    got = ops.attn_prefill(q, k, v, cu, max(lens), H, Hkv, dh, logcred=lc, head_mask=0b11, kernel=SM100)
    want = ref_attention(q, k, v, cu, None, H, Hkv, dh, logcred=lc, head_mask=0b11)
    assert report("1025 prompts: no masks, every tile runs", got, want) <= 2e-2
    lens = lens[:1024]
    total = sum(lens)
    got = ops.attn_prefill(q[:total], k[:total], v[:total], cu[:1025], max(lens), H, Hkv, dh, logcred=lc[:total], head_mask=0b11,
                           kernel=SM100)
    assert report("1024 prompts: masks", got, want[:total]) <= 2e-2


@pytest.mark.parametrize("seed", list(range(12)))
def test_sm100_randomized_configurations(seed):
    """Seeded random draw over everything the entry point takes at once: ragged batches, GQA ratio, head dim, dtype, causal or
    not, chunked prefill (len_q < len_k) with or without bias_len, head masks with high bits, dead spans that may or may
    not cover whole tiles, an output buffer with a row stride of its own -- tcgen05 kernel vs the fp64 statement of the
    contract, and vs the CUDA-core kernel."""
    rng = np.random.default_rng(1000 + seed)
    dh = int(rng.choice([64, 128]))
    Hkv = int(rng.choice([1, 2, 4, 8]))
    H = Hkv * int(rng.choice([1, 2, 4, 7, 8]))
    dtype = torch.bfloat16 if rng.random() < 0.7 else torch.float16
    causal = bool(rng.random() < 0.8)
    n_seq = int(rng.integers(1, 7))
    lens_k = [int(x) for x in rng.integers(1, 1400, size=n_seq)]
    chunked = bool(rng.random() < 0.4)
    lens_q = [int(rng.integers(1, lk + 1)) for lk in lens_k] if chunked else list(lens_k)
    q, _, _, _ = make_qkv(sum(lens_q), H, Hkv, dh, dtype, DEV, seed=2000 + seed)
    _, k, v, _ = make_qkv(sum(lens_k), H, Hkv, dh, dtype, DEV, seed=3000 + seed)
    lc = make_logcred(lens_k, DEV, dtype, seed=4000 + seed)
    for s_ in range(n_seq):  # dead spans of random extent
        if rng.random() < 0.6 and lens_k[s_] > 10:
            a = int(rng.integers(0, lens_k[s_] - 1))
            b = int(rng.integers(a + 1, min(lens_k[s_], a + 700) + 1))
            off = sum(lens_k[:s_])
            lc[off + a:off + b] = float("-inf")
    mask = int(rng.integers(1, 1 << min(H, 62))) | (1 << (H - 1))
    bl = None
    if chunked and rng.random() < 0.5:
        bl = torch.tensor([int(rng.integers(0, lk + 1)) for lk in lens_k], dtype=torch.int32, device=DEV)
    cuq, cuk = cu_of(lens_q, DEV), cu_of(lens_k, DEV)
    pad = 8 * int(rng.integers(0, 3))
    buf = torch.full((sum(lens_q), H * dh + pad), float("nan"), dtype=dtype, device=DEV)
    out = buf[:, :H * dh]
    kw = dict(logcred=lc, head_mask=mask, cu_seqlens_k=cuk, max_seqlen_k=max(lens_k), bias_len=bl, causal=causal)
    got = ops.attn_prefill(q, k, v, cuq, max(lens_q), H, Hkv, dh, kernel=SM100, out=out, **kw)
    want = ref_attention(q, k, v, cuq, cuk, H, Hkv, dh, logcred=lc, head_mask=mask,
                         bias_len=None if bl is None else bl.tolist(), causal=causal)
    desc = f"random {seed}: H={H}/{Hkv} dh={dh} {dtype} causal={causal} chunked={chunked} lens_k={lens_k}"
    assert not torch.isnan(got).any(), desc
    assert report(desc, got, want) <= 2e-2
    if pad:
        assert torch.isnan(buf[:, H * dh:]).all(), "wrote outside the output columns"
    gen = ops.attn_prefill(q, k, v, cuq, max(lens_q), H, Hkv, dh, kernel=GENERIC, **kw)
    assert report(desc + " vs generic", got, gen.double()) <= 2e-2
