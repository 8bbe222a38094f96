"""GPU: kernel (3) -- the batched IE sweep vs the oracle's cal_logits, and sharded == unsharded."""
import numpy as np
import pytest
import torch

from cram_b200 import ops
from cram_b200.engine import CramEngine, ModelSpec
from cram_b200.sweep import IESweep, assemble_gathered
from oracle import cram_oracle as orc
from oracle.tiny_model import random_state_dict

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _setup(dtype, L=3, H=8, Hkv=4, dh=64, T=200, seed=5):
    ospec = orc.ModelSpec(L, H, Hkv, dh, H * dh // 2, 384, 500, rms_eps=1e-5, rope_theta=10000.0)
    sd = random_state_dict(ospec, seed=seed, dtype=dtype)
    spec = ModelSpec(L, H, Hkv, dh, ospec.hidden, ospec.intermediate, ospec.vocab, 1e-5, 10000.0)
    eng = CramEngine.from_state_dict(spec, sd, device=DEV)
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, 500, (T,), generator=g)
    cred = torch.ones(T, dtype=torch.float16)
    cred[6:70] = 0.0        # the fake document of the sweep (find_best_heads.py:45)
    ans = torch.randint(3, 500, (4,), generator=g)
    ids[-4:] = ans          # the prompt ends with the wrong answer (re_weighting.py:156)
    return ospec, sd, eng, ids, cred, ans


def test_sweep_fp32_matches_oracle_cal_logits():
    ospec, sd, eng, ids, cred, ans = _setup(torch.float32)
    with torch.no_grad():
        want = np.array(orc.cal_logits(sd, ospec, ids[None], torch.log(cred)[None].float(), ans))
    sw = IESweep(eng)
    lc = ops.log_cred(cred.to(DEV), torch.float32)
    got = sw.cal_ie(ids.to(DEV), lc, ans.to(DEV)).cpu().numpy()
    print("[sweep fp32] max_abs", np.abs(got - want).max(), "IE range", want.min(), want.max())
    assert np.abs(got - want).max() <= 5e-4
    k = 6
    assert set(np.argsort(-got.flatten())[:k].tolist()) == set(np.argsort(-want.flatten())[:k].tolist())


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_equals_unsharded_bit_exact(world):
    ospec, sd, eng, ids, cred, ans = _setup(torch.bfloat16, T=300)
    lc = ops.log_cred(cred.to(DEV), torch.bfloat16)
    ids_d, ans_d = ids.to(DEV), ans.to(DEV)
    chunk = min(2, ospec.num_heads // world)  # same GEMM M on every rank count (see IESweep.variant_chunk)
    full = IESweep(eng, variant_chunk=chunk).cal_ie(ids_d, lc, ans_d)
    parts = [IESweep(eng, rank=r, world_size=world, variant_chunk=chunk).local_ie(ids_d, lc, ans_d) for r in range(world)]
    sharded = assemble_gathered(parts, ospec.num_layers, ospec.num_heads, world)
    assert not torch.isnan(full).any()
    assert torch.equal(full, sharded)
    # IE values are model-dtype (bf16) numbers, like the reference's (value granularity)
    assert torch.equal(full, full.bfloat16().float())


def test_sweep_bf16_vs_oracle_bf16():
    """bf16 IE is noisy by construction (difference of two bf16 sums); compare against the oracle's bf16
    run with a tolerance of a few bf16 ulps of the logit sum, and against fp32 for the large effects."""
    ospec, sd, eng, ids, cred, ans = _setup(torch.bfloat16)
    with torch.no_grad():
        want = np.array(orc.cal_logits(sd, ospec, ids[None], torch.log(cred)[None].bfloat16(), ans))
        sd32 = {k: v.float() for k, v in sd.items()}
        truth = np.array(orc.cal_logits(sd32, ospec, ids[None], torch.log(cred)[None].float(), ans))
    got = IESweep(eng).cal_ie(ids.to(DEV), ops.log_cred(cred.to(DEV), torch.bfloat16), ans.to(DEV)).cpu().numpy()
    e_ours, e_ref = np.abs(got - truth).max(), np.abs(want - truth).max()
    print(f"[sweep bf16] ours-vs-fp32 {e_ours:.3e}  oracle-bf16-vs-fp32 {e_ref:.3e}")
    assert e_ours <= 2 * e_ref + 2e-2


def test_tail_only_last_layer_is_exact_work_skipping():
    """The last layer of every suffix run for the answer rows alone (K / V of all rows, Q / attention / o_proj / MLP of n
    rows) gives the IE of the full-row suffix: fp32 to rounding noise, and still equal to the oracle's cal_logits."""
    ospec, sd, eng, ids, cred, ans = _setup(torch.float32)
    lc = ops.log_cred(cred.to(DEV), torch.float32)
    full = IESweep(eng, tail_only_last_layer=False).cal_ie(ids.to(DEV), lc, ans.to(DEV)).cpu().numpy()
    tail = IESweep(eng, tail_only_last_layer=True).cal_ie(ids.to(DEV), lc, ans.to(DEV)).cpu().numpy()
    print("[sweep tail-only vs full rows] max_abs", np.abs(full - tail).max())
    assert np.abs(full - tail).max() <= 2e-5
    with torch.no_grad():
        want = np.array(orc.cal_logits(sd, ospec, ids[None], torch.log(cred)[None].float(), ans))
    assert np.abs(tail - want).max() <= 5e-4


def test_run_many_prompts_equals_one_by_one_and_reuses_buffers():
    """IESweep.run: several prompts of different lengths and answer lengths, one gather at the end == cal_ie per prompt,
    bit for bit; sharded over 4 simulated ranks too."""
    ospec, sd, eng, ids, cred, ans = _setup(torch.bfloat16, T=260)
    g = torch.Generator().manual_seed(9)
    prompts = []
    for T, n in ((260, 4), (200, 3), (260, 6)):
        i = torch.randint(3, 500, (T,), generator=g)
        a = torch.randint(3, 500, (n,), generator=g)
        i[-n:] = a
        c = torch.ones(T, dtype=torch.float16)
        c[5:60] = 0.0
        prompts.append((i.to(DEV), ops.log_cred(c.to(DEV), torch.bfloat16), a.to(DEV)))
    sw = IESweep(eng, variant_chunk=2)
    many = sw.run(prompts)
    assert many.shape == (3, ospec.num_layers, ospec.num_heads) and not torch.isnan(many).any()
    for p, pr in enumerate(prompts):
        assert torch.equal(many[p], IESweep(eng, variant_chunk=2).cal_ie(*pr))
    parts = [torch.stack([IESweep(eng, rank=r, world_size=4, variant_chunk=2).local_ie(*pr) for pr in prompts]) for r in range(4)]
    assert torch.equal(assemble_gathered(parts, ospec.num_layers, ospec.num_heads, 4), many)
