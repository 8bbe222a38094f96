"""CPU: oracle/hf_hooked.py (the reference's hooked HF path restated for the GPU box) against the fixture the UNMODIFIED
reference produced and against the oracle's own forward -- eager and sdpa (mask None -> hook's :39-48 branch)."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from cram_b200.engine import CramEngine, ModelSpec
from oracle import cram_oracle as orc
from oracle import hf_hooked
from oracle.tiny_model import TINY_SPEC, tiny_state_dict


def _fake_engine(sd, spec_o):
    spec = ModelSpec(spec_o.num_layers, spec_o.num_heads, spec_o.num_kv_heads, spec_o.head_dim, spec_o.hidden,
                     spec_o.intermediate, spec_o.vocab, spec_o.rms_eps, spec_o.rope_theta, qkv_bias=spec_o.qkv_bias)
    return SimpleNamespace(spec=spec, w=CramEngine.fuse_state_dict(spec, sd, "cpu"), _rope_len=8192, device="cpu")


@pytest.mark.parametrize("attn", ["eager", "sdpa"])
def test_hooked_hf_equals_reference_golden(golden_dir, attn):
    g = np.load(os.path.join(golden_dir, "tiny_model.npz"))
    model = hf_hooked.hf_from_engine(_fake_engine(tiny_state_dict(), TINY_SPEC), attn=attn)
    ids = torch.from_numpy(g["input_ids"])
    lc = orc.log_chain(g["cred_bits"].view(np.float16), torch.float16)  # the hook receives torch.log(fp16 weights)
    rows = g["logit_rows"]
    for name, l2h in (("none", {}), ("l0h1", {0: [1]}), ("all", {0: [0, 1, 2, 3], 1: [0, 1, 2, 3]})):
        lg = hf_hooked.hooked_logits(model, ids, lc, l2h)[0].numpy()[rows]
        assert np.abs(lg - g[f"logits_{name}"]).max() <= (2e-5 if attn == "eager" else 2e-4), (attn, name)


def test_cal_logits_heads_equals_reference_ie(golden_dir):
    g = np.load(os.path.join(golden_dir, "tiny_model.npz"))
    model = hf_hooked.hf_from_engine(_fake_engine(tiny_state_dict(), TINY_SPEC))
    ids = torch.from_numpy(g["ie_input_ids"])
    lc = orc.log_chain(g["ie_cred_bits"].view(np.float16), torch.float16)
    ans = torch.from_numpy(g["ie_answer_ids"])
    ori, ie = hf_hooked.cal_logits_heads(model, ids, lc, ans, [(0, 0), (1, 2), (1, 3)])
    assert abs(ori - float(g["ie_ori_logit_sum"])) <= 2e-4
    want = [g["ie_ref"][0, 0], g["ie_ref"][1, 2], g["ie_ref"][1, 3]]
    assert np.abs(np.array(ie) - np.array(want)).max() <= 2e-4
