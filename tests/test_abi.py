"""CPU: the C-ABI library builds, loads and exports every symbol include/cram_b200.h declares; host-only
entry points work; compute entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes
import json
import os
import re

import numpy as np
import pytest
import torch

from cram_b200 import _lib, build, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cram_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cram_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    path = build.build_library()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cram_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, "python binding and header disagree"
    assert _lib.load().cram_abi_version() == _lib.ABI_VERSION


def test_struct_layout_matches_header(tmp_path):
    """sizeof / offsetof as the C compiler sees the header == the ctypes Structures (guards field drift)."""
    import subprocess
    fields = {"cram_attn_prefill_params": _lib.AttnPrefillParams, "cram_attn_decode_params": _lib.AttnDecodeParams}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "cram_b200.h"', 'int main(void){']
    for cname, st in fields.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in st._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for cname, st in fields.items():
        assert int(out[cname]) == ctypes.sizeof(st), cname
        for fname, _ in st._fields_:
            assert int(out[f"{cname}.{fname}"]) == getattr(st, fname).offset, (cname, fname)


def test_rank_heads_host_entry_point(golden_dir):
    g = np.load(os.path.join(golden_dir, "ranking_kat.npz"))
    from cram_b200.find_best_heads import rank_mean_scores, select_top_k
    for key in ("nq_llama3", "trivia_qwen", "nq_llama2_13b_chat"):
        ranked = rank_mean_scores(g[key + "_scores"].astype(np.float64))
        vals = np.array([r[0] for r in ranked])
        assert np.array_equal(vals, g[key + "_mean_values"]), key
        idx = np.array([r[1] for r in ranked])
        diff = np.nonzero(~(idx == g[key + "_mean_index"]).all(axis=1))[0]
        for pos in diff:  # only inside exact-tie groups (documented tie rule: value desc, flat index asc)
            assert (vals == vals[pos]).sum() > 1
        sel = select_top_k(ranked, int(g[key + "_topk"]))
        want = json.loads(str(g[key + "_selected_json"]))
        assert {str(k): sorted(v) for k, v in sel.items()} == {k: sorted(v) for k, v in want.items()}
    # tie rule
    a = np.zeros((2, 2, 3)); a[:, 1, 1] = 1.0
    ranked = rank_mean_scores(a)
    assert [r[1] for r in ranked] == [(1, 1), (0, 0), (0, 1), (0, 2), (1, 0), (1, 2)]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_compute_calls_fail_loudly_without_gpu():
    lib = _lib.load()
    buf = (ctypes.c_int64 * 4)()
    cu = (ctypes.c_int32 * 2)(0, 4)
    rc = lib.cram_span_index(ctypes.addressof(buf), ctypes.addressof(cu), 1, None, None, None, ctypes.addressof(cu), 2,
                             None, None, None, None, None)
    assert rc == -3, rc  # CRAM_ERR_CUDA
    assert b"span_index_kernel" in lib.cram_last_error()
    with pytest.raises(ops.CramError):
        ops.rmsnorm(torch.zeros(2, 8), torch.ones(8), 1e-5)  # CPU tensors are rejected, never computed
    from cram_b200.re_weighting import Re_Weighting_Strategy
    with pytest.raises(ops.CramError):
        Re_Weighting_Strategy("tiny-llama", tokenizer=object(), engine=object())


def test_invalid_arguments_are_reported():
    lib = _lib.load()
    p = _lib.AttnPrefillParams()
    p.n_heads, p.n_kv_heads, p.head_dim = 6, 4, 64
    rc = lib.cram_attn_prefill(ctypes.byref(p), None)
    assert rc == -1 and b"head counts" in lib.cram_last_error()
