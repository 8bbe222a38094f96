"""PyTorch-tensor front end of the C ABI (include/cram_b200.h). Torch is plumbing here: it owns the
device memory and the stream; every op below is one call into libcram_b200.so on the current stream.

All tensors must live on a CUDA device -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import CRAM_BF16, CRAM_F16, CRAM_F32, AttnDecodeParams, AttnPrefillParams, CramError, check

_DT = {torch.float32: CRAM_F32, torch.float16: CRAM_F16, torch.bfloat16: CRAM_BF16}

# counts launches of OUR kernels (bench.py reports it as gpu_launches)
launch_count = 0
# when a list, attn_prefill appends a (start, end) CUDA-event pair per launch, recorded on the launching stream
attn_events = None
# same for attn_decode (split-KV kernel + combine)
decode_events = None


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise CramError(f"unsupported dtype {dt}") from None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _cuda(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise CramError(f"{name} must be a CUDA tensor: cram_b200 has no CPU fallback")
    return t


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def heads_to_mask(heads: Sequence[int]) -> int:
    m = 0
    for h in heads:
        h = int(h)
        if not 0 <= h < 64:
            raise CramError(f"head index {h} outside [0, 64)")
        m |= 1 << h
    return m


# ------------------------------------------------------------------------------------------------
def span_index(offset_starts: torch.Tensor, cu_seqlens: torch.Tensor, span_begin: torch.Tensor,
               span_end: torch.Tensor, scores: torch.Tensor, cu_docs: torch.Tensor,
               logcred_dtype: torch.dtype = torch.bfloat16):
    """Kernel (2). Returns (doc_id int32 [T], cred fp16 [T], logcred [T] of logcred_dtype, err_flags int32 [B])."""
    global launch_count
    lib = _lib.load()
    _cuda(offset_starts, "offset_starts")
    T = offset_starts.numel()
    B = cu_seqlens.numel() - 1
    dev = offset_starts.device
    assert offset_starts.dtype == torch.int64 and cu_seqlens.dtype == torch.int32 and cu_docs.dtype == torch.int32
    assert span_begin.dtype == torch.int32 and span_end.dtype == torch.int32 and scores.dtype == torch.float32
    doc_id = torch.empty(T, dtype=torch.int32, device=dev)
    cred = torch.empty(T, dtype=torch.float16, device=dev)
    logcred = torch.empty(T, dtype=logcred_dtype, device=dev)
    err = torch.zeros(max(B, 1), dtype=torch.int32, device=dev)
    check(lib.cram_span_index(_ptr(offset_starts.contiguous()), _ptr(cu_seqlens), B, _ptr(span_begin), _ptr(span_end),
                              _ptr(scores), _ptr(cu_docs), dtype_code(logcred_dtype), _ptr(doc_id), _ptr(cred),
                              _ptr(logcred), _ptr(err), _stream()), "cram_span_index")
    launch_count += 1
    return doc_id, cred, logcred, err[:B]


def log_cred(cred_f16: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    global launch_count
    lib = _lib.load()
    _cuda(cred_f16, "cred")
    assert cred_f16.dtype == torch.float16
    c = cred_f16.contiguous()
    out = torch.empty(c.shape, dtype=dtype, device=c.device)
    check(lib.cram_log_cred(_ptr(c), c.numel(), dtype_code(dtype), _ptr(out), _stream()), "cram_log_cred")
    launch_count += 1
    return out


# ------------------------------------------------------------------------------------------------
def attn_prefill(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor, max_seqlen_q: int,
                 n_heads: int, n_kv_heads: int, head_dim: int, scale: Optional[float] = None,
                 logcred: Optional[torch.Tensor] = None, head_mask: int = 0,
                 cu_seqlens_k: Optional[torch.Tensor] = None, max_seqlen_k: Optional[int] = None,
                 bias_len: Optional[torch.Tensor] = None, causal: bool = True, out: Optional[torch.Tensor] = None,
                 return_lse: bool = False, kernel: int = 0):
    """Kernel (1), prefill. q/k/v are 2-D row views [tokens, heads*head_dim] (any row stride, unit
    column stride -- e.g. slices of one fused QKV buffer). Returns out [total_q, n_heads*head_dim]."""
    global launch_count
    lib = _lib.load()
    for name, t in (("q", q), ("k", k), ("v", v)):
        _cuda(t, name)
        if t.dim() != 2 or t.stride(1) != 1:
            raise CramError(f"{name} must be 2-D with unit column stride")
    total_q, total_k = q.shape[0], k.shape[0]
    if out is None:
        out = torch.empty(total_q, n_heads * head_dim, dtype=q.dtype, device=q.device)
    lse = torch.empty(n_heads, total_q, dtype=torch.float32, device=q.device) if return_lse else None
    p = AttnPrefillParams()
    p.q, p.k, p.v, p.out, p.lse = _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(lse)
    p.q_row_stride, p.k_row_stride, p.v_row_stride, p.out_row_stride = q.stride(0), k.stride(0), v.stride(0), out.stride(0)
    p.cu_seqlens_q, p.cu_seqlens_k = _ptr(cu_seqlens_q), _ptr(cu_seqlens_k)
    p.n_seq = cu_seqlens_q.numel() - 1
    p.total_q, p.total_k = total_q, total_k
    p.max_seqlen_q = int(max_seqlen_q)
    p.max_seqlen_k = int(max_seqlen_k if max_seqlen_k is not None else max_seqlen_q)
    p.n_heads, p.n_kv_heads, p.head_dim = n_heads, n_kv_heads, head_dim
    p.dtype = dtype_code(q.dtype)
    p.scale = float(scale if scale is not None else head_dim ** -0.5)
    if logcred is not None and head_mask:
        _cuda(logcred, "logcred")
        p.logcred, p.logcred_dtype = _ptr(logcred), dtype_code(logcred.dtype)
    else:
        p.logcred, p.logcred_dtype = None, CRAM_F32
    p.head_mask = int(head_mask) & 0xFFFFFFFFFFFFFFFF
    p.bias_len = _ptr(bias_len)
    p.causal = 1 if causal else 0
    p.kernel = kernel
    if attn_events is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
    check(lib.cram_attn_prefill(C.byref(p), _stream()), "cram_attn_prefill")
    if attn_events is not None:
        ev1.record()
        attn_events.append((ev0, ev1))
    launch_count += 1
    return (out, lse) if return_lse else out


def attn_decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                seqlens: torch.Tensor, max_seqlen: int, n_heads: int, n_kv_heads: int, head_dim: int,
                scale: Optional[float] = None, logcred: Optional[torch.Tensor] = None, head_mask: int = 0,
                bias_len: Optional[torch.Tensor] = None, n_splits: int = 0, workspace: Optional[torch.Tensor] = None,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Kernel (1), decode. k_cache/v_cache: [n_pages, page_size, n_kv_heads, head_dim]; q: [n_seq, n_heads*head_dim];
    logcred: [n_seq, T0max] (keys < bias_len[s] are biased)."""
    global launch_count
    lib = _lib.load()
    _cuda(q, "q")
    n_seq = q.shape[0]
    if out is None:
        out = torch.empty(n_seq, n_heads * head_dim, dtype=q.dtype, device=q.device)
    p = AttnDecodeParams()
    p.q, p.out = _ptr(q), _ptr(out)
    p.q_row_stride, p.out_row_stride = q.stride(0), out.stride(0)
    p.k_cache, p.v_cache = _ptr(k_cache), _ptr(v_cache)
    p.page_stride, p.token_stride = k_cache.stride(0), k_cache.stride(1)
    p.block_table, p.block_table_stride = _ptr(block_table), block_table.stride(0)
    p.seqlens = _ptr(seqlens)
    p.n_seq, p.page_size, p.max_seqlen = n_seq, k_cache.shape[1], int(max_seqlen)
    p.n_heads, p.n_kv_heads, p.head_dim = n_heads, n_kv_heads, head_dim
    p.dtype = dtype_code(q.dtype)
    p.scale = float(scale if scale is not None else head_dim ** -0.5)
    if logcred is not None and head_mask:
        p.logcred, p.logcred_stride, p.logcred_dtype = _ptr(logcred), logcred.stride(0), dtype_code(logcred.dtype)
    else:
        p.logcred, p.logcred_stride, p.logcred_dtype = None, 0, CRAM_F32
    p.head_mask = int(head_mask) & 0xFFFFFFFFFFFFFFFF
    p.bias_len = _ptr(bias_len)
    p.n_splits = n_splits
    if workspace is None:
        nbytes = lib.cram_attn_decode_workspace_bytes(n_seq, n_heads, head_dim, int(max_seqlen), n_splits)
        workspace = torch.empty(nbytes, dtype=torch.uint8, device=q.device)
    p.workspace, p.workspace_bytes = _ptr(workspace), workspace.numel() * workspace.element_size()
    if decode_events is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
    check(lib.cram_attn_decode(C.byref(p), _stream()), "cram_attn_decode")
    if decode_events is not None:
        ev1.record()
        decode_events.append((ev0, ev1))
    launch_count += 2
    return out


def attn_decode_static(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, block_table: torch.Tensor,
                       seqlens: torch.Tensor, max_seqlen: int, n_heads: int, n_kv_heads: int, head_dim: int, out: torch.Tensor,
                       scale: Optional[float] = None, logcred: Optional[torch.Tensor] = None, head_mask: int = 0,
                       bias_len: Optional[torch.Tensor] = None, n_splits: int = 0) -> torch.Tensor:
    """``attn_decode`` through cram_attn_decode_static: nothing is allocated, here or in the library -- ``out`` is the
    caller's, the split-KV partials live in the library's static device array (32 MiB; a launch that needs more raises
    CramError). No counterpart in the reference (include/cram_b200.h). Calls on different streams race on the array."""
    global launch_count
    lib = _lib.load()
    _cuda(q, "q")
    p = AttnDecodeParams()
    p.q, p.out = _ptr(q), _ptr(out)
    p.q_row_stride, p.out_row_stride = q.stride(0), out.stride(0)
    p.k_cache, p.v_cache = _ptr(k_cache), _ptr(v_cache)
    p.page_stride, p.token_stride = k_cache.stride(0), k_cache.stride(1)
    p.block_table, p.block_table_stride = _ptr(block_table), block_table.stride(0)
    p.seqlens = _ptr(seqlens)
    p.n_seq, p.page_size, p.max_seqlen = q.shape[0], k_cache.shape[1], int(max_seqlen)
    p.n_heads, p.n_kv_heads, p.head_dim = n_heads, n_kv_heads, head_dim
    p.dtype = dtype_code(q.dtype)
    p.scale = float(scale if scale is not None else head_dim ** -0.5)
    if logcred is not None and head_mask:
        p.logcred, p.logcred_stride, p.logcred_dtype = _ptr(logcred), logcred.stride(0), dtype_code(logcred.dtype)
    else:
        p.logcred, p.logcred_stride, p.logcred_dtype = None, 0, CRAM_F32
    p.head_mask = int(head_mask) & 0xFFFFFFFFFFFFFFFF
    p.bias_len = _ptr(bias_len)
    p.n_splits = n_splits
    p.workspace, p.workspace_bytes = None, 0
    check(lib.cram_attn_decode_static(C.byref(p), _stream()), "cram_attn_decode_static")
    launch_count += 2
    return out


def kv_append(k: torch.Tensor, v: torch.Tensor, slot_mapping: torch.Tensor, k_cache: torch.Tensor,
              v_cache: torch.Tensor, n_kv_heads: int, head_dim: int) -> None:
    global launch_count
    lib = _lib.load()
    _cuda(k, "k")
    assert slot_mapping.dtype == torch.int64 and k.stride(1) == 1 and v.stride(1) == 1
    check(lib.cram_kv_append(_ptr(k), _ptr(v), k.stride(0), v.stride(0), k.shape[0], n_kv_heads, head_dim,
                             dtype_code(k.dtype), _ptr(slot_mapping), _ptr(k_cache), _ptr(v_cache), k_cache.stride(0),
                             k_cache.stride(1), k_cache.shape[1], _stream()), "cram_kv_append")
    launch_count += 1


def rope_kv_append_(qkv: torch.Tensor, n_heads: int, n_kv_heads: int, head_dim: int, cos: torch.Tensor, sin: torch.Tensor,
                    positions: torch.Tensor, slot_mapping: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor) -> None:
    """rope_ on the Q and K heads of the fused QKV rows, then kv_append of rotated K and V, in ONE launch."""
    global launch_count
    lib = _lib.load()
    _cuda(qkv, "qkv")
    assert qkv.dim() == 2 and qkv.stride(1) == 1 and positions.dtype == torch.int32 and slot_mapping.dtype == torch.int64
    assert cos.is_contiguous() and sin.is_contiguous() and cos.dtype == qkv.dtype and cos.shape[1] == head_dim
    assert k_cache.dtype == qkv.dtype and k_cache.stride() == v_cache.stride()
    check(lib.cram_rope_kv_append(_ptr(qkv), qkv.stride(0), qkv.shape[0], n_heads, n_kv_heads, head_dim, _ptr(cos), _ptr(sin),
                                  _ptr(positions), _ptr(slot_mapping), _ptr(k_cache), _ptr(v_cache), k_cache.stride(0),
                                  k_cache.stride(1), k_cache.shape[1], dtype_code(qkv.dtype), _stream()),
          "cram_rope_kv_append")
    launch_count += 1


LINEAR_PLAIN, LINEAR_RESIDUAL, LINEAR_SWIGLU = 0, 1, 2

def linear_decode(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, mode: int = LINEAR_PLAIN,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Decode-step projection of <= 64 rows (cram_linear_decode). mode LINEAR_PLAIN: x @ w.T (+ bias);
    LINEAR_RESIDUAL: out += x @ w.T in place (out is the residual stream); LINEAR_SWIGLU: w = [gate; up] rows,
    returns silu(x @ gate.T) * (x @ up.T)."""
    global launch_count
    lib = _lib.load()
    _cuda(x, "x")
    assert x.dim() == 2 and w.dim() == 2 and x.stride(1) == 1 and w.stride(1) == 1 and x.dtype == w.dtype
    m, k = x.shape
    n_out = w.shape[0] // 2 if mode == LINEAR_SWIGLU else w.shape[0]
    assert w.shape[1] == k
    if out is None:
        assert mode != LINEAR_RESIDUAL, "LINEAR_RESIDUAL adds into `out`"
        out = torch.empty((m, n_out), dtype=x.dtype, device=x.device)
    assert out.shape == (m, n_out) and out.stride(1) == 1 and out.dtype == x.dtype
    check(lib.cram_linear_decode(_ptr(x), x.stride(0), m, _ptr(w), w.stride(0), n_out, k, _ptr(bias), _ptr(out),
                                 out.stride(0), mode, dtype_code(x.dtype), _stream()), "cram_linear_decode")
    launch_count += 1
    return out


# ------------------------------------------------------------------------------------------------
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out = weight * norm(x); with `residual`, first x += residual IN PLACE (fused residual add)."""
    global launch_count
    lib = _lib.load()
    _cuda(x, "x")
    assert x.is_contiguous() and x.dim() == 2 and (residual is None or residual.is_contiguous())
    if out is None:
        out = torch.empty_like(x)
    check(lib.cram_rmsnorm(_ptr(x), _ptr(residual), _ptr(weight), _ptr(out), x.shape[0], x.shape[1], float(eps),
                           dtype_code(x.dtype), _stream()), "cram_rmsnorm")
    launch_count += 1
    return out


def rope_(qkv: torch.Tensor, n_rot_heads: int, head_dim: int, cos: torch.Tensor, sin: torch.Tensor,
          positions: torch.Tensor) -> None:
    """In-place rotary embedding on the first n_rot_heads heads of every row of the fused QKV buffer."""
    global launch_count
    lib = _lib.load()
    _cuda(qkv, "qkv")
    assert qkv.dim() == 2 and qkv.stride(1) == 1 and positions.dtype == torch.int32
    assert cos.is_contiguous() and sin.is_contiguous() and cos.dtype == qkv.dtype and cos.shape[1] == head_dim
    check(lib.cram_rope(_ptr(qkv), qkv.stride(0), qkv.shape[0], n_rot_heads, head_dim, _ptr(cos), _ptr(sin),
                        _ptr(positions), dtype_code(qkv.dtype), _stream()), "cram_rope")
    launch_count += 1


def silu_mul(gate_up: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    global launch_count
    lib = _lib.load()
    _cuda(gate_up, "gate_up")
    assert gate_up.is_contiguous() and gate_up.dim() == 2
    inter = gate_up.shape[1] // 2
    if out is None:
        out = torch.empty(gate_up.shape[0], inter, dtype=gate_up.dtype, device=gate_up.device)
    check(lib.cram_silu_mul(_ptr(gate_up), _ptr(out), gate_up.shape[0], inter, dtype_code(gate_up.dtype), _stream()),
          "cram_silu_mul")
    launch_count += 1
    return out


def sweep_splice_heads(clean: torch.Tensor, biased: torch.Tensor, heads: torch.Tensor, n_heads: int,
                       head_dim: int) -> torch.Tensor:
    global launch_count
    lib = _lib.load()
    _cuda(clean, "clean")
    assert clean.is_contiguous() and biased.is_contiguous() and heads.dtype == torch.int32
    n_var = heads.numel()
    out = torch.empty(n_var, clean.shape[0], clean.shape[1], dtype=clean.dtype, device=clean.device)
    check(lib.cram_sweep_splice_heads(_ptr(clean), _ptr(biased), _ptr(heads), n_var, clean.shape[0], n_heads, head_dim,
                                      dtype_code(clean.dtype), _ptr(out), _stream()), "cram_sweep_splice_heads")
    launch_count += 1
    return out


def sweep_answer_ie(x: torch.Tensor, norm_weight: torch.Tensor, eps: float, lm_head: torch.Tensor,
                    answer_ids: torch.Tensor, ori: Optional[torch.Tensor]):
    """x: [n_var, T, hidden] final residual stream. Returns (cur f32 [n_var], ie f32 [n_var] or None)."""
    global launch_count
    lib = _lib.load()
    _cuda(x, "x")
    assert x.is_contiguous() and x.dim() == 3 and answer_ids.dtype == torch.int64 and lm_head.is_contiguous()
    n_var, T, hidden = x.shape
    cur = torch.empty(n_var, dtype=torch.float32, device=x.device)
    ie = torch.empty(n_var, dtype=torch.float32, device=x.device) if ori is not None else None
    check(lib.cram_sweep_answer_ie(_ptr(x), n_var, T, hidden, _ptr(norm_weight), float(eps), _ptr(lm_head),
                                   _ptr(answer_ids), answer_ids.numel(), dtype_code(x.dtype), _ptr(ori), _ptr(cur),
                                   _ptr(ie), _stream()), "cram_sweep_answer_ie")
    launch_count += 1
    return cur, ie


def rank_heads(scores: np.ndarray):
    """Host ranking (utils/find_best_heads.py:113-127). scores [N, L, H] float64 -> (order int32 [L*H], mean float64 [L*H])."""
    lib = _lib.load()
    a = np.ascontiguousarray(scores, dtype=np.float64)
    n, L, H = a.shape
    order = np.empty(L * H, dtype=np.int32)
    mean = np.empty(L * H, dtype=np.float64)
    check(lib.cram_rank_heads(a.ctypes.data, n, L, H, order.ctypes.data, mean.ctypes.data), "cram_rank_heads")
    return order, mean
