"""Kernel (3): batched head-intervention sweep for the indirect effect (IE) of every attention head.

The reference (Find_Best_Heads.cal_logits, utils/re_weighting.py:152-175) runs 1 + L*H full forwards
per prompt. Here one clean forward is cached per prompt (SweepCapture); for layer l the H single-head
variants share that layer's input, so ONE all-heads-biased attention launch + a head-splice kernel
builds them and the suffix layers l+1..L-1 run once over the batch of variants; the answer-token
logits come from a fused epilogue (final norm + the answer rows of lm_head + sum + IE) and a single
device->host copy per prompt replaces L*H ``.item()`` syncs.

Multi-GPU: every rank holds a model replica and the clean forward; for EVERY
layer rank r takes heads {h : h mod N == r} (cost is proportional to L - l, so splitting inside each
layer balances exactly); one all_gather of f32[L*H/N] per prompt assembles the matrix. The reduction
order does not depend on N, so sharded == unsharded bit for bit.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from . import ops
from .engine import CramEngine, SweepCapture


def heads_of_rank(n_heads: int, rank: int, world_size: int) -> List[int]:
    """Heads rank `rank` sweeps in every layer: h mod world_size == rank."""
    return [h for h in range(n_heads) if h % world_size == rank]


def assemble_gathered(gathered: Sequence[torch.Tensor], n_layers: int, n_heads: int, world_size: int) -> torch.Tensor:
    """Inverse of the head partition: gathered[r] is rank r's [L, len(heads_of_rank(r))] slice (padded to the
    widest rank with NaN when H % N != 0)."""
    lead = tuple(gathered[0].shape[:-2])  # () for one prompt, (P,) for a stack of prompts
    out = torch.full(lead + (n_layers, n_heads), float("nan"), dtype=gathered[0].dtype, device=gathered[0].device)
    for r in range(world_size):
        hs = heads_of_rank(n_heads, r, world_size)
        if hs:
            out[..., hs] = gathered[r][..., :len(hs)]
    return out


class IESweep:
    """Kernel (3) as a component: one object per engine, pre-allocated variant buffers, one entry per prompt
    (``cal_ie``) or per list of prompts (``run``, one all_gather for all of them)."""

    def __init__(self, engine: CramEngine, rank: int = 0, world_size: int = 1, group=None,
                 variant_chunk: int = 4, max_rows: int = 1 << 17, attn_kernel: int = 0, tail_only_last_layer: bool = True):
        """variant_chunk: how many single-head variants go through the suffix layers per batch. It is
        FIXED (not "all of this rank's heads") so every GEMM sees the same M whatever the rank count;
        that is what makes the sharded sweep bit-identical to the unsharded one.
        tail_only_last_layer: exact work skipping -- the answer logits only read the last n rows, so the last layer of
        every suffix projects Q, attends, and runs o_proj / MLP for those rows alone (engine.suffix_from_attn)."""
        self.engine = engine
        self.rank, self.world_size, self.group = rank, world_size, group
        self.variant_chunk = variant_chunk
        self.max_rows = max_rows
        self.attn_kernel = attn_kernel
        self.tail_only = tail_only_last_layer
        self.heads = heads_of_rank(engine.spec.num_heads, rank, world_size)
        self._bufs = {}     # (T, n_var) -> dict of reusable device tensors
        self._heads_t = {}  # n_var chunking of self.heads -> int32 device tensors (constant across prompts)
        self.last_ori = None

    # ---- reusable device state ------------------------------------------------------------------------
    def _chunks(self, T: int):
        per = max(1, min(self.variant_chunk, self.max_rows // T))
        key = per
        if key not in self._heads_t:
            dev = self.engine.device
            self._heads_t[key] = [(hs, torch.tensor(hs, dtype=torch.int32, device=dev))
                                  for hs in (self.heads[c:c + per] for c in range(0, len(self.heads), per))]
        return self._heads_t[key]

    def _buffers(self, T: int, nv: int):
        key = (T, nv)
        if key not in self._bufs:
            if len(self._bufs) > 8:   # prompts of many different lengths: keep the footprint bounded
                self._bufs.clear()
            eng, dev = self.engine, self.engine.device
            self._bufs[key] = dict(
                x=torch.empty(nv * T, eng.spec.hidden, dtype=eng.dtype, device=dev),
                cu=torch.arange(nv + 1, dtype=torch.int32, device=dev) * T,
                pos=torch.arange(T, dtype=torch.int32, device=dev).repeat(nv),
                cu1=torch.tensor([0, T], dtype=torch.int32, device=dev),
                tails={})
        return self._bufs[key]

    @torch.no_grad()
    def clean_pass(self, input_ids: torch.Tensor, answer_ids: torch.Tensor):
        """One clean forward; returns (capture, ori) with ori the device scalar of re_weighting.py:160."""
        eng = self.engine
        T = input_ids.numel()
        cu = self._buffers(T, 1)["cu1"]
        cap = SweepCapture()
        x = eng.prefill(input_ids, cu, T, capture=cap, attn_kernel=self.attn_kernel)
        n = answer_ids.numel()
        ori, _ = ops.sweep_answer_ie(x[T - n:].contiguous()[None], eng.w["norm"], eng.spec.rms_eps, eng.w["lm_head"],
                                     answer_ids, None)
        return cap, ori

    @torch.no_grad()
    def sweep_layer(self, cap: SweepCapture, l: int, heads: Sequence[int], logcred: torch.Tensor,
                    answer_ids: torch.Tensor, ori: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """IE of `heads` of layer l -> f32 [len(heads)] on the device. No host-built tensors and no allocation of
        variant buffers inside the loop: they are cached per (prompt length, chunk size)."""
        eng, s = self.engine, self.engine.spec
        T = cap.x_in[l].shape[0]
        n = answer_ids.numel()
        chunks = self._chunks(T) if list(heads) == self.heads else [
            (list(heads[c:c + self.variant_chunk]), torch.tensor(list(heads[c:c + self.variant_chunk]), dtype=torch.int32,
                                                                 device=logcred.device))
            for c in range(0, len(heads), self.variant_chunk)]
        q, k, v = eng.split_qkv(cap.qkv[l])
        cu1 = self._buffers(T, 1)["cu1"]
        biased = ops.attn_prefill(q, k, v, cu1, T, s.num_heads, s.num_kv_heads, s.head_dim, logcred=logcred,
                                  head_mask=ops.heads_to_mask(heads), kernel=self.attn_kernel)
        if out is None:
            out = torch.empty(len(heads), dtype=torch.float32, device=logcred.device)
        c0 = 0
        for hs, hs_t in chunks:
            nv = len(hs)
            buf = self._buffers(T, nv)
            variants = ops.sweep_splice_heads(cap.attn[l], biased, hs_t, s.num_heads, s.head_dim)
            x = buf["x"]
            x.view(nv, T, s.hidden).copy_(cap.x_in[l])  # broadcast over the variants (the suffix updates x in place)
            tail = None
            if self.tail_only:
                if n not in buf["tails"]:
                    from .engine import SuffixTail
                    buf["tails"][n] = SuffixTail.make(nv, T, n, x.device)
                tail = buf["tails"][n]
            y = eng.suffix_from_attn(l, x, variants.view(nv * T, -1), buf["cu"], T, buf["pos"], attn_kernel=self.attn_kernel,
                                     tail=tail)
            rows = y.view(nv, n, s.hidden) if tail is not None else y.view(nv, T, s.hidden)[:, T - n:].contiguous()
            _, ie = ops.sweep_answer_ie(rows, eng.w["norm"], s.rms_eps, eng.w["lm_head"], answer_ids, ori)
            out[c0:c0 + nv] = ie
            c0 += nv
        return out

    @torch.no_grad()
    def local_ie(self, input_ids: torch.Tensor, logcred: torch.Tensor, answer_ids: torch.Tensor,
                 layers: Optional[Sequence[int]] = None) -> torch.Tensor:
        """This rank's slice [L, ceil(H/N)] (f32, device): column i is head rank + N*i; NaN where not swept."""
        s = self.engine.spec
        layers = list(range(s.num_layers)) if layers is None else list(layers)
        cap, ori = self.clean_pass(input_ids, answer_ids)
        self.last_ori = ori
        mine = self.heads
        width = (s.num_heads + self.world_size - 1) // self.world_size
        local = torch.full((s.num_layers, width), float("nan"), dtype=torch.float32, device=input_ids.device)
        if mine:
            for l in layers:
                self.sweep_layer(cap, l, mine, logcred, answer_ids, ori, out=local[l, :len(mine)])
        return local

    def gather(self, local: torch.Tensor) -> torch.Tensor:
        """The path's only collective: one all_gather of the per-rank IE slices -> [..., L, H] on every rank."""
        s = self.engine.spec
        return gather_ie(local, s.num_layers, s.num_heads, self.world_size, self.group)

    @torch.no_grad()
    def cal_ie(self, input_ids: torch.Tensor, logcred: torch.Tensor, answer_ids: torch.Tensor,
               layers: Optional[Sequence[int]] = None) -> torch.Tensor:
        """IE matrix [L, H] (f32, device) for one prompt; rows of layers not in `layers` are NaN."""
        return self.gather(self.local_ie(input_ids, logcred, answer_ids, layers))

    @torch.no_grad()
    def run(self, prompts: Sequence, layers: Optional[Sequence[int]] = None) -> torch.Tensor:
        """Sweep a list of prompts, each (input_ids, logcred, answer_ids) on the device -> IE [P, L, H] on every rank
        with ONE all_gather of f32[P, L, ceil(H/N)] at the end (512 KB per rank for P = 1000 at N = 8).
        Nothing is read back to the host between prompts, so the next prompt's launches queue behind the current one's."""
        locals_ = [self.local_ie(ids, lc, ans, layers) for ids, lc, ans in prompts]
        return self.gather(torch.stack(locals_))


def gather_ie(local: torch.Tensor, n_layers: int, n_heads: int, world_size: int, group=None) -> torch.Tensor:
    if world_size == 1:
        return assemble_gathered([local], n_layers, n_heads, 1)
    import torch.distributed as dist
    gathered = [torch.empty_like(local) for _ in range(world_size)]
    dist.all_gather(gathered, local.contiguous(), group=group)
    return assemble_gathered(gathered, n_layers, n_heads, world_size)
