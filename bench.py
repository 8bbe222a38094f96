#!/usr/bin/env python
"""bench.py -- CrAM credibility-reweighted prefill (BASELINE.json configs[1]) on N B200s of one node,
plus the 32x32-head IE sweep (configs[2]) as a sub-measurement.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one batch: span index (kernel 2) -> 32-layer Llama-3-8B-shape
prefill with the credibility bias on the selected heads (kernel 1 inside every layer) -> next-token
argmax. Synthetic NQ-shaped prompts (question + 5 docs, ~2048 tokens, batch 16 per GPU), random-init
bf16 weights. Prefill does not shard: --gpus N runs N independent replicas ("weak").
The sweep is the part that shards; its heads/s at this N is reported under "sweep".

--impl reference times the reference's own CPU path (the oracle port; the reference is Python and
/root/reference does not travel to the GPU box) on the host cores for the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "cram_reweighted_prefill_tokens_per_s"
UNIT = "tokens/s"
WORKLOAD = "llama3-8b-shape bf16 CrAM prefill, question+5 docs ~2048 tok, batch 16/GPU, top-894 heads"
N_TOPK_HEADS = 894  # run.sh:4 of the reference (llama3, NQ)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="prefill", choices=["prefill", "decode", "longctx"],
                    help="prefill = BASELINE configs[1] (the default bench line); decode = configs[3] (Qwen2-7B shapes, paged "
                         "KV, batch 64, 256 new tokens); longctx = configs[4] (Llama-3-70B shapes, 20 docs / 32k tokens)")
    ap.add_argument("--batch", type=int, default=0, help="0 = the workload's own (16 prefill, 64 decode, 1 longctx)")
    ap.add_argument("--seq", type=int, default=0, help="0 = the workload's own (2048, 2048, 32768)")
    ap.add_argument("--docs", type=int, default=0, help="0 = the workload's own (5, 5, 20)")
    ap.add_argument("--new-tokens", type=int, default=256, help="decode workload: tokens generated per sequence per step")
    ap.add_argument("--page-size", type=int, default=16)
    ap.add_argument("--layers", type=int, default=0, help="debug only: fewer layers (marks the line invalid)")
    ap.add_argument("--sweep-prompts", type=int, default=1, help="prompts for the IE-sweep sub-measurement (0 = skip)")
    ap.add_argument("--sweep-layers", type=int, default=0, help="debug only: sweep only the last k layers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--attn-kernel", type=int, default=0, help="0 auto (tcgen05), 1 generic, 2 tcgen05")
    a = ap.parse_args()
    dflt = {"prefill": (16, 2048, 5), "decode": (64, 2048, 5), "longctx": (1, 32768, 20)}[a.workload]
    a.batch, a.seq, a.docs = a.batch or dflt[0], a.seq or dflt[1], a.docs or dflt[2]
    return a


# ---------------------------------------------------------------------------------------------------
# synthetic workload (shared by both arms)
# ---------------------------------------------------------------------------------------------------
def seeded_head_set(n_layers: int, n_heads: int, k: int, seed: int = 0):
    """top-k of a seeded random IE matrix -> {layer: [heads]}."""
    rng = np.random.default_rng(seed)
    ie = rng.standard_normal((1, n_layers, n_heads))
    order = np.lexsort((np.arange(n_layers * n_heads), -ie[0].flatten()))[:k]
    out = {}
    for f in order:
        out.setdefault(int(f) // n_heads, []).append(int(f) % n_heads)
    return out


def make_prompts(tok, batch: int, seq: int, docs: int, seed0: int, sweep: bool = False):
    from cram_b200 import synth
    from cram_b200.prompt import get_prompt
    qs, ps, ss, wr, lens = [], [], [], [], []
    for i in range(batch):
        q, paras, scores, wrong, n = synth.sample_with_token_target(
            tok, seed0 + i, docs, seq - (8 if sweep else 0), lambda q_, p_: get_prompt(p_, q_), sweep_scores=sweep)
        qs.append(q); ps.append(paras); ss.append(scores); wr.append(wrong); lens.append(n)
    return qs, ps, ss, wr, lens


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.index, self.rows, self._stop = index, [], threading.Event()
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(len(r) > 3 + j and r[3 + j].lower().startswith("active") for r in self.rows)]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return p.get("bf16_tflops", 1590.0), p.get("bf16_tflops_sustained", 1400.0), p.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1400.0, 6650.0, "fallback"


# ---------------------------------------------------------------------------------------------------
# CPU reference path (oracle port), used by --impl reference and by the cpu_baseline leg
# ---------------------------------------------------------------------------------------------------
def cpu_reference_sample(seq: int, n_layers_sample: int, steps: int, warmup: int, heads_per_layer: int = 28):
    """Times `n_layers_sample` Llama-3-8B-shape decoder layers of ONE sequence through the oracle's
    hooked-eager forward (dense [1,H,T,T] mask rewrite + eager attention + projections), bf16, all host
    threads. tokens/s is extrapolated to 32 layers (layers cost the same; the lm_head is last-row only)."""
    from oracle import cram_oracle as orc
    torch.set_grad_enabled(False)
    dtype = torch.bfloat16
    spec = orc.ModelSpec(n_layers_sample, 32, 8, 128, 4096, 14336, 1024, rms_eps=1e-5, rope_theta=500000.0)
    g = torch.Generator().manual_seed(0)

    def n(*s):
        return (torch.randn(*s, generator=g) * 0.02).to(dtype)

    sd = {"model.embed_tokens.weight": n(1024, 4096), "model.norm.weight": torch.ones(4096, dtype=dtype),
          "lm_head.weight": n(1024, 4096)}
    for l in range(n_layers_sample):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = torch.ones(4096, dtype=dtype)
        sd[p + "post_attention_layernorm.weight"] = torch.ones(4096, dtype=dtype)
        sd[p + "self_attn.q_proj.weight"] = n(4096, 4096)
        sd[p + "self_attn.k_proj.weight"] = n(1024, 4096)
        sd[p + "self_attn.v_proj.weight"] = n(1024, 4096)
        sd[p + "self_attn.o_proj.weight"] = n(4096, 4096)
        sd[p + "mlp.gate_proj.weight"] = n(14336, 4096)
        sd[p + "mlp.up_proj.weight"] = n(14336, 4096)
        sd[p + "mlp.down_proj.weight"] = n(4096, 14336)
    ids = torch.randint(0, 1024, (1, seq), generator=g)
    lc = torch.zeros(1, seq, dtype=dtype)
    per = max(1, (seq - 48) // 5)
    for d, sc in enumerate((0.1, 0.3, 0.5, 0.8, 1.0)):
        lc[:, 16 + d * per:16 + (d + 1) * per] = float(np.log(np.float16(sc)))
    hooks = {l: (lc, list(range(heads_per_layer))) for l in range(n_layers_sample)}
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        orc.llama_forward(sd, spec, ids, hooks)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    tokens_per_s = seq / ((ms / 1e3) * (32.0 / n_layers_sample))
    return tokens_per_s, ms, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to every rank; rank 0 alone works here, so give it the host's cores back
    # (the count an un-launched `python bench.py --impl reference` gets: physical cores).
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        try:
            import psutil
            ncores = psutil.cpu_count(logical=False) or os.cpu_count()
        except ImportError:
            ncores = os.cpu_count()
        torch.set_num_threads(max(1, min(ncores or 1, len(os.sched_getaffinity(0)))))
    nl = 2
    tps, ms, cores = cpu_reference_sample(args.seq, nl, max(1, args.steps), max(0, args.warmup))
    sample = (f"oracle port (reference is Python; /root/reference does not travel): {nl} of 32 layers x 1 of {args.batch} "
              f"sequences of {args.seq} tokens per step, hooked eager attention + projections, bf16, extrapolated x{32 // nl}")
    line = {"impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": args.batch, "seq_len": args.seq, "docs": args.docs,
                       "device": "host cpu"},
            "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
class Dist:
    """One process per GPU (torchrun env); barrier + max/sum over ranks for the timing contract."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py --impl ours needs a CUDA device: cram_b200 has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()

    def _reduce(self, v: float, op) -> float:
        if self.world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def allmax(self, v: float) -> float:
        return self._reduce(v, self.dist.ReduceOp.MAX)

    def allsum(self, v: float) -> float:
        return self._reduce(v, self.dist.ReduceOp.SUM)

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def timed_steps(D: Dist, fn, steps: int, sampler=None) -> float:
    """EXACTLY `steps` calls of fn between barrier + synchronize, CUDA events, max over ranks -> ms per step."""
    D.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.__enter__()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if sampler is not None:
        sampler.__exit__()
    D.barrier()
    return D.allmax(e0.elapsed_time(e1) / steps)


def run_ours(args):
    from cram_b200 import ops, synth
    from cram_b200.engine import LLAMA3_8B, CramEngine, ModelSpec
    from cram_b200.re_weighting import Find_Best_Heads, Re_Weighting_Strategy, _heads_dict_to_masks

    D = Dist()
    dist = D.dist
    world, rank, local, dev = D.world, D.rank, D.local, D.dev
    barrier, allmax, allsum = D.barrier, D.allmax, D.allsum

    spec = LLAMA3_8B
    if args.layers:
        spec = ModelSpec(args.layers, 32, 8, 128, 4096, 14336, 128256, 1e-5, 500000.0)
    engine = CramEngine.random_init(spec, seed=0, dtype=torch.bfloat16, device=dev, max_positions=args.seq + 256)
    tok = synth.load_tokenizer()
    heads = seeded_head_set(spec.num_layers, spec.num_heads, min(N_TOPK_HEADS, spec.num_layers * spec.num_heads))
    rw = Re_Weighting_Strategy("llama3-8b-shape-random-init", layers_to_be_modified=heads, tokenizer=tok, engine=engine)
    masks = _heads_dict_to_masks(heads, spec.num_layers)

    qs, ps, ss, _, lens = make_prompts(tok, args.batch, args.seq, args.docs, seed0=1000 * rank)
    prompts, ids, offs, sb, se, sc, cu, cud = rw._encode_batch(qs, ps, ss, [""] * args.batch)
    lens = [len(x) for x in ids]
    total_tokens = int(sum(lens))
    max_len = int(max(lens))

    # HOST buffers (pinned) of one batch: what the reference-facing call receives
    pin = lambda a, dt: torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=dt))).pin_memory()
    h = dict(ids=pin(np.concatenate(ids), np.int64), off=pin(np.concatenate(offs), np.int64), cu=pin(cu, np.int32),
             cud=pin(cud, np.int32), sb=pin(sb, np.int32), se=pin(se, np.int32), sc=pin(sc, np.float32))
    h2d_bytes = int(sum(t.numel() * t.element_size() for t in h.values()))
    d = {k: v.to(dev) for k, v in h.items()}
    last_rows = (d["cu"][1:] - 1).to(torch.int64)
    out_host = torch.empty(args.batch, dtype=torch.int64).pin_memory()

    def step_device(dd):
        _, _, logcred, _ = ops.span_index(dd["off"], dd["cu"], dd["sb"], dd["se"], dd["sc"], dd["cud"], torch.bfloat16)
        x = engine.prefill(dd["ids"], dd["cu"], max_len, logcred=logcred, head_masks=masks, attn_kernel=args.attn_kernel)
        return engine.logits(x[last_rows]).argmax(-1)

    def step_e2e():
        dd = {k: v.to(dev, non_blocking=True) for k, v in h.items()}
        nxt = step_device(dd)
        out_host.copy_(nxt, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the caller holds the answer tokens on the host
        return out_host

    with torch.no_grad():
        for _ in range(args.warmup):
            step_device(d)
        torch.cuda.synchronize()

        # ---- timed: device-resident inputs ------------------------------------------------------------
        ops.attn_events = []
        l0 = ops.launch_count
        barrier(); torch.cuda.synchronize()
        with ClockSampler(local) as clocks:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.steps):
                step_device(d)
            e1.record()
            torch.cuda.synchronize()
        barrier()
        launches = ops.launch_count - l0
        ms_step = allmax(e0.elapsed_time(e1) / args.steps)
        attn_ms = [a.elapsed_time(b) for a, b in ops.attn_events]
        ops.attn_events = None
        all_tokens = allsum(float(total_tokens))
        value = all_tokens / (ms_step / 1e3)

        # ---- timed: end to end from pinned host buffers -----------------------------------------------
        step_e2e()
        barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step_e2e()
        e1.record()
        torch.cuda.synchronize()
        barrier()
        ms_e2e = allmax(e0.elapsed_time(e1) / args.steps)
        e2e_value = all_tokens / (ms_e2e / 1e3)

    # ---- roofline of the dominant OWN kernel: the tcgen05 attention launch ----------------------------
    peak, peak_sus, hbm, which = measured_peaks()
    H, dh = spec.num_heads, spec.head_dim
    attn_flops = 4.0 * dh * H * sum(t * (t + 1) / 2 for t in lens)  # causal QK^T + PV, 2 FLOP/MAC
    attn_avg_ms = float(np.mean(attn_ms)) if attn_ms else float("nan")
    achieved = attn_flops / (attn_avg_ms / 1e3) / 1e12 if attn_ms else None
    # whole-step algorithmic FLOPs: 2*P_nonemb*tokens + attention + last-row lm_head
    p_layer = spec.hidden * (H * dh + 2 * spec.num_kv_heads * dh) + H * dh * spec.hidden + 3 * spec.hidden * spec.intermediate
    step_flops = 2.0 * p_layer * spec.num_layers * total_tokens + attn_flops * spec.num_layers + 2.0 * spec.vocab * spec.hidden * args.batch
    model_tflops = step_flops / (ms_step / 1e3) / 1e12
    # the launches are timed with events INSIDE the long step (SM clock under the power cap): the sustained figure is
    # the denominator; the burst one (the kernel alone does 842-853 TFLOP/s, scripts/profile_attn.py) is kept beside it
    roofline = {"bound": "tensor", "kernel": "attn_prefill_sm100_kernel<bf16,128>", "achieved": achieved, "peak": peak_sus,
                "unit": "TFLOP/s", "frac": (achieved / peak_sus) if achieved else None,
                "peak_burst": peak, "frac_of_burst": (achieved / peak) if achieved else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch at this shape, ncu --set full
                # (profiles/r1_attn_v4_ncu_and_trace.md); only recorded for the default shape
                "traffic": 1.079e9 if (args.batch, args.seq, args.layers) == (16, 2048, 0) else None, "traffic_unit": "bytes/launch",
                "algorithmic_bytes": 2.0 * total_tokens * (2 * spec.num_heads + 2 * spec.num_kv_heads) * spec.head_dim,
                "peak_source": f"{which} bf16_tflops_sustained (kernel timed inside the long step, between library GEMMs)",
                "launch_ms": attn_avg_ms, "launches_timed": len(attn_ms),
                "attn_share_of_step": (attn_avg_ms * spec.num_layers / ms_step) if attn_ms else None,
                "model_level": {"achieved": model_tflops, "peak": peak_sus, "frac": model_tflops / peak_sus,
                                "note": "whole prefill step (library GEMMs + own kernels) vs sustained bf16 peak"}}

    # ---- CPU baseline (rank 0, N = 1 only) -------------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tps, ms_cpu, cores = cpu_reference_sample(args.seq, 2, 2, 1)
        cpu_baseline = {"value": tps, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"oracle port, 2 of 32 layers x 1 sequence of {args.seq} tokens (hooked eager, bf16), "
                                  f"{ms_cpu:.0f} ms per sample, extrapolated x16"}

    # ---- the reference's hooked eager attention executed by PyTorch on THIS GPU (context: the bar a user of the
    # reference sees on a B200). One layer, one sequence: dense [1,H,T,T] mask rewrite
    # (utils/re_weighting.py:36-61) + HF eager attention (modeling_llama.py:199-221), via the oracle's restatement.
    ref_on_gpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            from oracle import cram_oracle as orc
            T0 = int(lens[0])
            gq = torch.randn(1, spec.num_heads, T0, spec.head_dim, device=dev).bfloat16()
            gk = torch.randn(1, spec.num_kv_heads, T0, spec.head_dim, device=dev).bfloat16()
            gv = torch.randn_like(gk)
            base = torch.triu(torch.full((T0, T0), torch.finfo(torch.bfloat16).min, dtype=torch.bfloat16, device=dev), 1)[None, None]
            lcg = torch.zeros(1, T0, dtype=torch.bfloat16, device=dev)
            lcg[:, 16:T0 - 32] = -0.6931
            hooked = list(range(28))

            def ref_layer():
                m = orc.edit_attention_mask(base, lcg, hooked, spec.num_heads, T0, T0, torch.bfloat16)
                return orc.eager_attention(gq, gk, gv, m, spec.head_dim ** -0.5)

            with torch.no_grad():
                ref_layer(); torch.cuda.synchronize()
                r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                r0.record()
                for _ in range(3):
                    ref_layer()
                r1.record(); torch.cuda.synchronize()
            ref_ms = r0.elapsed_time(r1) / 3
            ref_on_gpu = {"what": "reference hook + HF eager attention through PyTorch on this GPU, one layer x one sequence (28 heads hooked)",
                          "ms_per_layer_per_sequence": ref_ms, "ours_ms_per_layer_per_sequence": attn_avg_ms / args.batch,
                          "ratio": ref_ms / (attn_avg_ms / args.batch) if attn_ms else None}
        except Exception as e:  # context only: never fail the bench line on it
            ref_on_gpu = {"error": repr(e)[:200]}

    # ---- IE sweep sub-measurement (the part that shards) ------------------------------------------------
    sweep = None
    if args.sweep_prompts > 0:
        sweep = run_sweep(args, engine, tok, rank, world, dev, barrier, allmax, peak)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic",
                "config": {"workload": WORKLOAD, "batch_per_gpu": args.batch, "seq_len": args.seq, "docs": args.docs,
                           "tokens_per_step_per_gpu": total_tokens, "layers": spec.num_layers, "heads_biased": sum(len(v) for v in heads.values()),
                           "parallelism": f"replicas x{world} (prefill does not shard)",
                           "l2": "inputs larger than L2 (16 GB of weights streamed every step)",
                           "valid": args.layers == 0},
                "clocks": clocks.summary(),
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": int(out_host.numel() * out_host.element_size())},
                "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu_baseline,
                "reference_attention_on_this_gpu": ref_on_gpu, "sweep": sweep}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_sweep(args, engine, tok, rank, world, dev, barrier, allmax, peak):
    """Find_Best_Heads.cal_logits over all L x H heads of `sweep_prompts` prompts, heads sharded over ranks."""
    from cram_b200 import ops
    from cram_b200.re_weighting import Find_Best_Heads
    import torch.distributed as dist
    spec = engine.spec
    fb = Find_Best_Heads("llama3-8b-shape-random-init", tokenizer=tok, engine=engine, rank=rank, world_size=world)
    qs, ps, ss, wr, lens = make_prompts(tok, args.sweep_prompts, args.seq, args.docs, seed0=77, sweep=True)
    layers = list(range(spec.num_layers))
    if args.sweep_layers:
        layers = layers[-args.sweep_layers:]
    prepared = []
    for q, p, s, w in zip(qs, ps, ss, wr):
        ans = torch.as_tensor(tok([w])["input_ids"][0][1:], dtype=torch.int64, device=dev)
        inputs, cred = fb.decode_with_special_attention(question=q, paras=p, scores=s, answer=w)
        prepared.append((inputs["input_ids"][0], ops.log_cred(cred[0], torch.bfloat16), ans))
    with torch.no_grad():
        fb.sweep.cal_ie(*prepared[0], layers=[spec.num_layers - 1])  # warm-up: one cheap layer + the collective
        barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for ids, lc, ans in prepared:
            ie = fb.sweep.cal_ie(ids, lc, ans, layers=layers)
        ie_host = ie.cpu()
        e1.record()
        torch.cuda.synchronize()
        barrier()
    ms = allmax(e0.elapsed_time(e1))
    n_heads_swept = len(layers) * spec.num_heads * len(prepared)
    T = int(prepared[0][0].numel())
    # minimal-necessary FLOPs with clean-prefix reuse, per prompt
    H, dh, hid, mlp, L = spec.num_heads, spec.head_dim, spec.hidden, spec.intermediate, spec.num_layers
    p_layer = hid * (H * dh + 2 * spec.num_kv_heads * dh) + H * dh * hid + 3 * hid * mlp
    f_layer = 2.0 * p_layer * T + 4.0 * T * T * dh * H / 2
    f_1head = 4.0 * T * T * dh / 2 + 2.0 * T * H * dh * hid + 2.0 * T * 3 * hid * mlp
    flops = sum(H * (f_1head + (L - 1 - l) * f_layer) for l in layers) * len(prepared)
    tf = flops / (ms / 1e3) / 1e12
    return {"metric": "ie_sweep_heads_per_s", "value": n_heads_swept / (ms / 1e3), "unit": "heads/s", "n_gpus": world,
            "scaling": "strong", "prompts": len(prepared), "prompt_tokens": T, "layers_swept": len(layers),
            "heads_per_layer": spec.num_heads, "ms": ms, "finite": bool(torch.isfinite(ie_host[layers]).all()),
            "roofline": {"bound": "tensor", "achieved": tf, "peak": peak * world, "unit": "TFLOP/s", "frac": tf / (peak * world),
                         "note": "F of SURVEY.md 8(d) (clean-prefix reuse) / wall / (N x burst bf16 peak)"},
            "collective": "one all_gather of f32[L*H/N] per prompt" if world > 1 else "none (1 rank)",
            "valid": args.sweep_layers == 0 and args.layers == 0}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "decode":
        from bench_workloads import run_decode
        run_decode(a)
    elif a.workload == "longctx":
        from bench_workloads import run_longctx
        run_longctx(a)
    else:
        run_ours(a)
