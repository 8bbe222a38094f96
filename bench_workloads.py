"""bench.py's other two workloads (BASELINE.json configs[3] and configs[4]); same JSON contract as the default line.

    python bench.py --workload decode    Qwen2-7B shapes, CrAM greedy decode over the paged KV cache, batch 64, 256 new tokens
    python bench.py --workload longctx   Llama-3-70B shapes, 20 docs / 32k tokens, re-weighted prefill, top-k head set

Prefill and decode do not shard: --gpus N runs N independent replicas ("weak").
"""
from __future__ import annotations

import json
import time

import numpy as np
import torch

from bench import (ClockSampler, Dist, make_prompts, measured_peaks, ncu_traffic, seeded_head_set, timed_steps)


def _pin(a, dt):
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=dt))).pin_memory()


def _encode(rw, tok, args, seed0):
    qs, ps, ss, _, _ = make_prompts(tok, args.batch, args.seq, args.docs, seed0=seed0)
    prompts, ids, offs, sb, se, sc, cu, cud = rw._encode_batch(qs, ps, ss, [""] * args.batch)
    h = dict(ids=_pin(np.concatenate(ids), np.int64), off=_pin(np.concatenate(offs), np.int64), cu=_pin(cu, np.int32),
             cud=_pin(cud, np.int32), sb=_pin(sb, np.int32), se=_pin(se, np.int32), sc=_pin(sc, np.float32))
    return h, [len(x) for x in ids]


# ---------------------------------------------------------------------------------------------------
# CPU samples (oracle port) for the two workloads
# ---------------------------------------------------------------------------------------------------
def _random_layers_sd(spec, n_layers, dtype, vocab=1024, seed=0):
    g = torch.Generator().manual_seed(seed)

    def n(*s):
        return (torch.randn(*s, generator=g) * 0.02).to(dtype)

    H, Hkv, dh, d, f = spec.num_heads, spec.num_kv_heads, spec.head_dim, spec.hidden, spec.intermediate
    sd = {"model.embed_tokens.weight": n(vocab, d), "model.norm.weight": torch.ones(d, dtype=dtype), "lm_head.weight": n(vocab, d)}
    for l in range(n_layers):
        p = f"model.layers.{l}."
        sd[p + "input_layernorm.weight"] = torch.ones(d, dtype=dtype)
        sd[p + "post_attention_layernorm.weight"] = torch.ones(d, dtype=dtype)
        sd[p + "self_attn.q_proj.weight"] = n(H * dh, d)
        sd[p + "self_attn.k_proj.weight"] = n(Hkv * dh, d)
        sd[p + "self_attn.v_proj.weight"] = n(Hkv * dh, d)
        if spec.qkv_bias:
            sd[p + "self_attn.q_proj.bias"] = n(H * dh)
            sd[p + "self_attn.k_proj.bias"] = n(Hkv * dh)
            sd[p + "self_attn.v_proj.bias"] = n(Hkv * dh)
        sd[p + "self_attn.o_proj.weight"] = n(d, H * dh)
        sd[p + "mlp.gate_proj.weight"] = n(f, d)
        sd[p + "mlp.up_proj.weight"] = n(f, d)
        sd[p + "mlp.down_proj.weight"] = n(d, f)
    return sd


def cpu_decode_sample(spec, prompt_len: int, n_layers_sample: int = 2, steps: int = 8, warmup: int = 2):
    """The reference decodes one sequence at a time (b = 1): time `steps` cached decode steps of
    `n_layers_sample` Qwen2-7B-shape layers through the oracle's hooked cached step (bf16, all host
    threads) over a `prompt_len`-token cache; tokens/s extrapolated to the full layer count."""
    from oracle import cram_oracle as orc
    torch.set_grad_enabled(False)
    dtype = torch.bfloat16
    ospec = orc.ModelSpec(n_layers_sample, spec.num_heads, spec.num_kv_heads, spec.head_dim, spec.hidden, spec.intermediate,
                          1024, rms_eps=spec.rms_eps, rope_theta=spec.rope_theta, qkv_bias=spec.qkv_bias)
    sd = _random_layers_sd(spec, n_layers_sample, dtype)
    g = torch.Generator().manual_seed(1)
    kc = [(torch.randn(1, spec.num_kv_heads, prompt_len, spec.head_dim, generator=g) * 0.5).to(dtype) for _ in range(n_layers_sample)]
    vc = [(torch.randn(1, spec.num_kv_heads, prompt_len, spec.head_dim, generator=g) * 0.5).to(dtype) for _ in range(n_layers_sample)]
    lc = torch.zeros(1, prompt_len, dtype=dtype)
    lc[:, 16:prompt_len - 32] = float(np.log(np.float16(0.5)))
    hooks = {l: (lc, list(range(spec.num_heads - 4))) for l in range(n_layers_sample)}
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        orc.decode_step_cached(sd, ospec, torch.tensor([5]), prompt_len + i, kc, vc, hooks)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    return 1.0 / ((ms / 1e3) * (spec.num_layers / n_layers_sample)), ms, torch.get_num_threads()


def cpu_longctx_sample(spec, seq_sample: int = 2048, steps: int = 2, warmup: int = 1):
    """One Llama-3-70B-shape layer of a `seq_sample`-token prompt through the oracle's hooked eager
    forward. The reference cannot run the 32k config at all (its [1,H,T,T] mask is 128 GiB at T=32k), so
    tokens/s is extrapolated from the short sample assuming cost linear in T (which flatters the CPU)."""
    from oracle import cram_oracle as orc
    torch.set_grad_enabled(False)
    dtype = torch.bfloat16
    ospec = orc.ModelSpec(1, spec.num_heads, spec.num_kv_heads, spec.head_dim, spec.hidden, spec.intermediate, 1024,
                          rms_eps=spec.rms_eps, rope_theta=spec.rope_theta)
    sd = _random_layers_sd(spec, 1, dtype)
    ids = torch.randint(0, 1024, (1, seq_sample), generator=torch.Generator().manual_seed(0))
    lc = torch.zeros(1, seq_sample, dtype=dtype)
    lc[:, 16:seq_sample - 32] = float(np.log(np.float16(0.5)))
    hooks = {0: (lc, list(range(spec.num_heads - 8)))}
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        orc.llama_forward(sd, ospec, ids, hooks)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    return seq_sample / ((ms / 1e3) * spec.num_layers), ms, torch.get_num_threads()


# ---------------------------------------------------------------------------------------------------
# configs[3]: greedy decode with the credibility bias over the paged KV cache
# ---------------------------------------------------------------------------------------------------
def run_decode(args, D=None):
    """Standalone (D is None): prints the JSON line. As a sub-measurement of the default bench: returns it."""
    from cram_b200 import ops, synth
    from cram_b200.engine import QWEN2_7B, CramEngine, DecodeSession, ModelSpec, PagedKV
    from cram_b200.re_weighting import Re_Weighting_Strategy, _heads_dict_to_masks

    standalone = D is None
    D = D or Dist()
    dev, rank, world = D.dev, D.rank, D.world
    spec = QWEN2_7B
    if args.layers:
        spec = ModelSpec(args.layers, 28, 4, 128, 3584, 18944, 152064, 1e-6, 1000000.0, qkv_bias=True)
    B, new, ps = args.batch, args.new_tokens, args.page_size
    engine = CramEngine.random_init(spec, seed=0, dtype=torch.bfloat16, device=dev, max_positions=args.seq + new + 64)
    tok = synth.load_tokenizer()
    n_sel = int(round(0.873 * spec.num_layers * spec.num_heads))  # same fraction as run.sh:4's 894 of 1024
    heads = seeded_head_set(spec.num_layers, spec.num_heads, n_sel)
    rw = Re_Weighting_Strategy("qwen2-7b-shape-random-init", layers_to_be_modified=heads, tokenizer=tok, engine=engine)
    masks = _heads_dict_to_masks(heads, spec.num_layers)
    h, lens = _encode(rw, tok, args, seed0=5000 + 1000 * rank)
    h2d_bytes = int(sum(t.numel() * t.element_size() for t in h.values()))
    max_total = max(lens) + new + 1
    pages_per = (max_total + ps - 1) // ps
    kv = PagedKV(spec, B * pages_per, ps, torch.bfloat16, dev)
    # pages of a sequence are scattered over the pool like a real allocator leaves them
    block_table = torch.randperm(B * pages_per, device=dev, generator=torch.Generator(device=dev).manual_seed(3)).to(torch.int32).view(B, pages_per)
    cu_np = np.concatenate([[0], np.cumsum(lens)])

    def stage_and_prefill(dd):
        """H2D-resident inputs -> span index -> re-weighted prefill into the cache -> first tokens + decode bias."""
        _, _, logcred, _ = ops.span_index(dd["off"], dd["cu"], dd["sb"], dd["se"], dd["sc"], dd["cud"], torch.bfloat16)
        prompts = [dd["ids"][cu_np[b]:cu_np[b + 1]] for b in range(B)]
        lcs = [logcred[cu_np[b]:cu_np[b + 1]] for b in range(B)]
        logits = engine.prefill_into_cache(prompts, kv, block_table, lcs, masks, chunk=16)
        lc_dec = torch.zeros(B, max(lens), dtype=torch.bfloat16, device=dev)
        for b in range(B):
            lc_dec[b, :lens[b]] = lcs[b]
        return logits.argmax(-1), lc_dec

    with torch.no_grad():
        d = {k: v.to(dev) for k, v in h.items()}
        t0 = time.perf_counter()
        first, lc_dec = stage_and_prefill(d)
        torch.cuda.synchronize()
        prefill_s = time.perf_counter() - t0
        bias_len = torch.tensor(lens, dtype=torch.int32, device=dev)
        sess = DecodeSession(engine, kv, block_table, lens, new + 1, logcred=lc_dec, bias_len=bias_len, head_masks=masks)

        def step_device():
            sess.start(first)
            for _ in range(new):
                sess.step()

        for _ in range(max(args.warmup, 1)):
            step_device()
        torch.cuda.synchronize()
        clocks = ClockSampler(D.local)
        ms_step = timed_steps(D, step_device, args.steps, sampler=clocks)
        tokens_step = D.allsum(float(B * new))
        value = tokens_step / (ms_step / 1e3)
        tokens_dev = sess.out_tokens[1:new + 1].clone()

        # ---- end to end: pinned host buffers -> H2D -> span index -> prefill -> 256 decode steps -> D2H tokens
        out_host = torch.empty(new + 1, B, dtype=torch.int64).pin_memory()

        def step_e2e():
            dd = {k: v.to(dev, non_blocking=True) for k, v in h.items()}
            f, lcd = stage_and_prefill(dd)
            sess.logcred.copy_(lcd)
            sess.start(f)
            for _ in range(new):
                sess.step()
            out_host.copy_(sess.out_tokens, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        step_e2e()
        e2e_steps = max(1, min(args.steps, 2))
        ms_e2e = timed_steps(D, step_e2e, e2e_steps)
        e2e_value = tokens_step / (ms_e2e / 1e3)
        same = bool(torch.equal(out_host[1:new + 1].to(dev), tokens_dev))

        # ---- launches per token step (counted on one eager step) -------------------------------------------
        sess_e = DecodeSession(engine, kv, block_table, lens, 4, logcred=lc_dec, bias_len=bias_len, head_masks=masks,
                               use_graph=False)
        sess_e.start(first)
        l0 = ops.launch_count
        sess_e.step()
        torch.cuda.synchronize()
        launches_per_decode_step = ops.launch_count - l0

        # ---- roofline of the decode-attention kernel -----------------------------------------------------------
        # Event pairs cannot be recorded inside the replayed step graph and an eager launch is shorter than its
        # own Python call, so the kernel is timed in a graph of exactly the step's attention launches: all
        # layers, the real caches / block table / bias, sequence lengths of mid-generation.
        mid = torch.tensor([n + new // 2 for n in lens], dtype=torch.int32, device=dev)
        q_s = torch.randn(B, spec.num_heads * spec.head_dim, device=dev).bfloat16()
        out_s = torch.empty_like(q_s)
        ws = torch.empty(ops._lib.load().cram_attn_decode_workspace_bytes(B, spec.num_heads, spec.head_dim, max_total, 0),
                         dtype=torch.uint8, device=dev)

        def attn_only():
            for l in range(spec.num_layers):
                ops.attn_decode(q_s, kv.k[l], kv.v[l], block_table, mid, max_total, spec.num_heads, spec.num_kv_heads,
                                spec.head_dim, logcred=lc_dec, head_mask=masks[l], bias_len=bias_len, workspace=ws, out=out_s)

        attn_only()
        torch.cuda.synchronize()
        g_attn = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g_attn):
            attn_only()
        g_attn.replay()
        reps = 20
        ms_attn_graph = timed_steps(D, g_attn.replay, reps)
        dec_avg_ms = ms_attn_graph / spec.num_layers

        # ---- the decode MLP kernel (cram_linear_decode), timed the same way: every layer's gate/up + SwiGLU and
        # down + residual launches on the engine's own weights, in one graph
        mlp = None
        if engine.streaming_mlp and B <= 64:
            h_s = torch.randn(B, spec.hidden, device=dev).bfloat16()
            x_s = torch.zeros(B, spec.hidden, device=dev, dtype=torch.bfloat16)
            act_s = torch.empty(B, spec.intermediate, device=dev, dtype=torch.bfloat16)

            def mlp_only():
                for l in range(spec.num_layers):
                    ops.linear_decode(h_s, engine.w[f"wgu.{l}"], mode=ops.LINEAR_SWIGLU, out=act_s)
                    ops.linear_decode(act_s, engine.w[f"wdown.{l}"], mode=ops.LINEAR_RESIDUAL, out=x_s)

            mlp_only()
            torch.cuda.synchronize()
            g_mlp = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_mlp):
                mlp_only()
            g_mlp.replay()
            ms_mlp = timed_steps(D, g_mlp.replay, reps) / spec.num_layers
            mlp_bytes = 3.0 * spec.hidden * spec.intermediate * 2
            mlp = {"kernel": "linear_decode_sm100_kernel<bf16> (gate/up + SwiGLU, down + residual add)",
                   "ms_per_layer": ms_mlp, "weight_bytes_per_layer": mlp_bytes, "achieved": mlp_bytes / (ms_mlp / 1e3) / 1e9,
                   "unit": "GB/s", "launches_timed": reps * spec.num_layers * 2,
                   "timing": "CUDA events around replays of a graph holding the step's two MLP launches of every layer"}

    peak, peak_sus, hbm, which = measured_peaks()
    if mlp is not None:
        mlp["peak"] = hbm
        mlp["frac"] = mlp["achieved"] / hbm
        mlp["share_of_step"] = mlp["ms_per_layer"] * spec.num_layers / (ms_step / new)
    kv_bytes = float(sum(n + new // 2 for n in lens)) * 2 * spec.num_kv_heads * spec.head_dim * 2
    achieved = kv_bytes / (dec_avg_ms / 1e3) / 1e9
    p_layer = spec.hidden * (spec.num_heads + 2 * spec.num_kv_heads) * spec.head_dim + spec.num_heads * spec.head_dim * spec.hidden \
        + 3 * spec.hidden * spec.intermediate
    w_bytes = 2.0 * (p_layer * spec.num_layers + spec.vocab * spec.hidden)
    kv_step_bytes = B * (float(np.mean(lens)) + new / 2) * 2 * spec.num_kv_heads * spec.head_dim * 2 * spec.num_layers
    ms_token_step = ms_step / new
    model_gbs = (w_bytes + kv_step_bytes) / (ms_token_step / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": "attn_decode_mma_kernel<bf16,128> (+combine when split)", "achieved": achieved,
                "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                "traffic": ncu_traffic("attn_decode_mma", f"b{B}_t{args.seq}_p{ps}") if not args.layers else None,
                "traffic_unit": "bytes/launch",
                "peak_source": f"{which} hbm_gbs (copy bandwidth)", "launch_ms": dec_avg_ms, "launches_timed": reps * spec.num_layers,
                "bytes_per_launch": kv_bytes,
                "timing": "CUDA events around replays of a graph holding the step's attention launches (all layers, real caches)",
                "attn_share_of_step": dec_avg_ms * spec.num_layers / ms_token_step,
                "decode_mlp": mlp,
                "model_level": {"achieved": model_gbs, "peak": hbm, "frac": model_gbs / hbm, "ms_per_token_step": ms_token_step,
                                "note": "weights once + KV once per token step / graph-replayed step time"}}
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tps, ms_cpu, cores = cpu_decode_sample(spec, args.seq)
        cpu_baseline = {"value": tps, "unit": "tokens/s", "cores": cores, "kind": "port",
                        "sample": f"oracle port, cached hooked decode step, b=1 (the reference's batch), 2 of {spec.num_layers} layers, "
                                  f"{args.seq}-token cache, {ms_cpu:.1f} ms per sample, extrapolated x{spec.num_layers // 2}"}
    if rank == 0:
        line = {"metric": "cram_reweighted_decode_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"qwen2-7b-shape bf16 CrAM greedy decode, paged KV (page {ps}), batch {B}/GPU, prompt ~{args.seq} tok "
                                       f"(question+{args.docs} docs), {new} new tokens, top-{n_sel} heads",
                           "batch_per_gpu": B, "prompt_len": args.seq, "new_tokens": new, "page_size": ps, "layers": spec.num_layers,
                           "heads_biased": n_sel, "step": f"{new} CUDA-graph replays of the token step (prefill excluded from `value`, "
                           "included in `e2e`)", "prefill_s_once": prefill_s,
                           "parallelism": f"replicas x{world} (decode does not shard)",
                           "l2": "inputs larger than L2 (15 GB of weights + 8 GB of KV streamed every token step)",
                           "valid": args.layers == 0},
                "clocks": clocks.summary(),
                "e2e": {"value": e2e_value, "unit": "tokens/s", "ms_per_step": ms_e2e, "steps": e2e_steps, "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": int(out_host.numel() * out_host.element_size()), "tokens_equal_device_run": same,
                        "includes": "H2D, span index, re-weighted prefill into the paged cache, decode, D2H of the tokens"},
                "gpu_launches": launches_per_decode_step * new * args.steps, "roofline": roofline, "cpu_baseline": cpu_baseline}
        if standalone:
            print(json.dumps(line), flush=True)
    else:
        line = None
    del sess, kv, engine
    torch.cuda.empty_cache()
    if standalone:
        D.close()
    return line


# ---------------------------------------------------------------------------------------------------
# configs[4]: long-context prefill, Llama-3-70B shapes, 20 docs / 32k tokens
# ---------------------------------------------------------------------------------------------------
def run_longctx(args, D=None):
    """Standalone (D is None): prints the JSON line. As a sub-measurement of the default bench: returns it."""
    from cram_b200 import ops, synth
    from cram_b200.engine import LLAMA3_70B, CramEngine, ModelSpec
    from cram_b200.re_weighting import Re_Weighting_Strategy, _heads_dict_to_masks

    standalone = D is None
    D = D or Dist()
    dev, rank, world = D.dev, D.rank, D.world
    spec = LLAMA3_70B
    if args.layers:
        spec = ModelSpec(args.layers, 64, 8, 128, 8192, 28672, 128256, 1e-5, 500000.0)
    engine = CramEngine.random_init(spec, seed=0, dtype=torch.bfloat16, device=dev, max_positions=args.seq + 1024)
    tok = synth.load_tokenizer()
    n_sel = int(round(0.873 * spec.num_layers * spec.num_heads))
    heads = seeded_head_set(spec.num_layers, spec.num_heads, n_sel)
    rw = Re_Weighting_Strategy("llama3-70b-shape-random-init", layers_to_be_modified=heads, tokenizer=tok, engine=engine)
    masks = _heads_dict_to_masks(heads, spec.num_layers)
    h, lens = _encode(rw, tok, args, seed0=9000 + 1000 * rank)
    h2d_bytes = int(sum(t.numel() * t.element_size() for t in h.values()))
    total_tokens, max_len = int(sum(lens)), int(max(lens))
    d = {k: v.to(dev) for k, v in h.items()}
    last_rows = (d["cu"][1:] - 1).to(torch.int64)
    out_host = torch.empty(args.batch, dtype=torch.int64).pin_memory()

    def step_device(dd=d):
        _, _, logcred, _ = ops.span_index(dd["off"], dd["cu"], dd["sb"], dd["se"], dd["sc"], dd["cud"], torch.bfloat16)
        x = engine.prefill(dd["ids"], dd["cu"], max_len, logcred=logcred, head_masks=masks)
        return engine.logits(x[last_rows]).argmax(-1)

    def step_e2e():
        dd = {k: v.to(dev, non_blocking=True) for k, v in h.items()}
        out_host.copy_(step_device(dd), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    with torch.no_grad():
        for _ in range(args.warmup):
            step_device()
        torch.cuda.synchronize()
        ops.attn_events = []
        l0 = ops.launch_count
        clocks = ClockSampler(D.local)
        ms_step = timed_steps(D, step_device, args.steps, sampler=clocks)
        launches = ops.launch_count - l0
        attn_ms = [a.elapsed_time(b) for a, b in ops.attn_events]
        ops.attn_events = None
        step_e2e()
        ms_e2e = timed_steps(D, step_e2e, args.steps)
    all_tokens = D.allsum(float(total_tokens))
    peak, peak_sus, hbm, which = measured_peaks()
    H, dh = spec.num_heads, spec.head_dim
    attn_flops = 4.0 * dh * H * sum(t * (t + 1) / 2 for t in lens)
    attn_avg_ms = float(np.mean(attn_ms))
    achieved = attn_flops / (attn_avg_ms / 1e3) / 1e12
    p_layer = spec.hidden * (H * dh + 2 * spec.num_kv_heads * dh) + H * dh * spec.hidden + 3 * spec.hidden * spec.intermediate
    step_flops = 2.0 * p_layer * spec.num_layers * total_tokens + attn_flops * spec.num_layers + 2.0 * spec.vocab * spec.hidden * args.batch
    model_tflops = step_flops / (ms_step / 1e3) / 1e12
    roofline = {"bound": "tensor", "kernel": "attn_prefill_sm100_kernel<bf16,128>", "achieved": achieved, "peak": peak_sus,
                "unit": "TFLOP/s", "frac": achieved / peak_sus, "peak_burst": peak, "frac_of_burst": achieved / peak, "traffic": None,
                "peak_source": f"{which} bf16_tflops_sustained (kernel timed inside the long step)", "launch_ms": attn_avg_ms,
                "launches_timed": len(attn_ms),
                "attn_share_of_step": attn_avg_ms * spec.num_layers / ms_step,
                "model_level": {"achieved": model_tflops, "peak": peak_sus, "frac": model_tflops / peak_sus,
                                "note": "whole prefill step (library GEMMs + own kernels) vs sustained bf16 peak"}}
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        tps, ms_cpu, cores = cpu_longctx_sample(spec)
        cpu_baseline = {"value": tps, "unit": "tokens/s", "cores": cores, "kind": "port",
                        "sample": f"oracle port, 1 of {spec.num_layers} layers x 2048 tokens (hooked eager, bf16), {ms_cpu:.0f} ms, extrapolated "
                                  f"x{spec.num_layers} and assumed linear in T; the reference itself cannot run T=32k (128 GiB mask)"}
    if rank == 0:
        line = {"metric": "cram_reweighted_prefill_tokens_per_s", "value": all_tokens / (ms_step / 1e3), "unit": "tokens/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"llama3-70b-shape bf16 long-context CrAM prefill, question+{args.docs} docs ~{args.seq} tok, batch "
                                       f"{args.batch}/GPU, top-{n_sel} heads", "batch_per_gpu": args.batch, "seq_len": args.seq,
                           "docs": args.docs, "tokens_per_step_per_gpu": total_tokens, "layers": spec.num_layers, "heads_biased": n_sel,
                           "parallelism": f"replicas x{world} (prefill does not shard)",
                           "l2": "inputs larger than L2 (141 GB of weights streamed every step)", "valid": args.layers == 0},
                "clocks": clocks.summary(),
                "e2e": {"value": all_tokens / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": int(out_host.numel() * out_host.element_size())},
                "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu_baseline}
        if standalone:
            print(json.dumps(line), flush=True)
    else:
        line = None
    del engine, rw
    torch.cuda.empty_cache()
    if standalone:
        D.close()
    return line
