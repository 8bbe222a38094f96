"""One decode-attention launch at the cfg4 shape (Qwen2-7B layer, batch 64, ~2176 cached tokens, page 16), for ncu and timing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cram_b200 import ops

B, H, Hkv, dh, page = 64, 28, 4, 128, 16
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
splits_list = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
lens = [2048 + 128 + (i % 7) for i in range(B)]
pages_per = (max(lens) + page - 1) // page
n_layers = 6  # rotate over several layers' caches so the working set (6 x 285 MB) exceeds the 126 MB L2
kc = [torch.randn(B * pages_per, page, Hkv, dh, device=dev, generator=g).bfloat16() for _ in range(n_layers)]
vc = [torch.randn(B * pages_per, page, Hkv, dh, device=dev, generator=g).bfloat16() for _ in range(n_layers)]
bt = torch.randperm(B * pages_per, device=dev, generator=g).to(torch.int32).view(B, pages_per)
q = torch.randn(B, H * dh, device=dev, generator=g).bfloat16()
seqlens = torch.tensor(lens, dtype=torch.int32, device=dev)
lc = torch.zeros(B, 2048, device=dev).bfloat16(); lc[:, 100:900] = -0.69
bl = torch.full((B,), 2048, dtype=torch.int32, device=dev)
kv_bytes = sum(lens) * 2 * Hkv * dh * 2
ws = torch.empty(ops._lib.load().cram_attn_decode_workspace_bytes(B, H, dh, max(lens), 0), dtype=torch.uint8, device=dev)
out = torch.empty(B, H * dh, dtype=torch.bfloat16, device=dev)
n_in_graph = 24
for splits in splits_list:
    for mask, name in ((0, "clean"), (0x0FFFFFF, "biased24")):
        def launches():
            for i in range(n_in_graph):  # the Python call costs more than the kernel: time a CUDA graph of launches
                ops.attn_decode(q, kc[i % n_layers], vc[i % n_layers], bt, seqlens, max(lens), H, Hkv, dh, logcred=lc,
                                head_mask=mask, bias_len=bl, n_splits=splits, workspace=ws, out=out)
        launches()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            launches()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / (reps * n_in_graph)
        print(f"splits={splits} {name}: {ms*1e3:.1f} us/launch  {kv_bytes / ms / 1e6:.0f} GB/s of KV ({kv_bytes/1e6:.0f} MB)")
