"""One attention launch at the bench shape (Llama-3-8B layer, 16 x ~2048 tokens), for ncu and quick timing."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cram_b200 import ops

B, T, H, Hkv, dh = 16, 2048, 32, 8, 128
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
lens = [T - 3 * i for i in range(B)]
total = sum(lens)
qkv = torch.randn(total, (H + 2 * Hkv) * dh, device=dev, generator=g).bfloat16()
q, k, v = qkv[:, :H * dh], qkv[:, H * dh:(H + Hkv) * dh], qkv[:, (H + Hkv) * dh:]
cu = torch.zeros(B + 1, dtype=torch.int32); cu[1:] = torch.tensor(lens).cumsum(0); cu = cu.to(dev)
lc = torch.zeros(total, device=dev).bfloat16(); lc[100:900] = -0.69
flops = 4.0 * dh * H * sum(t * (t + 1) / 2 for t in lens)
for mask, name in ((0, "clean"), (0x0FFFFFFF, "biased28")):
    for _ in range(2):
        ops.attn_prefill(q, k, v, cu, T, H, Hkv, dh, logcred=lc, head_mask=mask, kernel=2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.attn_prefill(q, k, v, cu, T, H, Hkv, dh, logcred=lc, head_mask=mask, kernel=2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name}: {ms:.3f} ms/launch  {flops / ms / 1e9:.1f} TFLOP/s")
