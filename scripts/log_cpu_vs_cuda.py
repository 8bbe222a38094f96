"""List the fp16 inputs on which torch.log differs between CPU and CUDA (the log chain of
utils/re_weighting.py:122 follows torch-CPU here). Writes profiles/r2_log_cpu_vs_cuda_fp16.txt; run on the GPU box."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cram_b200 import ops

bits = np.arange(65536, dtype=np.uint16)
x = bits.view(np.float16)
ok = ~(np.isnan(x) | (x < 0))
cpu = torch.log(torch.from_numpy(x.copy())).view(torch.int16).numpy().view(np.uint16)
gpu_t = torch.from_numpy(x.copy()).cuda()
cuda = torch.log(gpu_t).cpu().view(torch.int16).numpy().view(np.uint16)
ours = ops.log_cred(gpu_t, torch.float16).cpu().view(torch.int16).numpy().view(np.uint16)
assert np.array_equal(ours[ok], cpu[ok]), "cram_log_cred must follow torch-CPU"
d = np.nonzero(ok & (cpu != cuda))[0]
lines = [f"torch {torch.__version__}, {torch.cuda.get_device_name(0)}: torch.log(fp16) differs between CPU and CUDA on {d.size} of "
         f"{int(ok.sum())} non-negative, non-NaN fp16 inputs (always by one unit in the last place of the fp16 result);",
         "cram_log_cred equals torch-CPU on all of them. Columns: input bits, input value, CPU result bits / value, CUDA result bits / value",
         "The credibilities the reference's pipeline produces (k/10 for k = 0..10, RAG_test.py:92-103) are not in the list."]
k10 = {int(np.array([np.float32(k / 10)]).astype(np.float16).view(np.uint16)[0]) for k in range(11)}
assert not (k10 & set(int(i) for i in d))
for i in d:
    lines.append(f"0x{int(i):04x} {float(x[i])!r:>14} 0x{int(cpu[i]):04x} {float(cpu[i:i+1].view(np.float16)[0])!r:>12} "
                 f"0x{int(cuda[i]):04x} {float(cuda[i:i+1].view(np.float16)[0])!r:>12}")
open(os.path.join(ROOT, "gpurun_out", "r2_log_cpu_vs_cuda_fp16.txt"), "w").write("\n".join(lines) + "\n")
print(lines[0])
