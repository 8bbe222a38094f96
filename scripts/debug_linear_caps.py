import ctypes, sys, os
sys.path.insert(0, "/root/repo")
import torch
torch.zeros(1, device="cuda")
from cram_b200 import _lib
lib = _lib.load()
caps = (ctypes.c_int * 4)()
sp = ctypes.c_int()
for (n, k, mode) in ((4608, 3584, 0), (3584, 3584, 1), (18944, 3584, 2), (3584, 18944, 1), (4096, 4096, 1), (14336, 4096, 2), (4096, 14336, 1)):
    lib.cram_debug_linear_caps(caps, n, k, mode, ctypes.byref(sp))
    print(n, k, mode, "caps", list(caps), "split", sp.value)
