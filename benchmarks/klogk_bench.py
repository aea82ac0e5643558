"""Micro-benchmark: throughput of the device k*log2(k) evaluation (mtb_klogk_eval), CUDA events."""
import ctypes

import numpy as np
import torch

from mpitree_b200.ops import _lib
from mpitree_b200.ops.cuda_engine import klogk_table

lib = _lib.load()
dev = torch.device("cuda", 0)
n = 1 << 26
k = torch.randint(0, 1 << 20, (n,), dtype=torch.int32, device=dev)
out = torch.empty(n, dtype=torch.float64, device=dev)
tab = klogk_table(dev)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(3):
    lib.mtb_klogk_eval(ctypes.c_void_p(k.data_ptr()), ctypes.c_void_p(out.data_ptr()), ctypes.c_int64(n), ctypes.c_void_p(tab.data_ptr()), st)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5):
    lib.mtb_klogk_eval(ctypes.c_void_p(k.data_ptr()), ctypes.c_void_p(out.data_ptr()), ctypes.c_int64(n), ctypes.c_void_p(tab.data_ptr()), st)
e.record()
torch.cuda.synchronize()
ms = s.elapsed_time(e) / 5
print(f"klogk_eval: {n} values in {ms:.3f} ms = {n / ms / 1e6:.2f} G evals/s; bytes {12 * n / ms / 1e9:.2f} TB/s")
# fp64 elementwise reference: torch fma chain
x = torch.rand(n, dtype=torch.float64, device=dev)
torch.cuda.synchronize()
s.record()
y = x
for _ in range(8):
    y = torch.addcmul(x, y, y)
e.record()
torch.cuda.synchronize()
print(f"torch fp64 addcmul x8: {s.elapsed_time(e):.3f} ms")
