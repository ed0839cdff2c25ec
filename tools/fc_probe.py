#!/usr/bin/env python
"""GPU probe: the c_fc forward GEMM (M=32768, N=3072, K=768) with its epilogue variants."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
def bench(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3
M, C = 32768, 768
x = torch.randn(M, C, device="cuda").bfloat16(); w = torch.randn(4 * C, C, device="cuda").bfloat16() * 0.02
b4 = torch.zeros(4 * C, device="cuda"); o4 = torch.empty(M, 4 * C, device="cuda", dtype=torch.bfloat16); p4 = torch.empty_like(o4)
fl = 2 * M * 4 * C * C
for name, fn in (("plain", lambda: ops.linear_fwd(x, w, o4, bias=b4)),
                 ("gelu", lambda: ops.linear_fwd(x, w, o4, bias=b4, act=1)),
                 ("gelu + preact", lambda: ops.linear_fwd(x, w, o4, bias=b4, act=1, preact=p4)),
                 ("gelu + stored derivative", lambda: ops.linear_fwd(x, w, o4, bias=b4, act=1 | ops.ACT_STORE_GRAD, preact=p4))):
    t = bench(fn)
    print(f"{name:28s} {t:7.1f} us {fl / t / 1e6:6.0f} TF/s", flush=True)
