#!/usr/bin/env python
"""GPU probe: cost of each epilogue feature on the fc-shaped GEMM (M=32768, N=3072, K=768)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops, _lib
dev = "cuda"
def bench(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters
M, C = 32768, 768
x = torch.randn(M, C, device=dev).bfloat16()
w = torch.randn(4 * C, C, device=dev).bfloat16() * 0.02
b4 = torch.zeros(4 * C, device=dev)
o4 = torch.empty(M, 4 * C, device=dev, dtype=torch.bfloat16); p4 = torch.empty_like(o4); r4 = torch.randn_like(o4)
fl = 2 * M * 4 * C * C
cases = {
  "plain": lambda: ops.linear_fwd(x, w, o4),
  "bias": lambda: ops.linear_fwd(x, w, o4, bias=b4),
  "bias+gelu_tanh": lambda: ops.linear_fwd(x, w, o4, bias=b4, act=1),
  "bias+gelu_erf": lambda: ops.linear_fwd(x, w, o4, bias=b4, act=2),
  "bias+preact": lambda: ops.linear_fwd(x, w, o4, bias=b4, preact=p4),
  "bias+gelu_tanh+preact": lambda: ops.linear_fwd(x, w, o4, bias=b4, act=1, preact=p4),
  "bias+residual": lambda: ops.linear_fwd(x, w, o4, bias=b4, residual=r4),
}
wt = torch.randn(C, 4 * C, device=dev).bfloat16() * 0.02
cases["dgrad plain"] = lambda: ops.linear_dgrad(x, wt, o4)
cases["dgrad gate tanh"] = lambda: ops.linear_dgrad(x, wt, o4, gate=p4, gate_kind=1)
cases["dgrad gate erf"] = lambda: ops.linear_dgrad(x, wt, o4, gate=p4, gate_kind=2)
for direct in (0, 1):
    _lib.set_option("gemm_direct_store", direct)
    print("direct_store =", direct)
    for name, fn in cases.items():
        ms = bench(fn)
        print(f"  {name:28s} {ms*1000:8.1f}us {fl/ms/1e9:7.0f} TF/s", flush=True)
