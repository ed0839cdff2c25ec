#!/usr/bin/env python
"""One GEMM shape, a few launches: the driver for an ncu source-level capture.  usage: gemm_one.py [fc|qkv|proj|fcplain]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
which = sys.argv[1] if len(sys.argv) > 1 else "fc"
M, C = 32768, 768
x = torch.randn(M, C, device="cuda").bfloat16()
w = (torch.randn(4 * C, C, device="cuda") * 0.02).bfloat16()
b = torch.zeros(4 * C, device="cuda")
o = torch.empty(M, 4 * C, device="cuda", dtype=torch.bfloat16); p = torch.empty_like(o)
for _ in range(4):
    if which == "fc": ops.linear_fwd(x, w, o, bias=b, act=1 | ops.ACT_STORE_GRAD, preact=p)
    elif which == "fcplain": ops.linear_fwd(x, w, o, bias=b)
    elif which == "qkv": ops.linear_fwd(x, w[:3 * C], o[:, :3 * C], bias=b[:3 * C])
    elif which == "proj": ops.linear_fwd(x, w[:C], o[:, :C], bias=b[:C], residual=x)
torch.cuda.synchronize()
print("ok")
