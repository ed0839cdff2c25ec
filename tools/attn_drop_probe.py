#!/usr/bin/env python
"""GPU probe: attention forward / backward with attention-probability dropout (p = 0.1) vs without, gpt2-124M shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
def bench(fn, iters=6):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3
B, T, H, d = 32, 1024, 12, 64
C = H * d
qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16(); dout = torch.randn(B, T, C, device="cuda").bfloat16()
out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B, H, T, device="cuda")
dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device="cuda")
seed = torch.tensor([7], dtype=torch.int64, device="cuda")
print(f"fwd p=0   {bench(lambda: ops.attention_fwd(qkv, out, lse, B, T, H, d, 1)):8.1f} us")
print(f"fwd p=0.1 {bench(lambda: ops.attention_fwd_dropout(qkv, out, lse, B, T, H, d, 1, 0.1, seed, 0)):8.1f} us")
print(f"bwd p=0   {bench(lambda: ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, 1)):8.1f} us")
print(f"bwd p=0.1 {bench(lambda: ops.attention_bwd_dropout(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, 1, 0.1, seed, 0)):8.1f} us")
x = torch.randn(B * T, C, device="cuda").bfloat16(); y = torch.empty_like(x)
print(f"elementwise dropout + residual [32768x768] {bench(lambda: ops.dropout(x, x, y, 0.1, seed, 1)):8.1f} us")
