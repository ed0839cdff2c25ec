#!/usr/bin/env python
"""Minimal driver for ncu captures of the attention kernels at the gpt2-124M shape (B reduced to 8)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
B, T, H, d = int(os.environ.get("ATTN_B", "8")), 1024, 12, 64
C = H * d
qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
dout = torch.randn(B, T, C, device="cuda").bfloat16()
out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B, H, T, device="cuda")
dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device="cuda")
for _ in range(3):
    ops.attention_fwd(qkv, out, lse, B, T, H, d, 1)
    ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, 1)
torch.cuda.synchronize()
print("ok")
