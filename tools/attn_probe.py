#!/usr/bin/env python
"""GPU probe: attention forward/backward time at the gpt2-124M shape, tcgen05 vs mma.sync kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops, _lib
def bench(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters
B, T, H, d = 32, 1024, 12, 64
C = H * d
qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
dout = torch.randn(B, T, C, device="cuda").bfloat16()
out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B, H, T, device="cuda")
dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device="cuda")
for mode, name in ((1, "causal"), (2, "add_tril"), (0, "none")):
    fl = 4 * T * T * d * B * H * (0.5 if mode == 1 else 1.0)
    for mma in (0, 1):
        _lib.set_option("attn_mma", mma)
        f = bench(lambda: ops.attention_fwd(qkv, out, lse, B, T, H, d, mode))
        bw = bench(lambda: ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, mode))
        print(f"{name:9s} {'mma.sync' if mma else 'tcgen05 '}  fwd {f*1000:8.1f}us {fl/f/1e9:6.0f} TF/s   bwd {bw*1000:8.1f}us {2.5*fl/bw/1e9:6.0f} TF/s", flush=True)
_lib.set_option("attn_mma", 0)
