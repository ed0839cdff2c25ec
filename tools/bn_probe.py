#!/usr/bin/env python
"""GPU probe: tile width (block_n) sweep on the N = 768 GEMMs of a gpt2-124M step (auto = the library's choice)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
dev = "cuda"
def bench(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3
M, C = 32768, 768
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
x, x3, x4 = bf(M, C), bf(M, 3 * C), bf(M, 4 * C)
w_proj, w_fc, w_fc2, w_qkv = bf(C, C), bf(4 * C, C), bf(C, 4 * C), bf(3 * C, C)
b1 = torch.zeros(C, device=dev); o1 = torch.empty_like(x)
cases = {
    "fwd proj+res (M,768,768)": lambda bn: ops.linear_fwd(x, w_proj, o1, bias=b1, residual=x, block_n=bn),
    "fwd fc2+res (M,768,3072)": lambda bn: ops.linear_fwd(x4, w_fc2, o1, bias=b1, residual=x, block_n=bn),
    "dgrad fc (M,768,3072)": lambda bn: ops.linear_dgrad(x4, w_fc, o1, block_n=bn),
    "dgrad proj (M,768,768)": lambda bn: ops.linear_dgrad(x, w_proj, o1, block_n=bn),
    "dgrad qkv (M,768,2304)": lambda bn: ops.linear_dgrad(x3, w_qkv, o1, block_n=bn),
}
print(f"{'case':30s}" + "".join(f"{str(b):>10s}" for b in ("auto", 128, 192, 256)))
for name, fn in cases.items():
    row = f"{name:30s}"
    for bn in (0, 128, 192, 256):
        try:
            row += f"{bench(lambda: fn(bn)):9.1f}u"
        except Exception as e:
            row += f"{'n/a':>10s}"
    print(row, flush=True)
