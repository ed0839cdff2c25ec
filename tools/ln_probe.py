#!/usr/bin/env python
"""GPU probe: LayerNorm forward / backward and the attention delta kernel at gpt2-124M shapes; LN backward with rows
staged through shared memory (bulk copies) vs held in registers. A 512 MB write between launches evicts L2."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops, _lib
dev = "cuda"
M, C = 32768, 768
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
x, dy, dres, y = bf(M, C), bf(M, C), bf(M, C), bf(M, C)
g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
ops.layernorm_fwd(x, g, b, y, mean, rstd)
def bench(fn, iters=10, cold=True):
    ts = []
    for _ in range(iters + 2):
        if cold: flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts = sorted(ts[2:])
    return ts[len(ts) // 2]
for cold in (True, False):
    tag = "cold L2" if cold else "warm L2"
    print(f"ln fwd                [{tag}] {bench(lambda: ops.layernorm_fwd(x, g, b, y, mean, rstd), cold=cold):7.1f} us")
    for opt in (0, 1):
        _lib.set_option("ln_unstaged", opt)
        t = bench(lambda: ops.layernorm_bwd(dy, x, g, mean, rstd, y, dg, db, dres=dres), cold=cold)
        print(f"ln bwd {'registers' if opt else 'staged   '}      [{tag}] {t:7.1f} us   {4 * M * C * 2 / t / 1e6:6.2f} TB/s")
    _lib.set_option("ln_unstaged", 0)
    cs = torch.zeros(C, device=dev)
    t = bench(lambda: ops.layernorm_bwd(dy, x, g, mean, rstd, y, dg, db, dres=dres, dres_colsum=cs), cold=cold)
    print(f"ln bwd staged + colsum [{tag}] {t:7.1f} us   (separate colsum [32768x768]: {bench(lambda: ops.colsum(dres, cs), cold=cold):5.1f} us)")
B, T, H, d = 32, 1024, 12, 64
qkv = bf(B, T, 3 * C); dout = bf(B, T, C)
out = torch.empty(B, T, C, dtype=torch.bfloat16, device=dev); lse = torch.empty(B, H, T, device=dev)
dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device=dev)
ops.attention_fwd(qkv, out, lse, B, T, H, d, 1)
print(f"attention bwd total   [cold L2] {bench(lambda: ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, 1)):7.1f} us")
rows, V, ld = 16384, 50257, 50304
lg = torch.randn(rows, ld, device=dev).bfloat16(); tg = torch.randint(0, V, (rows,), device=dev)
rl = torch.empty(rows, device=dev); nv = torch.full((1,), rows, dtype=torch.int32, device=dev)
for opt in (0, 1):
    _lib.set_option("ce_unstaged", opt)
    t = bench(lambda: ops.cross_entropy_fwd_bwd(lg, tg, rl, nv, V), iters=6)
    print(f"cross entropy {'streaming' if opt else 'staged   '} [16384 x 50257] {t:7.1f} us   {2 * rows * ld * 2 / t / 1e6:6.2f} TB/s")
_lib.set_option("ce_unstaged", 0)
