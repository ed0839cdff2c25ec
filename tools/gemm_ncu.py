#!/usr/bin/env python
"""Driver for ncu captures of every GEMM shape of a gpt2-124M step (B=32, T=1024), in a fixed order:
each case is launched 3 times (profile all; the third launch is the warm one).
    ncu --set full --clock-control none --profile-from-start off -k regex:gemm_pair_kernel -o gpurun_out/prof_gemm_shapes python tools/gemm_ncu.py
Prints the case list (index -> name, flops, algorithmic bytes) as JSON on stdout."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops
dev = "cuda"
M, C, V, Mc = 32768, 768, 50257, 16384
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
x, x3, x4 = bf(M, C), bf(M, 3 * C), bf(M, 4 * C)
w_qkv, w_proj, w_fc, w_fc2 = bf(3 * C, C), bf(C, C), bf(4 * C, C), bf(C, 4 * C)
b1, b3, b4 = torch.zeros(C, device=dev), torch.zeros(3 * C, device=dev), torch.zeros(4 * C, device=dev)
o1, o3, o4, p4 = torch.empty_like(x), torch.empty_like(x3), torch.empty_like(x4), torch.empty_like(x4)
dw1, dw3 = torch.zeros(C, C, device=dev), torch.zeros(3 * C, C, device=dev)
dw4, dw2 = torch.zeros(4 * C, C, device=dev), torch.zeros(C, 4 * C, device=dev)
hw = bf(V, C); lg = torch.empty(Mc, 50304, device=dev, dtype=torch.bfloat16); hdw = torch.zeros(V, C, device=dev)
cases = [
    ("fwd qkv + bias", lambda: ops.linear_fwd(x, w_qkv, o3, bias=b3), (M, 3 * C, C), 2 * (M * C + 3 * C * C + M * 3 * C)),
    ("fwd proj + bias + residual", lambda: ops.linear_fwd(x, w_proj, o1, bias=b1, residual=x), (M, C, C), 2 * (M * C + C * C + 2 * M * C)),
    ("fwd fc + bias + gelu + stored gelu'", lambda: ops.linear_fwd(x, w_fc, o4, bias=b4, act=1 | ops.ACT_STORE_GRAD, preact=p4), (M, 4 * C, C), 2 * (M * C + 4 * C * C + 2 * M * 4 * C)),
    ("fwd fc2 + bias + residual", lambda: ops.linear_fwd(x4, w_fc2, o1, bias=b1, residual=x), (M, C, 4 * C), 2 * (M * 4 * C + 4 * C * C + 2 * M * C)),
    ("dgrad fc2 * stored gelu'", lambda: ops.linear_dgrad(x, w_fc2, o4, gate=p4, gate_kind=ops.GATE_MUL), (M, 4 * C, C), 2 * (M * C + 4 * C * C + 2 * M * 4 * C)),
    ("dgrad fc", lambda: ops.linear_dgrad(x4, w_fc, o1), (M, C, 4 * C), 2 * (M * 4 * C + 4 * C * C + M * C)),
    ("dgrad proj", lambda: ops.linear_dgrad(x, w_proj, o1), (M, C, C), 2 * (M * C + C * C + M * C)),
    ("dgrad qkv", lambda: ops.linear_dgrad(x3, w_qkv, o1), (M, C, 3 * C), 2 * (M * 3 * C + 3 * C * C + M * C)),
    ("wgrad fc2", lambda: ops.linear_wgrad(x, x4, dw2), (C, 4 * C, M), 2 * (M * C + M * 4 * C) + 4 * 4 * C * C),
    ("wgrad fc", lambda: ops.linear_wgrad(x4, x, dw4), (4 * C, C, M), 2 * (M * C + M * 4 * C) + 4 * 4 * C * C),
    ("wgrad proj", lambda: ops.linear_wgrad(x, x, dw1), (C, C, M), 2 * (2 * M * C) + 4 * C * C),
    ("wgrad qkv", lambda: ops.linear_wgrad(x3, x, dw3), (3 * C, C, M), 2 * (M * 3 * C + M * C) + 4 * 3 * C * C),
    ("head fwd (16384 rows)", lambda: ops.linear_fwd(x[:Mc], hw, lg[:, :V]), (Mc, V, C), 2 * (Mc * C + V * C + Mc * V)),
    ("head dgrad (16384 rows)", lambda: ops.linear_dgrad(lg[:, :V], hw, o1[:Mc]), (Mc, C, V), 2 * (Mc * V + V * C + Mc * C)),
    ("head wgrad (16384 rows)", lambda: ops.linear_wgrad(lg[:, :V], x[:Mc], hdw), (V, C, Mc), 2 * (Mc * V + Mc * C) + 4 * V * C),
]
meta = []
for name, fn, (m, n, k), bytes_ in cases:
    fn(); fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()        # ncu --profile-from-start off: only this (warm) launch is captured
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    meta.append(dict(name=name, m=m, n=n, k=k, flops=2 * m * n * k, algorithmic_bytes=bytes_))
print(json.dumps(meta))
