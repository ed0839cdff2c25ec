#!/usr/bin/env python
"""Same-box A/B probe: times the kernels of a gpt2-124M step (B=32, T=1024) one by one. Run it twice in ONE gpurun
call -- once with MGB_LIB_PATH pointing at an older build of the library, once without -- and compare the columns;
numbers from different boxes differ by several percent (power-capped clocks).
usage: python tools/ab_probe.py [tag] [key=value library options ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mingpt_distributed_b200 import ops, _lib

tag = sys.argv[1] if len(sys.argv) > 1 else "run"
for kv in sys.argv[2:]:
    k, v = kv.split("=")
    try:
        _lib.set_option(k, int(v))
    except Exception as e:
        print(f"# option {k} not available in this build: {e}")
dev = "cuda"

def bench(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3

B, T, H, d = 32, 1024, 12, 64
M, C = B * T, H * d
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)
x = rn(M, C).bfloat16(); x4 = rn(M, 4 * C).bfloat16(); x3 = rn(M, 3 * C).bfloat16()
w_qkv = (rn(3 * C, C) * 0.02).bfloat16(); w_proj = (rn(C, C) * 0.02).bfloat16()
w_fc = (rn(4 * C, C) * 0.02).bfloat16(); w_fc2 = (rn(C, 4 * C) * 0.02).bfloat16()
b3 = torch.zeros(3 * C, device=dev); b1 = torch.zeros(C, device=dev); b4 = torch.zeros(4 * C, device=dev)
o3 = torch.empty(M, 3 * C, device=dev, dtype=torch.bfloat16); o1 = torch.empty(M, C, device=dev, dtype=torch.bfloat16)
o4 = torch.empty(M, 4 * C, device=dev, dtype=torch.bfloat16); p4 = torch.empty_like(o4)
dw4 = torch.zeros(4 * C, C, device=dev); dw2 = torch.zeros(C, 4 * C, device=dev); dw1 = torch.zeros(C, C, device=dev); dw3 = torch.zeros(3 * C, C, device=dev)
cs4 = torch.zeros(4 * C, device=dev)
SG = ops.ACT_STORE_GRAD
cases = [
    ("gemm fwd qkv +bias", lambda: ops.linear_fwd(x, w_qkv, o3, bias=b3), 2 * M * 3 * C * C),
    ("gemm fwd proj +bias+res", lambda: ops.linear_fwd(x, w_proj, o1, bias=b1, residual=x), 2 * M * C * C),
    ("gemm fwd fc +gelu+gelu'", lambda: ops.linear_fwd(x, w_fc, o4, bias=b4, act=1 | SG, preact=p4), 2 * M * 4 * C * C),
    ("gemm fwd fc plain", lambda: ops.linear_fwd(x, w_fc, o4, bias=b4), 2 * M * 4 * C * C),
    ("gemm fwd fc2 +bias+res", lambda: ops.linear_fwd(x4, w_fc2, o1, bias=b1, residual=x), 2 * M * 4 * C * C),
    ("gemm dgrad fc2 *gelu'+colsum", lambda: ops.linear_dgrad(x, w_fc2, o4, gate=p4, gate_kind=ops.GATE_MUL, colsum=cs4), 2 * M * 4 * C * C),
    ("gemm dgrad fc", lambda: ops.linear_dgrad(x4, w_fc, o1), 2 * M * 4 * C * C),
    ("gemm dgrad proj", lambda: ops.linear_dgrad(x, w_proj, o1), 2 * M * C * C),
    ("gemm dgrad qkv", lambda: ops.linear_dgrad(x3, w_qkv, o1), 2 * M * 3 * C * C),
    ("gemm wgrad fc2", lambda: ops.linear_wgrad(x, x4, dw2), 2 * M * 4 * C * C),
    ("gemm wgrad fc", lambda: ops.linear_wgrad(x4, x, dw4), 2 * M * 4 * C * C),
    ("gemm wgrad proj", lambda: ops.linear_wgrad(x, x, dw1), 2 * M * C * C),
    ("gemm wgrad qkv", lambda: ops.linear_wgrad(x3, x, dw3), 2 * M * 3 * C * C),
]
V, Mc = 50257, 16384
hw = (rn(V, C) * 0.02).bfloat16()
lg = torch.empty(Mc, 50304, device=dev, dtype=torch.bfloat16)
hdw = torch.zeros(V, C, device=dev)
cases += [("gemm head fwd", lambda: ops.linear_fwd(x[:Mc], hw, lg[:, :V]), 2 * Mc * V * C),
          ("gemm head dgrad", lambda: ops.linear_dgrad(lg[:, :V], hw, o1[:Mc]), 2 * Mc * V * C),
          ("gemm head wgrad", lambda: ops.linear_wgrad(lg[:, :V], x[:Mc], hdw), 2 * Mc * V * C)]
# cuBLAS (torch.matmul) on the same shapes, plain epilogue: the reference point for the tcgen05 mainloop, not a code path
of = torch.empty(4 * C, C, device=dev, dtype=torch.bfloat16)
cases += [("cublas fwd qkv", lambda: torch.matmul(x, w_qkv.t(), out=o3), 2 * M * 3 * C * C),
          ("cublas fwd fc", lambda: torch.matmul(x, w_fc.t(), out=o4), 2 * M * 4 * C * C),
          ("cublas fwd fc2", lambda: torch.matmul(x4, w_fc2.t(), out=o1), 2 * M * 4 * C * C),
          ("cublas dgrad fc2", lambda: torch.matmul(x, w_fc2, out=o4), 2 * M * 4 * C * C),
          ("cublas wgrad fc (bf16 out)", lambda: torch.matmul(x4.t(), x, out=of), 2 * M * 4 * C * C),
          ("cublas head fwd", lambda: torch.matmul(x[:Mc], hw.t(), out=lg[:, :V]) if False else torch.matmul(x[:Mc], hw.t()), 2 * Mc * V * C)]
# attention
qkv = rn(B, T, 3 * C).bfloat16(); dout = rn(B, T, C).bfloat16()
out = torch.empty(B, T, C, dtype=torch.bfloat16, device=dev); lse = torch.empty(B, H, T, device=dev)
dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device=dev)
csq = torch.zeros(3 * C, device=dev)
fl = 4 * T * T * d * B * H * 0.5
cases += [("attn fwd causal", lambda: ops.attention_fwd(qkv, out, lse, B, T, H, d, 1), fl),
          ("attn bwd causal +colsum", lambda: ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, 1, dqkv_colsum=csq), 2.5 * fl)]
# row kernels (bytes instead of flops)
gam = torch.ones(C, device=dev); bet = torch.zeros(C, device=dev); mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
dg = torch.zeros(C, device=dev); db = torch.zeros(C, device=dev); cs1 = torch.zeros(C, device=dev)
ops.layernorm_fwd(x, gam, bet, o1, mean, rstd)
y2 = torch.empty_like(x)
cases += [("ln fwd", lambda: ops.layernorm_fwd(x, gam, bet, o1, mean, rstd), -(2 * M * C * 2 + 8 * M)),
          ("ln bwd +dres+colsum", lambda: ops.layernorm_bwd(x, x, gam, mean, rstd, y2, dg, db, dres=o1, dres_colsum=cs1), -(4 * M * C * 2))]
tg = torch.randint(0, V, (Mc,), device=dev); rl = torch.empty(Mc, device=dev); nv = torch.zeros(1, dtype=torch.int32, device=dev)
ops.count_valid_targets(tg, nv)
cases += [("cross entropy 16384 rows", lambda: ops.cross_entropy_fwd_bwd(lg, tg, rl, nv, V), -(2 * Mc * 50304 * 2))]
print(f"# {tag}: lib={os.environ.get('MGB_LIB_PATH', 'current')}")
only = [t for t in os.environ.get("AB_ONLY", "").split(",") if t]
for name, fn, work in cases:
    if only and not any(name.startswith(t) for t in only):
        continue
    t = bench(fn)
    rate = f"{work / t / 1e6:7.0f} TF/s" if work > 0 else f"{-work / t / 1e3:7.0f} GB/s"
    print(f"{tag:10s} {name:32s} {t:8.1f} us {rate}", flush=True)
