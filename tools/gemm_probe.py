#!/usr/bin/env python
"""GPU probe: times the GEMM shapes of a gpt2-124M step on both tile engines and store paths, and checks
the padded-vocab boundary. usage (on the box): python tools/gemm_probe.py > gpurun_out/gemm_probe.log"""
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
x4 = torch.randn(M, 4 * C, device=dev).bfloat16()
x3 = torch.randn(M, 3 * C, device=dev).bfloat16()
w_qkv = torch.randn(3 * C, C, device=dev).bfloat16() * 0.02
w_proj = torch.randn(C, C, device=dev).bfloat16() * 0.02
w_fc = torch.randn(4 * C, C, device=dev).bfloat16() * 0.02
w_fc2 = torch.randn(C, 4 * C, device=dev).bfloat16() * 0.02
b3 = torch.zeros(3 * C, device=dev); b1 = torch.zeros(C, device=dev); b4 = torch.zeros(4 * C, device=dev)
o3 = torch.empty(M, 3 * C, device=dev, dtype=torch.bfloat16); o1 = torch.empty(M, C, device=dev, dtype=torch.bfloat16)
o4 = torch.empty(M, 4 * C, device=dev, dtype=torch.bfloat16); p4 = torch.empty_like(o4)
dw4 = torch.zeros(4 * C, C, device=dev); dw2 = torch.zeros(C, 4 * C, device=dev); dw1 = torch.zeros(C, C, device=dev); dw3 = torch.zeros(3 * C, C, device=dev)

cases = {
    "fwd qkv  (M,2304,768)": (lambda: ops.linear_fwd(x, w_qkv, o3, bias=b3), 2 * M * 3 * C * C),
    "fwd proj (M,768,768)+res": (lambda: ops.linear_fwd(x, w_proj, o1, bias=b1, residual=x), 2 * M * C * C),
    "fwd fc   (M,3072,768)+gelu+pre": (lambda: ops.linear_fwd(x, w_fc, o4, bias=b4, act=1, preact=p4), 2 * M * 4 * C * C),
    "fwd fc   (M,3072,768) plain": (lambda: ops.linear_fwd(x, w_fc, o4, bias=b4), 2 * M * 4 * C * C),
    "fwd fc2  (M,768,3072)+res": (lambda: ops.linear_fwd(x4, w_fc2, o1, bias=b1, residual=x), 2 * M * 4 * C * C),
    "dgrad fc2 (M,3072,768)+gate": (lambda: ops.linear_dgrad(x, w_fc2, o4, gate=p4, gate_kind=1), 2 * M * 4 * C * C),
    "dgrad fc2 (M,3072,768) plain": (lambda: ops.linear_dgrad(x, w_fc2, o4), 2 * M * 4 * C * C),
    "dgrad fc  (M,768,3072)": (lambda: ops.linear_dgrad(x4, w_fc, o1), 2 * M * 4 * C * C),
    "dgrad qkv (M,768,2304)": (lambda: ops.linear_dgrad(x3, w_qkv, o1), 2 * M * 3 * C * C),
    "wgrad fc2 (768,3072,M)": (lambda: ops.linear_wgrad(x, x4, dw2), 2 * M * 4 * C * C),
    "wgrad fc  (3072,768,M)": (lambda: ops.linear_wgrad(x4, x, dw4), 2 * M * 4 * C * C),
    "wgrad proj (768,768,M)": (lambda: ops.linear_wgrad(x, x, dw1), 2 * M * C * C),
    "wgrad qkv (2304,768,M)": (lambda: ops.linear_wgrad(x3, x, dw3), 2 * M * 3 * C * C),
}
V, Mc = 50257, 8192
hw = torch.randn(V, C, device=dev).bfloat16() * 0.02
lg = torch.empty(Mc, 50304, device=dev, dtype=torch.bfloat16)
hdw = torch.zeros(V, C, device=dev)
cases["head fwd (8192,50257,768)"] = (lambda: ops.linear_fwd(x[:Mc], hw, lg[:, :V]), 2 * Mc * V * C)
cases["head dgrad (8192,768,50257)"] = (lambda: ops.linear_dgrad(lg[:, :V], hw, o1[:Mc]), 2 * Mc * V * C)
cases["head wgrad (50257,768,8192)"] = (lambda: ops.linear_wgrad(lg[:, :V], x[:Mc], hdw), 2 * Mc * V * C)

modes = [("pair+tma", 0, 0), ("pair+direct", 0, 1), ("single", 1, 0)]
print(f"{'case':36s}" + "".join(f"{m[0]:>22s}" for m in modes))
for name, (fn, flops) in cases.items():
    row = f"{name:36s}"
    for _, single, direct in modes:
        _lib.set_option("gemm_single", single); _lib.set_option("gemm_direct_store", direct)
        ms = bench(fn)
        row += f"{ms*1000:9.1f}us {flops/ms/1e9:7.0f}TF/s"
    print(row, flush=True)
_lib.set_option("gemm_single", 0); _lib.set_option("gemm_direct_store", 0)

# boundary check: padded vocab
Mq, Vq, Cq, Vpad = 256, 1001, 128, 1024
xq = torch.randn(Mq, Cq, device=dev).bfloat16(); wq = (torch.randn(Vq, Cq, device=dev) * 0.2).bfloat16()
for direct in (0, 1):
    _lib.set_option("gemm_direct_store", direct)
    buf = torch.full((Mq, Vpad), 7.0, dtype=torch.bfloat16, device=dev)
    ops.linear_fwd(xq, wq, buf[:, :Vq]); torch.cuda.synchronize()
    bad = (buf[:, Vq:] != 7.0).nonzero()
    ok = torch.allclose(buf[:, :Vq].float(), xq.float() @ wq.float().t(), rtol=2e-2, atol=2e-2)
    print("boundary direct=%d: values ok=%s, cells touched beyond N: %d" % (direct, ok, bad.shape[0]),
          (bad[:5].tolist(), buf[:, Vq:][tuple(bad[0])].item()) if bad.shape[0] else "")

# layout isolation at one cubic shape, and split-K sweep for a wgrad shape
print("\nlayout sweep (4096^3, bf16 out / fp32 accumulate for mnmn):")
n = 4096
a = torch.randn(n, n, device=dev).bfloat16(); b = torch.randn(n, n, device=dev).bfloat16()
ob = torch.empty(n, n, device=dev, dtype=torch.bfloat16); of = torch.zeros(n, n, device=dev)
for name, kw, out in [("kk", {}, ob), ("k,mn", dict(b_mn=True), ob), ("mn,k", dict(a_mn=True), ob), ("mn,mn", dict(a_mn=True, b_mn=True), ob),
                      ("mn,mn f32 acc", dict(a_mn=True, b_mn=True, accumulate=True), of), ("kk f32 acc", dict(accumulate=True), of)]:
    row = f"{name:20s}"
    for _, single, direct in modes:
        _lib.set_option("gemm_single", single); _lib.set_option("gemm_direct_store", direct)
        ms = bench(lambda: ops.gemm(a, b, out, **kw))
        row += f"{ms*1000:9.1f}us {2*n**3/ms/1e9:7.0f}TF/s"
    print(row, flush=True)
_lib.set_option("gemm_single", 0); _lib.set_option("gemm_direct_store", 0)
print("\nwgrad fc2 (768,3072,M=32768) split-K sweep (pair+tma):")
for sk in (1, 2, 4, 6, 8, 12, 16):
    ms = bench(lambda: ops.linear_wgrad(x, x4, dw2, split_k=sk))
    print(f"  split {sk:2d}: {ms*1000:8.1f}us {2*M*4*C*C/ms/1e9:7.0f}TF/s", flush=True)
for bn in (128, 256):
    ms = bench(lambda: ops.linear_wgrad(x, x4, dw2, block_n=bn))
    print(f"  block_n {bn}: {ms*1000:8.1f}us {2*M*4*C*C/ms/1e9:7.0f}TF/s", flush=True)
