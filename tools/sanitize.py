#!/usr/bin/env python
"""(compute-sanitizer is closed on this GPU pool -- gpurun refuses it -- so this runs as a plain tiny-shape smoke of every
kernel family; it is kept in the form the sanitizer would take.)
Tiny-shape driver for `compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize.py`: one call of
every kernel family through the C ABI (tcgen05 GEMM both engines and every epilogue specialisation, tcgen05 attention
forward / backward, mma.sync attention, LayerNorm, cross-entropy both kernels, embedding, AdamW, decode, token windows)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mingpt_distributed_b200 as m
from mingpt_distributed_b200 import ops

dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)
# GEMMs: pair engine (M = 512), every epilogue specialisation + the generic one, and the single-CTA engine (M = 64)
M, N, K = 512, 256, 128
a = rn(M, K).bfloat16(); b = rn(N, K).bfloat16(); bias = rn(N); res = rn(M, N).bfloat16()
o = torch.empty(M, N, dtype=torch.bfloat16, device=dev); p = torch.empty_like(o); f = torch.zeros(N, K, device=dev)
ops.linear_fwd(a, b, o, bias=bias)                                              # mode 1
ops.linear_fwd(a, b, o, bias=bias, residual=res)                                # mode 1 + aux
ops.linear_fwd(a, b, o, bias=bias, act=1 | ops.ACT_STORE_GRAD, preact=p)        # mode 2
ops.linear_fwd(a, b, o, bias=bias, act=2, preact=p)                             # generic (erf)
ops.linear_dgrad(o, b, torch.empty(M, K, dtype=torch.bfloat16, device=dev))     # MN-major B
ops.linear_dgrad(o, b, torch.empty(M, K, dtype=torch.bfloat16, device=dev), gate=a, gate_kind=ops.GATE_MUL, colsum=torch.zeros(K, device=dev))
ops.linear_wgrad(o, a, f)                                                       # mode 3, split-K reduce-add
ops.linear_fwd(a[:64], b, o[:64], bias=bias)                                    # single-CTA engine
# attention: tcgen05 (d = 64) and mma.sync (d = 32)
for d, H in ((64, 2), (32, 2)):
    B, T = 2, 256
    C = H * d
    qkv = rn(B, T, 3 * C).bfloat16(); dout = rn(B, T, C).bfloat16()
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device=dev); lse = torch.empty(B, H, T, device=dev)
    dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device=dev)
    for mode in (1, 2):
        ops.attention_fwd(qkv, out, lse, B, T, H, d, mode)
        ops.attention_bwd(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, mode, dqkv_colsum=torch.zeros(3 * C, device=dev))
    ops.attention_decode(qkv, out[:, 0], B, T, T - 1, H, d)
# whole model step (embedding, LayerNorm fwd/bwd, cross entropy, AdamW, dropout) on a d = 64 model
torch.manual_seed(0)
cfg = m.GPTConfig(model_type=None, n_layer=2, n_head=2, n_embed=128, vocab_size=211, block_size=128, embed_drop=0.1, resid_drop=0.1, attn_drop=0.1)
model = m.GPT(cfg, verbose=False).cuda()
opt = m.create_optimizer(model, m.OptimizerConfig())
x = torch.randint(0, 211, (4, 128), device=dev); y = torch.randint(0, 211, (4, 128), device=dev)
for _ in range(2):
    m.train_step(model, opt, x, y, 1.0, zero_grad=True)
model.generate(x[:, :8], 4)
# staged cross-entropy (needs >= 2 rows per SM) and the streaming one
V, ld, rows = 1000, 1000, 400
lg = rn(rows, ld).bfloat16(); tg = torch.randint(0, V, (rows,), device=dev); rl = torch.empty(rows, device=dev)
nv = torch.zeros(1, dtype=torch.int32, device=dev); ops.count_valid_targets(tg, nv)
ops.cross_entropy_fwd_bwd(lg, tg, rl, nv, V)
ops.cross_entropy_fwd_bwd(lg[:8], tg[:8], rl[:8], nv, V)
codes = torch.arange(0, 200, dtype=torch.uint8, device=dev)
xs = torch.empty(3, 16, dtype=torch.int64, device=dev); ys = torch.empty_like(xs)
ops.token_windows(codes, torch.tensor([0, 5, 100], device=dev), 16, xs, ys)
torch.cuda.synchronize()
print("sanitize driver ok")
