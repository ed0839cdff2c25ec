#!/usr/bin/env python
"""Driver for ncu captures of the HBM-bound kernels at gpt2-124M shapes (M = 32768 rows, C = 768):
LayerNorm fwd/bwd, colsum, cross-entropy, AdamW, sumsq, embedding, dropout, attention delta/convert.
Each case: 2 warm launches, then one inside cudaProfilerStart/Stop (use ncu --profile-from-start off).
Prints the case list (name, algorithmic bytes) as JSON."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mingpt_distributed_b200 as m
from mingpt_distributed_b200 import ops
dev = "cuda"
M, C, V = 32768, 768, 50257
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
x, dy, dres, y = bf(M, C), bf(M, C), bf(M, C), bf(M, C)
g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
x4 = bf(M, 4 * C); db4 = torch.zeros(4 * C, device=dev)
rows = 16384
lg = bf(rows, 50304); tg = torch.randint(0, V, (rows,), device=dev); rl = torch.empty(rows, device=dev)
nv = torch.full((1,), rows, dtype=torch.int32, device=dev)
torch.manual_seed(0)
model = m.GPT(m.GPTConfig(model_type="gpt2", embed_drop=0, resid_drop=0, attn_drop=0), verbose=False).cuda()
opt = m.create_optimizer(model, m.OptimizerConfig())
model._ensure_device_state(); model._flat_grad.normal_()
for n_, p_ in model.named_parameters():
    o, k = model._spans[n_]; p_.grad = model._flat_grad[o:o + k].view(p_.shape)
P = model._total
idx = torch.randint(0, V, (32, 1024), device=dev)
emb = torch.empty(32, 1024, C, dtype=torch.bfloat16, device=dev)
seed = torch.tensor([1], dtype=torch.int64, device=dev)
kbits = torch.empty(ops.attention_dropout_bits_words(32, 1024, 12), dtype=torch.int32, device=dev)
ops.layernorm_fwd(x, g, b, y, mean, rstd)
cases = [
    ("layernorm fwd [32768x768]", lambda: ops.layernorm_fwd(x, g, b, y, mean, rstd), 2 * M * C * 2 + 8 * M),
    ("layernorm bwd [32768x768] + residual grad", lambda: ops.layernorm_bwd(dy, x, g, mean, rstd, y, dg, db, dres=dres), 4 * M * C * 2 + 8 * M),
    ("colsum [32768x768]", lambda: ops.colsum(x, dg), M * C * 2),
    ("colsum [32768x3072]", lambda: ops.colsum(x4, db4), M * 4 * C * 2),
    ("cross entropy in place [16384x50257]", lambda: ops.cross_entropy_fwd_bwd(lg, tg, rl, nv, V), 2 * rows * V * 2),
    ("adamw (163M params, both groups)", lambda: opt.step(), 30 * P),
    ("sumsq (163M grads)", lambda: opt.grad_norm_sq(), 4 * P),
    ("embedding fwd [32x1024x768]", lambda: ops.embedding_fwd(idx, model._w("wte", True), model._w("wpe", True).view(-1, C), emb), M * 8 + 2 * M * C * 2),
    ("dropout + residual [32768x768]", lambda: ops.dropout(x, dres, y, 0.1, seed, 3), 3 * M * C * 2),
    ("dropout backward + bias gradient [32768x768]", lambda: ops.dropout_bwd_colsum(x, y, 0.1, seed, 3, dg), 2 * M * C * 2),
    ("attention keep words, causal [B32 H12 T1024]", lambda: ops.attention_dropout_bits(kbits, 32, 1024, 12, 1, 0.1, seed, 0),
     2 * 32 * 12 * (1024 * 1024 // 2) // 8),
]
meta = []
for name, fn, nbytes in cases:
    fn(); fn(); torch.cuda.synchronize()
    torch.cuda.profiler.start(); fn(); torch.cuda.synchronize(); torch.cuda.profiler.stop()
    meta.append(dict(name=name, algorithmic_bytes=nbytes))
print(json.dumps(meta))
