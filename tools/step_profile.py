#!/usr/bin/env python
"""Kernel table of one training step from torch.profiler (CUPTI activity records: no replay, warm caches) -- a cheap
companion to the ncu launch list. usage: python tools/step_profile.py [--dropout P] [--batch B] [--model gpt2]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import mingpt_distributed_b200 as m

ap = argparse.ArgumentParser()
ap.add_argument("--dropout", type=float, default=0.0)
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=1024)
ap.add_argument("--model", default="gpt2")
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
torch.manual_seed(0)
cfg = m.GPTConfig(model_type=a.model, vocab_size=50257, block_size=a.seq, embed_drop=a.dropout, resid_drop=a.dropout, attn_drop=a.dropout)
model = m.GPT(cfg, verbose=False).cuda().train()
opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=3e-4))
x = torch.randint(0, 50257, (a.batch, a.seq), device="cuda"); y = torch.randint(0, 50257, (a.batch, a.seq), device="cuda")
def step():
    loss = model.forward_backward(x, y)
    opt.step(gradnorm_sq=opt.grad_norm_sq(), max_norm=1.0)
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps): step()
    torch.cuda.synchronize()
rows = sorted(((e.device_time_total, e.count, e.key) for e in prof.key_averages() if e.device_time_total > 0), reverse=True)
tot = sum(r[0] for r in rows)
print(f"# dropout={a.dropout} batch={a.batch}: {tot / a.steps / 1e3:.2f} ms of kernel time per step")
print("| share | us/step | launches/step | avg us | kernel |\n|---:|---:|---:|---:|---|")
for t, c, k in rows[:28]:
    print(f"| {100 * t / tot:.2f}% | {t / a.steps:.0f} | {c / a.steps:.0f} | {t / c:.1f} | `{k[:70]}` |")
