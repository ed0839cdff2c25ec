#!/usr/bin/env python
"""Kernel table of one training step from torch.profiler (CUPTI activity records: no replay, warm caches) -- a cheap
companion to the ncu launch list. usage: python tools/step_profile.py [--dropout P] [--batch B] [--model gpt2]"""
import argparse, os, sys
os.environ["MGB_DROP_SIDE_STREAM"] = "0"     # torch.profiler hangs on replayed graphs with parallel branches
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
ap.add_argument("--graph", action="store_true", help="replay the step as a CUDA graph (what bench.py times)")
a = ap.parse_args()
torch.manual_seed(0)
cfg = m.GPTConfig(model_type=a.model, vocab_size=50257, block_size=a.seq, embed_drop=a.dropout, resid_drop=a.dropout, attn_drop=a.dropout)
model = m.GPT(cfg, verbose=False).cuda().train()
opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=3e-4))
x = torch.randint(0, 50257, (a.batch, a.seq), device="cuda"); y = torch.randint(0, 50257, (a.batch, a.seq), device="cuda")
if a.graph:
    stepper = m.GraphedTrainStep(model, opt, 1.0, None, use_graph=True)
    def step():
        stepper(x, y)
else:
    def step():
        loss = model.forward_backward(x, y)
        opt.step(gradnorm_sq=opt.grad_norm_sq(), max_norm=1.0)
for _ in range(5): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): step()
e1.record(); torch.cuda.synchronize()
print(f"# {'graph' if a.graph else 'eager'}: {e0.elapsed_time(e1) / 10:.2f} ms per step by CUDA events")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps): step()
    torch.cuda.synchronize()
rows = sorted(((e.device_time_total, e.count, e.key) for e in prof.key_averages() if e.device_time_total > 0), reverse=True)
tot = sum(r[0] for r in rows)
print(f"# dropout={a.dropout} batch={a.batch}: {tot / a.steps / 1e3:.2f} ms of kernel time per step")
print("| share | us/step | launches/step | avg us | kernel |\n|---:|---:|---:|---:|---|")
for t, c, k in rows[:28]:
    print(f"| {100 * t / tot:.2f}% | {t / a.steps:.0f} | {c / a.steps:.0f} | {t / c:.1f} | `{k[:70]}` |")
# idle time in front of each kernel (start minus the previous kernel's end), summed by kernel name
try:
    from torch.autograd import DeviceType
    evs = sorted((e for e in prof.events() if e.device_type == DeviceType.CUDA), key=lambda e: e.time_range.start)
    gaps = {}
    for prev, cur in zip(evs, evs[1:]):
        g = cur.time_range.start - prev.time_range.end
        if 0 < g < 200:         # (the pause between steps is not a gap inside the step)
            k = cur.name[:60]
            t = gaps.setdefault(k, [0.0, 0])
            t[0] += g; t[1] += 1
    tg = sum(v[0] for v in gaps.values())
    print(f"# idle gaps inside the step: {tg / a.steps:.0f} us per step")
    for k, (t, c) in sorted(gaps.items(), key=lambda kv: -kv[1][0])[:12]:
        print(f"#   {t / a.steps:7.0f} us/step  {c / a.steps:5.0f}x  avg {t / c:5.1f} us  before `{k}`")
except Exception as ex:
    print("# gap analysis unavailable:", ex)
