#!/usr/bin/env python
"""GPU probe: step time / tokens/s / MFU of the other BASELINE.json configs on one B200 (synthetic tokens, graph replay):
gpt-mini B=64 T=256 V=65; gpt2-medium B=32 T=1024; gpt2-xl B=8 T=1024 with k=4 accumulation micro-steps (eager)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mingpt_distributed_b200 as m
dev = "cuda"
def run(name, model_type, B, T, V, accum=1, steps=6):
    torch.manual_seed(0)
    kw = dict(model_type=model_type, embed_drop=0, resid_drop=0, attn_drop=0, block_size=T)
    if V: kw["vocab_size"] = V
    cfg = m.GPTConfig(**kw)
    model = m.GPT(cfg, verbose=False).cuda()
    opt = m.create_optimizer(model, m.OptimizerConfig())
    tok = torch.randint(0, cfg.vocab_size, (B * accum, T + 1), device=dev)
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].contiguous()
    if accum == 1:
        stepper = m.GraphedTrainStep(model, opt, 1.0, None, use_graph=True, warmup=2)
        fn = lambda: stepper(x, y)
    else:
        gs = torch.full((), 1.0 / accum, device=dev)      # device scalar: d(loss)/d(micro-loss)
        def fn():
            for p in model.parameters(): p.grad = None
            loss = None
            for k in range(accum):            # k micro-batches accumulate in place (config 5), one optimizer step
                loss = model.forward_backward(x[k * B:(k + 1) * B], y[k * B:(k + 1) * B], grad_scale=gs)
            opt.step(gradnorm_sq=opt.grad_norm_sq(), max_norm=1.0)
            return loss
    for _ in range(4): l = fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps): l = fn()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / steps
    tps = B * accum * T / (ms / 1e3)
    f = 6 * (12 * cfg.n_layer * cfg.n_embed ** 2 + cfg.vocab_size * cfg.n_embed) + 6 * cfg.n_layer * cfg.n_embed * T
    print(json.dumps(dict(config=name, B=B, T=T, accum=accum, ms_per_step=round(ms, 2), tokens_per_s=round(tps), mfu_nominal=round(tps * f / 2.25e15, 4),
                          loss=round(float(l), 4), mem_gb=round(torch.cuda.max_memory_allocated() / 2**30, 1))), flush=True)
    del model, opt; torch.cuda.empty_cache(); torch.cuda.reset_peak_memory_stats()
run("gpt-mini", "gpt-mini", 64, 256, 65)
run("gpt2-medium", "gpt2-medium", 32, 1024, None)
run("gpt2-xl", "gpt2-xl", 8, 1024, None, accum=4, steps=3)
