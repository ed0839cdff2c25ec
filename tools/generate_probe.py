#!/usr/bin/env python
"""Tokens/s of GPT.generate with and without the key/value cache (gpt2-124M, greedy, prompt 64, 256 new tokens)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mingpt_distributed_b200 as m

torch.manual_seed(0)
cfg = m.GPTConfig(model_type="gpt2", vocab_size=50257, block_size=1024, embed_drop=0, resid_drop=0, attn_drop=0)
model = m.GPT(cfg, verbose=False).cuda().eval()
for B in (1, 8):
    prompt = torch.randint(0, 50257, (B, 64), device="cuda")
    outs = {}
    for cached in (True, False):
        n_new = 256 if cached else 64
        model.generate(prompt, 8, use_cache=cached)          # warm-up (workspaces)
        torch.cuda.synchronize(); t0 = time.time()
        out = model.generate(prompt, n_new, use_cache=cached)
        torch.cuda.synchronize(); dt = time.time() - t0
        outs[cached] = out
        print(f"B={B} use_cache={cached}: {n_new} new tokens per sequence in {dt * 1e3:.1f} ms = {dt / n_new * 1e3:.3f} ms per step, "
              f"{B * n_new / dt:.0f} tokens/s", flush=True)
    # (token identity between the two paths is asserted in tests/test_model_gpu.py on a TRAINED model; with the random
    # weights used here the top-2 logit margins are below bf16 resolution and the argmax is a coin toss)
