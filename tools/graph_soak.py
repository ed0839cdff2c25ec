"""Soak test: 300 CUDA-graph replays of the gpt2-124M training step with dropout 0.1 at all four sites (the path with
the most kernels and, with MGB_DROP_SIDE_STREAM=1, a forked side stream inside the graph). Prints the loss every 50 steps;
run it under `timeout`. env: HB = batch per step (32), HN = steps (300)."""
import sys, os, time
sys.path.insert(0, "/root/repo")
import torch
import mingpt_distributed_b200 as m
torch.manual_seed(0)
cfg = m.GPTConfig(model_type="gpt2", vocab_size=50257, block_size=1024, embed_drop=0.1, resid_drop=0.1, attn_drop=0.1)
model = m.GPT(cfg, verbose=False).cuda().train()
opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=3e-4))
B = int(os.environ.get("HB", "32"))
x = torch.randint(0, 50257, (B, 1024), device="cuda"); y = torch.randint(0, 50257, (B, 1024), device="cuda")
stepper = m.GraphedTrainStep(model, opt, 1.0, None, use_graph=True)
t0 = time.time()
for i in range(int(os.environ.get("HN", "300"))):
    loss = stepper(x, y)
    if i % 50 == 0:
        print(i, float(loss), f"{time.time() - t0:.1f}s", flush=True)
torch.cuda.synchronize()
print("done", f"{time.time() - t0:.1f}s")
