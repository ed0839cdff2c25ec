#!/usr/bin/env python
"""2+ GPU check of the data-parallel step (run under torchrun on the GPU box):
 (1) after backward + bucketed all-reduce, every rank's averaged gradient equals the gradient of the
     concatenated global batch computed on one GPU;  (2) after the optimizer step all ranks hold identical
     parameters;  (3) launches overlap: all-reduces are issued from the backward hooks.
Prints one JSON line from rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import mingpt_distributed_b200 as m

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(0)
cfg = m.GPTConfig(model_type=None, n_layer=4, n_head=4, n_embed=256, vocab_size=1000, block_size=128, embed_drop=0, resid_drop=0, attn_drop=0)
model = m.GPT(cfg, verbose=False).cuda()
opt = m.create_optimizer(model, m.OptimizerConfig())
dp = m.DataParallel(model, bucket_cap_mb=1.0)
g = torch.Generator().manual_seed(7)
tok = torch.randint(0, 1000, (4 * world, 129), generator=g)
xs, ys = tok[:, :-1].contiguous().cuda(), tok[:, 1:].contiguous().cuda()
# reference: the whole global batch on this GPU (mean over all tokens == mean of per-rank means, equal sizes)
_, loss_all = model(xs, ys); loss_all.backward()
want = model._flat_grad.clone()
for p in model.parameters(): p.grad = None
# data parallel: this rank's shard, SUM all-reduce in buckets, scale 1/world folded into the optimizer
_, loss = dp(xs[4 * rank:4 * rank + 4], ys[4 * rank:4 * rank + 4]); loss.backward()
dp.finish_backward(); torch.cuda.synchronize()
got = model._flat_grad * dp.grad_scale
rel = ((got - want).norm() / want.norm()).item()
opt.step(grad_scale=dp.grad_scale, gradnorm_sq=opt.grad_norm_sq(), max_norm=1.0)
torch.cuda.synchronize()
ref = model._flat.clone(); dist.broadcast(ref, src=0)
same = bool(torch.equal(ref, model._flat))
flags = torch.tensor([rel, float(same)], device="cuda"); dist.all_reduce(flags, op=dist.ReduceOp.MAX if False else dist.ReduceOp.MIN)

# (4) the segmented CUDA-graph step (graphs cut at bucket boundaries, all-reduces launched between them) follows
#     the eager data-parallel step
def run(use_graph, steps=6):
    torch.manual_seed(0)
    mdl = m.GPT(cfg, verbose=False).cuda()
    op = m.create_optimizer(mdl, m.OptimizerConfig(learning_rate=1e-3))
    d = m.DataParallel(mdl, bucket_cap_mb=1.0)
    stepper = m.GraphedTrainStep(mdl, op, 1.0, d, use_graph=use_graph, warmup=2)
    losses = []
    for s_ in range(steps):
        gg = torch.Generator().manual_seed(100 + s_)
        tk = torch.randint(0, 1000, (4 * world, 129), generator=gg)[4 * rank:4 * rank + 4]
        losses.append(float(stepper(tk[:, :-1].contiguous().pin_memory(), tk[:, 1:].contiguous().pin_memory())))
    torch.cuda.synchronize()
    return losses, mdl._flat.clone(), stepper
le, pe, _ = run(False)
lg, pg, stp = run(True)
graph_rel = ((pg - pe).norm() / pe.norm()).item()
loss_dev = max(abs(a - b) / abs(b) for a, b in zip(lg, le))
refp = pg.clone(); dist.broadcast(refp, src=0)
graph_same = bool(torch.equal(refp, pg))
nseg = len([k for k, _ in stp._states[(4, 128)]["graph"] if k == "graph"])
if rank == 0:
    print(json.dumps(dict(world=world, graph_segments=nseg, graph_vs_eager_param_rel=graph_rel, graph_vs_eager_loss_dev=loss_dev,
                          graph_params_identical_across_ranks=graph_same)))

# (5) bf16 on the wire (grad_wire_dtype) and reserved SMs: gradient still the global-batch gradient to bf16 rounding,
#     parameters still identical on every rank; (6) k = 2 accumulation micro-steps inside the segmented graph step
model._layer_done_hook = None
for p_ in model.parameters(): p_.grad = None
_, loss_all2 = model(xs, ys); loss_all2.backward()          # the parameters have moved since `want`: a fresh reference
want = model._flat_grad.clone()
for p_ in model.parameters(): p_.grad = None
dpw = m.DataParallel(model, bucket_cap_mb=1.0, broadcast_parameters=False, grad_wire_dtype=torch.bfloat16, comm_sms=16)
_, lw = dpw(xs[4 * rank:4 * rank + 4], ys[4 * rank:4 * rank + 4]); lw.backward()
dpw.finish_backward(); torch.cuda.synchronize()
wire_rel = ((model._flat_grad * dpw.grad_scale - want).norm() / want.norm()).item()
model._layer_done_hook = None
def run_accum(k, steps=5):
    torch.manual_seed(0)
    mdl = m.GPT(cfg, verbose=False).cuda()
    op = m.create_optimizer(mdl, m.OptimizerConfig(learning_rate=1e-3))
    d = m.DataParallel(mdl, bucket_cap_mb=1.0, grad_wire_dtype=torch.bfloat16 if k > 1 else None, comm_sms=16 if k > 1 else 0)
    stepper = m.GraphedTrainStep(mdl, op, 1.0, d, use_graph=True, warmup=2, accum=k)
    losses = []
    for s_ in range(steps):
        gg = torch.Generator().manual_seed(300 + s_)
        tk = torch.randint(0, 1000, (8 * world, 129), generator=gg)[8 * rank:8 * rank + 8]
        losses.append(float(stepper(tk[:, :-1].contiguous().pin_memory(), tk[:, 1:].contiguous().pin_memory())))
    torch.cuda.synchronize()
    return losses, mdl._flat.clone()
l1, p1 = run_accum(1)
l2, p2 = run_accum(2)
acc_rel = ((p2 - p1).norm() / p1.norm()).item()
acc_loss = max(abs(a - b) / abs(b) for a, b in zip(l2, l1))
refa = p2.clone(); dist.broadcast(refa, src=0)
if rank == 0:
    print(json.dumps(dict(world=world, bf16_wire_grad_rel_err_vs_global_batch=wire_rel, accum2_bf16wire_vs_accum1_fp32_param_rel=acc_rel,
                          accum2_vs_accum1_loss_dev=acc_loss, accum2_params_identical_across_ranks=bool(torch.equal(refa, p2)))))
if rank == 0:
    print(json.dumps(dict(world=world, grad_rel_err_vs_global_batch=rel, params_identical_after_step=same,
                          buckets=len(dp.buckets.buckets), allreduces_launched=dp.buckets.launched)))
dist.destroy_process_group()
