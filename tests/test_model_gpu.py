"""End-to-end parity of the fused CUDA training step against the oracle (oracle/gpt_oracle.py) and the
committed reference fixtures (tests/golden, minted from the repaired reference): loss, logits, every
parameter gradient, the optimizer trajectory, greedy generation, the trainer front-ends, and -- at the
full GPT-2 124M shape -- size-independent properties."""
import math
import os

import pytest
import torch

import mingpt_distributed_b200 as m
from conftest import load_golden, report
from oracle import gpt_oracle as go

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _model_from(sd, cfg, variant):
    c = m.GPTConfig(model_type=None, embed_drop=0.0, resid_drop=0.0, attn_drop=0.0, variant=variant, **cfg)
    model = m.GPT(c, verbose=False)
    model.load_state_dict(sd)
    return model.cuda()


@pytest.mark.parametrize("name", ["ref_nano", "ref_d32", "ref_d64"])
def test_reference_fixture_loss_logits_grads(name):
    """variant='reference' on the reference's own weights and batch: the numbers the reference produced"""
    fx = load_golden(name)
    model = _model_from(fx["state_dict"], fx["config"], "reference")
    x, y = fx["x"].cuda(), fx["y"].cuda()
    logits, loss = model(x, y)
    assert isinstance(logits, m.LazyLogits) and tuple(logits.shape) == tuple(fx["logits"].shape)
    lg = logits.materialize()
    loss.backward()
    torch.cuda.synchronize()
    report(f"{name}_loss", got=loss.item(), want=fx["loss"])
    assert abs(loss.item() - fx["loss"]) < 1e-2 * fx["loss"]
    assert (lg.cpu() - fx["logits"]).abs().max().item() < 2e-2 * max(1.0, fx["logits"].abs().max().item())
    worst = 0.0
    for n, p in model.named_parameters():
        r = _rel(p.grad, fx["grads"][n])
        if fx["grads"][n].norm() < 1e-6:      # k-bias gradient is analytically zero
            assert p.grad.float().norm().item() < 1e-3, n
            continue
        worst = max(worst, r)
        assert r < 2e-2, f"{n}: relative gradient error {r:.4f}"
    report(f"{name}_worst_grad_rel", value=worst)


@pytest.mark.parametrize("variant", ["mingpt", "reference"])
@pytest.mark.parametrize("shape", [(3, 3, 48, 65, 128, 4, 128), (2, 4, 128, 211, 64, 3, 50), (2, 2, 128, 300, 256, 2, 256)])
def test_forward_backward_vs_oracle(variant, shape):
    L, H, C, V, blk, B, T = shape
    torch.manual_seed(1)
    cfg = dict(n_layer=L, n_head=H, n_embed=C, vocab_size=V, block_size=blk)
    model = m.GPT(m.GPTConfig(model_type=None, embed_drop=0, resid_drop=0, attn_drop=0, variant=variant, **cfg), verbose=False)
    with torch.no_grad():      # make biases / norms non-trivial
        for p in model.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model = model.cuda()
    g = torch.Generator().manual_seed(2)
    tok = torch.randint(0, V, (B, T + 1), generator=g)
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].clone()
    y[torch.rand(B, T, generator=g) < 0.1] = -1
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ol, oloss = go.forward(osd, H, x, y, variant)
    oloss.backward()
    logits, loss = model(x.cuda(), y.cuda())
    lg = logits.materialize()
    loss.backward()
    torch.cuda.synchronize()
    tag = f"{variant}_L{L}C{C}T{T}"
    report(tag + "_loss", got=loss.item(), want=oloss.item())
    assert abs(loss.item() - oloss.item()) < 1e-2 * oloss.item()
    assert (lg.cpu() - ol.detach()).abs().max().item() < 3e-2 * max(1.0, ol.abs().max().item())
    worst = 0.0
    for n, p in model.named_parameters():
        ref = osd[n].grad
        if ref.norm() < 1e-6:
            continue
        r = _rel(p.grad, ref)
        worst = max(worst, r)
        assert r < 2e-2, f"{n}: relative gradient error {r:.4f}"
    report(tag + "_worst_grad_rel", value=worst)
    # inference path: full fp32 logits, no loss
    lg2, none = model(x.cuda())
    assert none is None and (lg2.cpu() - ol.detach()).abs().max().item() < 3e-2 * max(1.0, ol.abs().max().item())
    # eval loss without gradients
    with torch.no_grad():
        _, l3 = model(x.cuda(), y.cuda())
    assert abs(l3.item() - oloss.item()) < 1e-2 * oloss.item()


def test_gradient_accumulation_and_zero_grad_semantics():
    torch.manual_seed(3)
    cfg = m.GPTConfig(model_type="gpt-nano", vocab_size=65, block_size=64, embed_drop=0, resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda()
    x = torch.randint(0, 65, (4, 64), device="cuda"); y = torch.randint(0, 65, (4, 64), device="cuda")
    _, loss = model(x, y); loss.backward()
    g1 = {n: p.grad.clone() for n, p in model.named_parameters()}
    _, loss = model(x, y); (loss * 0.5).backward()           # accumulates 0.5x on top
    for n, p in model.named_parameters():
        assert torch.allclose(p.grad, 1.5 * g1[n], rtol=2e-2, atol=1e-5 * (1 + g1[n].abs().max().item())), n
    for p in model.parameters():
        p.grad = None                                        # optimizer.zero_grad(set_to_none=True)
    _, loss = model(x, y); loss.backward()
    for n, p in model.named_parameters():
        assert torch.allclose(p.grad, g1[n], rtol=2e-2, atol=1e-5 * (1 + g1[n].abs().max().item())), n
    # all targets ignored -> zero loss, zero gradients
    for p in model.parameters():
        p.grad = None
    _, loss = model(x, torch.full_like(y, -1)); loss.backward()
    assert loss.item() == 0.0 and all(p.grad.abs().max().item() == 0 for p in model.parameters())


def _periodic_batch(step, B, T, V):
    g = torch.Generator().manual_seed(1000 + step)
    start = torch.randint(0, V, (B, 1), generator=g)
    seq = (start + torch.arange(T + 1)[None, :] * 3) % V      # learnable: next = cur + 3 mod V
    return seq[:, :-1].contiguous(), seq[:, 1:].contiguous()


@pytest.mark.parametrize("variant", ["reference", "mingpt"])
def test_loss_trajectory_200_steps_and_greedy_generate(variant):
    """north_star: loss trajectory within 1 % over 200 steps from identical init; greedy tokens bit-exact"""
    L, H, C, V, blk, B, T = 3, 3, 48, 65, 64, 8, 64
    torch.manual_seed(0)
    model = m.GPT(m.GPTConfig(model_type=None, n_layer=L, n_head=H, n_embed=C, vocab_size=V, block_size=blk,
                              embed_drop=0, resid_drop=0, attn_drop=0, variant=variant), verbose=False)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=1e-3))
    model = model.cuda()
    oracle = go.OracleTrainer(sd, H, variant, lr=1e-3)
    ours, theirs = [], []
    for step in range(200):
        x, y = _periodic_batch(step, B, T, V)
        theirs.append(oracle.run_batch(x, y))
        ours.append(m.train_step(model, opt, x.cuda(), y.cuda(), 1.0, zero_grad=True))
    ours = [float(t) for t in ours]
    dev = max(abs(a - b) / max(abs(b), 0.05) for a, b in zip(ours, theirs))
    report(f"traj200_{variant}", max_rel_dev=dev, first=ours[0], last=ours[-1], oracle_last=theirs[-1])
    assert theirs[-1] < 0.5 * theirs[0], "the oracle itself must learn the pattern"
    assert dev < 1e-2, f"trajectory deviates {dev:.4f}"        # measured: 0.2 % for both variants
    # greedy continuation of a trained (peaked) model: tokens must be identical
    prompt, _ = _periodic_batch(9999, 4, 8, V)
    osd = {k: v.detach() for k, v in oracle.sd.items()}
    idx = prompt.clone()
    for _ in range(24):
        lg, _ = go.forward(osd, H, idx[:, -blk:], None, variant)
        idx = torch.cat([idx, lg[:, -1].argmax(-1, keepdim=True)], 1)
    got = model.generate(prompt.cuda(), 24).cpu()
    report(f"generate_{variant}", match=float((got == idx).float().mean()))
    assert torch.equal(got, idx), f"greedy tokens differ from the oracle's ({variant})"
    # KV-cached incremental decode (variant "mingpt"; the as-written graph is not causal and never caches) vs the
    # full-window recompute the reference does, and across the block_size boundary where the window starts to slide
    uncached = model.generate(prompt.cuda(), 24, use_cache=False).cpu()
    assert torch.equal(uncached, idx)
    long_prompt, _ = _periodic_batch(4242, 3, blk - 5, V)
    idx2 = long_prompt.clone()
    for _ in range(12):
        lg, _ = go.forward(osd, H, idx2[:, -blk:], None, variant)
        idx2 = torch.cat([idx2, lg[:, -1].argmax(-1, keepdim=True)], 1)
    a = model.generate(long_prompt.cuda(), 12).cpu()
    b = model.generate(long_prompt.cuda(), 12, use_cache=False).cpu()
    report(f"generate_sliding_{variant}", cached_match=float((a == idx2).float().mean()), uncached_match=float((b == idx2).float().mean()))
    assert torch.equal(a, idx2) and torch.equal(b, idx2)


def test_fixture_trajectory_and_generate_reference_variant():
    """the reference's own 10-step trajectory (mingpt/trainer.py:118-133 restated) and generate() output"""
    fx = load_golden("ref_nano")
    cfg = fx["config"]
    model = m.GPT(m.GPTConfig(model_type=None, embed_drop=0, resid_drop=0, attn_drop=0, variant="reference", **cfg), verbose=False)
    model.load_state_dict(fx["state_dict"])
    opt = m.create_optimizer(model, m.OptimizerConfig())
    model = model.cuda()
    B, T = fx["x"].shape
    losses = []
    for step in range(10):
        g = torch.Generator().manual_seed(100 + step)
        tk = torch.randint(0, cfg["vocab_size"], (B, T + 1), generator=g)
        losses.append(float(m.train_step(model, opt, tk[:, :-1].contiguous().cuda(), tk[:, 1:].contiguous().cuda(), 1.0, zero_grad=True)))
    dev = max(abs(a - b) / b for a, b in zip(losses, fx["traj"]))
    report("fixture_traj10", max_rel_dev=dev)
    assert dev < 1e-2
    C = cfg["n_embed"]
    worst = 0.0
    for n, p in model.named_parameters():
        want, got = fx["state_dict_after"][n], p.detach().cpu()
        if n.endswith("in_proj_bias"):
            # the key bias has an analytically ZERO gradient (softmax is shift invariant), so Adam turns pure
            # rounding noise into +-lr steps there, in the reference as much as here: compare q and v parts
            want, got = torch.cat([want[:C], want[2 * C:]]), torch.cat([got[:C], got[2 * C:]])
        if want.norm() > 0:
            r = _rel(got, want)
            report("fixture_after10_" + n, rel=r, norm=want.norm().item())
            # tensors that start at ZERO (biases, pos_embedding) are nothing but 10 Adam steps of size ~lr:
            # m / sqrt(v) turns the bf16 noise of small gradient entries into O(lr) differences
            starts_at_zero = fx["state_dict"][n].abs().max().item() == 0
            assert r < (0.12 if starts_at_zero else 2e-2), f"{n}: {r:.4f}"
            worst = max(worst, r)
    report("fixture_params_after10", worst_rel=worst)
    out = model.generate(fx["prompt"].cuda(), 12)
    assert out.shape == fx["generated"].shape and torch.equal(out[:, :8].cpu(), fx["prompt"])
    # The reference's own generate() output (fx["generated"], from ITS weights after the 10 steps). Ten steps on
    # random tokens leave the logits nearly flat (top-2 margins down to 1e-4), so a free-running comparison would
    # only measure rounding luck. Teacher-forced instead: load the reference's weights, feed the reference's own
    # prefix, and demand the same greedy token wherever the fp32 top-2 margin exceeds the bf16 error of a logit.
    ref_model = _model_from(fx["state_dict_after"], cfg, "reference")
    sd_after = {k: v for k, v in fx["state_dict_after"].items()}
    gen = fx["generated"]
    decided = agree = 0
    for t in range(8, gen.shape[1]):
        prefix = gen[:, :t]
        ol, _ = go.forward(sd_after, cfg["n_head"], prefix, None, "reference")
        top2 = ol[:, -1].topk(2, dim=-1).values
        margin = top2[:, 0] - top2[:, 1]
        assert torch.equal(ol[:, -1].argmax(-1), gen[:, t]), "the oracle itself must reproduce the reference's generate()"
        ours = ref_model.generate(prefix.cuda(), 1)[:, -1].cpu()
        tol = 2e-2 * max(1.0, ol[:, -1].abs().max().item())
        sure = margin > tol
        decided += int(sure.sum()); agree += int((ours[sure] == gen[:, t][sure]).sum())
        assert torch.equal(ours[sure], gen[:, t][sure]), f"step {t}: greedy token differs from the reference's generate()"
    report("fixture_generate_teacher_forced", decided=decided, agree=agree, total=int(gen[:, 8:].numel()))


def test_gpttrainer_end_to_end_with_chardataset(tmp_path, monkeypatch):
    """config 1/2 path: CharDataset -> GPTTrainer.train() (mingpt/train.py:11-58 flow), snapshot + resume"""
    g = torch.Generator().manual_seed(7)
    text = bytes((97 + (i * 7 + i // 5) % 26) for i in range(6000))
    path = tmp_path / "corpus.txt"
    path.write_bytes(text)
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("LOCAL_RANK", "0"); monkeypatch.setenv("RANK", "0")
    ds = m.CharDataset(m.DataConfig(path=str(path), block_size=32, train_split=0.9, truncate=1.0))
    gcfg = m.GPTConfig(model_type="gpt-nano", vocab_size=ds.vocab_size, block_size=32, embed_drop=0, resid_drop=0, attn_drop=0)
    model = m.GPT(gcfg, verbose=False)
    opt = m.create_optimizer(model, m.OptimizerConfig(learning_rate=3e-3))
    tcfg = m.GPTTrainerConfig(max_epochs=1, batch_size=64, grad_norm_clip=1.0, dl_num_workers=0,
                              snapshot_path=str(tmp_path / "snap.pt"), save_every=1)
    tr = m.GPTTrainer(tcfg, model, opt, ds)
    x0, y0 = ds[0]
    with torch.no_grad():
        _, l0 = tr.model(x0[None].cuda(), y0[None].cuda())
    tr.train()
    # the reference-API trainer runs the benchmarked path: every full batch after the 2 warm-up steps is a graph replay
    n_full = len(ds) // 64
    assert tr.graph_replays >= n_full - 2 > 10, (tr.graph_replays, n_full)
    with torch.no_grad():
        _, l1 = tr.model(x0[None].cuda(), y0[None].cuda())
    report("gpttrainer", loss_before=l0.item(), loss_after=l1.item())
    assert l1.item() < 0.7 * l0.item()
    snap = torch.load(tmp_path / "snap.pt", map_location="cpu", weights_only=False)
    assert set(snap) == {"model_state", "optimizer_state", "final_epoch"} and snap["final_epoch"] == 0
    model2 = m.GPT(gcfg, verbose=False)
    opt2 = m.create_optimizer(model2, m.OptimizerConfig(learning_rate=3e-3))
    tr2 = m.GPTTrainer(tcfg, model2, opt2, ds)           # resumes from the snapshot
    assert tr2.last_epoch == 0
    for (n, a), (_, b) in zip(tr.raw_model.named_parameters(), tr2.raw_model.named_parameters()):
        assert torch.equal(a.detach().cpu(), b.detach().cpu()), n
    # README-style Trainer.run with a callback
    cfg = m.Trainer.get_default_config(); cfg.max_iters = 5; cfg.batch_size = 16; cfg.num_workers = 0
    seen = []
    t3 = m.Trainer(cfg, model2, ds)
    t3.set_callback("on_batch_end", lambda t: seen.append(float(t.loss)))
    t3.run()
    assert len(seen) == 5 and all(math.isfinite(v) for v in seen)
    assert t3.graph_replays == 3             # 2 eager warm-up steps, then the captured graph


def test_gpt2_124m_full_size_properties():
    """BASELINE config 3 shape (vocab 50257, T = 1024, 12 x 12 x 768): initial loss = ln(V), gradient
    linearity in the upstream gradient, finite gradients everywhere, one optimizer step lowers the loss."""
    torch.manual_seed(0)
    cfg = m.GPTConfig(model_type="gpt2", embed_drop=0, resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda()
    assert sum(p.numel() for p in model.parameters()) == 163037184       # SURVEY section 8 (mingpt variant)
    opt = m.create_optimizer(model, m.OptimizerConfig())
    g = torch.Generator().manual_seed(1234)
    tok = torch.randint(0, 50257, (4, 1025), generator=g).cuda()
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].contiguous()
    _, loss = model(x, y)
    # logits at init have std ~ 0.02 * sqrt(768) = 0.55, so E[loss] = ln V + sigma^2 / 2 ~ 10.82 + 0.15
    assert abs(loss.item() - (math.log(50257) + 0.15)) < 0.1
    loss.backward()
    g1 = model._flat_grad.clone()
    assert torch.isfinite(g1).all() and g1.abs().max().item() > 0
    for p in model.parameters():
        p.grad = None
    _, loss2 = model(x, y)
    (loss2 * 0.5).backward()
    rel = ((model._flat_grad - 0.5 * g1).norm() / (0.5 * g1).norm()).item()
    report("gpt2_linearity", rel=rel, loss=loss.item())
    assert rel < 2e-2        # only fp32-atomic summation order differs
    for p in model.parameters():
        p.grad = None
    l_prev = float(m.train_step(model, opt, x, y, 1.0, zero_grad=True))
    for _ in range(3):
        l_now = float(m.train_step(model, opt, x, y, 1.0, zero_grad=True))
    assert l_now < l_prev
