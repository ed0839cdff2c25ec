"""The other configurations BASELINE.json names, at their real widths / head counts but few layers and short
sequences so the CPU oracle finishes in seconds: gpt2-medium (C=1024, H=16), gpt2-xl (C=1600, H=25: ragged
GEMM tile tails 4800 / 6400, LayerNorm width 1600), gpt-mini (d=32) -- loss and every gradient vs the oracle,
plus gradient accumulation (config 5's k micro-steps) against one big batch."""
import pytest
import torch

import mingpt_distributed_b200 as m
from conftest import report
from oracle import gpt_oracle as go

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,L,H,C,V,T,B", [("gpt2-medium", 2, 16, 1024, 1000, 192, 2),
                                               ("gpt2-xl", 2, 25, 1600, 777, 160, 2),
                                               ("gpt2-large", 1, 20, 1280, 512, 130, 1),
                                               ("gpt-mini", 6, 6, 192, 65, 256, 4)])
def test_preset_widths_vs_oracle(name, L, H, C, V, T, B):
    torch.manual_seed(3)
    cfg = dict(n_layer=L, n_head=H, n_embed=C, vocab_size=V, block_size=T)
    model = m.GPT(m.GPTConfig(model_type=None, embed_drop=0, resid_drop=0, attn_drop=0, **cfg), verbose=False)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model = model.cuda()
    g = torch.Generator().manual_seed(4)
    tok = torch.randint(0, V, (B, T + 1), generator=g)
    x, y = tok[:, :-1].contiguous(), tok[:, 1:].contiguous()
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    _, oloss = go.forward(osd, H, x, y, "mingpt")
    oloss.backward()
    _, loss = model(x.cuda(), y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert abs(loss.item() - oloss.item()) < 1e-2 * oloss.item()
    worst = 0.0
    for n, p in model.named_parameters():
        ref = osd[n].grad
        if ref.norm() < 1e-6:
            continue
        r = ((p.grad.cpu() - ref).norm() / ref.norm()).item()
        worst = max(worst, r)
        assert r < 2e-2, f"{name} {n}: relative gradient error {r:.4f}"
    report(f"preset_{name}", loss=loss.item(), oracle=oloss.item(), worst_grad_rel=worst)


def test_gradient_accumulation_equals_big_batch():
    """k micro-batches with loss/k accumulated in place == one batch of all the sequences"""
    torch.manual_seed(5)
    cfg = m.GPTConfig(model_type=None, n_layer=2, n_head=4, n_embed=256, vocab_size=300, block_size=64, embed_drop=0,
                      resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda()
    x = torch.randint(0, 300, (8, 64), device="cuda"); y = torch.randint(0, 300, (8, 64), device="cuda")
    _, loss = model(x, y); loss.backward()
    big = model._flat_grad.clone()
    for p in model.parameters():
        p.grad = None
    for k in range(4):
        _, l = model(x[2 * k:2 * k + 2], y[2 * k:2 * k + 2])
        (l / 4).backward()
    rel = ((model._flat_grad - big).norm() / big.norm()).item()
    report("grad_accumulation", rel=rel)
    assert rel < 1e-2
