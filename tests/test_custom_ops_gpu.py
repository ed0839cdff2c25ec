"""torch.ops.mingpt_b200.* -- the kernels registered as torch custom ops with autograd (ops.py) -- forward and
backward against the torch call each one replaces: F.linear (+ GELU / residual epilogues),
the attention core of nn.MultiheadAttention, nn.Embedding + positional add, F.cross_entropy(ignore_index=-1),
F.layer_norm. bf16 activations, fp32 bias / norm parameters; tolerance 2e-2 relative (north_star, bf16 inputs)."""
import math

import pytest
import torch
import torch.nn.functional as F

import mingpt_distributed_b200.ops  # noqa: F401  (registers torch.ops.mingpt_b200)
from conftest import report

pytestmark = pytest.mark.gpu
OPS = torch.ops.mingpt_b200


def _rel(name, got, want, bound=2e-2):
    got, want = got.float(), want.float()
    r = ((got - want).norm() / (want.norm() + 1e-12)).item()
    report("customop_" + name, rel_l2=r)
    assert r < bound, f"{name}: relative error {r:.4f}"


@pytest.mark.parametrize("act,approx", [(0, None), (1, "tanh"), (2, "none")])
@pytest.mark.parametrize("with_res", [False, True])
def test_linear_op_fwd_bwd(act, approx, with_res):
    g = torch.Generator(device="cuda").manual_seed(40 + act)
    B, T, K, N = 3, 96, 192, 320
    x = torch.randn(B, T, K, device="cuda", generator=g).bfloat16().requires_grad_(True)
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16().requires_grad_(True)
    b = torch.randn(N, device="cuda", generator=g).requires_grad_(True)
    res = torch.randn(B, T, N, device="cuda", generator=g).bfloat16().requires_grad_(True) if with_res else None
    up = torch.randn(B, T, N, device="cuda", generator=g).bfloat16()
    y, _ = OPS.linear(x, w, b, res, act)
    (y.float() * up.float()).sum().backward()
    xf, wf, bf = (t.detach().float().requires_grad_(True) for t in (x, w, b))
    z = F.linear(xf, wf, bf)
    if approx is not None:
        z = F.gelu(z, approximate=approx)
    rf = None
    if with_res:
        rf = res.detach().float().requires_grad_(True)
        z = z + rf
    (z * up.float()).sum().backward()
    tag = f"linear_a{act}_r{int(with_res)}"
    _rel(tag + "_y", y, z)
    _rel(tag + "_dx", x.grad, xf.grad)
    _rel(tag + "_dw", w.grad, wf.grad)
    _rel(tag + "_db", b.grad, bf.grad)
    if with_res:
        _rel(tag + "_dres", res.grad, rf.grad)


@pytest.mark.parametrize("d,mode", [(64, 1), (64, 2), (32, 1), (16, 2)])
def test_attention_op_fwd_bwd(d, mode):
    B, T, H = 2, 160, 3
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(50 + d)
    qkv = torch.randn(B, T, 3 * C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    up = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    out, lse = OPS.attention(qkv, H, mode)
    (out.float() * up.float()).sum().backward()
    qf = qkv.detach().float().requires_grad_(True)
    q, k, v = (t.view(B, T, H, d).transpose(1, 2) for t in qf.chunk(3, dim=-1))
    att = (q * (1.0 / math.sqrt(d))) @ k.transpose(-2, -1)
    tril = torch.tril(torch.ones(T, T, device="cuda"))
    att = att + tril if mode == 2 else att.masked_fill(tril == 0, float("-inf"))
    want = (torch.softmax(att, -1) @ v).transpose(1, 2).reshape(B, T, C)
    (want * up.float()).sum().backward()
    _rel(f"attention_d{d}_m{mode}_out", out, want)
    _rel(f"attention_d{d}_m{mode}_lse", lse, torch.logsumexp(att, -1).detach(), 1e-3)
    _rel(f"attention_d{d}_m{mode}_dqkv", qkv.grad, qf.grad)


def test_embedding_op_fwd_bwd():
    B, T, C, V, blk = 3, 40, 96, 211, 64
    g = torch.Generator(device="cuda").manual_seed(60)
    idx = torch.randint(0, V, (B, T), device="cuda", generator=g)
    wte = torch.randn(V, C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    wpe = torch.randn(blk, C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    up = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    y = OPS.embedding(idx, wte, wpe)
    (y.float() * up.float()).sum().backward()
    wf, pf = wte.detach().float().requires_grad_(True), wpe.detach().float().requires_grad_(True)
    want = F.embedding(idx, wf) + pf[:T]
    (want * up.float()).sum().backward()
    _rel("embedding_y", y, want)
    _rel("embedding_dwte", wte.grad, wf.grad)
    _rel("embedding_dwpe", wpe.grad, pf.grad)
    assert (wpe.grad[T:] == 0).all()


@pytest.mark.parametrize("rows,V", [(96, 211), (40, 50257)])
def test_cross_entropy_op_fwd_bwd(rows, V):
    g = torch.Generator(device="cuda").manual_seed(70)
    logits = (torch.randn(rows, V, device="cuda", generator=g) * 2).bfloat16().requires_grad_(True)
    tg = torch.randint(0, V, (rows,), device="cuda", generator=g)
    tg[::7] = -1
    loss, _ = OPS.cross_entropy(logits, tg)
    (loss * 3.0).backward()
    lf = logits.detach().float().requires_grad_(True)
    want = F.cross_entropy(lf, tg, ignore_index=-1)
    (want * 3.0).backward()
    report(f"customop_ce_{rows}x{V}", got=loss.item(), want=want.item())
    assert abs(loss.item() - want.item()) < 1e-3 * max(1.0, abs(want.item()))
    _rel(f"cross_entropy_{rows}x{V}_dlogits", logits.grad, lf.grad)
    assert (logits.grad[::7] == 0).all()


def test_ops_chain_like_a_block():
    """the ops compose under autograd the way a transformer block chains them: LN -> linear(qkv) -> attention ->
    linear(+residual) -> LN -> linear(gelu) -> linear(+residual) -> cross entropy"""
    B, T, H, d, V = 2, 128, 2, 64, 152       # (the linear op's backward reads dY through TMA: N must be a multiple of 8)
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(80)
    mk = lambda *s, sc=0.05: (torch.randn(*s, device="cuda", generator=g) * sc)
    x0 = mk(B, T, C, sc=1.0).bfloat16()
    P = dict(wq=mk(3 * C, C).bfloat16(), bq=mk(3 * C), wo=mk(C, C).bfloat16(), bo=mk(C), wf=mk(4 * C, C).bfloat16(),
             bf=mk(4 * C), wp=mk(C, 4 * C).bfloat16(), bp=mk(C), wh=mk(V, C).bfloat16(),
             g1=torch.ones(C, device="cuda"), b1=torch.zeros(C, device="cuda"), g2=torch.ones(C, device="cuda"), b2=torch.zeros(C, device="cuda"))
    tg = torch.randint(0, V, (B * T,), device="cuda", generator=g)

    def run(ours):
        p = {k: (v.detach().clone() if ours else v.detach().float()).requires_grad_(True) for k, v in P.items()}
        x = x0.detach().clone() if ours else x0.float()
        if ours:
            h, _, _ = OPS.layernorm(x, p["g1"], p["b1"], 1e-5)
            qkv, _ = OPS.linear(h, p["wq"], p["bq"], None, 0)
            a, _ = OPS.attention(qkv, H, 1)
            x1, _ = OPS.linear(a, p["wo"], p["bo"], x, 0)
            h2, _, _ = OPS.layernorm(x1, p["g2"], p["b2"], 1e-5)
            f, _ = OPS.linear(h2, p["wf"], p["bf"], None, 1)
            x2, _ = OPS.linear(f, p["wp"], p["bp"], x1, 0)
            lg, _ = OPS.linear(x2, p["wh"], None, None, 0)
            loss, _ = OPS.cross_entropy(lg.reshape(B * T, V), tg)
        else:
            h = F.layer_norm(x, (C,), p["g1"], p["b1"])
            qkv = F.linear(h, p["wq"], p["bq"])
            q, k, v = (t.view(B, T, H, d).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
            a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, C)
            x1 = x + F.linear(a, p["wo"], p["bo"])
            h2 = F.layer_norm(x1, (C,), p["g2"], p["b2"])
            x2 = x1 + F.linear(F.gelu(F.linear(h2, p["wf"], p["bf"]), approximate="tanh"), p["wp"], p["bp"])
            loss = F.cross_entropy(F.linear(x2, p["wh"]).reshape(B * T, V), tg)
        loss.backward()
        return loss, p

    lo, po = run(True)
    lt, pt = run(False)
    assert abs(lo.item() - lt.item()) < 1e-2 * lt.item()
    for k in P:
        _rel("chain_d" + k, po[k].grad, pt[k].grad, 3e-2)
