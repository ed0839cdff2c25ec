"""Per-op parity of the HBM-bound kernels and the fused attention against the torch call each replaces
(SURVEY.md 2b), on the same bf16-rounded inputs, through the C ABI. Tolerances (north_star): 2e-2 relative
for bf16 results, 1e-3 for fp32 statistics / reductions."""
import math

import pytest
import torch
import torch.nn.functional as F

from conftest import report

pytestmark = pytest.mark.gpu


def _close(name, got, want, rtol):
    """|got - want| <= rtol * (|want| + 0.1 * max|want|): relative, with a floor for near-zero entries"""
    got, want = got.float(), want.float()
    scale = want.abs().max().item() + 1e-12
    err = (got - want).abs()
    tol = rtol * (want.abs() + 0.1 * scale)
    bad = (err > tol).nonzero()
    report(name, max_abs_err=err.max().item(), scale=scale, bad=int(bad.shape[0]))
    assert bad.shape[0] == 0, f"{name}: {bad.shape[0]}/{got.numel()} off, max err {err.max().item():.4g} scale {scale:.3g} first {bad[:4].tolist()}"


# ------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,cols", [(512, 768), (33, 48), (257, 192), (64, 1600), (128, 1024), (5, 50), (1, 8)])
def test_layernorm_fwd_bwd(rows, cols):
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    x = (torch.randn(rows, cols, device="cuda", generator=g) * 2 + 0.5).bfloat16()
    gamma = torch.randn(cols, device="cuda", generator=g)
    beta = torch.randn(cols, device="cuda", generator=g)
    dy = torch.randn(rows, cols, device="cuda", generator=g).bfloat16()
    dres = torch.randn(rows, cols, device="cuda", generator=g).bfloat16()
    y = torch.empty_like(x); mean = torch.empty(rows, device="cuda"); rstd = torch.empty(rows, device="cuda")
    ops.layernorm_fwd(x, gamma, beta, y, mean, rstd)
    xf = x.float().requires_grad_(True); gf = gamma.clone().requires_grad_(True); bf = beta.clone().requires_grad_(True)
    want = F.layer_norm(xf, (cols,), gf, bf, 1e-5)
    want.backward(dy.float())
    _close(f"ln_fwd_{rows}x{cols}", y, want, 2e-2)
    _close(f"ln_mean_{rows}x{cols}", mean, x.float().mean(1), 1e-3)
    _close(f"ln_rstd_{rows}x{cols}", rstd, (x.float().var(1, unbiased=False) + 1e-5).rsqrt(), 1e-3)
    dx = torch.empty_like(x); dg = torch.zeros(cols, device="cuda"); db = torch.zeros(cols, device="cuda")
    ops.layernorm_bwd(dy, x, gamma, mean, rstd, dx, dg, db, dres=dres)
    torch.cuda.synchronize()
    _close(f"ln_dx_{rows}x{cols}", dx, xf.grad + dres.float(), 2e-2)
    _close(f"ln_dgamma_{rows}x{cols}", dg, gf.grad, 1e-3)
    _close(f"ln_dbeta_{rows}x{cols}", db, bf.grad, 1e-3)
    # accumulation into dgamma / dbeta, and dx aliasing dres
    ops.layernorm_bwd(dy, x, gamma, mean, rstd, dres, dg, db, dres=dres)
    torch.cuda.synchronize()
    _close(f"ln_dgamma_acc_{rows}x{cols}", dg, 2 * gf.grad, 1e-3)
    _close(f"ln_dx_alias_{rows}x{cols}", dres, dx, 1e-6)
    # fused column sum of dres (the bias gradient of the layer above): accumulates, leaves everything else unchanged
    dres2 = torch.randn(rows, cols, device="cuda", generator=g).bfloat16()
    dx2 = torch.empty_like(x); dx3 = torch.empty_like(x)
    dg2 = torch.zeros(cols, device="cuda"); db2 = torch.zeros(cols, device="cuda"); cs = torch.ones(cols, device="cuda")
    ops.layernorm_bwd(dy, x, gamma, mean, rstd, dx2, dg2, db2, dres=dres2, dres_colsum=cs)
    ops.layernorm_bwd(dy, x, gamma, mean, rstd, dx3, torch.zeros_like(dg2), torch.zeros_like(db2), dres=dres2)
    torch.cuda.synchronize()
    _close(f"ln_dres_colsum_{rows}x{cols}", cs, 1 + dres2.float().sum(0), 1e-3)
    assert torch.equal(dx2, dx3)
    _close(f"ln_dgamma_colsum_{rows}x{cols}", dg2, gf.grad, 1e-3)


def test_layernorm_custom_op_autograd():
    import mingpt_distributed_b200.ops  # registers torch.ops.mingpt_b200
    x = torch.randn(4, 16, 64, device="cuda").bfloat16().requires_grad_(True)
    g = torch.ones(64, device="cuda", requires_grad=True); b = torch.zeros(64, device="cuda", requires_grad=True)
    w = torch.randn(4, 16, 64, device="cuda").bfloat16()
    y, _, _ = torch.ops.mingpt_b200.layernorm(x, g, b, 1e-5)
    (y.float() * w.float()).sum().backward()
    xf = x.detach().float().requires_grad_(True)
    gf = g.detach().clone().requires_grad_(True)
    (F.layer_norm(xf, (64,), gf, b.detach()) * w.float()).sum().backward()
    _close("ln_customop_dx", x.grad, xf.grad, 3e-2)
    _close("ln_customop_dgamma", g.grad, gf.grad, 1e-2)


# ------------------------------------------------------------------ embedding
@pytest.mark.parametrize("B,T,C,V", [(4, 64, 48, 65), (2, 1024, 768, 50257), (3, 17, 50, 11)])
def test_embedding_fwd_bwd(B, T, C, V):
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(1)
    idx = torch.randint(0, V, (B, T), device="cuda", generator=g)
    wte = torch.randn(V, C, device="cuda", generator=g).bfloat16()
    wpe = torch.randn(T + 3, C, device="cuda", generator=g).bfloat16()
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.embedding_fwd(idx, wte, wpe, out, flag)
    want = wte.float()[idx] + wpe.float()[:T]
    _close(f"emb_fwd_{B}x{T}x{C}", out, want, 1e-2)
    assert flag.item() == 0
    dx = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    dwte = torch.zeros(V, C, device="cuda"); dwpe = torch.zeros(T + 3, C, device="cuda")
    ops.embedding_bwd(idx, dx, dwte, dwpe[:T])
    torch.cuda.synchronize()
    ref_wte = torch.zeros(V, C, device="cuda").index_add_(0, idx.reshape(-1), dx.float().reshape(-1, C))
    _close(f"emb_dwte_{B}x{T}x{C}", dwte, ref_wte, 1e-3)
    _close(f"emb_dwpe_{B}x{T}x{C}", dwpe[:T], dx.float().sum(0), 1e-3)
    assert (dwpe[T:] == 0).all()
    bad = idx.clone(); bad[0, 0] = V
    ops.embedding_fwd(bad, wte, wpe, out, flag)
    assert flag.item() == 1 and (out[0, 0] == 0).all()


# ------------------------------------------------------------------ attention
def _attn_ref(qkv, H, mode):
    B, T, C3 = qkv.shape
    C = C3 // 3
    d = C // H
    q, k, v = (t.view(B, T, H, d).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
    att = (q * (1.0 / math.sqrt(d))) @ k.transpose(-2, -1)
    tril = torch.tril(torch.ones(T, T, device=qkv.device))
    if mode == 2:
        att = att + tril          # the reference's float mask is ADDED (torch/nn/functional.py:6638-6643)
    elif mode == 1:
        att = att.masked_fill(tril == 0, float("-inf"))
    lse = torch.logsumexp(att, dim=-1)
    out = (torch.softmax(att, dim=-1) @ v).transpose(1, 2).reshape(B, T, C)
    return out, lse


@pytest.mark.parametrize("d", [16, 32, 64])
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("B,T,H", [(2, 128, 3), (1, 200, 2), (2, 64, 1), (1, 37, 2)])
def test_attention_fwd_bwd(d, mode, B, T, H):
    from mingpt_distributed_b200 import ops
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(2)
    qkv = torch.randn(B, T, 3 * C, device="cuda", generator=g).bfloat16()
    dout = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda")
    lse = torch.empty(B, H, T, device="cuda")
    ops.attention_fwd(qkv, out, lse, B, T, H, d, mode)
    qf = qkv.float().requires_grad_(True)
    want, want_lse = _attn_ref(qf, H, mode)
    want.backward(dout.float())
    tag = f"d{d}_m{mode}_{B}x{T}x{H}"
    _close("attn_out_" + tag, out, want, 2e-2)
    _close("attn_lse_" + tag, lse, want_lse, 1e-3)
    dqkv = torch.full_like(qkv, float("nan"))
    delta = torch.empty(B, H, T, device="cuda"); dq_acc = torch.empty(B, T, C, device="cuda")
    ops.attention_bwd(qkv, out, dout, lse, delta, dq_acc, dqkv, B, T, H, d, mode)
    torch.cuda.synchronize()
    assert torch.isfinite(dqkv.float()).all()
    _close("attn_dqkv_" + tag, dqkv, qf.grad, 3e-2)
    # the same call with the qkv bias gradient (column sums of dqkv) produced on the way: accumulates, dqkv unchanged
    dqkv2 = torch.full_like(qkv, float("nan")); cs = torch.ones(3 * C, device="cuda")
    ops.attention_bwd(qkv, out, dout, lse, delta, dq_acc, dqkv2, B, T, H, d, mode, dqkv_colsum=cs)
    torch.cuda.synchronize()
    # (not bit-equal: dQ is summed with fp32 atomics, whose order -- and so the last bf16 bit -- changes run to run)
    assert torch.allclose(dqkv2.float(), dqkv.float(), rtol=1.6e-2, atol=1e-3 * dqkv.float().abs().max().item())
    want_cs = qf.grad.reshape(B * T, 3 * C).sum(0)
    assert (cs - 1 - want_cs).abs().max().item() <= 2e-2 * want_cs.abs().max().item() + 2e-2, tag
    # delta handed in by the caller (the out-projection's dgrad GEMM computes it in its epilogue): same result
    want_delta = (out.float() * dout.float()).view(B, T, H, d).sum(-1).transpose(1, 2).contiguous()
    _close("attn_delta_" + tag, delta, want_delta, 1e-3)
    dqkv3 = torch.full_like(qkv, float("nan"))
    ops.attention_bwd(qkv, out, dout, lse, want_delta, dq_acc, dqkv3, B, T, H, d, mode, delta_given=True)
    torch.cuda.synchronize()
    assert torch.allclose(dqkv3.float(), dqkv.float(), rtol=1.6e-2, atol=1e-3 * dqkv.float().abs().max().item())


def test_attention_gpt2_shape_property():
    """full-size shape (T=1024, H=12, d=64): rows of the causal output only depend on the past"""
    from mingpt_distributed_b200 import ops
    B, T, H, d = 2, 1024, 12, 64
    C = H * d
    qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B, H, T, device="cuda")
    ops.attention_fwd(qkv, out, lse, B, T, H, d, 1)
    qkv2 = qkv.clone(); qkv2[:, 700:] = torch.randn_like(qkv2[:, 700:])
    out2 = torch.empty_like(out)
    ops.attention_fwd(qkv2, out2, lse, B, T, H, d, 1)
    torch.cuda.synchronize()
    assert torch.equal(out[:, :700], out2[:, :700]) and not torch.equal(out[:, 700:], out2[:, 700:])
    want, _ = _attn_ref(qkv[:1].float(), H, 1)
    _close("attn_gpt2_shape", out[:1], want, 2e-2)


# ------------------------------------------------------------------ cross entropy
# the last three shapes have enough rows for the shared-memory-staged kernel (one persistent CTA per SM)
@pytest.mark.parametrize("rows,V,ld", [(64, 65, 72), (33, 50257, 50304), (128, 1000, 1000), (7, 130, 131),
                                       (1200, 1000, 1000), (515, 50257, 50304), (700, 2003, 2008)])
def test_cross_entropy_in_place(rows, V, ld):
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    logits = torch.zeros(rows, ld, device="cuda").bfloat16()
    logits[:, :V] = (torch.randn(rows, V, device="cuda", generator=g) * 3).bfloat16()
    tg = torch.randint(0, V, (rows,), device="cuda", generator=g)
    tg[::5] = -1
    lf = logits[:, :V].float().requires_grad_(True)
    want = F.cross_entropy(lf, tg, ignore_index=-1)
    want.backward()
    row_loss = torch.empty(rows, device="cuda"); nv = torch.zeros(1, dtype=torch.int32, device="cuda")
    loss = torch.empty((), device="cuda")
    ops.count_valid_targets(tg, nv)
    keep = logits.clone()
    ops.cross_entropy_fwd_bwd(logits, tg, row_loss, nv, V, write_grad=False)
    assert torch.equal(logits, keep)
    ops.cross_entropy_fwd_bwd(logits, tg, row_loss, nv, V, grad_scale=1.0, write_grad=True)
    ops.loss_finalize(row_loss, nv, loss)
    torch.cuda.synchronize()
    assert nv.item() == int((tg != -1).sum())
    assert abs(loss.item() - want.item()) < 1e-3 * max(1.0, abs(want.item()))
    _close(f"ce_dlogits_{rows}x{V}", logits[:, :V], lf.grad, 2e-2)
    assert (logits[:, V:] == 0).all() and (logits[::5] == 0).all()
    report(f"ce_loss_{rows}x{V}", got=loss.item(), want=want.item())


def test_cross_entropy_staged_matches_streaming():
    """the two kernels (mgb_set_option "ce_unstaged") agree to bf16 rounding of the stored exponentials"""
    from mingpt_distributed_b200 import ops, _lib
    g = torch.Generator(device="cuda").manual_seed(13)
    rows, V, ld = 600, 50257, 50304
    base = torch.zeros(rows, ld, device="cuda").bfloat16()
    base[:, :V] = (torch.randn(rows, V, device="cuda", generator=g) * 2).bfloat16()
    tg = torch.randint(0, V, (rows,), device="cuda", generator=g); tg[::7] = -1
    nv = torch.zeros(1, dtype=torch.int32, device="cuda"); ops.count_valid_targets(tg, nv)
    outs = []
    for unstaged in (0, 1):
        prev = _lib.set_option("ce_unstaged", unstaged)
        lg = base.clone(); rl = torch.empty(rows, device="cuda")
        ops.cross_entropy_fwd_bwd(lg, tg, rl, nv, V, grad_scale=1.0, write_grad=True)
        torch.cuda.synchronize()
        _lib.set_option("ce_unstaged", prev)
        outs.append((lg.float(), rl))
    assert torch.allclose(outs[0][1], outs[1][1], rtol=1e-5, atol=1e-5)
    scale = outs[1][0].abs().max().item()
    assert (outs[0][0] - outs[1][0]).abs().max().item() <= 1e-2 * scale


def test_cross_entropy_all_ignored_and_bad_target():
    from mingpt_distributed_b200 import ops
    logits = torch.randn(8, 64, device="cuda").bfloat16()
    tg = torch.full((8,), -1, device="cuda", dtype=torch.long)
    row_loss = torch.empty(8, device="cuda"); nv = torch.zeros(1, dtype=torch.int32, device="cuda"); loss = torch.empty((), device="cuda")
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    ops.count_valid_targets(tg, nv)
    ops.cross_entropy_fwd_bwd(logits, tg, row_loss, nv, 64, err_flag=flag)
    ops.loss_finalize(row_loss, nv, loss)
    assert loss.item() == 0.0 and (logits == 0).all() and flag.item() == 0
    tg[3] = 64
    ops.cross_entropy_fwd_bwd(logits, tg, row_loss, nv, 64, err_flag=flag)
    assert flag.item() == 2


# ------------------------------------------------------------------ small helpers
def test_colsum_act_cast_axpy_sumsq():
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(4)
    for rows, cols in [(1000, 768), (257, 65), (4096, 3072)]:
        x = torch.randn(rows, cols, device="cuda", generator=g).bfloat16()
        out = torch.ones(cols, device="cuda")
        ops.colsum(x, out)
        _close(f"colsum_{rows}x{cols}", out, 1 + x.float().sum(0), 1e-3)
    z = torch.randn(1000, 96, device="cuda", generator=g).bfloat16(); dy = torch.randn_like(z)
    for act, approx in ((1, "tanh"), (2, "none")):
        zf = z.float().requires_grad_(True)
        yf = F.gelu(zf, approximate=approx); yf.backward(dy.float())
        dz = torch.empty_like(z); y = torch.empty_like(z)
        ops.act_bwd(dy, z, dz, act); ops.act_fwd(z, dy, y, act)
        _close(f"act_bwd_{act}", dz, zf.grad, 2e-2)
        _close(f"act_fwd_{act}", y, yf + dy.float(), 2e-2)
    src = torch.randn(100003 * 8, device="cuda", generator=g)
    dst = torch.empty_like(src, dtype=torch.bfloat16)
    ops.cast_f32_to_bf16(src, dst)
    assert torch.equal(dst, src.bfloat16())
    a = torch.randn(4099 * 4, device="cuda", generator=g); b = torch.randn_like(a); a0 = a.clone()
    ops.axpy(a, b, 0.5)
    alpha = torch.tensor(0.25, device="cuda")
    ops.axpy(a, b, alpha)
    _close("axpy", a, a0 + 0.75 * b, 1e-5)
    s = torch.zeros(1, device="cuda")
    ops.sumsq(src, s)
    assert abs(s.item() - src.double().pow(2).sum().item()) < 1e-3 * s.item()
    xb = torch.randn(8 * 1001, device="cuda", generator=g).bfloat16(); xb0 = xb.clone()
    ops.scale_bf16_dev(xb, alpha)
    _close("scale_dev", xb, xb0.float() * 0.25, 1e-2)


# ------------------------------------------------------------------ AdamW + clipping
@pytest.mark.parametrize("clip", [None, 1.0, 1e-3])
def test_fused_adamw_matches_torch_adamw(clip):
    """5 steps of FusedAdamW on the flat buffers == torch.optim.AdamW (+ clip_grad_norm_) on copies"""
    import mingpt_distributed_b200 as m
    torch.manual_seed(0)
    cfg = m.GPTConfig(model_type=None, n_layer=2, n_head=2, n_embed=64, vocab_size=97, block_size=32, embed_drop=0,
                      resid_drop=0, attn_drop=0, variant="reference")
    model = m.GPT(cfg, verbose=False).cuda()
    opt = m.create_optimizer(model, m.OptimizerConfig())
    model._ensure_device_state()
    names_d, names_nd = m.param_groups(model)
    ref = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters()}
    topt = torch.optim.AdamW([{"params": [ref[n] for n in names_d], "weight_decay": 0.1},
                              {"params": [ref[n] for n in names_nd], "weight_decay": 0.0}], lr=3e-4, betas=(0.9, 0.95))
    g = torch.Generator(device="cuda").manual_seed(5)
    for step in range(5):
        model._flat_grad.zero_()
        for n, p in model.named_parameters():
            gr = torch.randn(p.shape, device="cuda", generator=g) * (10.0 if step == 2 else 0.1)
            o, k = model._spans[n]
            model._flat_grad[o:o + k].view(p.shape).copy_(gr)
            p.grad = model._flat_grad[o:o + k].view(p.shape)
            ref[n].grad = gr.clone() * 0.5        # grad_scale = 0.5 (e.g. world size 2)
        if clip is not None:
            torch.nn.utils.clip_grad_norm_(list(ref.values()), clip)
            opt.step(grad_scale=0.5, gradnorm_sq=opt.grad_norm_sq(), max_norm=clip)
        else:
            opt.step(grad_scale=0.5)
        topt.step()
    torch.cuda.synchronize()
    worst = 0.0
    for n, p in model.named_parameters():
        worst = max(worst, (p.detach() - ref[n].detach()).abs().max().item())
        assert torch.allclose(p.detach(), ref[n].detach(), rtol=1e-4, atol=2e-6), n
        o, k = model._spans[n]
        assert torch.equal(model._flat_bf16[o:o + k].view(p.shape), p.detach().bfloat16()), n   # shadow refreshed
    report(f"adamw_clip{clip}", max_abs_err=worst)
    sd = opt.state_dict()
    assert float(sd["state"][0]["step"]) == 5.0 and sd["state"][0]["exp_avg"].shape == opt.param_groups[0]["params"][0].shape


def test_clip_grad_norm_standalone():
    import mingpt_distributed_b200 as m
    cfg = m.GPTConfig(model_type="gpt-nano", vocab_size=65, block_size=32, embed_drop=0, resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda()
    model._ensure_device_state()
    model._flat_grad.normal_()
    for n, p in model.named_parameters():     # padding between tensors must stay zero
        o, k = model._spans[n]
    want = model._flat_grad.clone()
    norm = m.clip_grad_norm_(model, 1.0)
    assert abs(norm.item() - want.norm().item()) < 1e-3 * norm.item()
    assert torch.allclose(model._flat_grad, want / (want.norm() + 1e-6), rtol=1e-4, atol=1e-7)
