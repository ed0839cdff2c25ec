"""Dropout (the reference trains with p = 0.1 at four sites, mingpt/model.py:49-51,140,150,183,215): the
counter-based masks have the right statistics, forward and backward use the SAME mask, and a whole training
forward/backward with dropout equals the oracle run with the exported masks."""
import math

import pytest
import torch

import mingpt_distributed_b200 as m
from conftest import report
from oracle import gpt_oracle as go

pytestmark = pytest.mark.gpu


def _seed(v):
    return torch.tensor([v], dtype=torch.int64, device="cuda")


def test_elementwise_dropout_statistics_and_determinism():
    from mingpt_distributed_b200 import ops
    n, p = 1 << 21, 0.1
    x = (torch.rand(n, device="cuda") + 0.5).bfloat16()
    res = torch.randn(n, device="cuda").bfloat16()
    y = torch.empty_like(x)
    ops.dropout(x, None, y, p, _seed(1234), 7)
    keep = y != 0
    rate = keep.float().mean().item()
    assert abs(rate - (1 - p)) < 4 * math.sqrt(p * (1 - p) / n) + 1e-4, rate
    assert torch.allclose(y[keep].float(), (x[keep].float() / (1 - p)).bfloat16().float(), rtol=1e-2)
    y2 = torch.empty_like(x)
    ops.dropout(x, None, y2, p, _seed(1234), 7)
    assert torch.equal(y, y2)                                   # same (seed, site) -> same mask
    ops.dropout(x, None, y2, p, _seed(1235), 7)
    assert (((y2 != 0) ^ keep).float().mean().item()) > 0.1     # new seed -> new mask
    ops.dropout(x, None, y2, p, _seed(1234), 8)
    assert (((y2 != 0) ^ keep).float().mean().item()) > 0.1     # other site -> other mask
    # fused residual add, and the backward (= the same call on the gradient) sees the same mask
    ops.dropout(x, res, y2, p, _seed(1234), 7)
    assert torch.allclose(y2.float(), y.float() + res.float(), rtol=2e-2, atol=2e-2)
    g = torch.ones_like(x)
    dg = torch.empty_like(x)
    ops.dropout(g, None, dg, p, _seed(1234), 7)
    assert torch.equal(dg != 0, keep)
    # ragged length
    ops.dropout(x[:1003], None, y2[:1003], p, _seed(1234), 7)
    assert torch.equal(y2[:1003], y[:1003])
    report("dropout_keep_rate", rate=rate)


@pytest.mark.parametrize("d,B,T,H", [(16, 2, 136, 3), (32, 2, 136, 3), (64, 2, 136, 3), (64, 3, 384, 2), (64, 40, 256, 4)])
@pytest.mark.parametrize("mode", [1, 2])
def test_attention_dropout_matches_torch_with_the_same_mask(d, mode, B, T, H):
    """head size 64 runs the tcgen05 kernels (mask bits regenerated per 8 keys in the forward, shared across a
    warp by shuffles in the transposed backward); (40, 256, 4) gives every persistent backward CTA several items"""
    from mingpt_distributed_b200 import ops
    p = 0.2
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(5)
    qkv = torch.randn(B, T, 3 * C, device="cuda", generator=g).bfloat16()
    dout = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    seed = _seed(99)
    keep = torch.empty(B, H, T, T, dtype=torch.uint8, device="cuda")
    ops.attention_dropout_mask(keep, B, T, H, p, seed, 12)
    rate = keep.float().mean().item()
    assert abs(rate - (1 - p)) < 0.01
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B, H, T, device="cuda")
    ops.attention_fwd_dropout(qkv, out, lse, B, T, H, d, mode, p, seed, 12)
    qf = qkv.float().requires_grad_(True)
    q, k, v = (t.view(B, T, H, d).transpose(1, 2) for t in qf.chunk(3, dim=-1))
    want = go._attention_core(q, k, v, "causal" if mode == 1 else "add_tril_ones", keep, p).transpose(1, 2).reshape(B, T, C)
    want.backward(dout.float())
    err = (out.float() - want).abs().max().item()
    assert err < 3e-2 * max(1.0, want.abs().max().item()), err
    dqkv = torch.empty_like(qkv); delta = torch.empty_like(lse); dq = torch.empty(B, T, C, device="cuda")
    ops.attention_bwd_dropout(qkv, out, dout, lse, delta, dq, dqkv, B, T, H, d, mode, p, seed, 12)
    torch.cuda.synchronize()
    gerr = (dqkv.float() - qf.grad).abs()
    scale = qf.grad.abs().max().item()
    # per element: 3 % of its own size plus 0.45 % of the largest gradient (bf16 P / dS operands, p = 0.2 scaling)
    assert (gerr <= 3e-2 * (qf.grad.abs() + 0.15 * scale)).all(), gerr.max().item()
    report(f"attn_dropout_d{d}_m{mode}_{B}x{T}x{H}", out_err=err, grad_err=gerr.max().item(), keep_rate=rate)


def _unpack_keep_words(bits, B, T, H):
    """keep[b,h,q,k] twice, from the two layouts of mgb_attention_dropout_bits (include/mingpt_b200.h, csrc/dropout.cuh):
    query-major words [(bh, k // 32, q)] with key k at bit 16 * (k & 1) + 15 - (k % 32) // 2, then key-major words
    [(bh, q // 32, k)] with query q at bit q % 32"""
    kw, tp = (T + 31) // 32, (T + 15) // 16 * 16
    nq = B * H * kw * tp
    w = bits[:nq].view(B, H, kw, tp)[..., :T].to(torch.int64) & 0xffffffff                    # [B,H,kw,q]
    k = torch.arange(kw * 32, device=bits.device)
    pos = 16 * (k & 1) + 15 - (k % 32) // 2
    keep_q = ((w[:, :, k // 32, :] >> pos.view(1, 1, -1, 1)) & 1).permute(0, 1, 3, 2)[..., :T]  # [B,H,q,k]
    wt = bits[nq:nq + B * H * kw * kw * 32].view(B, H, kw, kw * 32).to(torch.int64) & 0xffffffff   # [B,H,qw,k]
    q = torch.arange(kw * 32, device=bits.device)
    keep_t = ((wt[:, :, q // 32, :] >> (q % 32).view(1, 1, -1, 1)) & 1)[:, :, :T, :T]           # [B,H,q,k]
    return keep_q.to(torch.uint8), keep_t.to(torch.uint8)


@pytest.mark.parametrize("B,T,H", [(2, 136, 3), (3, 384, 2), (40, 256, 4), (2, 1024, 2)])
@pytest.mark.parametrize("mode", [1, 2])
def test_attention_dropout_from_pregenerated_keep_words(mode, B, T, H):
    """the keep words generated ahead of the kernels hold exactly the mask the seed-taking entry points draw, and the
    head-size-64 kernels reading them give the same forward (bit for bit) and backward (fp32 atomics: to rounding)"""
    from mingpt_distributed_b200 import ops
    d, p, site = 64, 0.2, 12
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(6)
    qkv = torch.randn(B, T, 3 * C, device="cuda", generator=g).bfloat16()
    dout = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    seed = _seed(77)
    keep = torch.empty(B, H, T, T, dtype=torch.uint8, device="cuda")
    ops.attention_dropout_mask(keep, B, T, H, p, seed, site)
    words = ops.attention_dropout_bits_words(B, T, H)
    kw = (T + 31) // 32
    assert words == B * H * kw * ((T + 15) // 16 * 16) + B * H * kw * kw * 32
    bits = torch.zeros(words, dtype=torch.int32, device="cuda")
    ops.attention_dropout_bits(bits, B, T, H, mode, p, seed, site)
    for got in _unpack_keep_words(bits, B, T, H):
        if mode == 1:    # causal: words right of the diagonal are not generated; compare what the kernels read
            qi = torch.arange(T, device="cuda").view(T, 1); ki = torch.arange(T, device="cuda").view(1, T)
            vis = ki <= qi
            assert torch.equal(got[..., vis], keep[..., vis])
        else:
            assert torch.equal(got, keep)
    out_a = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda"); lse_a = torch.empty(B, H, T, device="cuda")
    out_b = torch.empty_like(out_a); lse_b = torch.empty_like(lse_a)
    ops.attention_fwd_dropout(qkv, out_a, lse_a, B, T, H, d, mode, p, seed, site)
    ops.attention_fwd_dropout_bits(qkv, out_b, lse_b, B, T, H, d, mode, p, bits)
    assert torch.equal(out_a, out_b) and torch.equal(lse_a, lse_b)
    dq = torch.empty(B, T, C, device="cuda"); delta = torch.empty_like(lse_a)
    g_a = torch.empty_like(qkv); g_b = torch.empty_like(qkv)
    ops.attention_bwd_dropout(qkv, out_a, dout, lse_a, delta, dq, g_a, B, T, H, d, mode, p, seed, site)
    cs = torch.zeros(3 * C, device="cuda")
    ops.attention_bwd_dropout_bits(qkv, out_a, dout, lse_a, delta, dq, g_b, B, T, H, d, mode, p, bits, dqkv_colsum=cs)
    torch.cuda.synchronize()
    err = (g_a.float() - g_b.float()).abs().max().item()
    scale = g_a.float().abs().max().item()
    report(f"attn_dropout_bits_m{mode}_{B}x{T}x{H}", grad_dev=err, grad_max=scale)
    assert err <= 1e-2 * scale
    want_cs = g_b.float().sum(dim=(0, 1))
    assert torch.allclose(cs, want_cs, rtol=2e-2, atol=2e-2 * want_cs.abs().max().item())


@pytest.mark.parametrize("dims", [(2, 2, 64, 97, 64, 3), (2, 2, 128, 97, 96, 3)], ids=["d32", "d64"])
@pytest.mark.parametrize("variant", ["mingpt", "reference"])
def test_training_step_with_dropout_equals_oracle_with_exported_masks(variant, dims):
    """d64 (B*T = 288 rows): the tcgen05 attention kernels read keep words generated on a side stream, the residual
    dropouts sit in the c_proj GEMM epilogues and their backward carries the bias gradients"""
    from mingpt_distributed_b200 import ops
    L, H, C, V, T, B = dims
    pe, pr, pa = 0.1, 0.15, 0.2
    torch.manual_seed(11)
    cfg = m.GPTConfig(model_type=None, n_layer=L, n_head=H, n_embed=C, vocab_size=V, block_size=T, embed_drop=pe,
                      resid_drop=pr, attn_drop=pa, variant=variant)
    model = m.GPT(cfg, verbose=False)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model = model.cuda().train()
    x = torch.randint(0, V, (B, T)); y = torch.randint(0, V, (B, T))
    _, loss = model(x.cuda(), y.cuda())
    loss.backward()
    seed = model._rng_seed.clone()
    # rebuild every mask of that step from (seed, site)
    ones = torch.ones(B * T * C, device="cuda").bfloat16(); buf = torch.empty_like(ones)
    def ew_mask(p, site):
        ops.dropout(ones, None, buf, p, seed, site)
        return (buf != 0).view(B, T, C).cpu()
    masks = {"embed": ew_mask(pe, m.GPT.SITE_EMBED)}
    for l in range(L):
        keep = torch.empty(B, H, T, T, dtype=torch.uint8, device="cuda")
        ops.attention_dropout_mask(keep, B, T, H, pa, seed, 4 * l)
        masks[("attn", l)] = keep.cpu()
        masks[("resid1", l)] = ew_mask(pr, 4 * l + 1)
        masks[("resid2", l)] = ew_mask(pr, 4 * l + 2)
    q = lambda p: 1.0 - round(p * 65536) / 65536       # the kernels quantise the keep probability to 16 bits
    osd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    _, oloss = go.forward(osd, H, x, y, variant, masks, (1 - q(pe), 1 - q(pr), 1 - q(pa)))
    oloss.backward()
    report(f"dropout_step_{variant}_C{C}", loss=loss.item(), oracle=oloss.item())
    assert abs(loss.item() - oloss.item()) < 1e-2 * oloss.item()
    for n, p in model.named_parameters():
        ref = osd[n].grad
        if ref.norm() < 1e-6:
            continue
        r = ((p.grad.cpu() - ref).norm() / ref.norm()).item()
        assert r < 2e-2, f"{n}: {r:.4f}"
    # eval mode switches dropout off; a second training forward draws a new seed
    model.eval()
    with torch.no_grad():
        _, le = model(x.cuda(), y.cuda())
    _, oe = go.forward(sd, H, x, y, variant)
    assert abs(le.item() - oe.item()) < 1e-2 * oe.item()
    model.train()
    _, l2 = model(x.cuda(), y.cuda())
    assert int(model._rng_seed.item()) == int(seed.item()) + 1 and abs(l2.item() - loss.item()) > 0


@pytest.mark.parametrize("M,N,K", [(256, 768, 768), (1000, 768, 3072), (512, 64, 192)])
def test_gemm_dropout_epilogue_equals_gemm_then_dropout(M, N, K):
    """D = residual + dropout(A W^T + bias) in the CTA-pair GEMM's epilogue == the GEMM followed by the elementwise
    dropout kernel with the same (seed, site): the same mask exactly, values to bf16 rounding of the intermediate"""
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(8)
    p, site, seed = 0.15, 9, _seed(4242)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    fused = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.linear_fwd(a, w, fused, bias=bias, residual=res, dropout=(p, seed, site))
    tmp = torch.empty_like(fused); two = torch.empty_like(fused)
    ops.linear_fwd(a, w, tmp, bias=bias)
    ops.dropout(tmp, res, two, p, seed, site)
    ones = torch.ones(M * N, device="cuda").bfloat16(); mk = torch.empty_like(ones)
    ops.dropout(ones, None, mk, p, seed, site)
    keep = (mk != 0).view(M, N)
    want = res.float() + torch.where(keep, (a.float() @ w.float().t() + bias) / (1 - round(p * 65536) / 65536), torch.zeros((), device="cuda"))
    dropped = ~keep
    assert torch.equal(fused[dropped], res[dropped])                      # dropped entries: the residual alone, exactly
    err = (fused.float() - want).abs().max().item()
    err2 = (two.float() - want).abs().max().item()
    report(f"gemm_dropout_epilogue_{M}x{N}x{K}", fused_err=err, two_pass_err=err2, keep_rate=keep.float().mean().item())
    assert err <= 2e-2 * want.abs().max().item() and err <= 1.5 * err2 + 1e-3


def test_dropout_backward_with_fused_bias_gradient():
    from mingpt_distributed_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(9)
    for rows, cols in [(4096, 768), (1000, 192), (77, 1600), (300, 136)]:
        dy = torch.randn(rows, cols, device="cuda", generator=g).bfloat16()
        a = torch.empty_like(dy); b = torch.empty_like(dy)
        cs = torch.zeros(cols, device="cuda")
        ops.dropout(dy, None, a, 0.1, _seed(5), 3)
        ops.dropout_bwd_colsum(dy, b, 0.1, _seed(5), 3, cs)
        assert torch.equal(a, b)
        want = a.float().sum(0)
        assert torch.allclose(cs, want, rtol=1e-3, atol=1e-3 * want.abs().max().item())


def test_kernel_masks_equal_the_numpy_restatement():
    """every way the library draws a mask -- the elementwise kernel, the attention mask export, the pre-generated keep
    words -- against oracle/dropout_mask.py, bit for bit"""
    import numpy as np
    from mingpt_distributed_b200 import ops
    from oracle import dropout_mask as dm
    seed_v, site, p = 0x1234567890AB, 9, 0.3
    seed = _seed(seed_v)
    n = 100_003
    ones = torch.ones(n, device="cuda").bfloat16(); y = torch.empty_like(ones)
    ops.dropout(ones, None, y, p, seed, site)
    assert np.array_equal((y != 0).cpu().numpy(), dm.elementwise_keep(seed_v, site, n, p))
    for B, H, T in [(2, 2, 40), (1, 3, 136)]:
        keep = torch.empty(B, H, T, T, dtype=torch.uint8, device="cuda")
        ops.attention_dropout_mask(keep, B, T, H, p, seed, site)
        want = dm.attention_keep(seed_v, site, B, H, T, p)
        assert np.array_equal(keep.cpu().numpy().astype(bool), want)
        bits = torch.zeros(ops.attention_dropout_bits_words(B, T, H), dtype=torch.int32, device="cuda")
        ops.attention_dropout_bits(bits, B, T, H, 2, p, seed, site)
        for got in _unpack_keep_words(bits, B, T, H):
            assert np.array_equal(got.cpu().numpy().astype(bool), want)
