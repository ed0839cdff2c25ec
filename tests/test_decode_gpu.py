"""Incremental decode: the decode attention kernel against torch, and one cached decode step of the
whole stack against the same position of a full-window forward."""
import math

import pytest
import torch

import mingpt_distributed_b200 as m
from conftest import report

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("d,H,B,cap,pos", [(64, 12, 3, 1024, 1023), (64, 2, 2, 256, 0), (32, 6, 4, 256, 100), (16, 3, 5, 128, 77)])
def test_attention_decode_vs_torch(d, H, B, cap, pos):
    from mingpt_distributed_b200 import ops
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(90 + d)
    cache = torch.randn(B, cap, 3 * C, device="cuda", generator=g).bfloat16()
    out = torch.empty(B, C, dtype=torch.bfloat16, device="cuda")
    ops.attention_decode(cache, out, B, cap, pos, H, d)
    q = cache[:, pos, :C].float().view(B, H, 1, d)
    k = cache[:, :pos + 1, C:2 * C].float().view(B, pos + 1, H, d).transpose(1, 2)
    v = cache[:, :pos + 1, 2 * C:].float().view(B, pos + 1, H, d).transpose(1, 2)
    att = torch.softmax((q @ k.transpose(-2, -1)) / math.sqrt(d), -1)
    want = (att @ v).transpose(1, 2).reshape(B, C)
    err = (out.float() - want).abs().max().item()
    report(f"attn_decode_d{d}_pos{pos}", max_abs_err=err, scale=want.abs().max().item())
    assert err < 1e-2 * max(1.0, want.abs().max().item())


@pytest.mark.parametrize("preset,V,blk", [("gpt-nano", 65, 64), ("gpt-mini", 65, 128), (None, 300, 192)])
def test_cached_decode_step_equals_full_forward(preset, V, blk):
    """logits of position t from prefill(0..t-1) + one decode step == logits of position t from a full forward"""
    torch.manual_seed(0)
    if preset:
        cfg = m.GPTConfig(model_type=preset, vocab_size=V, block_size=blk, embed_drop=0, resid_drop=0, attn_drop=0)
    else:
        cfg = m.GPTConfig(model_type=None, n_layer=2, n_head=2, n_embed=128, vocab_size=V, block_size=blk, embed_drop=0, resid_drop=0, attn_drop=0)
    model = m.GPT(cfg, verbose=False).cuda().eval()
    with torch.no_grad():
        for p in model.parameters():          # init has zero biases / positions: make them matter
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    B, T0 = 3, blk // 2
    tok = torch.randint(0, V, (B, blk), device="cuda")
    full, _ = model(tok)                                          # fp32 logits of every position
    from mingpt_distributed_b200.model import _DecodeState
    model._ensure_device_state()
    ws = model._ws(B, T0)
    ws.generation += 1
    model._trunk(ws, tok[:, :T0].contiguous())
    st = _DecodeState(model, B)
    for l in range(model.n_layer):
        st.cache[l][:, :T0, :].copy_(ws.qkv[l].view(B, T0, -1))
    worst = 0.0
    for pos in range(T0, blk):
        lg = model._decode_step(st, tok[:, pos].contiguous(), pos)
        err = (lg - full[:, pos]).abs().max().item() / max(1.0, full[:, pos].abs().max().item())
        worst = max(worst, err)
    report(f"decode_step_{preset}", worst_rel_err=worst)
    assert worst < 2e-2
