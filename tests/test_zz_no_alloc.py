"""mgb_attention_bwd_noalloc: the attention backward for callers that cannot hand over scratch memory
(a plugin in a serving process whose device pool is fixed at start-up; a ctypes / cgo style stub that only
forwards pointers). Called through the C ABI and compared with the regular entry point, which gets its
scratch from the caller, on two calls in a row -- the second with different data and a smaller shape, so
stale partial sums left in the library's built-in array by the first call would show up.
Collected last on purpose (file name), so a run that stops at the first failure has run everything else."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run_both(B, T, H, d, mode, seed):
    from mingpt_distributed_b200 import ops
    C = H * d
    g = torch.Generator(device="cuda").manual_seed(seed)
    qkv = torch.randn(B, T, 3 * C, device="cuda", generator=g).bfloat16()
    dout = torch.randn(B, T, C, device="cuda", generator=g).bfloat16()
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device="cuda")
    lse = torch.empty(B, H, T, device="cuda")
    ops.attention_fwd(qkv, out, lse, B, T, H, d, mode)
    want = torch.full_like(qkv, float("nan"))
    delta = torch.empty(B, H, T, device="cuda")
    dq_acc = torch.empty(B, T, C, device="cuda")
    ops.attention_bwd(qkv, out, dout, lse, delta, dq_acc, want, B, T, H, d, mode)
    got = torch.full_like(qkv, float("nan"))
    ops.attention_bwd_noalloc(qkv, out, dout, lse, got, B, T, H, d, mode)
    torch.cuda.synchronize()
    return got.float(), want.float()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    for (B, T, H, d, mode, seed) in [(2, 256, 4, 64, 1, 11), (1, 100, 3, 32, 2, 12)]:
        got, want = _run_both(B, T, H, d, mode, seed)
        assert torch.isfinite(got).all()
        scale = want.abs().max().item()
        err = (got - want).abs()
        # same kernels, same inputs: only the order of the fp32 atomic adds into dQ differs; the bound is the
        # one the attention parity test uses for bf16 gradients
        assert (err <= 3e-2 * (want.abs() + 0.1 * scale)).all(), f"max err {err.max().item():.3g} (scale {scale:.3g})"


def test_noalloc_rejects_shapes_beyond_its_scratch():
    from mingpt_distributed_b200 import ops
    B, T, H, d = 9, 1024, 8, 64          # 4.7 M elements > 4194304
    C = H * d
    z = torch.zeros(B, T, 3 * C, dtype=torch.bfloat16, device="cuda")
    o = torch.zeros(B, T, C, dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros(B, H, T, device="cuda")
    with pytest.raises(RuntimeError, match="exceeds the built-in scratch"):
        ops.attention_bwd_noalloc(z, o, o, lse, torch.empty_like(z), B, T, H, d, 1)
