"""tcgen05 GEMM (mgb_gemm_bf16 through the C ABI) against torch fp32 matmul on the same bf16 inputs:
the three operand layouts of a linear layer (forward, dgrad, wgrad), ragged shapes, every epilogue,
split-K accumulation. Tolerance: bf16 outputs 2e-2 relative (north_star, BF16 inputs), fp32 outputs 1e-3."""
import pytest
import torch

from conftest import report

pytestmark = pytest.mark.gpu

LAYOUTS = {"fwd_kk": (False, False), "dgrad_kmn": (False, True), "wgrad_mnmn": (True, True)}
SHAPES = [(128, 64, 64), (128, 256, 64), (256, 256, 128), (128, 128, 256), (300, 200, 136), (1024, 768, 768),
          (520, 2304, 768), (64, 48, 48), (77, 65, 48), (2048, 768, 3072)]


def _mk(M, N, K, a_mn, b_mn, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = (torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g) * 0.5).bfloat16()
    b = (torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g) * 0.5).bfloat16()
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    return a, b, A @ B.t()


def _tma_ok(M, N, K, a_mn, b_mn):
    return (M if a_mn else K) % 8 == 0 and (N if b_mn else K) % 8 == 0


def _check(name, got, want, rtol):
    got, want = got.float(), want.float()
    scale = want.abs().max().item() + 1e-12
    err = (got - want).abs()
    bad = (err > rtol * (want.abs() + 0.05 * scale)).nonzero()
    report(name, max_abs_err=err.max().item(), scale=scale, bad=int(bad.shape[0]))
    assert bad.shape[0] == 0, (f"{name}: {bad.shape[0]}/{got.numel()} elements off, max err {err.max().item():.4g} "
                               f"(scale {scale:.3g}); first bad {bad[:6].tolist()} got {got[tuple(bad[0])].item()} want {want[tuple(bad[0])].item()}")


@pytest.mark.parametrize("layout", list(LAYOUTS))
@pytest.mark.parametrize("shape", SHAPES[:5])
def test_gemm_small(layout, shape):
    from mingpt_distributed_b200 import ops
    M, N, K = shape
    a_mn, b_mn = LAYOUTS[layout]
    if not _tma_ok(M, N, K, a_mn, b_mn):
        pytest.skip("TMA needs 16-byte row pitch")
    a, b, want = _mk(M, N, K, a_mn, b_mn)
    out = torch.empty(M, N, dtype=torch.float32, device="cuda")
    ops.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    _check(f"gemm_{layout}_{M}x{N}x{K}", out, want, 1e-3)


@pytest.mark.parametrize("layout", list(LAYOUTS))
@pytest.mark.parametrize("shape", SHAPES[5:])
@pytest.mark.parametrize("block_n", [0, 64, 128, 192, 256])
def test_gemm_large_and_tile_sizes(layout, shape, block_n):
    from mingpt_distributed_b200 import ops
    M, N, K = shape
    a_mn, b_mn = LAYOUTS[layout]
    if not _tma_ok(M, N, K, a_mn, b_mn):
        pytest.skip("TMA needs 16-byte row pitch")
    if block_n and shape != SHAPES[5] and shape != SHAPES[8]:
        pytest.skip("tile sweep on two shapes only")
    a, b, want = _mk(M, N, K, a_mn, b_mn, seed=1)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, block_n=block_n)
    torch.cuda.synchronize()
    _check(f"gemm_{layout}_{M}x{N}x{K}_bn{block_n}", out, want, 2e-2)


@pytest.mark.parametrize("act", [0, 1, 2])
@pytest.mark.parametrize("N", [768, 65])
def test_gemm_forward_epilogues(act, N):
    """bias + {none, NewGELU(tanh), GELU(erf)} + residual + saved pre-activation: mingpt/model.py:179-188"""
    from mingpt_distributed_b200 import ops
    M, K = 384, 256
    a, b, acc = _mk(M, N, K, False, False, seed=2)
    bias = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    pre = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(a, b, out, bias=bias, act=act, residual=res, preact=pre)
    torch.cuda.synchronize()
    z = acc + bias
    y = z if act == 0 else torch.nn.functional.gelu(z, approximate="tanh" if act == 1 else "none")
    _check(f"gemm_epi_pre_act{act}_N{N}", pre, z, 2e-2)
    _check(f"gemm_epi_out_act{act}_N{N}", out, y + res.float(), 2e-2)


@pytest.mark.parametrize("engine", [1024, 2048])
@pytest.mark.parametrize("act", [1, 2])
@pytest.mark.parametrize("N", [768, 65])
def test_gemm_forward_stores_activation_derivative(engine, act, N):
    """MGB_ACT_STORE_GRAD: the preact output receives GELU'(z); MGB_GATE_MUL then multiplies by it -- together the
    same forward/backward as saving z and evaluating GELU' in the backward GEMM."""
    from mingpt_distributed_b200 import ops
    M, K = 384, 256
    a, b, acc = _mk(M, N, K, False, False, seed=5)
    bias = torch.randn(N, device="cuda")
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    dact = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(a, b, out, bias=bias, act=act | ops.ACT_STORE_GRAD, preact=dact, block_n=engine)
    z = (acc + bias).requires_grad_(True)
    y = torch.nn.functional.gelu(z, approximate="tanh" if act == 1 else "none")
    y.sum().backward()
    _check(f"gemm_storegrad_out_act{act}_N{N}_e{engine}", out, y.detach(), 2e-2)
    _check(f"gemm_storegrad_dact_act{act}_N{N}_e{engine}", dact, z.grad, 2e-2)
    if N % 8 == 0:      # gate operands feed TMA: 16-byte row pitch
        g = torch.Generator(device="cuda").manual_seed(6)
        dy = (torch.randn(M, 128, device="cuda", generator=g) * 0.5).bfloat16()
        w = (torch.randn(128, N, device="cuda", generator=g) * 0.1).bfloat16()
        dx = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        ops.linear_dgrad(dy, w, dx, gate=dact, gate_kind=ops.GATE_MUL, block_n=engine)
        torch.cuda.synchronize()
        _check(f"gemm_gatemul_act{act}_e{engine}", dx, (dy.float() @ w.float()) * dact.float(), 2e-2)
    with pytest.raises(RuntimeError):       # needs both an activation and somewhere to store its derivative
        ops.gemm(a, b, out, bias=bias, act=ops.ACT_STORE_GRAD, preact=dact)
    with pytest.raises(RuntimeError):
        ops.gemm(a, b, out, bias=bias, act=act | ops.ACT_STORE_GRAD)


@pytest.mark.parametrize("engine", [0, 1024, 2048])
@pytest.mark.parametrize("N", [768, 200, 65])
def test_gemm_column_sums_of_the_output(engine, N):
    """colsum[n] += sum_m D[m,n]: the bias gradient that goes with an output-gradient GEMM (fused into the CTA-pair
    kernel's epilogue when D leaves through TMA, a separate pass otherwise)"""
    from mingpt_distributed_b200 import ops
    M, K = 1000, 256          # ragged last row tile
    a, b, acc = _mk(M, N, K, False, False, seed=8)
    gate = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    cs = torch.ones(N, device="cuda")
    ops.gemm(a, b, out, gate=gate, gate_kind=ops.GATE_MUL, colsum=cs, block_n=engine)
    ops.gemm(a, b, out, gate=gate, gate_kind=ops.GATE_MUL, colsum=cs, block_n=engine)      # accumulates
    torch.cuda.synchronize()
    want = acc * gate.float()
    _check(f"gemm_colsum_out_N{N}_e{engine}", out, want, 2e-2)
    got, ref = cs - 1, 2 * want.sum(0)
    assert (got - ref).abs().max().item() <= 2e-2 * ref.abs().max().item() + 1e-3, (got - ref).abs().max().item()
    with pytest.raises(RuntimeError):
        ops.gemm(a, b, torch.empty(M, N, device="cuda"), colsum=cs)          # fp32 D


@pytest.mark.parametrize("kind", [1, 2])
def test_gemm_dgrad_gate_epilogue(kind):
    """dX = (dY W) * gelu'(pre): the fused backward of c_fc's activation"""
    from mingpt_distributed_b200 import ops
    M, N, K = 256, 512, 384     # dY [M,N], W [N,K] -> dX [M,K]
    g = torch.Generator(device="cuda").manual_seed(3)
    dy = (torch.randn(M, N, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * 0.1).bfloat16()
    pre = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    out = torch.empty(M, K, dtype=torch.bfloat16, device="cuda")
    ops.linear_dgrad(dy, w, out, gate=pre, gate_kind=kind)
    torch.cuda.synchronize()
    p = pre.float().requires_grad_(True)
    torch.nn.functional.gelu(p, approximate="tanh" if kind == 1 else "none").sum().backward()
    _check(f"gemm_gate{kind}", out, (dy.float() @ w.float()) * p.grad, 2e-2)


@pytest.mark.parametrize("split_k", [0, 1, 3, 8])
def test_gemm_wgrad_split_k_accumulates(split_k):
    """dW += dY^T X with fp32 atomics: two calls must add up (gradient accumulation), any split"""
    from mingpt_distributed_b200 import ops
    M, N, K = 4096, 768, 256    # dY [M,N], X [M,K] -> dW [N,K]
    g = torch.Generator(device="cuda").manual_seed(4)
    dy = (torch.randn(M, N, device="cuda", generator=g) * 0.1).bfloat16()
    x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    dw = torch.zeros(N, K, dtype=torch.float32, device="cuda")
    ops.linear_wgrad(dy, x, dw, split_k=split_k)
    ops.linear_wgrad(dy, x, dw, split_k=split_k)
    torch.cuda.synchronize()
    _check(f"gemm_wgrad_split{split_k}", dw, 2 * (dy.float().t() @ x.float()), 1e-3)


@pytest.mark.parametrize("engine", [1024, 2048])          # MGB_GEMM_ENGINE_SINGLE / _PAIR
def test_gemm_accumulate_with_epilogue_applies_it_once(engine):
    """D += bias + residual + A B^T at K = 4096: with `accumulate` the automatic split over K must not run the
    epilogue once per split (it is forced to 1 when an epilogue term is present); an explicit split is refused"""
    from mingpt_distributed_b200 import ops
    M, N, K = 256, 256, 4096
    a, b, acc = _mk(M, N, K, False, False, seed=9)
    bias = torch.randn(N, device="cuda") * 5
    res = (torch.randn(M, N, device="cuda") * 5).bfloat16()
    d0 = torch.randn(M, N, device="cuda")
    out = d0.clone()
    ops.gemm(a, b, out, bias=bias, residual=res, accumulate=True, block_n=engine)
    torch.cuda.synchronize()
    _check(f"gemm_accumulate_bias_residual_e{engine}", out, d0 + acc + bias + res.float(), 1e-3)
    out = d0.clone()
    ops.gemm(a, b, out, accumulate=True, block_n=engine)        # no epilogue: split-K is free to kick in
    torch.cuda.synchronize()
    _check(f"gemm_accumulate_plain_e{engine}", out, d0 + acc, 1e-3)
    with pytest.raises(RuntimeError, match="split_k"):
        ops.gemm(a, b, out, bias=bias, accumulate=True, split_k=4, block_n=engine)


@pytest.mark.parametrize("M,H,T", [(1024, 4, 256), (520, 2, 130), (4096, 12, 1024)])
def test_gemm_rowdot_side_output_is_flash_attention_delta(M, H, T):
    """dgrad of the attention out-projection with the rowdot epilogue: D = dY W (bf16) and, per row and head,
    delta = sum_d D * O from the fp32 accumulators -- what attn_bwd_delta_kernel computes from the rounded D"""
    from mingpt_distributed_b200 import ops
    C = H * 64
    g = torch.Generator(device="cuda").manual_seed(6)
    dy = torch.randn(M, C, device="cuda", generator=g).bfloat16()
    w = (torch.randn(C, C, device="cuda", generator=g) * 0.05).bfloat16()
    o = torch.randn(M, C, device="cuda", generator=g).bfloat16()
    d = torch.empty(M, C, dtype=torch.bfloat16, device="cuda")
    delta = torch.full((M // T, H, T), float("nan"), device="cuda")
    ops.linear_dgrad(dy, w, d, rowdot=(o, delta, T))
    torch.cuda.synchronize()
    want_d = dy.float() @ w.float()
    _check(f"gemm_rowdot_D_{M}x{C}", d, want_d, 2e-2)
    want = (want_d * o.float()).view(M // T, T, H, 64).sum(-1).transpose(1, 2)
    _check(f"gemm_rowdot_delta_{M}x{C}", delta, want, 2e-3)
    with pytest.raises(RuntimeError, match="rowdot"):
        ops.linear_dgrad(dy, w, d, rowdot=(o, delta, T), residual=o)


def test_gemm_strided_views_and_padded_vocab():
    """the head: logits chunk written into a wider (padded) buffer, then used as a K-major A operand whose
    reduction length (the vocabulary) is not a multiple of anything"""
    from mingpt_distributed_b200 import ops
    M, V, C, Vpad = 256, 1000 + 1, 128, 1024
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(M, C, device="cuda", generator=g).bfloat16()
    w = (torch.randn(V, C, device="cuda", generator=g) * 0.2).bfloat16()
    buf = torch.full((M, Vpad), 7.0, dtype=torch.bfloat16, device="cuda")
    ops.linear_fwd(x, w, buf[:, :V])
    torch.cuda.synchronize()
    _check("gemm_head_logits", buf[:, :V], x.float() @ w.float().t(), 2e-2)
    V8 = (V + 7) // 8 * 8
    assert (buf[:, V8:] == 7.0).all()         # D is written in 16-byte units: nothing beyond the unit holding N-1
    dx = torch.empty(M, C, dtype=torch.bfloat16, device="cuda")
    ops.linear_dgrad(buf[:, :V], w, dx)        # K = V = 1001: TMA zero-fills the ragged tail
    dw = torch.zeros(V, C, dtype=torch.float32, device="cuda")
    ops.linear_wgrad(buf[:, :V], x, dw)
    torch.cuda.synchronize()
    lg = buf[:, :V].float()
    _check("gemm_head_dgrad", dx, lg @ w.float(), 2e-2)
    _check("gemm_head_wgrad", dw, lg.t() @ x.float(), 1e-3)


def test_gemm_rejects_bad_arguments():
    from mingpt_distributed_b200 import ops
    a = torch.zeros(128, 60, dtype=torch.bfloat16, device="cuda")   # row pitch 120 B: not TMA-able
    b = torch.zeros(64, 60, dtype=torch.bfloat16, device="cuda")
    out = torch.empty(128, 64, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="16-byte|multiples of 8"):
        ops.gemm(a, b, out)
    with pytest.raises(RuntimeError, match="accumulate"):
        ops.gemm(torch.zeros(128, 64, dtype=torch.bfloat16, device="cuda"), torch.zeros(64, 64, dtype=torch.bfloat16, device="cuda"),
                 out, accumulate=True)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.gemm(a.cpu(), b.cpu(), out.cpu())


@pytest.mark.parametrize("engine", [1024, 2048])          # MGB_GEMM_ENGINE_SINGLE / _PAIR
@pytest.mark.parametrize("layout", list(LAYOUTS))
def test_gemm_both_tile_engines_agree(engine, layout):
    """the single-CTA and the CTA-pair (cta_group::2) engines on the same problems, incl. ragged M and N"""
    from mingpt_distributed_b200 import ops
    for (M, N, K) in [(520, 776, 320), (1024, 768, 768), (136, 72, 64)]:
        a_mn, b_mn = LAYOUTS[layout]
        if not _tma_ok(M, N, K, a_mn, b_mn):
            continue
        a, b, want = _mk(M, N, K, a_mn, b_mn, seed=7)
        out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        ops.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, block_n=engine)
        torch.cuda.synchronize()
        _check(f"gemm_engine{engine}_{layout}_{M}x{N}x{K}", out, want, 2e-2)


@pytest.mark.parametrize("engine", [1024, 2048])
def test_gemm_single_side_input_epilogues(engine):
    """bias + residual (block output) and gelu' gate (fused c_fc backward), one side input each: on the
    CTA-pair engine the side tile arrives by TMA and is updated in place in shared memory"""
    from mingpt_distributed_b200 import ops
    M, N, K = 768, 512, 192
    a, b, acc = _mk(M, N, K, False, False, seed=8)
    bias = torch.randn(N, device="cuda")
    side = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(a, b, out, bias=bias, residual=side, block_n=engine)
    torch.cuda.synchronize()
    _check(f"gemm_residual_engine{engine}", out, acc + bias + side.float(), 2e-2)
    for kind, approx in ((1, "tanh"), (2, "none")):
        ops.gemm(a, b, out, gate=side, gate_kind=kind, block_n=engine)
        torch.cuda.synchronize()
        s = side.float().requires_grad_(True)
        torch.nn.functional.gelu(s, approximate=approx).sum().backward()
        _check(f"gemm_gate{kind}_engine{engine}", out, acc * s.grad, 2e-2)
    # many tiles per CTA pair: exercises the box ring / barrier phases across tiles
    M2 = 256 * 74 * 2 + 256
    a2 = (torch.randn(M2, 64, device="cuda") * 0.5).bfloat16()
    b2 = (torch.randn(128, 64, device="cuda") * 0.5).bfloat16()
    side2 = torch.randn(M2, 128, device="cuda").bfloat16()
    out2 = torch.empty(M2, 128, dtype=torch.bfloat16, device="cuda")
    ops.gemm(a2, b2, out2, residual=side2, block_n=engine)
    torch.cuda.synchronize()
    _check(f"gemm_residual_manytiles_engine{engine}", out2, a2.float() @ b2.float().t() + side2.float(), 2e-2)
