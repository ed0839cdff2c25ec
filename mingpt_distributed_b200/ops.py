"""Tensor-level wrappers over the C ABI (no allocation policy of their own: the caller passes outputs)
plus the torch custom ops (`torch.ops.mingpt_b200.*`) with autograd that mirror the torch calls the
reference makes: F.layer_norm, F.linear (+GELU/residual epilogues), the attention core of
nn.MultiheadAttention, nn.Embedding + positional add, F.cross_entropy(ignore_index=-1).

Every function here launches CUDA kernels from libmingpt_b200.so on the current torch stream. Nothing
falls back to PyTorch math: a missing library or a non-CUDA tensor is an error.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch

from . import _lib
from ._lib import AdamWChunk, GemmArgs, check

ACT_NONE, ACT_GELU_TANH, ACT_GELU_ERF = 0, 1, 2
GATE_MUL = 3            # gate_kind / act_bwd kind: the tensor holds the derivative itself
ACT_STORE_GRAD = 256    # OR-ed into act: the preact output receives act'(v) instead of v
MASK_NONE, MASK_CAUSAL, MASK_ADD_TRIL_ONES = 0, 1, 2
ACT_BY_NAME = {None: 0, "none": 0, "gelu_tanh": 1, "gelu_erf": 2}
MASK_BY_NAME = {None: 0, "none": 0, "causal": 1, "add_tril_ones": 2}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _chk_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("mingpt_b200 ops run on CUDA tensors only (there is no CPU path)")


def _chk2d(t: torch.Tensor, dtype, name: str):
    if t.dim() != 2 or t.stride(1) != 1:
        raise RuntimeError(f"{name}: expected a 2-D tensor with unit inner stride, got shape {tuple(t.shape)} strides {t.stride()}")
    if t.dtype != dtype:
        raise RuntimeError(f"{name}: expected {dtype}, got {t.dtype}")


# ----------------------------------------------------------------------------------------------------
# GEMM
# ----------------------------------------------------------------------------------------------------
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         bias: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None, act: int = 0,
         preact: Optional[torch.Tensor] = None, gate: Optional[torch.Tensor] = None, gate_kind: int = 0,
         accumulate: bool = False, split_k: int = 0, block_n: int = 0, alpha: float = 1.0,
         colsum: Optional[torch.Tensor] = None, rowdot=None, dropout=None) -> torch.Tensor:
    """out[M,N] = epilogue(alpha * A B^T).  a is [M,K] (or [K,M] if a_mn), b is [N,K] (or [K,N] if b_mn).
    dropout=(p, seed_dev, site): out = residual + dropout_p(alpha * A B^T + bias), the mask of dropout(..., site)."""
    _chk_cuda(a, b, out, bias, residual, preact, gate, colsum)
    _chk2d(a, torch.bfloat16, "gemm a")
    _chk2d(b, torch.bfloat16, "gemm b")
    if out.dim() != 2 or out.stride(1) != 1 or out.dtype not in (torch.bfloat16, torch.float32):
        raise RuntimeError("gemm out: expected 2-D bf16/fp32 with unit inner stride")
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    if K != Kb or tuple(out.shape) != (M, N):
        raise RuntimeError(f"gemm: shape mismatch M={M} N={N} K={K}/{Kb} out={tuple(out.shape)}")
    args = GemmArgs()
    args.a, args.lda, args.a_mn_major = a.data_ptr(), a.stride(0), int(a_mn)
    args.b, args.ldb, args.b_mn_major = b.data_ptr(), b.stride(0), int(b_mn)
    args.d, args.ldd, args.d_f32, args.accumulate = out.data_ptr(), out.stride(0), int(out.dtype == torch.float32), int(accumulate)
    args.m, args.n, args.k = M, N, K
    if bias is not None:
        if bias.dtype != torch.float32 or bias.numel() != N or not bias.is_contiguous():
            raise RuntimeError("gemm bias: expected contiguous fp32 [N]")
        args.bias = bias.data_ptr()
    if residual is not None:
        _chk2d(residual, torch.bfloat16, "gemm residual")
        args.residual, args.ldr = residual.data_ptr(), residual.stride(0)
    args.act = act
    if preact is not None:
        _chk2d(preact, torch.bfloat16, "gemm preact")
        args.preact, args.ldp = preact.data_ptr(), preact.stride(0)
    if gate is not None:
        _chk2d(gate, torch.bfloat16, "gemm gate")
        args.gate, args.ldg, args.gate_kind = gate.data_ptr(), gate.stride(0), gate_kind
    args.split_k, args.block_n, args.alpha = split_k, block_n, alpha
    if colsum is not None:
        if colsum.dtype != torch.float32 or colsum.numel() != N or not colsum.is_contiguous():
            raise RuntimeError("gemm colsum: expected contiguous fp32 [N]")
        args.colsum = colsum.data_ptr()
    if rowdot is not None:     # (x bf16 [M,N], out fp32 [B, N/64, T], T): out = per-head row sums of D * x (see mgb_gemm_args)
        rx, rout, rt = rowdot
        _chk_cuda(rx, rout)
        _chk2d(rx, torch.bfloat16, "gemm rowdot x")
        if tuple(rx.shape) != (M, N) or rout.dtype != torch.float32 or rout.numel() != M * (N // 64) or not rout.is_contiguous():
            raise RuntimeError("gemm rowdot: x must be [M,N] bf16 and out contiguous fp32 with M * N/64 elements")
        args.rowdot_x, args.ld_rowdot, args.rowdot_out, args.rowdot_t = rx.data_ptr(), rx.stride(0), rout.data_ptr(), int(rt)
    if dropout is not None and dropout[0] > 0:
        dp, dseed, dsite = dropout
        _chk_cuda(dseed)
        args.drop_seed, args.drop_p, args.drop_site = dseed.data_ptr(), float(dp), int(dsite)
    check(_lib.lib().mgb_gemm_bf16(C.byref(args), _stream()), "gemm")
    return out


def linear_fwd(x, w, out, bias=None, act=0, residual=None, preact=None, **kw):
    """Y = act(X W^T + b) (+ residual): X [M,K], W [N,K]  (nn.Linear forward)"""
    return gemm(x, w, out, bias=bias, act=act, residual=residual, preact=preact, **kw)


def linear_dgrad(dy, w, out, gate=None, gate_kind=0, residual=None, **kw):   # kw: colsum=fp32 [K] += column sums of out
    """dX = dY W: dY [M,N], W [N,K] read as the MN-major B operand"""
    return gemm(dy, w, out, b_mn=True, gate=gate, gate_kind=gate_kind, residual=residual, **kw)


def linear_wgrad(dy, x, dw, accumulate=True, **kw):
    """dW (+)= dY^T X: dY [M,N], X [M,K], dW fp32 [N,K]; both operands MN-major, split-K over M"""
    return gemm(dy, x, dw, a_mn=True, b_mn=True, accumulate=accumulate, **kw)


# ----------------------------------------------------------------------------------------------------
# LayerNorm / embedding / attention / cross entropy / helpers
# ----------------------------------------------------------------------------------------------------
def layernorm_fwd(x, gamma, beta, y, mean, rstd, eps=1e-5):
    _chk_cuda(x, gamma, beta, y, mean, rstd)
    rows, cols = x.shape
    check(_lib.lib().mgb_layernorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), rows, cols, eps, _stream()), "layernorm_fwd")
    return y


def layernorm_bwd(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dres=None, dres_colsum=None):
    """dres_colsum (fp32 [cols], accumulated): column sums of dres, i.e. the bias gradient of the linear layer whose
    output gradient dres is, taken from the read this kernel makes anyway."""
    _chk_cuda(dy, x, gamma, mean, rstd, dx, dgamma, dbeta, dres, dres_colsum)
    rows, cols = x.shape
    if dres_colsum is not None:
        check(_lib.lib().mgb_layernorm_bwd_colsum(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx), _p(dgamma),
                                                  _p(dbeta), _p(dres_colsum), rows, cols, _stream()), "layernorm_bwd_colsum")
    else:
        check(_lib.lib().mgb_layernorm_bwd(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx), _p(dgamma), _p(dbeta), rows, cols, _stream()), "layernorm_bwd")
    return dx


def embedding_fwd(idx, wte, wpe, out, err_flag=None):
    _chk_cuda(idx, wte, wpe, out)
    B, T = idx.shape
    vocab, Cdim = wte.shape
    check(_lib.lib().mgb_embedding_fwd(_p(idx), _p(wte), _p(wpe), _p(out), B, T, Cdim, vocab, _p(err_flag), _stream()), "embedding_fwd")
    return out


def embedding_bwd(idx, dx, dwte, dwpe):
    _chk_cuda(idx, dx, dwte, dwpe)
    B, T = idx.shape
    vocab, Cdim = dwte.shape
    check(_lib.lib().mgb_embedding_bwd(_p(idx), _p(dx), _p(dwte), _p(dwpe), B, T, Cdim, vocab, _stream()), "embedding_bwd")


def token_windows(codes, starts, T, x, y, err_flag=None):
    """x[b] = codes[starts[b] : +T], y[b] = codes[starts[b]+1 : +T] (int64) from a device-resident uint8 corpus"""
    _chk_cuda(codes, starts, x, y, err_flag)
    if codes.dtype != torch.uint8 or starts.dtype != torch.int64 or x.dtype != torch.int64 or y.dtype != torch.int64:
        raise RuntimeError("token_windows: codes uint8, starts / x / y int64")
    B = starts.numel()
    if tuple(x.shape) != (B, T) or tuple(y.shape) != (B, T) or not x.is_contiguous() or not y.is_contiguous():
        raise RuntimeError("token_windows: x and y must be contiguous [len(starts), T]")
    check(_lib.lib().mgb_token_windows(_p(codes), codes.numel(), _p(starts), B, T, _p(x), _p(y), _p(err_flag), _stream()), "token_windows")
    return x, y


def attention_fwd(qkv, out, lse, B, T, H, d, mask_mode, scale=None):
    _chk_cuda(qkv, out, lse)
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    check(_lib.lib().mgb_attention_fwd(_p(qkv), _p(out), _p(lse), B, T, H, d, mask_mode, scale, _stream()), "attention_fwd")
    return out


ATTN_DELTA_GIVEN = 0x100


def attention_bwd(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, d, mask_mode, scale=None, dqkv_colsum=None,
                  delta_given=False):
    """dqkv_colsum (fp32 [3*H*d], accumulated): column sums of dqkv, i.e. the qkv projection's bias gradient.
    delta_given: `delta` already holds rowsum(dout * out) (the dgrad GEMM that produced dout wrote it: gemm rowdot)."""
    _chk_cuda(qkv, out, dout, lse, delta, dq_accum, dqkv, dqkv_colsum)
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    if delta_given:
        mask_mode |= ATTN_DELTA_GIVEN
    if dqkv_colsum is not None:
        check(_lib.lib().mgb_attention_bwd_colsum(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dq_accum), _p(dqkv),
                                                  _p(dqkv_colsum), B, T, H, d, mask_mode, scale, _stream()), "attention_bwd_colsum")
    else:
        check(_lib.lib().mgb_attention_bwd(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dq_accum), _p(dqkv), B, T, H, d, mask_mode, scale, _stream()), "attention_bwd")
    return dqkv


def attention_decode(qkv_cache, out, B, Tcap, pos, H, d, scale=None):
    """one new query per sequence (row `pos` of the cache) against cached keys / values 0..pos; out bf16 [B, H*d]"""
    _chk_cuda(qkv_cache, out)
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    check(_lib.lib().mgb_attention_decode(_p(qkv_cache), _p(out), B, Tcap, pos, H, d, scale, _stream()), "attention_decode")
    return out


def attention_bwd_noalloc(qkv, out, dout, lse, dqkv, B, T, H, d, mask_mode, scale=None):
    """attention_bwd for callers with no scratch to offer (a plugin in a process with a fixed device pool, a
    pointer-forwarding FFI stub): delta / dQ accumulator come from a fixed array inside the library, so B*T*H*d
    is capped at 4194304 and overlapping calls on one GPU must be serialised (see mingpt_b200.h)."""
    _chk_cuda(qkv, out, dout, lse, dqkv)
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    check(_lib.lib().mgb_attention_bwd_noalloc(_p(qkv), _p(out), _p(dout), _p(lse), _p(dqkv), B, T, H, d, mask_mode, scale, _stream()), "attention_bwd_noalloc")
    return dqkv


def dropout(x, res, y, p, seed_dev, site):
    """y = (res or 0) + dropout_p(x); mask = f(*seed_dev, site, element). Also its own backward (res=None)."""
    _chk_cuda(x, res, y, seed_dev)
    check(_lib.lib().mgb_dropout_bf16(_p(x), _p(res), _p(y), x.numel(), float(p), _p(seed_dev), int(site), _stream()), "dropout")
    return y


def dropout_bwd_colsum(dy, y, p, seed_dev, site, colsum):
    """y = dropout_p mask applied to the 2-D gradient dy (the mask of dropout(..., site)), colsum += y.sum(0)"""
    _chk_cuda(dy, y, seed_dev, colsum)
    _chk2d(dy, torch.bfloat16, "dropout_bwd_colsum dy")
    if dy.stride(0) != dy.shape[1] or y.stride(0) != dy.shape[1]:
        raise ValueError("dropout_bwd_colsum: contiguous rows only")
    check(_lib.lib().mgb_dropout_bf16_colsum(_p(dy), _p(y), dy.shape[0], dy.shape[1], float(p), _p(seed_dev), int(site), _p(colsum), _stream()), "dropout_colsum")
    return y


def attention_fwd_dropout(qkv, out, lse, B, T, H, d, mask_mode, p, seed_dev, site, scale=None):
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    check(_lib.lib().mgb_attention_fwd_dropout(_p(qkv), _p(out), _p(lse), B, T, H, d, mask_mode, scale, float(p), _p(seed_dev), int(site), _stream()), "attention_fwd_dropout")
    return out


def attention_bwd_dropout(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, d, mask_mode, p, seed_dev, site, scale=None,
                          delta_given=False):
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    if delta_given:
        mask_mode |= ATTN_DELTA_GIVEN
    check(_lib.lib().mgb_attention_bwd_dropout(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dq_accum), _p(dqkv), B, T, H, d, mask_mode, scale, float(p), _p(seed_dev), int(site), _stream()), "attention_bwd_dropout")
    return dqkv


def attention_dropout_bits_words(B, T, H) -> int:
    """uint32 words of the keep-bit buffer of one attention dropout site (mgb_attention_dropout_bits_words)"""
    import ctypes
    n = ctypes.c_int64(0)
    check(_lib.lib().mgb_attention_dropout_bits_words(B, T, H, ctypes.addressof(n)), "attention_dropout_bits_words")
    return int(n.value)


def attention_dropout_bits(bits, B, T, H, mask_mode, p, seed_dev, site):
    """fills `bits` (int32 [attention_dropout_bits_words]) with the keep words of one site: the mask of
    attention_fwd_dropout / attention_dropout_mask, laid out for the head-size-64 kernels"""
    if bits.numel() < attention_dropout_bits_words(B, T, H) or bits.element_size() != 4:
        raise ValueError("attention_dropout_bits: buffer too small")
    check(_lib.lib().mgb_attention_dropout_bits(_p(bits), B, T, H, mask_mode, float(p), _p(seed_dev), int(site), _stream()), "attention_dropout_bits")
    return bits


def attention_fwd_dropout_bits(qkv, out, lse, B, T, H, d, mask_mode, p, bits, scale=None):
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    check(_lib.lib().mgb_attention_fwd_dropout_bits(_p(qkv), _p(out), _p(lse), B, T, H, d, mask_mode, scale, float(p), _p(bits), _stream()), "attention_fwd_dropout_bits")
    return out


def attention_bwd_dropout_bits(qkv, out, dout, lse, delta, dq_accum, dqkv, B, T, H, d, mask_mode, p, bits, scale=None,
                               delta_given=False, dqkv_colsum=None):
    scale = 1.0 / math.sqrt(d) if scale is None else scale
    if delta_given:
        mask_mode |= ATTN_DELTA_GIVEN
    check(_lib.lib().mgb_attention_bwd_dropout_bits(_p(qkv), _p(out), _p(dout), _p(lse), _p(delta), _p(dq_accum), _p(dqkv), B, T, H, d, mask_mode, scale, float(p), _p(bits), _p(dqkv_colsum), _stream()), "attention_bwd_dropout_bits")
    return dqkv


def attention_dropout_mask(keep, B, T, H, p, seed_dev, site):
    check(_lib.lib().mgb_attention_dropout_mask(_p(keep), B, T, H, float(p), _p(seed_dev), int(site), _stream()), "attention_dropout_mask")
    return keep


def cross_entropy_fwd_bwd(logits, targets, row_loss, n_valid, vocab, grad_scale=1.0, write_grad=True, err_flag=None):
    """in place on bf16 logits [rows, ld]; see mgb_cross_entropy_fwd_bwd"""
    _chk_cuda(logits, targets, row_loss, n_valid)
    rows = logits.shape[0]
    check(_lib.lib().mgb_cross_entropy_fwd_bwd(_p(logits), logits.stride(0), _p(targets), _p(row_loss), _p(n_valid), rows, vocab, grad_scale, int(write_grad), _p(err_flag), _stream()), "cross_entropy")


def count_valid_targets(targets, n_valid):
    check(_lib.lib().mgb_count_valid_targets(_p(targets), targets.numel(), _p(n_valid), _stream()), "count_valid")


def loss_finalize(row_loss, n_valid, loss):
    check(_lib.lib().mgb_loss_finalize(_p(row_loss), row_loss.numel(), _p(n_valid), _p(loss), _stream()), "loss_finalize")


def colsum(x, out):
    _chk2d(x, torch.bfloat16, "colsum x")
    check(_lib.lib().mgb_colsum_bf16(_p(x), x.stride(0), x.shape[0], x.shape[1], _p(out), _stream()), "colsum")


def act_bwd(dy, z, dz, act):
    check(_lib.lib().mgb_act_bwd_bf16(_p(dy), _p(z), _p(dz), dy.numel(), act, _stream()), "act_bwd")


def act_fwd(z, res, y, act):
    check(_lib.lib().mgb_act_fwd_bf16(_p(z), _p(res), _p(y), z.numel(), act, _stream()), "act_fwd")


def cast_f32_to_bf16(src, dst):
    check(_lib.lib().mgb_cast_f32_to_bf16(_p(src), _p(dst), src.numel(), _stream()), "cast")


def cast_bf16_to_f32(src, dst):
    check(_lib.lib().mgb_cast_bf16_to_f32(_p(src), _p(dst), src.numel(), _stream()), "cast")


def axpy(dst, src, alpha):
    if isinstance(alpha, torch.Tensor):
        check(_lib.lib().mgb_axpy_f32_dev(_p(dst), _p(src), _p(alpha), dst.numel(), _stream()), "axpy")
    else:
        check(_lib.lib().mgb_axpy_f32(_p(dst), _p(src), float(alpha), dst.numel(), _stream()), "axpy")


def scale_bf16_dev(x, alpha_dev):
    check(_lib.lib().mgb_scale_bf16_dev(_p(x), x.numel(), _p(alpha_dev), _stream()), "scale")


def sumsq(x, out):
    check(_lib.lib().mgb_sumsq_f32(_p(x), x.numel(), _p(out), _stream()), "sumsq")


def clip_scale(g, gradnorm_sq, max_norm):
    check(_lib.lib().mgb_clip_scale_f32(_p(g), g.numel(), _p(gradnorm_sq), float(max_norm), _stream()), "clip_scale")


def adamw_multi(params, grads, exp_avg, exp_avg_sq, params_bf16, table_dev, num_chunks, lr, beta1, beta2, eps, step,
                grad_scale=1.0, gradnorm_sq=None, max_norm=0.0, step_dev=None):
    check(_lib.lib().mgb_adamw_multi(_p(params), _p(grads), _p(exp_avg), _p(exp_avg_sq), _p(params_bf16), _p(table_dev),
                                     num_chunks, lr, beta1, beta2, eps, step, grad_scale, _p(gradnorm_sq), max_norm,
                                     _p(step_dev), _stream()), "adamw")


def make_adamw_table(entries, chunk: int = 8192) -> torch.Tensor:
    """entries: iterable of (offset, numel, weight_decay) -> uint8 host tensor holding mgb_adamw_chunk[]"""
    rows = []
    for off, numel, wd in entries:
        done = 0
        while done < numel:
            n = min(chunk, numel - done)
            rows.append((off + done, n, wd))
            done += n
    arr = (AdamWChunk * len(rows))()
    for i, (o, n, wd) in enumerate(rows):
        arr[i].offset, arr[i].numel, arr[i].weight_decay = o, n, wd
    buf = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone() if rows else torch.empty(0, dtype=torch.uint8)
    return buf, len(rows)


# ----------------------------------------------------------------------------------------------------
# torch custom ops (torch.ops.mingpt_b200.*) with autograd: the per-op drop-ins and the units the
# parity tests exercise.  bf16 activations, fp32 parameters / statistics.
# ----------------------------------------------------------------------------------------------------
_lib_def = torch.library.Library("mingpt_b200", "DEF")
_lib_def.define("layernorm(Tensor x, Tensor gamma, Tensor beta, float eps) -> (Tensor, Tensor, Tensor)")
_lib_def.define("layernorm_backward(Tensor dy, Tensor x, Tensor gamma, Tensor mean, Tensor rstd) -> (Tensor, Tensor, Tensor)")
_lib_def.define("linear(Tensor x, Tensor w, Tensor? bias, Tensor? residual, int act) -> (Tensor, Tensor)")
_lib_def.define("linear_backward(Tensor dy, Tensor x, Tensor w, Tensor? preact, int act, bool has_bias) -> (Tensor, Tensor, Tensor)")
_lib_def.define("attention(Tensor qkv, int n_head, int mask_mode) -> (Tensor, Tensor)")
_lib_def.define("attention_backward(Tensor dout, Tensor qkv, Tensor out, Tensor lse, int n_head, int mask_mode) -> Tensor")
_lib_def.define("embedding(Tensor idx, Tensor wte, Tensor wpe) -> Tensor")
_lib_def.define("embedding_backward(Tensor dx, Tensor idx, int vocab, int block_size) -> (Tensor, Tensor)")
_lib_def.define("cross_entropy(Tensor logits, Tensor targets) -> (Tensor, Tensor)")


def _impl(name):
    def deco(fn):
        _lib_def.impl(name, fn, "CUDA")
        return fn
    return deco


@_impl("layernorm")
def _layernorm(x, gamma, beta, eps):
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    y = torch.empty_like(x2)
    mean = torch.empty(x2.shape[0], dtype=torch.float32, device=x.device)
    rstd = torch.empty_like(mean)
    layernorm_fwd(x2, gamma.contiguous(), beta.contiguous(), y, mean, rstd, eps)
    return y.view(x.shape), mean, rstd


@_impl("layernorm_backward")
def _layernorm_backward(dy, x, gamma, mean, rstd):
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    dy2 = dy.reshape(-1, x.shape[-1]).contiguous()
    dx = torch.empty_like(x2)
    dg = torch.zeros_like(gamma, dtype=torch.float32)
    db = torch.zeros_like(gamma, dtype=torch.float32)
    layernorm_bwd(dy2, x2, gamma.contiguous(), mean, rstd, dx, dg, db)
    return dx.view(x.shape), dg, db


@_impl("linear")
def _linear(x, w, bias, residual, act):
    x2 = x.reshape(-1, x.shape[-1])
    out = torch.empty(x2.shape[0], w.shape[0], dtype=torch.bfloat16, device=x.device)
    pre = torch.empty_like(out) if act != 0 else out
    res2 = residual.reshape(-1, w.shape[0]) if residual is not None else None
    linear_fwd(x2, w, out, bias=bias, act=act, residual=res2, preact=pre if act != 0 else None)
    return out.view(*x.shape[:-1], w.shape[0]), pre.view(*x.shape[:-1], w.shape[0])


@_impl("linear_backward")
def _linear_backward(dy, x, w, preact, act, has_bias):
    N, K = w.shape
    dy2 = dy.reshape(-1, N).contiguous()
    x2 = x.reshape(-1, K)
    if act != 0:
        dz = torch.empty_like(dy2)
        act_bwd(dy2, preact.reshape(-1, N).contiguous(), dz, act)
        dy2 = dz
    dx = torch.empty(x2.shape, dtype=torch.bfloat16, device=x.device)
    linear_dgrad(dy2, w, dx)
    dw = torch.zeros(N, K, dtype=torch.float32, device=x.device)
    linear_wgrad(dy2, x2, dw)
    db = torch.zeros(N, dtype=torch.float32, device=x.device)
    if has_bias:
        colsum(dy2, db)
    return dx.view(x.shape), dw, db


@_impl("attention")
def _attention(qkv, n_head, mask_mode):
    B, T, C3 = qkv.shape
    Cdim = C3 // 3
    d = Cdim // n_head
    qkv = qkv.contiguous()
    out = torch.empty(B, T, Cdim, dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty(B, n_head, T, dtype=torch.float32, device=qkv.device)
    attention_fwd(qkv, out, lse, B, T, n_head, d, mask_mode)
    return out, lse


@_impl("attention_backward")
def _attention_backward(dout, qkv, out, lse, n_head, mask_mode):
    B, T, C3 = qkv.shape
    Cdim = C3 // 3
    d = Cdim // n_head
    dqkv = torch.empty_like(qkv)
    delta = torch.empty_like(lse)
    dq_acc = torch.empty(B, T, Cdim, dtype=torch.float32, device=qkv.device)
    attention_bwd(qkv.contiguous(), out, dout.contiguous(), lse, delta, dq_acc, dqkv, B, T, n_head, d, mask_mode)
    return dqkv


@_impl("embedding")
def _embedding(idx, wte, wpe):
    B, T = idx.shape
    out = torch.empty(B, T, wte.shape[1], dtype=torch.bfloat16, device=idx.device)
    embedding_fwd(idx.contiguous(), wte.contiguous(), wpe.contiguous(), out)
    return out


@_impl("embedding_backward")
def _embedding_backward(dx, idx, vocab, block_size):
    Cdim = dx.shape[-1]
    dwte = torch.zeros(vocab, Cdim, dtype=torch.float32, device=dx.device)
    dwpe = torch.zeros(block_size, Cdim, dtype=torch.float32, device=dx.device)
    embedding_bwd(idx.contiguous(), dx.contiguous(), dwte, dwpe)
    return dwte, dwpe


@_impl("cross_entropy")
def _cross_entropy(logits, targets):
    """returns (mean loss over targets != -1, dlogits for an upstream gradient of 1)"""
    V = logits.shape[-1]
    l2 = logits.reshape(-1, V)
    ld = (V + 7) // 8 * 8
    work = torch.zeros(l2.shape[0], ld, dtype=torch.bfloat16, device=logits.device)
    work[:, :V] = l2
    tg = targets.reshape(-1).contiguous()
    row_loss = torch.empty(l2.shape[0], dtype=torch.float32, device=logits.device)
    n_valid = torch.zeros(1, dtype=torch.int32, device=logits.device)
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    count_valid_targets(tg, n_valid)
    cross_entropy_fwd_bwd(work, tg, row_loss, n_valid, V)
    loss_finalize(row_loss, n_valid, loss)
    return loss, work[:, :V].reshape(logits.shape)


def _fake(name, fn):
    torch.library.register_fake(f"mingpt_b200::{name}", fn, lib=_lib_def)


_fake("layernorm", lambda x, g, b, eps: (torch.empty_like(x), x.new_empty(x.numel() // x.shape[-1], dtype=torch.float32), x.new_empty(x.numel() // x.shape[-1], dtype=torch.float32)))
_fake("linear", lambda x, w, bias, residual, act: (x.new_empty(*x.shape[:-1], w.shape[0]), x.new_empty(*x.shape[:-1], w.shape[0])))
_fake("attention", lambda qkv, h, m: (qkv.new_empty(qkv.shape[0], qkv.shape[1], qkv.shape[2] // 3), qkv.new_empty(qkv.shape[0], h, qkv.shape[1], dtype=torch.float32)))
_fake("embedding", lambda idx, wte, wpe: wte.new_empty(*idx.shape, wte.shape[1]))
_fake("cross_entropy", lambda logits, targets: (logits.new_empty((), dtype=torch.float32), torch.empty_like(logits)))


def _ln_setup(ctx, inputs, output):
    x, gamma, beta, eps = inputs
    _, mean, rstd = output
    ctx.save_for_backward(x, gamma, mean, rstd)


def _ln_bwd(ctx, dy, dmean, drstd):
    x, gamma, mean, rstd = ctx.saved_tensors
    dx, dg, db = torch.ops.mingpt_b200.layernorm_backward(dy, x, gamma, mean, rstd)
    return dx, dg, db, None


torch.library.register_autograd("mingpt_b200::layernorm", _ln_bwd, setup_context=_ln_setup, lib=_lib_def)


def _lin_setup(ctx, inputs, output):
    x, w, bias, residual, act = inputs
    _, pre = output
    ctx.act, ctx.has_bias, ctx.has_res = act, bias is not None, residual is not None
    ctx.save_for_backward(x, w, pre if act != 0 else None)


def _lin_bwd(ctx, dy, dpre):
    x, w, pre = ctx.saved_tensors
    dx, dw, db = torch.ops.mingpt_b200.linear_backward(dy, x, w, pre, ctx.act, ctx.has_bias)
    return dx, dw.to(w.dtype), (db if ctx.has_bias else None), (dy if ctx.has_res else None), None


torch.library.register_autograd("mingpt_b200::linear", _lin_bwd, setup_context=_lin_setup, lib=_lib_def)


def _att_setup(ctx, inputs, output):
    qkv, n_head, mask_mode = inputs
    out, lse = output
    ctx.n_head, ctx.mask_mode = n_head, mask_mode
    ctx.save_for_backward(qkv, out, lse)


def _att_bwd(ctx, dout, dlse):
    qkv, out, lse = ctx.saved_tensors
    return torch.ops.mingpt_b200.attention_backward(dout, qkv, out, lse, ctx.n_head, ctx.mask_mode), None, None


torch.library.register_autograd("mingpt_b200::attention", _att_bwd, setup_context=_att_setup, lib=_lib_def)


def _emb_setup(ctx, inputs, output):
    idx, wte, wpe = inputs
    ctx.vocab, ctx.block = wte.shape[0], wpe.shape[0]
    ctx.save_for_backward(idx)


def _emb_bwd(ctx, dx):
    (idx,) = ctx.saved_tensors
    dwte, dwpe = torch.ops.mingpt_b200.embedding_backward(dx, idx, ctx.vocab, ctx.block)
    return None, dwte.to(dx.dtype), dwpe.to(dx.dtype)


torch.library.register_autograd("mingpt_b200::embedding", _emb_bwd, setup_context=_emb_setup, lib=_lib_def)


def _ce_setup(ctx, inputs, output):
    _, dlogits = output
    ctx.save_for_backward(dlogits)


def _ce_bwd(ctx, dloss, _unused):
    (dlogits,) = ctx.saved_tensors
    return dlogits * dloss.to(dlogits.dtype), None


torch.library.register_autograd("mingpt_b200::cross_entropy", _ce_bwd, setup_context=_ce_setup, lib=_lib_def)
