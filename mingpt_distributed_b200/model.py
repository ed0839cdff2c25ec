"""GPT language model with the reference's API (mingpt/model.py) and a B200-native training step.

Public surface kept from the reference (file:line = /root/reference/mingpt/model.py):
  GPTConfig :38-51, OptimizerConfig :54-59, create_optimizer :62-122, print_model_size :21-33,
  GPT.__init__ :237-259, GPT.forward(inputs, targets=None) -> (logits, loss) :309-320, GPT.generate :322-356
and the spellings its README (README.md:29-54) and the task's north_star use:
  GPT.get_default_config(), GPT(config), model.configure_optimizers(train_config).

Two module graphs share one set of kernels ("fidelity decision"):
  variant="reference"  the graph mingpt/model.py really builds: float tril mask ADDED to the scores
                       (nothing is masked), MHA out_proj followed by a second c_proj, MLP = Linear ->
                       Linear -> GELU(erf); parameter names and initialisation identical to the reference,
                       so its snapshots load and `torch.manual_seed(s); GPT(cfg)` gives the same weights.
  variant="mingpt"     the graph the README / north_star describe: causal attention, c_attn / c_proj,
                       c_fc -> NewGELU(tanh) -> c_proj, lm_head; parameter names of upstream minGPT.

nn.Parameters stay ordinary fp32 leaves (state_dict, parameters(), foreign optimizers all work) but live
as views of ONE flat fp32 buffer, with a flat bf16 shadow for the GEMM operands and a flat fp32 gradient
buffer the backward pass accumulates into -- the layout the fused AdamW and the bucketed all-reduce use.
The forward/backward themselves are hand-scheduled launches of the C-ABI kernels (ops.py); there is no
PyTorch-math path on the device side.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import ops

PRESETS = {
    # names follow the huggingface naming conventions (mingpt/model.py:269-294)
    "openai-gpt": dict(n_layer=12, n_head=12, n_embed=768),
    "gpt2": dict(n_layer=12, n_head=12, n_embed=768),
    "gpt2-medium": dict(n_layer=24, n_head=16, n_embed=1024),
    "gpt2-large": dict(n_layer=36, n_head=20, n_embed=1280),
    "gpt2-xl": dict(n_layer=48, n_head=25, n_embed=1600),
    "gopher-44m": dict(n_layer=8, n_head=16, n_embed=512),
    "gpt-mini": dict(n_layer=6, n_head=6, n_embed=192),
    "gpt-micro": dict(n_layer=4, n_head=4, n_embed=128),
    "gpt-nano": dict(n_layer=3, n_head=3, n_embed=48),
}


def print_model_size(model: nn.Module):
    """params + buffers in MB (mingpt/model.py:21-33)"""
    size = sum(p.numel() * p.element_size() for p in model.parameters())
    size += sum(b.numel() * b.element_size() for b in model.buffers())
    print(f"Model size (MB): {size / 1024**2:.3f}")


@dataclass
class GPTConfig:
    """mingpt/model.py:38-51. `n_embd` is accepted as an alias of `n_embed` (the reference's YAML and preset
    table spell it that way, mingpt/gpt2_config.yaml:4 / model.py:273-293, and crash on it: SURVEY D4, D6)."""
    model_type: Optional[str] = "gpt2"
    n_layer: Optional[int] = None
    n_head: Optional[int] = None
    n_embed: Optional[int] = None
    vocab_size: int = 50257
    block_size: int = 1024
    embed_drop: float = 0.1
    resid_drop: float = 0.1
    attn_drop: float = 0.1
    n_embd: Optional[int] = None
    variant: str = "mingpt"

    def __post_init__(self):
        if self.n_embed is None and self.n_embd is not None:
            self.n_embed = self.n_embd
        self.n_embd = self.n_embed


@dataclass
class OptimizerConfig:
    """mingpt/model.py:54-59"""
    learning_rate: float = 3e-4
    weight_decay: float = 0.1
    betas: tuple = (0.9, 0.95)


class CfgNode:
    """Minimal attribute bag in the spirit of upstream minGPT's CfgNode (README.md:31-38 usage)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def merge_from_dict(self, d):
        self.__dict__.update(d)

    def to_dict(self):
        return dict(self.__dict__)

    def __repr__(self):
        return "CfgNode(" + ", ".join(f"{k}={v!r}" for k, v in self.__dict__.items()) + ")"


def _to_gpt_config(config) -> GPTConfig:
    if isinstance(config, GPTConfig):
        return config
    d = config.to_dict() if hasattr(config, "to_dict") else dict(vars(config))
    return GPTConfig(
        model_type=d.get("model_type", None),
        n_layer=d.get("n_layer"), n_head=d.get("n_head"),
        n_embed=d.get("n_embed", d.get("n_embd")),
        vocab_size=d.get("vocab_size", 50257), block_size=d.get("block_size", 1024),
        embed_drop=d.get("embed_drop", d.get("embd_pdrop", 0.1)),
        resid_drop=d.get("resid_drop", d.get("resid_pdrop", 0.1)),
        attn_drop=d.get("attn_drop", d.get("attn_pdrop", 0.1)),
        variant=d.get("variant", "mingpt"),
    )


# ---------------------------------------------------------------------------------------------------
# parameter containers.  They are real torch modules of the same classes the reference instantiates, so
# construction consumes the RNG exactly as the reference does and module-type based code (the reference's
# own create_optimizer, mingpt/model.py:70-93) classifies the parameters identically.  Their forward()
# is never used: the compute goes through the engine below.
# ---------------------------------------------------------------------------------------------------
class _RefEmbedding(nn.Module):          # GPTEmbedding, mingpt/model.py:193-231
    def __init__(self, cfg):
        super().__init__()
        self.token_embedding = nn.Embedding(cfg.vocab_size, cfg.n_embed)
        self.pos_embedding = nn.Parameter(torch.zeros(1, cfg.block_size, cfg.n_embed))
        self.drop = nn.Dropout(cfg.embed_drop)


class _RefAttention(nn.Module):          # MultiHeadSelfAttention, mingpt/model.py:125-168
    def __init__(self, cfg):
        super().__init__()
        assert cfg.n_embed % cfg.n_head == 0
        self.c_proj = nn.Linear(cfg.n_embed, cfg.n_embed)
        self.resid_drop = nn.Dropout(cfg.resid_drop)
        # the reference also registers a (1,1,T,T) float `mask` buffer per layer (model.py:142-145); the
        # attention kernel applies it implicitly, so it is not materialised (nor broadcast by DDP).
        self.attn = nn.MultiheadAttention(embed_dim=cfg.n_embed, num_heads=cfg.n_head, dropout=cfg.attn_drop,
                                          batch_first=True)


class _RefBlock(nn.Module):              # Block, mingpt/model.py:171-189
    def __init__(self, cfg):
        super().__init__()
        self.ln_1 = nn.LayerNorm(cfg.n_embed)
        self.attn = _RefAttention(cfg)
        self.ln_2 = nn.LayerNorm(cfg.n_embed)
        self.mlp = nn.Sequential(nn.Linear(cfg.n_embed, 4 * cfg.n_embed), nn.Linear(4 * cfg.n_embed, cfg.n_embed),
                                 nn.GELU(), nn.Dropout(cfg.resid_drop))


class _MinAttention(nn.Module):          # upstream minGPT CausalSelfAttention (README.md:10-15)
    def __init__(self, cfg):
        super().__init__()
        assert cfg.n_embed % cfg.n_head == 0
        self.c_attn = nn.Linear(cfg.n_embed, 3 * cfg.n_embed)
        self.c_proj = nn.Linear(cfg.n_embed, cfg.n_embed)
        self.attn_dropout = nn.Dropout(cfg.attn_drop)
        self.resid_dropout = nn.Dropout(cfg.resid_drop)


class _MinBlock(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.ln_1 = nn.LayerNorm(cfg.n_embed)
        self.attn = _MinAttention(cfg)
        self.ln_2 = nn.LayerNorm(cfg.n_embed)
        self.mlp = nn.ModuleDict(dict(c_fc=nn.Linear(cfg.n_embed, 4 * cfg.n_embed),
                                      c_proj=nn.Linear(4 * cfg.n_embed, cfg.n_embed),
                                      dropout=nn.Dropout(cfg.resid_drop)))


_ALIGN = 64  # elements: every parameter starts 256-byte (fp32) / 128-byte (bf16 shadow) aligned


_DROP_SIDE_STREAM = os.environ.get("MGB_DROP_SIDE_STREAM", "0") == "1"


class _Nvtx:
    """NVTX ranges around the phases of a step (embedding, each block's forward / backward, head + loss) so that an
    nsys / ncu timeline reads like the model. Off unless MGB_NVTX=1: a range is two host calls per phase."""

    def __init__(self):
        import os
        self.on = os.environ.get("MGB_NVTX", "0") == "1"

    def push(self, name: str):
        if self.on:
            torch.cuda.nvtx.range_push(name)

    def pop(self):
        if self.on:
            torch.cuda.nvtx.range_pop()


_nvtx = _Nvtx()


class LazyLogits:
    """What forward() returns as `logits` when a loss is requested on the fused path: the full
    (B, T, vocab) tensor is only produced if somebody asks (mingpt/trainer.py:122 discards it)."""

    def __init__(self, model, ws, generation):
        self._model, self._ws, self._gen = model, ws, generation
        self.shape = torch.Size((ws.B, ws.T, model.vocab_size))
        self.dtype = torch.float32

    def size(self, dim=None):
        return self.shape if dim is None else self.shape[dim]

    def materialize(self) -> torch.Tensor:
        if self._ws.generation != self._gen:
            raise RuntimeError("LazyLogits is stale: the model ran another forward since it was returned")
        self._model.check_device_errors()
        return self._model._head_logits(self._ws)

    def view(self, *shape):
        return self.materialize().view(*shape)

    def __getitem__(self, item):
        return self.materialize()[item]

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        conv = lambda a: a.materialize() if isinstance(a, LazyLogits) else a
        return func(*[conv(a) for a in args], **{k: conv(v) for k, v in kwargs.items()})


class _Workspace:
    """Activation / gradient scratch for one (B, T): allocated once through torch's caching allocator and
    reused every step, so the step performs no allocation (and can be captured in a CUDA graph)."""

    def __init__(self, model: "GPT", B: int, T: int):
        dev = model._flat.device
        C, L, H, V = model.n_embed, model.n_layer, model.n_head, model.vocab_size
        M = B * T
        self.B, self.T, self.M = B, T, M
        self.generation = 0
        bf = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        ref = model.variant == "reference"
        self.x = [bf(M, C) for _ in range(L + 1)]
        self.xmid = [bf(M, C) for _ in range(L)]
        self.ln1 = [bf(M, C) for _ in range(L)]
        self.ln2 = [bf(M, C) for _ in range(L)]
        self.stats = [[f32(M) for _ in range(4)] for _ in range(L)]  # ln1 mean/rstd, ln2 mean/rstd
        self.qkv = [bf(M, 3 * C) for _ in range(L)]
        self.att = [bf(M, C) for _ in range(L)]
        self.att_o = [bf(M, C) for _ in range(L)] if ref else None
        self.lse = [f32(B, H, T) for _ in range(L)]
        self.fc_pre = [bf(M, 4 * C) for _ in range(L)]
        self.fc_act = [bf(M, 4 * C) for _ in range(L)] if not ref else None
        self.z = [bf(M, C) for _ in range(L)] if ref else None
        self.lnf = bf(M, C)
        self.lnf_stats = [f32(M), f32(M)]
        # head / loss
        self.chunk = min(M, model.head_chunk_rows)
        self.vpad = (V + 63) // 64 * 64
        self.logits = bf(self.chunk, self.vpad)
        self.row_loss = f32(M)
        self.n_valid = torch.zeros(1, dtype=torch.int32, device=dev)
        self.loss = f32(())
        self.err_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        # backward scratch
        self.dlnf = bf(M, C)
        self.head_dw = f32(V, C)
        self.dx_a, self.dx_b = bf(M, C), bf(M, C)
        self.dln = bf(M, C)
        self.datt = bf(M, C)
        self.datt2 = bf(M, C) if ref else None
        self.dz = bf(M, C) if ref else None
        self.dfc = bf(M, 4 * C)
        self.dqkv = bf(M, 3 * C)
        self.dq_acc = f32(M, C)
        self.delta = f32(B, H, T)
        self.idx = None
        self.has_head_grads = False
        self.drop = (0.0, 0.0, 0.0)      # dropout probabilities used by the forward that filled this workspace
        self.seed = torch.zeros(1, dtype=torch.int64, device=dev)   # dropout seed of that forward (backward regenerates its masks)
        self.tmp = self.dtmp = None
        self.drop_bits = None            # per layer: keep words of the attention-probability dropout (head size 64)
        self._mc = (M, C, dev)

    def ensure_dropout_buffers(self):
        """GEMM output before residual dropout, and the masked gradient of it (training with dropout only)"""
        if self.tmp is None:
            M, C, dev = self._mc
            self.tmp = torch.empty(M, C, dtype=torch.bfloat16, device=dev)
            self.dtmp = torch.empty(M, C, dtype=torch.bfloat16, device=dev)

    def ensure_drop_bits(self, L, B, T, H):
        """one keep-word buffer per layer: written once by the forward's generator launch, read by the attention
        forward AND backward of that layer (B*H*T*T/8 bytes each; 50 MB at B=32, H=12, T=1024)"""
        if self.drop_bits is None:
            words = ops.attention_dropout_bits_words(B, T, H)
            dev = self._mc[2]
            self.drop_bits = [torch.empty(words, dtype=torch.int32, device=dev) for _ in range(L)]
            self.drop_stream = torch.cuda.Stream(device=dev)
            self.drop_fork = torch.cuda.Event()
            self.drop_ready = [torch.cuda.Event() for _ in range(L)]


class _DecodeState:
    """Key / value cache and one-row activations for incremental decoding (GPT.generate with use_cache): per layer a
    packed qkv buffer [B, block_size, 3C] -- the layout the attention kernels read -- whose row t holds position t."""

    def __init__(self, model: "GPT", B: int):
        dev = model._flat.device
        C, L = model.n_embed, model.n_layer
        bf = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.empty(*s, dtype=torch.float32, device=dev)
        self.B, self.cap = B, model.block_size
        self.cache = [bf(B, self.cap, 3 * C) for _ in range(L)]
        self.x, self.xmid, self.ln, self.att = bf(B, C), bf(B, C), bf(B, C), bf(B, C)
        self.fc = bf(B, 4 * C)
        self.mean, self.rstd = f32(B), f32(B)
        self.logits = f32(B, model.vocab_size)
        self.err_flag = torch.zeros(1, dtype=torch.int32, device=dev)


class _LossFn(torch.autograd.Function):
    """Ties the fused forward to autograd: loss.backward() (mingpt/trainer.py:127) lands here and runs
    the hand-scheduled backward, which writes into the flat gradient buffer and sets p.grad views."""

    @staticmethod
    def forward(ctx, anchor, model, ws):
        ctx.model, ctx.ws, ctx.gen = model, ws, ws.generation
        return ws.loss.clone()

    @staticmethod
    def backward(ctx, grad_loss):
        if ctx.ws.generation != ctx.gen:
            raise RuntimeError("backward() after another forward reused the activation workspace")
        ctx.model._backward(ctx.ws, grad_loss.contiguous().float())
        return None, None, None


class _LossFnParams(torch.autograd.Function):
    """The same, with every parameter as an autograd INPUT: backward returns the parameter gradients instead of
    publishing them itself, so they reach p.grad through autograd's AccumulateGrad nodes -- which is where
    torch.nn.parallel.DistributedDataParallel hangs its reducer hooks (mingpt/trainer.py:71). Used when the model
    finds itself wrapped in torch's DDP (or when GPT.autograd_param_grads is set): DDP then buckets, all-reduces
    and averages the gradients exactly as it does for the reference. The returned tensors are views of the flat
    gradient buffer, so AccumulateGrad normally adopts them without a copy."""

    @staticmethod
    def forward(ctx, anchor, model, ws, *params):
        ctx.model, ctx.ws, ctx.gen = model, ws, ws.generation
        return ws.loss.clone()

    @staticmethod
    def backward(ctx, grad_loss):
        if ctx.ws.generation != ctx.gen:
            raise RuntimeError("backward() after another forward reused the activation workspace")
        grads = ctx.model._backward_returning_grads(ctx.ws, grad_loss.contiguous().float())
        return (None, None, None) + tuple(grads)


def _called_from_torch_ddp() -> bool:
    import sys
    f = sys._getframe(2)
    for _ in range(16):
        if f is None:
            return False
        if f.f_code.co_filename.endswith("torch/nn/parallel/distributed.py"):
            return True
        f = f.f_back
    return False


class GPT(nn.Module):
    """GPT Language Model (mingpt/model.py:234-356)."""

    @staticmethod
    def get_default_config() -> CfgNode:
        """README.md:31-38 spelling of the configuration object."""
        return CfgNode(model_type="gpt", n_layer=None, n_head=None, n_embd=None, vocab_size=None, block_size=None,
                       embd_pdrop=0.1, resid_pdrop=0.1, attn_pdrop=0.1, variant="mingpt")

    def __init__(self, config, variant: Optional[str] = None, head_chunk_rows: int = 16384, verbose: bool = True):
        super().__init__()
        config = _to_gpt_config(config)
        if variant is not None:
            config.variant = variant
        config = self._set_model_config(config, verbose)
        if config.variant not in ("reference", "mingpt"):
            raise ValueError("variant must be 'reference' or 'mingpt'")
        self.config = config
        self.variant = config.variant
        self.block_size = config.block_size
        self.vocab_size = config.vocab_size
        self.n_layer, self.n_head, self.n_embed = config.n_layer, config.n_head, config.n_embed
        self.head_dim = config.n_embed // config.n_head
        self.head_chunk_rows = head_chunk_rows
        self.drop_p = (config.embed_drop, config.resid_drop, config.attn_drop)
        L = config.n_layer
        if self.variant == "reference":
            # construction order == mingpt/model.py:243-249
            self.embedding = _RefEmbedding(config)
            self.blocks = nn.Sequential(*[_RefBlock(config) for _ in range(L)])
            self.ln_ff = nn.LayerNorm(config.n_embed)
            self.head = nn.Linear(config.n_embed, config.vocab_size, bias=False)
        else:
            self.transformer = nn.ModuleDict(dict(
                wte=nn.Embedding(config.vocab_size, config.n_embed),
                wpe=nn.Embedding(config.block_size, config.n_embed),
                drop=nn.Dropout(config.embed_drop),
                h=nn.ModuleList([_MinBlock(config) for _ in range(L)]),
                ln_f=nn.LayerNorm(config.n_embed)))
            self.lm_head = nn.Linear(config.n_embed, config.vocab_size, bias=False)
        # mingpt/model.py:251-256
        self.apply(self._init_weights)
        for name, param in self.named_parameters():
            if name.endswith("c_proj.weight"):
                param.data.normal_(mean=0.0, std=0.02 / math.sqrt(2 * L))
        if verbose:
            print(f"Number of parameters: {sum(p.numel() for p in self.parameters())}")
            print_model_size(self)
        # engine state
        self._roles: Dict[str, str] = self._role_table()
        self._flat = self._flat_bf16 = self._flat_grad = None
        self._spans: Dict[str, Tuple[int, int]] = {}
        self._shadow_token = None
        self._workspaces: Dict[Tuple[int, int], _Workspace] = {}
        self._anchor = None
        self._layer_done_hook: Optional[Callable[[int], None]] = None
        self._rng_seed: Optional[torch.Tensor] = None   # device int64: dropout seed, advanced once per training forward
        self.autograd_param_grads = False     # True: gradients reach p.grad through autograd (torch DDP's hooks fire)
        self._grad_target: Optional[torch.Tensor] = None
        self._warned_ddp = False
        self.loss_scale_hint = 1.0
        self._flatten()

    # ------------------------------------------------------------------ configuration
    @staticmethod
    def _set_model_config(config: GPTConfig, verbose=True) -> GPTConfig:
        """mingpt/model.py:261-296 with the intended semantics: explicit n_layer/n_head/n_embed win,
        otherwise the preset named by model_type fills them in (the reference does neither: SURVEY D4/D5)."""
        params_given = all(v is not None for v in (config.n_layer, config.n_head, config.n_embed))
        if not params_given:
            if config.model_type not in PRESETS:
                raise ValueError("give n_layer/n_head/n_embed or a known model_type: " + ", ".join(PRESETS))
            for k, v in PRESETS[config.model_type].items():
                if getattr(config, k) is None:
                    setattr(config, k, v)
        config.n_embd = config.n_embed
        if config.vocab_size is None or config.block_size is None:
            raise ValueError("vocab_size and block_size must be set")
        if verbose:
            print(config)
        return config

    @staticmethod
    def _init_weights(module):
        """mingpt/model.py:298-307"""
        if isinstance(module, nn.Linear):
            module.weight.data.normal_(mean=0.0, std=0.02)
            if module.bias is not None:
                module.bias.data.zero_()
        elif isinstance(module, nn.Embedding):
            module.weight.data.normal_(mean=0.0, std=0.02)
        elif isinstance(module, nn.LayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)

    def _role_table(self) -> Dict[str, str]:
        """engine role -> parameter name"""
        r = {}
        if self.variant == "reference":
            r["wte"], r["wpe"] = "embedding.token_embedding.weight", "embedding.pos_embedding"
            for i in range(self.n_layer):
                b = f"blocks.{i}."
                r.update({f"{i}.ln1_w": b + "ln_1.weight", f"{i}.ln1_b": b + "ln_1.bias",
                          f"{i}.qkv_w": b + "attn.attn.in_proj_weight", f"{i}.qkv_b": b + "attn.attn.in_proj_bias",
                          f"{i}.o_w": b + "attn.attn.out_proj.weight", f"{i}.o_b": b + "attn.attn.out_proj.bias",
                          f"{i}.proj_w": b + "attn.c_proj.weight", f"{i}.proj_b": b + "attn.c_proj.bias",
                          f"{i}.ln2_w": b + "ln_2.weight", f"{i}.ln2_b": b + "ln_2.bias",
                          f"{i}.fc_w": b + "mlp.0.weight", f"{i}.fc_b": b + "mlp.0.bias",
                          f"{i}.fc2_w": b + "mlp.1.weight", f"{i}.fc2_b": b + "mlp.1.bias"})
            r.update({"lnf_w": "ln_ff.weight", "lnf_b": "ln_ff.bias", "head_w": "head.weight"})
        else:
            r["wte"], r["wpe"] = "transformer.wte.weight", "transformer.wpe.weight"
            for i in range(self.n_layer):
                b = f"transformer.h.{i}."
                r.update({f"{i}.ln1_w": b + "ln_1.weight", f"{i}.ln1_b": b + "ln_1.bias",
                          f"{i}.qkv_w": b + "attn.c_attn.weight", f"{i}.qkv_b": b + "attn.c_attn.bias",
                          f"{i}.proj_w": b + "attn.c_proj.weight", f"{i}.proj_b": b + "attn.c_proj.bias",
                          f"{i}.ln2_w": b + "ln_2.weight", f"{i}.ln2_b": b + "ln_2.bias",
                          f"{i}.fc_w": b + "mlp.c_fc.weight", f"{i}.fc_b": b + "mlp.c_fc.bias",
                          f"{i}.fc2_w": b + "mlp.c_proj.weight", f"{i}.fc2_b": b + "mlp.c_proj.bias"})
            r.update({"lnf_w": "transformer.ln_f.weight", "lnf_b": "transformer.ln_f.bias", "head_w": "lm_head.weight"})
        return r

    # ------------------------------------------------------------------ flat storage
    def _flatten(self, device=None):
        """(Re)pack every parameter into one flat fp32 buffer; Parameter objects keep their identity."""
        named = list(self.named_parameters())
        device = device if device is not None else named[0][1].device
        off, spans = 0, {}
        for name, p in named:
            spans[name] = (off, p.numel())
            off += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        flat = torch.zeros(off, dtype=torch.float32, device=device)
        for name, p in named:
            o, n = spans[name]
            view = flat[o:o + n].view(p.shape)
            view.copy_(p.data)
            p.data = view
        self._flat, self._spans, self._total = flat, spans, off
        self._flat_bf16 = self._flat_grad = None
        self._shadow_token = None
        self._workspaces = {}
        self._layer_spans = self._compute_layer_spans()

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        first = next(self.parameters(), None)
        if first is not None and self._flat is not None:
            if any(p.dtype != torch.float32 for p in self.parameters()):
                raise RuntimeError("GPT keeps fp32 master parameters (bf16 compute copies are internal); "
                                   "do not cast the module")
            self._flatten(first.device)
        return out

    def _compute_layer_spans(self) -> List[Tuple[int, int]]:
        """[start, end) element ranges of the flat buffers: embeddings, each block, final norm + head."""
        names = [n for n, _ in self.named_parameters()]
        def group(n):
            parts = n.split(".")
            if self.variant == "reference":
                return 1 + int(parts[1]) if parts[0] == "blocks" else (0 if parts[0] == "embedding" else self.n_layer + 1)
            if parts[0] == "transformer" and parts[1] == "h":
                return 1 + int(parts[2])
            return 0 if (parts[0] == "transformer" and parts[1] in ("wte", "wpe")) else self.n_layer + 1
        spans: Dict[int, List[int]] = {}
        for n in names:
            o, k = self._spans[n]
            g = group(n)
            end = o + (k + _ALIGN - 1) // _ALIGN * _ALIGN
            s = spans.setdefault(g, [o, end])
            s[0], s[1] = min(s[0], o), max(s[1], end)
        return [tuple(spans[g]) for g in sorted(spans)]

    def _param_token(self):
        return sum(p._version for p in self.parameters())

    def mark_params_dirty(self):
        """Call after modifying parameters through `.data` (invisible to version counters)."""
        self._shadow_token = None

    def _ensure_device_state(self):
        if not self._flat.is_cuda:
            raise RuntimeError("this GPT runs on a CUDA (B200) device only: call model.to('cuda') first "
                               "(there is no CPU compute path)")
        if self._flat_bf16 is None:
            self._flat_bf16 = torch.empty(self._total, dtype=torch.bfloat16, device=self._flat.device)
            self._shadow_token = None
        if self._flat_grad is None:
            self._flat_grad = torch.zeros(self._total, dtype=torch.float32, device=self._flat.device)
        if self._anchor is None or self._anchor.device != self._flat.device:
            self._anchor = torch.zeros((), device=self._flat.device, requires_grad=True)
            self._one = torch.ones((), device=self._flat.device)
        tok = self._param_token()
        if self._shadow_token != tok:
            ops.cast_f32_to_bf16(self._flat, self._flat_bf16)
            self._shadow_token = tok

    def _w(self, role: str, bf16: bool = False) -> torch.Tensor:
        name = self._roles[role]
        o, n = self._spans[name]
        src = self._flat_bf16 if bf16 else self._flat
        shape = self._shape_of[name]
        return src[o:o + n].view(shape)

    def _g(self, role: str) -> torch.Tensor:
        name = self._roles[role]
        o, n = self._spans[name]
        buf = self._flat_grad if self._grad_target is None else self._grad_target
        return buf[o:o + n].view(self._shape_of[name])

    @property
    def _shape_of(self):
        cache = self.__dict__.get("_shape_cache")
        if cache is None:
            cache = {n: p.shape for n, p in self.named_parameters()}
            self.__dict__["_shape_cache"] = cache
        return cache

    def check_device_errors(self):
        """Raise if a kernel flagged an out-of-range token id (embedding) or target (cross entropy) since the last
        check -- the device assert torch's nn.Embedding / F.cross_entropy would have hit (mingpt/model.py:226,316).
        The kernels themselves only set a flag (and produce zeros / ignore the row) so the step never syncs with
        the host; this reads the flags back, so call it where the host syncs anyway (the trainer does, whenever
        it prints the loss)."""
        for ws in self._workspaces.values():
            flag = int(ws.err_flag.item())
            if flag:
                ws.err_flag.zero_()
                what = {1: "token id outside [0, vocab_size)", 2: "target outside [0, vocab_size) and != -1"}.get(flag, f"flag {flag}")
                raise IndexError(f"GPT.forward: {what} (batch {ws.B} x {ws.T})")

    def _ws(self, B: int, T: int) -> _Workspace:
        ws = self._workspaces.get((B, T))
        if ws is None:
            ws = _Workspace(self, B, T)
            self._workspaces[(B, T)] = ws
        return ws

    # ------------------------------------------------------------------ forward
    SITE_EMBED = 0xFFFF

    def _dropout_now(self, ws: "_Workspace") -> Tuple[float, float, float]:
        """(embed, resid, attn) dropout probabilities in force for this forward: the config's values in
        training mode (the reference never leaves it: mingpt/trainer.py:118-122,180-183), zeros in eval mode.
        Draws a fresh seed for the step; backward regenerates the masks from it (the head-size-64 attention reads the
        keep words its forward generated instead)."""
        if not self.training or not any(p > 0 for p in self.drop_p):
            ws.drop = (0.0, 0.0, 0.0)
            return ws.drop
        if self._rng_seed is None or self._rng_seed.device != self._flat.device:
            # follows torch.manual_seed(): drawn from the default CPU generator once, then advanced on the device
            # (data-parallel ranks usually call torch.manual_seed with the same value so that they build the same model:
            # the rank is mixed in, or every rank would draw the same masks for its different shard of the batch)
            seed0 = int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())
            if torch.distributed.is_available() and torch.distributed.is_initialized():
                seed0 = (seed0 + torch.distributed.get_rank() * 0x9E3779B97F4A7C15) % (2 ** 62)
            self._rng_seed = torch.tensor([seed0], dtype=torch.int64, device=self._flat.device)
        self._rng_seed += 1
        # the seed travels with the workspace: a training forward of ANOTHER shape between this forward and its
        # backward advances the model-wide seed, and backward must still regenerate THIS forward's masks
        ws.seed.copy_(self._rng_seed)
        ws.ensure_dropout_buffers()
        ws.drop = self.drop_p
        return ws.drop

    def _trunk(self, ws: _Workspace, idx: torch.Tensor):
        """embeddings -> blocks -> final LayerNorm (mingpt/model.py:310-312 up to ln_ff)"""
        B, T = idx.shape
        C, H, d, L = self.n_embed, self.n_head, self.head_dim, self.n_layer
        ref = self.variant == "reference"
        mask = ops.MASK_ADD_TRIL_ONES if ref else ops.MASK_CAUSAL
        W = lambda r: self._w(r, bf16=True)
        P = lambda r: self._w(r)
        wpe = W("wpe").view(-1, C)
        pe, pr, pa = self._dropout_now(ws)
        seed = ws.seed
        _nvtx.push("embedding")
        ops.embedding_fwd(idx, W("wte"), wpe, ws.x[0], ws.err_flag)
        if pe > 0:   # nn.Dropout(embed_drop), mingpt/model.py:215,231
            ops.dropout(ws.x[0], None, ws.x[0], pe, seed, self.SITE_EMBED)
        _nvtx.pop()
        if pa > 0 and d == 64:
            # The keep words of every layer's attention dropout depend on the seed alone, so all of them are generated
            # here, ahead of the blocks. MGB_DROP_SIDE_STREAM=1 moves the generator to a side stream (layer l then waits
            # for its own event only): it is pure ALU work in small CTAs without shared memory and does share the SMs
            # with the tensor-bound GEMMs of the first layers, but on a power-capped part the step is bound by energy
            # rather than by issue slots and the overlap measured 0 (33.35 vs 33.21 ms), and torch.profiler / CUPTI has
            # been seen to hang on replayed CUDA graphs with parallel branches -- so it is off by default.
            ws.ensure_drop_bits(self.n_layer, B, T, H)
            side = ws.drop_stream if _DROP_SIDE_STREAM else None
            if side is not None:
                ws.drop_fork.record(torch.cuda.current_stream())
                side.wait_event(ws.drop_fork)
            with torch.cuda.stream(side):
                for l in range(L):
                    ops.attention_dropout_bits(ws.drop_bits[l], B, T, H, mask, pa, seed, 4 * l)
                    if side is not None:
                        ws.drop_ready[l].record(side)
        for l in range(L):
            _nvtx.pop() if l else None
            _nvtx.push(f"block{l}.fwd")
            s = ws.stats[l]
            ops.layernorm_fwd(ws.x[l], P(f"{l}.ln1_w"), P(f"{l}.ln1_b"), ws.ln1[l], s[0], s[1])
            ops.linear_fwd(ws.ln1[l], W(f"{l}.qkv_w"), ws.qkv[l], bias=P(f"{l}.qkv_b"))
            if pa > 0 and d == 64:   # dropout on the attention probabilities, mingpt/model.py:150
                if _DROP_SIDE_STREAM:
                    torch.cuda.current_stream().wait_event(ws.drop_ready[l])
                ops.attention_fwd_dropout_bits(ws.qkv[l], ws.att[l], ws.lse[l], B, T, H, d, mask, pa, ws.drop_bits[l])
            elif pa > 0:
                ops.attention_fwd_dropout(ws.qkv[l], ws.att[l], ws.lse[l], B, T, H, d, mask, pa, seed, 4 * l)
            else:
                ops.attention_fwd(ws.qkv[l], ws.att[l], ws.lse[l], B, T, H, d, mask)
            a_in = ws.att[l]
            if ref:  # nn.MultiheadAttention's own out_proj, then the extra c_proj (model.py:147-154,167)
                ops.linear_fwd(ws.att[l], W(f"{l}.o_w"), ws.att_o[l], bias=P(f"{l}.o_b"))
                a_in = ws.att_o[l]
            # resid_drop after c_proj (model.py:140,167), then the residual add (:187): in the projection's epilogue on
            # the CTA-pair GEMM engine (M > 128), a separate pass otherwise
            fuse_drop = pr > 0 and B * T > 128 and C % 16 == 0
            if fuse_drop:
                ops.linear_fwd(a_in, W(f"{l}.proj_w"), ws.xmid[l], bias=P(f"{l}.proj_b"), residual=ws.x[l],
                               dropout=(pr, seed, 4 * l + 1))
            elif pr > 0:
                ops.linear_fwd(a_in, W(f"{l}.proj_w"), ws.tmp, bias=P(f"{l}.proj_b"))
                ops.dropout(ws.tmp, ws.x[l], ws.xmid[l], pr, seed, 4 * l + 1)
            else:
                ops.linear_fwd(a_in, W(f"{l}.proj_w"), ws.xmid[l], bias=P(f"{l}.proj_b"), residual=ws.x[l])
            ops.layernorm_fwd(ws.xmid[l], P(f"{l}.ln2_w"), P(f"{l}.ln2_b"), ws.ln2[l], s[2], s[3])
            if ref:  # Linear -> Linear -> GELU(erf) -> Dropout (model.py:179-184)
                ops.linear_fwd(ws.ln2[l], W(f"{l}.fc_w"), ws.fc_pre[l], bias=P(f"{l}.fc_b"))
                if pr > 0:
                    ops.linear_fwd(ws.fc_pre[l], W(f"{l}.fc2_w"), ws.tmp, bias=P(f"{l}.fc2_b"),
                                   act=ops.ACT_GELU_ERF | ops.ACT_STORE_GRAD, preact=ws.z[l])
                    ops.dropout(ws.tmp, ws.xmid[l], ws.x[l + 1], pr, seed, 4 * l + 2)
                else:
                    ops.linear_fwd(ws.fc_pre[l], W(f"{l}.fc2_w"), ws.x[l + 1], bias=P(f"{l}.fc2_b"),
                                   act=ops.ACT_GELU_ERF | ops.ACT_STORE_GRAD, preact=ws.z[l], residual=ws.xmid[l])
            else:    # c_fc + NewGELU epilogue, then c_proj (+ dropout) + residual
                # the "pre-activation" buffers (fc_pre here, z above) receive GELU'(pre-activation): all backward needs
                ops.linear_fwd(ws.ln2[l], W(f"{l}.fc_w"), ws.fc_act[l], bias=P(f"{l}.fc_b"),
                               act=ops.ACT_GELU_TANH | ops.ACT_STORE_GRAD, preact=ws.fc_pre[l])
                if fuse_drop:
                    ops.linear_fwd(ws.fc_act[l], W(f"{l}.fc2_w"), ws.x[l + 1], bias=P(f"{l}.fc2_b"), residual=ws.xmid[l],
                                   dropout=(pr, seed, 4 * l + 2))
                elif pr > 0:
                    ops.linear_fwd(ws.fc_act[l], W(f"{l}.fc2_w"), ws.tmp, bias=P(f"{l}.fc2_b"))
                    ops.dropout(ws.tmp, ws.xmid[l], ws.x[l + 1], pr, seed, 4 * l + 2)
                else:
                    ops.linear_fwd(ws.fc_act[l], W(f"{l}.fc2_w"), ws.x[l + 1], bias=P(f"{l}.fc2_b"), residual=ws.xmid[l])
        _nvtx.pop() if L else None
        ops.layernorm_fwd(ws.x[L], P("lnf_w"), P("lnf_b"), ws.lnf, ws.lnf_stats[0], ws.lnf_stats[1])

    def _head_loss(self, ws: _Workspace, targets: torch.Tensor, need_grad: bool, direct: bool = False):
        """head + cross_entropy (mingpt/model.py:312,316-318) in row chunks; the bf16 logits of a chunk are
        turned into dlogits in place and consumed by the head's dgrad/wgrad right away, so the full
        (B*T, vocab) logits / dlogits never exist. `direct`: the upstream gradient of the loss is known to be 1
        (forward_backward), so the head's weight gradient accumulates straight into the flat gradient buffer
        instead of a scratch copy that backward would scale and add."""
        V, M = self.vocab_size, ws.M
        tg = targets.reshape(-1)
        hw = self._w("head_w", bf16=True)
        _nvtx.push("head+loss")
        ops.count_valid_targets(tg, ws.n_valid)
        head_dw = self._g("head_w") if direct else ws.head_dw
        if need_grad and not direct:
            head_dw.zero_()
        for r0 in range(0, M, ws.chunk):
            rows = min(ws.chunk, M - r0)
            lg_full = ws.logits[:rows]
            lg = lg_full[:, :V]
            ops.linear_fwd(ws.lnf[r0:r0 + rows], hw, lg)
            ops.cross_entropy_fwd_bwd(lg_full, tg[r0:r0 + rows], ws.row_loss[r0:r0 + rows], ws.n_valid, V,
                                      write_grad=need_grad, err_flag=ws.err_flag)
            if need_grad:
                ops.linear_dgrad(lg, hw, ws.dlnf[r0:r0 + rows])
                ops.linear_wgrad(lg, ws.lnf[r0:r0 + rows], head_dw, accumulate=True)
        ops.loss_finalize(ws.row_loss, ws.n_valid, ws.loss)
        _nvtx.pop()
        ws.has_head_grads = need_grad
        ws.head_direct = direct

    def _head_logits(self, ws: _Workspace, last_only: bool = False) -> torch.Tensor:
        V, C = self.vocab_size, self.n_embed
        hw = self._w("head_w", bf16=True)
        if last_only:
            x = ws.lnf.view(ws.B, ws.T, C)[:, -1, :].contiguous()
            out = torch.empty(ws.B, V, dtype=torch.float32, device=x.device)
            ops.linear_fwd(x, hw, out)
            return out
        out = torch.empty(ws.M, V, dtype=torch.float32, device=ws.lnf.device)
        ops.linear_fwd(ws.lnf, hw, out)
        return out.view(ws.B, ws.T, V)

    def forward(self, inputs: torch.Tensor, targets: Optional[torch.Tensor] = None):
        """mingpt/model.py:309-320: returns (logits, loss); loss is None without targets.
        With targets the logits come back as a LazyLogits handle (materialised on first use)."""
        if inputs.dim() != 2:
            raise ValueError("inputs must be (batch, seq)")
        B, T = inputs.shape
        assert T <= self.block_size, "Sequence length must be less than or equal to block size for positional embeddings."
        self._ensure_device_state()
        if not inputs.is_cuda:
            raise RuntimeError("inputs must be on the model's CUDA device")
        idx = inputs.contiguous()
        if idx.dtype != torch.int64:
            idx = idx.long()
        ws = self._ws(B, T)
        ws.generation += 1
        ws.idx = idx
        self._trunk(ws, idx)
        if targets is None:
            return self._head_logits(ws), None
        tg = targets.contiguous()
        if tg.dtype != torch.int64:
            tg = tg.long()
        need_grad = torch.is_grad_enabled()
        self._head_loss(ws, tg, need_grad)
        logits = LazyLogits(self, ws, ws.generation)
        if need_grad and not self.autograd_param_grads and _called_from_torch_ddp():
            # torch's DistributedDataParallel wraps this model (mingpt/trainer.py:71): its reducer only sees
            # gradients that arrive through autograd, so switch to that path (DataParallel is the fast one)
            self.autograd_param_grads = True
            self.mark_params_dirty()           # DDP's construction-time broadcast wrote the parameters in place
            self._ensure_device_state()
            if not self._warned_ddp:
                import warnings
                warnings.warn("GPT is wrapped in torch.nn.parallel.DistributedDataParallel: gradients now go through "
                              "autograd so DDP's hooks fire. mingpt_distributed_b200.DataParallel (what GPTTrainer "
                              "uses) overlaps the all-reduce without the extra copies.")
                self._warned_ddp = True
        if need_grad and self.autograd_param_grads:
            loss = _LossFnParams.apply(self._anchor, self, ws, *self.parameters())
        elif need_grad:
            loss = _LossFn.apply(self._anchor, self, ws)
        else:
            loss = ws.loss.clone()
        return logits, loss

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor,
                         grad_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """forward(inputs, targets) immediately followed by loss.backward(), without going through autograd:
        the fast path of one training step (mingpt/trainer.py:122-127). Gradients land in p.grad exactly as
        backward() would leave them. Performs no allocation and no host sync, so it can be captured in a CUDA
        graph. Returns the loss as a device scalar that is OVERWRITTEN by the next call."""
        B, T = inputs.shape
        assert T <= self.block_size, "Sequence length must be less than or equal to block size for positional embeddings."
        self._ensure_device_state()
        if inputs.dtype != torch.int64 or targets.dtype != torch.int64 or not inputs.is_contiguous() or not targets.is_contiguous():
            raise RuntimeError("forward_backward expects contiguous int64 CUDA tensors")
        ws = self._ws(B, T)
        ws.generation += 1
        ws.idx = inputs
        self._trunk(ws, inputs)
        direct = grad_scale is None
        if direct:
            self._prepare_grads()        # the head's wgrad below already accumulates into the gradient buffer
        self._head_loss(ws, targets, True, direct=direct)
        self._backward(ws, self._one if grad_scale is None else grad_scale, publish=not direct)
        return ws.loss

    # ------------------------------------------------------------------ backward
    def _prepare_grads(self):
        """Gradient accumulation semantics of .grad: tensors whose .grad is None (fresh model, or
        optimizer.zero_grad(set_to_none=True) as in mingpt/trainer.py:125) start from zero."""
        params = list(self.named_parameters())
        missing = [n for n, p in params if p.grad is None]
        if len(missing) == len(params):
            self._flat_grad.zero_()
        else:
            for n in missing:
                o, k = self._spans[n]
                self._flat_grad[o:o + k].zero_()
        for n, p in params:
            if p.grad is None:
                o, k = self._spans[n]
                p.grad = self._flat_grad[o:o + k].view(p.shape)

    def _backward_returning_grads(self, ws: _Workspace, grad_loss: torch.Tensor) -> List[torch.Tensor]:
        """backward for the autograd-visible path: computes THIS backward's gradients (not accumulated onto
        p.grad -- autograd's AccumulateGrad does that) and returns one tensor per parameter."""
        params = list(self.named_parameters())
        fresh = all(p.grad is None for _, p in params)
        # gradients already live in the flat buffer (accumulation, e.g. DDP.no_sync): write this pass elsewhere
        target = self._flat_grad if fresh else torch.zeros_like(self._flat_grad)
        if fresh:
            target.zero_()
        self._grad_target = target
        try:
            self._backward(ws, grad_loss, publish=False)
        finally:
            self._grad_target = None
        return [target[o:o + k].view(p.shape) for (o, k), p in ((self._spans[n], p) for n, p in params)]

    def _backward(self, ws: _Workspace, grad_loss: torch.Tensor, publish: bool = True):
        if not ws.has_head_grads:
            raise RuntimeError("backward() needs a forward that ran with gradients enabled")
        B, T, M = ws.B, ws.T, ws.M
        C, H, d, L = self.n_embed, self.n_head, self.head_dim, self.n_layer
        ref = self.variant == "reference"
        mask = ops.MASK_ADD_TRIL_ONES if ref else ops.MASK_CAUSAL
        W = lambda r: self._w(r, bf16=True)
        P = lambda r: self._w(r)
        G = self._g
        if publish:
            self._prepare_grads()
        if not getattr(ws, "head_direct", False):
            # upstream gradient of the scalar loss, applied on the device (no host sync)
            ops.scale_bf16_dev(ws.dlnf, grad_loss)
            ops.axpy(G("head_w"), ws.head_dw, grad_loss)
        ops.layernorm_bwd(ws.dlnf, ws.x[L], P("lnf_w"), ws.lnf_stats[0], ws.lnf_stats[1], ws.dx_a, G("lnf_w"), G("lnf_b"))
        if self._layer_done_hook is not None:
            self._layer_done_hook(L + 1)
        dx = ws.dx_a
        pe, pr, pa = ws.drop
        seed = ws.seed
        for l in range(L - 1, -1, -1):
            _nvtx.push(f"block{l}.bwd")
            s = ws.stats[l]
            dmlp = dx          # gradient of the MLP branch output: through its dropout mask when there is one
            if pr > 0 and ref:
                ops.dropout(dx, None, ws.dtmp, pr, seed, 4 * l + 2)
                dmlp = ws.dtmp
            elif pr > 0:       # the mask, and the c_proj bias gradient = column sums of the masked gradient, in one pass
                ops.dropout_bwd_colsum(dx.view(-1, C), ws.dtmp, pr, seed, 4 * l + 2, G(f"{l}.fc2_b"))
                dmlp = ws.dtmp
            if ref:
                ops.act_bwd(dmlp, ws.z[l], ws.dz, ops.GATE_MUL)
                ops.colsum(ws.dz, G(f"{l}.fc2_b"))
                ops.linear_wgrad(ws.dz, ws.fc_pre[l], G(f"{l}.fc2_w"))
                ops.linear_dgrad(ws.dz, W(f"{l}.fc2_w"), ws.dfc)
            else:
                # (the c_proj bias gradient: from the dropout backward above, or without dropout from the LayerNorm
                # backward below, which reads dx anyway)
                ops.linear_wgrad(dmlp, ws.fc_act[l], G(f"{l}.fc2_w"))
                # d(c_fc output) and, in the same epilogue, its column sums = the c_fc bias gradient
                ops.linear_dgrad(dmlp, W(f"{l}.fc2_w"), ws.dfc, gate=ws.fc_pre[l], gate_kind=ops.GATE_MUL, colsum=G(f"{l}.fc_b"))
            if ref:
                ops.colsum(ws.dfc, G(f"{l}.fc_b"))
            ops.linear_wgrad(ws.dfc, ws.ln2[l], G(f"{l}.fc_w"))
            ops.linear_dgrad(ws.dfc, W(f"{l}.fc_w"), ws.dln)
            ops.layernorm_bwd(ws.dln, ws.xmid[l], P(f"{l}.ln2_w"), s[2], s[3], ws.dx_b, G(f"{l}.ln2_w"), G(f"{l}.ln2_b"), dres=dx,
                              dres_colsum=G(f"{l}.fc2_b") if (pr == 0 and not ref) else None)
            dxm = ws.dx_b
            a_in = ws.att_o[l] if ref else ws.att[l]
            dproj = dxm
            if pr > 0:         # (without dropout the bias gradient is fused into the ln_1 backward below, whose dres is this tensor)
                ops.dropout_bwd_colsum(dxm.view(-1, C), ws.dtmp, pr, seed, 4 * l + 1, G(f"{l}.proj_b"))
                dproj = ws.dtmp
            ops.linear_wgrad(dproj, a_in, G(f"{l}.proj_w"))
            # flash attention's delta = rowsum(dO * O) comes out of the epilogue of the GEMM that produces dO (head size
            # 64 = one 64-column output slice per head; M > 128 = the CTA-pair engine) instead of a pass over O and dO
            fuse_delta = d == 64 and M > 128
            rd = (ws.att[l], ws.delta, T) if fuse_delta else None
            ops.linear_dgrad(dproj, W(f"{l}.proj_w"), ws.datt, rowdot=None if ref else rd)
            datt = ws.datt
            if ref:
                ops.colsum(ws.datt, G(f"{l}.o_b"))
                ops.linear_wgrad(ws.datt, ws.att[l], G(f"{l}.o_w"))
                ops.linear_dgrad(ws.datt, W(f"{l}.o_w"), ws.datt2, rowdot=rd)
                datt = ws.datt2
            if pa > 0 and d == 64:   # the forward's keep words are still in the workspace
                ops.attention_bwd_dropout_bits(ws.qkv[l], ws.att[l], datt, ws.lse[l], ws.delta, ws.dq_acc, ws.dqkv, B, T, H, d,
                                               mask, pa, ws.drop_bits[l], delta_given=fuse_delta, dqkv_colsum=G(f"{l}.qkv_b"))
            elif pa > 0:
                ops.attention_bwd_dropout(ws.qkv[l], ws.att[l], datt, ws.lse[l], ws.delta, ws.dq_acc, ws.dqkv, B, T, H, d,
                                          mask, pa, seed, 4 * l, delta_given=fuse_delta)
                ops.colsum(ws.dqkv, G(f"{l}.qkv_b"))
            else:      # the qkv bias gradient (column sums of dqkv) comes out of the attention backward itself
                ops.attention_bwd(ws.qkv[l], ws.att[l], datt, ws.lse[l], ws.delta, ws.dq_acc, ws.dqkv, B, T, H, d, mask,
                                  dqkv_colsum=G(f"{l}.qkv_b"), delta_given=fuse_delta)
            ops.linear_wgrad(ws.dqkv, ws.ln1[l], G(f"{l}.qkv_w"))
            ops.linear_dgrad(ws.dqkv, W(f"{l}.qkv_w"), ws.dln)
            ops.layernorm_bwd(ws.dln, ws.x[l], P(f"{l}.ln1_w"), s[0], s[1], ws.dx_a, G(f"{l}.ln1_w"), G(f"{l}.ln1_b"), dres=dxm,
                              dres_colsum=G(f"{l}.proj_b") if pr == 0 else None)
            dx = ws.dx_a
            _nvtx.pop()
            if self._layer_done_hook is not None:
                self._layer_done_hook(l + 1)
        gwpe = G("wpe").view(-1, C)
        if pe > 0:
            ops.dropout(dx, None, dx, pe, seed, self.SITE_EMBED)
        ops.embedding_bwd(ws.idx, dx, G("wte"), gwpe)
        if self._layer_done_hook is not None:
            self._layer_done_hook(0)
        ws.has_head_grads = False

    # ------------------------------------------------------------------ optimizer / generate
    def configure_optimizers(self, train_config):
        """README.md / upstream spelling of create_optimizer (mingpt/model.py:62-122)."""
        oc = OptimizerConfig(learning_rate=getattr(train_config, "learning_rate", 3e-4),
                             weight_decay=getattr(train_config, "weight_decay", 0.1),
                             betas=tuple(getattr(train_config, "betas", (0.9, 0.95))))
        return create_optimizer(self, oc)

    def _decode_step(self, st: _DecodeState, tok: torch.Tensor, pos: int) -> torch.Tensor:
        """One new token per sequence at position `pos` through the whole stack against the cache: every layer is
        M = B rows (LayerNorm, the single-CTA tcgen05 GEMM, the decode attention kernel). Returns fp32 logits [B, V]."""
        C, H, d, L, B = self.n_embed, self.n_head, self.head_dim, self.n_layer, st.B
        W = lambda r: self._w(r, bf16=True)
        P = lambda r: self._w(r)
        wpe = W("wpe").view(-1, C)
        ops.embedding_fwd(tok.view(B, 1), W("wte"), wpe[pos:pos + 1], st.x.view(B, 1, C), st.err_flag)
        for l in range(L):
            ops.layernorm_fwd(st.x, P(f"{l}.ln1_w"), P(f"{l}.ln1_b"), st.ln, st.mean, st.rstd)
            ops.linear_fwd(st.ln, W(f"{l}.qkv_w"), st.cache[l][:, pos, :], bias=P(f"{l}.qkv_b"))   # q | k | v of the new token
            ops.attention_decode(st.cache[l], st.att, B, st.cap, pos, H, d)
            ops.linear_fwd(st.att, W(f"{l}.proj_w"), st.xmid, bias=P(f"{l}.proj_b"), residual=st.x)
            ops.layernorm_fwd(st.xmid, P(f"{l}.ln2_w"), P(f"{l}.ln2_b"), st.ln, st.mean, st.rstd)
            ops.linear_fwd(st.ln, W(f"{l}.fc_w"), st.fc, bias=P(f"{l}.fc_b"), act=ops.ACT_GELU_TANH)
            ops.linear_fwd(st.fc, W(f"{l}.fc2_w"), st.x, bias=P(f"{l}.fc2_b"), residual=st.xmid)
        ops.layernorm_fwd(st.x, P("lnf_w"), P("lnf_b"), st.ln, st.mean, st.rstd)
        ops.linear_fwd(st.ln, W("head_w"), st.logits)
        return st.logits

    @torch.no_grad()
    def generate(self, idx, max_new_tokens, temperature=1.0, do_sample=False, top_k=None, use_cache=True):
        """mingpt/model.py:322-356. The reference re-runs the whole window for every new token (:334-339); here the
        prompt is run once (prefill: the training forward's kernels, keeping every layer's keys / values) and each
        further token is ONE row through the stack against that cache (`_decode_step`) -- as long as the sequence
        fits block_size. Beyond it the reference crops to the last block_size tokens (:336-337), which moves every
        token to a new learned position: the cache is then void by construction and each token is a full-window
        forward, as in the reference. `variant="reference"` never uses the cache: its attention is not causal
        (SURVEY Q1), so a new token changes every earlier position's output. Only the last position goes through
        the head GEMM either way."""
        was_training = self.training
        if any(p > 0 for p in self.drop_p):
            self.eval()
        cached = use_cache and self.variant == "mingpt"
        state: Optional[_DecodeState] = None

        def pick(logits):
            logits = logits / temperature
            if top_k is not None:
                v, _ = torch.topk(logits, top_k)
                logits[logits < v[:, [-1]]] = -float("Inf")
            probs = torch.softmax(logits, dim=-1)
            if do_sample:
                return torch.multinomial(probs, num_samples=1)
            return torch.topk(probs, k=1, dim=-1)[1]

        try:
            for _ in range(max_new_tokens):
                self._ensure_device_state()
                T_all = idx.size(1)
                if cached and state is not None and T_all <= self.block_size:
                    logits = self._decode_step(state, idx[:, -1].contiguous().long(), T_all - 1).clone()
                else:
                    idx_cond = idx if T_all <= self.block_size else idx[:, -self.block_size:]
                    B, T = idx_cond.shape
                    ws = self._ws(B, T)
                    ws.generation += 1
                    self._trunk(ws, idx_cond.contiguous().long())
                    logits = self._head_logits(ws, last_only=True)
                    state = None
                    if cached and T_all < self.block_size:      # prefill: keep the keys / values of every layer
                        state = _DecodeState(self, B)
                        for l in range(self.n_layer):
                            state.cache[l][:, :T, :].copy_(ws.qkv[l].view(B, T, -1))
                idx = torch.cat((idx, pick(logits)), dim=1)
        finally:
            self.train(was_training)
        return idx

    # ------------------------------------------------------------------ snapshots
    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        """Reference snapshots carry one `blocks.N.attn.mask` buffer per layer (mingpt/model.py:142-145,
        saved by mingpt/trainer.py:155); they are implicit here and dropped on load."""
        sd = {k: v for k, v in state_dict.items() if not k.endswith(".attn.mask")}
        out = super().load_state_dict(sd, strict=strict, assign=False)
        self._shadow_token = None
        return out


def param_groups(model: nn.Module) -> Tuple[List[str], List[str]]:
    """Names in the decay / no-decay groups, by the rules of mingpt/model.py:68-104."""
    decay, no_decay = set(), set()
    for mn, m in model.named_modules():
        for pn, _ in m.named_parameters():
            fpn = f"{mn}.{pn}" if mn else pn
            if pn.endswith("bias"):
                no_decay.add(fpn)
            elif pn.endswith("weight") and isinstance(m, nn.Linear):
                decay.add(fpn)
            elif pn.endswith("in_proj_weight"):
                decay.add(fpn)
            elif pn.endswith("weight") and isinstance(m, (nn.LayerNorm, nn.Embedding)):
                no_decay.add(fpn)
            elif pn.endswith("pos_embedding"):
                no_decay.add(fpn)
    names = {n for n, _ in model.named_parameters()}
    assert not (decay & no_decay), f"parameters {decay & no_decay} made it into both decay/no_decay sets!"
    assert not (names - (decay | no_decay)), f"parameters {names - (decay | no_decay)} were not separated into either decay/no_decay set!"
    return sorted(decay), sorted(no_decay)


def create_optimizer(model: nn.Module, optimizer_config: OptimizerConfig):
    """mingpt/model.py:62-122: two groups (decay / no decay), AdamW(lr, betas). Returns the fused
    multi-tensor AdamW (same param_groups / state_dict layout as torch.optim.AdamW)."""
    from .optim import FusedAdamW
    decay, no_decay = param_groups(model)
    pd = dict(model.named_parameters())
    groups = [
        {"params": [pd[n] for n in decay], "weight_decay": optimizer_config.weight_decay},
        {"params": [pd[n] for n in no_decay], "weight_decay": 0.0},
    ]
    return FusedAdamW(groups, lr=optimizer_config.learning_rate, betas=tuple(optimizer_config.betas), model=model)
