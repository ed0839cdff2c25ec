"""ORACLE -- test infrastructure only (tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this; the product package must never do so).

CPU restatement, in plain fp32 PyTorch, of the reference's training step:
  * forward of the GPT the reference really builds (variant "reference"; /root/reference/mingpt/model.py,
    cited per function) and of the graph its README / the task's north_star describe (variant "mingpt";
    the reference has no code for that one: README.md:29-54 documents the API only);
  * the optimizer grouping of create_optimizer (mingpt/model.py:62-122);
  * one step of GPTTrainer._run_batch (mingpt/trainer.py:118-133).

PINNING.  The reference ships no tests, golden vectors or expected numbers, so the pins are
minted from the reference ITSELF: oracle/reference_loader.py imports /root/reference/mingpt/model.py with
the three crash repairs D1-D3 applied in memory, oracle/make_golden.py runs it on seeded inputs and
commits the results under tests/golden/, and tests/test_oracle.py checks this restatement against both
(the live import where /root/reference exists, the committed fixtures everywhere).  The "mingpt" variant
has no reference code to pin against: parity for it is "unpinned" by the reference and rests on this
restatement alone (it differs from the pinned variant only in mask, projection count and MLP order).

The arithmetic of both lives in torch (un-vendored, unpinned by the reference: requirements.txt:1); the
fixtures record the torch version that produced them.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

StateDict = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------------------------------
# forward
# ----------------------------------------------------------------------------------------------------
def _layer_norm(x, sd, prefix):
    """nn.LayerNorm(n_embed), eps 1e-5 (mingpt/model.py:176,178,248)"""
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".weight"], sd[prefix + ".bias"], 1e-5)


def _drop(x, keep, p):
    """nn.Dropout with a GIVEN keep mask (the CUDA path exports its counter-based masks for the tests)"""
    return x if keep is None else x * keep.to(x.dtype) / (1.0 - p)


def _attention_core(q, k, v, mode: str, keep=None, p: float = 0.0):
    """q,k,v (B,H,T,d).  mode 'add_tril_ones': what nn.MultiheadAttention does with the FLOAT tril mask the
    reference passes (mingpt/model.py:142-145,164 -> torch/nn/functional.py:6630-6647): q is scaled, the
    mask is ADDED to the scores, softmax, P@V -- nothing is masked.  mode 'causal': -inf above the diagonal."""
    T, d = q.shape[-2], q.shape[-1]
    att = (q * (1.0 / math.sqrt(d))) @ k.transpose(-2, -1)
    tril = torch.tril(torch.ones(T, T, dtype=att.dtype, device=att.device))
    if mode == "add_tril_ones":
        att = att + tril
    elif mode == "causal":
        att = att.masked_fill(tril == 0, float("-inf"))
    att = torch.softmax(att, dim=-1)
    att = _drop(att, keep, p)          # dropout on the probabilities (torch/nn/functional.py:6644-6645)
    return att @ v


def _split_heads(t, n_head):
    B, T, C = t.shape
    return t.view(B, T, n_head, C // n_head).transpose(1, 2)


def forward_reference(sd: StateDict, n_head: int, idx: torch.Tensor, targets: Optional[torch.Tensor] = None,
                      masks: Optional[dict] = None, drop=(0.0, 0.0, 0.0)):
    """GPT.forward of the repaired reference (mingpt/model.py:309-320). Dropout (embed, resid, attn = `drop`)
    is applied with the keep masks given in `masks` ({"embed", ("attn", l), ("resid1", l), ("resid2", l)}).
    Keys are the reference's state_dict names."""
    B, T = idx.shape
    masks = masks or {}
    pe, pr, pa = drop
    # GPTEmbedding.forward, mingpt/model.py:222-231 (with repairs D2/D3: pos_embedding[:, :T, :])
    x = sd["embedding.token_embedding.weight"][idx] + sd["embedding.pos_embedding"][:, :T, :]
    x = _drop(x, masks.get("embed"), pe)
    n_layer = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))
    for i in range(n_layer):
        p = f"blocks.{i}."
        # Block.forward, mingpt/model.py:186-189; MultiHeadSelfAttention.forward :156-168
        h = _layer_norm(x, sd, p + "ln_1")
        qkv = F.linear(h, sd[p + "attn.attn.in_proj_weight"], sd[p + "attn.attn.in_proj_bias"])
        q, k, v = (_split_heads(t, n_head) for t in qkv.chunk(3, dim=-1))
        y = _attention_core(q, k, v, "add_tril_ones", masks.get(("attn", i)), pa).transpose(1, 2).reshape(B, T, -1)
        y = F.linear(y, sd[p + "attn.attn.out_proj.weight"], sd[p + "attn.attn.out_proj.bias"])
        y = F.linear(y, sd[p + "attn.c_proj.weight"], sd[p + "attn.c_proj.bias"])
        x = x + _drop(y, masks.get(("resid1", i)), pr)
        # mlp = Linear -> Linear -> GELU(erf) -> Dropout, mingpt/model.py:179-184
        h = _layer_norm(x, sd, p + "ln_2")
        m = F.linear(h, sd[p + "mlp.0.weight"], sd[p + "mlp.0.bias"])
        m = F.linear(m, sd[p + "mlp.1.weight"], sd[p + "mlp.1.bias"])
        x = x + _drop(F.gelu(m), masks.get(("resid2", i)), pr)
    logits = F.linear(_layer_norm(x, sd, "ln_ff"), sd["head.weight"])
    loss = None
    if targets is not None:
        loss = F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)
    return logits, loss


def forward_mingpt(sd: StateDict, n_head: int, idx: torch.Tensor, targets: Optional[torch.Tensor] = None,
                   masks: Optional[dict] = None, drop=(0.0, 0.0, 0.0)):
    """The north_star graph: causal attention, c_attn/c_proj, c_fc -> NewGELU(tanh) -> c_proj, lm_head.
    Keys are upstream minGPT's state_dict names. Dropout as in forward_reference."""
    B, T = idx.shape
    masks = masks or {}
    pe, pr, pa = drop
    x = sd["transformer.wte.weight"][idx] + sd["transformer.wpe.weight"][:T]
    x = _drop(x, masks.get("embed"), pe)
    n_layer = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("transformer.h."))
    for i in range(n_layer):
        p = f"transformer.h.{i}."
        h = _layer_norm(x, sd, p + "ln_1")
        qkv = F.linear(h, sd[p + "attn.c_attn.weight"], sd[p + "attn.c_attn.bias"])
        q, k, v = (_split_heads(t, n_head) for t in qkv.chunk(3, dim=-1))
        y = _attention_core(q, k, v, "causal", masks.get(("attn", i)), pa).transpose(1, 2).reshape(B, T, -1)
        x = x + _drop(F.linear(y, sd[p + "attn.c_proj.weight"], sd[p + "attn.c_proj.bias"]), masks.get(("resid1", i)), pr)
        h = _layer_norm(x, sd, p + "ln_2")
        m = F.gelu(F.linear(h, sd[p + "mlp.c_fc.weight"], sd[p + "mlp.c_fc.bias"]), approximate="tanh")
        x = x + _drop(F.linear(m, sd[p + "mlp.c_proj.weight"], sd[p + "mlp.c_proj.bias"]), masks.get(("resid2", i)), pr)
    logits = F.linear(_layer_norm(x, sd, "transformer.ln_f"), sd["lm_head.weight"])
    loss = None
    if targets is not None:
        loss = F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)
    return logits, loss


def forward(sd: StateDict, n_head: int, idx, targets=None, variant: Optional[str] = None, masks=None,
            drop=(0.0, 0.0, 0.0)):
    variant = variant or ("reference" if "head.weight" in sd else "mingpt")
    fn = forward_reference if variant == "reference" else forward_mingpt
    return fn(sd, n_head, idx, targets, masks, drop)


# ----------------------------------------------------------------------------------------------------
# optimizer grouping and one training step
# ----------------------------------------------------------------------------------------------------
def decay_split(names: List[str]) -> Tuple[List[str], List[str]]:
    """create_optimizer's sets (mingpt/model.py:68-104) restated on parameter NAMES: every bias, LayerNorm
    and embedding weight and the positional table are not decayed; nn.Linear weights and MHA's
    in_proj_weight are."""
    decay, no_decay = [], []
    for n in names:
        leaf = n.rsplit(".", 1)[-1]
        is_norm = ".ln_1." in n or ".ln_2." in n or n.startswith("ln_ff.") or ".ln_f." in n
        is_emb = "token_embedding" in n or "pos_embedding" in n or ".wte." in n or ".wpe." in n
        if leaf.endswith("bias") or is_norm or is_emb:
            no_decay.append(n)
        else:
            decay.append(n)
    return sorted(decay), sorted(no_decay)


class OracleTrainer:
    """GPTTrainer._run_batch (mingpt/trainer.py:118-133) on CPU: forward, zero_grad(set_to_none), backward,
    clip_grad_norm (coef = c / (norm + 1e-6), clamped to 1), AdamW.step (lr 3e-4, betas (0.9, 0.95),
    eps 1e-8, weight_decay 0.1 on the decay group: mingpt/model.py:54-59,107-121)."""

    def __init__(self, sd: StateDict, n_head: int, variant: Optional[str] = None, lr=3e-4, weight_decay=0.1,
                 betas=(0.9, 0.95), grad_norm_clip: Optional[float] = 1.0):
        self.sd = {k: v.detach().clone().float().requires_grad_(True) for k, v in sd.items() if not k.endswith(".attn.mask")}
        self.n_head, self.variant, self.clip = n_head, variant, grad_norm_clip
        decay, no_decay = decay_split(list(self.sd))
        self.opt = torch.optim.AdamW(
            [{"params": [self.sd[n] for n in decay], "weight_decay": weight_decay},
             {"params": [self.sd[n] for n in no_decay], "weight_decay": 0.0}], lr=lr, betas=betas)

    def run_batch(self, idx, targets) -> float:
        _, loss = forward(self.sd, self.n_head, idx, targets, self.variant)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        if self.clip is not None:
            torch.nn.utils.clip_grad_norm_(list(self.sd.values()), self.clip)
        self.opt.step()
        return loss.item()

    def grads(self) -> StateDict:
        return {k: v.grad.detach().clone() for k, v in self.sd.items() if v.grad is not None}


def flops_per_token(n_layer: int, n_embed: int, vocab: int, T: int, variant: str = "mingpt") -> int:
    """ALGORITHMIC training flops per token: 6*N_mm + attention term.
    mingpt: N_mm = 12 L C^2 + V C, causal attention 6 L C T.  reference: N_mm = 13 L C^2 + V C and the
    full-square 12 L C T (its attention is not causal)."""
    if variant == "reference":
        return 6 * (13 * n_layer * n_embed ** 2 + vocab * n_embed) + 12 * n_layer * n_embed * T
    return 6 * (12 * n_layer * n_embed ** 2 + vocab * n_embed) + 6 * n_layer * n_embed * T
