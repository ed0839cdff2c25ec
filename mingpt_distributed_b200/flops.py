"""Algorithmic training flops per token and MFU -- what bench.py reports and what the trainer
prints next to tokens/s (the reference prints the loss only: mingpt/trainer.py:144-147).

    F_tok = 6 * N_mm + attention term
    mingpt    graph: N_mm = 12 L C^2 + V C, causal attention      6 L C T
    reference graph: N_mm = 13 L C^2 + V C (two output projections), full-square attention 12 L C T
Embeddings, LayerNorm, biases and the recompute inside the attention / loss backward are not counted (MFU, not HFU).
"""
from __future__ import annotations

NOMINAL_BF16_FLOPS = 2.25e15      # B200 dense bf16, per GPU


def flops_per_token(n_layer: int, n_embed: int, vocab: int, T: int, variant: str = "mingpt") -> int:
    if variant == "reference":
        return 6 * (13 * n_layer * n_embed ** 2 + vocab * n_embed) + 12 * n_layer * n_embed * T
    return 6 * (12 * n_layer * n_embed ** 2 + vocab * n_embed) + 6 * n_layer * n_embed * T


def model_flops_per_token(model, T: int) -> int:
    return flops_per_token(model.n_layer, model.n_embed, model.vocab_size, T, model.variant)


def mfu(tokens_per_s: float, f_tok: float, n_gpus: int = 1, peak_flops: float = NOMINAL_BF16_FLOPS) -> float:
    return tokens_per_s * f_tok / (n_gpus * peak_flops)
