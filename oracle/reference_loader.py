"""ORACLE -- test infrastructure only. Imports the UNMODIFIED reference model file from /root/reference
and makes its forward runnable by applying the three crash repairs of SURVEY.md (D1-D3) in memory:
  D1  GPT.forward reads self.embeddings, the attribute is self.embedding          (mingpt/model.py:310 vs :243)
  D2  GPTEmbedding.forward reads self.block_size, the attribute is self.block_sz  (:225 vs :217)
  D3  self.pos_embedding(x) calls an nn.Parameter; intended pos_embedding[:, :T]  (:228)
Nothing is copied from the reference and nothing is written to it. Only usable where /root/reference is
mounted (the build container); the GPU box uses the fixtures under tests/golden/ instead."""
from __future__ import annotations

import contextlib
import importlib.util
import io
import os

REFERENCE_ROOT = os.environ.get("MINGPT_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.exists(os.path.join(REFERENCE_ROOT, "mingpt", "model.py"))


def load_reference_module():
    spec = importlib.util.spec_from_file_location("_reference_mingpt_model", os.path.join(REFERENCE_ROOT, "mingpt", "model.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    def embedding_forward(self, x):  # D2 + D3, following mingpt/model.py:222-231
        batch, seq_len = x.size()
        assert seq_len <= self.block_sz
        return self.drop(self.token_embedding(x) + self.pos_embedding[:, :seq_len, :])

    mod.GPTEmbedding.forward = embedding_forward
    mod.GPT.embeddings = property(lambda self: self.embedding)  # D1
    return mod


def build_reference_gpt(n_layer, n_head, n_embed, vocab_size, block_size, seed=0, dropout=0.0):
    """Constructs the reference's GPT with explicit sizes and model_type=None (avoids D4/D5)."""
    import torch
    mod = load_reference_module()
    cfg = mod.GPTConfig(model_type=None, n_layer=n_layer, n_head=n_head, n_embed=n_embed, vocab_size=vocab_size,
                        block_size=block_size, embed_drop=dropout, resid_drop=dropout, attn_drop=dropout)
    torch.manual_seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):
        model = mod.GPT(cfg)
    return mod, model
