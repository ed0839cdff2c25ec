"""Reads the reference's run configuration (mingpt/gpt2_config.yaml:1-23) without hydra/omegaconf
(neither is a dependency here): four sections -> four dataclasses, `hydra:` section ignored, hydra-style
`section.key=value` overrides supported. Accepts both `n_embd` (what the YAML says) and `n_embed` (what
the reference's dataclass calls it; the mismatch crashes the reference)."""
from __future__ import annotations

from dataclasses import fields
from typing import Iterable, Optional, Tuple

import yaml

from .char_dataset import DataConfig
from .model import GPTConfig, OptimizerConfig
from .trainer import GPTTrainerConfig

SECTIONS = {"gpt_config": GPTConfig, "optimizer_config": OptimizerConfig, "data_config": DataConfig,
            "trainer_config": GPTTrainerConfig}


def _coerce(text: str):
    return yaml.safe_load(text)


def load_config(path: str, overrides: Optional[Iterable[str]] = None) -> Tuple[GPTConfig, OptimizerConfig, DataConfig, GPTTrainerConfig]:
    with open(path) as f:
        raw = yaml.safe_load(f) or {}
    for ov in overrides or ():
        key, _, val = ov.partition("=")
        sec, _, name = key.partition(".")
        raw.setdefault(sec, {})[name] = _coerce(val)
    out = []
    for sec, cls in SECTIONS.items():
        d = dict(raw.get(sec) or {})
        known = {f.name for f in fields(cls)}
        unknown = set(d) - known
        if unknown:
            raise TypeError(f"{sec}: unexpected keys {sorted(unknown)} (known: {sorted(known)})")
        if "betas" in d:
            d["betas"] = tuple(d["betas"])
        out.append(cls(**d))
    return tuple(out)
