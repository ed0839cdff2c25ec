"""Byte-level language-model dataset with the reference's interface (mingpt/char_dataset.py:12-47).

Same semantics as the reference: the file is read as BYTES (fsspec.open defaults to 'rb'), the
vocabulary is the sorted set of byte values, item i is (codes[i:i+block], codes[i+1:i+block+1]) as int64.
The per-item Python list comprehension of the reference (char_dataset.py:38-47) is replaced by one
vectorised encode of the whole corpus at construction; __getitem__ is then two tensor slices.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
from torch.utils.data import Dataset


@dataclass
class DataConfig:
    path: Optional[str] = None
    block_size: Optional[int] = None
    train_split: Optional[float] = None
    truncate: float = 1.0


def _read_bytes(path: str) -> bytes:
    if "://" in path:
        import fsspec  # remote URLs only; local files need nothing beyond open()
        with fsspec.open(path) as f:
            return f.read()
    with open(path, "rb") as f:
        return f.read()


class CharDataset(Dataset):
    def __init__(self, config: DataConfig, data: Optional[bytes] = None):
        if data is None:
            data = _read_bytes(config.path)
        if isinstance(data, str):
            data = data.encode("utf-8")
        data = data[: int(len(data) * config.truncate)]
        chars = sorted(set(data))
        print(f"Data has {len(data)} chars and {len(chars)} unique characters")
        self.stoi = {ch: i for i, ch in enumerate(chars)}
        self.itos = {i: ch for i, ch in enumerate(chars)}
        self.block_size = config.block_size
        self.vocab_size = len(chars)
        self.data = data
        lut = torch.zeros(256, dtype=torch.int64)
        for ch, i in self.stoi.items():
            lut[ch] = i
        raw = torch.frombuffer(bytearray(data), dtype=torch.uint8).long() if len(data) else torch.zeros(0, dtype=torch.int64)
        self.codes = lut[raw]

    def __len__(self):
        return len(self.data) - self.block_size

    def __getitem__(self, idx):
        if idx < 0:
            idx += len(self)
        block = self.codes[idx: idx + self.block_size + 1]
        return block[:-1].clone(), block[1:].clone()

    def decode(self, indices) -> bytes:
        return bytes(self.itos[int(i)] for i in indices)
