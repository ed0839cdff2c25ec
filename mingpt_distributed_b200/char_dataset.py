"""Byte-level language-model dataset with the reference's interface (mingpt/char_dataset.py:12-47).

Same semantics as the reference: the file is read as BYTES (fsspec.open defaults to 'rb'), the
vocabulary is the sorted set of byte values, item i is (codes[i:i+block], codes[i+1:i+block+1]) as int64.
The per-item Python list comprehension of the reference (char_dataset.py:38-47) is replaced by one
vectorised encode of the whole corpus at construction; __getitem__ is then two tensor slices.

DeviceWindowLoader is the training-time input pipeline for it: the encoded corpus is put on the GPU once (one byte
per token) and every batch is cut out of it by one gather kernel (mgb_token_windows) from the window starts the
sampler dictates -- no per-item Python, no collation, no per-batch host->device copy. The order of the windows
is torch's DistributedSampler's, i.e. exactly what the reference's DataLoader would deliver
(mingpt/trainer.py:73-81).
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

    def device_codes(self, device) -> torch.Tensor:
        """the encoded corpus as a uint8 tensor on `device` (uploaded once, cached)"""
        device = torch.device(device)
        cache = self.__dict__.setdefault("_device_codes", {})
        if device not in cache:
            cache[device] = self.codes.to(torch.uint8).to(device)
        return cache[device]


def _unwrap_subset(dataset):
    """(CharDataset, index map or None) for a CharDataset or (nested) torch Subsets of one (random_split,
    mingpt/train.py:20-22); (None, None) for anything else"""
    from torch.utils.data import Subset
    index_map = None
    while isinstance(dataset, Subset):
        idx = torch.as_tensor(dataset.indices, dtype=torch.int64)
        index_map = idx if index_map is None else idx[index_map]
        dataset = dataset.dataset
    if isinstance(dataset, CharDataset):
        return dataset, index_map
    return None, None


class DeviceWindowLoader:
    """Iterates over (x, y) int64 CUDA batches of a CharDataset (or a random_split part of one) in the order
    `sampler` gives -- the drop-in for `DataLoader(dataset, batch_size, sampler=DistributedSampler(dataset))`
    (mingpt/trainer.py:73-81). Per epoch: one small host->device copy of the window starts; per batch: one kernel."""

    def __init__(self, dataset, batch_size: int, device, sampler):
        base, index_map = _unwrap_subset(dataset)
        if base is None:
            raise TypeError("DeviceWindowLoader needs a CharDataset (or torch Subsets of one)")
        self.dataset, self.base, self.index_map = dataset, base, index_map
        self.batch_size, self.device, self.sampler = batch_size, torch.device(device), sampler
        self.block_size = base.block_size
        self._err = None

    def __len__(self):
        n = len(self.sampler)
        return (n + self.batch_size - 1) // self.batch_size

    def window_starts(self) -> torch.Tensor:
        """corpus offsets of this epoch's windows for this rank, in sampler order (CPU int64)"""
        idx = torch.tensor(list(iter(self.sampler)), dtype=torch.int64)
        return idx if self.index_map is None else self.index_map[idx]

    def __iter__(self):
        from . import ops
        codes = self.base.device_codes(self.device)
        starts = self.window_starts().to(self.device, non_blocking=True)
        if self._err is None:
            self._err = torch.zeros(1, dtype=torch.int32, device=self.device)
        T, B = self.block_size, self.batch_size
        for i in range(0, starts.numel(), B):
            s = starts[i:i + B]
            x = torch.empty(s.numel(), T, dtype=torch.int64, device=self.device)
            y = torch.empty_like(x)
            ops.token_windows(codes, s, T, x, y, self._err)
            yield x, y
