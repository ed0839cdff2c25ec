"""Entry point mirroring mingpt/train.py:11-61 (hydra replaced by config.load_config):

    torchrun --standalone --nproc-per-node N -m mingpt_distributed_b200.train \
        --config mingpt_distributed_b200/gpt2_config.yaml data_config.path=corpus.txt
"""
from __future__ import annotations

import argparse
import os

import torch
from torch.distributed import destroy_process_group, init_process_group
from torch.utils.data import random_split

from .char_dataset import CharDataset, DataConfig
from .config import load_config
from .model import GPT, GPTConfig, OptimizerConfig, create_optimizer
from .trainer import GPTTrainer


def get_resources(gpt_config: GPTConfig, optimizer_config: OptimizerConfig, data_config: DataConfig):
    """mingpt/train.py:11-27"""
    data = CharDataset(data_config)
    train_size = int(len(data) * data_config.train_split)
    train_split, test_split = random_split(data, [train_size, len(data) - train_size])
    gpt_config.vocab_size = data.vocab_size
    gpt_config.block_size = data.block_size
    model = GPT(gpt_config)
    optimizer = create_optimizer(model, optimizer_config)
    return model, optimizer, train_split, test_split


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default=os.path.join(os.path.dirname(__file__), "gpt2_config.yaml"))
    ap.add_argument("overrides", nargs="*")
    args = ap.parse_args(argv)
    distributed = "RANK" in os.environ and "WORLD_SIZE" in os.environ
    if distributed:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        init_process_group(backend="nccl")
    gpt_config, optimizer_config, data_config, trainer_config = load_config(args.config, args.overrides)
    model, optimizer, train, test = get_resources(gpt_config, optimizer_config, data_config)
    trainer = GPTTrainer(trainer_config, model, optimizer, train, test)
    trainer.train()
    if distributed:
        destroy_process_group()


if __name__ == "__main__":
    main()
