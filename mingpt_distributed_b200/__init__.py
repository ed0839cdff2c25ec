"""mingpt_distributed_b200: the training step of aponte411/minGPT-distributed, rebuilt for B200 (sm_100a).

Drop-in names (reference spelling and README / upstream spelling):
    GPT, GPTConfig, OptimizerConfig, create_optimizer, print_model_size        (mingpt/model.py)
    GPTTrainer, GPTTrainerConfig, ModelSnapshot, Trainer                       (mingpt/trainer.py)
    CharDataset, DataConfig                                                    (mingpt/char_dataset.py)
    load_config                                                                (mingpt/gpt2_config.yaml)
The compute path is CUDA-only (libmingpt_b200.so, built in-tree by csrc/build.sh); importing the package
and building models / optimizers / datasets works anywhere.
"""
from .model import (GPT, CfgNode, GPTConfig, LazyLogits, OptimizerConfig, create_optimizer, param_groups,
                    print_model_size)
from .char_dataset import CharDataset, DataConfig, DeviceWindowLoader
from .optim import FusedAdamW, clip_grad_norm_
from .parallel import DataParallel, GradBuckets, plan_buckets
from .trainer import GPTTrainer, GPTTrainerConfig, GraphedTrainStep, ModelSnapshot, Trainer, train_step
from .config import load_config
from .flops import flops_per_token, mfu, model_flops_per_token

__all__ = ["GPT", "CfgNode", "GPTConfig", "LazyLogits", "OptimizerConfig", "create_optimizer", "param_groups",
           "print_model_size", "CharDataset", "DataConfig", "DeviceWindowLoader", "FusedAdamW", "clip_grad_norm_", "DataParallel",
           "GradBuckets", "plan_buckets", "GPTTrainer", "GraphedTrainStep", "GPTTrainerConfig", "ModelSnapshot", "Trainer",
           "train_step", "load_config", "flops_per_token", "model_flops_per_token", "mfu"]
