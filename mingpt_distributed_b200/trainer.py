"""Training loop with the reference's interface (mingpt/trainer.py) on the fused B200 step.

GPTTrainer(config, model, optimizer, train_dataset, test_dataset=None).train() -- mingpt/trainer.py:40-183
Trainer(config, model, train_dataset).run()                                   -- README.md:47-53 spelling

One step (mingpt/trainer.py:118-133) = forward -> zero_grad(set_to_none=True) -> backward -> clip ->
optimizer.step -> loss.  Kept: that order and its numerics.  Changed for B200: the clip's norm reduction
and coefficient are fused into the AdamW launch, gradients are all-reduced per bucket on a side stream
while backward runs (parallel.DataParallel instead of torch DDP), the two H2D copies are non-blocking from
pinned memory, and the loss is only read back (a device->host sync) when it is printed, not every batch.
"""
from __future__ import annotations

import os
import time
import warnings
from collections import defaultdict
from dataclasses import asdict, dataclass
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader, Dataset
from torch.utils.data.distributed import DistributedSampler

from .char_dataset import DeviceWindowLoader, _unwrap_subset
from .flops import mfu, model_flops_per_token
from .model import CfgNode
from .optim import FusedAdamW, clip_grad_norm_
from .parallel import DataParallel


@dataclass
class GPTTrainerConfig:
    """mingpt/trainer.py:21-29"""
    max_epochs: Optional[int] = None
    batch_size: Optional[int] = None
    learning_rate: Optional[float] = None   # unused by the reference as well (LR comes from OptimizerConfig)
    grad_norm_clip: Optional[float] = None
    dl_num_workers: Optional[int] = None
    snapshot_path: Optional[str] = None
    save_every: Optional[int] = None
    # additions (defaults keep the reference's behaviour):
    shuffle_each_epoch: bool = False       # True: call sampler.set_epoch(epoch); the reference never does, so every
                                           # epoch replays the same order (mingpt/trainer.py:80,135-147,174)
    device_loader: Optional[bool] = None   # None/True: CharDataset batches are cut on the GPU from a device-resident
                                           # corpus (DeviceWindowLoader); False: torch DataLoader as in the reference


@dataclass
class ModelSnapshot:
    """mingpt/trainer.py:33-37"""
    model_state: "Dict[str, torch.Tensor]"
    optimizer_state: Dict[str, Any]
    final_epoch: int


def train_step(model, optimizer, inputs, labels, grad_norm_clip: Optional[float], dp: Optional[DataParallel] = None,
               accumulate: bool = False, loss_div: float = 1.0, zero_grad: bool = False) -> torch.Tensor:
    """The hot path: mingpt/trainer.py:122-132 on device tensors. Returns the loss as a DEVICE scalar.
    `accumulate=True` runs forward/backward only (no all-reduce, no update): gradient accumulation.
    `zero_grad=True` drops the gradients between forward and backward, where the reference does (:125)."""
    fwd = dp if dp is not None else model
    logits, loss = fwd(inputs, labels)
    if zero_grad:
        optimizer.zero_grad(set_to_none=True)
    if accumulate and dp is not None:
        with dp.no_sync():
            (loss / loss_div if loss_div != 1.0 else loss).backward()
        return loss.detach()
    (loss / loss_div if loss_div != 1.0 else loss).backward()
    if accumulate:
        return loss.detach()
    scale = 1.0
    if dp is not None:
        dp.finish_backward()
        scale = dp.grad_scale
    if isinstance(optimizer, FusedAdamW):
        clip = grad_norm_clip is not None and grad_norm_clip > 0
        norm_sq = optimizer.grad_norm_sq() if clip else None
        optimizer.step(grad_scale=scale, gradnorm_sq=norm_sq, max_norm=float(grad_norm_clip) if clip else 0.0)
    else:
        if scale != 1.0:
            model._flat_grad.mul_(scale)
        if grad_norm_clip is not None and grad_norm_clip > 0:
            clip_grad_norm_(model, grad_norm_clip)
        optimizer.step()
    return loss.detach()


class GraphedTrainStep:
    """One training step (mingpt/trainer.py:118-133) as a replayable CUDA graph.

    The step launches ~320 kernels of 20-600 us each; issued one by one from Python their launch overhead
    shows up as ~15 % of idle GPU time on a B200. After `warmup` eager steps the whole step -- forward,
    backward, norm, clip + AdamW -- is captured once per (batch, seq) shape and replayed. Inputs are copied
    into static device buffers; the returned loss is a static device scalar (read it before the next call).
    With use_graph=False the same body runs eagerly.

    Data parallel: NCCL stays OUT of the graphs. The step is captured as a chain of graph SEGMENTS cut where a
    gradient bucket completes; on replay each bucket's all-reduce is launched eagerly on the side stream
    between two segments, so it overlaps the rest of the backward exactly like the eager path, and the last
    segment (norm, clip, AdamW) starts after the side stream has drained.

    Gradient accumulation (`accum` = k > 1, BASELINE config 5): a call takes k micro-batches stacked along dim 0
    ((k*B, T) tokens), runs k forward/backward passes that add into the same flat gradient buffer and ONE
    optimizer step; the 1/k of the mean is folded into the AdamW kernel's gradient scale (no extra pass), and under
    data parallelism only the LAST micro-step exchanges gradients (DDP.no_sync for the others). The returned loss
    is the mean over the k micro-batches."""

    def __init__(self, model, optimizer, grad_norm_clip: Optional[float] = None, dp: Optional[DataParallel] = None,
                 use_graph: bool = True, warmup: int = 2, accum: int = 1):
        if not isinstance(optimizer, FusedAdamW):
            raise TypeError("GraphedTrainStep needs the FusedAdamW returned by create_optimizer")
        if accum < 1:
            raise ValueError("accum must be >= 1")
        self.model, self.opt, self.clip, self.dp = model, optimizer, grad_norm_clip, dp
        self.use_graph, self.warmup, self.accum = use_graph, warmup, accum
        self._states = {}
        self.replays = 0          # graph replays so far (the trainers' tests assert the fast path really ran)

    def _optimizer_part(self, scale: float, graph_step: bool):
        clip = self.clip is not None and self.clip > 0
        norm_sq = self.opt.grad_norm_sq() if clip else None
        self.opt.step(grad_scale=scale, gradnorm_sq=norm_sq, max_norm=float(self.clip) if clip else 0.0,
                      graph_step=graph_step)

    def _micro_steps(self, st, last_hook=None):
        """the k forward/backward passes of one step; `last_hook` (data parallel) is installed for the last one only"""
        model, k = self.model, self.accum
        for p in model.parameters():          # optimizer.zero_grad(set_to_none=True), mingpt/trainer.py:125
            p.grad = None
        if k == 1:
            if last_hook is not None:
                model._layer_done_hook = last_hook
            return model.forward_backward(st["x"], st["y"])
        B = st["x"].shape[0] // k
        st["loss_acc"].zero_()
        saved = model._layer_done_hook
        for i in range(k):
            model._layer_done_hook = (last_hook if last_hook is not None else saved) if i == k - 1 else None
            loss = model.forward_backward(st["x"][i * B:(i + 1) * B], st["y"][i * B:(i + 1) * B])
            st["loss_acc"].add_(loss, alpha=1.0 / k)
        model._layer_done_hook = saved
        return st["loss_acc"]

    def _body(self, st, graph_step: bool):
        if self.dp is not None:
            self.dp._ensure()
        loss = self._micro_steps(st)
        scale = 1.0 / self.accum
        if self.dp is not None:
            self.dp.finish_backward()
            scale *= self.dp.grad_scale
        self._optimizer_part(scale, graph_step)
        return loss

    # ---- data parallel: a chain of graph segments with the all-reduces in between ----
    def _capture_segments(self, st):
        model, dp = self.model, self.dp
        dp._ensure()
        buckets = dp.buckets
        segs = []
        pool = torch.cuda.graph_pool_handle()
        cur = {"g": None}

        def begin():
            g = torch.cuda.CUDAGraph()
            # thread_local: NCCL's watchdog thread polls CUDA events; in the default (global) mode that invalidates
            # a capture in progress on this thread
            g.capture_begin(pool=pool, capture_error_mode="thread_local")
            cur["g"] = g

        def end():
            with warnings.catch_warnings():       # a cut right after another cut leaves an (harmless) empty graph
                warnings.simplefilter("ignore")
                cur["g"].capture_end()
            segs.append(("graph", cur["g"]))
            cur["g"] = None

        def hook(group: int):
            span = buckets._by_trigger.get(group)
            if span is None or not buckets.enabled:      # (world size 1 still cuts: its all-reduces are no-ops)
                return
            end()
            segs.append(("allreduce", span))
            buckets.reserve_sms(True)      # the segments that run beside an all-reduce are captured with smaller grids
            begin()

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        saved = model._layer_done_hook
        with torch.cuda.stream(side):
            begin()
            try:
                st["loss"] = self._micro_steps(st, last_hook=hook)
                end()
                segs.append(("drain", None))
                buckets.reserve_sms(False)
                begin()
                self._optimizer_part(dp.grad_scale / self.accum, True)
                end()
            finally:
                model._layer_done_hook = saved
                buckets.reserve_sms(False)
        torch.cuda.current_stream().wait_stream(side)
        return segs

    def _replay_segments(self, segs):
        buckets = self.dp.buckets
        cur = torch.cuda.current_stream()
        for kind, obj in segs:
            if kind == "graph":
                obj.replay()
            elif kind == "allreduce":
                buckets.reduce_span(*obj)
            else:
                cur.wait_stream(buckets.stream)

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        key = tuple(x.shape)
        st = self._states.get(key)
        if st is None:
            if key[0] % self.accum:
                raise ValueError(f"batch dimension {key[0]} is not a multiple of accum={self.accum}")
            dev = self.model._flat.device
            st = dict(x=torch.empty(key, dtype=torch.int64, device=dev), y=torch.empty(key, dtype=torch.int64, device=dev),
                      calls=0, graph=None, loss=None, launches=0,
                      loss_acc=torch.zeros((), dtype=torch.float32, device=dev))
            self._states[key] = st
        st["x"].copy_(x, non_blocking=True)
        st["y"].copy_(y, non_blocking=True)
        if not self.use_graph or st["calls"] < self.warmup:
            st["loss"] = self._body(st, graph_step=False)
        else:
            if st["graph"] is None:
                torch.cuda.synchronize()
                self.opt.begin_graph_mode()
                from . import _lib
                n0 = _lib.launch_count()
                if self.dp is not None:
                    st["graph"] = self._capture_segments(st)
                else:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        st["loss"] = self._body(st, graph_step=True)
                    st["graph"] = g
                st["launches"] = _lib.launch_count() - n0      # kernels of ours inside one replay
            if self.dp is not None:
                self._replay_segments(st["graph"])
            else:
                st["graph"].replay()
            self.opt.note_replay()
            self.replays += 1
        st["calls"] += 1
        return st["loss"]

    def launches_per_replay(self, shape) -> int:
        st = self._states.get(tuple(shape))
        return st["launches"] if st else 0


class GPTTrainer:
    """Same constructor and methods as the reference's GPTTrainer (mingpt/trainer.py:40-183)."""

    def __init__(self, config: GPTTrainerConfig, model: torch.nn.Module, optimizer: Any, train_dataset: Dataset,
                 test_dataset: Dataset = None):
        # torchrun / srun environment contract (mingpt/trainer.py:53-54, mingpt/slurm/slurm_run.sh:17-23)
        self.local_rank = int(os.environ.get("LOCAL_RANK", 0))
        self.global_rank = int(os.environ.get("RANK", 0))
        self.config = config
        self.train_dataset = train_dataset
        self.test_dataset = test_dataset
        self.device = torch.device("cuda", self.local_rank)
        torch.cuda.set_device(self.device)
        self.train_loader = self._prepare_dataloader(train_dataset)
        self.test_loader = self._prepare_dataloader(test_dataset) if test_dataset else None
        self.model = model.to(self.device)
        self.optimizer = optimizer
        self.save_every = config.save_every if config.save_every else 1
        self.last_epoch = 0
        if self.config.snapshot_path is None:
            self.config.snapshot_path = "gpt_snapshot.pt"
        self._load_snapshot()
        self.raw_model = self.model
        self.dp = None
        if dist.is_initialized() and dist.get_world_size() > 1:
            self.dp = DataParallel(self.model, device_ids=[self.local_rank])
            self.model = self.dp
        self.tokens_seen = 0
        # the fast path: the whole step (forward, backward, clip + AdamW) replayed as a CUDA graph (graph segments with
        # the all-reduces between them under data parallelism). Needs the fused optimizer and a static batch shape;
        # anything else (a foreign optimizer, the ragged last batch of an epoch) runs the same kernels eagerly.
        self._stepper = None
        if isinstance(optimizer, FusedAdamW) and os.environ.get("MGB_EAGER_STEP", "0") != "1":
            self._stepper = GraphedTrainStep(self.raw_model, optimizer, config.grad_norm_clip, self.dp)
        self._f_tok = None

    @property
    def graph_replays(self) -> int:
        return self._stepper.replays if self._stepper is not None else 0

    def _prepare_dataloader(self, dataset: Dataset):
        if dist.is_initialized():
            sampler = DistributedSampler(dataset)
        else:
            sampler = DistributedSampler(dataset, num_replicas=1, rank=0)
        if self.config.device_loader is not False and _unwrap_subset(dataset)[0] is not None:
            return DeviceWindowLoader(dataset, self.config.batch_size, self.device, sampler)
        workers = self.config.dl_num_workers or 0
        return DataLoader(dataset, batch_size=self.config.batch_size, pin_memory=True, shuffle=False,
                          num_workers=workers, sampler=sampler)

    # ---------------------------------------------------------------- snapshots (mingpt/trainer.py:83-116,149-167)
    def _load_snapshot(self):
        path = self.config.snapshot_path
        try:
            if "://" in path:
                import fsspec
                with fsspec.open(path) as f:
                    snapshot_data = torch.load(f, map_location="cpu")
            else:
                snapshot_data = torch.load(path, map_location="cpu")
        except FileNotFoundError:
            print("Model snapshot not found. You need to train your model from scratch.")
            return
        snap = ModelSnapshot(**snapshot_data)
        self.model.load_state_dict(snap.model_state)
        self.optimizer.load_state_dict(snap.optimizer_state)
        self.last_epoch = snap.final_epoch
        print(f"Resuming training from epoch: {self.last_epoch}")

    def _save_snapshot(self, epoch: int) -> None:
        model = self.model.module if hasattr(self.model, "module") else self.model
        snapshot = dict(model_state=model.state_dict(), optimizer_state=self.optimizer.state_dict(), final_epoch=epoch)
        path = self.config.snapshot_path
        if "://" in path:
            import fsspec
            with fsspec.open(path, "wb") as f:
                torch.save(snapshot, f)
        else:
            torch.save(snapshot, path)
        print(f"Model snapshot taken and saved at epoch {epoch}")

    # ---------------------------------------------------------------- step / epoch
    def _run_batch(self, inputs, labels, train: bool = True):
        """mingpt/trainer.py:118-133. Returns the loss as a device scalar (float() it to sync)."""
        if not train:
            with torch.no_grad():
                _, loss = self.model(inputs, labels)
            return loss
        if self._stepper is not None and inputs.shape[0] == self.config.batch_size:
            return self._stepper(inputs, labels)       # static device scalar: read it before the next step
        return train_step(self.raw_model, self.optimizer, inputs, labels, self.config.grad_norm_clip, self.dp,
                          zero_grad=True)

    def _run_epoch(self, epoch: int, dataloader: DataLoader, train: bool = True) -> None:
        t0, tok0 = time.time(), self.tokens_seen
        for idx, (x, y) in enumerate(dataloader):
            if train and self._stepper is not None and x.shape[0] == self.config.batch_size:
                inputs, labels = x, y      # pinned host batch: the graphed step copies it straight into its static buffers
            else:
                inputs = x.to(self.device, non_blocking=True)
                labels = y.to(self.device, non_blocking=True)
            loss = self._run_batch(inputs, labels, train)
            self.tokens_seen += inputs.numel()
            if idx % 100 == 0:
                lv = float(loss)   # the only device->host sync
                self.raw_model.check_device_errors()    # out-of-range token ids / targets flagged by the kernels
                dt = max(time.time() - t0, 1e-9)
                tps = (self.tokens_seen - tok0) / dt
                if self._f_tok is None:
                    self._f_tok = model_flops_per_token(self.raw_model, inputs.shape[1])
                # forward only = a third of the training flops
                util = mfu(tps, self._f_tok if train else self._f_tok / 3.0)
                print(f"[GPU{self.local_rank}] Epoch {epoch},{'Training' if train else 'Test'} loss {lv} "
                      f"| {tps:.0f} tok/s/rank | MFU {100.0 * util:.1f}% of 2.25 PFLOP/s bf16")

    def train(self) -> None:
        for epoch in range(self.last_epoch, self.config.max_epochs):
            if self.config.shuffle_each_epoch:      # the DistributedSampler.set_epoch call the reference lacks
                for loader in (self.train_loader, self.test_loader):
                    if loader is not None and hasattr(loader.sampler, "set_epoch"):
                        loader.sampler.set_epoch(epoch)
            self._run_epoch(epoch, self.train_loader, True)
            if self.local_rank == 0 and epoch % self.save_every == 0:
                self._save_snapshot(epoch)
            if self.test_loader is not None:
                self._run_epoch(epoch, self.test_loader, False)


class Trainer:
    """The README.md:47-53 / upstream minGPT spelling: Trainer(config, model, dataset).run(), iteration
    based, with callbacks. Uses the same fused step."""

    @staticmethod
    def get_default_config() -> CfgNode:
        return CfgNode(device="auto", num_workers=4, max_iters=None, batch_size=64, learning_rate=3e-4,
                       betas=(0.9, 0.95), weight_decay=0.1, grad_norm_clip=1.0)

    def __init__(self, config, model, train_dataset):
        self.config = config
        self.model = model
        self.train_dataset = train_dataset
        self.callbacks = defaultdict(list)
        self.optimizer = None
        if config.device == "auto":
            self.device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        else:
            self.device = torch.device(config.device)
        self.model = self.model.to(self.device)
        print("running on device", self.device)
        self.iter_num, self.iter_time, self.iter_dt = 0, 0.0, 0.0
        self.loss = None
        self.graph_replays = 0

    def add_callback(self, onevent: str, callback):
        self.callbacks[onevent].append(callback)

    def set_callback(self, onevent: str, callback):
        self.callbacks[onevent] = [callback]

    def trigger_callbacks(self, onevent: str):
        for cb in self.callbacks.get(onevent, []):
            cb(self)

    def run(self):
        model, config = self.model, self.config
        self.optimizer = model.configure_optimizers(config)
        dp = DataParallel(model) if dist.is_initialized() and dist.get_world_size() > 1 else None
        sampler = torch.utils.data.RandomSampler(self.train_dataset, replacement=True, num_samples=int(1e10))
        loader = DataLoader(self.train_dataset, sampler=sampler, shuffle=False, pin_memory=True,
                            batch_size=config.batch_size, num_workers=config.num_workers)
        model.train()
        stepper = None
        if isinstance(self.optimizer, FusedAdamW) and os.environ.get("MGB_EAGER_STEP", "0") != "1":
            stepper = GraphedTrainStep(model, self.optimizer, config.grad_norm_clip, dp)
        self.iter_num, self.iter_time = 0, time.time()
        it = iter(loader)
        while True:
            try:
                batch = next(it)
            except StopIteration:
                it = iter(loader)
                batch = next(it)
            x, y = batch
            if stepper is not None and x.shape[0] == config.batch_size:
                self.loss = stepper(x, y)          # copies the pinned host batch into its static device buffers
                self.graph_replays = stepper.replays
            else:
                x, y = x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)
                self.loss = train_step(model, self.optimizer, x, y, config.grad_norm_clip, dp, zero_grad=True)
            self.trigger_callbacks("on_batch_end")
            self.iter_num += 1
            tnow = time.time()
            self.iter_dt, self.iter_time = tnow - self.iter_time, tnow
            if config.max_iters is not None and self.iter_num >= config.max_iters:
                break
