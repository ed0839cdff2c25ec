"""Data parallelism for the fused training step: bucketed gradient all-reduce over NCCL, launched from a
side stream while the backward pass is still running.

Replaces torch.nn.parallel.DistributedDataParallel as the reference uses it (mingpt/trainer.py:71, all
defaults: 25 MiB buckets, reverse registration order, SUM then divide by world size) -- see SURVEY.md 2c:
  C2  one broadcast of the flat parameter buffer from rank 0 at construction,
  C3  the per-forward broadcast of the L (1,1,T,T) mask buffers disappears (the mask is implicit),
  C4  gradient all-reduce per bucket as soon as the last layer of the bucket has finished its wgrads.
The gradients live in ONE flat fp32 buffer laid out by layer (model.GPT._layer_spans), so a bucket is a
contiguous slice: no flatten/unflatten copies, and the optimizer reads the reduced buffer directly.
The 1/world_size average is not a separate pass: it is folded into FusedAdamW.step(grad_scale=...).

Two knobs for the NVLink / SM budget (both off by default, so the default exchange is torch DDP's: fp32 SUM):
  grad_wire_dtype=torch.bfloat16  the bucket is rounded to bf16 for the wire (cast -> all-reduce -> cast back, all on
                                  the side stream): half the bytes, half the time NCCL's CTAs share the SMs with the
                                  backward GEMMs. Master gradients, AdamW and its state stay fp32.
  comm_sms=n                      while all-reduces are in flight the persistent grids of the compute stream (GEMM,
                                  attention backward, ...) are sized for n fewer SMs (mgb_set_option "reserved_sms"),
                                  so no CTA pair waits a whole tile wave for an SM that NCCL occupies.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def plan_buckets(layer_spans: Sequence[Tuple[int, int]], cap_elems: int) -> List[Tuple[int, int, int]]:
    """Group layer spans (forward order: embeddings, block 0..L-1, final norm + head) into buckets in
    REVERSE order -- the order backward finishes them. Returns [(start, end, trigger_group)]: the bucket
    [start, end) of the flat buffer is complete once `trigger_group` (its earliest layer) is done."""
    buckets: List[Tuple[int, int, int]] = []
    cur_start = cur_end = None
    for g in range(len(layer_spans) - 1, -1, -1):
        s, e = layer_spans[g]
        if cur_start is None:
            cur_start, cur_end = s, e
        else:
            cur_start, cur_end = min(cur_start, s), max(cur_end, e)
        if cur_end - cur_start >= cap_elems or g == 0:
            buckets.append((cur_start, cur_end, g))
            cur_start = cur_end = None
    return buckets


class GradBuckets:
    """All-reduces slices of a flat gradient buffer as layer groups complete.

    Works on CUDA (NCCL, side stream + events) and on CPU tensors (Gloo, synchronous) -- the latter is
    what the world_size-2 CPU tests drive."""

    def __init__(self, flat_grad: torch.Tensor, layer_spans, bucket_cap_mb: float = 25.0, process_group=None,
                 wire_dtype: Optional[torch.dtype] = None, comm_sms: int = 0):
        self.flat = flat_grad
        self.group = process_group
        if wire_dtype not in (None, torch.float32, torch.bfloat16):
            raise ValueError("grad_wire_dtype must be None / torch.float32 / torch.bfloat16")
        self.wire = None
        if wire_dtype == torch.bfloat16 and flat_grad.is_cuda:
            self.wire = torch.empty(flat_grad.numel(), dtype=torch.bfloat16, device=flat_grad.device)
        self.comm_sms = int(comm_sms)
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        cap = max(1, int(bucket_cap_mb * 1024 * 1024 / flat_grad.element_size()))
        self.buckets = plan_buckets(layer_spans, cap)
        self._by_trigger = {g: (s, e) for s, e, g in self.buckets}
        self.cuda = flat_grad.is_cuda
        self.stream = torch.cuda.Stream(device=flat_grad.device) if self.cuda else None
        self.launched = 0
        self.enabled = True   # False = gradient accumulation micro-step (DDP.no_sync)

    def reduce_span(self, s: int, e: int):
        """SUM all-reduce of flat[s:e] on the side stream, ordered after everything the compute stream has queued"""
        chunk = self.flat[s:e]
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                if self.wire is not None:
                    from . import ops
                    w = self.wire[s:e]
                    ops.cast_f32_to_bf16(chunk, w)
                    dist.all_reduce(w, op=dist.ReduceOp.SUM, group=self.group)
                    ops.cast_bf16_to_f32(w, chunk)
                else:
                    dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)
        else:
            dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)
        self.launched += 1

    def reserve_sms(self, on: bool):
        """size the compute stream's persistent grids for `comm_sms` fewer SMs while all-reduces are in flight"""
        if self.cuda and self.comm_sms > 0:
            from . import _lib
            _lib.set_option("reserved_sms", self.comm_sms if on else 0)

    def on_group_done(self, g: int):
        if not self.enabled or self.world == 1:
            return
        span = self._by_trigger.get(g)
        if span is None:
            return
        self.reduce_span(*span)
        self.reserve_sms(True)

    def finish(self):
        """Make the compute stream wait for every outstanding all-reduce."""
        if self.cuda and self.world > 1:
            torch.cuda.current_stream().wait_stream(self.stream)
        self.reserve_sms(False)

    @property
    def grad_scale(self) -> float:
        return 1.0 / self.world


class DataParallel(torch.nn.Module):
    """Drop-in for `DDP(model, device_ids=[local_rank])` (mingpt/trainer.py:71) around a fused GPT:
    same .module attribute, forward delegates, gradients are averaged across ranks by the time
    `finish_backward()` returns (the Trainer calls it before clipping / stepping)."""

    def __init__(self, module, device_ids=None, bucket_cap_mb: float = 25.0, process_group=None,
                 broadcast_parameters: bool = True, grad_wire_dtype: Optional[torch.dtype] = None, comm_sms: int = 0):
        super().__init__()
        self.module = module
        self.process_group = process_group
        self.bucket_cap_mb = bucket_cap_mb
        import os
        if grad_wire_dtype is None and os.environ.get("MGB_GRAD_WIRE", "").lower() in ("bf16", "bfloat16"):
            grad_wire_dtype = torch.bfloat16          # environment defaults, for launch scripts that cannot pass kwargs
        if not comm_sms:
            comm_sms = int(os.environ.get("MGB_COMM_SMS", "0"))
        self.grad_wire_dtype, self.comm_sms = grad_wire_dtype, comm_sms
        self.buckets: Optional[GradBuckets] = None
        if broadcast_parameters and dist.is_initialized() and dist.get_world_size(process_group) > 1:
            dist.broadcast(module._flat, src=0, group=process_group)      # C2, one message
            module.mark_params_dirty()

    def _ensure(self):
        m = self.module
        m._ensure_device_state()
        if self.buckets is None or self.buckets.flat.data_ptr() != m._flat_grad.data_ptr():
            self.buckets = GradBuckets(m._flat_grad, m._layer_spans, self.bucket_cap_mb, self.process_group,
                                       wire_dtype=self.grad_wire_dtype, comm_sms=self.comm_sms)
            m._layer_done_hook = self.buckets.on_group_done

    def forward(self, *args, **kwargs):
        self._ensure()
        return self.module(*args, **kwargs)

    def no_sync(self):
        dp = self

        class _Ctx:
            def __enter__(self_inner):
                dp._ensure()
                dp.buckets.enabled = False

            def __exit__(self_inner, *exc):
                dp.buckets.enabled = True
        return _Ctx()

    def finish_backward(self):
        if self.buckets is not None:
            self.buckets.finish()

    @property
    def grad_scale(self) -> float:
        return self.buckets.grad_scale if self.buckets is not None else 1.0
