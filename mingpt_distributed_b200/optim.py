"""Fused multi-tensor AdamW + global-norm gradient clipping for the flat parameter layout of model.GPT.

Replaces, on the training step of mingpt/trainer.py:118-133,
  torch.nn.utils.clip_grad_norm(model.parameters(), c)   (:129-130; an extra pass over every gradient)
  optimizer.step() of torch.optim.AdamW with two groups   (:132; mingpt/model.py:107-121)
with: one sum-of-squares reduction over the flat gradient buffer and ONE AdamW launch per parameter
group that applies the clip coefficient on the fly (read on the device, no host sync), honours the
group's weight decay and refreshes the bf16 compute copy of the parameters in the same pass.

`FusedAdamW` is a torch.optim.Optimizer: param_groups, state (step / exp_avg / exp_avg_sq per parameter)
and state_dict()/load_state_dict() have torch.optim.AdamW's layout, so the reference's snapshots
(ModelSnapshot.optimizer_state, mingpt/trainer.py:33-37,156) round-trip.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, model=None):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("invalid AdamW hyper-parameter")
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self._model = model
        self._m = self._v = None
        self._step_t = torch.zeros((), dtype=torch.float32)   # shared by every parameter's state["step"]
        self._tables = {}
        self._norm_sq = None
        self._step_dev = None      # device-side step counter (CUDA-graph replay)

    # ------------------------------------------------------------------ helpers
    def _flat_ready(self) -> bool:
        m = self._model
        return m is not None and getattr(m, "_flat", None) is not None and m._flat.is_cuda

    def _names(self):
        m = self._model
        cache = self.__dict__.get("_name_of")
        if cache is None:
            cache = {id(p): n for n, p in m.named_parameters()}
            self.__dict__["_name_of"] = cache
        return cache

    def _ensure_state(self):
        """flat exp_avg / exp_avg_sq buffers, adopting per-tensor state loaded by load_state_dict"""
        m = self._model
        if self._m is not None and self._m.device == m._flat.device and self._m.numel() == m._flat.numel():
            return
        self._m = torch.zeros_like(m._flat)
        self._v = torch.zeros_like(m._flat)
        names = self._names()
        step = None
        for group in self.param_groups:
            for p in group["params"]:
                o, n = m._spans[names[id(p)]]
                mv, vv = self._m[o:o + n].view(p.shape), self._v[o:o + n].view(p.shape)
                st = self.state[p]
                if "exp_avg" in st:
                    mv.copy_(st["exp_avg"])
                    vv.copy_(st["exp_avg_sq"])
                    step = float(st["step"]) if step is None else step
                st["exp_avg"], st["exp_avg_sq"], st["step"] = mv, vv, self._step_t
        if step is not None:
            self._step_t.fill_(step)
        self._tables = {}

    def _table(self, gi: int, group) -> tuple:
        m = self._model
        names = self._names()
        live = tuple(id(p) for p in group["params"] if p.grad is not None)
        key = (gi, live, group["weight_decay"], m._flat.data_ptr())
        hit = self._tables.get(gi)
        if hit is not None and hit[0] == key:
            return hit[1], hit[2]
        entries = []
        for p in group["params"]:
            if p.grad is None:
                continue
            o, n = m._spans[names[id(p)]]
            entries.append((o, n, float(group["weight_decay"])))
        host, count = ops.make_adamw_table(entries)
        dev = host.to(m._flat.device)
        self._tables[gi] = (key, dev, count)
        return dev, count

    def _sync_foreign_grads(self):
        """If somebody replaced p.grad by a tensor that is not the flat-buffer view, copy it in."""
        m = self._model
        names = self._names()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                o, n = m._spans[names[id(p)]]
                want = m._flat_grad.data_ptr() + 4 * o
                if p.grad.data_ptr() != want:
                    m._flat_grad[o:o + n].view(p.shape).copy_(p.grad)

    # ------------------------------------------------------------------ clipping
    @torch.no_grad()
    def grad_norm_sq(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """sum(g^2) over the flat gradient buffer as a device scalar (padding between tensors is zero)"""
        m = self._model
        if self._norm_sq is None or self._norm_sq.device != m._flat.device:
            self._norm_sq = torch.zeros(1, dtype=torch.float32, device=m._flat.device)
        out = self._norm_sq if out is None else out
        out.zero_()
        ops.sumsq(m._flat_grad, out)
        return out

    # ------------------------------------------------------------------ step
    def begin_graph_mode(self):
        """Call right before capturing a CUDA graph that contains step(graph_step=True): moves the step count
        to a device counter the captured kernels read (host arguments are frozen at capture time)."""
        m = self._model
        self._step_dev = torch.tensor([int(self._step_t.item())], dtype=torch.int32, device=m._flat.device)

    def note_replay(self):
        """Call after each replay of such a graph: keeps the host-side step (state_dict) in sync."""
        self._step_t += 1

    @torch.no_grad()
    def step(self, closure=None, *, grad_scale: float = 1.0, gradnorm_sq: Optional[torch.Tensor] = None,
             max_norm: float = 0.0, graph_step: bool = False):
        """One AdamW update. grad_scale multiplies every gradient (1/world_size after a SUM all-reduce,
        1/accumulation_steps, ...); gradnorm_sq (device scalar from grad_norm_sq()) + max_norm > 0 apply
        clip_grad_norm's coefficient inside the same kernel. graph_step=True: the step count is taken from
        (and advanced in) device memory so the call can be captured in a CUDA graph and replayed."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not self._flat_ready():
            raise RuntimeError("FusedAdamW needs a mingpt_distributed_b200 GPT on a CUDA device "
                               "(there is no CPU optimizer path)")
        m = self._model
        m._ensure_device_state()
        if m._flat_grad is None:
            return loss
        self._ensure_state()
        self._sync_foreign_grads()
        step_dev = None
        if graph_step:
            if self._step_dev is None:
                self.begin_graph_mode()
            self._step_dev += 1            # a device kernel: part of the captured graph
            step_dev, step = self._step_dev, 1
        else:
            self._step_t += 1
            step = int(self._step_t.item())
            self._step_dev = None
        for gi, group in enumerate(self.param_groups):
            table, count = self._table(gi, group)
            if count == 0:
                continue
            b1, b2 = group["betas"]
            ops.adamw_multi(m._flat, m._flat_grad, self._m, self._v, m._flat_bf16, table, count, float(group["lr"]),
                            float(b1), float(b2), float(group["eps"]), step, grad_scale, gradnorm_sq, float(max_norm),
                            step_dev=step_dev)
        # the kernel refreshed the bf16 shadow: no recast needed before the next forward
        m._shadow_token = m._param_token()
        return loss

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._m = self._v = None   # adopt the loaded per-tensor state on the next step
        self.__dict__.pop("_name_of", None)


def clip_grad_norm_(model, max_norm: float) -> torch.Tensor:
    """In-place global-norm clipping of the flat gradient buffer (torch.nn.utils.clip_grad_norm_ semantics,
    mingpt/trainer.py:129-130) for callers that keep a foreign optimizer. Returns the norm (device scalar)."""
    norm_sq = torch.zeros(1, dtype=torch.float32, device=model._flat_grad.device)
    ops.sumsq(model._flat_grad, norm_sq)
    ops.clip_scale(model._flat_grad, norm_sq, max_norm)
    return norm_sq.sqrt()
