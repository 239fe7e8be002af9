"""SGD with momentum and weight decay as one fused kernel over the flat arena.

API and ``state_dict()`` layout follow ``torch.optim.SGD`` (the reference builds
``optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)``, main.py:103-104), so checkpoints
interchange.  On CUDA:

* parameters adopted by a :class:`~distributed_pytorch_b200.utils.arena.FlatArena` whose gradients live in the
  gradient arena are updated by a single launch over the whole arena (replaces 4 x 34 elementwise ops of the
  reference's optimizer step);
* any other CUDA parameter is updated by the same kernel, one launch per tensor.

CPU parameters (tests, the reference's CPU configuration) use the equivalent PyTorch expressions.
"""
from __future__ import annotations

import torch
from torch.optim import Optimizer

from ..ops import kernels as K

__all__ = ["SGD"]


class SGD(Optimizer):
    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if momentum < 0.0:
            raise ValueError(f"Invalid momentum value: {momentum}")
        if weight_decay < 0.0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if dampening != 0.0:
            raise NotImplementedError("dampening != 0 is not implemented (the reference uses the default 0)")
        if nesterov and momentum <= 0:
            raise ValueError("Nesterov momentum requires a momentum")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        super().__init__(params, defaults)
        # sharded=True is set by the fused all-reduce+SGD sync: the update already happened inside the collective
        self.skip_next_step = False

    # ------------------------------------------------------------------------------------------------ helpers
    def _momentum_buffer(self, p: torch.Tensor) -> torch.Tensor:
        st = self.state[p]
        buf = st.get("momentum_buffer")
        if buf is None:
            arena = getattr(p, "_dpb_arena", None)
            if arena is not None:
                buf = arena.mom_view(p._dpb_index)
            else:
                buf = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["momentum_buffer"] = buf
        return buf

    def load_state_dict(self, state_dict):
        """torch's loader replaces state tensors with copies; re-home momentum buffers into the arena."""
        super().load_state_dict(state_dict)
        with torch.no_grad():
            for group in self.param_groups:
                for p in group["params"]:
                    arena = getattr(p, "_dpb_arena", None)
                    buf = self.state.get(p, {}).get("momentum_buffer")
                    if arena is not None and buf is not None:
                        view = arena.mom_view(p._dpb_index)
                        if buf.data_ptr() != view.data_ptr():
                            view.copy_(buf)
                            self.state[p]["momentum_buffer"] = view

    @staticmethod
    def _arena_of(group):
        """The arena if this group is exactly one arena's parameters with in-place gradients."""
        ps = group["params"]
        if not ps:
            return None
        arena = getattr(ps[0], "_dpb_arena", None)
        if arena is None or len(ps) != len(arena.params):
            return None
        # the same set of parameters in any order (DDP lays its arena out in gradient-ready order): the update is
        # elementwise, so one launch over the whole arena is the same thing
        if {id(p) for p in ps} != {id(p) for p in arena.params}:
            return None
        return arena if arena.grads_in_place() else None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if self.skip_next_step:
            # the fused all-reduce+SGD kernel already applied this step; just make the state visible
            self.skip_next_step = False
            for group in self.param_groups:
                if group["momentum"] != 0:
                    for p in group["params"]:
                        self._momentum_buffer(p)
            return loss
        for group in self.param_groups:
            lr, mom, wd, nest = group["lr"], group["momentum"], group["weight_decay"], group["nesterov"]
            arena = self._arena_of(group) if group["params"] and group["params"][0].is_cuda else None
            if arena is not None:
                if mom != 0:
                    for p in group["params"]:
                        self._momentum_buffer(p)
                K.sgd_step(arena.param_buf, arena.grad_buf, arena.mom_buf if mom != 0 else None, arena.total, lr, mom,
                           wd, 1.0, nest)
                continue
            for p in group["params"]:
                g = p.grad
                if g is None:
                    continue
                if p.is_cuda:
                    if g.stride() != p.stride():
                        g = torch.empty_strided(p.shape, p.stride(), device=p.device, dtype=p.dtype).copy_(g)
                    buf = self._momentum_buffer(p) if mom != 0 else None
                    if buf is not None and buf.stride() != p.stride():
                        raise RuntimeError("momentum buffer layout differs from its parameter")
                    K.sgd_step(p, g, buf, p.numel(), lr, mom, wd, 1.0, nest)
                else:
                    d = g.add(p, alpha=wd) if wd != 0 else g
                    if mom != 0:
                        buf = self._momentum_buffer(p)
                        buf.mul_(mom).add_(d)
                        d = d.add(buf, alpha=mom) if nest else buf
                    p.add_(d, alpha=-lr)
        return loss
