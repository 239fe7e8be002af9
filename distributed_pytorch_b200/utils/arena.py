"""Flat fp32 arenas for parameters, gradients and SGD momentum.

All three buffers share one layout: parameter ``i`` occupies ``[offset[i], offset[i] + numel[i])`` (offsets aligned to
64 floats = 256 bytes so TMA descriptors and float4 kernels can address every tensor directly).  Module parameters
become strided views into the parameter buffer (shapes and strides unchanged, so ``state_dict()`` still yields the
reference's 58 keys with OIHW shapes), ``param.grad`` becomes a view into the gradient buffer, and the optimizer's
``momentum_buffer`` a view into the momentum buffer.

Why: the reference synchronises 34 separate gradient tensors (main_gather.py:42, main_all_reduce.py:45) and DDP copies
them in and out of flat buckets.  With one arena the whole model is a single contiguous NVLink message, a DDP bucket is
just a sub-range, and the optimizer is one kernel.

The gradient buffer (and, for the fused all-reduce+SGD path, the other two) may be backed by NVLink-symmetric memory
from :mod:`distributed_pytorch_b200.parallel.comm`.
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Sequence

import torch

ALIGN = 64  # floats


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


def dense_span(t: torch.Tensor) -> int:
    """Number of storage elements spanned by a dense (possibly permuted) tensor."""
    if t.numel() == 0:
        return 0
    return 1 + sum((s - 1) * st for s, st in zip(t.shape, t.stride()))


class FlatArena:
    def __init__(self, params: Sequence[torch.nn.Parameter], names: Optional[Sequence[str]] = None,
                 allocator: Optional[Callable[[int, str], torch.Tensor]] = None, tail_multiple: int = 4 * 8 * 64):
        """``allocator(numel, kind)`` returns a zero-initialised 1-D fp32 tensor (kind in {"param","grad","mom"});
        default is ``torch.zeros`` on the parameters' device."""
        params = list(params)
        if not params:
            raise ValueError("FlatArena needs at least one parameter")
        dev = params[0].device
        for p in params:
            if p.dtype != torch.float32:
                raise TypeError("FlatArena holds fp32 parameters only")
            if p.device != dev:
                raise ValueError("all parameters must be on one device")
            if dense_span(p) != p.numel():
                raise ValueError("parameters must be dense in memory")
        self.device = dev
        self.params = params
        self.names = list(names) if names is not None else [f"param{i}" for i in range(len(params))]
        self.offsets: list[int] = []
        self.numels: list[int] = []
        off = 0
        for p in params:
            self.offsets.append(off)
            self.numels.append(p.numel())
            off += _round_up(p.numel(), ALIGN)
        self.total = _round_up(off, tail_multiple)  # divisible into float4 slices for up to 8 ranks
        alloc = allocator or (lambda n, kind: torch.zeros(n, device=dev, dtype=torch.float32))
        self.param_buf = alloc(self.total, "param")
        self.grad_buf = alloc(self.total, "grad")
        self.mom_buf = alloc(self.total, "mom")
        self._adopt()

    # -------------------------------------------------------------------------------------------------- views
    def view(self, buf: torch.Tensor, i: int) -> torch.Tensor:
        p = self.params[i]
        return torch.as_strided(buf, p.shape, p.stride(), self.offsets[i])

    def param_view(self, i: int) -> torch.Tensor:
        return self.view(self.param_buf, i)

    def grad_view(self, i: int) -> torch.Tensor:
        return self.view(self.grad_buf, i)

    def mom_view(self, i: int) -> torch.Tensor:
        return self.view(self.mom_buf, i)

    def _adopt(self) -> None:
        with torch.no_grad():
            for i, p in enumerate(self.params):
                v = self.param_view(i)
                v.copy_(p.data)
                p.data = v
                p.grad = None
                # a fresh view per call: autograd adopts the returned tensor object as param.grad
                p._dpb_grad_slot = (lambda i=i: self.grad_view(i))
                p._dpb_arena = self
                p._dpb_index = i

    # -------------------------------------------------------------------------------------------------- invariants
    def grads_in_place(self) -> bool:
        for i, p in enumerate(self.params):
            g = p.grad
            if g is None or g.data_ptr() != self.grad_buf.data_ptr() + 4 * self.offsets[i]:
                return False
        return True

    def ensure_grads_in_place(self) -> int:
        """Make every ``param.grad`` alias the arena (copying if some producer handed autograd a private tensor).
        Returns the number of gradients that had to be copied."""
        moved = 0
        with torch.no_grad():
            for i, p in enumerate(self.params):
                g = p.grad
                want = self.grad_buf.data_ptr() + 4 * self.offsets[i]
                if g is None:
                    v = self.grad_view(i)
                    v.zero_()
                    p.grad = v
                    moved += 1
                elif g.data_ptr() != want:
                    v = self.grad_view(i)
                    v.copy_(g)
                    p.grad = v
                    moved += 1
        return moved

    def bytes(self) -> int:
        return 4 * self.total


def adopt_module(module: torch.nn.Module, allocator=None) -> FlatArena:
    """Move every parameter of ``module`` into a :class:`FlatArena` (in ``named_parameters()`` order, the order the
    reference's sync loops walk, main_gather.py:42)."""
    existing = getattr(module, "_dpb_arena", None)
    if existing is not None:
        return existing
    names, params = zip(*[(n, p) for n, p in module.named_parameters()])
    arena = FlatArena(params, names, allocator)
    module._dpb_arena = arena
    return arena
