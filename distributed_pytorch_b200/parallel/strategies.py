"""Gradient-synchronisation strategies: the one thing that differs between the reference's training scripts
(SURVEY.md §1 layer L2).

================  ======================================  ==================================================
strategy          reference                               here (CUDA)
================  ======================================  ==================================================
``NoSync``        main.py (single process)                nothing
``GatherScatter`` main_gather.py:42-59, 34x gather +      ONE kernel: rank 0 pulls every peer's gradient arena
                  stack + mean + scatter, serial          over NVLink, reduces in rank order, x1/N in registers,
                                                          pushes the mean back with peer stores
``AllReduce``     main_all_reduce.py:45-48, 34x           ONE kernel over the arena: one-shot / two-shot P2P
                  all_reduce(SUM) + ``/= N``              all-reduce with the 1/N folded in; optionally the SGD
                                                          update fused into the reduce-scatter epilogue
================  ======================================  ==================================================

(Bucketed DDP lives in :mod:`distributed_pytorch_b200.parallel.ddp`.)

On CPU tensors (gloo) the strategies execute the reference's per-parameter collectives verbatim in semantics: that is
the oracle the tests compare the CUDA kernels against, and what the CPU configuration of BASELINE.json runs.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from ..utils.arena import FlatArena, adopt_module
from .comm import Communicator, SymmetricBuffer


class SyncStrategy:
    name = "none"
    fuses_optimizer = False

    def __init__(self, model: torch.nn.Module, world_size: int = 1, rank: int = 0):
        self.model = model
        self.world_size = world_size
        self.rank = rank

    def sync(self) -> None:
        """Average ``param.grad`` across ranks, in place."""

    def close(self) -> None:
        pass


class NoSync(SyncStrategy):
    name = "single"


def _is_cuda_model(model: torch.nn.Module) -> bool:
    return next(model.parameters()).is_cuda


class _ArenaSync(SyncStrategy):
    """Shared set-up: communicator + parameters adopted into a flat arena whose gradient buffer is symmetric."""

    def __init__(self, model, world_size=None, rank=None, comm: Optional[Communicator] = None, group=None,
                 symmetric_params: bool = False):
        world_size = dist.get_world_size(group) if world_size is None else world_size
        rank = dist.get_rank(group) if rank is None else rank
        super().__init__(model, world_size, rank)
        self.group = group
        self.cuda = _is_cuda_model(model)
        self.comm = comm
        self.arena: Optional[FlatArena] = None
        self._bufs: dict[str, SymmetricBuffer] = {}
        self._owns_comm = False
        if self.cuda:
            if self.comm is None:
                dev = next(model.parameters()).device
                self.comm = Communicator(rank, world_size, dev, group)
                self._owns_comm = True

            def allocator(numel: int, kind: str) -> torch.Tensor:
                if kind == "grad" or symmetric_params:
                    b = self.comm.alloc(numel)
                    self._bufs[kind] = b
                    return b.tensor
                return torch.zeros(numel, device=self.comm.device, dtype=torch.float32)

            existing = getattr(model, "_dpb_arena", None)
            if existing is not None and "grad" not in self._bufs:
                raise RuntimeError("model already adopted by an arena without symmetric gradient memory")
            self.arena = adopt_module(model, allocator)

    def close(self) -> None:
        if self.cuda and self._owns_comm and self.comm is not None:
            self.comm.close()
            self.comm = None


class GatherScatterSync(_ArenaSync):
    name = "gather"

    def sync(self) -> None:
        if self.cuda:
            self.arena.ensure_grads_in_place()
            self.comm.gather_mean_scatter(self._bufs["grad"], 0, self.arena.total, 1.0 / self.world_size, root=0)
            return
        # CPU oracle: reference main_gather.py:42-59
        n, rank = self.world_size, self.rank
        for _, p in self.model.named_parameters():
            g = p.grad.contiguous()  # gloo moves raw storage: stage KRSC-strided gradients through a dense tensor
            gl = [torch.zeros_like(g) for _ in range(n)] if rank == 0 else None
            dist.gather(g, gather_list=gl, dst=0, group=self.group)
            sl = None
            if rank == 0:
                mean = torch.mean(torch.stack(gl, dim=0), dim=0)
                sl = [mean for _ in range(n)]
            dist.scatter(g, scatter_list=sl, src=0, group=self.group)
            if g.data_ptr() != p.grad.data_ptr():
                p.grad.copy_(g)


class AllReduceSync(_ArenaSync):
    name = "all_reduce"

    def __init__(self, model, world_size=None, rank=None, comm=None, group=None, fuse_optimizer: bool = False,
                 optimizer=None):
        super().__init__(model, world_size, rank, comm, group, symmetric_params=fuse_optimizer)
        self.fuses_optimizer = bool(fuse_optimizer and self.cuda)
        self.optimizer = optimizer
        if self.fuses_optimizer and optimizer is None:
            raise ValueError("fuse_optimizer=True needs the optimizer")

    def sync(self) -> None:
        if self.cuda:
            self.arena.ensure_grads_in_place()
            if self.fuses_optimizer:
                g = self.optimizer.param_groups[0]
                self.comm.all_reduce_sgd(self._bufs["grad"], self._bufs["param"], self._bufs["mom"],
                                         1.0 / self.world_size, g["lr"], g["momentum"], g["weight_decay"],
                                         g["nesterov"], write_mean=False)
                self.optimizer.skip_next_step = True
            else:
                self.comm.all_reduce(self._bufs["grad"], 0, self.arena.total, 1.0 / self.world_size)
            return
        # CPU oracle: reference main_all_reduce.py:45-48
        for _, p in self.model.named_parameters():
            if p.grad is not None:
                dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group)
                p.grad /= self.world_size

    def gather_momentum(self) -> None:
        """With the fused optimizer each rank holds the momentum of its own slice only; before a checkpoint, pull
        every slice so ``optimizer.state_dict()`` is complete on every rank."""
        if not self.fuses_optimizer or self.world_size == 1:
            return
        n4 = self.arena.total // 4
        sl = (n4 + self.world_size - 1) // self.world_size * 4
        mom = self._bufs["mom"]
        self.comm.barrier()
        for r in range(self.world_size):
            lo, hi = r * sl, min((r + 1) * sl, self.arena.total)
            if hi > lo:
                self.comm.broadcast(mom, lo, hi - lo, root=r)
