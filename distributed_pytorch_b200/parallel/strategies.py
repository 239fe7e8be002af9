"""Gradient-synchronisation strategies: the one thing that differs between the reference's training scripts
(SURVEY.md §1 layer L2).

================  ======================================  ==================================================
strategy          reference                               here (CUDA)
================  ======================================  ==================================================
``NoSync``        main.py (single process)                nothing
``GatherScatter`` main_gather.py:42-59, 34x gather +      ONE kernel: rank 0 pulls every peer's gradient arena
                  stack + mean + scatter, serial          over NVLink, reduces in rank order, x1/N in registers,
                                                          pushes the mean back with peer stores
``AllReduce``     main_all_reduce.py:45-48, 34x           two-shot all-reduce over the arena -- in-switch (NVLS
                  all_reduce(SUM) + ``/= N``              multimem) from 4 GPUs, peer loads / stores otherwise -- with
                                                          the 1/N and the SGD update fused in; by default cut into
                                                          buckets that are reduced while backward is still running
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
                # gradients always live in symmetric memory; parameters too when the optimizer is fused into the
                # collective (the slice owner pushes new parameters).  Momentum is only ever touched by its owner, but
                # the P2P fused kernel addresses it through the same peer table, so it is symmetric as well.
                if kind == "grad" or symmetric_params:
                    b = self.comm.alloc(numel, multicast=(kind != "mom"))
                    self._bufs[kind] = b
                    return b.tensor
                return torch.zeros(numel, device=self.comm.device, dtype=torch.float32)

            existing = getattr(model, "_dpb_arena", None)
            if existing is not None and "grad" not in self._bufs:
                raise RuntimeError("model already adopted by an arena without symmetric gradient memory")
            self.arena = adopt_module(model, allocator)

    def close(self) -> None:
        red = getattr(self, "reducer", None)
        if red is not None:
            red.close()
            self.reducer = None
        if self.cuda and self._owns_comm and self.comm is not None:
            self.comm.close()
            self.comm = None


class GatherScatterSync(_ArenaSync):
    name = "gather"

    def sync(self) -> None:
        if self.cuda:
            self.arena.ensure_grads_in_place()
            if self.world_size == 1:
                return  # the mean over one rank is the gradient itself
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
                 optimizer=None, overlap: bool = False, bucket_cap_mb: float = 25.0):
        """``fuse_optimizer``: the SGD update runs inside the collective (each rank updates its slice, pushes the new
        parameters).  ``overlap``: instead of one collective after backward, the arena is cut into buckets (the DDP
        rule) and each bucket is reduced on a side stream as soon as its last gradient exists, while backward goes on
        -- nothing in the reference's loop (main_all_reduce.py:45-48) depends on the order in which parameters are
        reduced, only on the result."""
        super().__init__(model, world_size, rank, comm, group, symmetric_params=fuse_optimizer)
        self.fuses_optimizer = bool(fuse_optimizer and self.cuda and self.world_size > 1)
        self.optimizer = optimizer
        if self.fuses_optimizer and optimizer is None:
            raise ValueError("fuse_optimizer=True needs the optimizer")
        self.reducer = None
        if overlap and self.cuda and self.world_size > 1:
            from .ddp import BucketReducer, plan_reverse_buckets
            params = list(model.parameters())
            # the arena is in named_parameters() order: buckets are cut back to front so each is one contiguous range
            bucket_of = plan_reverse_buckets([p.numel() * 4 for p in params], 1024 * 1024,
                                             int(bucket_cap_mb * 1024 * 1024))
            self.reducer = BucketReducer(self.comm, self.arena, self._bufs["grad"], bucket_of, max(bucket_of) + 1,
                                         1.0 / self.world_size, params_buf=self._bufs.get("param"),
                                         mom_buf=self._bufs.get("mom"))
            self.reducer.attach(params, optimizer if self.fuses_optimizer else None)

    def nvls(self) -> bool:
        """True if the collective runs inside the NVSwitch (multicast-bound arenas)."""
        if not self.cuda:
            return False
        need = [self._bufs["grad"]] + ([self._bufs["param"]] if self.fuses_optimizer else [])
        return self.comm.nvls_ready(*need)

    def sync(self) -> None:
        if self.cuda:
            if self.reducer is not None:
                # buckets were launched from the gradient hooks during backward; the end-of-backward callback has
                # already made this stream wait for them
                if self.fuses_optimizer:
                    self.optimizer.skip_next_step = True
                return
            self.arena.ensure_grads_in_place()
            if self.world_size == 1:
                return  # nothing to average: optimizer.step() runs the plain fused SGD kernel over the arena
            scale = 1.0 / self.world_size
            if self.fuses_optimizer:
                g = self.optimizer.param_groups[0]
                if self.nvls():
                    self.comm.all_reduce_nvls(self._bufs["grad"], scale, self._bufs["param"], self._bufs["mom"].tensor,
                                              g["lr"], g["momentum"], g["weight_decay"], g["nesterov"])
                else:
                    self.comm.all_reduce_sgd(self._bufs["grad"], self._bufs["param"], self._bufs["mom"], scale,
                                             g["lr"], g["momentum"], g["weight_decay"], g["nesterov"],
                                             write_mean=False)
                self.optimizer.skip_next_step = True
            elif self.nvls():
                self.comm.all_reduce_nvls(self._bufs["grad"], scale)
            else:
                self.comm.all_reduce(self._bufs["grad"], 0, self.arena.total, scale)
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
        from .ddp import gather_sharded_momentum
        ranges = self.reducer.ranges if self.reducer is not None else [(0, self.arena.total)]
        gather_sharded_momentum(self.comm, self._bufs["mom"], ranges)
