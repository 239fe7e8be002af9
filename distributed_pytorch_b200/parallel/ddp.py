"""Bucketed, backward-overlapped data parallelism: this framework's ``DistributedDataParallel``.

Stands in for ``torch.nn.parallel.DistributedDataParallel(model)`` as the reference uses it with all defaults
(main_ddp.py:137, main_part3.py:122; behaviour catalogued in SURVEY.md §3.4):

* construction: parameter shapes are verified across ranks, then parameters and buffers are broadcast from rank 0;
* every training forward: BatchNorm buffers are broadcast from rank 0 (``broadcast_buffers=True``);
* backward: gradients are grouped into buckets in gradient-ready order (first bucket capped at 1 MiB, then 25 MiB; a
  tensor that overflows the cap closes the bucket after being added -- for VGG-11 that gives the reference's rebuilt
  layout of 9 461 800 / 27 157 504 / 305 152 bytes); when the last gradient of a bucket has been accumulated the bucket
  is averaged (pre-divide by world size, SUM) while the rest of backward keeps running;
* ``state_dict()`` keys carry the ``module.`` prefix.

On CUDA the bucket *is* a contiguous range of the NVLink-symmetric gradient arena (the arena is laid out in
gradient-ready order, so no copy-in / copy-out), the bucket planner and ready-count reducer are C++
(``dpb_plan_buckets`` / ``dpb_reducer_*`` in libdpb200), and each bucket is reduced by the two-shot P2P kernel on a side
stream, fenced with events against the main stream.  On CPU (gloo) the same planner drives flat-bucket async
``all_reduce`` calls -- the oracle for tests.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .. import _native as N
from ..ops import kernels as K
from ..utils.arena import FlatArena
from .comm import Communicator, SymmetricBuffer

FIRST_BUCKET_BYTES = 1024 * 1024
DEFAULT_BUCKET_MB = 25        # torch's bucket_cap_mb: reproduces its rebuilt layout (9 461 800 / 27 157 504 / 305 152 bytes)
CUDA_BUCKET_MB = 25           # the CUDA path keeps torch's layout too.  Measured at 8 GPUs / 32 images per rank
                              # (profiles/mode_sweep_8.md): every collective costs ~20 us of peer barriers whatever its
                              # size, so three buckets (0.474 ms/step) beat five 8 MB ones (0.494) and nine 4 MB ones
                              # (0.495), although the small ones start earlier.


def expected_ready_order(module: nn.Module) -> List[int]:
    """Indices into ``list(module.parameters())`` in the order their gradients become ready during backward.
    The network is a chain, so modules finish in reverse; inside a module autograd produces weight then bias, except
    for ``Linear`` (addmm) where the bias gradient comes first -- VGG-11: 33,32,30,31,28,29,...,2,3,0,1."""
    index = {id(p): i for i, p in enumerate(module.parameters())}
    order: List[int] = []
    leaves = [m for m in module.modules() if len(list(m.parameters(recurse=False))) > 0]
    for m in reversed(leaves):
        ps = [p for _, p in m.named_parameters(recurse=False)]
        if isinstance(m, nn.Linear):
            ps = list(reversed(ps))
        order.extend(index[id(p)] for p in ps)
    return order


def plan_buckets(bytes_in_ready_order: List[int], first_cap: int = FIRST_BUCKET_BYTES,
                 cap: int = DEFAULT_BUCKET_MB * 1024 * 1024) -> List[int]:
    """Bucket id per tensor (ready order).  C++ planner when the native library is loadable, same rule in Python
    otherwise (CPU-only environments without nvcc)."""
    n = len(bytes_in_ready_order)
    lib = N.load(build_if_missing=False)
    if lib is not None:
        arr = (ctypes.c_longlong * n)(*bytes_in_ready_order)
        out = (ctypes.c_int * n)()
        lib.dpb_plan_buckets(n, arr, first_cap, cap, out)
        return list(out)
    ids, bucket, fill, limit = [], 0, 0, first_cap
    for i, b in enumerate(bytes_in_ready_order):
        ids.append(bucket)
        fill += b
        if fill >= limit and i + 1 < n:
            bucket, fill, limit = bucket + 1, 0, cap
    return ids


def plan_reverse_buckets(bytes_in_named_order: List[int], first_cap: int = FIRST_BUCKET_BYTES,
                         cap: int = DEFAULT_BUCKET_MB * 1024 * 1024) -> List[int]:
    """Bucket id per parameter (``named_parameters()`` order) for an arena laid out in that order: the DDP cap rule applied
    walking the parameters back to front -- gradients become ready roughly in that order -- so that every bucket is one
    contiguous arena range.  (Cutting in true gradient-ready order would split a layer's conv.weight / conv.bias /
    bn.weight / bn.bias across two buckets whose ranges then interleave.)"""
    n = len(bytes_in_named_order)
    ids_rev = plan_buckets(list(reversed(bytes_in_named_order)), first_cap, cap)
    out = [0] * n
    for pos, b in enumerate(ids_rev):
        out[n - 1 - pos] = b
    return out


def gather_sharded_momentum(comm: Communicator, mom_buf: SymmetricBuffer, ranges) -> None:
    """Momentum of a range lives on the rank that owns the corresponding slice of the two-shot all-reduce (slice r of
    ``ceil(n4 / world)`` float4s); broadcast every slice from its owner."""
    world = comm.world_size
    if world == 1:
        return
    comm.barrier()
    for lo, hi in ranges:
        n4 = (hi - lo) // 4
        sl = (n4 + world - 1) // world * 4
        for r in range(world):
            a, b = lo + r * sl, min(lo + (r + 1) * sl, hi)
            if b > a:
                comm.broadcast(mom_buf, a, b - a, root=r)


class BucketReducer:
    """Reduces contiguous ranges ("buckets") of a gradient arena on a high-priority side stream as soon as the last
    gradient of a bucket has been produced, while backward keeps running on the main stream (the C++ ready-count
    reducer ``dpb_reducer_*``).  With an optimizer attached the SGD update is fused into each bucket's collective (the
    slice owner updates parameters and momentum and pushes the new parameters), so no separate optimizer pass is
    left; with multicast-bound arenas the reduction happens inside the NVSwitch.  Used by
    :class:`DistributedDataParallel` and by ``AllReduceSync(overlap=True)``.

    The collective of a bucket uses few CTAs (``DPB_DDP_BLOCKS``, default 24): it has to fit next to the convolution
    kernels it overlaps instead of competing with them for SM slots.
    """

    def __init__(self, comm: Communicator, arena: FlatArena, grad_buf: SymmetricBuffer, bucket_of_param: List[int],
                 num_buckets: int, scale: float, params_buf: Optional[SymmetricBuffer] = None,
                 mom_buf: Optional[SymmetricBuffer] = None):
        """``bucket_of_param`` is indexed like the parameter list handed to :meth:`attach`."""
        self.comm, self.arena, self.grad_buf = comm, arena, grad_buf
        self.params_buf, self.mom_buf = params_buf, mom_buf
        self.num_buckets = num_buckets
        self.scale = scale
        self._handle = ctypes.c_void_p()
        self._callback_queued = False
        self._optimizer = None
        self._arena_pos = {id(p): pos for pos, p in enumerate(arena.params)}
        self._bucket_of_param = bucket_of_param
        self._params: List[torch.nn.Parameter] = []
        self.on_finalize = None
        self.ranges: List[tuple] = []  # (first float, one past the last float) of every bucket in the arena

    def attach(self, params: List[torch.nn.Parameter], optimizer=None) -> None:
        lib = N.require()
        self._params = list(params)
        n = len(self.arena.params)
        pos_of = [self._arena_pos[id(p)] for p in self._params]
        bucket_by_pos = [0] * n
        for i, pos in enumerate(pos_of):
            bucket_by_pos[pos] = self._bucket_of_param[i]
        offs = (ctypes.c_longlong * n)(*self.arena.offsets)
        cnts = (ctypes.c_longlong * n)(*[(x + 63) // 64 * 64 for x in self.arena.numels])
        bks = (ctypes.c_int * n)(*bucket_by_pos)
        N.check(lib.dpb_reducer_create(self.comm.handle, n, offs, cnts, bks, self.num_buckets, self.grad_buf.table,
                                       self.scale, ctypes.byref(self._handle)), "dpb_reducer_create")
        self.ranges = []
        for b in range(self.num_buckets):
            members = [pos for pos in range(n) if bucket_by_pos[pos] == b]
            self.ranges.append((min(self.arena.offsets[m] for m in members),
                                max(self.arena.offsets[m] + (self.arena.numels[m] + 63) // 64 * 64 for m in members)))
        mc_g = self.grad_buf.mc_ptr if self.comm.nvls_ready(self.grad_buf) else None
        mc_p = self.params_buf.mc_ptr if self.params_buf is not None else None
        if mc_g:
            N.check(lib.dpb_reducer_set_multicast(
                self._handle, mc_g, mc_p, self.params_buf.tensor.data_ptr() if self.params_buf is not None else None,
                self.mom_buf.tensor.data_ptr() if self.mom_buf is not None else None), "dpb_reducer_set_multicast")
        self._pos_of_param = {i: pos for i, pos in enumerate(pos_of)}
        if self.comm.world_size > 1:  # a world of one has nothing to reduce: no hooks, the optimizer does its own step
            for i, p in enumerate(self._params):
                if p.requires_grad:
                    p.register_post_accumulate_grad_hook(self._make_hook(i))
        if optimizer is not None:
            self.fuse_optimizer(optimizer)
        self.prepare()

    # ------------------------------------------------------------------------------------------------ optimizer
    def fuse_optimizer(self, optimizer) -> None:
        """From now on every bucket's collective also applies ``optimizer``'s SGD update to the bucket's parameters."""
        if self.params_buf is None or self.mom_buf is None:
            raise RuntimeError("fusing the optimizer needs symmetric parameter and momentum arenas")
        if self.comm.world_size == 1:
            return
        self._optimizer = optimizer
        self._push_hyper()

    @property
    def fuses_optimizer(self) -> bool:
        return self._optimizer is not None

    def _push_hyper(self) -> None:
        if self._optimizer is None:
            return
        g = self._optimizer.param_groups[0]
        N.check(N.require().dpb_reducer_set_sgd(self._handle, self.params_buf.table, self.mom_buf.table, g["lr"],
                                                g["momentum"], g["weight_decay"], int(g.get("nesterov", False))),
                "dpb_reducer_set_sgd")

    # ------------------------------------------------------------------------------------------------ per iteration
    def prepare(self) -> None:
        N.check(N.require().dpb_reducer_prepare(self._handle), "dpb_reducer_prepare")

    def _make_hook(self, i: int):
        def hook(param):
            self._on_grad_ready(i, param)
        return hook

    def _on_grad_ready(self, i: int, param: torch.Tensor) -> None:
        lib = N.require()
        if not self._callback_queued:  # first gradient of this backward pass
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
            self._push_hyper()  # the learning rate may have changed since the last step
        pos = self._pos_of_param[i]
        want = self.arena.grad_buf.data_ptr() + 4 * self.arena.offsets[pos]
        if param.grad.data_ptr() != want:  # producer did not write in place: move it into the bucket memory
            v = self.arena.grad_view(pos)
            v.copy_(param.grad)
            param.grad = v
        lane = K.active_wgrad_lane(param.device)
        if lane is not None:  # weight gradients computed on the second compute stream: the reduction waits for it too
            N.check(lib.dpb_reducer_wait_stream(self._handle, lane.stream.cuda_stream), "dpb_reducer_wait_stream")
        N.check(lib.dpb_reducer_mark_ready(self._handle, pos, N.stream()), "dpb_reducer_mark_ready")

    def _finalize_backward(self) -> None:
        lib = N.require()
        self._callback_queued = False
        launched = lib.dpb_reducer_launched(self._handle)
        N.check(lib.dpb_reducer_finalize(self._handle, N.stream()), "dpb_reducer_finalize")
        N.count_launches(launched)
        if self._optimizer is not None and launched > 0:
            self._optimizer.skip_next_step = True
        self.prepare()  # ready counts for the next backward pass
        if self.on_finalize is not None:
            self.on_finalize()

    def gather_momentum(self) -> None:
        """With the fused optimizer each rank holds the momentum of its own slice of every bucket only; before a
        checkpoint, pull every slice so ``optimizer.state_dict()`` is complete on every rank (collective)."""
        if self._optimizer is None:
            return
        gather_sharded_momentum(self.comm, self.mom_buf, self.ranges)

    def side_stream(self) -> int:
        out = ctypes.c_void_p()
        N.check(N.require().dpb_reducer_side_stream(self._handle, ctypes.byref(out)), "dpb_reducer_side_stream")
        return out.value

    def close(self) -> None:
        if self._handle:
            torch.cuda.synchronize()
            N.require().dpb_reducer_destroy(self._handle)
            self._handle = ctypes.c_void_p()


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, bucket_cap_mb: Optional[float] = None, broadcast_buffers: bool = True,
                 process_group=None, comm: Optional[Communicator] = None):
        """Same call as ``torch.nn.parallel.DistributedDataParallel(model)`` (reference main_ddp.py:137).  After
        building the optimizer, ``ddp.fuse_optimizer(optimizer)`` (CUDA) moves the SGD update into each bucket's
        collective, which removes the separate optimizer pass from the step.  ``bucket_cap_mb``: None = 25 (torch's default:
        the reference's bucket layout 9 461 800 / 27 157 504 / 305 152 bytes); bucket boundaries do not change any result."""
        super().__init__()
        if bucket_cap_mb is None:
            bucket_cap_mb = CUDA_BUCKET_MB if next(module.parameters()).is_cuda else DEFAULT_BUCKET_MB
        self.module = module
        self.process_group = process_group
        self.broadcast_buffers = broadcast_buffers
        self.world_size = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self._params = list(module.parameters())
        self._cuda = self._params[0].is_cuda
        self._verify_shapes()

        ready = expected_ready_order(module)
        sizes = [self._params[i].numel() * 4 for i in ready]
        bucket_ready = plan_buckets(sizes, FIRST_BUCKET_BYTES, int(bucket_cap_mb * 1024 * 1024))
        self.ready_order = ready
        self.bucket_of_param = [0] * len(self._params)
        for pos, i in enumerate(ready):
            self.bucket_of_param[i] = bucket_ready[pos]
        self.num_buckets = max(bucket_ready) + 1
        self.bucket_sizes = [sum(s for s, b in zip(sizes, bucket_ready) if b == k) for k in range(self.num_buckets)]

        self._callback_queued = False
        self.reducer: Optional[BucketReducer] = None
        self.comm = comm
        self._owns_comm = False
        self._bufs: dict[str, SymmetricBuffer] = {}
        self._buffer_block: Optional[SymmetricBuffer] = None
        if self._cuda:
            self._setup_cuda()
        else:
            self._pending = [0] * self.num_buckets
            self._works = []
        self._sync_module_states()
        if not self._cuda:
            for i, p in enumerate(self._params):
                if p.requires_grad:
                    p.register_post_accumulate_grad_hook(self._make_hook(i))

    # ------------------------------------------------------------------------------------------------ set-up
    def _verify_shapes(self) -> None:
        if self.world_size == 1:
            return
        mine = [tuple(p.shape) for p in self._params]
        allv = [None] * self.world_size
        dist.all_gather_object(allv, mine, group=self.process_group)
        for r, other in enumerate(allv):
            if other != mine:
                raise RuntimeError(f"DDP: parameter shapes on rank {r} differ from rank {self.rank}")

    def _setup_cuda(self) -> None:
        lib = N.require()
        dev = self._params[0].device
        if self.comm is None:
            self.comm = Communicator(self.rank, self.world_size, dev, self.process_group)
            self._owns_comm = True
        if getattr(self.module, "_dpb_arena", None) is not None:
            raise RuntimeError("wrap the model in DistributedDataParallel before any other arena adoption")

        def allocator(numel: int, kind: str) -> torch.Tensor:
            # all three arenas are symmetric: gradients are reduced, parameters are broadcast at construction and pushed
            # by the slice owner when the optimizer is fused, momentum is addressed through the same peer table
            b = self.comm.alloc(numel, multicast=(kind != "mom"))
            self._bufs[kind] = b
            return b.tensor

        # arena laid out in gradient-ready order => every bucket is one contiguous range
        names = [n for n, _ in self.module.named_parameters()]
        order = self.ready_order
        self.arena = FlatArena([self._params[i] for i in order], [names[i] for i in order], allocator)
        self.module._dpb_arena = self.arena
        self._arena_pos = {i: pos for pos, i in enumerate(order)}
        self.reducer = BucketReducer(self.comm, self.arena, self._bufs["grad"], self.bucket_of_param, self.num_buckets,
                                     1.0 / self.world_size, params_buf=self._bufs["param"], mom_buf=self._bufs["mom"])
        self.reducer.attach(self._params)
        # BatchNorm buffers into one symmetric block: floats first, then the int64 counters
        fbufs = [(m, k, b) for m in self.module.modules() for k, b in m._buffers.items()
                 if b is not None and b.dtype == torch.float32]
        ibufs = [(m, k, b) for m in self.module.modules() for k, b in m._buffers.items()
                 if b is not None and b.dtype == torch.int64]
        fl = sum((b.numel() + 3) // 4 * 4 for _, _, b in fbufs)
        il = sum((2 * b.numel() + 3) // 4 * 4 for _, _, b in ibufs)
        if fl + il > 0:
            blk = self.comm.alloc(fl + il)
            off = 0
            with torch.no_grad():
                for m, k, b in fbufs:
                    v = blk.tensor[off:off + b.numel()].view(b.shape)
                    v.copy_(b)
                    m._buffers[k] = v
                    off += (b.numel() + 3) // 4 * 4
                for m, k, b in ibufs:
                    v = blk.tensor[off:off + 2 * b.numel()].view(torch.int64).view(b.shape)
                    v.copy_(b)
                    m._buffers[k] = v
                    off += (2 * b.numel() + 3) // 4 * 4
            self._buffer_block = blk

    def _sync_module_states(self) -> None:
        """Broadcast parameters and buffers from rank 0 (torch DDP's ``_sync_module_states``)."""
        if self.world_size == 1:
            return
        if self._cuda:
            self.comm.broadcast(self._bufs["param"], 0, self.arena.total, root=0)
            if self._buffer_block is not None:
                self.comm.broadcast(self._buffer_block, root=0)
            torch.cuda.synchronize()
        else:
            with torch.no_grad():
                for t in list(self.module.parameters()) + list(self.module.buffers()):
                    dist.broadcast(t, src=0, group=self.process_group)

    def _sync_buffers(self) -> None:
        if self.world_size == 1:
            return
        if self._cuda:
            if self._buffer_block is not None:
                # On the reducer's side stream: the first consumer / producer of the buffers is the first BatchNorm
                # kernel (it updates the running statistics), not the first convolution, so the ~9 us broadcast runs
                # underneath that convolution.  The BatchNorm launcher waits for the event (ops.kernels.buffer_sync_*).
                main = torch.cuda.current_stream()
                side = torch.cuda.ExternalStream(self.reducer.side_stream(), device=self._params[0].device)
                ev = torch.cuda.Event()
                ev.record(main)
                side.wait_event(ev)     # after everything that touched the buffers so far (previous step, eval, load)
                with torch.cuda.stream(side):
                    self.comm.broadcast(self._buffer_block, root=0)
                    done = torch.cuda.Event()
                    done.record(side)
                K.set_buffer_sync_event(self._params[0].device, done)
        else:
            with torch.no_grad():
                for b in self.module.buffers():
                    dist.broadcast(b, src=0, group=self.process_group)

    # ------------------------------------------------------------------------------------------------ forward/backward
    def _pre_forward(self) -> None:
        if self.training and torch.is_grad_enabled():
            if self.broadcast_buffers:
                self._sync_buffers()
            self._prepare_for_backward()

    def forward(self, *args, **kwargs):
        self._pre_forward()
        return self.module(*args, **kwargs)

    def forward_loss(self, x, target):
        """Forward + fused classifier/loss of the wrapped model (see ``VGG.forward_loss``), with DDP's pre-forward
        work (buffer broadcast, reducer reset)."""
        self._pre_forward()
        inner = getattr(self.module, "forward_loss", None)
        if inner is None:
            from ..ops import functional as Fn
            return Fn.cross_entropy(self.module(x), target)
        return inner(x, target)

    def fuse_optimizer(self, optimizer) -> None:
        """CUDA: apply ``optimizer``'s SGD update inside each bucket's collective (``optimizer.step()`` then only
        publishes the state).  No-op on CPU."""
        if self._cuda and self.world_size >= 1:
            self.reducer.fuse_optimizer(optimizer)

    @property
    def fuses_optimizer(self) -> bool:
        return bool(self._cuda and self.reducer is not None and self.reducer.fuses_optimizer)

    def nvls(self) -> bool:
        return bool(self._cuda and self.comm.nvls_ready(self._bufs["grad"]))

    def _prepare_for_backward(self) -> None:
        self._callback_queued = False
        if self._cuda:
            self.reducer.prepare()
        else:
            counts = [0] * self.num_buckets
            for i, p in enumerate(self._params):
                if p.requires_grad:
                    counts[self.bucket_of_param[i]] += 1
            self._pending = counts
            self._works = []

    def _make_hook(self, i: int):
        def hook(param):
            self._on_grad_ready(i, param)
        return hook

    def _on_grad_ready(self, i: int, param: torch.Tensor) -> None:
        # CPU (gloo) path; on CUDA the BucketReducer owns the hooks
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._callback_queued = True
        b = self.bucket_of_param[i]
        self._pending[b] -= 1
        if self._pending[b] == 0 and self.world_size > 1:
            members = [j for j in self.ready_order if self.bucket_of_param[j] == b and self._params[j].grad is not None]
            flat = torch.cat([self._params[j].grad.reshape(-1) for j in members]).div_(self.world_size)
            work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.process_group, async_op=True)
            self._works.append((work, flat, members))

    def _finalize_backward(self) -> None:
        for work, flat, members in self._works:
            work.wait()
            off = 0
            with torch.no_grad():
                for j in members:
                    g = self._params[j].grad
                    g.copy_(flat[off:off + g.numel()].view_as(g))
                    off += g.numel()
        self._works = []

    def close(self) -> None:
        if self._cuda:
            torch.cuda.synchronize()
            if self.reducer is not None:
                self.reducer.close()
                self.reducer = None
            if self._owns_comm and self.comm is not None:
                self.comm.close()
                self.comm = None
