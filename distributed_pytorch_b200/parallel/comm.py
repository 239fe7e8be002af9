"""NVLink symmetric-memory communicator.

Replaces the reference's Gloo/TCP data plane (``init_process_group(backend='gloo', init_method='tcp://ip:6585')``,
main_gather.py:107) for everything on the hot path.  The control plane stays ``torch.distributed`` (TCPStore
rendezvous, same ``--master-ip`` and port): it is used once, to exchange CUDA IPC handles.

* :class:`Communicator` owns this rank's signal pad, the peer mappings of every rank's pad, and the native
  ``dpb_comm`` handle the collective kernels take.
* :class:`SymmetricBuffer` is a device buffer every rank can address: ``tensor`` is the local 1-D fp32 view,
  ``peer_ptrs[r]`` is rank r's copy as mapped into this process.

``Communicator.emulated(n)`` builds an n-rank world inside ONE process on ONE GPU (n buffers, n pads, cooperative
launch with the rank in ``blockIdx.y``): that is how the collective kernels are unit-tested on a single B200.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from .. import _native as N

CH_ALLREDUCE, CH_DDP, CH_BCAST, CH_BARRIER = 0, 1, 2, 3


class _CudaArray:
    """Minimal ``__cuda_array_interface__`` holder so a raw device pointer can become a torch tensor."""

    def __init__(self, ptr: int, numel: int, typestr: str = "<f4"):
        self.__cuda_array_interface__ = {
            "shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None,
        }


def _wrap(ptr: int, numel: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_CudaArray(ptr, numel), device=device)


class SymmetricBuffer:
    def __init__(self, comm: "Communicator", numel: int):
        lib = N.require()
        self.comm = comm
        self.numel = int(numel)
        nbytes = 4 * self.numel
        self._local_ptrs: List[int] = []   # emulated: one per virtual rank; real: just ours
        self._imported: List[int] = []
        nlocal = comm.world_size if comm.emulate else 1
        for _ in range(nlocal):
            p = ctypes.c_void_p()
            N.check(lib.dpb_symm_alloc(nbytes, ctypes.byref(p)), "dpb_symm_alloc")
            self._local_ptrs.append(p.value)
        if comm.emulate or comm.world_size == 1:
            self.peer_ptrs = list(self._local_ptrs) if comm.emulate else [self._local_ptrs[0]]
        else:
            handle = ctypes.create_string_buffer(64)
            N.check(lib.dpb_symm_export(self._local_ptrs[0], handle), "dpb_symm_export")
            handles = comm.exchange(bytes(handle.raw))
            self.peer_ptrs = []
            for r, h in enumerate(handles):
                if r == comm.rank:
                    self.peer_ptrs.append(self._local_ptrs[0])
                else:
                    q = ctypes.c_void_p()
                    N.check(lib.dpb_symm_import(ctypes.create_string_buffer(h, 64), ctypes.byref(q)),
                            f"dpb_symm_import(rank {r})")
                    self.peer_ptrs.append(q.value)
                    self._imported.append(q.value)
        self.table = (ctypes.c_void_p * len(self.peer_ptrs))(*self.peer_ptrs)
        # local views (emulated: one per virtual rank)
        self.tensors = [_wrap(p, self.numel, comm.device) for p in self._local_ptrs]
        self.tensor = self.tensors[comm.rank if comm.emulate else 0]

    def free(self) -> None:
        lib = N.require()
        for q in self._imported:
            lib.dpb_symm_close(q)
        self._imported = []
        for p in self._local_ptrs:
            lib.dpb_symm_free(p)
        self._local_ptrs = []


class Communicator:
    def __init__(self, rank: int, world_size: int, device: torch.device, group=None, emulate: bool = False):
        lib = N.require()
        self.rank, self.world_size, self.device, self.group, self.emulate = rank, world_size, device, group, emulate
        self._buffers: List[SymmetricBuffer] = []
        self._handle = ctypes.c_void_p()
        torch.cuda.set_device(device)
        pad_floats = (lib.dpb_signal_pad_bytes() + 3) // 4
        self._pad = SymmetricBuffer(self, pad_floats)
        N.check(lib.dpb_comm_create(rank, world_size, self._pad.table, int(emulate), ctypes.byref(self._handle)),
                "dpb_comm_create")
        if not emulate and world_size > 1:
            self.host_barrier()

    @classmethod
    def from_process_group(cls, device: Optional[torch.device] = None, group=None) -> "Communicator":
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        device = device or torch.device("cuda", torch.cuda.current_device())
        return cls(rank, world, device, group)

    @classmethod
    def emulated(cls, world_size: int, device: Optional[torch.device] = None) -> "Communicator":
        device = device or torch.device("cuda", torch.cuda.current_device())
        return cls(0, world_size, device, None, emulate=True)

    # ---------------------------------------------------------------------------------------------- control plane
    def exchange(self, payload: bytes) -> List[bytes]:
        import torch.distributed as dist
        out: List[Optional[bytes]] = [None] * self.world_size
        dist.all_gather_object(out, payload, group=self.group)
        return out  # type: ignore[return-value]

    def host_barrier(self) -> None:
        import torch.distributed as dist
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    # ---------------------------------------------------------------------------------------------- buffers
    def alloc(self, numel: int) -> SymmetricBuffer:
        buf = SymmetricBuffer(self, numel)
        self._buffers.append(buf)
        return buf

    @property
    def handle(self):
        return self._handle

    # ---------------------------------------------------------------------------------------------- collectives
    def all_reduce(self, buf: SymmetricBuffer, offset: int = 0, count: Optional[int] = None, scale: float = 1.0,
                   algo: int = 0, blocks: int = 0, channel: int = CH_ALLREDUCE) -> None:
        """In-place sum over ranks times ``scale`` (fixed rank order, bit-identical on every rank)."""
        lib = N.require()
        count = buf.numel - offset if count is None else count
        N.check(lib.dpb_allreduce(self._handle, channel, buf.table, offset, count, scale, algo, blocks, N.stream()),
                "dpb_allreduce")
        N.count_launches(1)

    def all_reduce_sgd(self, grads: SymmetricBuffer, params: SymmetricBuffer, mom: SymmetricBuffer, scale: float,
                       lr: float, momentum: float, wd: float, nesterov: bool = False, write_mean: bool = False,
                       offset: int = 0, count: Optional[int] = None, blocks: int = 0) -> None:
        lib = N.require()
        count = grads.numel - offset if count is None else count
        N.check(lib.dpb_allreduce_sgd(self._handle, CH_ALLREDUCE, grads.table, params.table, mom.table, offset, count,
                                      scale, lr, momentum, wd, int(nesterov), int(write_mean), blocks, N.stream()),
                "dpb_allreduce_sgd")
        N.count_launches(1)

    def gather_mean_scatter(self, buf: SymmetricBuffer, offset: int = 0, count: Optional[int] = None,
                            scale: Optional[float] = None, root: int = 0, blocks: int = 0) -> None:
        lib = N.require()
        count = buf.numel - offset if count is None else count
        scale = 1.0 / self.world_size if scale is None else scale
        N.check(lib.dpb_gather_mean_scatter(self._handle, CH_ALLREDUCE, buf.table, offset, count, scale, root, blocks,
                                            N.stream()), "dpb_gather_mean_scatter")
        N.count_launches(1)

    def broadcast(self, buf: SymmetricBuffer, offset: int = 0, count: Optional[int] = None, root: int = 0,
                  blocks: int = 0) -> None:
        """offset / count in floats (multiples of 4)."""
        lib = N.require()
        count = buf.numel - offset if count is None else count
        if offset % 4 or count % 4:
            raise ValueError("broadcast works in 16-byte units")
        N.check(lib.dpb_broadcast(self._handle, CH_BCAST, buf.table, offset // 4, count // 4, root, blocks,
                                  N.stream()), "dpb_broadcast")
        N.count_launches(1)

    def barrier(self) -> None:
        lib = N.require()
        N.check(lib.dpb_barrier(self._handle, CH_BARRIER, N.stream()), "dpb_barrier")
        N.count_launches(1)

    def close(self) -> None:
        lib = N.require()
        torch.cuda.synchronize(self.device)
        if not self.emulate and self.world_size > 1:
            try:
                self.host_barrier()
            except Exception:  # noqa: BLE001 - best effort during teardown
                pass
        for b in self._buffers:
            b.free()
        self._buffers = []
        if self._handle:
            lib.dpb_comm_destroy(self._handle)
            self._handle = ctypes.c_void_p()
        self._pad.free()
