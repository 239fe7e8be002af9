"""Collective kernels on ONE GPU with 8 emulated ranks (cooperative launch, rank = blockIdx.y) over the VGG-11 arena
size, for `ncu --set full`: a multi-rank job must never run under ncu, so this is the only way to get per-kernel
hardware counters for them.  NVLink counters are of course absent (all "peers" are local HBM); what the capture shows is
the memory-side behaviour of the kernels -- bytes moved per element, DRAM / L2 throughput, stall reasons."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_b200.parallel.comm import Communicator  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = 36924464 // 4
comm = Communicator.emulated(world)
buf = comm.alloc(n)
for t in buf.tensors:
    t.normal_()
for _ in range(3):
    comm.all_reduce(buf, 0, n, 1.0 / world, algo=2)
    comm.gather_mean_scatter(buf, 0, n, 1.0 / world, root=0)
    comm.broadcast(buf, 0, n, root=0)
    comm.all_reduce(buf, 0, 65536, 1.0 / world, algo=1)
torch.cuda.synchronize()
print("ok")
comm.close()
