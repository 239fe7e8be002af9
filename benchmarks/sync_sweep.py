"""Gradient-sync bandwidth sweep (BASELINE.json config 5): our NVLink P2P kernels vs the reference's NCCL calls, 4 KB-64 MB
plus the VGG-11 tensor / DDP bucket / whole-arena sizes.  Run under torchrun, one process per GPU:

    torchrun --standalone --nproc-per-node 8 benchmarks/sync_sweep.py [--out gpurun_out/sync_sweep_8.json]

Device-timed with CUDA events, max over ranks.  Roofline per message of S bytes on N ranks, against the NOMINAL
900 GB/s per direction per GPU that BASELINE.json fixes (the fraction against the 770 GB/s measured peer copy of
B200_PROFILING.md is printed next to it):  all-reduce 2(N-1)/N * S / link;  gather->scatter (N-1) * S / link (rank-0
ingress bound).  The NVLS arm (multimem.ld_reduce / multimem.st, reduction inside the NVSwitch) moves only ~S bytes per
direction per GPU, so against the two-shot roofline it can exceed 1.0; its own floor S / link is reported too.
Reference arms: `all_reduce(SUM)` + `/= N` (main_all_reduce.py:47-48) and
gather -> stack -> mean -> scatter (main_gather.py:42-59), on NCCL."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_b200.parallel.comm import Communicator  # noqa: E402

LINK_GBS = 900.0       # nominal NVLink 5 per direction per GPU (BASELINE.json)
LINK_MEASURED_GBS = 770.0  # measured peer copy (B200_PROFILING.md)


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) * 1e3 / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = Communicator(rank, world, torch.device("cuda", local))
    sizes = [4 << 10, 16 << 10, 64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20]
    named = {"fc1.bias(40B->64B)": 64, "bn(2KB)": 2048, "conv512(9.4MB)": 9437184, "ddp_bucket0": 9461800 // 16 * 16,
             "ddp_bucket1": 27157504, "ddp_bucket2": 305152, "whole_arena": 36924456 // 16 * 16 + 16}
    todo = [(f"{s >> 10}KB" if s < (1 << 20) else f"{s >> 20}MB", s) for s in sizes] + list(named.items())
    buf = comm.alloc(max(s for _, s in todo) // 4, multicast=True)
    buf.tensor.normal_()
    nvls = comm.nvls_ready(buf)
    if rank == 0:
        print(f"world {world}: transport {'VMM' if comm.use_vmm else 'CUDA IPC'}, multicast "
              f"{'bound' if nvls else 'unavailable (' + str(comm.multicast_error) + ')'}", flush=True)
    rows = []
    for label, nbytes in todo:
        n = nbytes // 4
        t = torch.randn(n, device="cuda")
        ours_ar = timed(lambda: comm.all_reduce(buf, 0, n, 1.0 / world), args.iters)
        ours_nvls = timed(lambda: comm.all_reduce_nvls(buf, 1.0 / world, offset=0, count=n), args.iters) if nvls else None
        ours_gs = timed(lambda: comm.gather_mean_scatter(buf, 0, n, 1.0 / world, root=0), args.iters)

        def nccl_ar():
            dist.all_reduce(t)
            t.div_(world)

        def nccl_gs():
            gl = [torch.zeros_like(t) for _ in range(world)] if rank == 0 else None
            dist.gather(t, gather_list=gl, dst=0)
            sl = None
            if rank == 0:
                m = torch.mean(torch.stack(gl, 0), 0)
                sl = [m for _ in range(world)]
            dist.scatter(t, scatter_list=sl, src=0)

        ref_ar = timed(nccl_ar, args.iters)
        ref_gs = timed(nccl_gs, max(args.iters // 2, 5))
        roof_ar = 2.0 * (world - 1) / world * nbytes / (LINK_GBS * 1e3)  # us
        roof_gs = (world - 1) * nbytes / (LINK_GBS * 1e3)
        rows.append({"label": label, "bytes": nbytes, "ours_allreduce_us": ours_ar, "nccl_allreduce_us": ref_ar,
                     "ours_gather_scatter_us": ours_gs, "nccl_gather_scatter_us": ref_gs,
                     "allreduce_busbw_gbs": 2.0 * (world - 1) / world * nbytes / ours_ar / 1e3,
                     "allreduce_roofline_frac": roof_ar / ours_ar, "gather_roofline_frac": roof_gs / ours_gs,
                     "allreduce_roofline_frac_770": roof_ar / ours_ar * LINK_GBS / LINK_MEASURED_GBS,
                     "gather_roofline_frac_770": roof_gs / ours_gs * LINK_GBS / LINK_MEASURED_GBS,
                     "nvls_allreduce_us": ours_nvls,
                     "nvls_roofline_frac": (roof_ar / ours_nvls) if ours_nvls else None,
                     "nvls_floor_frac": (nbytes / (LINK_GBS * 1e3) / ours_nvls) if ours_nvls else None})
        if rank == 0:
            r = rows[-1]
            nv = f"nvls {ours_nvls:9.1f} us roof {r['nvls_roofline_frac']:.2f} | " if ours_nvls else ""
            print(f"{label:>20s} {nbytes:>10d} B | all-reduce {nv}p2p {ours_ar:9.1f} us  nccl {ref_ar:9.1f} us  "
                  f"x{ref_ar / ours_ar:5.2f}  roof {r['allreduce_roofline_frac']:.2f} | gather-scatter ours {ours_gs:9.1f} us  "
                  f"nccl {ref_gs:9.1f} us  x{ref_gs / ours_gs:5.2f}  roof {r['gather_roofline_frac']:.2f}", flush=True)
    # grid sweep on the whole arena: how many CTAs each all-reduce needs to fill the wire
    n = 36924464 // 4
    grid_rows = []
    for blocks in (8, 16, 32, 64, 128):
        p2p = timed(lambda: comm.all_reduce(buf, 0, n, 1.0 / world, algo=2, blocks=blocks), args.iters)
        nv = timed(lambda: comm.all_reduce_nvls(buf, 1.0 / world, offset=0, count=n, blocks=blocks), args.iters) if nvls else None
        grid_rows.append({"blocks": blocks, "p2p_us": p2p, "nvls_us": nv})
        if rank == 0:
            print(f"  whole arena, {blocks:3d} CTAs: p2p {p2p:8.1f} us" + (f"   nvls {nv:8.1f} us" if nv else ""), flush=True)
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as fh:
            json.dump({"world": world, "link_gbs": LINK_GBS, "link_measured_gbs": LINK_MEASURED_GBS, "nvls": nvls,
                       "transport": "vmm" if comm.use_vmm else "ipc", "rows": rows, "grid_sweep": grid_rows}, fh,
                      indent=1)
    torch.cuda.synchronize()
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
