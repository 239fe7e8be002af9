"""Timeline of ONE CUDA-graph replay of the training step, from %globaltimer stamps written by the kernels themselves
(experiments build): shows whether the per-bucket gradient collectives (side stream) really run while the convolution
kernels of earlier layers are still executing on the main stream.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 tools/overlap_timeline.py [--mode ddp|all_reduce]
                                                                    [--batch 128] [--out gpurun_out/overlap_ddp_2.txt]

Rank 0 prints one line per kernel (start / end in microseconds relative to the first kernel of the step) and, for every
collective, which compute kernels overlapped it.  A stamp is taken by thread 0 of block 0 at kernel entry and exit: for
the collectives that is the true extent (block 0 leaves after the closing peer barrier); for wide compute kernels the
grid may run a little longer than its block 0.
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

os.environ["DPB_EXPERIMENTS"] = "1"  # every rank loads (rank 0 builds) the library with the in-kernel time stamps
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_b200 import _native  # noqa: E402

NAMES = {1: "conv fprop", 2: "conv dgrad", 3: "conv wgrad", 4: "splitk reduce", 5: "bn+relu+pool fwd", 6: "bn bwd reduce+finalize",
         7: "bn bwd apply", 8: "first conv fprop", 9: "first conv wgrad", 10: "head fwd", 11: "head bwd", 12: "sgd",
         20: "ALLREDUCE two-shot (P2P)", 21: "ALLREDUCE NVLS (multimem)", 22: "GATHER-MEAN-SCATTER", 23: "BROADCAST",
         24: "ALLREDUCE one-shot"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="ddp", choices=("ddp", "all_reduce", "all_reduce_overlap"))
    ap.add_argument("--batch", type=int, default=None, help="per-rank batch (default 256 / world)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    if rank == 0:
        _native.build(force=True, experiments=True)
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dist.barrier()
    import distributed_pytorch_b200 as dpb
    lib = _native.require()
    dev = torch.device("cuda", local)
    batch = args.batch or max(256 // world, 1)
    torch.manual_seed(1)
    model = dpb.VGG11().to(dev)
    if args.mode == "ddp":
        model = dpb.DistributedDataParallel(model)
    opt = dpb.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    if args.mode == "ddp":
        model.fuse_optimizer(opt)
        sync = dpb.NoSync(model)
    else:
        sync = dpb.AllReduceSync(model, world, rank, fuse_optimizer=True, optimizer=opt,
                                 overlap=(args.mode == "all_reduce_overlap"))
    step = dpb.TrainStep(model, opt, dpb.CrossEntropyLoss(), sync, use_graph=True, warmup=2)
    x = torch.randn(batch, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (batch,), device=dev)
    trace = torch.zeros(1 + 3 * 4096, dtype=torch.int64, device=dev)
    for _ in range(6):
        step(x, y)
    torch.cuda.synchronize()
    dist.barrier()
    _native.check(lib.dpb_debug_set_trace(trace.data_ptr()), "dpb_debug_set_trace")
    torch.cuda.synchronize()
    step(x, y)
    torch.cuda.synchronize()
    lib.dpb_debug_set_trace(None)
    t = trace.cpu().tolist()
    n = min(t[0], 4096)
    recs = sorted(((t[1 + 3 * i], t[2 + 3 * i], t[3 + 3 * i]) for i in range(n)), key=lambda r: r[1])
    if rank == 0 and recs:
        t0 = recs[0][1]
        lines = [f"# one graph replay, mode {args.mode}, {world} GPUs, {batch} images per rank, {n} kernels "
                 f"(replays so far: {step.replays})", f"# {'start us':>9s} {'end us':>9s} {'dur us':>7s}  kernel"]
        comm = [r for r in recs if r[0] >= 20]
        comp = [r for r in recs if r[0] < 20]
        for k, a, b in recs:
            lines.append(f"  {(a - t0) / 1e3:9.2f} {(b - t0) / 1e3:9.2f} {(b - a) / 1e3:7.2f}  {NAMES.get(k, k)}")
        lines.append("#")
        end_compute = max(b for _, _, b in comp) if comp else t0
        for k, a, b in comm:
            under = [NAMES.get(kk, kk) for kk, aa, bb in comp if aa < b and bb > a]
            hidden = max(0, min(b, end_compute) - a)
            lines.append(f"# {NAMES.get(k, k)} [{(a - t0) / 1e3:.1f}, {(b - t0) / 1e3:.1f}] us: {len(under)} compute kernels ran "
                         f"during it ({', '.join(under[:6])}{' ...' if len(under) > 6 else ''}); "
                         f"{100.0 * hidden / max(b - a, 1):.0f} % of it lies before the last compute kernel ends")
        lines.append(f"# step extent {(max(b for _, _, b in recs) - t0) / 1e3:.1f} us; last compute kernel ends at "
                     f"{(end_compute - t0) / 1e3:.1f} us")
        text = "\n".join(lines)
        print(text, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as fh:
                fh.write(text + "\n")
    torch.cuda.synchronize()
    dist.barrier()
    if hasattr(model, "close"):
        model.close()
    sync.close()
    dist.destroy_process_group()
    if rank == 0:
        del os.environ["DPB_EXPERIMENTS"]
        _native.build(force=True)  # back to the production build


if __name__ == "__main__":
    main()
