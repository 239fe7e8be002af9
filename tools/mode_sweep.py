"""A/B of the gradient-sync variants inside ONE torchrun job (same box, same processes, interleaved rounds), because
separate bench.py runs of a 0.5 ms step differ by several per cent from run to run:

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 tools/mode_sweep.py [--batch 32] [--out file.json]

Every variant is built once, warmed up (graph capture), then timed `--rounds` times for `--steps` steps each, the
variants taking turns; device-timed with CUDA events, max over ranks, the minimum and the median over rounds reported."""
import argparse
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402
from distributed_pytorch_b200 import _native  # noqa: E402

VARIANTS = [  # name, mode, kwargs
    ("all_reduce (one collective after backward)", "all_reduce", dict(overlap=False)),
    ("all_reduce overlap cap25 b16", "all_reduce", dict(overlap=True, cap=25, blocks=16)),
    ("all_reduce overlap cap25 b16 p2p", "all_reduce", dict(overlap=True, cap=25, blocks=16, nvls=False)),
    ("all_reduce overlap cap25 b8", "all_reduce", dict(overlap=True, cap=25, blocks=8)),
    ("all_reduce overlap cap25 b24", "all_reduce", dict(overlap=True, cap=25, blocks=24)),
    ("ddp cap25 b16", "ddp", dict(cap=25, blocks=16)),
    ("ddp cap25 b16 p2p", "ddp", dict(cap=25, blocks=16, nvls=False)),
    ("ddp cap25 b8", "ddp", dict(cap=25, blocks=8)),
    ("ddp cap25 b16 unfused", "ddp", dict(cap=25, blocks=16, fuse=False)),
    ("gather", "gather", dict()),
]


def build(mode, rank, world, kw):
    lib = _native.require()
    lib.dpb_set_tuning(14, 1 if kw.get("nvls", True) else 0)
    lib.dpb_set_tuning(13, kw.get("blocks", 16))
    torch.manual_seed(1)
    dev = torch.device("cuda", torch.cuda.current_device())
    model = dpb.VGG11().to(dev)
    if mode == "ddp":
        model = dpb.DistributedDataParallel(model, bucket_cap_mb=kw.get("cap"))
    opt = dpb.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    if mode == "ddp":
        if kw.get("fuse", True):
            model.fuse_optimizer(opt)
        sync = dpb.NoSync(model)
    elif mode == "gather":
        sync = dpb.GatherScatterSync(model, world, rank)
    else:
        sync = dpb.AllReduceSync(model, world, rank, fuse_optimizer=True, optimizer=opt, overlap=kw.get("overlap", False),
                                 bucket_cap_mb=kw.get("cap", 25))
    step = dpb.TrainStep(model, opt, dpb.CrossEntropyLoss(), sync, use_graph=True, warmup=2)
    nvls = bool((hasattr(sync, "nvls") and sync.nvls()) or (hasattr(model, "nvls") and model.nvls()))
    return model, sync, step, nvls


def timed(step, batches, steps):
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        x, y = batches[i % len(batches)]
        step(x, y)
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / steps], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--rounds", type=int, default=4)
    ap.add_argument("--only", default=None, help="comma-separated substrings of variant names")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    batch = args.batch or max(256 // world, 1)
    g = torch.Generator().manual_seed(1234 + rank)
    batches = [(torch.randn(batch, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (batch,), generator=g).cuda())
               for _ in range(8)]
    todo = [v for v in VARIANTS if args.only is None or any(s in v[0] for s in args.only.split(","))]
    built = []
    for name, mode, kw in todo:
        model, sync, step, nvls = build(mode, rank, world, kw)
        for i in range(6):
            step(*batches[i % 8])
        built.append((name, kw, model, sync, step, nvls, []))
    lib = _native.require()
    for _ in range(args.rounds):
        for name, kw, model, sync, step, nvls, ts in built:
            lib.dpb_set_tuning(14, 1 if kw.get("nvls", True) else 0)  # (baked into the captured graphs already)
            ts.append(timed(step, batches, args.steps))
    rows = []
    for name, kw, model, sync, step, nvls, ts in built:
        rows.append({"variant": name, "nvls": nvls, "ms_min": min(ts), "ms_median": statistics.median(ts), "ms_all": ts,
                     "images_per_sec": batch * world / (min(ts) / 1e3)})
        if rank == 0:
            print(f"N={world} b={batch:3d} {name:30s} nvls={int(nvls)} min {min(ts):.4f} median {statistics.median(ts):.4f} ms/step "
                  f"({batch * world / (min(ts) / 1e3):9.0f} img/s)  rounds {[round(t, 4) for t in ts]}", flush=True)
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump({"world": world, "per_rank_batch": batch, "steps": args.steps, "rows": rows}, open(args.out, "w"), indent=1)
    torch.cuda.synchronize()
    dist.barrier()
    for name, kw, model, sync, step, nvls, ts in built:
        if hasattr(model, "close"):
            model.close()
        sync.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
