"""Multi-GPU correctness check, one process per GPU (run under torchrun; also driven by tests/test_gpu_multi.py):

  torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 tools/multi_gpu_check.py

For each sync mode (gather, all_reduce, all_reduce+fused SGD, ddp): a few eager steps and a few CUDA-graph steps on
different data per rank, then
  * the synced gradient equals NCCL's all_reduce(SUM)/N of the local gradients (oracle = the library the reference uses),
  * parameters are bit-identical on every rank after every phase,
  * DDP repairs deliberately diverged initial weights through its rank-0 broadcast.
Rank 0 prints one JSON line and every rank exits non-zero on failure.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402


def params_identical(model) -> bool:
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(flat, ref))], device=flat.device)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    return bool(same.item())


def run_mode(mode: str, rank: int, world: int, batch: int, report: dict):
    torch.manual_seed(1)
    dev = torch.device("cuda", torch.cuda.current_device())
    model = dpb.VGG11().to(dev)
    if mode == "ddp":
        if rank != 0:
            with torch.no_grad():
                for p in model.parameters():
                    p.add_(0.25)
        model = dpb.DistributedDataParallel(model)
    opt = dpb.SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-4)
    if mode == "gather":
        sync = dpb.GatherScatterSync(model, world, rank)
    elif mode in ("all_reduce", "all_reduce_fused"):
        sync = dpb.AllReduceSync(model, world, rank, fuse_optimizer=(mode == "all_reduce_fused"), optimizer=opt)
    else:
        sync = dpb.NoSync(model)
    crit = dpb.CrossEntropyLoss()
    g = torch.Generator().manual_seed(100 + rank)
    info = {}
    info["identical_at_start"] = params_identical(model)

    # --- one hand-rolled step: check the synced gradient against NCCL
    x = torch.randn(batch, 3, 32, 32, generator=g).to(dev)
    y = torch.randint(0, 10, (batch,), generator=g).to(dev)
    if mode in ("gather", "all_reduce"):
        opt.zero_grad()
        crit(model(x), y).backward()
        local = [p.grad.detach().clone() for p in model.parameters()]
        sync.sync()
        torch.cuda.synchronize()
        worst = 0.0
        for p, lg in zip(model.parameters(), local):
            tot = lg.contiguous()
            dist.all_reduce(tot)
            tot /= world
            worst = max(worst, float((p.grad.contiguous() - tot).abs().max() / (tot.abs().max() + 1e-20)))
        info["grad_vs_nccl_rel_err"] = worst
        opt.step()
        info["identical_after_manual_step"] = params_identical(model)

    # --- eager steps then graph steps through TrainStep
    step = dpb.TrainStep(model, opt, crit, sync, use_graph=True, warmup=2)
    losses = []
    for it in range(6):
        x = torch.randn(batch, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (batch,), generator=g).pin_memory()
        losses.append(float(step(x, y).item()))
        if it in (1, 5):
            info[f"identical_after_step{it}"] = params_identical(model)
    info["losses"] = losses
    info["graph_replays"] = step.replays
    info["finite"] = all(l == l and abs(l) < 1e4 for l in losses)
    if mode == "all_reduce_fused":
        sync.gather_momentum()
        torch.cuda.synchronize()
        arena = sync.arena
        m = arena.mom_buf.clone()
        ref = m.clone()
        dist.broadcast(ref, src=0)
        info["momentum_identical_after_gather"] = bool(torch.equal(m, ref))
    torch.cuda.synchronize()
    dist.barrier()
    if hasattr(model, "close"):
        model.close()
    sync.close()
    report[mode] = info
    ok = info["finite"] and all(v for k, v in info.items() if k.startswith("identical") or k.startswith("momentum"))
    if "grad_vs_nccl_rel_err" in info:
        ok = ok and info["grad_vs_nccl_rel_err"] < 1e-5
    return ok and step.replays >= 3


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    batch = int(os.environ.get("CHECK_BATCH", "32"))
    modes = os.environ.get("CHECK_MODES", "gather,all_reduce,all_reduce_fused,ddp").split(",")
    report, ok = {"world": world, "batch": batch}, True
    for mode in modes:
        ok = run_mode(mode, rank, world, batch, report) and ok
    report["ok"] = ok
    if rank == 0:
        print("MULTI_GPU_CHECK " + json.dumps(report), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
