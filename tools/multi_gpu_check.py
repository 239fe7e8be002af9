"""Multi-GPU correctness check, one process per GPU (run under torchrun; also driven by tests/test_gpu_multi.py):

  torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 tools/multi_gpu_check.py

Every sync mode -- gather, all_reduce, all_reduce with the fused optimizer, the same bucketed and overlapped with
backward, ddp, ddp with the fused optimizer -- is trained for a few eager and a few CUDA-graph steps on different data
per rank NEXT TO AN ORACLE: a second replica with the same initial weights and the same data whose gradients are
averaged by NCCL (``dist.all_reduce`` + divide, what the reference does on a GPU) and whose weights are updated by
``torch.optim.SGD``.  Checked per mode:

  * step 1: the synced gradient equals the oracle's (modes that materialise it), bit for bit at 2 ranks;
  * EVERY step on its own: before each step the oracle is set to the tested replica's parameters and momentum, both
    take the step on the same batch, and the resulting parameters must agree to rounding (the fused SGD kernels use
    FMAs, torch does not).  A free-running twin is useless as an oracle here: on random labels the trajectory is
    chaotic, one flipped ReLU in step 3 is a 10 % parameter difference by step 6 -- between ANY two implementations;
  * parameters are bit-identical on every rank after every step;
  * DDP repairs deliberately diverged initial weights through its rank-0 broadcast;
  * with the fused optimizer, the momentum gathered for a checkpoint equals the oracle's.

Where the arenas are multicast-bound the modes run twice: reduced inside the NVSwitch (multimem) and with NVLS switched
off (P2P loads / stores).  Rank 0 prints one JSON line; every rank exits non-zero on failure.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

os.environ.setdefault("DPB_NVLS_MIN_WORLD", "2")  # exercise the in-switch path wherever multicast binds, even at 2 GPUs
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402
from distributed_pytorch_b200 import _native  # noqa: E402

MODES = ("gather", "all_reduce", "all_reduce_fused", "all_reduce_overlap", "ddp", "ddp_fused")
LR, MOM, WD = 0.02, 0.9, 1e-4


def params_identical(model) -> bool:
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([int(torch.equal(flat, ref))], device=flat.device)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    return bool(same.item())


def max_rel_err(a, b) -> float:
    """max over tensors of max|x - y| / max|y| (tensors that are still ~0, e.g. BatchNorm biases, count absolutely)."""
    worst = 0.0
    for x, y in zip(a, b):
        x, y = x.detach().contiguous().reshape(-1), y.detach().contiguous().reshape(-1)
        worst = max(worst, float((x - y).abs().max() / (y.abs().max() + 1e-3)))
    return worst


class Oracle:
    """Replica trained the reference's way on a GPU: our model for the local math, NCCL + torch.optim.SGD for the rest."""

    def __init__(self, dev, world):
        torch.manual_seed(1)
        self.model = dpb.VGG11().to(dev)
        self.opt = torch.optim.SGD(self.model.parameters(), lr=LR, momentum=MOM, weight_decay=WD)
        self.crit = dpb.CrossEntropyLoss()
        self.world = world

    def adopt(self, params, momenta):
        """Become an exact copy of the replica under test (parameters, and momentum once it exists)."""
        with torch.no_grad():
            for p, q in zip(self.model.parameters(), params):
                p.copy_(q)
            for p, m in zip(self.model.parameters(), momenta):
                if m is not None and "momentum_buffer" in self.opt.state.get(p, {}):
                    self.opt.state[p]["momentum_buffer"].copy_(m)

    def step(self, x, y):
        self.model.train()
        self.opt.zero_grad()
        loss = self.model.forward_loss(x, y)  # the same kernels as the replica under test: local gradients are bit-equal
        loss.backward()
        for p in self.model.parameters():  # reference main_all_reduce.py:45-48 with backend nccl
            g = p.grad.contiguous()
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            g /= self.world
            p.grad.copy_(g)
        grads = [p.grad.detach().clone() for p in self.model.parameters()]
        self.opt.step()
        return float(loss.item()), grads


def run_mode(mode: str, rank: int, world: int, batch: int, nvls_wanted: bool):
    lib = _native.require()
    lib.dpb_set_tuning(14, 1 if nvls_wanted else 0)
    torch.manual_seed(1)
    dev = torch.device("cuda", torch.cuda.current_device())
    model = dpb.VGG11().to(dev)
    is_ddp = mode.startswith("ddp")
    if is_ddp:
        if rank != 0:
            with torch.no_grad():
                for p in model.parameters():
                    p.add_(0.25)
        model = dpb.DistributedDataParallel(model)
    opt = dpb.SGD(model.parameters(), lr=LR, momentum=MOM, weight_decay=WD)
    fused = mode in ("all_reduce_fused", "all_reduce_overlap", "ddp_fused")
    if mode == "gather":
        sync = dpb.GatherScatterSync(model, world, rank)
    elif mode.startswith("all_reduce"):
        sync = dpb.AllReduceSync(model, world, rank, fuse_optimizer=fused, optimizer=opt,
                                 overlap=(mode == "all_reduce_overlap"))
    else:
        sync = dpb.NoSync(model)
        if fused:
            model.fuse_optimizer(opt)
    crit = dpb.CrossEntropyLoss()
    oracle = Oracle(dev, world)
    raw = model.module if is_ddp else model
    info = {"nvls": bool(model.nvls() if is_ddp else (sync.nvls() if hasattr(sync, "nvls") else False))}
    info["identical_at_start"] = params_identical(model)
    info["equals_oracle_at_start"] = max_rel_err(raw.parameters(), oracle.model.parameters()) == 0.0

    g = torch.Generator().manual_seed(100 + rank)
    step = dpb.TrainStep(model, opt, crit, sync, use_graph=True, warmup=2)
    losses, errs = [], []
    def gather_momentum():
        if fused:
            (model.reducer if is_ddp else sync).gather_momentum()
            torch.cuda.synchronize()

    for it in range(6):
        x = torch.randn(batch, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (batch,), generator=g).pin_memory()
        if it > 0:  # every step is checked on its own: the oracle starts it from the tested replica's exact state
            gather_momentum()
            ours_p = list(raw.parameters())
            oracle.adopt(ours_p, [opt.state.get(p, {}).get("momentum_buffer") for p in ours_p])
        o_loss, o_grads = oracle.step(x.to(dev), y.to(dev))
        losses.append(float(step(x, y).item()))
        torch.cuda.synchronize()
        if it == 0 and not fused:  # the averaged gradient is in param.grad: compare it with NCCL's
            info["grad_vs_oracle_rel_err"] = max_rel_err([p.grad for p in raw.parameters()], o_grads)
        errs.append(max_rel_err(raw.parameters(), oracle.model.parameters()))
        info[f"identical_after_step{it}"] = params_identical(model)
        info.setdefault("loss_vs_oracle", []).append(abs(losses[-1] - o_loss))
    info["param_vs_oracle_rel_err"] = errs
    info["losses"] = losses
    info["graph_replays"] = step.replays
    info["finite"] = all(l == l and abs(l) < 1e4 for l in losses)
    if fused:
        gather_momentum()
        for p in raw.parameters():
            opt._momentum_buffer(p)
        ours = [opt.state[p]["momentum_buffer"] for p in raw.parameters()]
        theirs = [oracle.opt.state[p]["momentum_buffer"] for p in oracle.model.parameters()]
        info["momentum_vs_oracle_rel_err"] = max_rel_err(ours, theirs)
    torch.cuda.synchronize()
    dist.barrier()
    if hasattr(model, "close"):
        model.close()
    sync.close()
    ok = info["finite"] and info["equals_oracle_at_start"]
    ok = ok and all(v for k, v in info.items() if k.startswith("identical"))
    ok = ok and max(errs) < 1e-5 and max(info["loss_vs_oracle"]) < 1e-4 and step.replays >= 3
    if "grad_vs_oracle_rel_err" in info:
        ok = ok and info["grad_vs_oracle_rel_err"] <= (0.0 if world == 2 else 1e-5)
    if "momentum_vs_oracle_rel_err" in info:
        ok = ok and info["momentum_vs_oracle_rel_err"] < 1e-5
    info["ok"] = bool(ok)
    return info


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    batch = int(os.environ.get("CHECK_BATCH", "32"))
    modes = os.environ.get("CHECK_MODES", ",".join(MODES)).split(",")
    report, ok = {"world": world, "batch": batch}, True
    for mode in modes:
        info = run_mode(mode, rank, world, batch, nvls_wanted=True)
        report[mode] = info
        ok = ok and info["ok"]
        if info["nvls"] and os.environ.get("CHECK_P2P_TOO", "1") == "1":  # the same again over plain peer loads / stores
            info2 = run_mode(mode, rank, world, batch, nvls_wanted=False)
            report[mode + "_p2p"] = info2
            ok = ok and info2["ok"] and not info2["nvls"]
    _native.require().dpb_set_tuning(14, 1)
    report["ok"] = bool(ok)
    if rank == 0:
        print("MULTI_GPU_CHECK " + json.dumps(report), flush=True)
        out = os.environ.get("CHECK_OUT")
        if out:
            with open(out, "w") as fh:
                json.dump(report, fh, indent=1)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
