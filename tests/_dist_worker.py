"""Worker for the multi-process CPU tests: python _dist_worker.py <mode> <rank> <world> <port> <outdir> <iters>."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402


def main():
    mode, rank, world, port, outdir, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5], int(sys.argv[6])
    torch.set_num_threads(2)
    torch.manual_seed(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", world_size=world, rank=rank)
    model = dpb.VGG11()
    if mode == "ddp" and rank != 0:
        with torch.no_grad():  # DDP must repair this through its rank-0 broadcast
            for p in model.parameters():
                p.add_(0.5)
    if mode == "ddp":
        model = dpb.DistributedDataParallel(model)
    elif mode == "torch_ddp":
        model = torch.nn.parallel.DistributedDataParallel(model)
    opt = dpb.SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    sync = {"gather": dpb.GatherScatterSync, "all_reduce": dpb.AllReduceSync}.get(mode, lambda m, w, r: dpb.NoSync(m))(model, world, rank)
    crit = dpb.CrossEntropyLoss()
    g = torch.Generator().manual_seed(100 + rank)
    local_grads = None
    for it in range(iters):
        x = torch.randn(4, 3, 32, 32, generator=g)
        y = torch.randint(0, 10, (4,), generator=g)
        opt.zero_grad()
        crit(model(x), y).backward()
        if it == 0 and mode in ("gather", "all_reduce"):
            local_grads = [p.grad.clone() for p in model.parameters()]
        sync.sync()
        if it == 0 and local_grads is not None:
            # synced gradient == mean over ranks of the local gradients
            for p, lg in zip(model.parameters(), local_grads):
                tot = lg.clone()
                dist.all_reduce(tot)
                assert torch.allclose(p.grad, tot / world, rtol=1e-6, atol=1e-7)
        opt.step()
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    torch.save(sd, os.path.join(outdir, f"{mode}_{rank}.pt"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
