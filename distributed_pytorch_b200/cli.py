"""Entry points behind ``main.py``, ``main_gather.py``, ``main_all_reduce.py``, ``main_ddp.py``, ``main_part3.py``.

Flags are the reference's, verbatim (``--master-ip``, ``--num-nodes``, ``--rank``; main_gather.py:100-102) and
required where the reference requires them; ``main_ddp`` reads the torchrun environment instead
(main_ddp.py:93-104); ``main`` takes none.  Rendezvous is ``tcp://<master-ip>:6585`` like the reference
(main_gather.py:107).  Reference constants (batch 256 per rank, lr 0.1, momentum 0.9, wd 1e-4, one epoch, seed 1,
sampler seed 0, 4 CPU threads) are the defaults; the extra flags below only exist because there is no network for
CIFAR-10 and because benchmarks want bounded runs.
"""
from __future__ import annotations

import argparse
import os
import sys
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from .data.datasets import cifar10
from .data.loader import BatchLoader, ShardedSampler
from .models.vgg import VGG11
from .optim.sgd import SGD
from .parallel.ddp import DistributedDataParallel
from .parallel.strategies import AllReduceSync, GatherScatterSync, NoSync
from .trainer import CrossEntropyLoss, TrainStep, test_model, train_model
from .utils.arena import adopt_module
from .utils.checkpoint import load_checkpoint, save_checkpoint

PORT = 6585
BATCH_SIZE = 256  # per rank ("batch for one node", reference main_gather.py:21)
MODES = ("single", "gather", "all_reduce", "ddp", "part3")


def build_parser(mode: str) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="Input arguments", description="gather ip, nunber of workers, rank")
    if mode in ("gather", "all_reduce", "part3"):
        p.add_argument("--master-ip", required=True)
        p.add_argument("--num-nodes", required=True, type=int)
        p.add_argument("--rank", required=True, type=int)
    x = p.add_argument_group("extensions (defaults reproduce the reference)")
    x.add_argument("--device", default="auto", choices=("auto", "cpu", "cuda"))
    x.add_argument("--master-port", type=int, default=PORT)
    x.add_argument("--batch-size", type=int, default=BATCH_SIZE, help="per-rank batch (reference: 256)")
    x.add_argument("--global-batch", type=int, default=None, help="split this batch over the ranks instead")
    x.add_argument("--epochs", type=int, default=1)
    x.add_argument("--iters", type=int, default=None, help="stop each epoch after this many iterations")
    x.add_argument("--test-batches", type=int, default=None, help="evaluate only this many test batches")
    x.add_argument("--data-root", default="./data")
    x.add_argument("--synthetic", action="store_true", help="force synthetic CIFAR-10-shaped data")
    x.add_argument("--train-size", type=int, default=None)
    x.add_argument("--test-size", type=int, default=None)
    x.add_argument("--no-graph", action="store_true", help="run the step eagerly instead of as a CUDA graph")
    x.add_argument("--fuse-optimizer", dest="fuse_optimizer", action="store_true", default=True,
                   help="all_reduce / ddp modes on CUDA: SGD update inside the collective kernel (default)")
    x.add_argument("--no-fuse-optimizer", dest="fuse_optimizer", action="store_false",
                   help="separate all-reduce and optimizer kernels")
    x.add_argument("--overlap", dest="overlap", action="store_true", default=True,
                   help="all_reduce mode on CUDA: reduce bucket by bucket on a side stream while backward is still "
                        "running (default)")
    x.add_argument("--no-overlap", dest="overlap", action="store_false",
                   help="all_reduce mode: one collective over the whole gradient arena after backward")
    x.add_argument("--host-loader", action="store_true",
                   help="augment on the host and copy batches (default on CUDA: dataset resident on the GPU, one "
                        "augmentation kernel per batch)")
    x.add_argument("--save", default=None, help="write model state_dict (+ .optim) here at the end")
    x.add_argument("--load", default=None, help="load model state_dict (+ .optim) before training")
    x.add_argument("--quiet-nonzero", action="store_true", help="only rank 0 prints (the reference prints on all)")
    return p


def _pick_device(choice: str, local_rank: int) -> torch.device:
    if choice == "cpu" or (choice == "auto" and not torch.cuda.is_available()):
        return torch.device("cpu")
    torch.cuda.set_device(local_rank % max(torch.cuda.device_count(), 1))
    return torch.device("cuda", torch.cuda.current_device())


def run(mode: str, argv: Optional[Sequence[str]] = None) -> dict:
    assert mode in MODES
    torch.manual_seed(1)
    args = build_parser(mode).parse_args(argv)
    rank, world = 0, 1
    if mode in ("gather", "all_reduce", "part3"):
        rank, world = args.rank, args.num_nodes
        device = _pick_device(args.device, int(os.environ.get("LOCAL_RANK", rank)))
        dist.init_process_group(backend="nccl" if device.type == "cuda" else "gloo",
                                init_method=f"tcp://{args.master_ip}:{args.master_port}", world_size=world, rank=rank,
                                **({"device_id": device} if device.type == "cuda" else {}))
        if mode == "part3":
            print("test")
    elif mode == "ddp":
        env_dict = {k: os.environ[k] for k in
                    ("MASTER_ADDR", "MASTER_PORT", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "LOCAL_RANK", "RANK")}
        print(f"[{os.getpid()}] Initializing process group with: {env_dict}")
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        device = _pick_device(args.device, int(os.environ["LOCAL_RANK"]))
        dist.init_process_group(backend="nccl" if device.type == "cuda" else "gloo",
                                **({"device_id": device} if device.type == "cuda" else {}))
    else:
        device = _pick_device(args.device, 0)
    if device.type == "cpu":
        torch.set_num_threads(4)
    log = print if (rank == 0 or not args.quiet_nonzero) else (lambda *a, **k: None)

    batch = args.batch_size if args.global_batch is None else max(args.global_batch // world, 1)
    train_set = cifar10(args.data_root, True, True if args.synthetic else None, args.train_size)
    test_set = cifar10(args.data_root, False, True if args.synthetic else None, args.test_size)
    sampler = None
    if mode != "single":
        sampler = ShardedSampler(len(train_set), world, rank, shuffle=True, seed=0, drop_last=False)
    loader_dev = device if (device.type == "cuda" and not args.host_loader) else None
    train_loader = BatchLoader(train_set, batch, sampler=sampler, shuffle=(sampler is None), pin_memory=True,
                               device=loader_dev)
    test_loader = BatchLoader(test_set, batch, shuffle=False, pin_memory=True, max_batches=args.test_batches,
                              device=loader_dev)

    criterion = CrossEntropyLoss()
    model = VGG11().to(device)
    sync = None
    if mode in ("ddp", "part3"):
        model = DistributedDataParallel(model)
    optimizer = SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    fuse = args.fuse_optimizer and device.type == "cuda"
    if mode in ("ddp", "part3") and fuse:
        model.fuse_optimizer(optimizer)
    if mode == "gather":
        sync = GatherScatterSync(model, world, rank)
    elif mode == "all_reduce":
        sync = AllReduceSync(model, world, rank, fuse_optimizer=fuse, optimizer=optimizer, overlap=args.overlap)
    else:
        sync = NoSync(model)
        if device.type == "cuda" and mode == "single":
            adopt_module(model)  # one flat arena => one fused SGD launch
    if args.load:
        load_checkpoint(args.load, model, optimizer)

    step = TrainStep(model, optimizer, criterion, sync, use_graph=not args.no_graph)
    result = {}
    for epoch in range(args.epochs):
        if sampler is not None and args.epochs > 1:
            sampler.set_epoch(epoch)
        train_model(model, train_loader, optimizer, criterion, epoch, args, sync=sync, step=step,
                    max_iters=args.iters, log=log)
        result["test_loss"], result["correct"] = test_model(model, test_loader, criterion, log=log)
    if args.save:
        if isinstance(sync, AllReduceSync):
            sync.gather_momentum()  # collective: every rank
        if isinstance(model, DistributedDataParallel) and model.reducer is not None:
            model.reducer.gather_momentum()
        if rank == 0:
            save_checkpoint(args.save, model, optimizer)
    if device.type == "cuda":
        torch.cuda.synchronize()
    if isinstance(model, DistributedDataParallel):
        model.close()
    if sync is not None:
        sync.close()
    if dist.is_initialized():
        dist.destroy_process_group()
    return result


def main(mode: str) -> None:
    run(mode, sys.argv[1:])
