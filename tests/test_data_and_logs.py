"""Sampler parity with DistributedSampler, loader, reference log-line protocol, CLI flags."""
import re

import pytest
import torch

import distributed_pytorch_b200 as dpb
from distributed_pytorch_b200.cli import build_parser
from distributed_pytorch_b200.data.datasets import synthetic_cifar10
from distributed_pytorch_b200.data.loader import BatchLoader, ShardedSampler


@pytest.mark.parametrize("n,world", [(50000, 4), (1003, 8), (17, 2), (5, 8)])
def test_sampler_matches_distributed_sampler(n, world):
    from torch.utils.data.distributed import DistributedSampler
    ds = list(range(n))
    for rank in range(world):
        ref = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=True, seed=0, drop_last=False)
        ours = ShardedSampler(n, world, rank, shuffle=True, seed=0, drop_last=False)
        assert list(ref) == list(ours)
        assert len(ref) == len(ours)
        ref.set_epoch(3)
        ours.set_epoch(3)
        assert list(ref) == list(ours)


def test_loader_shapes_and_iteration_count():
    ds = synthetic_cifar10(True, size=600)
    sampler = ShardedSampler(len(ds), 4, 1)
    loader = BatchLoader(ds, 64, sampler=sampler, pin_memory=False)
    batches = list(loader)
    assert len(batches) == len(loader) == 3  # ceil(150/64)
    x, y = batches[0]
    assert x.shape == (64, 3, 32, 32) and x.dtype == torch.float32 and y.dtype == torch.int64
    assert batches[-1][0].shape[0] == 150 - 128
    test = BatchLoader(synthetic_cifar10(False, size=100), 64, pin_memory=False)
    assert test.num_examples == 100 and len(list(test)) == 2


def test_reference_log_lines():
    torch.manual_seed(1)
    model = dpb.VGG11()
    opt = dpb.SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    ds = synthetic_cifar10(True, size=4 * 41)
    loader = BatchLoader(ds, 4, pin_memory=False, prefetch=0)
    lines = []
    dpb.train_model(model, loader, opt, dpb.CrossEntropyLoss(), 0, log=lines.append)
    assert re.fullmatch(r"Epoch: 1, Iteration: 1-20, Average Loss: \d+\.\d{3}", lines[0])
    assert re.fullmatch(r"Epoch: 1, Iteration: 21-40, Average Loss: \d+\.\d{3}", lines[1])
    assert re.fullmatch(r"Avg Time for iteration 2-40: [\d.e-]+ seconds\.", lines[2])
    test = BatchLoader(synthetic_cifar10(False, size=10), 4, pin_memory=False)
    out = []
    loss, correct = dpb.test_model(model, test, dpb.CrossEntropyLoss(), log=out.append)
    assert re.fullmatch(r"Test set: Average loss: \d+\.\d{4}, Accuracy: \d+/10 \(\d+%\)\n", out[0])


def test_cli_flags_are_the_reference_ones():
    for mode in ("gather", "all_reduce", "part3"):
        p = build_parser(mode)
        a = p.parse_args(["--master-ip", "172.18.0.2", "--num-nodes", "4", "--rank", "2"])
        assert (a.master_ip, a.num_nodes, a.rank) == ("172.18.0.2", 4, 2)
        assert a.batch_size == 256 and a.master_port == 6585 and a.epochs == 1
        with pytest.raises(SystemExit):
            p.parse_args(["--master-ip", "x"])
    assert build_parser("single").parse_args([]).batch_size == 256
    assert build_parser("ddp").parse_args([]).global_batch is None
