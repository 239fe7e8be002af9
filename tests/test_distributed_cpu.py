"""Two-rank gloo runs of the three sync modes on CPU: the implied correctness contract of SURVEY.md §4.2 --
parameters bit-identical across ranks after every step, synced gradient = mean of local gradients, DDP repairs
diverged initial weights by its rank-0 broadcast, our DDP == torch DDP, `module.` key prefix."""
import os
import socket
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _run(mode, outdir, world=2, iters=3):
    port = _free_port()
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_dist_worker.py"), mode, str(r), str(world),
                               str(port), str(outdir), str(iters)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(world)]
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    return [torch.load(os.path.join(outdir, f"{mode}_{r}.pt")) for r in range(world)]


@pytest.mark.timeout(900)
def test_gather_allreduce_ddp_contract(tmp_path):
    res = {m: _run(m, tmp_path) for m in ("gather", "all_reduce", "ddp", "torch_ddp")}
    for mode, (a, b) in res.items():
        for k in a:
            if "running_" in k or "num_batches" in k:
                continue
            assert torch.equal(a[k], b[k]), f"{mode}: {k} differs between ranks"
    # at world 2 the modes agree with each other bit for bit (mean of two numbers is exact in any order)
    for k in res["gather"][0]:
        if "running_" in k or "num_batches" in k:
            continue
        assert torch.equal(res["gather"][0][k], res["all_reduce"][0][k]), k
        assert torch.equal(res["ddp"][0]["module." + k], res["torch_ddp"][0]["module." + k]), k
        assert torch.equal(res["ddp"][0]["module." + k], res["all_reduce"][0][k]), k
    assert all(k.startswith("module.") for k in res["ddp"][0])
    # BatchNorm statistics stay per rank in gather / all_reduce modes (no SyncBatchNorm in the reference)
    assert not torch.equal(res["gather"][0]["layers.1.running_mean"], res["gather"][1]["layers.1.running_mean"])
