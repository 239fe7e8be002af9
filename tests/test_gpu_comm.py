"""NVLink collective kernels, emulated on ONE GPU: N virtual ranks = N buffers + N signal pads on the same device, the
kernels launched cooperatively with the rank in blockIdx.y (mutually-waiting kernels must not be separate launches on
one GPU).  Oracle: closed-form mean computed on the CPU; results must be bit-identical across ranks."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SIZES = [4, 1024, 4096 + 4, 65536, 262144 + 64, 2_359_296]  # floats: 16 B ... 9.4 MB (largest VGG-11 tensor)


@pytest.fixture(scope="module")
def Comm():
    from distributed_pytorch_b200 import _native
    _native.require()
    from distributed_pytorch_b200.parallel.comm import Communicator
    return Communicator


def _fill(buf, world, n, seed):
    g = torch.Generator().manual_seed(seed)
    vals = [torch.randn(n, generator=g) for _ in range(world)]
    for r in range(world):
        buf.tensors[r][:n].copy_(vals[r])
    return vals


def _mean_rank_order(vals):
    acc = torch.zeros_like(vals[0])
    for v in vals:
        acc = acc + v
    return acc * (1.0 / len(vals))


@pytest.mark.parametrize("world", [2, 4, 8, 3])
def test_allreduce_oneshot_twoshot_and_gather(Comm, world):
    comm = Comm.emulated(world)
    try:
        buf = comm.alloc(max(SIZES))
        for n in SIZES:
            for algo in (1, 2, "gather"):
                if algo == 1 and n * 4 > 128 * 512 * 4 * 16:
                    continue
                vals = _fill(buf, world, n, seed=n + world)
                for rep in range(2):  # twice: epochs advance, flags are never reset
                    if rep == 1:
                        vals = _fill(buf, world, n, seed=n + world + 77)
                    if algo == "gather":
                        comm.gather_mean_scatter(buf, 0, n, 1.0 / world, root=0)
                    else:
                        comm.all_reduce(buf, 0, n, 1.0 / world, algo=algo)
                    torch.cuda.synchronize()
                    want = _mean_rank_order(vals)
                    got0 = buf.tensors[0][:n].cpu()
                    assert torch.equal(got0, want), (world, n, algo, float((got0 - want).abs().max()))
                    for r in range(1, world):
                        assert torch.equal(buf.tensors[r][:n].cpu(), got0), (world, n, algo, r)
    finally:
        comm.close()


def test_allreduce_subrange_and_offsets(Comm):
    comm = Comm.emulated(4)
    try:
        buf = comm.alloc(4096)
        vals = _fill(buf, 4, 4096, 1)
        comm.all_reduce(buf, 1024, 2048, 0.25, algo=2)
        torch.cuda.synchronize()
        want = _mean_rank_order(vals)
        for r in range(4):
            t = buf.tensors[r].cpu()
            assert torch.equal(t[1024:3072], want[1024:3072])
            assert torch.equal(t[:1024], vals[r][:1024]) and torch.equal(t[3072:], vals[r][3072:])
    finally:
        comm.close()


@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_fused_sgd(Comm, world):
    comm = Comm.emulated(world)
    try:
        n = 64 * 1024 + 128
        grads, params, mom = comm.alloc(n), comm.alloc(n), comm.alloc(n)
        g = torch.Generator().manual_seed(3)
        p0 = torch.randn(n, generator=g)
        for r in range(world):
            params.tensors[r].copy_(p0)
        ref_p = torch.nn.Parameter(p0.clone())
        opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=1e-4)
        for it in range(3):
            vals = _fill(grads, world, n, seed=10 + it)
            ref_p.grad = _mean_rank_order(vals)
            opt.step()
            comm.all_reduce_sgd(grads, params, mom, 1.0 / world, 0.1, 0.9, 1e-4)
            torch.cuda.synchronize()
        for r in range(world):
            assert torch.allclose(params.tensors[r].cpu(), ref_p.detach(), rtol=1e-5, atol=1e-6)
            assert torch.equal(params.tensors[r].cpu(), params.tensors[0].cpu())
        # momentum is sharded: rank r holds slice r
        n4 = n // 4
        sl = (n4 + world - 1) // world * 4
        full = torch.cat([mom.tensors[r][r * sl:min((r + 1) * sl, n)].cpu() for r in range(world)])
        assert torch.allclose(full, opt.state[ref_p]["momentum_buffer"], rtol=1e-5, atol=1e-6)
    finally:
        comm.close()


def test_broadcast_and_barrier(Comm):
    comm = Comm.emulated(4)
    try:
        buf = comm.alloc(8192)
        vals = _fill(buf, 4, 8192, 5)
        comm.barrier()
        comm.broadcast(buf, 0, 8192, root=0)
        comm.broadcast(buf, 4096, 1024, root=2)
        torch.cuda.synchronize()
        want = vals[0].clone()
        for r in range(4):
            assert torch.equal(buf.tensors[r].cpu(), want)
    finally:
        comm.close()
