"""Out-of-bounds and unwritten-output checks of our own (compute-sanitizer is closed on this GPU pool, see
profiles/sanitizer_closed.txt).  Every output of a kernel lives in the middle of a larger allocation
whose head and tail are guard zones filled with a sentinel bit pattern, and the output itself starts out as the same
sentinel: after the kernel the guards must be untouched (no stray store) and no sentinel may be left inside the output
(every element written).  Scratch the kernels share (split-K workspace) gets the same guards."""
import pytest
import torch

pytestmark = pytest.mark.gpu
SENTINEL = -1.2345678e30
GUARD = 4096  # floats on either side


@pytest.fixture(scope="module")
def K():
    from distributed_pytorch_b200 import _native
    _native.require()
    from distributed_pytorch_b200.ops import kernels
    return kernels


class Guarded:
    """A tensor of `shape` (optionally channels_last) inside a sentinel-filled allocation."""

    def __init__(self, shape, channels_last=False, dtype=torch.float32):
        n = 1
        for s in shape:
            n *= s
        self.raw = torch.full((n + 2 * GUARD,), SENTINEL, device="cuda", dtype=torch.float32)
        body = self.raw[GUARD:GUARD + n]
        if channels_last:
            b, c, h, w = shape
            self.t = body.view(b, h, w, c).permute(0, 3, 1, 2)
        else:
            self.t = body.view(*shape)
        self.n = n

    def check(self, what):
        torch.cuda.synchronize()
        head, tail = self.raw[:GUARD], self.raw[GUARD + self.n:]
        assert bool((head == SENTINEL).all()) and bool((tail == SENTINEL).all()), f"{what}: store outside the output"
        body = self.raw[GUARD:GUARD + self.n]
        assert not bool((body == SENTINEL).any()), f"{what}: {int((body == SENTINEL).sum())} output elements never written"
        assert bool(torch.isfinite(body).all()), f"{what}: non-finite output"


def nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("N,H,Cin,Cout,nt,splits", [(4, 8, 256, 256, 128, 1), (4, 8, 256, 256, 128, 8), (4, 8, 256, 256, 64, 4),
                                                   (4, 8, 256, 256, 128, 3), (3, 16, 64, 128, 0, 0), (5, 2, 512, 512, 0, 0),
                                                   (32, 4, 512, 512, 0, 0)])
def test_conv_outputs_are_fully_written_and_stay_in_bounds(K, N, H, Cin, Cout, nt, splits):
    torch.manual_seed(0)
    x = nhwc(torch.randn(N, Cin, H, H, device="cuda"))
    w = nhwc(torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05)
    dy = nhwc(torch.randn(N, Cout, H, H, device="cuda"))
    b = torch.randn(Cout, device="cuda")
    y, dx, dw = Guarded((N, Cout, H, H), True), Guarded((N, Cin, H, H), True), Guarded((Cout, Cin, 3, 3), True)
    acc = torch.zeros(2 * Cout, dtype=torch.int64, device="cuda")
    K.conv3x3_gemm(0, N, H, H, Cin, Cout, x, None, w, y.t, b, stat_acc=acc, nt=nt, splits=splits)
    K.conv3x3_gemm(1, N, H, H, Cin, Cout, dy, None, w, dx.t, nt=nt, splits=splits)
    K.conv3x3_gemm(2, N, H, H, Cin, Cout, dy, x, w, dw.t, nt=nt, splits=splits)
    y.check("fprop")
    dx.check("dgrad")
    dw.check("wgrad")


def test_first_layer_outputs(K):
    torch.manual_seed(1)
    for n in (1, 5, 32):
        x = torch.randn(n, 3, 32, 32, device="cuda")
        w = nhwc(torch.randn(64, 3, 3, 3, device="cuda") * 0.2)
        y, dw = Guarded((n, 64, 32, 32), True), Guarded((64, 3, 3, 3), True)
        K.first_conv_fprop(x, w, torch.randn(64, device="cuda"), y.t, False)
        K.first_conv_wgrad(x, nhwc(torch.randn(n, 64, 32, 32, device="cuda")), dw.t)
        y.check("first conv fprop")
        dw.check("first conv wgrad")


@pytest.mark.parametrize("N,C,H,pool", [(4, 64, 32, True), (3, 128, 16, True), (5, 256, 8, False), (7, 512, 2, True)])
def test_batchnorm_kernels_stay_in_bounds(K, N, C, H, pool):
    torch.manual_seed(2)
    y = nhwc(torch.randn(N, C, H, H, device="cuda"))
    gamma, beta = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda")
    sc = K.scratch(y.device)
    acc, ticket = sc.next_acc()
    flat = y.permute(0, 2, 3, 1).reshape(-1, C).double()
    acc[:C] = (flat.sum(0) * (1 << 20)).round().long()
    acc[C:2 * C] = ((flat * flat).sum(0) * (1 << 20)).round().long()
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    out, stats = K.bn_relu_pool_fwd_train(y, acc, ticket, gamma, beta, rm, rv, None, 0.1, 1e-5, pool)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out).all()) and int(acc.abs().max()) == 0
    dg, db, dcb = Guarded((C,)), Guarded((C,)), Guarded((C,))
    dy = K.bn_relu_pool_bwd(y, nhwc(torch.randn_like(out)), stats, gamma, dg.t, db.t, dcb.t, pool)
    dg.check("dgamma")
    db.check("dbeta")
    dcb.check("conv dbias")
    assert bool(torch.isfinite(dy).all())


def test_head_and_augmentation_stay_in_bounds(K):
    torch.manual_seed(3)
    for B in (1, 9, 32, 250):
        x, w, b = torch.randn(B, 512, device="cuda"), torch.randn(10, 512, device="cuda") * 0.05, torch.randn(10, device="cuda")
        tgt = torch.randint(0, 10, (B,), device="cuda")
        loss, logits, dz = K.head_fwd(x, w, b, tgt)
        dx, dw, db = Guarded((B, 512)), Guarded((10, 512)), Guarded((10,))
        K.head_bwd(dz, x, w, torch.ones((), device="cuda"), dx.t, dw.t, db.t)
        dx.check("head dx")
        dw.check("head dW")
        db.check("head db")
        assert bool(torch.isfinite(loss)) and bool(torch.isfinite(logits).all())
    from distributed_pytorch_b200.data.datasets import synthetic_cifar10
    ds = synthetic_cifar10(True, size=64)
    images, labels = ds.to_device(torch.device("cuda"))
    idx = torch.arange(63, -1, -3, dtype=torch.int32, device="cuda")
    xb, yb = K.augment_batch(images, labels, idx, True, 0, 0)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(xb).all()) and xb.shape == (idx.numel(), 3, 32, 32) and bool((yb == labels[idx.long()]).all())
