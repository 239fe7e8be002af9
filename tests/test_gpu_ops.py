"""Memory-bound sm_100a kernels against plain PyTorch fp32 CPU references of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from distributed_pytorch_b200 import _native
    _native.require()
    from distributed_pytorch_b200.ops import kernels
    return kernels


def nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


def test_sgd_kernel_matches_torch(K):
    torch.manual_seed(0)
    for n in (10, 4096 + 3, 1 << 20):
        p = torch.randn(n)
        ref = torch.nn.Parameter(p.clone())
        opt = torch.optim.SGD([ref], lr=0.1, momentum=0.9, weight_decay=1e-4)
        pd, md = p.cuda(), torch.zeros(n).cuda()
        for _ in range(3):
            g = torch.randn(n)
            ref.grad = g.clone()
            opt.step()
            K.sgd_step(pd, g.cuda(), md, n, 0.1, 0.9, 1e-4)
        assert torch.allclose(pd.cpu(), ref.detach(), rtol=1e-5, atol=1e-6)
        assert torch.allclose(md.cpu(), opt.state[ref]["momentum_buffer"], rtol=1e-5, atol=1e-6)


def test_linear_and_cross_entropy(K):
    torch.manual_seed(1)
    for B in (1, 32, 80, 256):
        x = torch.randn(B, 512, requires_grad=True)
        w = torch.randn(10, 512, requires_grad=True) * 0.05
        w.retain_grad()
        b = torch.randn(10, requires_grad=True)
        y = torch.randint(0, 10, (B,))
        logits = F.linear(x, w, b)
        loss = F.cross_entropy(logits, y)
        loss.backward()
        xd, wd, bd, yd = x.detach().cuda(), w.detach().cuda(), b.detach().cuda(), y.cuda()
        out = K.linear_fwd(xd, wd, bd)
        assert torch.allclose(out.cpu(), logits.detach(), rtol=1e-4, atol=1e-4)
        metrics = torch.zeros(2, device="cuda")
        l = K.cross_entropy_fwd(out, yd, metrics)
        assert torch.allclose(l.cpu(), loss.detach(), rtol=1e-5, atol=1e-5)
        assert int(metrics[1].item()) == int((logits.argmax(1) == y).sum())
        dz = K.cross_entropy_bwd(out, yd, torch.ones((), device="cuda"))
        dx, dw, db = torch.empty_like(xd), torch.empty_like(wd), torch.empty_like(bd)
        K.linear_bwd(dz, xd, wd, dx, dw, db)
        assert torch.allclose(dx.cpu(), x.grad, rtol=1e-4, atol=1e-6)
        assert torch.allclose(dw.cpu(), w.grad, rtol=1e-4, atol=1e-5)
        assert torch.allclose(db.cpu(), b.grad, rtol=1e-4, atol=1e-6)


def test_fused_head_forward_backward(K):
    """Linear + cross-entropy as one forward and one backward launch against the CPU oracle (loss, logits, arg-max hits,
    dx / dW / db including a non-unit upstream gradient); twice per size so the ticket counter must have re-armed."""
    torch.manual_seed(5)
    for B in (1, 7, 32, 80, 256):
        for g in (1.0, 0.37):
            x = torch.randn(B, 512, requires_grad=True)
            w = (torch.randn(10, 512) * 0.05).requires_grad_(True)
            b = torch.randn(10, requires_grad=True)
            y = torch.randint(0, 10, (B,))
            logits = F.linear(x, w, b)
            loss = F.cross_entropy(logits, y)
            (loss * g).backward()
            xd, wd, bd, yd = x.detach().cuda(), w.detach().cuda(), b.detach().cuda(), y.cuda()
            metrics = torch.zeros(2, device="cuda")
            l, lg, dz = K.head_fwd(xd, wd, bd, yd, metrics)
            assert torch.allclose(lg.cpu(), logits.detach(), rtol=1e-4, atol=1e-4)
            assert torch.allclose(l.cpu(), loss.detach(), rtol=1e-5, atol=1e-5)
            assert int(metrics[1].item()) == int((logits.argmax(1) == y).sum())
            assert torch.allclose(metrics[0].cpu(), loss.detach(), rtol=1e-5, atol=1e-5)
            dx, dw, db = torch.empty_like(xd), torch.empty_like(wd), torch.empty_like(bd)
            K.head_bwd(dz, xd, wd, torch.full((), g, device="cuda"), dx, dw, db)
            assert torch.allclose(dx.cpu(), x.grad, rtol=1e-4, atol=1e-6)
            assert torch.allclose(dw.cpu(), w.grad, rtol=1e-4, atol=1e-5)
            assert torch.allclose(db.cpu(), b.grad, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("N,C,H,pool", [(4, 64, 32, True), (8, 128, 16, True), (32, 512, 2, True), (8, 512, 4, False)])
def test_bn_backward_separate_finalize_kernel(K, N, C, H, pool):
    """Same oracle through the three-launch path (partial rows, finalize kernel, apply) that callers without
    accumulators get."""
    from distributed_pytorch_b200 import _native
    lib = _native.require()
    assert lib.dpb_bn_bwd_launches(C, 0) == 3 and lib.dpb_bn_bwd_launches(C, 1) == 2
    test_bn_relu_pool_forward_backward(K, N, C, H, pool, use_acc=False)


@pytest.mark.parametrize("train", [True, False])
def test_device_augmentation_matches_host_transform(K, train):
    """The augmentation kernel (gather + RandomCrop(pad 4) + flip + Normalize) against the host transform of
    data/datasets.py for the same indices: hashed draws, and explicit draws covering every corner of the crop range."""
    import numpy as np
    from distributed_pytorch_b200.data.datasets import augment_draws, synthetic_cifar10
    ds = synthetic_cifar10(train, size=300, seed=3)
    ds.set_epoch(2)
    idx = np.array([5, 17, 0, 299, 123, 42, 42, 7, 250, 11, 1, 2, 3], dtype=np.int64)
    images, labels = ds.to_device(torch.device("cuda"))
    idx_dev = torch.from_numpy(idx.astype(np.int32)).cuda()
    x, y = K.augment_batch(images, labels, idx_dev, train, ds.seed, ds.epoch)
    xr, yr = ds.get_batch(idx)
    assert torch.equal(y.cpu(), yr)
    assert torch.allclose(x.cpu(), xr, rtol=0, atol=1e-6)
    if train:
        draws = augment_draws(idx, ds.seed, ds.epoch)
        assert len({tuple(d) for d in draws.tolist()}) > 3  # the hash does vary
        corner = np.array([[0, 0, 0], [8, 8, 1], [0, 8, 1], [8, 0, 0], [4, 4, 1]] * 3, dtype=np.int32)[:len(idx)]
        x2, _ = K.augment_batch(images, labels, idx_dev, True, 0, 0, torch.from_numpy(corner).cuda())
        xr2, _ = ds.get_batch(idx, corner)
        assert torch.allclose(x2.cpu(), xr2, rtol=0, atol=1e-6)


@pytest.mark.parametrize("N,C,H,pool", [(4, 64, 32, True), (3, 128, 16, True), (5, 256, 8, False), (32, 512, 2, True),
                                        (7, 512, 4, False)])
def test_bn_relu_pool_forward_backward(K, N, C, H, pool, use_acc=True):
    torch.manual_seed(2)
    y = torch.randn(N, C, H, H) * 1.5 + 0.3
    gamma = torch.rand(C) + 0.5
    beta = torch.randn(C) * 0.2
    rm, rv = torch.zeros(C), torch.ones(C)
    yr = y.clone().requires_grad_(True)
    g_, b_ = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    z = F.relu(F.batch_norm(yr, rm, rv, g_, b_, True, 0.1, 1e-5))
    if pool:
        z = F.max_pool2d(z, 2, 2)
    dout = torch.randn_like(z)
    z.backward(dout)
    # device: stats as the conv epilogue would leave them (per 128-pixel tile partials)
    yd = nhwc(y.cuda())
    flat = yd.permute(0, 2, 3, 1).reshape(-1, C)
    tiles = (flat.shape[0] + 127) // 128
    pad = torch.zeros(tiles * 128, C, device="cuda")
    pad[: flat.shape[0]] = flat
    ps = pad.view(tiles, 128, C).sum(1).contiguous()
    pq = (pad * pad).view(tiles, 128, C).sum(1).contiguous()
    rmd, rvd = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    stats = K.bn_finalize(ps, pq, tiles, C, float(N * H * H), gamma.cuda(), beta.cuda(), rmd, rvd, nbt, 0.1, 1e-5)
    out = K.bn_relu_pool_fwd(yd, stats[0], stats[1], pool)
    assert out.shape == z.shape
    assert torch.allclose(out.cpu(), z.detach(), rtol=1e-4, atol=1e-4)
    assert torch.allclose(rmd.cpu(), rm, rtol=1e-4, atol=1e-5) and torch.allclose(rvd.cpu(), rv, rtol=1e-4, atol=1e-5)
    assert int(nbt.item()) == 1
    dg, db, dcb = torch.empty(C, device="cuda"), torch.empty(C, device="cuda"), torch.ones(C, device="cuda")
    dy = K.bn_relu_pool_bwd(yd, nhwc(dout.cuda()), stats, gamma.cuda(), dg, db, dcb, pool, use_acc=use_acc)
    if use_acc:  # and once more: the apply pass must have cleared the accumulator slot it used
        sc = K.scratch(yd.device)
        torch.cuda.synchronize()
        assert int(sc.bn_acc.abs().max().item()) == 0 and int(sc.bn_tickets.abs().max().item()) == 0
    scale = max(1.0, float(yr.grad.abs().max()))
    assert torch.allclose(dy.cpu(), yr.grad, rtol=2e-3, atol=2e-4 * scale)
    assert torch.allclose(dg.cpu(), g_.grad, rtol=2e-3, atol=2e-3)
    assert torch.allclose(db.cpu(), b_.grad, rtol=2e-3, atol=2e-3)
    assert float(dcb.abs().max()) == 0.0


def test_bn_eval_mode(K):
    torch.manual_seed(3)
    C = 128
    y = torch.randn(2, C, 8, 8)
    gamma, beta, rm, rv = torch.rand(C) + 0.5, torch.randn(C), torch.randn(C) * 0.1, torch.rand(C) + 0.5
    ref = F.max_pool2d(F.relu(F.batch_norm(y, rm, rv, gamma, beta, False, 0.1, 1e-5)), 2, 2)
    st = K.bn_eval_scale_shift(C, gamma.cuda(), beta.cuda(), rm.cuda(), rv.cuda(), 1e-5)
    out = K.bn_relu_pool_fwd(nhwc(y.cuda()), st[0], st[1], True)
    assert torch.allclose(out.cpu(), ref, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("tensor_cores", [0, 1])
@pytest.mark.parametrize("N,HW", [(1, 32), (8, 32), (33, 32), (5, 16), (3, 64), (4, 30)])
def test_first_conv_fprop_and_wgrad(K, N, HW, tensor_cores):
    """Cin = 3 entry layer vs the fp32 oracle: the FP32-pipe kernels to fp32 accuracy, the tcgen05 kernels (TF32
    operands, fp32 accumulate; 30 x 30 images do not tile into 128 pixels and take the FP32-pipe path) to TF32 accuracy."""
    from distributed_pytorch_b200 import _native
    torch.manual_seed(4)
    x = torch.randn(N, 3, HW, HW)
    w = (torch.randn(64, 3, 3, 3) * 0.2).requires_grad_(True)
    b = torch.randn(64)
    ref = F.conv2d(x, w, b, padding=1)
    dy = torch.randn_like(ref)
    ref.backward(dy)
    wd = nhwc(w.detach().cuda())
    y = torch.empty((N, 64, HW, HW), device="cuda").contiguous(memory_format=torch.channels_last)
    lib = _native.require()
    lib.dpb_set_tuning(7, tensor_cores)
    try:
        ps, pq, parts = K.first_conv_fprop(x.cuda(), wd, b.cuda(), y, True)
        if K.first_conv_on_tensor_cores(HW, HW):  # the same statistics through the fixed-point accumulators
            acc = torch.zeros(128, dtype=torch.int64, device="cuda")
            K.first_conv_fprop(x.cuda(), wd, b.cuda(), torch.empty_like(y), False, stat_acc=acc)
            ssum, ssq = K.acc_to_float(acc, 64)
            assert torch.allclose(ssum.float(), ps.sum(0), rtol=1e-4, atol=1e-2)
            assert torch.allclose(ssq.float(), pq.sum(0), rtol=1e-4, atol=1e-2)
        dw = torch.empty_like(wd)
        K.first_conv_wgrad(x.cuda(), nhwc(dy.cuda()), dw)
        torch.cuda.synchronize()
    finally:
        lib.dpb_set_tuning(7, 1)
    tf32 = tensor_cores and HW != 30
    tol = 2e-3 if tf32 else 1e-4
    refd = ref.detach()
    assert float((y.cpu() - refd).abs().max()) <= tol * max(1.0, float(refd.abs().max()))
    cnt = N * HW * HW
    assert torch.allclose(ps.sum(0).cpu(), refd.sum((0, 2, 3)), rtol=1e-3, atol=1e-2 + (2e-4 * cnt if tf32 else 0))
    assert torch.allclose(pq.sum(0).cpu(), (refd ** 2).sum((0, 2, 3)), rtol=2e-3 if tf32 else 1e-3, atol=1e-2)
    assert torch.allclose(dw.cpu(), w.grad, rtol=1e-3, atol=(3e-3 if tf32 else 1e-3) * float(w.grad.abs().max()))
