"""The allocation-free entry point ``dpb_bn_relu_pool_bwd_noalloc``.

Who needs it: a caller that cannot get device memory of its own -- a plain C host program that uses nothing but this
library's C ABI, or a component that is handed its input and output buffers by a framework and may not call
``cudaMalloc`` (or any framework allocator) on its hot path.  The ordinary entry point makes such a caller supply the
scratch for the BatchNorm-backward partial sums; this variant keeps that scratch inside the library (a fixed
``__device__`` array, about 1.2 MB per GPU, enough for 512 channels).

What is checked, straight through the C ABI with ctypes: on calls in a row (different inputs, each finds the previous
call's leftovers in the shared scratch) the results are bit-equal to those of ``dpb_bn_relu_pool_bwd`` given caller
scratch -- both run the same deterministic kernels on the same grid, so any difference would mean a wrong scratch
layout; one case is wide and long enough (512 channels, all 296 reduce blocks) to use the whole array; on a box with
several GPUs the check is repeated on a second device (the array's address is looked up per device); a channel count
beyond the limit is refused with a readable message.  This file is collected last on purpose, so a
failure here cannot hide a result of any other test in a stop-at-first-failure run; it does not depend on any other
test having run (or not run) before it in the same process.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(n, c, h, pool, seed, dev="cuda"):
    g = torch.Generator().manual_seed(seed)
    y = (torch.randn(n, c, h, h, generator=g) * 1.3 + 0.2).to(dev).contiguous(memory_format=torch.channels_last)
    ho = h // 2 if pool else h
    dout = torch.randn(n, c, ho, ho, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
    gamma = (torch.rand(c, generator=g) + 0.5).to(dev)
    beta = (torch.randn(c, generator=g) * 0.1).to(dev)
    flat = y.permute(0, 2, 3, 1).reshape(-1, c)
    mean = flat.mean(0)
    var = flat.var(0, unbiased=False)
    invstd = torch.rsqrt(var + 1e-5)
    scale = gamma * invstd
    shift = beta - mean * scale
    return y, dout, gamma, scale.contiguous(), shift.contiguous(), mean.contiguous(), invstd.contiguous()


def _run(lib, noalloc, y, dout, gamma, scale, shift, mean, invstd, pool):
    from distributed_pytorch_b200 import _native as N
    n, c, h, w = y.shape
    dev = y.device
    dg, db, dcb = (torch.full((c,), 7.0, device=dev) for _ in range(3))
    dy = torch.full_like(y, float("nan"))
    s = torch.cuda.current_stream(dev).cuda_stream
    if noalloc:
        rc = lib.dpb_bn_relu_pool_bwd_noalloc(y.data_ptr(), dout.data_ptr(), scale.data_ptr(), shift.data_ptr(),
                                              mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), dg.data_ptr(),
                                              db.data_ptr(), dcb.data_ptr(), dy.data_ptr(), n, h, w, c, int(pool), s)
    else:
        part = torch.empty(2 * lib.dpb_bn_bwd_blocks() * c, device=dev)
        coef = torch.empty(3 * c, device=dev)
        # no accumulators: the three separate launches (reduce, finalize, apply), which is what the no-alloc variant runs
        rc = lib.dpb_bn_relu_pool_bwd(y.data_ptr(), dout.data_ptr(), scale.data_ptr(), shift.data_ptr(),
                                      mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), part.data_ptr(),
                                      coef.data_ptr(), None, None, dg.data_ptr(), db.data_ptr(), dcb.data_ptr(),
                                      dy.data_ptr(), n, h, w, c, int(pool), s)
    N.check(rc, "bn_relu_pool_bwd" + ("_noalloc" if noalloc else ""))
    torch.cuda.synchronize(dev)
    return dy, dg, db, dcb


CASES = [(8, 512, 4, True), (5, 128, 16, False), (40, 512, 4, False)]  # the last: 640 pixels x 512 channels = 296 blocks


def _compare_on(dev):
    from distributed_pytorch_b200 import _native as N
    lib = N.require()
    with torch.cuda.device(dev):
        for seed, (n, c, h, pool) in enumerate(CASES):
            args = _inputs(n, c, h, pool, seed, dev)
            want = _run(lib, False, *args, pool)
            got = _run(lib, True, *args, pool)
            for a, b, name in zip(got, want, ("dy", "dgamma", "dbeta", "conv_dbias")):
                assert torch.isfinite(a).all(), (str(dev), name)
                assert torch.equal(a, b), (str(dev), (n, c, h, pool), name, float((a - b).abs().max()))


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    from distributed_pytorch_b200 import _native as N
    lib = N.require()
    assert lib.dpb_noalloc_max_channels() == 512
    # back-to-back calls with different shapes: each one reuses the scratch the previous one dirtied
    _compare_on(torch.device("cuda", torch.cuda.current_device()))


def test_noalloc_on_a_second_device():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs a second GPU")
    _compare_on(torch.device("cuda", 1))


def test_noalloc_rejects_sizes_beyond_its_scratch():
    from distributed_pytorch_b200 import _native as N
    lib = N.require()
    c = 1024
    t = torch.zeros(4 * c, device="cuda")
    rc = lib.dpb_bn_relu_pool_bwd_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(),
                                          t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), None, t.data_ptr(),
                                          1, 2, 2, c, 0, torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.dpb_last_error().decode()
    assert "1024" in msg and "512" in msg
    torch.cuda.synchronize()
