"""tcgen05 implicit-GEMM convolution (fprop / dgrad / wgrad, all N tiles, with and without split-K) against
F.conv2d and its autograd on the CPU in fp32, at the VGG-11 layer shapes.  Inputs are pre-rounded to TF32 so the
tensor-core products are exact and the comparison is tight (a descriptor or layout bug cannot hide in tolerance)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def tf32(t):
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


@pytest.fixture(scope="module")
def K():
    from distributed_pytorch_b200 import _native
    _native.require()
    from distributed_pytorch_b200.ops import kernels
    return kernels


def _case(N, H, Cin, Cout, seed=0):
    torch.manual_seed(seed)
    x = tf32(torch.randn(N, Cin, H, H))
    w = tf32(torch.randn(Cout, Cin, 3, 3) * 0.1).requires_grad_(True)
    b = torch.randn(Cout)
    xr = x.clone().requires_grad_(True)
    y = F.conv2d(xr, w, b, padding=1)
    dy = tf32(torch.randn_like(y))
    y.backward(dy)
    return x, w, b, y.detach(), dy, xr.grad, w.grad


def _check(a, ref, what, rtol=2e-4):
    err = float((a - ref).abs().max())
    scale = float(ref.abs().max())
    assert err <= rtol * scale + 1e-6, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


SHAPES = [  # (N, H, Cin, Cout): VGG-11 layers 4, 8, 11, 15, 18, 22 at small batch, plus ragged batches
    (2, 16, 64, 128), (4, 8, 128, 256), (2, 8, 256, 256), (8, 4, 256, 512), (8, 4, 512, 512), (32, 2, 512, 512),
    (3, 16, 64, 128), (5, 2, 512, 512), (1, 32, 64, 64),
]


@pytest.mark.parametrize("N,H,Cin,Cout", SHAPES)
def test_conv_all_modes_default_heuristics(K, N, H, Cin, Cout):
    x, w, b, y_ref, dy, dx_ref, dw_ref = _case(N, H, Cin, Cout)
    xd, wd, bd, dyd = nhwc(x.cuda()), nhwc(w.detach().cuda()), b.cuda(), nhwc(dy.cuda())
    y = K.nhwc_empty(N, Cout, H, H, "cuda")
    ps, pq, tiles = K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, y, bd, want_stats=True)
    torch.cuda.synchronize()
    _check(y.cpu(), y_ref, "fprop")
    _check(ps[:tiles].sum(0).cpu(), y_ref.sum((0, 2, 3)), "stat sum", 1e-3)
    _check(pq[:tiles].sum(0).cpu(), (y_ref ** 2).sum((0, 2, 3)), "stat sumsq", 1e-3)
    dx = K.nhwc_empty(N, Cin, H, H, "cuda")
    K.conv3x3_gemm(1, N, H, H, Cin, Cout, dyd, None, wd, dx)
    torch.cuda.synchronize()
    _check(dx.cpu(), dx_ref, "dgrad")
    dw = torch.empty_like(wd)
    K.conv3x3_gemm(2, N, H, H, Cin, Cout, dyd, xd, wd, dw)
    torch.cuda.synchronize()
    _check(dw.cpu(), dw_ref, "wgrad")


@pytest.fixture
def pair_reduce():
    """Two-way splits finished inside the GEMM kernel by whichever CTA arrives second (dpb_set_tuning key 8)."""
    from distributed_pytorch_b200 import _native
    lib = _native.require()
    lib.dpb_set_tuning(8, 1)
    yield
    lib.dpb_set_tuning(8, 0)


@pytest.mark.parametrize("nt", [64, 128, 256])
def test_conv_two_way_split_in_kernel(K, pair_reduce, nt):
    test_conv_tiles_and_split_k(K, nt, 2)


@pytest.fixture
def persistent():
    """All three GEMMs through the persistent kernel (one CTA per SM, TMEM double-buffered; dpb_set_tuning key 10)."""
    from distributed_pytorch_b200 import _native
    lib = _native.require()
    lib.dpb_set_tuning(10, 7)
    yield
    lib.dpb_set_tuning(10, 0)


@pytest.mark.parametrize("nt,splits", [(64, 1), (128, 1), (256, 1), (128, 3), (64, 5), (128, 0)])
def test_conv_persistent_tiles_and_split_k(K, persistent, nt, splits):
    test_conv_tiles_and_split_k(K, nt, splits)


@pytest.mark.parametrize("N,H,Cin,Cout", [(32, 16, 64, 128), (16, 8, 128, 256), (64, 4, 256, 512), (256, 2, 512, 512)])
def test_conv_persistent_layer_shapes(K, persistent, N, H, Cin, Cout):
    """VGG layer shapes: several work units per CTA, so the accumulator ping-pong and the cross-unit stage ring run."""
    test_conv_all_modes_default_heuristics(K, N, H, Cin, Cout)


@pytest.mark.parametrize("nt,splits", [(64, 1), (128, 1), (256, 1), (128, 3), (64, 5), (256, 2)])
def test_conv_tiles_and_split_k(K, nt, splits):
    N, H, Cin, Cout = 4, 8, 256, 256
    x, w, b, y_ref, dy, dx_ref, dw_ref = _case(N, H, Cin, Cout, seed=1)
    xd, wd, bd, dyd = nhwc(x.cuda()), nhwc(w.detach().cuda()), b.cuda(), nhwc(dy.cuda())
    for rep in range(2):  # twice: split-K counters must re-arm themselves
        y = K.nhwc_empty(N, Cout, H, H, "cuda")
        ps, pq, tiles = K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, y, bd, want_stats=True, nt=nt,
                                       splits=splits)
        dx = K.nhwc_empty(N, Cin, H, H, "cuda")
        K.conv3x3_gemm(1, N, H, H, Cin, Cout, dyd, None, wd, dx, nt=nt, splits=splits)
        dw = torch.empty_like(wd)
        K.conv3x3_gemm(2, N, H, H, Cin, Cout, dyd, xd, wd, dw, nt=nt, splits=splits)
        torch.cuda.synchronize()
        _check(y.cpu(), y_ref, f"fprop nt={nt} splits={splits} rep={rep}")
        _check(ps[:tiles].sum(0).cpu(), y_ref.sum((0, 2, 3)), "stat sum", 1e-3)
        _check(dx.cpu(), dx_ref, f"dgrad nt={nt} splits={splits}")
        _check(dw.cpu(), dw_ref, f"wgrad nt={nt} splits={splits}")
