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
    acc = torch.zeros(2 * Cout, dtype=torch.int64, device="cuda")  # fixed-point BatchNorm accumulators
    K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, y, bd, stat_acc=acc)
    torch.cuda.synchronize()
    _check(y.cpu(), y_ref, "fprop")
    ssum, ssq = K.acc_to_float(acc, Cout)
    _check(ssum.float().cpu(), y_ref.sum((0, 2, 3)), "stat sum", 1e-3)
    _check(ssq.float().cpu(), (y_ref ** 2).sum((0, 2, 3)), "stat sumsq", 1e-3)
    dx = K.nhwc_empty(N, Cin, H, H, "cuda")
    K.conv3x3_gemm(1, N, H, H, Cin, Cout, dyd, None, wd, dx)
    torch.cuda.synchronize()
    _check(dx.cpu(), dx_ref, "dgrad")
    dw = torch.empty_like(wd)
    K.conv3x3_gemm(2, N, H, H, Cin, Cout, dyd, xd, wd, dw)
    torch.cuda.synchronize()
    _check(dw.cpu(), dw_ref, "wgrad")


@pytest.fixture
def workspace_split_k():
    """Every split-K GEMM through the global workspace + reduce kernel (dpb_set_tuning key 5 = 1: no cluster reduce)."""
    from distributed_pytorch_b200 import _native
    lib = _native.require()
    lib.dpb_set_tuning(5, 1)
    yield
    lib.dpb_set_tuning(5, 8)


@pytest.mark.parametrize("N,H,Cin,Cout", SHAPES)
def test_conv_all_modes_workspace_path(K, workspace_split_k, N, H, Cin, Cout):
    test_conv_all_modes_default_heuristics(K, N, H, Cin, Cout)


@pytest.fixture(params=[8, 16])
def cluster_max(request):
    """Largest cluster the split-K heuristic may use (dpb_set_tuning key 5): the default and the non-portable 16."""
    from distributed_pytorch_b200 import _native
    lib = _native.require()
    lib.dpb_set_tuning(5, request.param)
    yield request.param
    lib.dpb_set_tuning(5, 8)


# layer shapes whose heuristic split factors exercise 2-, 4-, 8- and 16-CTA clusters (DSMEM reduce)
@pytest.mark.parametrize("N,H,Cin,Cout", [(32, 4, 512, 512), (32, 2, 512, 512), (256, 4, 256, 512), (32, 8, 256, 256),
                                          (64, 16, 64, 128), (32, 16, 64, 128)])
def test_conv_cluster_reduce_layer_shapes(K, cluster_max, N, H, Cin, Cout):
    test_conv_all_modes_default_heuristics(K, N, H, Cin, Cout)


@pytest.mark.parametrize("N,H,Cin,Cout,splits", [(2, 16, 64, 128, 0), (8, 4, 256, 512, 0), (32, 2, 512, 512, 0),
                                                 (4, 8, 128, 256, 3), (4, 8, 128, 256, 1), (64, 16, 64, 128, 0)])
def test_conv_then_batchnorm_from_accumulators(K, N, H, Cin, Cout, splits):
    """Conv (statistics into the fixed-point accumulators: cluster, workspace and unsplit paths) followed by the
    training-mode BatchNorm + ReLU kernel that derives scale / shift from them: output, saved statistics and running
    statistics against F.batch_norm on the CPU; twice on the same accumulator slot, which the consumer must have cleared."""
    x, w, b, y_ref, _, _, _ = _case(N, H, Cin, Cout, seed=3)
    gamma, beta = torch.rand(Cout) + 0.5, torch.randn(Cout) * 0.2
    rm, rv = torch.zeros(Cout), torch.ones(Cout)
    for _ in range(2):
        z_ref = F.relu(F.batch_norm(y_ref, rm, rv, gamma, beta, True, 0.1, 1e-5))
    mean = y_ref.mean((0, 2, 3))
    var = y_ref.var((0, 2, 3), unbiased=False)
    invstd = 1.0 / torch.sqrt(var + 1e-5)
    xd, wd, bd = nhwc(x.cuda()), nhwc(w.detach().cuda()), b.cuda()
    rmd, rvd = torch.zeros(Cout, device="cuda"), torch.ones(Cout, device="cuda")
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    acc = torch.zeros(2 * Cout, dtype=torch.int64, device="cuda")
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
    for rep in range(2):
        y = K.nhwc_empty(N, Cout, H, H, "cuda")
        K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, y, bd, stat_acc=acc, splits=splits)
        out, stats = K.bn_relu_pool_fwd_train(y, acc, ticket, gamma.cuda(), beta.cuda(), rmd, rvd, nbt, 0.1, 1e-5, False)
        torch.cuda.synchronize()
        _check(y.cpu(), y_ref, f"fprop rep={rep}")
        st = stats.cpu()
        assert torch.allclose(st[2], mean, rtol=1e-3, atol=1e-4), f"mean rep={rep}"
        assert torch.allclose(st[3], invstd, rtol=1e-3, atol=1e-5), f"invstd rep={rep}"
        assert torch.allclose(st[0], gamma * invstd, rtol=1e-3, atol=1e-5)
        assert torch.allclose(st[1], beta - mean * gamma * invstd, rtol=1e-3, atol=1e-4)
        assert torch.allclose(out.cpu(), z_ref, rtol=1e-3, atol=2e-3), f"bn+relu rep={rep}"
        assert int(acc.abs().max().item()) == 0 and int(ticket.item()) == 0, "consumer must clear the accumulators"
    assert int(nbt.item()) == 2
    assert torch.allclose(rmd.cpu(), rm, rtol=1e-3, atol=1e-5) and torch.allclose(rvd.cpu(), rv, rtol=1e-3, atol=1e-5)


# splits 1: none; 3, 5 and anything at NT = 256: global workspace + reduce kernel; 2 / 4 / 8 / 16 at NT <= 128: cluster
@pytest.mark.parametrize("nt,splits", [(64, 1), (128, 1), (256, 1), (128, 3), (64, 5), (256, 2), (128, 2), (128, 4),
                                       (128, 8), (128, 16), (64, 2), (64, 8), (64, 16)])
def test_conv_tiles_and_split_k(K, nt, splits):
    N, H, Cin, Cout = 4, 8, 256, 256
    x, w, b, y_ref, dy, dx_ref, dw_ref = _case(N, H, Cin, Cout, seed=1)
    xd, wd, bd, dyd = nhwc(x.cuda()), nhwc(w.detach().cuda()), b.cuda(), nhwc(dy.cuda())
    for rep in range(2):
        y = K.nhwc_empty(N, Cout, H, H, "cuda")
        acc = torch.zeros(2 * Cout, dtype=torch.int64, device="cuda")
        K.conv3x3_gemm(0, N, H, H, Cin, Cout, xd, None, wd, y, bd, stat_acc=acc, nt=nt, splits=splits)
        dx = K.nhwc_empty(N, Cin, H, H, "cuda")
        K.conv3x3_gemm(1, N, H, H, Cin, Cout, dyd, None, wd, dx, nt=nt, splits=splits)
        dw = torch.empty_like(wd)
        K.conv3x3_gemm(2, N, H, H, Cin, Cout, dyd, xd, wd, dw, nt=nt, splits=splits)
        torch.cuda.synchronize()
        _check(y.cpu(), y_ref, f"fprop nt={nt} splits={splits} rep={rep}")
        _check(K.acc_to_float(acc, Cout)[0].float().cpu(), y_ref.sum((0, 2, 3)), "stat sum", 1e-3)
        _check(dx.cpu(), dx_ref, f"dgrad nt={nt} splits={splits}")
        _check(dw.cpu(), dw_ref, f"wgrad nt={nt} splits={splits}")
