"""The reference's layer types used on their own (any arrangement other than conv -> bn -> relu [-> pool]) run on native
kernels too: forward and backward against the CPU fp32 oracle (reference model.py:16-25)."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from distributed_pytorch_b200.models.vgg import BatchNorm, Conv3x3, FusedSequential, MaxPool2x2, ReLU

pytestmark = pytest.mark.gpu


def nhwc(t):
    return t.contiguous(memory_format=torch.channels_last)


def _rel(a, b):
    a, b = a.detach(), b.detach()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-6))


def test_relu_and_maxpool_alone():
    torch.manual_seed(0)
    x = torch.randn(6, 64, 8, 8)
    x[0, :, :2, :2] = 0.5  # ties inside a pooling window: the first maximum must take the gradient (aten's rule)
    for mod, ref in ((ReLU(), F.relu), (MaxPool2x2(), lambda t: F.max_pool2d(t, 2, 2))):
        xc = x.clone().requires_grad_(True)
        yc = ref(xc)
        g = torch.randn_like(yc)
        yc.backward(g)
        xg = nhwc(x.cuda()).requires_grad_(True)
        yg = mod(xg)
        yg.backward(nhwc(g.cuda()))
        assert torch.equal(yg.cpu(), yc.detach())
        assert torch.equal(xg.grad.cpu(), xc.grad)


@pytest.mark.parametrize("training", [True, False])
def test_batchnorm_alone(training):
    torch.manual_seed(1)
    cpu = nn.BatchNorm2d(128)
    with torch.no_grad():
        cpu.weight.uniform_(0.5, 1.5)
        cpu.bias.normal_(0, 0.2)
        cpu.running_mean.normal_(0, 0.1)
        cpu.running_var.uniform_(0.5, 1.5)
    gpu = BatchNorm(128)
    gpu.load_state_dict(cpu.state_dict())
    gpu = gpu.cuda()
    cpu.train(training)
    gpu.train(training)
    x = torch.randn(8, 128, 4, 4) * 1.3 + 0.2
    xc = x.clone().requires_grad_(True)
    yc = cpu(xc)
    xg = nhwc(x.cuda()).requires_grad_(True)
    yg = gpu(xg)
    assert torch.allclose(yg.cpu(), yc.detach(), rtol=1e-4, atol=1e-5)
    if training:
        g = torch.randn_like(yc)
        yc.backward(g)
        yg.backward(nhwc(g.cuda()))
        assert torch.allclose(xg.grad.cpu(), xc.grad, rtol=1e-3, atol=1e-5)
        assert torch.allclose(gpu.weight.grad.cpu(), cpu.weight.grad, rtol=1e-3, atol=1e-4)
        assert torch.allclose(gpu.bias.grad.cpu(), cpu.bias.grad, rtol=1e-3, atol=1e-4)
        assert torch.allclose(gpu.running_mean.cpu(), cpu.running_mean, rtol=1e-4, atol=1e-6)
        assert torch.allclose(gpu.running_var.cpu(), cpu.running_var, rtol=1e-4, atol=1e-6)
        assert int(gpu.num_batches_tracked) == int(cpu.num_batches_tracked) == 1


@pytest.mark.parametrize("cin,cout,hw", [(64, 128, 8), (3, 64, 32)])
def test_conv_alone(cin, cout, hw):
    """Conv2d(3x3) + bias with no BatchNorm behind it: output, data / weight gradients (TF32 accuracy) and the bias
    gradient (a plain column sum, fp32 accuracy)."""
    torch.manual_seed(3)
    cpu = Conv3x3(cin, cout)
    gpu = copy.deepcopy(cpu).cuda()
    x = torch.randn(8, cin, hw, hw)
    xc = x.clone().requires_grad_(cin != 3)
    yc = cpu(xc)
    g = torch.randn_like(yc)
    yc.backward(g)
    xg = (nhwc(x.cuda()) if cin != 3 else x.cuda()).requires_grad_(cin != 3)
    yg = gpu(xg)
    yg.backward(nhwc(g.cuda()))
    assert _rel(yg.cpu(), yc) < 3e-3
    assert _rel(gpu.weight.grad.cpu(), cpu.weight.grad) < 3e-3
    assert torch.allclose(gpu.bias.grad.cpu(), cpu.bias.grad, rtol=1e-4, atol=1e-3)
    if cin != 3:
        assert _rel(xg.grad.cpu(), xc.grad) < 3e-3


def test_modules_in_another_arrangement():
    """conv -> relu -> pool -> conv -> bn -> pool -> relu: nothing here matches the fused block, every module runs alone;
    wrapped in a FusedSequential as well, to check that it only fuses what it should."""
    torch.manual_seed(2)
    def build(seq):
        return seq(Conv3x3(3, 64), ReLU(), MaxPool2x2(), Conv3x3(64, 128), BatchNorm(128), MaxPool2x2(), ReLU())
    cpu = build(nn.Sequential)
    x = torch.randn(8, 3, 16, 16)
    yc = cpu(x)
    g = torch.randn_like(yc)
    yc.backward(g)
    for seq in (nn.Sequential, FusedSequential):
        gpu = build(seq)
        gpu.load_state_dict(cpu.state_dict())
        gpu = gpu.cuda()
        yg = gpu(x.cuda())
        yg.backward(nhwc(g.cuda()))
        assert _rel(yg.cpu(), yc.detach()) < 5e-3  # TF32 operands in the two convolutions
        # TF32 rounding flips a few ReLU masks / pool arg-maxes on the way, so gradients agree in direction, not per element
        for (k, p), (_, q) in zip(gpu.named_parameters(), cpu.named_parameters()):
            a, b = p.grad.detach().cpu().flatten(), q.grad.detach().flatten()
            if float(b.abs().max()) < 1e-3:  # bias of the conv in front of the BatchNorm: zero up to rounding noise
                assert float(a.abs().max()) < 1e-2, k
                continue
            cos, rel = float(F.cosine_similarity(a, b, dim=0)), _rel(a, b)
            assert cos > 0.995 and rel < 0.2, (seq.__name__, k, cos, rel)
