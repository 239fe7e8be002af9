"""state_dict schema, initialisation parity and CPU numerics against the reference's model.py."""
import torch

import distributed_pytorch_b200 as dpb
from distributed_pytorch_b200.utils.checkpoint import load_checkpoint, save_checkpoint


def test_58_keys_and_shapes(ref_model_module):
    ref = ref_model_module.VGG11()
    ours = dpb.VGG11()
    a, b = ref.state_dict(), ours.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert len(b) == 58
    for k in a:
        assert a[k].shape == b[k].shape and a[k].dtype == b[k].dtype, k
    assert sum(p.numel() for p in ours.parameters()) == 9_231_114
    assert len(list(ours.parameters())) == 34 and len(list(ours.buffers())) == 24


def test_same_seed_same_init(ref_model_module):
    torch.manual_seed(1)
    ref = ref_model_module.VGG11()
    torch.manual_seed(1)
    ours = dpb.VGG11()
    for (k, a), (_, b) in zip(ref.state_dict().items(), ours.state_dict().items()):
        assert torch.equal(a, b), k


def test_cpu_forward_backward_matches_reference(ref_model_module):
    torch.manual_seed(1)
    ref = ref_model_module.VGG11()
    torch.manual_seed(1)
    ours = dpb.VGG11()
    x = torch.randn(8, 3, 32, 32)
    y = torch.randint(0, 10, (8,))
    la = torch.nn.CrossEntropyLoss()(ref(x), y)
    lb = dpb.CrossEntropyLoss()(ours(x), y)
    la.backward()
    lb.backward()
    assert torch.allclose(la, lb, rtol=1e-5, atol=1e-6)
    for (k, p), (_, q) in zip(ref.named_parameters(), ours.named_parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=1e-4, atol=1e-5), k
    for (k, p), (_, q) in zip(ref.named_buffers(), ours.named_buffers()):
        assert torch.allclose(p.float(), q.float(), rtol=1e-5, atol=1e-6), k


def test_checkpoint_roundtrip_with_reference(tmp_path, ref_model_module):
    torch.manual_seed(3)
    ours = dpb.VGG11()
    opt = dpb.SGD(ours.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 10, (4,))
    dpb.CrossEntropyLoss()(ours(x), y).backward()
    opt.step()
    path = str(tmp_path / "ckpt.pt")
    save_checkpoint(path, ours, opt)
    # the reference model loads our file ...
    ref = ref_model_module.VGG11()
    ref.load_state_dict(torch.load(path))
    for (k, a), (_, b) in zip(ref.state_dict().items(), ours.state_dict().items()):
        assert torch.equal(a, b), k
        assert a.is_contiguous()
    # ... and we load a file written from the reference model
    torch.save(ref.state_dict(), path + ".ref")
    again = dpb.VGG11()
    opt2 = dpb.SGD(again.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    load_checkpoint(path + ".ref", again)
    load_checkpoint(path, again, opt2)
    for (k, a), (_, b) in zip(again.state_dict().items(), ours.state_dict().items()):
        assert torch.equal(a, b), k
    sa, sb = opt.state_dict(), opt2.state_dict()
    assert sa["param_groups"][0]["lr"] == sb["param_groups"][0]["lr"]
    for i in sa["state"]:
        assert torch.equal(sa["state"][i]["momentum_buffer"], sb["state"][i]["momentum_buffer"])


def test_other_vgg_variants_build():
    for f, n_conv in ((dpb.VGG13, 10), (dpb.VGG16, 13), (dpb.VGG19, 16)):
        m = f()
        assert sum(isinstance(x, torch.nn.Conv2d) for x in m.modules()) == n_conv
    out = dpb.VGG13().eval()(torch.randn(2, 3, 32, 32))
    assert out.shape == (2, 10)


def test_modules_one_by_one_equal_the_fused_sequence_on_cpu():
    """Walking `model.layers` module by module (what code written against the reference's nn.Sequential may do) gives
    the same result as the FusedSequential forward; the functional entry points agree with torch.nn.functional."""
    import torch.nn.functional as F
    from distributed_pytorch_b200.ops import functional as Fn
    torch.manual_seed(3)
    model = dpb.VGG11()
    x = torch.randn(4, 3, 32, 32)
    fused = model.layers(x)
    y = x
    for m in model.layers:
        y = m(y)
    assert torch.allclose(fused, y, rtol=1e-5, atol=1e-6)
    conv, bn = model.layers[0], model.layers[1]
    assert torch.equal(Fn.conv3x3(x, conv.weight, conv.bias), F.conv2d(x, conv.weight, conv.bias, padding=1))
    z = torch.randn(4, 64, 8, 8)
    assert torch.equal(Fn.relu(z), F.relu(z)) and torch.equal(Fn.max_pool2x2(z), F.max_pool2d(z, 2, 2))
    rm, rv, nbt = torch.zeros(64), torch.ones(64), torch.zeros((), dtype=torch.long)
    out = Fn.batch_norm(z, bn.weight, bn.bias, rm, rv, nbt, True)
    assert torch.allclose(out, F.batch_norm(z, None, None, bn.weight, bn.bias, True), rtol=1e-5, atol=1e-6)
    assert int(nbt) == 1 and float(rm.abs().max()) > 0
