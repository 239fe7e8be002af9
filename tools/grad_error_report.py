"""Per-parameter gradient error of (a) our sm_100a path, (b) torch CUDA with TF32 convs, (c) torch CUDA fp32, all
against the CPU fp32 oracle, for one VGG-11 forward/backward.  Tells TF32 rounding apart from a kernel bug."""
import copy
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402


def rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


def main(batch=32):
    torch.manual_seed(1)
    cpu = dpb.VGG11()
    ours = copy.deepcopy(cpu).cuda()
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = next(p for p in (os.path.join(root, "baseline", "_ref", "model.py"), "/root/reference/model.py") if os.path.exists(p))
    spec = importlib.util.spec_from_file_location("_reference_model", path)
    ref_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_mod)
    tref = ref_mod.VGG11()
    tref.load_state_dict(cpu.state_dict())
    torch.manual_seed(5)
    x, y = torch.randn(batch, 3, 32, 32), torch.randint(0, 10, (batch,))
    crit = torch.nn.CrossEntropyLoss()
    crit(cpu(x), y).backward()
    dpb.CrossEntropyLoss()(ours(x.cuda()), y.cuda()).backward()
    grads = {}
    for name, tf32 in (("torch_tf32", True), ("torch_fp32", False)):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        m = copy.deepcopy(tref).cuda()
        crit(m(x.cuda()), y.cuda()).backward()
        grads[name] = [p.grad.cpu() for p in m.parameters()]
    print(f"batch {batch}: max-abs error / max-abs value vs CPU fp32")
    print(f"{'param':24s} {'ours':>10s} {'torch_tf32':>10s} {'torch_fp32':>10s}")
    worst = 0.0
    for i, ((k, p), q) in enumerate(zip(cpu.named_parameters(), ours.parameters())):
        e = rel(q.grad.cpu(), p.grad)
        et, ef = rel(grads["torch_tf32"][i], p.grad), rel(grads["torch_fp32"][i], p.grad)
        print(f"{k:24s} {e:10.2e} {et:10.2e} {ef:10.2e}")
        if p.grad.abs().max() > 1e-6:
            worst = max(worst, e / max(et, 1e-4))
    print(f"worst ours/torch_tf32 error ratio: {worst:.2f}")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 32)
