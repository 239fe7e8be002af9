"""A few eager (no CUDA graph) training steps of VGG-11 for kernel-level profiling under ncu:
   python tools/step_profile.py BATCH STEPS
Every kernel of ours lives in namespace dpb::, so `ncu -k regex:dpb` selects exactly the step's kernels; one step is
70 launches (single process)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_pytorch_b200 as dpb  # noqa: E402
from distributed_pytorch_b200 import _native  # noqa: E402
from distributed_pytorch_b200.utils.arena import adopt_module  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
torch.manual_seed(1)
model = dpb.VGG11().cuda()
adopt_module(model)
opt = dpb.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
step = dpb.TrainStep(model, opt, dpb.CrossEntropyLoss(), use_graph=False)
x, y = torch.randn(batch, 3, 32, 32).cuda(), torch.randint(0, 10, (batch,)).cuda()
for i in range(steps):
    l0 = _native.launches()
    loss = step(x, y)
    torch.cuda.synchronize()
    print(f"step {i}: loss {float(loss):.4f}, launches {_native.launches() - l0}")
