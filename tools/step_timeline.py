"""In-graph timeline of one training step on ONE GPU (experiments build: every kernel stamps %globaltimer at entry and
exit of its block 0):  python tools/step_timeline.py BATCH [out.txt]
Shows, per kernel of the replayed CUDA graph, start / duration and the gap to the previous kernel -- i.e. where a
latency-bound step (32 images per GPU) actually spends its time, which ncu's serialised cold-cache numbers cannot."""
import os
import sys

import torch

os.environ["DPB_EXPERIMENTS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_b200 import _native  # noqa: E402
from tools.overlap_timeline import NAMES  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 32
out = sys.argv[2] if len(sys.argv) > 2 else None
_native.build(experiments=True)
import distributed_pytorch_b200 as dpb  # noqa: E402
from distributed_pytorch_b200.utils.arena import adopt_module  # noqa: E402

lib = _native.require()
torch.manual_seed(1)
model = dpb.VGG11().cuda()
adopt_module(model)
opt = dpb.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
step = dpb.TrainStep(model, opt, dpb.CrossEntropyLoss(), use_graph=True, warmup=2)
x, y = torch.randn(batch, 3, 32, 32).cuda(), torch.randint(0, 10, (batch,)).cuda()
for _ in range(8):
    step(x, y)
torch.cuda.synchronize()
trace = torch.zeros(1 + 3 * 4096, dtype=torch.int64, device="cuda")
_native.check(lib.dpb_debug_set_trace(trace.data_ptr()), "dpb_debug_set_trace")
torch.cuda.synchronize()
step(x, y)
torch.cuda.synchronize()
lib.dpb_debug_set_trace(None)
t = trace.cpu().tolist()
n = min(t[0], 4096)
recs = sorted(((t[1 + 3 * i], t[2 + 3 * i], t[3 + 3 * i]) for i in range(n)), key=lambda r: r[1])
t0 = recs[0][1]
lines = [f"# one graph replay, batch {batch}, {n} kernels; tuning env: " +
         " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("DPB_") and k != "DPB_EXPERIMENTS"),
         f"# {'start us':>9s} {'dur us':>7s} {'gap us':>7s}  kernel"]
prev_end, busy, per = t0, 0.0, {}
for k, a, b in recs:
    lines.append(f"  {(a - t0) / 1e3:9.2f} {(b - a) / 1e3:7.2f} {(a - prev_end) / 1e3:7.2f}  {NAMES.get(k, k)}")
    busy += (b - a) / 1e3
    d = per.setdefault(NAMES.get(k, k), [0, 0.0, 0.0])
    d[0] += 1; d[1] += (b - a) / 1e3; d[2] += max(0.0, (a - prev_end) / 1e3)
    prev_end = b
total = (max(b for _, _, b in recs) - t0) / 1e3
lines.append(f"# extent {total:.1f} us, sum of block-0 durations {busy:.1f} us, gaps {total - busy:.1f} us")
for name, (c, d, g) in sorted(per.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"#   {name:28s} x{c:3d}  {d:8.1f} us in kernel  {g:7.1f} us gaps before")
text = "\n".join(lines)
print(text)
if out:
    os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
    open(out, "w").write(text + "\n")
