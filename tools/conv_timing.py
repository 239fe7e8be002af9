"""Where does a tensor-core conv CTA spend its time?  Uses the bring-up instrumentation of conv_gemm_kernel (a debug
buffer makes CTA (grid.x/2, 0, 0) report cycle counters): issuer cycles waiting for operands vs main-loop cycles, and
the epilogue's wait-for-accumulator vs its own duration.  VGG-11 layer shapes at a given batch."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_b200 import _native  # noqa: E402
from distributed_pytorch_b200.ops import kernels as K  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
os.environ["DPB_EXPERIMENTS"] = "1"
_native.build(experiments=True)  # the in-kernel tracing is compiled in only with -DDPB_EXPERIMENTS
lib = _native.require()
dump = torch.zeros(16384, device="cuda")
lib.dpb_debug_set_conv_dump(dump.data_ptr())
layers = [(16, 64, 128), (8, 128, 256), (8, 256, 256), (4, 256, 512), (4, 512, 512), (2, 512, 512)]
print(f"batch {batch}: cycles of CTA (grid.x/2,0,0); wait = issuer thread blocked on the TMA 'full' barrier")
for h, cin, cout in layers:
    x = torch.randn(batch, cin, h, h, device="cuda").contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 3, 3, device="cuda") * 0.05).contiguous(memory_format=torch.channels_last)
    dy = torch.randn(batch, cout, h, h, device="cuda").contiguous(memory_format=torch.channels_last)
    for mode, name in ((0, "fprop"), (1, "dgrad"), (2, "wgrad")):
        if mode == 0:
            out = K.nhwc_empty(batch, cout, h, h, "cuda")
            args = (0, batch, h, h, cin, cout, x, None, w, out)
        elif mode == 1:
            out = K.nhwc_empty(batch, cin, h, h, "cuda")
            args = (1, batch, h, h, cin, cout, dy, None, w, out)
        else:
            out = torch.empty_like(w)
            args = (2, batch, h, h, cin, cout, dy, x, w, out)
        if len(sys.argv) > 2:  # force the split-K factor (1 = few CTAs with long K loops on an otherwise idle chip)
            args = args + (None, False, 0, int(sys.argv[2]))
        for _ in range(2):
            K.conv3x3_gemm(*args)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dump.zero_()
        s.record()
        K.conv3x3_gemm(*args)
        e.record()
        torch.cuda.synchronize()
        d = dump[12288:12296].cpu().tolist()
        wait, loop, kc, acc_wait, epi, setup, cyc_total, ns_total = d
        tr = dump[12800:12800 + 3072].view(torch.int32).cpu().view(-1, 3).long() & 0xFFFFFFFF
        tr = tr[tr[:, 0] != 0]
        if len(tr) == 0:  # persistent kernel (DPB_CONV_PERSIST): no per-CTA trace, but two more counters
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(20):
                    K.conv3x3_gemm(*args)
            g.replay()
            torch.cuda.synchronize()
            s.record()
            for _ in range(5):
                g.replay()
            e.record()
            torch.cuda.synchronize()
            print(f"{h:2d}x{h:<2d} {cin:3d}->{cout:3d} {name}: in-graph {s.elapsed_time(e) * 1e3 / 100:6.1f} us/launch | persistent, {cyc_total:.0f} units/CTA, "
                  f"{int(kc)} k-iters: issue loop {loop:8.0f} cyc ({loop / max(kc, 1):5.0f}/iter), waiting for operands {100 * wait / max(loop, 1):4.1f}%, "
                  f"for a free accumulator {100 * setup / max(loop, 1):4.1f}% | epilogue warps: waited {acc_wait:7.0f}, drained {epi:7.0f} cyc")
            continue
        t0 = int(tr[:, 0].min())
        st, en = (tr[:, 0] - t0) & 0xFFFFFFFF, (tr[:, 1] - t0) & 0xFFFFFFFF
        life = (en - st).float()
        late = int((st > 2000).sum())  # CTAs that started more than 2 us after the first: a second wave
        sm_ids, sm_counts = tr[:, 2].unique(return_counts=True)
        per_sm = dict(zip(sm_ids.tolist(), sm_counts.tolist()))
        shared = torch.tensor([per_sm[int(s_)] > 1 for s_ in tr[:, 2]])
        q = torch.quantile(life, torch.tensor([0.1, 0.5, 0.9, 1.0])) / 1e3
        lone = life[~shared].median().item() / 1e3 if (~shared).any() else float("nan")
        pair = life[shared].median().item() / 1e3 if shared.any() else float("nan")
        end_q = torch.quantile(en.float(), torch.tensor([0.1, 0.5, 0.9, 1.0])) / 1e3
        dist = (f"life p10/p50/p90/max {q[0]:.1f}/{q[1]:.1f}/{q[2]:.1f}/{q[3]:.1f} us; alone-on-SM median {lone:.1f}, sharing median {pair:.1f}; "
                f"end time p10/p50/p90/max {end_q[0]:.1f}/{end_q[1]:.1f}/{end_q[2]:.1f}/{end_q[3]:.1f} us")
        trace = (f"{dist}\n        timed CTA: prologue {setup:.0f} cyc, entry->exit {cyc_total:.0f} cyc in {ns_total / 1e3:.1f} us "
                 f"= {cyc_total / max(ns_total, 1) * 1e3:.0f} MHz | {len(tr)} CTAs on {len(tr[:, 2].unique())} SMs: span {int(en.max()) / 1e3:5.1f} us, last start +{int(st.max()) / 1e3:4.1f} us "
                 f"({late} late), CTA life median {life.median().item() / 1e3:4.1f} max {life.max().item() / 1e3:4.1f} us")
        # steady-state cost of the launch inside a CUDA graph: 20 back-to-back copies, replayed 5 times
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                K.conv3x3_gemm(*args)
        g.replay()
        torch.cuda.synchronize()
        s.record()
        for _ in range(5):
            g.replay()
        e.record()
        torch.cuda.synchronize()
        print(f"{h:2d}x{h:<2d} {cin:3d}->{cout:3d} {name}: in-graph {s.elapsed_time(e) * 1e3 / 100:6.1f} us/launch | k-iters {int(kc):4d} "
              f"main loop {loop:8.0f} cyc ({loop / max(kc, 1):6.0f}/iter, MMA floor 256) operand wait {100 * wait / max(loop, 1):5.1f}% | "
              f"epilogue: waited {acc_wait:8.0f} cyc, ran {epi:7.0f} cyc\n        {trace}")

# cost of a (nearly) empty dependent kernel node in a graph: the floor every one of the step's ~86 launches pays
c = 64
one, zero = torch.ones(c, device="cuda"), torch.zeros(c, device="cuda")
K.bn_eval_scale_shift(c, one, zero, zero, one, 1e-5)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(100):
        K.bn_eval_scale_shift(c, one, zero, zero, one, 1e-5)
g.replay()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5):
    g.replay()
e.record()
torch.cuda.synchronize()
print(f"empty dependent kernel node in a graph: {s.elapsed_time(e) * 1e3 / 500:.2f} us")
