#!/usr/bin/env python
"""Headline benchmark: VGG-11 training throughput (images/s, whole job) at global batch 256 on N B200s, per gradient-sync
mode -- the metric BASELINE.json names ("VGG-11 avg sec/iter & images/sec per sync mode at 1/2/4/8 B200 vs ref NCCL").

    python bench.py --gpus 1 --steps 30 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 30 --warmup 5
    python bench.py --impl reference ...       # the UNMODIFIED reference (baseline/_ref) on cuda + NCCL + cuDNN

Rank 0 prints ONE JSON line.  Protocol: W untimed warm-up steps, then exactly K steps between barrier +
torch.cuda.synchronize() on both sides, CUDA-event time, max over ranks; synthetic CIFAR-10-shaped data, random-init
weights; a full optimizer step is inside every timed step.  `value` is device-timed with device-resident inputs;
`e2e` repeats the measurement through the public API with a pinned-host -> device copy of every step's batch and a
device -> host read of every step's loss.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
GLOBAL_BATCH = 256
MODES = ("all_reduce", "gather", "ddp", "single")
DEFAULT_OVERLAP = True   # all_reduce mode: bucketed reduction under backward (profiles/bench_multi_gpu.md: 0.64 vs 0.70 ms at 2 GPUs)


def parse_args(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=30)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", choices=("ours", "reference"), default="ours")
    p.add_argument("--mode", choices=MODES, default="all_reduce")
    p.add_argument("--global-batch", type=int, default=GLOBAL_BATCH)
    p.add_argument("--per-rank-batch", type=int, default=None,
                   help="weak scaling: fixed per-rank batch (the reference's own convention is 256 per rank)")
    p.add_argument("--bucket-cap-mb", type=float, default=None,
                   help="ddp mode: gradient bucket size (default: the framework's, see DistributedDataParallel)")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-fuse-optimizer", action="store_true")
    p.add_argument("--overlap", choices=("auto", "on", "off"), default="auto",
                   help="all_reduce mode: bucketed reduction on a side stream during backward (auto = the default of "
                        "the framework for this world size)")
    p.add_argument("--no-nvls", action="store_true", help="P2P loads/stores even where NVSwitch multicast is available")
    p.add_argument("--all-modes", action="store_true", help="also time the other sync modes (extra keys)")
    return p.parse_args(argv)


# ---------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md recipe).  Polls NVML every few
    milliseconds from a thread (the timed region of a fast run is shorter than nvidia-smi's sampling period); falls
    back to an `nvidia-smi -lms` subprocess when NVML cannot be loaded."""

    def __init__(self, gpu_index: int, period_s: float = 0.004):
        self.gpu, self.period = gpu_index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._fallback = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001 - any NVML problem -> nvidia-smi
            self._nv = None
            self._fallback = _SmiClockSampler(gpu_index)

    def _poll(self):
        nv = self._nv
        names = (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap", "nvmlClocksThrottleReasonSwPowerCap"))
        masks = [(n, getattr(nv, a, None) or getattr(nv, b, 0)) for n, a, b in names]
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(
            nv, "nvmlDeviceGetCurrentClocksThrottleReasons", None)
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                if get_reasons is not None:
                    r = int(get_reasons(self._h))
                    for n, m in masks:
                        if m and (r & m):
                            self.reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def start(self):
        if self._fallback is not None:
            return self._fallback.start()
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()

    def stop(self) -> dict:
        if self._fallback is not None:
            return self._fallback.stop()
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm = self.samples
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
                "reasons": sorted(self.reasons), "source": "nvml"}


class _SmiClockSampler:
    """nvidia-smi clocks / throttle reasons (fallback sampler)."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.lines:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------- distributed
def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def init_distributed(n_gpus: int):
    """(rank, world, local_rank).  Under torchrun the env is read; a bare `python bench.py --gpus N` with N > 1
    re-launches itself through torch.distributed.run."""
    import torch
    import torch.distributed as dist
    if "RANK" not in os.environ:
        if n_gpus > 1:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n_gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.abspath(__file__),
                   *sys.argv[1:]]
            raise SystemExit(subprocess.call(cmd))
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1",
                          MASTER_PORT=str(_free_port()))
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def max_over_ranks(x: float) -> float:
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def fence():
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()


SETTLE_S = 0.25


def settle_steps(sec_per_step: float) -> int:
    """Untimed steps after the W warm-up steps so that the GPU has been under load for SETTLE_S before the timed
    region (clock ramp); agreed across ranks (they must all take the same number of steps)."""
    import torch
    import torch.distributed as dist
    n = int(min(max(SETTLE_S / max(sec_per_step, 1e-5), 0), 2000))
    t = torch.tensor([n], dtype=torch.int64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return int(t.item())


def make_batches(batch: int, count: int = 8, seed: int = 0):
    """`count` distinct pinned host batches (synthetic 3x32x32 images, 10 classes), cycled by the benchmark."""
    import torch
    g = torch.Generator().manual_seed(1234 + seed)
    out = []
    for _ in range(count):
        x = torch.randn(batch, 3, 32, 32, generator=g).pin_memory()
        y = torch.randint(0, 10, (batch,), generator=g).pin_memory()
        out.append((x, y))
    return out


def _config(global_batch: int, batch: int, world: int, mode: str) -> dict:
    """Identical keys (and, for the same command line, identical values) in both arms."""
    return {"model": "VGG-11", "global_batch": global_batch, "per_rank_batch": batch, "seq_len": None,
            "input": "3x32x32", "parallelism": f"dp{world}", "sync_mode": mode,
            "optimizer": "SGD lr0.1 m0.9 wd1e-4",
            "l2": "no explicit flush: every step streams params+grads+momentum (111 MB) plus that step's "
                  "activations, a working set larger than the 126 MB L2"}


# ---------------------------------------------------------------------------------------------------- our arm
def build_ours(mode: str, rank: int, world: int, args):
    import torch
    import distributed_pytorch_b200 as dpb
    from distributed_pytorch_b200.utils.arena import adopt_module

    from distributed_pytorch_b200 import _native
    if args.no_nvls:
        _native.require().dpb_set_tuning(14, 0)
    torch.manual_seed(1)
    dev = torch.device("cuda", torch.cuda.current_device())
    model = dpb.VGG11().to(dev)
    if mode == "ddp":
        kw = {} if args.bucket_cap_mb is None else {"bucket_cap_mb": args.bucket_cap_mb}
        model = dpb.DistributedDataParallel(model, **kw)
    opt = dpb.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    if mode == "ddp" and not args.no_fuse_optimizer:
        model.fuse_optimizer(opt)
    if mode == "gather":
        sync = dpb.GatherScatterSync(model, world, rank)
    elif mode == "all_reduce":
        overlap = (args.overlap == "on") or (args.overlap == "auto" and DEFAULT_OVERLAP and world > 1)
        kw = {} if args.bucket_cap_mb is None else {"bucket_cap_mb": args.bucket_cap_mb}
        sync = dpb.AllReduceSync(model, world, rank, fuse_optimizer=not args.no_fuse_optimizer, optimizer=opt,
                                 overlap=overlap, **kw)
    else:
        sync = dpb.NoSync(model)
        if mode == "single":
            adopt_module(model)
    step = dpb.TrainStep(model, opt, dpb.CrossEntropyLoss(), sync, use_graph=not args.no_graph, warmup=2)
    return model, opt, sync, step


def time_ours(step, batches, steps: int, warmup: int, e2e: bool, gpu_index: int):
    """Returns (ms_per_step max over ranks, clocks, launches in the timed region, last loss)."""
    import torch
    from distributed_pytorch_b200 import _native
    dev_batches = batches if e2e else [(x.cuda(non_blocking=True), y.cuda(non_blocking=True)) for x, y in batches]
    loss = None
    sampler = ClockSampler(gpu_index)  # NVML is initialised here, not between warm-up and the timed region
    tw = time.perf_counter()
    for i in range(warmup):
        x, y = dev_batches[i % len(dev_batches)]
        loss = step(x, y)
        if e2e:
            loss.item()
    torch.cuda.synchronize()
    # W warm-up steps of a ~0.5 ms step are a few milliseconds of GPU work: not enough for the SM clock to come up from
    # idle.  Keep stepping (untimed) until the GPU has been busy for SETTLE_S; the same number of steps on every rank.
    # (decided chunk by chunk: the warm-up steps include eager runs and the graph capture, so their average says nothing
    # about how long a replayed step takes)
    settle, ts = 0, time.perf_counter()
    while settle < 4000:
        for i in range(25):
            x, y = dev_batches[i % len(dev_batches)]
            loss = step(x, y)
            if e2e:
                loss.item()
        settle += 25
        torch.cuda.synchronize()
        if max_over_ranks(time.perf_counter() - ts) >= SETTLE_S:  # the same decision on every rank
            break
    fence()
    sampler.start()
    l0 = _native.launches()
    t0 = time.perf_counter()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(steps):
        x, y = dev_batches[(warmup + i) % len(dev_batches)]
        loss = step(x, y)
        if e2e:
            loss.item()  # device -> host read of the step's result
    end.record()
    fence()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = start.elapsed_time(end)
    clocks = sampler.stop()
    launches = _native.launches() - l0
    ms = max_over_ranks((wall_ms if e2e else dev_ms) / steps)
    clocks["settle_steps"] = settle
    return ms, clocks, launches, float(loss.item())


def run_ours(args, rank, world, local):
    import torch
    batch = args.per_rank_batch or max(args.global_batch // world, 1)
    global_batch = batch * world
    batches = make_batches(batch, seed=rank)
    model, opt, sync, step = build_ours(args.mode, rank, world, args)
    ms, clocks, launches, loss = time_ours(step, batches, args.steps, args.warmup, False, local)
    e2e_ms, _, _, _ = time_ours(step, batches, args.steps, max(3, args.warmup // 2), True, local)
    x0, y0 = batches[0]
    out = {
        "metric": "vgg11_train_images_per_sec", "value": global_batch / (ms / 1e3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "sec_per_iter": ms / 1e3,
        "higher_is_better": True, "scaling": "weak" if args.per_rank_batch else "strong", "vs_baseline": None,
        "dtype": "tf32 tensor-core inputs, fp32 accumulate/storage (the reference's cuDNN default)",
        "data": "synthetic 3x32x32 images, random-init weights (no network for CIFAR-10)", "impl": "ours",
        "config": _config(global_batch, batch, world, args.mode),
        "impl_notes": {"fused_optimizer": bool(getattr(sync, "fuses_optimizer", False)
                                               or getattr(model, "fuses_optimizer", False)),
                       "overlap": bool(getattr(sync, "reducer", None) is not None or args.mode == "ddp"),
                       "nvls": bool((hasattr(sync, "nvls") and sync.nvls()) or (hasattr(model, "nvls") and model.nvls())),
                       "cuda_graph": not args.no_graph,
                       "host_sync_per_step": False,
                       "note": "device-timed arm: graph replays back to back, loss read once at the end; the e2e arm "
                               "copies a pinned host batch in and reads the loss every step (like the reference loop)"},
        "clocks": clocks,
        "e2e": {"value": global_batch / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": x0.numel() * 4 + y0.numel() * 8, "d2h_bytes_per_step": 4,
                "api": "distributed_pytorch_b200.TrainStep(pinned host batch) + loss.item() every step"},
        "gpu_launches": launches, "final_loss": loss,
    }
    if args.all_modes:
        extra = {}
        for m in MODES:
            if m == args.mode or (m == "single" and world > 1):
                continue
            a2 = argparse.Namespace(**vars(args))
            a2.mode = m
            _, _, s2, st2 = build_ours(m, rank, world, a2)
            ms2, _, _, _ = time_ours(st2, batches, args.steps, args.warmup, False, local)
            extra[m] = {"ms_per_step": ms2, "images_per_sec": global_batch / (ms2 / 1e3)}
        out["other_modes"] = extra
    return out


# ---------------------------------------------------------------------------------------------------- reference arm
REF_DIR = os.path.join(ROOT, "baseline", "_ref")
REF_SCRIPT = {"all_reduce": "main_all_reduce", "gather": "main_gather", "ddp": "main_ddp", "single": "main"}


class _TimedLoader:
    """Feeds the reference's own train_model() (its `for batch_idx, (data, target) in enumerate(train_loader)` loop,
    reference main.py:30) W + K synthetic batches and brackets the last K with barrier + synchronize + CUDA events.
    The reference code itself is untouched: it only ever sees an iterable."""

    def __init__(self, batches, warmup, steps, h2d, gpu_index):
        self.batches, self.warmup, self.steps, self.h2d, self.gpu = batches, warmup, steps, h2d, gpu_index
        self.dev_ms = self.wall_ms = None
        self.clocks = None
        self.settle = 0

    def __len__(self):
        return self.warmup + self.settle + self.steps

    def __iter__(self):
        import torch
        dev = [(x.cuda(), y.cuda()) for x, y in self.batches]
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(self.gpu)
        t0 = 0.0
        tw = time.perf_counter()
        settle, i = 0, 0
        while i < self.warmup + settle + self.steps:
            if i == self.warmup and settle == 0:  # same settle rule as our arm (clock ramp), see settle_steps()
                torch.cuda.synchronize()
                settle = settle_steps((time.perf_counter() - tw) / max(self.warmup, 1))
                self.settle = settle
            if i == self.warmup + settle:
                fence()
                sampler.start()
                t0 = time.perf_counter()
                start.record()
            if self.h2d:
                x, y = self.batches[i % len(self.batches)]
                yield x.cuda(non_blocking=True), y.cuda(non_blocking=True)
            else:
                yield dev[i % len(dev)]
            i += 1
        end.record()
        fence()
        self.wall_ms = (time.perf_counter() - t0) * 1e3
        self.dev_ms = start.elapsed_time(end)
        self.clocks = sampler.stop()
        self.clocks["settle_steps"] = self.settle


def run_reference(args, rank, world, local):
    import contextlib
    import importlib
    import torch
    if not os.path.exists(os.path.join(REF_DIR, "model.py")):
        return {"impl": "reference", "unavailable": "baseline/_ref is empty (run baseline/install_reference.py)"}
    sys.modules.pop("model", None)
    sys.path.insert(0, REF_DIR)   # the reference does `import model as mdl`: must resolve to ITS model.py
    try:
        with contextlib.redirect_stdout(sys.stderr):
            script = importlib.import_module(REF_SCRIPT[args.mode])
            mdl = importlib.import_module("model")
    finally:
        sys.path.remove(REF_DIR)
    assert os.path.dirname(os.path.abspath(mdl.__file__)) == REF_DIR
    batch = args.per_rank_batch or max(args.global_batch // world, 1)
    global_batch = batch * world
    batches = make_batches(batch, seed=rank)
    dev = torch.device("cuda", local)

    # exactly what the reference's main() builds (main_all_reduce.py:124-130), on the GPU
    torch.manual_seed(1)
    model = mdl.VGG11()
    model.to(dev)
    if args.mode == "ddp":
        model = torch.nn.parallel.DistributedDataParallel(model)
    criterion = torch.nn.CrossEntropyLoss().to(dev)
    optimizer = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    ns = argparse.Namespace(num_nodes=world, rank=rank, master_ip="127.0.0.1")

    def one_pass(h2d: bool, warmup: int):
        loader = _TimedLoader(batches, warmup, args.steps, h2d, local)
        with contextlib.redirect_stdout(sys.stderr):   # the reference prints progress lines; keep stdout = one JSON
            import inspect
            if len(inspect.signature(script.train_model).parameters) >= 6:  # every script but main.py takes `args`
                script.train_model(model, loader, optimizer, criterion, 0, ns)
            else:
                script.train_model(model, loader, optimizer, criterion, 0)
        return loader

    l1 = one_pass(False, args.warmup)
    ms = max_over_ranks(l1.dev_ms / args.steps)
    l2 = one_pass(True, max(3, args.warmup // 2))
    e2e_ms = max_over_ranks(l2.wall_ms / args.steps)
    x0, y0 = batches[0]
    return {
        "metric": "vgg11_train_images_per_sec", "value": global_batch / (ms / 1e3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "sec_per_iter": ms / 1e3,
        "higher_is_better": True, "scaling": "weak" if args.per_rank_batch else "strong", "vs_baseline": None,
        "dtype": "fp32 storage, cuDNN default (TF32 conv allowed), as the reference code runs on CUDA",
        "data": "synthetic 3x32x32 images, random-init weights", "impl": "reference",
        "config": _config(global_batch, batch, world, args.mode),
        "impl_notes": {"code": f"baseline/_ref/{REF_SCRIPT[args.mode]}.py train_model() unmodified, backend=nccl, "
                               "torch cuDNN/cuBLAS kernels",
                       "host_sync_per_step": True,
                       "note": "the reference loop calls loss.item() every step in both arms"},
        "clocks": l1.clocks,
        "e2e": {"value": global_batch / (e2e_ms / 1e3), "unit": "images/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": x0.numel() * 4 + y0.numel() * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": 0,
    }


def main():
    args = parse_args()
    if args.impl == "reference" and not os.path.exists(os.path.join(REF_DIR, "model.py")):
        print(json.dumps({"impl": "reference",
                          "unavailable": "baseline/_ref is empty: run `python baseline/install_reference.py`"}))
        return
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible"}))
        return
    rank, world, local = init_distributed(args.gpus)
    try:
        res = run_reference(args, rank, world, local) if args.impl == "reference" else run_ours(args, rank, world, local)
        if rank == 0:
            print(json.dumps(res), flush=True)
    finally:
        import torch.distributed as dist
        torch.cuda.synchronize()
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
