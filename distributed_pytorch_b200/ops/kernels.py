"""Thin Python launchers over the C ABI of ``libdpb200.so`` (one function per kernel group).

All tensors are fp32 CUDA tensors; activations are NHWC in memory (PyTorch ``channels_last`` tensors of logical shape
``(N, C, H, W)``), conv weights are KRSC in memory (``channels_last`` views of ``(Cout, Cin, 3, 3)``).  Every launcher
enqueues on ``torch.cuda.current_stream()`` and returns immediately.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from .. import _native as N

__all__ = [
    "Scratch", "scratch", "conv3x3_gemm", "first_conv_fprop", "first_conv_wgrad", "bn_finalize",
    "bn_eval_scale_shift", "bn_relu_pool_fwd", "bn_relu_pool_fwd_train", "bn_relu_pool_bwd", "linear_fwd",
    "linear_bwd", "cross_entropy_fwd", "cross_entropy_bwd", "head_supported", "head_fwd", "head_bwd", "augment_batch",
    "sgd_step", "is_nhwc", "nhwc_empty",
]


def is_nhwc(t: torch.Tensor) -> bool:
    """True if a 4-D tensor is dense NHWC in memory."""
    if t.dim() != 4:
        return False
    n, c, h, w = t.shape
    return t.stride() == (h * w * c, 1, w * c, c) or t.is_contiguous(memory_format=torch.channels_last)


def nhwc_empty(n: int, c: int, h: int, w: int, device) -> torch.Tensor:
    return torch.empty((n, c, h, w), device=device, dtype=torch.float32, memory_format=torch.channels_last)


class Scratch:
    """Per-device scratch the kernels share (stream-ordered reuse): split-K partial tiles of the workspace path, the ring
    of BatchNorm accumulator slots, BatchNorm-backward partial rows and coefficients of the three-launch path,
    first-layer wgrad partials."""

    def __init__(self, device: torch.device):
        lib = N.require()
        N.init_device(device.index)
        self.device = device
        self.conv_ws = torch.empty(lib.dpb_conv_workspace_bytes() // 4, device=device, dtype=torch.float32)
        # BatchNorm accumulators (bn_acc.cuh): a ring of zeroed [2][1024] int64 slots + one ticket each.  A producer
        # (conv / backward reduce) fills a slot, its consumer clears it again, so a slot is clean whenever its turn
        # comes round; 64 slots cover every BatchNorm of a forward + backward pass of the deepest VGG twice over.
        self.acc_slots = 64
        self.bn_acc = torch.zeros((self.acc_slots, 2 * 1024), device=device, dtype=torch.int64)
        self.bn_tickets = torch.zeros(self.acc_slots, device=device, dtype=torch.int32)
        self._acc_next = 0
        self.head_counter = torch.zeros(4, device=device, dtype=torch.int32)
        # second split-K workspace: weight-gradient GEMMs running on the side stream (WgradLane) must not share partial
        # tiles with the data-gradient GEMM that runs next to them
        self.conv_ws_side = None
        self.bn_blocks = lib.dpb_bn_bwd_blocks()
        self.bn_part = torch.empty(2 * self.bn_blocks * 1024, device=device, dtype=torch.float32)
        self.bn_coef = torch.empty(3 * 1024, device=device, dtype=torch.float32)
        self.fc_wgrad_part = torch.empty(lib.dpb_first_conv_wgrad_blocks() * 64 * 27, device=device, dtype=torch.float32)

    def next_acc(self):
        """(accumulator slot, its ticket) for one producer -> consumer pair of launches."""
        i = self._acc_next
        self._acc_next = (i + 1) % self.acc_slots
        return self.bn_acc[i], self.bn_tickets[i:i + 1]


_scratch: dict[int, Scratch] = {}


def scratch(device: torch.device) -> Scratch:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    s = _scratch.get(idx)
    if s is None:
        s = Scratch(torch.device("cuda", idx))
        _scratch[idx] = s
    return s


class WgradLane:
    """Second compute stream of a device for weight-gradient kernels.

    In backward, a layer's weight gradient depends only on that layer's ``dY`` and saved input, while the chain that
    continues backward is data-gradient GEMM -> previous layer's BatchNorm/ReLU/pool backward (memory-bound) -> ...
    Running the weight-gradient GEMM (+ its split-K reduce) on its own stream lets it fill the SMs the chain leaves idle:
    wave tails, epilogues, the elementwise kernels.  The reference's backward (cuDNN through autograd) is single-stream.

    ``fork()`` orders the lane after the current stream, ``hold()`` keeps the kernels' inputs alive (they were allocated
    on the main stream, so the caching allocator must not hand them out again before the lane is done), ``join()`` makes
    the current stream wait for the lane and drops the references.  ``join`` is queued as an autograd end-of-backward
    callback by the first fork of a backward pass, so ``.grad`` is safe to read on the main stream after ``backward()``;
    consumers that run earlier (the DDP reducer) call ``pending`` / ``stream`` themselves.
    """

    def __init__(self, device: torch.device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self._fork_ev = torch.cuda.Event()
        self._join_ev = torch.cuda.Event()
        self._held: list = []
        self.pending = False
        self._callback_queued = False
        sc = scratch(device)
        if sc.conv_ws_side is None:
            sc.conv_ws_side = torch.empty_like(sc.conv_ws)

    def fork(self) -> None:
        self._fork_ev.record(torch.cuda.current_stream(self.device))
        self.stream.wait_event(self._fork_ev)
        self.pending = True
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self.join)
            self._callback_queued = True

    def hold(self, *tensors) -> None:
        self._held.extend(tensors)

    def join(self) -> None:
        self._callback_queued = False
        if self.pending:
            self._join_ev.record(self.stream)
            torch.cuda.current_stream(self.device).wait_event(self._join_ev)
            self.pending = False
        self._held.clear()


_lanes: dict[int, WgradLane] = {}
_wgrad_overlap = os.environ.get("DPB_WGRAD_OVERLAP", "1") == "1"
# CTA count each of the two concurrent GEMMs (data gradient on the main stream, weight gradient on the lane) aims for:
# one per SM, so that together they fill the two CTA slots of every SM.
BWD_TARGET_CTAS = int(os.environ.get("DPB_BWD_TARGET_CTAS", "148"))


def set_wgrad_overlap(enabled: bool) -> None:
    """Weight-gradient kernels on a second stream next to the data-gradient GEMM (default on; ``DPB_WGRAD_OVERLAP=0``
    switches it off).  Round 1 measured this as neutral because each GEMM was split to fill both CTA slots of every SM
    on its own; with each of the two asking for one CTA per SM (``BWD_TARGET_CTAS``) they genuinely run side by side:
    0.952 -> 0.900 ms/step at batch 256, 0.483 -> 0.455 ms at batch 32 (profiles/step_variants.md)."""
    global _wgrad_overlap
    _wgrad_overlap = bool(enabled)


def wgrad_overlap() -> bool:
    return _wgrad_overlap


def wgrad_lane(device: torch.device) -> WgradLane:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    lane = _lanes.get(idx)
    if lane is None:
        lane = WgradLane(torch.device("cuda", idx))
        _lanes[idx] = lane
    return lane


def active_wgrad_lane(device: torch.device) -> Optional[WgradLane]:
    """The device's lane if it holds work the current stream has not been ordered after yet."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    lane = _lanes.get(idx)
    return lane if lane is not None and lane.pending else None


def conv3x3_gemm(mode: int, n: int, h: int, w: int, cin: int, cout: int, inp: torch.Tensor,
                 inp2: Optional[torch.Tensor], weight: torch.Tensor, out: torch.Tensor,
                 bias: Optional[torch.Tensor] = None, stat_acc: Optional[torch.Tensor] = None, nt: int = 0,
                 splits: int = 0, side: bool = False, target_ctas: int = 0) -> None:
    """tcgen05 implicit-GEMM conv. mode 0 fprop / 1 dgrad / 2 wgrad.  ``stat_acc`` (fprop): zeroed int64 [2*cout]
    accumulators that receive the BatchNorm batch statistics (fixed point, 2^-20).  ``side`` selects the second split-K
    workspace (see :class:`WgradLane`)."""
    lib = N.require()
    sc = scratch(out.device)
    conv_ws = sc.conv_ws_side if side else sc.conv_ws
    N.check(lib.dpb_conv3x3_gemm(mode, n, h, w, cin, cout, N.ptr(inp), N.ptr(inp2), N.ptr(weight), N.ptr(out),
                                 N.ptr(bias), N.ptr(stat_acc), N.ptr(conv_ws), nt, splits, target_ctas, N.stream()),
            "dpb_conv3x3_gemm")
    N.count_launches(lib.dpb_conv_last_launches())


ACC_SCALE_FWD = float(1 << 20)


def acc_to_float(acc: torch.Tensor, c: int):
    """(sum, sum of squares) per channel out of a forward accumulator slot (tests)."""
    return acc[:c].double() / ACC_SCALE_FWD, acc[c:2 * c].double() / ACC_SCALE_FWD


def first_conv_fprop(x_nchw: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], y: torch.Tensor,
                     want_stats: bool, stat_acc: Optional[torch.Tensor] = None):
    """First layer.  Statistics as partial rows (``want_stats``: returns (part_sum, part_sq, rows) for bn_finalize) or
    into ``stat_acc`` (tensor-core kernel only)."""
    lib = N.require()
    n, _, h, w = x_nchw.shape
    blocks = lib.dpb_first_conv_blocks(n, h, w)
    ps = pq = None
    if want_stats:
        ps = torch.empty((blocks, 64), device=y.device, dtype=torch.float32)
        pq = torch.empty((blocks, 64), device=y.device, dtype=torch.float32)
    N.check(lib.dpb_first_conv_fprop(N.ptr(x_nchw), N.ptr(weight), N.ptr(bias), N.ptr(y), N.ptr(ps), N.ptr(pq),
                                     N.ptr(stat_acc), n, h, w, N.stream()), "dpb_first_conv_fprop")
    N.count_launches(1)
    return ps, pq, blocks


def first_conv_on_tensor_cores(h: int, w: int) -> bool:
    return bool(N.require().dpb_first_conv_uses_tensor_cores(h, w))


def first_conv_wgrad(x_nchw: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor):
    lib = N.require()
    n, _, h, w = x_nchw.shape
    sc = scratch(dy.device)
    N.check(lib.dpb_first_conv_wgrad(N.ptr(x_nchw), N.ptr(dy), N.ptr(sc.fc_wgrad_part), N.ptr(dw), n, h, w,
                                     N.stream()), "dpb_first_conv_wgrad")
    N.count_launches(2)


def bn_finalize(part_sum, part_sq, parts: int, c: int, count: float, gamma, beta, running_mean, running_var, nbt,
                momentum: float, eps: float):
    lib = N.require()
    wait_buffer_sync(part_sum.device)
    dev = part_sum.device
    out = torch.empty((4, c), device=dev, dtype=torch.float32)  # scale, shift, mean, invstd
    N.check(lib.dpb_bn_finalize(N.ptr(part_sum), N.ptr(part_sq), parts, c, float(count), N.ptr(gamma), N.ptr(beta),
                                N.ptr(running_mean), N.ptr(running_var), N.ptr(nbt), momentum, eps, N.ptr(out[0]),
                                N.ptr(out[1]), N.ptr(out[2]), N.ptr(out[3]), N.stream()), "dpb_bn_finalize")
    N.count_launches(1)
    return out


def bn_eval_scale_shift(c: int, gamma, beta, running_mean, running_var, eps: float):
    lib = N.require()
    wait_buffer_sync(gamma.device)
    out = torch.empty((2, c), device=gamma.device, dtype=torch.float32)
    N.check(lib.dpb_bn_eval_scale_shift(c, N.ptr(gamma), N.ptr(beta), N.ptr(running_mean), N.ptr(running_var), eps,
                                        N.ptr(out[0]), N.ptr(out[1]), N.stream()), "dpb_bn_eval_scale_shift")
    N.count_launches(1)
    return out


def bn_relu_pool_fwd(y: torch.Tensor, scale, shift, pool: bool) -> torch.Tensor:
    lib = N.require()
    n, c, h, w = y.shape
    out = nhwc_empty(n, c, h // 2 if pool else h, w // 2 if pool else w, y.device)
    N.check(lib.dpb_bn_relu_pool_fwd(N.ptr(y), N.ptr(scale), N.ptr(shift), N.ptr(out), n, h, w, c, int(pool),
                                     N.stream()), "dpb_bn_relu_pool_fwd")
    N.count_launches(1)
    return out


# ---- DDP's per-forward buffer broadcast runs on a side stream; whoever touches BatchNorm buffers next waits for it ----
_buffer_sync_events: dict[int, "torch.cuda.Event"] = {}


def set_buffer_sync_event(device: torch.device, event) -> None:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    _buffer_sync_events[idx] = event


def wait_buffer_sync(device: torch.device) -> None:
    """Make the current stream wait for a pending BatchNorm-buffer broadcast on this device (once)."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    ev = _buffer_sync_events.pop(idx, None)
    if ev is not None:
        torch.cuda.current_stream(device).wait_event(ev)


def bn_relu_pool_fwd_train(y: torch.Tensor, acc: torch.Tensor, ticket: torch.Tensor, gamma, beta, running_mean,
                           running_var, nbt, momentum: float, eps: float, pool: bool):
    """Training-mode BatchNorm + ReLU (+ pool) straight from the accumulators the conv filled.  Returns (out, stats)
    with stats = (4, C): scale, shift, mean, invstd."""
    lib = N.require()
    n, c, h, w = y.shape
    wait_buffer_sync(y.device)  # running statistics are read-modify-written here
    out = nhwc_empty(n, c, h // 2 if pool else h, w // 2 if pool else w, y.device)
    stats = torch.empty((4, c), device=y.device, dtype=torch.float32)
    N.check(lib.dpb_bn_relu_pool_fwd_train(N.ptr(y), N.ptr(acc), N.ptr(ticket), N.ptr(gamma), N.ptr(beta),
                                           N.ptr(running_mean), N.ptr(running_var), N.ptr(nbt), momentum, eps,
                                           N.ptr(stats), N.ptr(out), n, h, w, c, int(pool), N.stream()),
            "dpb_bn_relu_pool_fwd_train")
    N.count_launches(1)
    return out, stats


def bn_relu_pool_bwd(y, dout, stats, gamma, dgamma, dbeta, conv_dbias, pool: bool, use_acc: bool = True) -> torch.Tensor:
    """stats = (4, C): scale, shift, mean, invstd. Writes dgamma/dbeta, returns dY (NHWC, conv resolution).
    ``use_acc``: two launches through the fixed-point accumulators; else reduce / finalize / apply with scratch rows."""
    lib = N.require()
    n, c, h, w = y.shape
    sc = scratch(y.device)
    dy = torch.empty_like(y)
    acc = ticket = None
    if use_acc and c % 32 == 0:
        acc, ticket = sc.next_acc()
    N.check(lib.dpb_bn_relu_pool_bwd(N.ptr(y), N.ptr(dout), N.ptr(stats[0]), N.ptr(stats[1]), N.ptr(stats[2]),
                                     N.ptr(stats[3]), N.ptr(gamma), N.ptr(sc.bn_part), N.ptr(sc.bn_coef),
                                     N.ptr(acc), N.ptr(ticket), N.ptr(dgamma), N.ptr(dbeta), N.ptr(conv_dbias),
                                     N.ptr(dy), n, h, w, c, int(pool), N.stream()),
            "dpb_bn_relu_pool_bwd")
    N.count_launches(lib.dpb_bn_bwd_launches(c, int(acc is not None)))
    return dy


# ---- the reference's layer types on their own (anything that is not conv -> bn -> relu [-> pool]) ----------------------
def relu_fwd(x: torch.Tensor) -> torch.Tensor:
    lib = N.require()
    y = torch.empty_like(x)
    N.check(lib.dpb_relu_fwd(N.ptr(x), N.ptr(y), x.numel(), N.stream()), "dpb_relu_fwd")
    N.count_launches(1)
    return y


def relu_bwd(y: torch.Tensor, dout: torch.Tensor) -> torch.Tensor:
    lib = N.require()
    dx = torch.empty_like(y)
    N.check(lib.dpb_relu_bwd(N.ptr(y), N.ptr(dout), N.ptr(dx), y.numel(), N.stream()), "dpb_relu_bwd")
    N.count_launches(1)
    return dx


def maxpool2x2_fwd(x: torch.Tensor) -> torch.Tensor:
    lib = N.require()
    n, c, h, w = x.shape
    out = nhwc_empty(n, c, h // 2, w // 2, x.device)
    N.check(lib.dpb_maxpool2x2_fwd(N.ptr(x), N.ptr(out), n, h, w, c, N.stream()), "dpb_maxpool2x2_fwd")
    N.count_launches(1)
    return out


def maxpool2x2_bwd(x: torch.Tensor, dout: torch.Tensor) -> torch.Tensor:
    lib = N.require()
    n, c, h, w = x.shape
    dx = torch.empty_like(x)
    N.check(lib.dpb_maxpool2x2_bwd(N.ptr(x), N.ptr(dout), N.ptr(dx), n, h, w, c, N.stream()), "dpb_maxpool2x2_bwd")
    N.count_launches(1)
    return dx


def column_stats(x: torch.Tensor):
    """Per-block partial sums / sums of squares over the pixels of an NHWC tensor: (part_sum, part_sq, rows)."""
    lib = N.require()
    n, c, h, w = x.shape
    blocks = lib.dpb_column_stats_blocks()
    ps = torch.empty((blocks, c), device=x.device, dtype=torch.float32)
    pq = torch.empty((blocks, c), device=x.device, dtype=torch.float32)
    rows = ctypes.c_int(0)
    N.check(lib.dpb_column_stats(N.ptr(x), N.ptr(ps), N.ptr(pq), n * h * w, c, ctypes.byref(rows), N.stream()),
            "dpb_column_stats")
    N.count_launches(1)
    return ps, pq, rows.value


def column_sum(x: torch.Tensor, out: torch.Tensor) -> None:
    """out[c] = sum over pixels of an NHWC tensor (bias gradient of a conv that is not followed by BatchNorm)."""
    lib = N.require()
    n, c, h, w = x.shape
    part = torch.empty((lib.dpb_column_stats_blocks(), c), device=x.device, dtype=torch.float32)
    N.check(lib.dpb_column_sum(N.ptr(x), N.ptr(part), N.ptr(out), n * h * w, c, N.stream()), "dpb_column_sum")
    N.count_launches(2)


def bn_apply(x: torch.Tensor, scale, shift) -> torch.Tensor:
    lib = N.require()
    n, c, h, w = x.shape
    y = torch.empty_like(x)
    N.check(lib.dpb_bn_apply(N.ptr(x), N.ptr(scale), N.ptr(shift), N.ptr(y), n * h * w, c, N.stream()), "dpb_bn_apply")
    N.count_launches(1)
    return y


def bn_plain_bwd(x, dout, stats, gamma, dgamma, dbeta) -> torch.Tensor:
    """BatchNorm backward without ReLU; stats = (4, C) tensor from bn_finalize."""
    lib = N.require()
    n, c, h, w = x.shape
    sc = scratch(x.device)
    dx = torch.empty_like(x)
    N.check(lib.dpb_bn_plain_bwd(N.ptr(x), N.ptr(dout), N.ptr(stats[2]), N.ptr(stats[3]), N.ptr(gamma), N.ptr(sc.bn_part),
                                 N.ptr(sc.bn_coef), N.ptr(dgamma), N.ptr(dbeta), N.ptr(dx), n * h * w, c, N.stream()),
            "dpb_bn_plain_bwd")
    N.count_launches(3)
    return dx


def linear_fwd(x, w, b) -> torch.Tensor:
    lib = N.require()
    bsz, k = x.shape
    o = w.shape[0]
    out = torch.empty((bsz, o), device=x.device, dtype=torch.float32)
    N.check(lib.dpb_linear_fwd(N.ptr(x), N.ptr(w), N.ptr(b), N.ptr(out), bsz, k, o, N.stream()), "dpb_linear_fwd")
    N.count_launches(1)
    return out


def linear_bwd(dz, x, w, dx, dw, db) -> None:
    lib = N.require()
    bsz, k = x.shape
    o = w.shape[0]
    N.check(lib.dpb_linear_bwd(N.ptr(dz), N.ptr(x), N.ptr(w), N.ptr(dx), N.ptr(dw), N.ptr(db), bsz, k, o, N.stream()),
            "dpb_linear_bwd")
    N.count_launches((1 if dx is not None else 0) + (1 if dw is not None else 0))


def cross_entropy_fwd(logits, target, metrics=None) -> torch.Tensor:
    lib = N.require()
    bsz, o = logits.shape
    loss = torch.empty((), device=logits.device, dtype=torch.float32)
    N.check(lib.dpb_cross_entropy_fwd(N.ptr(logits), N.ptr(target), N.ptr(loss), N.ptr(metrics), bsz, o, N.stream()),
            "dpb_cross_entropy_fwd")
    N.count_launches(1)
    return loss


def cross_entropy_bwd(logits, target, grad_out) -> torch.Tensor:
    lib = N.require()
    bsz, o = logits.shape
    dz = torch.empty_like(logits)
    N.check(lib.dpb_cross_entropy_bwd(N.ptr(logits), N.ptr(target), N.ptr(grad_out), N.ptr(dz), bsz, o, N.stream()),
            "dpb_cross_entropy_bwd")
    N.count_launches(1)
    return dz


def head_supported(k: int, o: int) -> bool:
    return bool(N.require().dpb_head_supported(k, o))


def head_fwd(x, w, b, target, metrics=None, want_dz: bool = True):
    """Linear + mean cross-entropy (+ dz = dLoss/dlogits for a unit upstream gradient) in ONE launch.
    Returns (loss, logits, dz)."""
    lib = N.require()
    bsz, k = x.shape
    o = w.shape[0]
    sc = scratch(x.device)
    logits = torch.empty((bsz, o), device=x.device, dtype=torch.float32)
    dz = torch.empty((bsz, o), device=x.device, dtype=torch.float32) if want_dz else None
    terms = torch.empty((bsz, 2), device=x.device, dtype=torch.float32)
    loss = torch.empty((), device=x.device, dtype=torch.float32)
    N.check(lib.dpb_head_fwd(N.ptr(x), N.ptr(w), N.ptr(b), N.ptr(target), N.ptr(logits), N.ptr(dz), N.ptr(terms),
                             N.ptr(loss), N.ptr(metrics), N.ptr(sc.head_counter), bsz, k, o, N.stream()), "dpb_head_fwd")
    N.count_launches(1)
    return loss, logits, dz


def head_bwd(dz, x, w, grad_out, dx, dw, db) -> None:
    lib = N.require()
    bsz, k = x.shape
    o = w.shape[0]
    N.check(lib.dpb_head_bwd(N.ptr(dz), N.ptr(x), N.ptr(w), N.ptr(grad_out), N.ptr(dx), N.ptr(dw), N.ptr(db), bsz, k, o,
                             N.stream()), "dpb_head_bwd")
    N.count_launches(1)


def augment_batch(images_u8: torch.Tensor, labels: torch.Tensor, indices: torch.Tensor, train: bool, seed: int,
                  epoch: int, draws: Optional[torch.Tensor] = None):
    """One launch: gather + RandomCrop(pad 4) + flip + Normalize on the device.  images_u8: (N, 3, H, W) uint8 on the
    GPU, indices: int32 device tensor.  Returns (x fp32 NCHW, y int64)."""
    from ..data.datasets import CIFAR_MEAN, CIFAR_STD
    lib = N.require()
    b = indices.numel()
    _, _, h, w = images_u8.shape
    x = torch.empty((b, 3, h, w), device=images_u8.device, dtype=torch.float32)
    y = torch.empty((b,), device=images_u8.device, dtype=torch.int64)
    inv = (1.0 / torch.tensor(CIFAR_STD, dtype=torch.float32)).tolist()  # the same rounding as the host transform
    mean = torch.tensor(CIFAR_MEAN, dtype=torch.float32).tolist()
    N.check(lib.dpb_augment_batch(N.ptr(images_u8), N.ptr(labels), N.ptr(indices), b, h, w, int(train),
                                  seed & 0xFFFFFFFF, epoch & 0xFFFFFFFF, N.ptr(draws), mean[0], mean[1], mean[2],
                                  inv[0], inv[1], inv[2], N.ptr(x), N.ptr(y), N.stream()), "dpb_augment_batch")
    N.count_launches(1)
    return x, y


def sgd_step(p, g, m, n: int, lr: float, momentum: float, wd: float, grad_scale: float = 1.0,
             nesterov: bool = False) -> None:
    lib = N.require()
    N.check(lib.dpb_sgd_step(N.ptr(p), N.ptr(g), N.ptr(m), n, lr, momentum, wd, grad_scale, int(nesterov),
                             N.stream()), "dpb_sgd_step")
    N.count_launches(1)
