"""Training / evaluation loops with the reference's logging protocol, plus the CUDA-graph train step.

``train_model`` / ``test_model`` keep the reference's signatures and printed lines (main.py:19-66):

* per-iteration wall time from the top of the loop body to after ``loss.item()``, iteration 0 excluded;
* every 20 iterations ``Epoch: E, Iteration: a-b, Average Loss: x.xxx``;
* every 40 iterations ``Avg Time for iteration a-b: t seconds.`` (first window divides by 39, later ones by 40);
* ``Test set: Average loss: x.xxxx, Accuracy: c/n (p%)`` with the loss averaged over batches.

:class:`TrainStep` is the B200 execution model for the loop body: zero_grad -> forward -> loss -> backward -> gradient
sync -> optimizer step is captured ONCE into a CUDA graph (every kernel of the step is ours and stream-ordered, the
DDP side stream included) and replayed per iteration, so the launch-bound small-batch regime (global batch 256 over 8
GPUs = 32 images per rank) costs one graph launch instead of ~80 kernel launches.
"""
from __future__ import annotations

import time
from datetime import datetime, timedelta
from typing import Callable, Optional

import torch

from .ops import functional as Fn
from .parallel.strategies import NoSync, SyncStrategy


class CrossEntropyLoss(torch.nn.Module):
    """``torch.nn.CrossEntropyLoss()`` (mean reduction) on our softmax-CE kernel (reference main.py:99)."""

    def forward(self, logits, target):
        return Fn.cross_entropy(logits, target)


class TrainStep:
    """One optimisation step as a callable: ``loss = step(data, target)`` (device scalar tensor).

    ``data`` / ``target`` may be host (pinned) or device tensors; they are copied into static device buffers, which
    is the host->device transfer of the step.  With ``use_graph`` (default on CUDA) the step body is captured after
    ``warmup`` eager iterations; batches of any other shape run eagerly.
    """

    def __init__(self, model: torch.nn.Module, optimizer, criterion: Optional[Callable] = None,
                 sync: Optional[SyncStrategy] = None, use_graph: Optional[bool] = None, warmup: int = 2):
        self.model = model
        self.optimizer = optimizer
        self.criterion = criterion or CrossEntropyLoss()
        self.sync = sync or NoSync(model)
        p = next(model.parameters())
        self.device = p.device
        self.cuda = p.is_cuda
        self.use_graph = self.cuda if use_graph is None else (use_graph and self.cuda)
        self.warmup = warmup
        self._eager_steps = 0
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._static_x = self._static_y = self._static_loss = None
        self._graph_key = None
        self.graph_launches = 0   # kernels of ours inside one replay
        self.replays = 0

    # -------------------------------------------------------------------------------------------------- body
    def _forward_loss(self, x, y):
        # model(x) -> criterion(out, y) like the reference loop (main.py:33-34); when the model offers it and the
        # criterion is our CrossEntropyLoss, the classifier and the loss run as one fused kernel each way
        fused = getattr(self.model, "forward_loss", None)
        if fused is not None and type(self.criterion) is CrossEntropyLoss and x.is_cuda:
            return fused(x, y)
        return self.criterion(self.model(x), y)

    def _body(self, x, y):
        nvtx = torch.cuda.nvtx if self.cuda else None  # ranges show up in nsys / ncu timelines
        self.optimizer.zero_grad(set_to_none=True)
        if nvtx:
            nvtx.range_push("dpb.forward")
        loss = self._forward_loss(x, y)
        if nvtx:
            nvtx.range_pop()
            nvtx.range_push("dpb.backward")
        loss.backward()
        if nvtx:
            nvtx.range_pop()
            nvtx.range_push("dpb.sync")
        self.sync.sync()
        if nvtx:
            nvtx.range_pop()
            nvtx.range_push("dpb.optimizer")
        self.optimizer.step()
        if nvtx:
            nvtx.range_pop()
        return loss

    def _hyper_key(self):
        g = self.optimizer.param_groups[0]
        return (g["lr"], g["momentum"], g["weight_decay"], g.get("nesterov", False))

    def _capture(self, x, y) -> None:
        from . import _native as N
        self._static_x = torch.empty_like(x, device=self.device)
        self._static_y = torch.empty_like(y, device=self.device)
        self._static_x.copy_(x)
        self._static_y.copy_(y)
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        before = N.launches()
        with torch.cuda.graph(graph):
            self._static_loss = self._body(self._static_x, self._static_y)
        self.graph_launches = N.launches() - before
        self._graph = graph
        self._graph_key = (tuple(x.shape), tuple(y.shape), self._hyper_key())

    def __call__(self, data: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        if not self.cuda:
            return self._body(data, target).detach()
        key = (tuple(data.shape), tuple(target.shape), self._hyper_key())
        if self.use_graph and self._graph is not None and key == self._graph_key:
            self._static_x.copy_(data, non_blocking=True)
            self._static_y.copy_(target, non_blocking=True)
            self._graph.replay()
            self.replays += 1
            from . import _native as N
            N.count_launches(self.graph_launches)
            return self._static_loss
        x = data.to(self.device, non_blocking=True)
        y = target.to(self.device, non_blocking=True)
        if self.use_graph and self._eager_steps >= self.warmup and (self._graph is None or key[2] != self._graph_key[2]
                                                                    and key[:2] == self._graph_key[:2]):
            # this very batch is consumed by the capture pass (capture executes nothing), then replayed once
            self._capture(x, y)
            self._graph.replay()
            self.replays += 1
            return self._static_loss
        self._eager_steps += 1
        return self._body(x, y).detach()


def train_model(model, train_loader, optimizer, criterion, epoch, args=None, sync: Optional[SyncStrategy] = None,
                step: Optional[TrainStep] = None, max_iters: Optional[int] = None, log=print):
    """One epoch in the reference's protocol (main.py:19-49; main_gather.py:22-75 adds ``args`` and the sync block)."""
    model.train()
    if step is None:
        step = TrainStep(model, optimizer, criterion, sync)
    running_loss = 0.0
    time_per_iteration = timedelta(seconds=0)
    for batch_idx, (data, target) in enumerate(train_loader):
        if max_iters is not None and batch_idx >= max_iters:
            break
        begin_time = datetime.now()
        try:
            loss = step(data, target)
            running_loss += loss.item()
        except RuntimeError as exc:
            # a kernel of ours that gave up (e.g. a peer that never reached the collective) leaves a record in host
            # memory: report that instead of the bare CUDA error
            from . import _native as N
            N.raise_if_device_error(exc)
            raise
        if batch_idx != 0:
            time_per_iteration += datetime.now() - begin_time
        if batch_idx % 20 == 19:
            log(f"Epoch: {epoch + 1}, Iteration: {batch_idx - 18}-{batch_idx + 1}, "
                f"Average Loss: {running_loss / 20:.3f}")
            running_loss = 0.0
        if batch_idx % 40 == 39:
            if batch_idx == 39:
                log(f"Avg Time for iteration {batch_idx - 37}-{batch_idx + 1}: "
                    f"{time_per_iteration.total_seconds() / 39} seconds.")
            else:
                log(f"Avg Time for iteration {batch_idx - 38}-{batch_idx + 1}: "
                    f"{time_per_iteration.total_seconds() / 40} seconds.")
            time_per_iteration = timedelta(seconds=0)
    return step


def test_model(model, test_loader, criterion=None, log=print):
    """Full-test-set evaluation, every rank on its own replica (main.py:51-66).  Loss is the mean of per-batch means;
    on CUDA the loss sum and hit count stay on the device until the end (one sync instead of two per batch)."""
    model.eval()
    p = next(model.parameters())
    metrics = torch.zeros(2, device=p.device, dtype=torch.float32)
    nb = 0
    with torch.no_grad():
        for data, target in test_loader:
            data = data.to(p.device, non_blocking=True)
            target = target.to(p.device, non_blocking=True)
            output = model(data)
            Fn.eval_metrics(output, target, metrics)
            nb += 1
    test_loss = metrics[0].item() / max(nb, 1)
    correct = int(round(metrics[1].item()))
    total = getattr(test_loader, "num_examples", None)
    if total is None:
        total = len(test_loader.dataset)
    log("Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(
        test_loss, correct, total, 100.0 * correct / total))
    return test_loss, correct


class DeviceTimer:
    """CUDA-event timing of a region on the current stream (device-timed, max over ranks)."""

    def __init__(self):
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.start.record()
        return self

    def __exit__(self, *exc):
        self.end.record()

    def ms(self) -> float:
        self.end.synchronize()
        return self.start.elapsed_time(self.end)


def wall() -> float:
    return time.perf_counter()
