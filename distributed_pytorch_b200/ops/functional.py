"""Differentiable ops of the framework.

On CUDA tensors every op is a ``torch.autograd.Function`` over the sm_100a kernels in ``libdpb200.so`` (no cuDNN, no
cuBLAS, no Triton).  On CPU tensors the same functions evaluate the plain PyTorch fp32 formulas: that path is the
numerical oracle of the tests and what ``main.py`` runs for the reference's CPU configuration; it is never taken for a
CUDA tensor.

The unit of fusion is the reference's repeating block ``Conv2d(3x3) -> BatchNorm2d -> ReLU [-> MaxPool2d(2,2)]``
(reference model.py:14-26).

Gradient placement: when a parameter has been adopted by a :class:`~distributed_pytorch_b200.utils.arena.FlatArena`,
its gradient is written straight into the flat (NVLink-symmetric) gradient arena and the returned tensor is a fresh
view of that memory.  Autograd's ``AccumulateGrad`` adopts such a tensor without copying when ``param.grad`` is None
(the reference calls ``optimizer.zero_grad()`` every iteration, main.py:32), so the "bucket" *is* the gradient memory.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import kernels as K

__all__ = ["conv_bn_relu_pool", "conv3x3", "batch_norm", "relu", "max_pool2x2", "linear", "cross_entropy",
           "linear_cross_entropy", "grad_destination"]


def grad_destination(param: torch.Tensor) -> torch.Tensor:
    """Where the gradient of ``param`` should be written by a backward kernel."""
    slot = getattr(param, "_dpb_grad_slot", None)
    if slot is not None:
        g = param.grad
        view = slot()
        if g is None or g.data_ptr() != view.data_ptr():
            return view
        # param.grad already lives in the arena (gradient accumulation): autograd will add, so use a temporary
    return torch.empty_strided(param.shape, param.stride(), device=param.device, dtype=param.dtype)


def _as_nhwc(t: torch.Tensor) -> torch.Tensor:
    return t if K.is_nhwc(t) else t.contiguous(memory_format=torch.channels_last)


def _as_krsc(w: torch.Tensor) -> torch.Tensor:
    return w if K.is_nhwc(w) else w.contiguous(memory_format=torch.channels_last)


class _ConvBnReluPool(torch.autograd.Function):
    """Fused block, CUDA.  Saves only X, the raw conv output Y and four per-channel vectors."""

    @staticmethod
    def forward(ctx, x, conv_w, conv_b, bn_w, bn_b, running_mean, running_var, nbt, training, momentum, eps, pool):
        n, cin, h, w = x.shape
        cout = conv_w.shape[0]
        first = cin == 3
        wk = _as_krsc(conv_w.detach())
        if first:
            if cout != 64:
                raise NotImplementedError("the Cin=3 entry layer kernel is built for 64 output channels")
            xin = x.detach().contiguous()  # NCHW as delivered by the loader
        else:
            xin = _as_nhwc(x.detach())
        y = K.nhwc_empty(n, cout, h, w, x.device)
        if training:
            # conv kernels add the per-channel sums to fixed-point accumulators; the BatchNorm kernel derives scale and
            # shift from them in its prologue (no partial rows, no finalize launch) and clears them again
            if first and not K.first_conv_on_tensor_cores(h, w):
                ps, pq, parts = K.first_conv_fprop(xin, wk, conv_b, y, True)
                stats = K.bn_finalize(ps, pq, parts, cout, float(n * h * w), bn_w, bn_b, running_mean, running_var, nbt,
                                      momentum, eps)
                out = K.bn_relu_pool_fwd(y, stats[0], stats[1], pool)
            else:
                acc, ticket = K.scratch(x.device).next_acc()
                if first:
                    K.first_conv_fprop(xin, wk, conv_b, y, False, stat_acc=acc)
                else:
                    K.conv3x3_gemm(0, n, h, w, cin, cout, xin, None, wk, y, conv_b, stat_acc=acc)
                out, stats = K.bn_relu_pool_fwd_train(y, acc, ticket, bn_w, bn_b, running_mean, running_var, nbt,
                                                      momentum, eps, pool)
        else:
            if first:
                K.first_conv_fprop(xin, wk, conv_b, y, False)
            else:
                K.conv3x3_gemm(0, n, h, w, cin, cout, xin, None, wk, y, conv_b)
            stats = K.bn_eval_scale_shift(cout, bn_w, bn_b, running_mean, running_var, eps)
            out = K.bn_relu_pool_fwd(y, stats[0], stats[1], pool)
        ctx.pool = pool
        ctx.first = first
        ctx.training = training
        ctx.params = (conv_w, conv_b, bn_w, bn_b)
        ctx.save_for_backward(xin, wk, y, stats)
        return out

    @staticmethod
    def backward(ctx, dout):
        xin, wk, y, stats = ctx.saved_tensors
        conv_w, conv_b, bn_w, bn_b = ctx.params
        if not ctx.training:
            raise RuntimeError("backward through an eval-mode BatchNorm block is not supported")
        n, cout, h, w = y.shape
        cin = conv_w.shape[1]
        dout = _as_nhwc(dout) if dout.dim() == 4 else dout.contiguous()
        d_bn_w = grad_destination(bn_w)
        d_bn_b = grad_destination(bn_b)
        d_conv_b = grad_destination(conv_b) if conv_b is not None else None
        dy = K.bn_relu_pool_bwd(y, dout, stats, bn_w, d_bn_w, d_bn_b, d_conv_b, ctx.pool)
        d_conv_w = grad_destination(conv_w)
        if not K.is_nhwc(d_conv_w):
            raise RuntimeError("conv weight gradients must be KRSC (channels_last) in memory")
        dx = None
        if ctx.first:
            K.first_conv_wgrad(xin, dy, d_conv_w)
        else:
            # the weight gradient feeds nothing further down the chain: run it next to the data gradient (and the
            # previous block's BatchNorm backward) on the device's second stream.  Only when autograd will adopt the
            # result without launching an accumulation kernel on this stream (param.grad is None).
            side_by_side = K.wgrad_overlap() and conv_w.grad is None and ctx.needs_input_grad[0]
            if side_by_side:
                lane = K.wgrad_lane(y.device)
                lane.fork()
                with torch.cuda.stream(lane.stream):
                    K.conv3x3_gemm(2, n, h, w, cin, cout, dy, xin, wk, d_conv_w, side=True,
                                   target_ctas=K.BWD_TARGET_CTAS)
                lane.hold(dy, xin)  # not d_conv_w: AccumulateGrad only adopts a gradient nobody else references
            else:
                K.conv3x3_gemm(2, n, h, w, cin, cout, dy, xin, wk, d_conv_w)
            if ctx.needs_input_grad[0]:
                dx = K.nhwc_empty(n, cin, h, w, y.device)
                K.conv3x3_gemm(1, n, h, w, cin, cout, dy, None, wk, dx,
                               target_ctas=K.BWD_TARGET_CTAS if side_by_side else 0)
        return dx, d_conv_w, d_conv_b, d_bn_w, d_bn_b, None, None, None, None, None, None, None


def conv_bn_relu_pool(x, conv_w, conv_b, bn_w, bn_b, running_mean, running_var, num_batches_tracked,
                      training: bool, momentum: float = 0.1, eps: float = 1e-5, pool: bool = False):
    """Conv3x3(s1,p1,bias) -> BatchNorm2d -> ReLU [-> MaxPool2d(2,2)].  (reference model.py:16-25)"""
    if x.is_cuda:
        return _ConvBnReluPool.apply(x, conv_w, conv_b, bn_w, bn_b, running_mean, running_var, num_batches_tracked,
                                     training, momentum, eps, pool)
    # CPU oracle
    y = F.conv2d(x, conv_w, conv_b, stride=1, padding=1)
    if training and num_batches_tracked is not None:
        num_batches_tracked += 1
    y = F.batch_norm(y, running_mean, running_var, bn_w, bn_b, training, momentum, eps)
    y = F.relu(y)
    if pool:
        y = F.max_pool2d(y, 2, 2)
    return y


# ---- the same layer types on their own (reference model.py:16-25 in any other arrangement) ----------------------------
class _Conv3x3(torch.autograd.Function):
    """Conv2d(3x3, stride 1, padding 1, bias) not followed by BatchNorm: tcgen05 fprop / dgrad / wgrad, bias gradient by a
    fixed-order column sum."""

    @staticmethod
    def forward(ctx, x, conv_w, conv_b):
        n, cin, h, w = x.shape
        cout = conv_w.shape[0]
        first = cin == 3
        wk = _as_krsc(conv_w.detach())
        y = K.nhwc_empty(n, cout, h, w, x.device)
        if first:
            if cout != 64:
                raise NotImplementedError("the Cin=3 entry layer kernel is built for 64 output channels")
            xin = x.detach().contiguous()
            K.first_conv_fprop(xin, wk, conv_b, y, False)
        else:
            xin = _as_nhwc(x.detach())
            K.conv3x3_gemm(0, n, h, w, cin, cout, xin, None, wk, y, conv_b)
        ctx.first = first
        ctx.params = (conv_w, conv_b)
        ctx.save_for_backward(xin, wk)
        return y

    @staticmethod
    def backward(ctx, dy):
        xin, wk = ctx.saved_tensors
        conv_w, conv_b = ctx.params
        dy = _as_nhwc(dy)
        n, cout, h, w = dy.shape
        cin = conv_w.shape[1]
        d_w = grad_destination(conv_w)
        if not K.is_nhwc(d_w):
            raise RuntimeError("conv weight gradients must be KRSC (channels_last) in memory")
        d_b = None
        if conv_b is not None:
            d_b = grad_destination(conv_b)
            K.column_sum(dy, d_b)
        dx = None
        if ctx.first:
            if ctx.needs_input_grad[0]:
                raise NotImplementedError("no data-gradient kernel for the Cin=3 entry layer (the network input)")
            K.first_conv_wgrad(xin, dy, d_w)
        else:
            K.conv3x3_gemm(2, n, h, w, cin, cout, dy, xin, wk, d_w)
            if ctx.needs_input_grad[0]:
                dx = K.nhwc_empty(n, cin, h, w, dy.device)
                K.conv3x3_gemm(1, n, h, w, cin, cout, dy, None, wk, dx)
        return dx, d_w, d_b


class _BatchNorm(torch.autograd.Function):
    """BatchNorm2d on its own: batch statistics by a fixed-order column reduction, then y = x * scale + shift."""

    @staticmethod
    def forward(ctx, x, bn_w, bn_b, running_mean, running_var, nbt, training, momentum, eps):
        xin = _as_nhwc(x.detach())
        n, c, h, w = xin.shape
        if training:
            ps, pq, rows = K.column_stats(xin)
            stats = K.bn_finalize(ps, pq, rows, c, float(n * h * w), bn_w, bn_b, running_mean, running_var, nbt,
                                  momentum, eps)
        else:
            stats = K.bn_eval_scale_shift(c, bn_w, bn_b, running_mean, running_var, eps)
        y = K.bn_apply(xin, stats[0], stats[1])
        ctx.training = training
        ctx.params = (bn_w, bn_b)
        ctx.save_for_backward(xin, stats)
        return y

    @staticmethod
    def backward(ctx, dout):
        if not ctx.training:
            raise RuntimeError("backward through an eval-mode BatchNorm is not supported")
        xin, stats = ctx.saved_tensors
        bn_w, bn_b = ctx.params
        d_w, d_b = grad_destination(bn_w), grad_destination(bn_b)
        dx = K.bn_plain_bwd(xin, _as_nhwc(dout), stats, bn_w, d_w, d_b)
        return dx, d_w, d_b, None, None, None, None, None, None


class _ReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xin = x.detach()
        if not (xin.is_contiguous() or K.is_nhwc(xin)):
            xin = xin.contiguous()
        y = K.relu_fwd(xin)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dout):
        (y,) = ctx.saved_tensors
        if dout.stride() != y.stride():
            dout = dout.contiguous(memory_format=torch.channels_last) if K.is_nhwc(y) else dout.contiguous()
        return K.relu_bwd(y, dout)


class _MaxPool2x2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xin = _as_nhwc(x.detach())
        ctx.save_for_backward(xin)
        return K.maxpool2x2_fwd(xin)

    @staticmethod
    def backward(ctx, dout):
        (xin,) = ctx.saved_tensors
        return K.maxpool2x2_bwd(xin, _as_nhwc(dout))


def conv3x3(x, conv_w, conv_b):
    """Conv2d(3x3, s1, p1) alone (reference model.py:18-23)."""
    if x.is_cuda:
        return _Conv3x3.apply(x, conv_w, conv_b)
    return F.conv2d(x, conv_w, conv_b, stride=1, padding=1)


def batch_norm(x, bn_w, bn_b, running_mean, running_var, num_batches_tracked, training: bool, momentum: float = 0.1,
               eps: float = 1e-5):
    """BatchNorm2d alone (reference model.py:24)."""
    if x.is_cuda:
        return _BatchNorm.apply(x, bn_w, bn_b, running_mean, running_var, num_batches_tracked, training, momentum, eps)
    if training and num_batches_tracked is not None:
        num_batches_tracked += 1
    return F.batch_norm(x, running_mean, running_var, bn_w, bn_b, training, momentum, eps)


def relu(x):
    """ReLU alone (reference model.py:25)."""
    if x.is_cuda:
        if x.numel() % 4:
            raise NotImplementedError("the CUDA ReLU kernel works on multiples of 4 elements")
        return _ReLU.apply(x)
    return F.relu(x)


def max_pool2x2(x):
    """MaxPool2d(2, 2) alone (reference model.py:16)."""
    if x.is_cuda:
        return _MaxPool2x2.apply(x)
    return F.max_pool2d(x, 2, 2)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        xc = x.detach().contiguous()
        wc = w.detach().contiguous()
        out = K.linear_fwd(xc, wc, b)
        ctx.params = (w, b)
        ctx.save_for_backward(xc, wc)
        return out

    @staticmethod
    def backward(ctx, dz):
        xc, wc = ctx.saved_tensors
        w, b = ctx.params
        dz = dz.contiguous()
        dx = torch.empty_like(xc) if ctx.needs_input_grad[0] else None
        dw = grad_destination(w)
        db = grad_destination(b) if b is not None else None
        K.linear_bwd(dz, xc, wc, dx, dw, db)
        return dx, dw, db


def linear(x, w, b):
    """y = x W^T + b  (reference model.py:40,45)."""
    if x.is_cuda:
        return _Linear.apply(x, w, b)
    return F.linear(x, w, b)


class _HeadLoss(torch.autograd.Function):
    """Linear -> mean cross-entropy as two launches (forward, backward) instead of five: the forward kernel also leaves
    dz = dLoss/dlogits for a unit upstream gradient, the backward kernel scales by autograd's grad_output and produces
    dx, dW and db together."""

    @staticmethod
    def forward(ctx, x, w, b, target):
        xc = x.detach().contiguous()
        wc = w.detach().contiguous()
        tg = target.contiguous()
        loss, logits, dz = K.head_fwd(xc, wc, b, tg)
        ctx.params = (w, b)
        ctx.save_for_backward(xc, wc, dz)
        ctx.mark_non_differentiable(logits)
        return loss, logits

    @staticmethod
    def backward(ctx, gout, _glogits):
        xc, wc, dz = ctx.saved_tensors
        w, b = ctx.params
        dx = torch.empty_like(xc) if ctx.needs_input_grad[0] else None
        dw = grad_destination(w)
        db = grad_destination(b) if b is not None else None
        K.head_bwd(dz, xc, wc, gout.contiguous(), dx, dw, db)
        return dx, dw, db, None


def linear_cross_entropy(x, w, b, target):
    """``cross_entropy(linear(x, w, b), target)`` fused (reference model.py:45 + main.py:34); returns (loss, logits)."""
    if x.is_cuda and b is not None and K.head_supported(x.shape[1], w.shape[0]):
        if target.dtype != torch.int64:
            target = target.long()
        return _HeadLoss.apply(x, w, b, target)
    logits = linear(x, w, b)
    return cross_entropy(logits, target), logits


class _CrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        lg = logits.detach().contiguous()
        tg = target.contiguous()
        loss = K.cross_entropy_fwd(lg, tg)
        ctx.save_for_backward(lg, tg)
        return loss

    @staticmethod
    def backward(ctx, gout):
        lg, tg = ctx.saved_tensors
        return K.cross_entropy_bwd(lg, tg, gout.contiguous()), None


def cross_entropy(logits, target):
    """Mean-reduced softmax cross-entropy (reference main.py:34,99 ``CrossEntropyLoss()``)."""
    if logits.is_cuda:
        if target.dtype != torch.int64:
            target = target.long()
        return _CrossEntropy.apply(logits, target)
    return F.cross_entropy(logits, target)


def eval_metrics(logits, target, metrics: Optional[torch.Tensor]):
    """Adds the batch-mean loss to ``metrics[0]`` and the number of arg-max hits to ``metrics[1]`` on the device
    (reference main.py:59-61 does this with two ``.item()`` syncs per batch)."""
    if logits.is_cuda:
        return K.cross_entropy_fwd(logits.detach().contiguous(), target.long().contiguous(), metrics)
    loss = F.cross_entropy(logits, target)
    if metrics is not None:
        metrics[0] += loss
        metrics[1] += (logits.argmax(1) == target).sum()
    return loss
