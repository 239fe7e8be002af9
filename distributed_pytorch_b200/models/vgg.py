"""VGG family for 3x32x32 inputs and 10 classes, with the reference's exact module tree and ``state_dict`` schema
(reference model.py:1-50: ``layers`` Sequential of Conv2d / BatchNorm2d / ReLU / MaxPool2d, then ``fc1``; 34 parameters,
24 buffers, 58 keys for VGG-11 -- SURVEY.md §2.7).

The module classes below subclass the ``torch.nn`` containers so parameter initialisation consumes the RNG exactly
like the reference (same seed => same initial weights, which the reference's gather/all-reduce modes rely on instead
of a broadcast, main_gather.py:96).  What differs is execution: :class:`FusedSequential` walks the tree in
``conv, bn, relu[, pool]`` groups and runs each group as one fused op of sm_100a kernels
(:func:`distributed_pytorch_b200.ops.functional.conv_bn_relu_pool`); conv weights are kept KRSC
(``channels_last``) in memory so the TMA descriptors of the tcgen05 kernels read them in place.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops import functional as Fn

__all__ = ["VGG", "VGG11", "VGG13", "VGG16", "VGG19", "Conv3x3", "BatchNorm", "ReLU", "MaxPool2x2", "Linear",
           "FusedSequential", "CONFIGS"]

# ints = conv output channels, 'M' = 2x2 max-pool (same recipes as reference model.py:3-8)
CONFIGS = {
    "VGG11": (64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG13": (64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"),
    "VGG16": (64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"),
    "VGG19": (64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"),
}


class Conv3x3(nn.Conv2d):
    """3x3, stride 1, padding 1, bias (reference model.py:18-23). Weight memory is KRSC."""

    def __init__(self, in_channels: int, out_channels: int):
        super().__init__(in_channels, out_channels, kernel_size=3, stride=1, padding=1, bias=True)
        self.weight.data = self.weight.data.contiguous(memory_format=torch.channels_last)

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        if not self.weight.data.is_contiguous(memory_format=torch.channels_last):
            self.weight.data = self.weight.data.contiguous(memory_format=torch.channels_last)
        return out

    def forward(self, x):
        # on its own (not as the head of a conv->bn->relu[->pool] run of a FusedSequential): tcgen05 conv + bias
        return Fn.conv3x3(x, self.weight, self.bias)


class BatchNorm(nn.BatchNorm2d):
    def forward(self, x):
        if not x.is_cuda:
            return super().forward(x)
        training = self.training or self.running_mean is None
        momentum = 0.1 if self.momentum is None else self.momentum
        return Fn.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var,
                             self.num_batches_tracked if training else None, training, momentum, self.eps)


class ReLU(nn.ReLU):
    def forward(self, x):
        return Fn.relu(x) if x.is_cuda else super().forward(x)


class MaxPool2x2(nn.MaxPool2d):
    def __init__(self):
        super().__init__(kernel_size=2, stride=2)

    def forward(self, x):
        return Fn.max_pool2x2(x) if x.is_cuda else super().forward(x)


class Linear(nn.Linear):
    def forward(self, x):
        return Fn.linear(x, self.weight, self.bias)


class FusedSequential(nn.Sequential):
    """``nn.Sequential`` whose forward executes conv/bn/relu[/pool] runs as single fused ops."""

    def forward(self, x):
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            if (isinstance(m, nn.Conv2d) and i + 2 < len(mods) and isinstance(mods[i + 1], nn.BatchNorm2d)
                    and isinstance(mods[i + 2], nn.ReLU)):
                bn = mods[i + 1]
                pool = i + 3 < len(mods) and isinstance(mods[i + 3], nn.MaxPool2d)
                training = bn.training or bn.running_mean is None
                momentum = 0.1 if bn.momentum is None else bn.momentum
                x = Fn.conv_bn_relu_pool(x, m.weight, m.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var,
                                         bn.num_batches_tracked if training else None, training, momentum, bn.eps,
                                         pool)
                i += 4 if pool else 3
            else:
                x = m(x)
                i += 1
        return x


def make_layers(cfg) -> FusedSequential:
    layers = []
    cin = 3
    for item in cfg:
        if item == "M":
            layers.append(MaxPool2x2())
        else:
            layers += [Conv3x3(cin, item), BatchNorm(item), ReLU(inplace=True)]
            cin = item
    return FusedSequential(*layers)


class VGG(nn.Module):
    def __init__(self, name: str = "VGG11"):
        super().__init__()
        if name not in CONFIGS:
            raise KeyError(f"unknown VGG variant {name!r}; choose from {sorted(CONFIGS)}")
        self.name = name
        self.layers = make_layers(CONFIGS[name])
        self.fc1 = Linear(512, 10)

    def forward(self, x):
        y = self.layers(x)
        y = y.reshape(y.size(0), -1)
        return self.fc1(y)

    def forward_loss(self, x, target):
        """``CrossEntropyLoss()(self(x), target)`` with the classifier and the loss fused into one kernel each way
        (reference main.py:33-35: ``output = model(data); loss = criterion(output, target)``).  Returns the loss."""
        y = self.layers(x)
        y = y.reshape(y.size(0), -1)
        loss, _ = Fn.linear_cross_entropy(y, self.fc1.weight, self.fc1.bias, target)
        return loss


def VGG11() -> VGG:
    return VGG("VGG11")


def VGG13() -> VGG:
    return VGG("VGG13")


def VGG16() -> VGG:
    return VGG("VGG16")


def VGG19() -> VGG:
    return VGG("VGG19")
