"""Checkpoints in the plain ``torch.save(model.state_dict())`` format.

The reference never saves anything but its module tree fixes the schema: 58 keys for VGG-11
(``layers.{i}.weight|bias|running_mean|running_var|num_batches_tracked``, ``fc1.weight|bias``; ``module.``-prefixed
under DDP).  Files written here load into the reference's ``model.py`` with ``load_state_dict`` and vice versa:
tensors are saved as ordinary contiguous CPU tensors (our in-memory KRSC weight layout and arena views are an
execution detail and never leak into a file).  The optimizer file follows ``torch.optim.SGD.state_dict()``.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Optional

import torch


def portable_state_dict(module: torch.nn.Module, strip_module_prefix: bool = False) -> "OrderedDict[str, torch.Tensor]":
    out = OrderedDict()
    for k, v in module.state_dict().items():
        if strip_module_prefix and k.startswith("module."):
            k = k[len("module."):]
        out[k] = v.detach().to("cpu").contiguous().clone()
    return out


def save_checkpoint(path: str, model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None,
                    strip_module_prefix: bool = False) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(portable_state_dict(model, strip_module_prefix), path)
    if optimizer is not None:
        sd = optimizer.state_dict()
        for st in sd["state"].values():
            for k, v in list(st.items()):
                if torch.is_tensor(v):
                    st[k] = v.detach().to("cpu").contiguous().clone()
        torch.save(sd, path + ".optim")


def load_checkpoint(path: str, model: torch.nn.Module, optimizer: Optional[torch.optim.Optimizer] = None) -> None:
    sd = torch.load(path, map_location="cpu")
    own = model.state_dict()
    has_prefix = any(k.startswith("module.") for k in own)
    fixed = OrderedDict()
    for k, v in sd.items():
        if has_prefix and not k.startswith("module."):
            k = "module." + k
        elif not has_prefix and k.startswith("module."):
            k = k[len("module."):]
        fixed[k] = v
    missing = set(own) - set(fixed)
    unexpected = set(fixed) - set(own)
    if missing or unexpected:
        raise KeyError(f"checkpoint mismatch: missing {sorted(missing)}, unexpected {sorted(unexpected)}")
    with torch.no_grad():
        for k, dst in own.items():  # copy in place: parameters and buffers may be arena / symmetric-memory views
            dst.copy_(fixed[k])
    if optimizer is not None and os.path.exists(path + ".optim"):
        optimizer.load_state_dict(torch.load(path + ".optim", map_location="cpu"))
