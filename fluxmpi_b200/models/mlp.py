"""Quick-Start MLP: ``Chain(Dense(1=>256,tanh), Dense(256=>512,tanh), Dense(512=>256,tanh),
Dense(256=>1))`` — /root/reference/README.md:37-38 (263,681 parameters; BASELINE config 1)."""
from __future__ import annotations

import torch
from torch import nn


def quickstart_mlp(in_dim: int = 1, widths=(256, 512, 256), out_dim: int = 1) -> nn.Sequential:
    dims = [in_dim, *widths]
    layers: list[nn.Module] = []
    for a, b in zip(dims[:-1], dims[1:]):
        layers += [nn.Linear(a, b), nn.Tanh()]
    layers.append(nn.Linear(dims[-1], out_dim))
    return nn.Sequential(*layers)


def functional_forward(params: dict, x: torch.Tensor) -> torch.Tensor:
    """Forward pass of :func:`quickstart_mlp` from an explicit parameter tree
    (``model(x, ps, st)`` in Lux, README.md:54)."""
    keys = sorted({k.split(".")[0] for k in params}, key=int)
    h = x
    for i, k in enumerate(keys):
        h = torch.nn.functional.linear(h, params[f"{k}.weight"], params[f"{k}.bias"])
        if i + 1 < len(keys):
            h = torch.tanh(h)
    return h
