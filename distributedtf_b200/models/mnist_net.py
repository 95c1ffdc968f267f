"""The MNIST CNN as a functional PyTorch model (oracle + restated baseline).

Parity: reference ``mnist_model.py:62-126`` — conv 5x5 SAME 1->32 + bias + ReLU,
max-pool 2x2/2, conv 5x5 SAME 32->64 + bias + ReLU, max-pool 2x2/2, flatten
(7*7*64 in H,W,C order), dense 1024 + ReLU, dropout 0.4 (train only), dense 10;
sparse softmax cross-entropy.  No BatchNorm, no regulariser, no LR decay.
Kernels are stored OHWI / [out, in] like the ResNet's.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from .resnet import ParamSpec

DROPOUT_RATE = 0.4


def param_specs() -> List[ParamSpec]:
    return [ParamSpec("conv0/kernel", (32, 5, 5, 1), "conv"), ParamSpec("conv0/bias", (32,), "dense_b"),
            ParamSpec("conv1/kernel", (64, 5, 5, 32), "conv"), ParamSpec("conv1/bias", (64,), "dense_b"),
            ParamSpec("dense0/kernel", (1024, 7 * 7 * 64), "dense_w"), ParamSpec("dense0/bias", (1024,), "dense_b"),
            ParamSpec("dense1/kernel", (10, 1024), "dense_w"), ParamSpec("dense1/bias", (10,), "dense_b")]


def forward(params: Dict[str, torch.Tensor], images_nchw: torch.Tensor, training: bool,
            dropout_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    p = params
    x = images_nchw
    x = F.relu(F.conv2d(x, p["conv0/kernel"].permute(0, 3, 1, 2).to(x.dtype), p["conv0/bias"].to(x.dtype), padding=2))
    x = F.max_pool2d(x, 2, 2)
    x = F.relu(F.conv2d(x, p["conv1/kernel"].permute(0, 3, 1, 2).to(x.dtype), p["conv1/bias"].to(x.dtype), padding=2))
    x = F.max_pool2d(x, 2, 2)
    x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)            # NHWC flatten order
    x = F.relu(F.linear(x, p["dense0/kernel"].to(x.dtype), p["dense0/bias"].to(x.dtype)))
    if training:
        if dropout_mask is not None:
            x = x * dropout_mask.to(x.dtype) / (1.0 - DROPOUT_RATE)
        else:
            x = F.dropout(x, DROPOUT_RATE, training=True)
    return F.linear(x, p["dense1/kernel"].to(x.dtype), p["dense1/bias"].to(x.dtype)).float()


def init_params(initializer: str, seed: int = 0, device="cpu") -> Dict[str, torch.Tensor]:
    from ..ops import reference
    gen = torch.Generator().manual_seed(seed)
    out = {}
    for s in param_specs():
        if s.kind in ("conv", "dense_w"):
            # conv1/conv2/dense-1024 take the HP initializer; the logits layer keeps the default
            init = initializer if s.name != "dense1/kernel" else "None"
            out[s.name] = reference.init_kernel(s.shape, init, gen).to(device)
        else:
            out[s.name] = torch.zeros(s.shape, device=device)
    return out
