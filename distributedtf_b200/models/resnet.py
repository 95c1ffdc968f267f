"""ResNet v1/v2 (building + bottleneck blocks) as a *functional* PyTorch model.

Parity: reference ``resnet/resnet_model.py`` — ``batch_norm`` (:45-52, decay 0.997,
eps 1e-5), ``fixed_padding``/``conv2d_fixed_padding`` (:55-92: SAME for stride 1,
explicit pad (k-1)//2 | rest then VALID for stride > 1, no bias), the four block
functions (:127-320), ``block_layer`` (:323-359: the first block of *every* stage
gets a 1x1 projection shortcut, even stage 1 at stride 1), ``Model.__call__``
(:487-554) and the CIFAR configuration ``cifar10_main.py:146-185`` (6n+2 layers,
16 filters, strides [1,2,2], final 64, no max-pool).

This file is (a) the numerical oracle for the sm_100a kernels, (b) the
"restated baseline" engine (torch eager + cuDNN) and (c) the single source of
truth for the parameter list/order that the flat arena uses.  Parameters are
plain tensors in a dict, stored **OHWI** (``[out, kh, kw, in]`` — the GEMM "B"
matrix ``[N][K]`` the implicit-GEMM kernels consume); activations are NCHW here
(torch) and NHWC in the native engine.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

BATCH_NORM_DECAY = 0.997
BATCH_NORM_EPSILON = 1e-5


@dataclass(frozen=True)
class ResNetConfig:
    resnet_size: int = 32
    bottleneck: bool = False
    num_classes: int = 10
    num_filters: int = 16
    kernel_size: int = 3
    conv_stride: int = 1
    first_pool_size: Optional[int] = None
    first_pool_stride: Optional[int] = None
    block_sizes: Tuple[int, ...] = (5, 5, 5)
    block_strides: Tuple[int, ...] = (1, 2, 2)
    final_size: int = 64
    resnet_version: int = 2
    in_channels: int = 3
    image_size: int = 32

    @property
    def pre_activation(self) -> bool:
        return self.resnet_version == 2


def cifar10_config(resnet_size: int = 32, resnet_version: int = 2, num_classes: int = 10,
                   bottleneck: bool = False) -> ResNetConfig:
    """The reference's CIFAR-10 model (``cifar10_main.py:167-199``: building blocks, 6n + 2 layers).  With
    ``bottleneck=True`` the same three stages are built from bottleneck blocks (``resnet_model.py:215-320``; 9n + 2
    layers, the dense layer sees 4 x 64 channels) -- the reference's ``Model`` takes the switch, its CIFAR main
    never sets it."""
    per_block = 9 if bottleneck else 6
    if resnet_size % per_block != 2:
        raise ValueError("resnet_size must be {}n + 2:".format(per_block), resnet_size)
    if resnet_version not in (1, 2):
        raise ValueError("Resnet version should be 1 or 2.")
    n = (resnet_size - 2) // per_block
    return ResNetConfig(resnet_size=resnet_size, bottleneck=bottleneck, block_sizes=(n, n, n),
                        final_size=256 if bottleneck else 64, resnet_version=resnet_version, num_classes=num_classes)


@dataclass(frozen=True)
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    kind: str                 # conv | dense_w | dense_b | bn_gamma | bn_beta | bn_mean | bn_var
    regularized: bool = False  # conv kernels only (resnet_model.py:91)
    trainable: bool = True

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


class _Builder:
    """Walks the architecture once; used both to list parameters and to run it."""

    def __init__(self, cfg: ResNetConfig):
        self.cfg = cfg
        self.specs: List[ParamSpec] = []
        self._bn = 0
        self._conv = 0

    def conv(self, cin: int, cout: int, k: int) -> str:
        name = f"conv{self._conv}"
        self._conv += 1
        self.specs.append(ParamSpec(name + "/kernel", (cout, k, k, cin), "conv", regularized=True))
        return name

    def bn(self, c: int) -> str:
        name = f"bn{self._bn}"
        self._bn += 1
        self.specs += [ParamSpec(name + "/gamma", (c,), "bn_gamma"),
                       ParamSpec(name + "/beta", (c,), "bn_beta"),
                       ParamSpec(name + "/moving_mean", (c,), "bn_mean", trainable=False),
                       ParamSpec(name + "/moving_variance", (c,), "bn_var", trainable=False)]
        return name


def _walk(cfg: ResNetConfig, ops: "_Ops") -> Any:
    """Architecture as a sequence of calls on ``ops`` (list-params or execute)."""
    x = ops.input()
    x = ops.conv(x, cfg.in_channels, cfg.num_filters, cfg.kernel_size, cfg.conv_stride)
    if cfg.resnet_version == 1:
        x = ops.relu(ops.bn(x, cfg.num_filters))
    if cfg.first_pool_size:
        x = ops.maxpool(x, cfg.first_pool_size, cfg.first_pool_stride)
    cin = cfg.num_filters
    for i, nblocks in enumerate(cfg.block_sizes):
        filters = cfg.num_filters * (2 ** i)
        fout = filters * 4 if cfg.bottleneck else filters
        for b in range(nblocks):
            stride = cfg.block_strides[i] if b == 0 else 1
            x = _block(cfg, ops, x, cin, filters, fout, stride, projection=(b == 0))
            cin = fout
    if cfg.pre_activation:
        x = ops.relu(ops.bn(x, cin))
    x = ops.mean_pool(x)
    return ops.dense(x, cfg.final_size, cfg.num_classes)


def _block(cfg, ops, x, cin, filters, fout, stride, projection):
    v2 = cfg.resnet_version == 2
    shortcut = x
    if v2:
        x = ops.relu(ops.bn(x, cin))
        if projection:
            shortcut = ops.conv(x, cin, fout, 1, stride)
    elif projection:
        shortcut = ops.bn(ops.conv(x, cin, fout, 1, stride), fout)
    if cfg.bottleneck:
        chain = [(cin, filters, 1, 1), (filters, filters, 3, stride), (filters, fout, 1, 1)]
    else:
        chain = [(cin, filters, 3, stride), (filters, filters, 3, 1)]
    for j, (ci, co, k, s) in enumerate(chain):
        last = j == len(chain) - 1
        if last and v2:
            ops.fused_add_next = True      # engine: conv epilogue adds the shortcut before rounding
        x = ops.conv(x, ci, co, k, s)
        if v2:
            if not last:
                x = ops.relu(ops.bn(x, co))
        else:
            x = ops.bn(x, co)
            if not last:
                x = ops.relu(x)
    x = ops.add(x, shortcut)
    return x if v2 else ops.relu(x)


class _Ops:
    def input(self): ...


class _SpecOps(_Ops):
    def __init__(self, cfg):
        self.b = _Builder(cfg)

    def input(self): return None
    def conv(self, x, cin, cout, k, s): self.b.conv(cin, cout, k); return None
    def bn(self, x, c): self.b.bn(c); return None
    def relu(self, x): return None
    def maxpool(self, x, k, s): return None
    def add(self, a, b): return None
    def mean_pool(self, x): return None

    def dense(self, x, cin, cout):
        self.b.specs += [ParamSpec("dense/kernel", (cout, cin), "dense_w"),
                         ParamSpec("dense/bias", (cout,), "dense_b")]
        return None


def param_specs(cfg: ResNetConfig) -> List[ParamSpec]:
    ops = _SpecOps(cfg)
    _walk(cfg, ops)
    return ops.b.specs


def conv2d_fixed_padding(x: torch.Tensor, w_ohwi: torch.Tensor, stride: int) -> torch.Tensor:
    k = w_ohwi.shape[1]
    w = w_ohwi.permute(0, 3, 1, 2)
    if stride > 1:
        total = k - 1
        beg = total // 2
        x = F.pad(x, (beg, total - beg, beg, total - beg))
        return F.conv2d(x, w, stride=stride)
    return F.conv2d(x, w, stride=1, padding=k // 2)


def batch_norm(x: torch.Tensor, gamma, beta, mean, var, training: bool,
               decay: float = BATCH_NORM_DECAY, eps: float = BATCH_NORM_EPSILON) -> torch.Tensor:
    """Fused-BN semantics: batch statistics (biased variance) when training, with the
    moving averages updated in place using the unbiased variance; moving statistics
    in inference."""
    return F.batch_norm(x, mean, var, gamma, beta, training=training, momentum=1.0 - decay, eps=eps)


class _ExecOps(_Ops):
    def __init__(self, cfg, params: Dict[str, torch.Tensor], x: torch.Tensor, training: bool,
                 taps: Optional[Dict[str, torch.Tensor]] = None, emulate_bf16: bool = False,
                 relu_masks: Optional[Sequence[torch.Tensor]] = None):
        self.cfg, self.p, self.x, self.training = cfg, params, x, training
        self._bn = self._conv = self._add = self._relu = 0
        self.taps = taps
        self.emulate_bf16 = emulate_bf16
        self.relu_masks = relu_masks

    def _tap(self, name, t):
        if self.emulate_bf16 and name.startswith("conv") and getattr(self, "fused_add_next", False):
            self.fused_add_next = False
        elif self.emulate_bf16 and (name.startswith("conv") or name.startswith("add")):
            # the native engine stores these tensors in bf16: round with a straight-through gradient
            t = t + (t.bfloat16().float() - t).detach()
        if self.taps is not None:
            self.taps[name] = t
        return t

    def input(self): return self.x

    def conv(self, x, cin, cout, k, s):
        name = f"conv{self._conv}"
        self._conv += 1
        w = self.p[name + "/kernel"].to(x.dtype)
        if self.emulate_bf16:   # bf16 operands, fp32 accumulation (activations after BN+ReLU are re-rounded)
            w = w + (w.bfloat16().float() - w).detach()
            x = x + (x.bfloat16().float() - x).detach()
        return self._tap(name, conv2d_fixed_padding(x, w, s))

    def bn(self, x, c):
        name = f"bn{self._bn}"
        self._bn += 1
        p = self.p
        y = batch_norm(x.float(), p[name + "/gamma"], p[name + "/beta"], p[name + "/moving_mean"],
                       p[name + "/moving_variance"], self.training)
        return self._tap(name, y.to(x.dtype))

    def relu(self, x):
        k = self._relu
        self._relu += 1
        if self.relu_masks is not None:      # masks taken from another implementation's activations (see forward)
            return x * self.relu_masks[k].to(x.dtype)
        return F.relu(x)

    def maxpool(self, x, k, s): return F.max_pool2d(x, k, s, padding=(k - 1) // 2)
    def add(self, a, b):
        name = f"add{self._add}"
        self._add += 1
        return self._tap(name, a + b)

    def mean_pool(self, x): return x.float().mean(dim=(2, 3))

    def dense(self, x, cin, cout):
        return self._tap("logits", F.linear(x, self.p["dense/kernel"], self.p["dense/bias"]))


def forward(cfg: ResNetConfig, params: Dict[str, torch.Tensor], images_nchw: torch.Tensor,
            training: bool, taps: Optional[Dict[str, torch.Tensor]] = None,
            emulate_bf16: bool = False, relu_masks: Optional[Sequence[torch.Tensor]] = None) -> torch.Tensor:
    """Logits ``[N, num_classes]`` (fp32).  ``params`` maps ``ParamSpec.name`` -> tensor.
    ``emulate_bf16`` rounds conv operands and stored activations to bf16 (straight-through
    gradients), reproducing where the native engine rounds.  ``relu_masks`` (one 0/1 NCHW tensor per ReLU, in call
    order) replaces every ``relu(x)`` by ``x * mask``: with the masks of the native engine's own activations the
    comparison of gradients is free of mask flips and measures the arithmetic alone."""
    return _walk(cfg, _ExecOps(cfg, params, images_nchw, training, taps, emulate_bf16, relu_masks)).float()


def init_params(cfg: ResNetConfig, initializer: str, seed: int = 0,
                device: str = "cpu") -> Dict[str, torch.Tensor]:
    from ..ops import reference
    gen = torch.Generator().manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    for s in param_specs(cfg):
        if s.kind in ("conv", "dense_w"):
            # the dense layer keeps the layer default (glorot_uniform): resnet_model.py:551-552
            t = reference.init_kernel(s.shape, initializer if s.kind == "conv" else "None", gen)
        elif s.kind in ("bn_gamma", "bn_var"):
            t = torch.ones(s.shape)
        else:
            t = torch.zeros(s.shape)
        out[s.name] = t.to(device)
    return out


# --------------------------------------------------------------------------- single blocks
def run_block(version: int, bottleneck: bool, projection: bool, filters: int, stride: int,
              params: Dict[str, torch.Tensor], x: torch.Tensor, training: bool = True) -> torch.Tensor:
    """One residual block in isolation — used against the golden block fixtures."""
    cin = x.shape[1]
    cfg = ResNetConfig(resnet_version=version, bottleneck=bottleneck)
    fout = filters * 4 if bottleneck else filters
    return _block(cfg, _ExecOps(cfg, params, x, training), x, cin, filters, fout, stride, projection)


def block_param_specs(version: int, bottleneck: bool, projection: bool, cin: int, filters: int
                      ) -> List[ParamSpec]:
    cfg = ResNetConfig(resnet_version=version, bottleneck=bottleneck)
    ops = _SpecOps(cfg)
    fout = filters * 4 if bottleneck else filters
    _block(cfg, ops, None, cin, filters, fout, 1, projection)
    return ops.b.specs
