"""Torch-eager training engine: the numerical oracle and the *restated baseline*.

It trains the same architectures with the same TF-1.x update rules as the native
engine, but every FLOP is a library call (cuDNN convolutions / fused BN, cuBLAS
GEMMs, ATen element-wise) driven by autograd — i.e. what the reference's
TensorFlow graph amounts to on this hardware ("B-train").

State is one flat fp32 tensor for parameters (leaf, requires_grad) with
per-parameter views, plus two flat slot tensors, so exploit/checkpoint paths see
the same "arena" shape as the native engine.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from ..ops import reference
from .resnet import ParamSpec


class EagerEngine:
    kind = "eager"

    def __init__(self, specs: Sequence[ParamSpec], forward: Callable[..., torch.Tensor],
                 init: Dict[str, torch.Tensor], hparams: Dict[str, Any], device, *,
                 compute_dtype: torch.dtype = torch.float32, lr_schedule: bool = False,
                 channels_last: bool = True, loss_scale: Optional[float] = None, dp_group=None):
        self.specs = list(specs)
        self.device = torch.device(device)
        self.forward_fn = forward
        self.compute_dtype = compute_dtype
        # static loss scale: loss x S before backward, gradients / S before the update; default 128 for fp16,
        # 1 otherwise (reference: _performance.py:30-33, resnet_run_loop.py:284-294)
        self.loss_scale = float(loss_scale) if loss_scale else (128.0 if compute_dtype == torch.float16 else 1.0)
        self.lr_schedule = lr_schedule
        # data-parallel baseline (BASELINE.md "B-allreduce-nccl"): NCCL all-reduce of the flat gradient, then the
        # separate optimizer step -- what the reference's dormant MirroredStrategy path amounts to
        # (official/utils/misc/distribution_utils.py:41-47)
        self.dp_group = dp_group
        self.channels_last = channels_last and self.device.type == "cuda"
        self.train_specs = [s for s in self.specs if s.trainable]
        self.state_specs = [s for s in self.specs if not s.trainable]
        n = sum(s.numel for s in self.train_specs)
        self.flat = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.slot0 = torch.zeros_like(self.flat)
        self.slot1 = torch.zeros_like(self.flat)
        self.reg_mask = torch.zeros_like(self.flat)
        self.params: Dict[str, torch.Tensor] = {}
        off = 0
        with torch.no_grad():
            for s in self.train_specs:
                self.flat[off: off + s.numel] = init[s.name].reshape(-1).to(self.device)
                if s.regularized:
                    self.reg_mask[off: off + s.numel] = 1.0
                off += s.numel
        self.flat.requires_grad_(True)
        for s in self.state_specs:
            self.params[s.name] = init[s.name].clone().to(self.device)
        self._refresh_views()
        self.step = 0
        self.hp: Dict[str, Any] = {}
        self.set_hparams(hparams, reset_slots=True)
        self.last_loss: Optional[torch.Tensor] = None

    def _refresh_views(self) -> None:
        """Per-parameter views of the flat leaf.  Re-made at the start of every training step: autograd runs a
        node's backward on the stream its forward ran on, and views made once at construction (on the legacy
        default stream) would tie every backward pass to that stream -- illegal inside a CUDA-graph capture."""
        off = 0
        for s in self.train_specs:
            self.params[s.name] = self.flat[off: off + s.numel].view(s.shape)
            off += s.numel

    # ------------------------------------------------------------------ hparams
    def set_hparams(self, hp: Dict[str, Any], reset_slots: bool = False) -> None:
        prev = self.hp.get("opt_case", {}).get("optimizer") if self.hp else None
        self.hp = hp
        if reset_slots or prev != hp["opt_case"]["optimizer"]:
            a, b = reference.slot_init_values(hp["opt_case"]["optimizer"])
            self.slot0.fill_(a)
            self.slot1.fill_(b)

    def current_lr(self) -> float:
        if self.lr_schedule:
            return reference.cifar_lr_at(self.step, self.hp)
        return float(self.hp["opt_case"]["lr"])

    # ------------------------------------------------------------------ steps
    def _prep(self, images_nhwc: torch.Tensor) -> torch.Tensor:
        x = images_nhwc.to(self.compute_dtype).permute(0, 3, 1, 2)
        return x.contiguous(memory_format=torch.channels_last) if self.channels_last else x.contiguous()

    def train_step(self, images_nhwc: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """One optimizer step; returns the (device) mean cross-entropy."""
        self.flat.grad = None
        self._refresh_views()
        logits = self.forward_fn(self.params, self._prep(images_nhwc), True)
        ce = F.cross_entropy(logits.float(), labels.long())
        # torch.autograd.grad instead of .backward(): no AccumulateGrad node, hence no synchronisation with the
        # leaf's (legacy default) stream -- which is what lets the whole step be captured in a CUDA graph
        # (bench.py --impl restated-graph)
        g, = torch.autograd.grad(ce * self.loss_scale if self.loss_scale != 1.0 else ce, self.flat)
        if self.loss_scale != 1.0:
            g = g / self.loss_scale
        oc = self.hp["opt_case"]
        reg, wd = self.hp.get("regularizer", "None"), float(self.hp.get("weight_decay", 0.0))
        with torch.no_grad():
            if self.dp_group is not None:
                import torch.distributed as dist
                g = g.contiguous()
                dist.all_reduce(g, group=self.dp_group)
                g = g / dist.get_world_size(self.dp_group)
            if reg != "None":
                g = g + self.reg_mask * reference.regularizer_grad(self.flat, reg, wd)
            lr = self.current_lr()
            self.step += 1
            reference.apply_update(oc["optimizer"], self.flat, g, self.slot0, self.slot1, lr=lr,
                                   step=self.step, momentum=float(oc.get("momentum", 0.0)),
                                   grad_decay=float(oc.get("grad_decay", 0.9)))
        self.last_loss = ce.detach()
        self.last_train_accuracy = (logits.detach().argmax(dim=1) == labels.long()).float().mean()
        return self.last_loss

    @torch.no_grad()
    def eval_batch(self, images_nhwc: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        logits = self.forward_fn(self.params, self._prep(images_nhwc), False)
        return (logits.argmax(dim=1) == labels.long()).sum()

    # ------------------------------------------------------------------ state
    def state_tensors(self) -> Dict[str, torch.Tensor]:
        out = {"params": self.flat.detach(), "slot0": self.slot0, "slot1": self.slot1}
        for s in self.state_specs:
            out[s.name] = self.params[s.name]
        return out

    @torch.no_grad()
    def load_state(self, tensors: Dict[str, Any], step: int) -> None:
        for k, dst in self.state_tensors().items():
            src = tensors[k]
            src = src if isinstance(src, torch.Tensor) else torch.from_numpy(src)
            dst.copy_(src.to(self.device).view_as(dst))
        self.step = int(step)

    def state_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.state_tensors().values())

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.current_stream(self.device).synchronize()
