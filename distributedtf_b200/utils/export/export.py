"""Inference export.

Parity: reference ``official/utils/export/export.py:24-49`` builds the serving input receiver used by
``classifier.export_savedmodel`` (``resnet_run_loop.py:510-514``).  Without TensorFlow the exported
artefact is an *inference bundle*: the member's parameters in the oracle's naming (fp32, checkpoint
format of ``ckpt.py``) plus ``signature.json`` describing the input tensor, and ``load_inference_bundle``
returns a callable that classifies raw image batches.
"""
from __future__ import annotations

import json
import os
from typing import Any, Callable, Dict, Optional, Sequence

import torch

from ... import ckpt


def build_tensor_serving_input_receiver_fn(shape: Sequence[int], dtype: torch.dtype = torch.float32,
                                           batch_size: Optional[int] = 1) -> Callable[[], Dict[str, Any]]:
    """Describes the serving input: a float tensor ``[batch_size] + shape`` named ``input_tensor``."""
    def serving_input_receiver_fn() -> Dict[str, Any]:
        full = [batch_size] + list(shape)
        return {"name": "input_tensor", "shape": full, "dtype": str(dtype).replace("torch.", ""),
                "placeholder": torch.zeros([1 if d is None else d for d in full], dtype=dtype)}
    return serving_input_receiver_fn


def export_inference_bundle(export_dir: str, named_params: Dict[str, torch.Tensor], model: str,
                            model_config: Dict[str, Any], shape: Sequence[int], batch_size: Optional[int] = None) -> str:
    os.makedirs(export_dir, exist_ok=True)
    ckpt.save_checkpoint(export_dir, named_params, step=0, optimizer="", meta={"model": model})
    spec = build_tensor_serving_input_receiver_fn(shape, batch_size=batch_size)()
    spec.pop("placeholder")
    with open(os.path.join(export_dir, "signature.json"), "w") as f:
        json.dump({"model": model, "config": model_config, "inputs": spec,
                   "outputs": {"classes": "int64 [batch]", "probabilities": "float32 [batch, num_classes]"}}, f, indent=1)
    return export_dir


def load_inference_bundle(export_dir: str, device="cpu") -> Callable[[torch.Tensor], Dict[str, torch.Tensor]]:
    with open(os.path.join(export_dir, "signature.json")) as f:
        sig = json.load(f)
    tensors, _ = ckpt.load_checkpoint(export_dir)
    params = {k: torch.from_numpy(v).to(device) for k, v in tensors.items()}
    if sig["model"] == "resnet":
        from ...models import resnet
        cfg = resnet.ResNetConfig(**{k: tuple(v) if isinstance(v, list) else v for k, v in sig["config"].items()})
        fwd = lambda x: resnet.forward(cfg, params, x.permute(0, 3, 1, 2), False)
    else:
        from ...models import mnist_net
        fwd = lambda x: mnist_net.forward(params, x.permute(0, 3, 1, 2), False)

    @torch.no_grad()
    def predict(images_nhwc: torch.Tensor) -> Dict[str, torch.Tensor]:
        logits = fwd(images_nhwc.to(device).float())
        return {"classes": logits.argmax(1), "probabilities": torch.softmax(logits, 1)}
    return predict
