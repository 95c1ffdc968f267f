"""CIFAR-10 ResNet population member.

Parity: reference ``cifar10_model.py:10-32`` (adapter) + ``resnet/cifar10_main.py``
(model_fn, LR schedule, data) + ``resnet/resnet_run_loop.py:357-516`` (train ->
eval cycle, ``learning_curve.csv`` row with the column order the plots index by
position).  ``resnet_size`` defaults to 32 (BASELINE target); the reference's own
default is '50' (``cifar10_main.py:294``), selectable with ``resnet_size=50``.
Hyper-parameters arrive in the constructor / ``set_values`` — no flag globals.
"""
from __future__ import annotations

from typing import Optional, Any, Dict, List

import torch

from ..utils.data import (CIFAR_EVAL, CIFAR_TRAIN, cifar_preprocess_torch, load_cifar10,
                          synthetic_dataset)
from . import resnet
from ._member import GpuMember

_DTYPES = {"fp32": torch.float32, "tf32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}


class Cifar10Model(GpuMember):
    train_examples = CIFAR_TRAIN
    eval_examples = CIFAR_EVAL
    csv_fields = ["epochs", "eval_accuracy", "optimizer", "learning_rate", "decay_rate", "decay_steps",
                  "initializer", "regularizer", "weight_decay", "batch_size", "model_id"]

    def __init__(self, cluster_id: int, hparams: Dict[str, Any], save_base_dir: str, *,
                 resnet_size: int = 32, resnet_version: int = 2, loss_scale: Optional[float] = None,
                 resnet_bottleneck: bool = False, **kw):
        self.net_cfg = resnet.cifar10_config(resnet_size, resnet_version, bottleneck=resnet_bottleneck)
        self.loss_scale = loss_scale            # None: 128 for fp16, 1 otherwise (flags.get_loss_scale)
        super().__init__(cluster_id, hparams, save_base_dir, **kw)

    def build_datasets(self):
        if self.synthetic:
            n_train = 8192 if self.max_train_steps else CIFAR_TRAIN
            n_eval = 2048 if self.eval_batches else CIFAR_EVAL
            tr = synthetic_dataset(n_train, 32, 32, 3, 10, self.synthetic_mode, 11, self.device)
            ev = synthetic_dataset(n_eval, 32, 32, 3, 10, self.synthetic_mode, 12, self.device)
            return tr, ev
        return load_cifar10(self.data_dir, True, self.device), load_cifar10(self.data_dir, False, self.device)

    def build_engine(self):
        seed = 7919 * (self.cluster_id + 1)
        if self.engine_kind == "native":
            from .resnet_native import NativeResNetEngine
            from .resnet_native_bottleneck import NativeResNetBottleneckEngine, NativeResNetBottleneckV1Engine
            from .resnet_native_v1 import NativeResNetV1Engine
            cls = {(2, False): NativeResNetEngine, (1, False): NativeResNetV1Engine,
                   (2, True): NativeResNetBottleneckEngine, (1, True): NativeResNetBottleneckV1Engine}[
                       (self.net_cfg.resnet_version, self.net_cfg.bottleneck)]
            try:
                return cls(self.net_cfg, self.hparams, self.device, seed=seed, dtype=self.dtype,
                           loss_scale=self.loss_scale,
                           arena_alloc=self.make_symmetric_alloc, use_graphs=self.use_graphs,
                           dp_group=self.dp_group if self.dp_size > 1 else None,
                           all_reduce_alg=self.all_reduce_alg if self.dp_size > 1 else None)
            except NotImplementedError:
                # fp32 / tf32 compute is served by the eager engine; an explicit request for the native
                # engine is an error, `auto` degrades.
                if self._engine_requested == "native":
                    raise
                self.engine_kind, self.symm = "eager", None
        from .engine_eager import EagerEngine
        cfg = self.net_cfg
        init = resnet.init_params(cfg, self.hparams.get("initializer", "None"), seed)
        return EagerEngine(resnet.param_specs(cfg), lambda p, x, t: resnet.forward(cfg, p, x, t), init,
                           self.hparams, self.device, compute_dtype=_DTYPES[self.dtype]
                           if self.device.type == "cuda" else torch.float32, lr_schedule=True,
                           loss_scale=self.loss_scale)

    export_model_name = "resnet"
    export_input_shape = [32, 32, 3]

    def export_config(self) -> Dict[str, Any]:
        import dataclasses
        return dataclasses.asdict(self.net_cfg)

    def preprocess(self, images_u8: torch.Tensor, training: bool) -> torch.Tensor:
        pre = getattr(self.engine, "preprocess", None)
        if pre is not None:
            return pre(images_u8, training)
        return cifar_preprocess_torch(images_u8, training)

    def csv_fields_now(self) -> List[str]:
        fields = list(self.csv_fields)
        name = self.hparams["opt_case"]["optimizer"]
        if name in ("Momentum", "RMSProp"):
            fields.append("momentum")
        if name == "RMSProp":
            fields.append("grad_decay")
        return fields

    def csv_row(self, epoch_index: int, accuracy: float) -> Dict[str, Any]:
        hp, oc = self.hparams, self.hparams["opt_case"]
        row = {"epochs": epoch_index, "eval_accuracy": accuracy, "optimizer": oc["optimizer"],
               "learning_rate": oc["lr"], "decay_rate": hp.get("decay_rate"), "decay_steps": hp.get("decay_steps"),
               "initializer": hp.get("initializer"), "regularizer": hp.get("regularizer"),
               "weight_decay": hp.get("weight_decay"), "batch_size": hp["batch_size"], "model_id": self.cluster_id}
        if "momentum" in self.csv_fields_now():
            row["momentum"] = oc.get("momentum")
        if "grad_decay" in self.csv_fields_now():
            row["grad_decay"] = oc.get("grad_decay")
        return row
