"""MNIST CNN population member.

Parity: reference ``mnist_model.py:128-201``.  ``learning_curve.csv`` columns:
``global_step`` (= epoch index, as in ``mnist_model.py:184``), ``eval_accuracy``,
``optimizer``, ``lr``.  Two reference defects are *not* reproduced:
an "epoch" is a full pass over the 60 000 training images rather than 10 debug
steps (``mnist_model.py:162-165``; cap it with ``max_train_steps`` if wanted), and
``epoches_trained`` advances by the number of epochs trained, not by 1.
Pixels stay unnormalised floats 0..255, as in the reference.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from ..utils.data import MNIST_EVAL, MNIST_TRAIN, load_mnist, synthetic_dataset
from . import mnist_net
from ._member import GpuMember

_DTYPES = {"fp32": torch.float32, "tf32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}


class MNISTModel(GpuMember):
    train_examples = MNIST_TRAIN
    eval_examples = MNIST_EVAL
    csv_fields = ["global_step", "eval_accuracy", "optimizer", "lr"]
    export_model_name = "mnist"
    export_input_shape = [28, 28, 1]

    def build_datasets(self):
        if self.synthetic:
            n_train = 8192 if self.max_train_steps else MNIST_TRAIN
            n_eval = 2048 if self.eval_batches else MNIST_EVAL
            return (synthetic_dataset(n_train, 28, 28, 1, 10, self.synthetic_mode, 21, self.device),
                    synthetic_dataset(n_eval, 28, 28, 1, 10, self.synthetic_mode, 22, self.device))
        return load_mnist(self.data_dir, True, self.device), load_mnist(self.data_dir, False, self.device)

    def build_engine(self):
        seed = 104729 * (self.cluster_id + 1)
        if self.engine_kind == "native":
            try:
                from .mnist_native import NativeMnistEngine
            except ImportError:
                if self._engine_requested == "native":
                    raise
                self.engine_kind = "eager"          # `auto`: the MNIST kernels are not built into this version
            else:
                return NativeMnistEngine(self.hparams, self.device, seed=seed, dtype=self.dtype,
                                         arena_alloc=self.make_symmetric_alloc, use_graphs=self.use_graphs,
                                         dp_group=self.dp_group if self.dp_size > 1 else None,
                                          all_reduce_alg=self.all_reduce_alg if self.dp_size > 1 else None)
        from .engine_eager import EagerEngine
        init = mnist_net.init_params(self.hparams.get("initializer", "None"), seed)
        return EagerEngine(mnist_net.param_specs(), mnist_net.forward, init, self.hparams, self.device,
                           compute_dtype=_DTYPES[self.dtype] if self.device.type == "cuda" else torch.float32,
                           lr_schedule=False)

    def preprocess(self, images_u8: torch.Tensor, training: bool) -> torch.Tensor:
        pre = getattr(self.engine, "preprocess", None)
        if pre is not None:
            return pre(images_u8, training)
        return images_u8.float()

    def csv_row(self, epoch_index: int, accuracy: float) -> Dict[str, Any]:
        oc = self.hparams["opt_case"]
        return {"global_step": epoch_index, "eval_accuracy": accuracy, "optimizer": oc["optimizer"], "lr": oc["lr"]}
