"""Stand-alone run loop of one CIFAR-10 ResNet member, driven by the harness flag groups.

Parity: reference ``resnet/cifar10_main.py:239-330`` (``define_cifar_flags``: the ResNet flag groups plus one
flag per hyper-parameter, with defaults taken from a member's hparams; ``run_cifar``; ``main``) and
``resnet/resnet_run_loop.py:357-516`` (``resnet_main``: clean, train -> eval cycles every
``epochs_between_evals`` epochs, ``stop_threshold`` early exit, ``export_dir``).  The reference wipes and
re-defines process-global absl flags on every member call; here ``define_cifar_flags`` returns a fresh
``FlagSet`` object and ``main(argv)`` can be called any number of times.

    python -m distributedtf_b200.models.resnet_run --model_dir /tmp/m --use_synthetic_data --max_train_steps 10

is the one-member equivalent of ``python cifar10_main.py``; the PBT runner reaches the same loop through
``Cifar10Model.train``.
"""
from __future__ import annotations

import os
import sys
from typing import Any, Dict, Optional, Sequence, Tuple

from ..hpspace import default_cifar_hparam
from ..utils.flags import core as flags_core
from ..utils.logs import logger as _logger
from ..utils.misc import distribution_utils, model_helpers

RESNET_SIZES = sorted({str(6 * n + 2) for n in range(1, 34)} | {str(9 * n + 2) for n in range(1, 23)}, key=int)


def define_cifar_flags(hp: Optional[Dict[str, Any]] = None, model_id: int = 0, model_dir: str = "/tmp/cifar10_model",
                       data_dir: str = "/tmp/cifar10_data", train_epochs: int = 250,
                       total_epochs: Optional[int] = None, epoch_index: int = 0) -> flags_core.FlagSet:
    hp = hp or default_cifar_hparam()
    oc = hp["opt_case"]
    fs = flags_core.define_resnet_flags(resnet_size_choices=RESNET_SIZES)
    fs.define("optimizer", "opt", oc["optimizer"], "The name of optimizer type")
    fs.define("momentum", "mm", float(oc.get("momentum", 0.0)), "The momentum of Momentum SGD or RMSProp", type=float)
    fs.define("grad_decay", "rmspd", float(oc.get("grad_decay", 0.9)), "The decay of RMSProp", type=float)
    fs.define("learning_rate", "lr", float(oc["lr"]), "The initial learning rate of optimizer", type=float)
    fs.define("decay_rate", "lrdr", float(hp["decay_rate"]), "The base term of learning rate decay function", type=float)
    fs.define("decay_steps", "lrds", int(hp["decay_steps"]), "Decay interval in percent of train_epochs; 0 turns "
              "decay off", type=int)
    fs.define("initializer", "initn", hp["initializer"], "Initialization method; None means glorot_uniform")
    fs.define("regularizer", "regn", hp["regularizer"], "Regularization method; None turns weight decay off")
    fs.define("weight_decay", "wd", float(hp["weight_decay"]), "The amount of regularization", type=float)
    fs.define("model_id", "mid", model_id, "The index of model in the population", type=int)
    fs.define("total_epochs", "ttep", total_epochs, "The total epochs the model will be trained (default: "
              "epoch_index + train_epochs as parsed; the reference takes the unparsed default, SURVEY 2.8)", type=int)
    fs.define("epoch_index", "epi", epoch_index, "The epoch index written to the csv", type=int)
    fs.define("resnet_bottleneck", "rb", False, "Bottleneck blocks (resnet_size is then 9n + 2).", type=bool)
    fs.define("engine", None, "auto", "native (sm_100a kernels) | eager (torch oracle) | auto",
              choices=["auto", "native", "eager"])
    fs.set_defaults(data_dir=data_dir, model_dir=model_dir, resnet_size="32", train_epochs=train_epochs,
                    epochs_between_evals=1, batch_size=int(hp["batch_size"]), dtype="bf16")
    return fs


def hparams_from_flags(ns) -> Dict[str, Any]:
    oc: Dict[str, Any] = {"optimizer": ns.optimizer, "lr": ns.learning_rate}
    if ns.optimizer in ("Momentum", "RMSProp"):
        oc["momentum"] = ns.momentum
    if ns.optimizer == "RMSProp":
        oc["grad_decay"] = ns.grad_decay
    return {"opt_case": oc, "decay_steps": ns.decay_steps, "decay_rate": ns.decay_rate,
            "weight_decay": ns.weight_decay, "regularizer": ns.regularizer, "initializer": ns.initializer,
            "batch_size": ns.batch_size}


def resnet_main(ns) -> Tuple[float, int]:
    """One member, ``train_epochs`` epochs in cycles of ``epochs_between_evals``; returns (eval accuracy, id)."""
    from .cifar10 import Cifar10Model
    model_helpers.apply_clean(ns)
    num_gpus = flags_core.get_num_gpus(ns)
    strategy = distribution_utils.get_distribution_strategy(num_gpus, ns.all_reduce_alg)
    if strategy.kind == "data_parallel":
        # a data-parallel member needs one process per GPU: that launch lives in the PBT runner (--dp-size)
        raise ValueError("resnet_run drives one process; use `python -m distributedtf_b200 --dp-size N` for num_gpus > 1")
    distribution_utils.per_device_batch_size(ns.batch_size, max(num_gpus, 1))
    save_base = os.path.join(ns.model_dir, "model_")          # the member's files land in <model_dir>/model_<id>/
    member = Cifar10Model(ns.model_id, hparams_from_flags(ns), save_base,
                          engine=ns.engine, dtype=ns.dtype, synthetic=bool(ns.use_synthetic_data),
                          data_dir=ns.data_dir, max_train_steps=ns.max_train_steps,
                          eval_batches=4 if ns.max_train_steps else None,
                          resnet_size=int(ns.resnet_size), resnet_version=int(ns.resnet_version),
                          resnet_bottleneck=bool(ns.resnet_bottleneck),
                          loss_scale=flags_core.get_loss_scale(ns) if ns.dtype == "fp16" or ns.loss_scale else None,
                          epochs_between_evals=ns.epochs_between_evals, stop_threshold=ns.stop_threshold,
                          export_dir=ns.export_dir, resume=os.path.isdir(save_base + str(ns.model_id)),
                          benchmark=(ns.benchmark_logger_type, ns.benchmark_log_dir))
    member.epoches_trained = ns.epoch_index
    try:
        member.train(ns.train_epochs, max(ns.total_epochs or 0, ns.epoch_index + ns.train_epochs))
        return float(member.accuracy), ns.model_id
    finally:
        member.close()


def run_cifar(ns) -> Tuple[float, int]:
    with _logger.benchmark_context(ns):
        return resnet_main(ns)


def main(argv: Optional[Sequence[str]] = None, hp: Optional[Dict[str, Any]] = None) -> Tuple[float, int]:
    ns = define_cifar_flags(hp).parse(list(argv) if argv is not None else sys.argv[1:])
    return run_cifar(ns)


if __name__ == "__main__":
    acc, mid = main()
    print("Return {}".format((acc, mid)))
