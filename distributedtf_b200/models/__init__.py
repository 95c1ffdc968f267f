"""Population member implementations.

``toy``        the PBT-paper two-parameter toy (reference ``toy_model.py``)
``mnist``      the MNIST CNN member (reference ``mnist_model.py``)
``cifar10``    the CIFAR-10 ResNet member (reference ``cifar10_model.py`` + ``resnet/``)
``resnet``     ResNet v1/v2 definition in plain PyTorch — numerical oracle and the
               "restated baseline" (torch eager + cuDNN)
``resnet_native``  the B200 engine: flat arena + hand-written sm_100a kernels
"""

def get_model_class(name: str):
    name = name.lower()
    if name in ("toy", "toymodel"):
        from .toy import ToyModel
        return ToyModel
    if name in ("mnist", "mnistmodel"):
        from .mnist import MNISTModel
        return MNISTModel
    if name in ("cifar10", "cifar", "cifar10model", "resnet"):
        from .cifar10 import Cifar10Model
        return Cifar10Model
    raise ValueError(f"unknown model '{name}' (toy | mnist | cifar10)")
