"""Input pipelines: synthetic generators, CIFAR-10 / MNIST readers, augmentation."""
from .datasets import (Dataset, synthetic_dataset, load_cifar10, load_mnist,  # noqa: F401
                       BatchIterator, cifar_preprocess_torch, CIFAR_TRAIN, CIFAR_EVAL,
                       MNIST_TRAIN, MNIST_EVAL)
