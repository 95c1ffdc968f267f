"""distributedtf_b200 — Blackwell-native population-based training (PBT).

A from-scratch framework with the capabilities of youzhenfei1995/DistributedTF
(master/worker PBT over a host control plane, toy / MNIST / CIFAR-10 ResNet
members, checkpoint + report layout), re-designed for B200:

* one process per GPU, host control messages over TCP / gloo / in-process queues;
* every member's state lives in one flat device arena (weights, optimizer slots,
  BN statistics, step counter, device-resident hyper-parameter block);
* the hot ops are hand-written sm_100a kernels (``csrc/``) reached through a
  C ABI (``include/dtf_b200.h``) — tcgen05/TMEM implicit-GEMM convolutions, fused
  BatchNorm/ReLU prologues and epilogues, softmax-CE head, a multi-tensor
  optimizer, and peer-memory exploit / all-reduce / score-board kernels.

Sub-packages
------------
``parallel``  control-plane transports, symmetric memory, DP groups, exploit paths
``models``    toy, MNIST CNN, CIFAR-10 ResNet members (+ torch-eager oracles)
``ops``       ctypes bindings to the native library and fp32 reference restatements
``utils``     flags, logging, hooks, testing helpers (official/utils equivalents)
"""

__version__ = "0.1.0"

from .protocol import WorkerInstruction  # noqa: F401
