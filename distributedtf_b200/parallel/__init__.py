"""Parallelism and communication.

Control plane (host, byte-sized messages — never tensors):
  ``comm_base.Comm``  mpi4py-shaped endpoint API (``Get_rank/Get_size/isend/send/recv``)
  ``local``           in-process fabric (threads + queues) — tests and 1-process runs
  ``tcp``             star-topology TCP transport (master <-> each worker)
  ``gloo``            torch.distributed (gloo) object transport
  ``mpi``             optional adapter when mpi4py is importable

Data plane (device):
  ``symm``            symmetric device arenas shared between the ranks of one node
  ``exploit``         winner -> loser state transfer paths (p2p kernel / nccl / host / file)
  ``scoreboard``      device-side eval-score board read by the master
  ``dp``              data-parallel member groups with the fused all-reduce + update

Strategies that the reference does not have (TP / PP / SP / EP / CP / ring
attention / ZeRO / FSDP) are intentionally absent: the models are small CNNs
without a sequence axis.
"""
from .comm_base import Comm, Request, CommTimeout  # noqa: F401
from .local import LocalFabric  # noqa: F401
