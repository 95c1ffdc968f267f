#!/usr/bin/env python
"""Single-GPU launches of the communication kernels for `ncu` (which must never wrap a multi-rank command):
the exploit copy of a ResNet-32 member state (5.6 MB, fused hyper-parameter write + cell reset + epoch release)
and the fused all-reduce + update kernel with a world of one replica (same code path, the peer is the GPU itself).
Peer traffic itself is measured by bench/exploit_sweep.py and bench.py --dp-size on real pairs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtf_b200.hpspace import default_cifar_hparam
from distributedtf_b200.models import resnet
from distributedtf_b200.models.resnet_native import NativeResNetEngine
from distributedtf_b200.parallel import exploit
from distributedtf_b200.parallel.dp import DataParallelGroup

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
hp = default_cifar_hparam()
hp["batch_size"] = 128
box = {}


def alloc(n):
    box["mb"] = exploit.MemberBuffer(n, dev)
    return box["mb"].alloc_state(n)


eng = NativeResNetEngine(resnet.cifar10_config(32), hp, dev, seed=1, use_graphs=False, arena_alloc=alloc)
mb = box["mb"]
g = torch.Generator().manual_seed(0)
imgs = torch.randint(0, 256, (128, 32, 32, 3), dtype=torch.uint8, generator=g).to(dev)
labels = torch.randint(0, 10, (128,), generator=g).to(dev)
eng.train_step_u8(imgs, labels, augment=False)
for rnd in range(6):
    exploit.publish_buffer(mb, rnd)                                        # exploit_copy_kernel: state -> snapshot
    exploit.pull_buffer(mb, mb.buf.ptr, rnd, hp_offset=eng.hp_state_offset, new_hp=eng.hparams_struct(hp),
                        zero=eng.round_cells)                             # snapshot -> state, fused writes
dp = DataParallelGroup(eng.zero_region.numel() * 4, dev)                  # world of one: peer == self
grads = dp.alloc_grads(eng.zero_region.numel() * 4).view(torch.float32)
grads[: eng.grads.numel()].copy_(eng.grads)
for _ in range(6):
    dp.allreduce_update(eng.params, eng.slot0, eng.slot1, eng.pack, eng.segs_dev, eng.chunks_dev, eng.n_chunks,
                        eng.hp_dev, eng.step_dev)
torch.cuda.synchronize()
print("comm targets ok", dp.comm_stats())
