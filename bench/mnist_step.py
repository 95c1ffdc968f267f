#!/usr/bin/env python
"""One native MNIST member: steps/s on a resident synthetic dataset (and a target for `ncu` launch lists).
Usage: python bench/mnist_step.py [--steps 200] [--no-graphs]"""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedtf_b200.hpspace import default_cifar_hparam
from distributedtf_b200.models.mnist_native import NativeMnistEngine

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--no-graphs", action="store_true")
a = ap.parse_args()
hp = default_cifar_hparam()
hp["opt_case"] = {"optimizer": "Momentum", "lr": 0.01, "momentum": 0.9}
e = NativeMnistEngine(hp, "cuda:0", seed=0, use_graphs=not a.no_graphs)
g = torch.Generator().manual_seed(0)
images = torch.randint(0, 256, (8192, 28, 28, 1), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (8192,), dtype=torch.int32, generator=g).cuda()
idx = [torch.randperm(8192, generator=g)[: a.batch].cuda() for _ in range(16)]
for i in range(5):
    e.train_step_indexed(images, labels, idx[i % 16])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(a.steps):
    e.train_step_indexed(images, labels, idx[i % 16])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
print(json.dumps({"model": "mnist cnn", "batch": a.batch, "ms_per_step": round(ms, 4), "images_per_s": round(a.batch / ms * 1e3),
                  "launches_per_step": e.launches_per_step, "loss": float(e.last_loss)}))
