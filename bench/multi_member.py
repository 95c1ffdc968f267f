#!/usr/bin/env python
"""Several population members on ONE GPU, trained concurrently (one CUDA stream + one host thread each) versus
one after the other.  The reference trains the members of a worker sequentially (training_worker.py:60-75);
with pop_size > #GPUs (its own sweep: pop 10-50 on 1-4 workers) overlapping the members' latency-bound steps
is free throughput.  Usage: python bench/multi_member.py [--members 1 2 4] [--steps 100] [--model resnet|mnist]"""
import argparse
import json
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedtf_b200.hpspace import default_cifar_hparam  # noqa: E402
from distributedtf_b200.utils.streams import new_stream  # noqa: E402


def make_engine(model, seed, dev):
    hp = default_cifar_hparam()
    if model == "resnet":
        from distributedtf_b200.models.resnet import ResNetConfig
        from distributedtf_b200.models.resnet_native import NativeResNetEngine
        return NativeResNetEngine(ResNetConfig(resnet_size=32), hp, dev, seed=seed), (32, 32, 3)
    from distributedtf_b200.models.mnist_native import NativeMnistEngine
    hp["opt_case"] = {"optimizer": "Momentum", "lr": 0.01, "momentum": 0.9}
    return NativeMnistEngine(hp, dev, seed=seed), (28, 28, 1)


def run(model, k, steps, bs=128, share=False):
    from distributedtf_b200.ops import native
    dev = torch.device("cuda:0")
    torch.zeros(1, device=dev)
    native.set_sm_limit(native.set_sm_limit(0) // k if share else 0)
    engines, streams, data = [], [], []
    for i in range(k):
        e, shape = make_engine(model, i, dev)
        s = new_stream(dev)          # not torch.cuda.Stream(): its 32 pooled streams repeat (utils/streams.py)
        x = torch.randint(0, 256, (bs, *shape), dtype=torch.uint8, device=dev)
        y = torch.randint(0, 10, (bs,), dtype=torch.int32, device=dev)
        with torch.cuda.stream(s):                       # plain run + graph capture, one member at a time
            for _ in range(3):
                e.train_step_u8(x, y)
        torch.cuda.synchronize()
        engines.append(e); streams.append(s); data.append((x, y))

    def work(i):
        with torch.cuda.stream(streams[i]):
            x, y = data[i]
            for _ in range(steps):
                engines[i].train_step_u8(x, y)

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    threads = [threading.Thread(target=work, args=(i,)) for i in range(k)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return k * steps * bs / dt, dt / steps * 1e3


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--members", type=int, nargs="+", default=[1, 2, 4])
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--model", default="resnet")
    ap.add_argument("--share-sms", action="store_true", help="give each member 1/k of the SMs")
    a = ap.parse_args()
    base = None
    for k in a.members:
        ips, ms = run(a.model, k, a.steps, share=a.share_sms)
        base = base or ips
        print(json.dumps({"model": a.model, "members_in_flight": k, "share_sms": a.share_sms, "images_per_s": round(ips), "ms_per_round_of_steps": round(ms, 3),
                          "speedup_vs_1": round(ips / base, 3)}))
