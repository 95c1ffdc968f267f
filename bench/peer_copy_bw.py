#!/usr/bin/env python
"""Measured NVLink peer-copy bandwidth of this box: the denominator of the exploit-transfer roofline.

cudaMemcpyPeerAsync (torch `copy_` between two devices) of 1 GiB from GPU 1 to GPU 0, best of 10, CUDA events on
the destination's stream; then the same with every GPU i copying from GPU (i + N/2) % N at once.  Prints one JSON
line.  Needs >= 2 GPUs in ONE process (no torchrun)."""
import json
import torch

n = torch.cuda.device_count()
assert n >= 2, "needs two GPUs"
nbytes = 1 << 30
bufs = [torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
for i in range(n):
    for j in range(n):
        if i != j and not torch.cuda.can_device_access_peer(i, j):
            raise SystemExit(f"no peer access {i}->{j}")
dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")


def one_pair():
    best = 1e9
    for _ in range(10):
        with torch.cuda.device(0):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            dst.copy_(bufs[1], non_blocking=True)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
    return nbytes / best / 1e6


def all_pairs():
    dsts = [torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
    best = 1e9
    for _ in range(5):
        evs = []
        for i in range(n):
            torch.cuda.synchronize(i)
        for i in range(n):
            with torch.cuda.device(i):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dsts[i].copy_(bufs[(i + n // 2) % n], non_blocking=True)
                e1.record()
                evs.append((e0, e1))
        for i in range(n):
            torch.cuda.synchronize(i)
        best = min(best, max(a.elapsed_time(b) for a, b in evs))
    return nbytes / best / 1e6


print(json.dumps({"what": "cudaMemcpyPeer 1 GiB, GB/s per direction per GPU", "gpus": n,
                  "one_pair_gbs": round(one_pair(), 1), "all_gpus_at_once_gbs": round(all_pairs(), 1)}))
