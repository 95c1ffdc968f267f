#!/usr/bin/env python
"""Time single launches of the generic (gather) convolution kernels on ResNet-32 / batch-128 shapes.
Usage: [DTF_LIB=path/to/other/libdtf_b200.so] python bench/conv_micro.py"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedtf_b200.ops import native as nat  # noqa: E402

dev = torch.device("cuda")
NB = 128


def bf(*shape):
    return (torch.randn(*shape, device=dev) * 0.5).to(torch.bfloat16)


def stats_of(x):
    f = x.float().reshape(-1, x.shape[-1])
    return torch.cat([f.sum(0), (f * f).sum(0)]).contiguous()


def timeit(fn, iters=20, reps=10):
    """Device time per launch: `iters` launches captured in one CUDA graph (no host launch overhead)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        with torch.cuda.graph(graph, stream=stream):
            for _ in range(iters):
                fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (iters * reps)


def fwd_case(H, Cin, Cout, k, stride, bn_pro):
    x = bf(NB, H, H, Cin)
    w = torch.randn(Cout, k, k, Cin, device=dev) / math.sqrt(k * k * Cin)
    wpk = nat.pack_weight_host(w)
    Ho = H // stride
    out = torch.empty(NB, Ho, Ho, Cout, device=dev, dtype=torch.bfloat16)
    st = stats_of(x)
    keep = [x, st, torch.ones(Cin, device=dev), torch.zeros(Cin, device=dev)]
    bn = nat.bn_ref(Cin, NB * H * H, stats=st, gamma=keep[2], beta=keep[3]) if bn_pro else None
    so = torch.zeros(2 * Cout, device=dev)
    g = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2,
                   mode=nat.PRO_BN_RELU if bn_pro else nat.PRO_NONE, bn=bn)
    return lambda: nat.conv_gemm(g, wpk, out, NB=NB, Ho=Ho, Wo=Ho, N=Cout, epi_stats_out=so), keep


def wgrad_case(H, Cin, Cout, k, stride, bn_pro):
    x = bf(NB, H, H, Cin)
    Ho = H // stride
    gy = bf(NB, Ho, Ho, Cout)
    st = stats_of(x)
    keep = [x, gy, st, torch.ones(Cin, device=dev), torch.zeros(Cin, device=dev)]
    bn = nat.bn_ref(Cin, NB * H * H, stats=st, gamma=keep[3], beta=keep[4]) if bn_pro else None
    dw = torch.zeros(Cout, k, k, Cin, device=dev)
    gx = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2,
                    mode=nat.PRO_BN_RELU if bn_pro else nat.PRO_NONE, bn=bn)
    gg = nat.gather(gy, Ho, Ho, Cout)
    return lambda: nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=Ho, Wo=Ho, Cout=Cout), keep


def dgrad_case(H, Cin, Cout, k, stride, bn_pro):
    """Input gradient of a conv Cin->Cout over an HxH input (gy is (H/stride)^2 x Cout)."""
    Ho = H // stride
    gy, c_out, x_in = bf(NB, Ho, Ho, Cout), bf(NB, Ho, Ho, Cout), bf(NB, H, H, Cin)
    w = torch.randn(Cout, k, k, Cin, device=dev) / math.sqrt(k * k * Cin)
    wpk = nat.pack_weight_host(w, transpose_for_dgrad=True)
    st_o, st_i = stats_of(c_out), stats_of(x_in)
    ones_o, zeros_o = torch.ones(Cout, device=dev), torch.zeros(Cout, device=dev)
    ones_i, zeros_i = torch.ones(Cin, device=dev), torch.zeros(Cin, device=dev)
    bwd = torch.zeros(2 * Cout, device=dev)
    keep = [gy, c_out, x_in, st_o, st_i, ones_o, zeros_o, ones_i, zeros_i, bwd]
    bn_o = nat.bn_ref(Cout, NB * Ho * Ho, stats=st_o, gamma=ones_o, beta=zeros_o) if bn_pro else None
    g = nat.gather(gy, Ho, Ho, Cout, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, transposed=True,
                   mode=nat.PRO_BN_BWD if bn_pro else nat.PRO_NONE, bn=bn_o, src2=c_out if bn_pro else None,
                   bwd_stats=bwd if bn_pro else None)
    out = torch.zeros(NB, H, H, Cin, device=dev, dtype=torch.bfloat16)
    so = torch.zeros(2 * Cin, device=dev)
    epi_bn = nat.bn_ref(Cin, NB * H * H, stats=st_i, gamma=ones_i, beta=zeros_i)
    return lambda: nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in,
                                 epi_bn=epi_bn, epi_stats_out=so, epi_accumulate=(k == 1)), keep


CASES = [
    ("fwd conv0 3x3 8->16", fwd_case, (32, 8, 16, 3, 1, False)),
    ("fwd proj 1x1 16->16", fwd_case, (32, 16, 16, 1, 1, True)),
    ("fwd proj 1x1 s2 16->32", fwd_case, (32, 16, 32, 1, 2, True)),
    ("fwd 3x3 s2 16->32", fwd_case, (32, 16, 32, 3, 2, True)),
    ("fwd 3x3 s2 32->64", fwd_case, (16, 32, 64, 3, 2, True)),
    ("dgrad proj 1x1 16->16", dgrad_case, (32, 16, 16, 1, 1, False)),
    ("dgrad 3x3 s2 16->32", dgrad_case, (32, 16, 32, 3, 2, True)),
    ("dgrad 1x1 s2 16->32", dgrad_case, (32, 16, 32, 1, 2, False)),
    ("dgrad 3x3 s2 32->64", dgrad_case, (16, 32, 64, 3, 2, True)),
    ("dgrad 1x1 s2 32->64", dgrad_case, (16, 32, 64, 1, 2, False)),
    ("wgrad conv0 3x3 8->16", wgrad_case, (32, 8, 16, 3, 1, False)),
    ("wgrad proj 1x1 16->16", wgrad_case, (32, 16, 16, 1, 1, True)),
    ("wgrad proj 1x1 s2 16->32", wgrad_case, (32, 16, 32, 1, 2, True)),
    ("wgrad 3x3 s2 16->32", wgrad_case, (32, 16, 32, 3, 2, True)),
    ("wgrad 3x3 s2 32->64", wgrad_case, (16, 32, 64, 3, 2, True)),
]

if __name__ == "__main__":
    print("lib:", nat.LIB_PATH)
    for name, mk, args in CASES:
        fn, keep = mk(*args)
        print(f"{name:28s} {timeit(fn):8.2f} us")
