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


def fwd_res_case(H, C):
    """3x3 / stride-1 layer as it runs in a block: BN+ReLU prologue, residual epilogue, statistics."""
    x, res = bf(NB, H, H, C), bf(NB, H, H, C)
    w = torch.randn(C, 3, 3, C, device=dev) / math.sqrt(9 * C)
    wpk = nat.pack_weight_host(w)
    out = torch.empty(NB, H, H, C, device=dev, dtype=torch.bfloat16)
    st = stats_of(x)
    ones, zeros = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    bn = nat.bn_ref(C, NB * H * H, stats=st, gamma=ones, beta=zeros)
    so = torch.zeros(2 * C, device=dev)
    g = nat.gather(x, H, H, C, KH=3, KW=3, stride=1, pad=1, mode=nat.PRO_BN_RELU, bn=bn)
    return lambda: nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=C, epi_mode=nat.EPI_RESIDUAL, epi_res=res,
                                 epi_stats_out=so), [x, res, st, ones, zeros]


def bwd_fused_case(H, C):
    """dgrad + wgrad of a 3x3 / stride-1 layer in one launch (BN-backward prologue on gy, ReLU-mask epilogue)."""
    x_in, c_out, gy = bf(NB, H, H, C), bf(NB, H, H, C), bf(NB, H, H, C)
    w = torch.randn(C, 3, 3, C, device=dev) / math.sqrt(9 * C)
    wd = nat.pack_weight_host(w, transpose_for_dgrad=True)
    st_i, st_o = stats_of(x_in), stats_of(c_out)
    ones, zeros = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    bwd = torch.zeros(2 * C, device=dev)
    bn_i = nat.bn_ref(C, NB * H * H, stats=st_i, gamma=ones, beta=zeros)
    bn_o = nat.bn_ref(C, NB * H * H, stats=st_o, gamma=ones, beta=zeros)
    out = torch.empty(NB, H, H, C, device=dev, dtype=torch.bfloat16)
    so = torch.zeros(2 * C, device=dev)
    dw = torch.zeros(C, 3, 3, C, device=dev)
    gd = nat.gather(gy, H, H, C, KH=3, KW=3, stride=1, pad=1, transposed=True, mode=nat.PRO_BN_BWD, bn=bn_o,
                    src2=c_out, bwd_stats=bwd)
    dgrad = nat.conv_gemm_params(gd, wd, out, NB=NB, Ho=H, Wo=H, N=C, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in,
                                 epi_bn=bn_i, epi_stats_out=so)
    gx = nat.gather(x_in, H, H, C, KH=3, KW=3, stride=1, pad=1, mode=nat.PRO_BN_RELU, bn=bn_i)
    gg = nat.gather(gy, H, H, C, mode=nat.PRO_BN_BWD, bn=bn_o, src2=c_out, bwd_stats=bwd)
    # the structs hold raw pointers: everything they point at must outlive the launches
    return lambda: nat.conv_bwd(dgrad, gx, gg, dw, NB=NB, Ho=H, Wo=H, Cout=C), \
        [x_in, c_out, gy, st_i, st_o, ones, zeros, bwd, wd, out, so, dw]


def bn_bwd_case(H, C):
    t, x, res = bf(NB, H, H, C), bf(NB, H, H, C), bf(NB, H, H, C)
    st = stats_of(x)
    ones, zeros = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    bwd = torch.zeros(2 * C, device=dev)
    bn = nat.bn_ref(C, NB * H * H, stats=st, gamma=ones, beta=zeros)
    out = torch.empty(NB, H, H, C, device=dev, dtype=torch.bfloat16)
    return lambda: nat.bn_bwd_apply(t, x, res, out, bn, bwd, NB * H * H, C), [t, x, res, st, ones, zeros, bwd, out]


CASES = [
    ("fwd 3x3 16ch 32x32 (x10)", fwd_res_case, (32, 16)),
    ("fwd 3x3 32ch 16x16 (x9)", fwd_res_case, (16, 32)),
    ("fwd 3x3 64ch 8x8 (x9)", fwd_res_case, (8, 64)),
    ("bwd fused 16ch 32x32 (x9)", bwd_fused_case, (32, 16)),
    ("bwd fused 32ch 16x16 (x9)", bwd_fused_case, (16, 32)),
    ("bwd fused 64ch 8x8 (x9)", bwd_fused_case, (8, 64)),
    ("bn_bwd_apply 16ch 32x32 (x5)", bn_bwd_case, (32, 16)),
    ("bn_bwd_apply 32ch 16x16 (x5)", bn_bwd_case, (16, 32)),
    ("bn_bwd_apply 64ch 8x8 (x5)", bn_bwd_case, (8, 64)),
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
