#!/usr/bin/env python
"""Where do the fp32 / tf32 conv kernels disagree with the reference?  (development aid for tests/test_tf32_gpu.py)"""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from distributedtf_b200.ops import native as nat
from test_kernels_gpu import bn_consts, conv_ref, rnd, stats_of
nat.require()
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def report(name, got, want, tol=1e-2):
    err = (got - want).abs()
    bad = err > tol + tol * want.abs()
    print(f"{name}: {int(bad.sum())}/{bad.numel()} off, max err {err.max().item():.4g}, nan {int(torch.isnan(got).sum())}")
    if bad.any():
        idx = bad.nonzero()
        print("   first bad idx:", idx[:6].tolist())
        print("   bad per channel:", bad.reshape(-1, bad.shape[-1]).sum(0).tolist()[:64])
        rows = bad.reshape(-1, bad.shape[-1]).any(1).nonzero().flatten()
        print("   bad rows (first 20):", rows[:20].tolist(), "count", rows.numel(), "of", bad.reshape(-1, bad.shape[-1]).shape[0])
        i = tuple(idx[0].tolist())
        print("   got", got[i[:-1]][:8].tolist(), "\n   want", want[i[:-1]][:8].tolist())


def fwd(NB, H, Cin, Cout, k, stride, pro, dtype):
    Ho = H // stride
    x = (rnd(NB, H, H, Cin, seed=1) + 0.3).to(dtype)
    w = rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2).to(dtype)
    gamma, beta = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
    a, bn, mode = x.float(), None, nat.PRO_NONE
    if pro == "bn":
        _, _, scale, shift = bn_consts(x, gamma, beta)
        a = torch.relu(x.float() * scale + shift).to(dtype).float()
        bn = nat.bn_ref(Cin, NB * H * H, stats=stats_of(x), gamma=gamma, beta=beta, run_mean=torch.zeros(Cin, device="cuda"),
                        run_var=torch.ones(Cin, device="cuda"))
        mode = nat.PRO_BN_RELU
    want = conv_ref(a, w.float(), stride)
    out = torch.full((NB, Ho, Ho, Cout), float("nan"), device="cuda", dtype=dtype)
    g = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mode, bn=bn)
    nat.conv_gemm(g, nat.pack_weight_host(w.float(), dtype=dtype), out, NB=NB, Ho=Ho, Wo=Ho, N=Cout)
    torch.cuda.synchronize()
    report(f"fwd {dtype} NB{NB} H{H} {Cin}->{Cout} k{k} s{stride} {pro}", out.float(), want, 1e-2 if dtype == torch.float32 else 3e-2)


for dt in (torch.float32, torch.bfloat16):
    fwd(4, 32, 16, 32, 1, 2, "bn", dt)
    fwd(4, 32, 16, 32, 1, 2, "none", dt)
    fwd(4, 32, 16, 32, 1, 1, "bn", dt)
    fwd(4, 16, 16, 32, 1, 1, "bn", dt)
    fwd(3, 8, 64, 256, 1, 1, "bn", dt)
    fwd(3, 8, 64, 256, 1, 1, "none", dt)
    fwd(3, 8, 64, 128, 1, 1, "bn", dt)
    fwd(2, 8, 16, 16, 1, 1, "bn", dt)
