"""fp32-storage / tf32-math path of the native kernels (``DTF_F32``: ``kind::tf32`` tcgen05 MMAs on 4-element chunks)
against plain PyTorch fp32 references, and the ResNet engine in that mode against the fp32 oracle.

Parity: the reference's default compute dtype is fp32 (``official/utils/flags/_performance.py:107``; the PBT members
never pass ``--dtype``).  Tolerances here are those of tf32 products with fp32 accumulation (about 1e-3 relative per
product), several times tighter than the bf16 tests in ``test_kernels_gpu.py``.
"""
import math
import os
import sys

import pytest
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))      # helpers of the bf16 kernel tests, any import mode
from test_kernels_gpu import _bn_bwd_ref, bn_consts, conv_ref, rnd, stats_of  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from distributedtf_b200.ops import native
    native.require()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return native


def close(got, want, rtol, atol, what):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    bad = (err > atol + rtol * want.abs()).sum().item()
    assert bad == 0, f"{what}: {bad}/{got.numel()} off, max err {err.max().item():.4g}, ref max {want.abs().max().item():.4g}"


FWD = [
    # NB, H, Cin, Cout, k, stride, prologue, epilogue
    (2, 8, 16, 16, 1, 1, "none", "store"),
    (4, 32, 16, 16, 3, 1, "bn", "residual"),
    (3, 8, 64, 64, 3, 1, "bn", "residual"),          # 144 weight chunks: the largest B operand of ResNet-32
    (4, 32, 16, 32, 3, 2, "bn", "store"),
    (4, 32, 16, 32, 1, 2, "bn", "store"),
    (4, 32, 8, 16, 3, 1, "none", "store"),           # conv0: 3 real channels in 8
    (3, 8, 256, 64, 1, 1, "bn", "store"),
    (3, 8, 64, 256, 1, 1, "bn", "residual"),
    (128, 16, 32, 32, 3, 1, "bn", "residual"),       # a benchmarked shape: several tiles per CTA
]


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride,pro,epi", FWD)
def test_conv_forward_tf32(nat, NB, H, Cin, Cout, k, stride, pro, epi):
    Ho = H // stride
    x = rnd(NB, H, H, Cin, seed=1) + 0.3
    if Cin == 8:
        x[..., 3:] = 0
    w = rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2)
    gamma, beta = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
    a, bn, mode = x, None, nat.PRO_NONE
    run_mean, run_var = torch.zeros(Cin, device="cuda"), torch.ones(Cin, device="cuda")
    if pro == "bn":
        _, _, scale, shift = bn_consts(x, gamma, beta)
        a = torch.relu(x * scale + shift)
        bn = nat.bn_ref(Cin, NB * H * H, stats=stats_of(x), gamma=gamma, beta=beta, run_mean=run_mean, run_var=run_var)
        mode = nat.PRO_BN_RELU
    want = conv_ref(a, w, stride)
    res = None
    if epi == "residual":
        res = rnd(NB, Ho, Ho, Cout, seed=5)
        want = want + res
    out = torch.full((NB, Ho, Ho, Cout), float("nan"), device="cuda")
    stats_out = torch.zeros(2 * Cout, device="cuda")
    g = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mode, bn=bn, update_running=pro == "bn")
    assert g.dtype == nat.F32
    nat.conv_gemm(g, nat.pack_weight_host(w, dtype=torch.float32), out, NB=NB, Ho=Ho, Wo=Ho, N=Cout,
                  epi_mode=nat.EPI_RESIDUAL if res is not None else nat.EPI_STORE, epi_res=res, epi_stats_out=stats_out)
    torch.cuda.synchronize()
    close(out, want, 4e-3, 4e-3, "conv out (tf32)")
    wf = want.reshape(-1, Cout)
    close(stats_out[:Cout], wf.sum(0), 2e-3, 0.01 * math.sqrt(wf.shape[0]), "sum")
    close(stats_out[Cout:], (wf * wf).sum(0), 5e-3, 0.2, "sumsq")
    if pro == "bn":
        xf = x.reshape(-1, Cin)
        close(run_mean, 0.003 * xf.mean(0), 1e-3, 1e-5, "moving mean")


DGRAD = [
    (2, 8, 16, 16, 3, 1),
    (4, 32, 16, 32, 3, 2),
    (4, 32, 16, 32, 1, 2),
    (3, 8, 64, 64, 3, 1),
    (3, 8, 64, 256, 1, 1),
    (3, 16, 128, 256, 1, 2),
]


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride", DGRAD)
@pytest.mark.parametrize("pro", ["none", "bn_bwd"])
def test_conv_dgrad_tf32(nat, NB, H, Cin, Cout, k, stride, pro):
    Ho = H // stride
    x_in = rnd(NB, H, H, Cin, seed=1) + 0.2
    w = rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2)
    g_in, b_in = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
    gy = rnd(NB, Ho, Ho, Cout, seed=5)
    gyp, src2, bwd, bn_out, mode = gy, None, None, None, nat.PRO_NONE
    if pro == "bn_bwd":
        c_out = rnd(NB, Ho, Ho, Cout, seed=6) + 0.1
        g_o, b_o = 1 + 0.1 * rnd(Cout, seed=7), 0.1 * rnd(Cout, seed=8)
        dx, s1, s2 = _bn_bwd_ref(gy.reshape(-1, Cout), c_out.reshape(-1, Cout), g_o)
        gyp = dx.reshape(NB, Ho, Ho, Cout)
        bwd, src2 = torch.cat([s1, s2]).contiguous(), c_out
        bn_out = nat.bn_ref(Cout, NB * Ho * Ho, stats=stats_of(c_out), gamma=g_o, beta=b_o)
        mode = nat.PRO_BN_BWD
    a = torch.zeros(NB, H, H, Cin, device="cuda", requires_grad=True)
    conv_ref(a, w, stride).backward(gyp)
    mean, invstd, scale, shift = bn_consts(x_in, g_in, b_in)
    pre = x_in * scale + shift
    mask = pre > 0
    want = a.grad * mask
    xhat = (x_in - mean) * invstd
    out = torch.full((NB, H, H, Cin), float("nan"), device="cuda")
    stats_out = torch.zeros(2 * Cin, device="cuda")
    g = nat.gather(gy, Ho, Ho, Cout, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, transposed=True, mode=mode,
                   bn=bn_out, src2=src2, bwd_stats=bwd)
    epi_bn = nat.bn_ref(Cin, NB * H * H, stats=stats_of(x_in), gamma=g_in, beta=b_in)
    wpk = nat.pack_weight_host(w, transpose_for_dgrad=True, dtype=torch.float32)
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in, epi_bn=epi_bn,
                  epi_stats_out=stats_out)
    torch.cuda.synchronize()
    sure = pre.abs() > 1e-4                          # elements whose mask does not hinge on the last fp32 bit
    close(out[sure], want[sure], 5e-3, 5e-3, "dgrad (tf32)")
    P = NB * H * H
    close(stats_out[:Cin], want.reshape(-1, Cin).sum(0), 5e-3, 0.01 * math.sqrt(P), "sum t")
    close(stats_out[Cin:], (want * xhat).reshape(-1, Cin).sum(0), 5e-3, 0.01 * math.sqrt(P), "sum t xhat")
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in, epi_bn=epi_bn,
                  epi_stats_out=stats_out, epi_accumulate=True)
    torch.cuda.synchronize()
    close(out[sure], 2 * want[sure], 5e-3, 1e-2, "dgrad accumulate (tf32)")


WGRAD = [
    (2, 8, 16, 16, 1, 1),
    (4, 32, 16, 16, 3, 1),
    (4, 32, 8, 16, 3, 1),           # conv0: 18 chunks of four channels, padded to 20
    (4, 32, 16, 32, 3, 2),
    (4, 32, 16, 32, 1, 2),
    (7, 8, 64, 64, 3, 1),           # 144 chunks = 9 column blocks
    (3, 8, 64, 256, 1, 1),          # two 128-channel halves
    (3, 8, 256, 64, 1, 1),
    (128, 16, 32, 32, 3, 1),
]


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride", WGRAD)
@pytest.mark.parametrize("pro", ["none", "both"])
def test_conv_wgrad_tf32(nat, NB, H, Cin, Cout, k, stride, pro):
    Ho = H // stride
    x = rnd(NB, H, H, Cin, seed=1) + 0.2
    gy = rnd(NB, Ho, Ho, Cout, scale=0.1, seed=5)
    a, gyp = x, gy
    bn_x = bn_o = src2 = bwd = None
    mx = mg = nat.PRO_NONE
    if pro == "both":
        g_in, b_in = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
        _, _, scale, shift = bn_consts(x, g_in, b_in)
        a = torch.relu(x * scale + shift)
        bn_x = nat.bn_ref(Cin, NB * H * H, stats=stats_of(x), gamma=g_in, beta=b_in)
        mx = nat.PRO_BN_RELU
        c_out = rnd(NB, Ho, Ho, Cout, seed=6) + 0.1
        g_o, b_o = 1 + 0.1 * rnd(Cout, seed=7), 0.1 * rnd(Cout, seed=8)
        dx, s1, s2 = _bn_bwd_ref(gy.reshape(-1, Cout), c_out.reshape(-1, Cout), g_o)
        gyp = dx.reshape(NB, Ho, Ho, Cout)
        bwd, src2 = torch.cat([s1, s2]).contiguous(), c_out
        bn_o = nat.bn_ref(Cout, NB * Ho * Ho, stats=stats_of(c_out), gamma=g_o, beta=b_o)
        mg = nat.PRO_BN_BWD
    w = torch.zeros(Cout, k, k, Cin, device="cuda", requires_grad=True)
    conv_ref(a, w, stride).backward(gyp)
    want = w.grad
    dw = torch.zeros(Cout, k, k, Cin, device="cuda")
    gx = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mx, bn=bn_x)
    gg = nat.gather(gy, Ho, Ho, Cout, mode=mg, bn=bn_o, src2=src2, bwd_stats=bwd)
    nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=Ho, Wo=Ho, Cout=Cout)
    torch.cuda.synchronize()
    ref = want.abs().max().item()
    close(dw, want, 5e-3, 3e-3 * ref + 1e-4, "wgrad (tf32)")
    nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=Ho, Wo=Ho, Cout=Cout)
    torch.cuda.synchronize()
    close(dw, 2 * want, 5e-3, 6e-3 * ref + 2e-4, "wgrad x2 (tf32)")


def test_bn_bwd_apply_and_preprocess_fp32(nat):
    P, Cc = 4 * 16 * 16, 32
    t, x, res = rnd(P, Cc, seed=1), rnd(P, Cc, seed=2) + 0.2, rnd(P, Cc, seed=3)
    gamma, beta = 1 + 0.1 * rnd(Cc, seed=4), 0.1 * rnd(Cc, seed=5)
    dx, s1, s2 = _bn_bwd_ref(t, x, gamma)
    out = torch.empty(P, Cc, device="cuda")
    bn = nat.bn_ref(Cc, P, stats=stats_of(x), gamma=gamma, beta=beta)
    nat.bn_bwd_apply(t, x, res, out, bn, torch.cat([s1, s2]).contiguous(), P, Cc)
    torch.cuda.synchronize()
    close(out, dx + res, 1e-4, 1e-4, "bn_bwd_apply fp32")
    g = torch.Generator().manual_seed(3)
    imgs = torch.randint(0, 256, (8, 32, 32, 3), dtype=torch.uint8, generator=g).cuda()
    o32 = torch.empty(8, 32, 32, 8, device="cuda")
    o16 = torch.empty(8, 32, 32, 8, device="cuda", dtype=torch.bfloat16)
    nat.preprocess(imgs, None, o32, mode=1)
    nat.preprocess(imgs, None, o16, mode=1)
    torch.cuda.synchronize()
    assert (o32[..., 3:] == 0).all()
    close(o32, o16.float(), 1e-2, 1e-2, "preprocess fp32 vs bf16")
    xf = imgs.float()
    mean = xf.mean(dim=(1, 2, 3), keepdim=True)
    std = xf.std(dim=(1, 2, 3), keepdim=True, unbiased=False)
    close(o32[..., :3], (xf - mean) / std, 1e-4, 1e-4, "per-image standardisation")


@pytest.mark.parametrize("version,bottleneck,size,nb", [(2, False, 8, 32), (2, False, 14, 64), (2, True, 11, 32),
                                                        (1, False, 8, 32), (1, True, 11, 32)])
def test_native_tf32_step_matches_fp32_oracle(nat, version, bottleneck, size, nb):
    """One full training step with fp32 tensors / tf32 MMAs against autograd on the fp32 oracle.  The gradient error
    is ReLU-mask noise, as in the bf16 test, but from tf32's 1e-3 instead of bf16's 4e-3 relative error: a
    pre-activation flips its mask when it lies that close to zero, each flip is a full-size error, so the relative
    gradient error goes like the square root of the operand precision (measured 3-10 % here against 13-33 % for the
    bf16 engine on the same networks) -- every tensor is held to 15 % and a cosine of 0.985."""
    from distributedtf_b200.models import resnet
    from distributedtf_b200.models.resnet_native import NativeResNetEngine
    from distributedtf_b200.models.resnet_native_bottleneck import (NativeResNetBottleneckEngine,
                                                                    NativeResNetBottleneckV1Engine)
    from distributedtf_b200.models.resnet_native_v1 import NativeResNetV1Engine
    from distributedtf_b200.utils.data import cifar_preprocess_torch
    cfg = resnet.cifar10_config(size, resnet_version=version, bottleneck=bottleneck)
    cls = {(2, False): NativeResNetEngine, (2, True): NativeResNetBottleneckEngine,
           (1, False): NativeResNetV1Engine, (1, True): NativeResNetBottleneckV1Engine}[(version, bottleneck)]
    hp = {"opt_case": {"optimizer": "Momentum", "lr": 0.1, "momentum": 0.9}, "decay_steps": 20, "decay_rate": 0.1,
          "weight_decay": 2e-4, "regularizer": "l2_regularizer", "initializer": "he_init", "batch_size": nb}
    eng = cls(cfg, hp, "cuda", seed=5, use_graphs=False, dtype="tf32")
    assert eng.image.dtype == torch.float32 and eng.pack.dtype == torch.float32
    named = {k: v.clone() for k, v in eng.export_named().items()}
    g = torch.Generator().manual_seed(1)
    imgs = torch.randint(0, 256, (nb, 32, 32, 3), dtype=torch.uint8, generator=g).cuda()
    labels = torch.randint(0, 10, (nb,), generator=g).cuda()
    loss = eng.train_step_u8(imgs, labels, augment=False)
    torch.cuda.synchronize()
    x = cifar_preprocess_torch(imgs, False).permute(0, 3, 1, 2)
    params = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "moving" not in k) for k, v in named.items()}
    logits = resnet.forward(cfg, params, x, True)
    ref_loss = F.cross_entropy(logits, labels)
    ref_loss.backward()
    close(eng.logits[:nb], logits.detach(), 1e-2, 1e-2, "logits")
    close(loss, ref_loss.detach().reshape(1), 3e-3, 3e-3, "loss")
    grads = eng.export_named(eng.grads)
    report, worst, min_cos = [], 0.0, 1.0
    for k, p in params.items():
        if p.requires_grad:
            err = ((grads[k] - p.grad).norm() / (p.grad.norm() + 1e-8)).item()
            worst = max(worst, err)
            min_cos = min(min_cos, F.cosine_similarity(grads[k].reshape(1, -1), p.grad.reshape(1, -1)).item())
            report.append(f"{k}:{err:.4f}")
    print(f"tf32 resnet-{size} relative gradient errors:", " ".join(report))
    for k in ("dense/kernel", "dense/bias"):
        assert ((grads[k] - params[k].grad).norm() / params[k].grad.norm()).item() < 2e-3, k
    assert worst < 0.15 and min_cos > 0.985, f"worst {worst:.4f} min cosine {min_cos:.4f}: " + " ".join(report)
    after = eng.export_named()
    for k in named:
        if "moving" in k:
            close(after[k], params[k], 5e-3, 5e-4, k)
    # graph replay keeps training; evaluation counts hits
    eng2 = cls(cfg, hp, "cuda", seed=5, use_graphs=True, dtype="fp32")
    losses = [eng2.train_step_u8(imgs, labels, augment=False).clone() for _ in range(25)]
    torch.cuda.synchronize()
    assert losses[-1].item() < 0.7 * losses[0].item(), [l.item() for l in losses]
    assert 0 <= eng2.eval_batch_u8(imgs, labels).item() <= nb
