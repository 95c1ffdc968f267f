"""sm_100a kernels vs plain PyTorch fp32 references of the same ops (run with -m gpu)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

EPS = 1e-5


@pytest.fixture(scope="module")
def nat():
    from distributedtf_b200.ops import native
    native.require()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return native


def bf(t):
    return t.to(torch.bfloat16)


def rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", generator=g) * scale


def stats_of(x_nhwc_bf16):
    xf = x_nhwc_bf16.float().reshape(-1, x_nhwc_bf16.shape[-1])
    return torch.cat([xf.sum(0), (xf * xf).sum(0)]).contiguous()


def bn_consts(x_bf16, gamma, beta):
    xf = x_bf16.float().reshape(-1, x_bf16.shape[-1])
    mean = xf.mean(0)
    var = xf.var(0, unbiased=False)
    invstd = torch.rsqrt(var + EPS)
    return mean, invstd, gamma * invstd, beta - mean * gamma * invstd


def conv_ref(a_nhwc_f32, w_ohwi_bf16, stride):
    k = w_ohwi_bf16.shape[1]
    x = a_nhwc_f32.permute(0, 3, 1, 2)
    w = w_ohwi_bf16.float().permute(0, 3, 1, 2)
    if stride > 1:
        tot = k - 1
        beg = tot // 2
        x = F.pad(x, (beg, tot - beg, beg, tot - beg))
        y = F.conv2d(x, w, stride=stride)
    else:
        y = F.conv2d(x, w, padding=k // 2)
    return y.permute(0, 2, 3, 1).contiguous()


def assert_close(got, want, rtol=2e-2, atol=2e-2, what=""):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    bad = (err > tol).sum().item()
    assert bad == 0, f"{what}: {bad}/{got.numel()} mismatches, max err {err.max().item():.4g}, ref max {want.abs().max().item():.4g}"


FWD_CASES = [
    # NB, H, Cin, Cout, k, stride, prologue, epilogue
    (2, 8, 16, 16, 1, 1, "none", "store"),
    (2, 8, 16, 16, 3, 1, "none", "store"),
    (4, 32, 16, 16, 3, 1, "bn", "residual"),
    (3, 8, 64, 64, 3, 1, "bn", "residual"),
    (4, 32, 16, 32, 3, 2, "bn", "store"),
    (4, 32, 16, 32, 1, 2, "bn", "store"),
    (4, 16, 32, 64, 3, 2, "bn", "store"),
    (5, 16, 32, 32, 3, 1, "bn", "residual"),
    (4, 32, 8, 16, 3, 1, "none", "store"),          # conv0: 3 real channels padded to 8, odd chunk count
    (2, 28, 8, 32, 5, 1, "none", "bias_relu"),      # MNIST conv1 (1 real channel)
    (2, 14, 32, 64, 5, 1, "none", "bias_relu"),     # MNIST conv2
    # the benchmarked shapes (batch 128 / 256): several tiles per CTA, both epilogue sets, every gather job shape
    (128, 32, 16, 16, 3, 1, "bn", "residual"),
    (128, 16, 32, 32, 3, 1, "bn", "residual"),
    (256, 8, 64, 64, 3, 1, "bn", "residual"),
    # bottleneck blocks: 1x1 reduce / expand with up to 256 channels, strided projection, 3x3 stride 2 with Cin == Cout
    (3, 8, 256, 64, 1, 1, "bn", "store"),
    (3, 8, 64, 256, 1, 1, "bn", "residual"),
    (3, 16, 128, 256, 1, 2, "bn", "store"),
    (4, 16, 32, 32, 3, 2, "bn", "store"),
    (128, 32, 16, 32, 3, 2, "bn", "store"),
    (128, 32, 16, 16, 1, 1, "bn", "store"),
]


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride,pro,epi", FWD_CASES)
def test_conv_forward(nat, NB, H, Cin, Cout, k, stride, pro, epi):
    Ho = H // stride
    x = bf(rnd(NB, H, H, Cin, seed=1) + 0.3)
    if Cin == 8:
        x[..., (3 if k == 3 else 1):] = 0
    w = bf(rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2))
    gamma, beta = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
    stats = stats_of(x)
    a = x.float()
    bn = None
    mode = nat.PRO_NONE
    run_mean, run_var = torch.zeros(Cin, device="cuda"), torch.ones(Cin, device="cuda")
    if pro == "bn":
        mean, invstd, scale, shift = bn_consts(x, gamma, beta)
        a = bf(torch.relu(x.float() * scale + shift)).float()
        bn = nat.bn_ref(Cin, NB * H * H, stats=stats, gamma=gamma, beta=beta, run_mean=run_mean, run_var=run_var)
        mode = nat.PRO_BN_RELU
    want = conv_ref(a, w, stride)
    res = bias = None
    epi_mode = nat.EPI_STORE
    if epi == "residual":
        res = bf(rnd(NB, Ho, Ho, Cout, seed=5))
        want = want + res.float()
        epi_mode = nat.EPI_RESIDUAL
    elif epi == "bias_relu":
        bias = 0.1 * rnd(Cout, seed=6)
        want = torch.relu(want + bias)
        epi_mode = nat.EPI_BIAS_RELU
    wpk = nat.pack_weight_host(w.float())
    out = torch.full((NB, Ho, Ho, Cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    stats_out = torch.zeros(2 * Cout, device="cuda")
    g = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mode, bn=bn,
                   update_running=(pro == "bn"))
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=Ho, Wo=Ho, N=Cout, epi_mode=epi_mode, epi_res=res, epi_bias=bias,
                  epi_stats_out=stats_out)
    torch.cuda.synchronize()
    assert_close(out, want, what="conv out")
    wf = want.reshape(-1, Cout)
    assert_close(stats_out[:Cout], wf.sum(0), rtol=1e-2, atol=0.05 * math.sqrt(wf.shape[0]), what="sum")
    assert_close(stats_out[Cout:], (wf * wf).sum(0), rtol=2e-2, atol=1.0, what="sumsq")
    if pro == "bn":
        xf = x.float().reshape(-1, Cin)
        assert_close(run_mean, 0.003 * xf.mean(0), rtol=1e-3, atol=1e-5, what="moving mean")
        assert_close(run_var, 0.997 + 0.003 * xf.var(0, unbiased=True), rtol=1e-3, atol=1e-5, what="moving var")


def test_conv_forward_inference_bn(nat):
    NB, H, Cc = 2, 16, 32
    x = bf(rnd(NB, H, H, Cc, seed=1))
    w = bf(rnd(Cc, 3, 3, Cc, scale=0.06, seed=2))
    gamma, beta = 1 + 0.1 * rnd(Cc, seed=3), 0.1 * rnd(Cc, seed=4)
    rm, rv = 0.2 * rnd(Cc, seed=5), 1 + 0.3 * torch.rand(Cc, device="cuda")
    scale = gamma * torch.rsqrt(rv + EPS)
    a = bf(torch.relu(x.float() * scale + (beta - rm * scale))).float()
    want = conv_ref(a, w, 1)
    out = torch.empty(NB, H, H, Cc, device="cuda", dtype=torch.bfloat16)
    bn = nat.bn_ref(Cc, NB * H * H, gamma=gamma, beta=beta, run_mean=rm, run_var=rv, use_running=True)
    g = nat.gather(x, H, H, Cc, KH=3, KW=3, stride=1, pad=1, mode=nat.PRO_BN_RELU, bn=bn)
    nat.conv_gemm(g, nat.pack_weight_host(w.float()), out, NB=NB, Ho=H, Wo=H, N=Cc)
    torch.cuda.synchronize()
    assert_close(out, want, what="inference conv")


DGRAD_CASES = [
    # NB, Hin, Cin, Cout, k, stride
    (2, 8, 16, 16, 3, 1),
    (4, 32, 16, 16, 3, 1),
    (4, 32, 16, 32, 3, 2),
    (4, 32, 16, 32, 1, 2),
    (3, 16, 32, 64, 3, 2),
    (5, 8, 32, 64, 3, 2),        # fewer pixels per parity class than one tile
    (5, 8, 32, 64, 1, 2),
    (3, 8, 64, 64, 3, 1),
    (128, 32, 16, 16, 3, 1),     # benchmarked shapes
    (128, 16, 32, 32, 3, 1),
    (256, 8, 64, 64, 3, 1),
    (128, 32, 16, 32, 3, 2),
    (3, 8, 256, 64, 1, 1),       # bottleneck blocks: wide 1x1 layers, strided projection, 3x3 stride 2 with Cin == Cout
    (3, 8, 64, 256, 1, 1),
    (3, 16, 128, 256, 1, 2),
    (4, 16, 32, 32, 3, 2),
    (4, 16, 64, 64, 3, 2),
]


def _bn_bwd_ref(t, x, gamma):
    """dL/dx of y = gamma*xhat + beta given t = dL/dy (all [P, C] fp32)."""
    mean = x.mean(0)
    invstd = torch.rsqrt(x.var(0, unbiased=False) + EPS)
    xhat = (x - mean) * invstd
    return gamma * invstd * (t - t.mean(0) - xhat * (t * xhat).mean(0)), t.sum(0), (t * xhat).sum(0)


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride", DGRAD_CASES)
@pytest.mark.parametrize("pro", ["none", "bn_bwd"])
def test_conv_dgrad(nat, NB, H, Cin, Cout, k, stride, pro):
    """gx = conv^T(gy') masked by the ReLU of the input BN, with (sum t, sum t*xhat) statistics."""
    Ho = H // stride
    x_in = bf(rnd(NB, H, H, Cin, seed=1) + 0.2)                 # pre-BN input of the conv
    w = bf(rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2))
    g_in, b_in = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
    gy = bf(rnd(NB, Ho, Ho, Cout, seed=5))
    gyp = gy.float()
    src2 = bwd = bn_out = None
    mode = nat.PRO_NONE
    if pro == "bn_bwd":                                          # gy' = BN-backward(t = gy, x = c_out)
        c_out = bf(rnd(NB, Ho, Ho, Cout, seed=6) + 0.1)
        g_o, b_o = 1 + 0.1 * rnd(Cout, seed=7), 0.1 * rnd(Cout, seed=8)
        dx, s1, s2 = _bn_bwd_ref(gy.float().reshape(-1, Cout), c_out.float().reshape(-1, Cout), g_o)
        gyp = bf(dx).float().reshape(NB, Ho, Ho, Cout)
        bwd = torch.cat([s1, s2]).contiguous()
        src2 = c_out
        st_o = stats_of(c_out)
        bn_out = nat.bn_ref(Cout, NB * Ho * Ho, stats=st_o, gamma=g_o, beta=b_o)
        mode = nat.PRO_BN_BWD
    # reference: autograd through the fp32 conv
    a = torch.zeros(NB, H, H, Cin, device="cuda", requires_grad=True)
    conv_ref(a, w, stride).backward(gyp)
    mean, invstd, scale, shift = bn_consts(x_in, g_in, b_in)
    mask = (x_in.float() * scale + shift) > 0
    want = a.grad * mask
    xhat = (x_in.float() - mean) * invstd
    wpk = nat.pack_weight_host(w.float(), transpose_for_dgrad=True)
    out = torch.full((NB, H, H, Cin), float("nan"), device="cuda", dtype=torch.bfloat16)
    stats_out = torch.zeros(2 * Cin, device="cuda")
    g = nat.gather(gy, Ho, Ho, Cout, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, transposed=True, mode=mode,
                   bn=bn_out, src2=src2, bwd_stats=bwd)
    st_in = stats_of(x_in)
    epi_bn = nat.bn_ref(Cin, NB * H * H, stats=st_in, gamma=g_in, beta=b_in)
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in, epi_bn=epi_bn,
                  epi_stats_out=stats_out)
    torch.cuda.synchronize()
    assert_close(out, want, what="dgrad")
    P = NB * H * H
    assert_close(stats_out[:Cin], want.reshape(-1, Cin).sum(0), rtol=2e-2, atol=0.05 * math.sqrt(P), what="sum t")
    assert_close(stats_out[Cin:], (want * xhat).reshape(-1, Cin).sum(0), rtol=2e-2, atol=0.05 * math.sqrt(P), what="sum t xhat")
    # accumulate: a second launch adds onto the stored tensor and onto the statistics
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x_in, epi_bn=epi_bn,
                  epi_stats_out=stats_out, epi_accumulate=True)
    torch.cuda.synchronize()
    assert_close(out, 2 * want, rtol=3e-2, atol=4e-2, what="dgrad accumulate")
    assert_close(stats_out[:Cin], 2 * want.reshape(-1, Cin).sum(0), rtol=2e-2, atol=0.1 * math.sqrt(P), what="sum t x2")


WGRAD_CASES = [
    # NB, Hin, Cin, Cout, k, stride
    (2, 8, 16, 16, 1, 1),
    (2, 8, 16, 16, 3, 1),
    (4, 32, 16, 16, 3, 1),
    (4, 32, 8, 16, 3, 1),
    (4, 32, 16, 32, 3, 2),
    (4, 32, 16, 32, 1, 2),
    (5, 16, 32, 32, 3, 1),
    (3, 16, 32, 64, 3, 2),
    (7, 8, 64, 64, 3, 1),
    (2, 14, 32, 64, 5, 1),
    (128, 32, 16, 16, 3, 1),     # benchmarked shapes
    (128, 16, 32, 32, 3, 1),
    (256, 8, 64, 64, 3, 1),
    (128, 16, 32, 64, 3, 2),
    (3, 8, 256, 64, 1, 1),       # bottleneck blocks; Cout 256 = two 128-channel halves (blockIdx.z)
    (3, 8, 64, 256, 1, 1),
    (3, 16, 128, 256, 1, 2),
    (4, 16, 64, 64, 3, 2),
]


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride", WGRAD_CASES)
@pytest.mark.parametrize("pro", ["none", "both"])
def test_conv_wgrad(nat, NB, H, Cin, Cout, k, stride, pro):
    Ho = H // stride
    x = bf(rnd(NB, H, H, Cin, seed=1) + 0.2)
    gy = bf(rnd(NB, Ho, Ho, Cout, scale=0.1, seed=5))
    a = x.float()
    gyp = gy.float()
    bn_x = bn_o = src2 = bwd = None
    mx = mg = nat.PRO_NONE
    if pro == "both":
        g_in, b_in = 1 + 0.1 * rnd(Cin, seed=3), 0.1 * rnd(Cin, seed=4)
        _, _, scale, shift = bn_consts(x, g_in, b_in)
        a = bf(torch.relu(x.float() * scale + shift)).float()
        st_x = stats_of(x)
        bn_x = nat.bn_ref(Cin, NB * H * H, stats=st_x, gamma=g_in, beta=b_in)
        mx = nat.PRO_BN_RELU
        c_out = bf(rnd(NB, Ho, Ho, Cout, seed=6) + 0.1)
        g_o, b_o = 1 + 0.1 * rnd(Cout, seed=7), 0.1 * rnd(Cout, seed=8)
        dx, s1, s2 = _bn_bwd_ref(gy.float().reshape(-1, Cout), c_out.float().reshape(-1, Cout), g_o)
        gyp = bf(dx).float().reshape(NB, Ho, Ho, Cout)
        bwd = torch.cat([s1, s2]).contiguous()
        src2 = c_out
        st_o = stats_of(c_out)
        bn_o = nat.bn_ref(Cout, NB * Ho * Ho, stats=st_o, gamma=g_o, beta=b_o)
        mg = nat.PRO_BN_BWD
    w = torch.zeros(Cout, k, k, Cin, device="cuda", requires_grad=True)
    k_ = k
    xin = a.permute(0, 3, 1, 2)
    if stride > 1:
        tot = k_ - 1
        beg = tot // 2
        y = F.conv2d(F.pad(xin, (beg, tot - beg, beg, tot - beg)), w.permute(0, 3, 1, 2), stride=stride)
    else:
        y = F.conv2d(xin, w.permute(0, 3, 1, 2), padding=k_ // 2)
    y.backward(gyp.permute(0, 3, 1, 2))
    want = w.grad
    dw = torch.zeros(Cout, k, k, Cin, device="cuda")
    gx = nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mx, bn=bn_x)
    gg = nat.gather(gy, Ho, Ho, Cout, mode=mg, bn=bn_o, src2=src2, bwd_stats=bwd)
    nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=Ho, Wo=Ho, Cout=Cout)
    torch.cuda.synchronize()
    scale_ref = want.abs().max().item()
    assert_close(dw, want, rtol=2e-2, atol=1e-2 * scale_ref + 1e-3, what="wgrad")
    # accumulation into an existing gradient buffer
    nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=Ho, Wo=Ho, Cout=Cout)
    torch.cuda.synchronize()
    assert_close(dw, 2 * want, rtol=2e-2, atol=2e-2 * scale_ref + 2e-3, what="wgrad x2")


GUARD = 4096            # elements of sentinel on each side of every output


def _guarded(shape, dtype, fill):
    """A tensor of `shape` embedded in a larger buffer whose margins hold a sentinel (compute-sanitizer is not
    available on the GPU pool: out-of-bounds writes are caught by canaries instead)."""
    n = 1
    for s in shape:
        n *= s
    buf = torch.full((n + 2 * GUARD,), fill, dtype=dtype, device="cuda")
    return buf, buf[GUARD: GUARD + n].view(*shape)


def _margins_intact(buf, fill):
    head, tail = buf[:GUARD].float(), buf[-GUARD:].float()
    return bool((head == fill).all()) and bool((tail == fill).all())


@pytest.mark.parametrize("NB,H,Cin,Cout,k,stride", [(3, 8, 16, 16, 3, 1), (3, 16, 16, 32, 3, 2), (3, 16, 16, 32, 1, 2),
                                                    (5, 8, 32, 64, 3, 2), (2, 14, 32, 64, 5, 1), (7, 8, 64, 64, 3, 1)])
def test_conv_kernels_stay_inside_their_outputs(nat, NB, H, Cin, Cout, k, stride):
    """Forward, input-gradient (halo, parity-class and generic paths) and weight-gradient kernels with odd batch
    sizes / partial tiles: every output sits between sentinel margins that must survive the launch."""
    Ho = H // stride
    x = bf(rnd(NB, H, H, Cin, seed=1))
    w = rnd(Cout, k, k, Cin, scale=1.0 / math.sqrt(k * k * Cin), seed=2)
    gy = bf(rnd(NB, Ho, Ho, Cout, seed=3))
    S = 7.0
    # forward with statistics
    obuf, out = _guarded((NB, Ho, Ho, Cout), torch.bfloat16, S)
    sbuf, so = _guarded((2 * Cout,), torch.float32, S)
    so.zero_()
    nat.conv_gemm(nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2), nat.pack_weight_host(w), out,
                  NB=NB, Ho=Ho, Wo=Ho, N=Cout, epi_stats_out=so)
    # input gradient (ReLU mask epilogue with statistics), then the accumulating variant
    st_in = stats_of(x)
    g_in, b_in = torch.ones(Cin, device="cuda"), torch.zeros(Cin, device="cuda")
    epi_bn = nat.bn_ref(Cin, NB * H * H, stats=st_in, gamma=g_in, beta=b_in)
    gbuf, gx = _guarded((NB, H, H, Cin), torch.bfloat16, S)
    gx.zero_()
    tbuf, ts = _guarded((2 * Cin,), torch.float32, S)
    ts.zero_()
    gg = nat.gather(gy, Ho, Ho, Cout, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, transposed=True)
    wd = nat.pack_weight_host(w, transpose_for_dgrad=True)
    for acc in (False, True):
        nat.conv_gemm(gg, wd, gx, NB=NB, Ho=H, Wo=H, N=Cin, epi_mode=nat.EPI_RELU_MASK, epi_x=x, epi_bn=epi_bn,
                      epi_stats_out=ts, epi_accumulate=acc)
    # weight gradient
    wbuf, dw = _guarded((Cout, k, k, Cin), torch.float32, S)
    dw.zero_()
    nat.conv_wgrad(nat.gather(x, H, H, Cin, KH=k, KW=k, stride=stride, pad=(k - 1) // 2), nat.gather(gy, Ho, Ho, Cout), dw,
                   NB=NB, Ho=Ho, Wo=Ho, Cout=Cout)
    torch.cuda.synchronize()
    for name, b in (("out", obuf), ("stats", sbuf), ("gx", gbuf), ("bwd stats", tbuf), ("dw", wbuf)):
        assert _margins_intact(b, S), f"{name}: a kernel wrote outside its output"
    assert torch.isfinite(out.float()).all() and torch.isfinite(gx.float()).all() and torch.isfinite(dw).all()


@pytest.mark.parametrize("mean_over_std", [1.0, 8.0, 30.0])
def test_bn_statistics_with_a_large_mean(nat, mean_over_std):
    """The conv epilogue accumulates (sum, sum of squares) in fp32 and the consumer forms var = E[x^2] - E[x]^2.
    Envelope of the cancellation error at the benchmarked size (131 072 values per channel): the variance of a
    tensor whose mean is 1x / 8x / 30x its standard deviation is still within 0.5 % / 0.5 % / 3 % of the fp64
    value (activations of this network sit below 3x)."""
    NB, H, Cc = 128, 32, 16
    x = bf(rnd(NB, H, H, Cc, seed=11) + mean_over_std)
    w = torch.zeros(Cc, 1, 1, Cc, device="cuda")
    w[torch.arange(Cc), 0, 0, torch.arange(Cc)] = 1.0                     # identity 1x1 conv: out == x exactly
    out = torch.empty_like(x)
    stats = torch.zeros(2 * Cc, device="cuda")
    g = nat.gather(x, H, H, Cc, KH=1, KW=1, stride=1, pad=0)
    nat.conv_gemm(g, nat.pack_weight_host(w), out, NB=NB, Ho=H, Wo=H, N=Cc, epi_stats_out=stats)
    torch.cuda.synchronize()
    assert torch.equal(out, x)
    M = NB * H * H
    mean = stats[:Cc].double() / M
    var = stats[Cc:].double() / M - mean * mean
    xf = x.double().reshape(-1, Cc)
    want_var = xf.var(0, unbiased=False)
    rel = ((var - want_var).abs() / want_var).max().item()
    print(f"mean/std = {mean_over_std}: worst relative variance error {rel:.2e}")
    assert ((mean - xf.mean(0)).abs() / xf.mean(0).abs()).max().item() < 1e-5
    assert rel < (0.005 if mean_over_std <= 8 else 0.03), rel
