"""The eight residual-block configurations of the reference's golden fixtures, run on the sm_100a kernels.

The fixtures themselves (reference ``official/utils/testing/reference_data/resnet/*``: batch 32, width 8, 4 channels,
{building, bottleneck} x {plain, projection (stride 2)} x {v1, v2}) live in the read-only reference tree and are
checked against the torch oracle on the CPU side (``tests/test_golden_fixtures.py``).  Here the SAME block function
(``models/resnet._block``) is walked with an ops object that issues the native kernels -- BatchNorm + ReLU in the
operand gather, BN statistics and the residual add in the epilogue, the v1 tail kernel -- with the 4 / 1 / 2 / 8
channels padded to 16, and its output is compared with the oracle on the same random parameters.  Together:
fixtures == oracle (CPU), oracle == native kernels (here).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

PAD = 16          # channel padding: operand chunks are 8 channels, output tiles multiples of 16


@pytest.fixture(scope="module")
def nat():
    from distributedtf_b200.ops import native
    native.require()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return native


class _Raw:
    """A stored bf16 NHWC tensor (padded channels) with the per-channel (sum, sum of squares) of its contents."""
    def __init__(self, t, stats, c_real):
        self.t, self.stats, self.c = t, stats, c_real


class _Bn:
    def __init__(self, raw, name):
        self.raw, self.name = raw, name


class _Relu:
    def __init__(self, bn):
        self.bn = bn


class _Pending:
    """The last convolution of a v2 block: issued by add() so that the shortcut rides in its epilogue."""
    def __init__(self, args):
        self.args = args


class _AddV1:
    def __init__(self, bn, shortcut):
        self.bn, self.shortcut = bn, shortcut


class NativeBlockOps:
    def __init__(self, nat, params, nb):
        self.nat, self.p, self.nb = nat, params, nb
        self._bn = self._conv = 0
        self.fused_add_next = False
        self.launches = 0

    # -- helpers
    def _pad_vec(self, v, fill):
        out = torch.full((PAD,), fill, device="cuda", dtype=torch.float32)
        out[: v.numel()] = v.cuda()
        return out

    def _bn_ref(self, node: _Bn):
        raw = node.raw
        h = raw.t.shape[1]
        gamma = self._pad_vec(self.p[node.name + "/gamma"], 1.0)
        beta = self._pad_vec(self.p[node.name + "/beta"], 0.0)
        keep = (gamma, beta, torch.zeros(PAD, device="cuda"), torch.ones(PAD, device="cuda"))
        self._keep = getattr(self, "_keep", []) + [keep]
        return self.nat.bn_ref(PAD, self.nb * h * h, stats=raw.stats, gamma=gamma, beta=beta, run_mean=keep[2],
                               run_var=keep[3])

    def _run_conv(self, x, w_ohwi, stride, res=None):
        nat = self.nat
        if isinstance(x, _Relu):
            src, mode, bn = x.bn.raw, nat.PRO_BN_RELU, self._bn_ref(x.bn)
        else:
            assert isinstance(x, _Raw), type(x)
            src, mode, bn = x, nat.PRO_NONE, None
        cout, k, _, cin = w_ohwi.shape
        w = torch.zeros(PAD, k, k, PAD, device="cuda")
        w[:cout, :, :, :cin] = w_ohwi.cuda()
        h = src.t.shape[1]
        ho = h // stride
        out = torch.full((self.nb, ho, ho, PAD), float("nan"), device="cuda", dtype=torch.bfloat16)
        stats = torch.zeros(2 * PAD, device="cuda")
        g = nat.gather(src.t, h, h, PAD, KH=k, KW=k, stride=stride, pad=(k - 1) // 2, mode=mode, bn=bn)
        nat.conv_gemm(g, nat.pack_weight_host(w), out, NB=self.nb, Ho=ho, Wo=ho, N=PAD,
                      epi_mode=nat.EPI_RESIDUAL if res is not None else nat.EPI_STORE,
                      epi_res=res.t if res is not None else None, epi_stats_out=stats)
        self.launches += 1
        return _Raw(out, stats, cout)

    # -- the interface models/resnet._block walks
    def conv(self, x, cin, cout, k, s):
        w = self.p[f"conv{self._conv}/kernel"]
        self._conv += 1
        if self.fused_add_next:
            self.fused_add_next = False
            return _Pending((x, w, s))
        return self._run_conv(x, w, s)

    def bn(self, x, c):
        name = f"bn{self._bn}"
        self._bn += 1
        return _Bn(x, name)

    def relu(self, x):
        if isinstance(x, _AddV1):                       # v1 tail: relu(bn(c) + shortcut) in one kernel
            nat, raw = self.nat, x.bn.raw
            h = raw.t.shape[1]
            out = torch.empty_like(raw.t)
            sc = x.shortcut
            if isinstance(sc, _Bn):
                nat.bn_act_apply(raw.t, self._bn_ref(x.bn), out, self.nb * h * h, PAD, sc=sc.raw.t,
                                 bn_sc=self._bn_ref(sc))
            else:
                nat.bn_act_apply(raw.t, self._bn_ref(x.bn), out, self.nb * h * h, PAD, sc=sc.t)
            self.launches += 1
            return _Raw(out, None, raw.c)
        return _Relu(x)

    def add(self, a, b):
        if isinstance(a, _Pending):                     # v2: residual add in the last convolution's epilogue
            x, w, s = a.args
            return self._run_conv(x, w, s, res=b)
        return _AddV1(a, b)


CASES = [(v, bott, proj) for v in (1, 2) for bott in (False, True) for proj in (False, True)]


@pytest.mark.parametrize("version,bottleneck,projection", CASES)
def test_fixture_block_on_native_kernels(nat, version, bottleneck, projection):
    from distributedtf_b200.models import resnet
    nb, width, channels = 32, 8, 4
    if projection:
        stride, filters = 2, (2 if bottleneck else 8)
    else:
        stride, filters = 1, (1 if bottleneck else 4)
    fout = filters * 4 if bottleneck else filters
    g = torch.Generator().manual_seed(17 + 4 * version + 2 * bottleneck + projection)
    params = {}
    for s in resnet.block_param_specs(version, bottleneck, projection, channels, filters):
        if s.kind == "conv":
            fan_in = s.shape[1] * s.shape[2] * s.shape[3]
            t = torch.randn(s.shape, generator=g) / fan_in ** 0.5
            t = t.bfloat16().float()                    # bf16-exact weights: the comparison is about the kernels
        elif s.kind == "bn_gamma":
            t = 1 + 0.2 * torch.randn(s.shape, generator=g)
        elif s.kind == "bn_beta":
            t = 0.2 * torch.randn(s.shape, generator=g)
        elif s.kind == "bn_var":
            t = torch.ones(s.shape)
        else:
            t = torch.zeros(s.shape)
        params[s.name] = t
    x = (torch.randn(nb, channels, width, width, generator=g) + 0.3).bfloat16().float()
    # oracle (fp32), the function the CPU test ties to the reference's fixtures
    want = resnet.run_block(version, bottleneck, projection, filters, stride,
                            {k: v.clone() for k, v in params.items()}, x, training=True)
    want = want.permute(0, 2, 3, 1).contiguous()
    # native kernels
    xp = torch.zeros(nb, width, width, PAD, device="cuda", dtype=torch.bfloat16)
    xp[..., :channels] = x.permute(0, 2, 3, 1).cuda()
    xf = xp.float().reshape(-1, PAD)
    x_raw = _Raw(xp, torch.cat([xf.sum(0), (xf * xf).sum(0)]).contiguous(), channels)
    cfg = resnet.ResNetConfig(resnet_version=version, bottleneck=bottleneck)
    ops = NativeBlockOps(nat, params, nb)
    y = resnet._block(cfg, ops, x_raw, channels, filters, fout, stride, projection)
    torch.cuda.synchronize()
    assert isinstance(y, _Raw)
    got = y.t.float().cpu()
    assert got.shape[:3] == want.shape[:3]
    assert torch.isfinite(got).all()
    assert got[..., fout:].abs().max().item() == 0.0, "padded channels must stay zero"
    err = (got[..., :fout] - want).abs()
    # bf16 storage of every intermediate tensor against an fp32 oracle: rounding errors of ~0.4 % per tensor, and
    # an occasional ReLU flip of a pre-activation that sits at zero
    tol = 3e-2 + 3e-2 * want.abs()
    frac_bad = (err > tol).float().mean().item()
    assert frac_bad < 0.005 and err.max().item() < 0.3, \
        f"{frac_bad:.4f} of the elements off, max err {err.max().item():.4f} (ref max {want.abs().max().item():.3f})"
    # same fingerprint the fixtures store: first / last element and the sum
    assert abs(got[..., :fout].sum().item() - want.sum().item()) < 2e-2 * want.abs().sum().item() + 1.0
    expected_launches = (3 if bottleneck else 2) + (1 if projection else 0) + (1 if version == 1 else 0)
    assert ops.launches == expected_launches
