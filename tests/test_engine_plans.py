"""CPU-side checks of the native engines' plans: parameter order and shapes equal the oracle's walk of the same
architecture (so `load_named_params` / `export_named` line up), pack arena sizes for both operand types, and the
configuration plumbing of the block-type / dtype options.  No GPU needed: the plans are built without allocating."""
import pytest

from distributedtf_b200.config import RunConfig, parse_args
from distributedtf_b200.models import resnet


def _plan(cls, cfg, E=8):
    eng = object.__new__(cls)
    eng.cfg, eng.E = cfg, E
    eng._finish_segments = lambda: None
    eng._check_config(cfg)
    eng._build_plan()
    return eng


def _engines():
    from distributedtf_b200.models.resnet_native import NativeResNetEngine
    from distributedtf_b200.models.resnet_native_bottleneck import (NativeResNetBottleneckEngine,
                                                                    NativeResNetBottleneckV1Engine)
    from distributedtf_b200.models.resnet_native_v1 import NativeResNetV1Engine
    return {(2, False): NativeResNetEngine, (1, False): NativeResNetV1Engine,
            (2, True): NativeResNetBottleneckEngine, (1, True): NativeResNetBottleneckV1Engine}


@pytest.mark.parametrize("version", [1, 2])
@pytest.mark.parametrize("bottleneck,size", [(False, 8), (False, 32), (True, 11), (True, 29)])
def test_plan_follows_the_oracle_walk(version, bottleneck, size):
    cfg = resnet.cifar10_config(size, version, bottleneck=bottleneck)
    eng = _plan(_engines()[(version, bottleneck)], cfg)
    specs = {s.name: s.shape for s in resnet.param_specs(cfg)}
    for i, c in enumerate(eng.convs):
        assert specs[f"conv{i}/kernel"] == (c.cout, c.k, c.k, c.cin if i else 3), i
    for i, b in enumerate(eng.bns):
        assert specs[f"bn{i}/gamma"] == (b.c,), i
    assert sum(k.startswith("conv") for k in specs) == len(eng.convs)
    assert sum(k.endswith("/gamma") for k in specs) == len(eng.bns)
    assert specs["dense/kernel"] == (10, eng.final_c) and eng.final_c == (256 if bottleneck else 64)
    # the flat layout covers every trainable element exactly once
    trainable = sum(s.numel for s in resnet.param_specs(cfg) if s.trainable)
    pad = eng.convs[0].cout * eng.convs[0].k ** 2 * (eng.convs[0].cin - 3)           # conv0 input channels 3 -> 8
    assert eng.n_params == trainable + pad
    offs = sorted((s.offset, s.count) for s in eng.segments)
    assert offs[0][0] == 0 and all(a[0] + a[1] == b[0] for a, b in zip(offs, offs[1:]))
    assert offs[-1][0] + offs[-1][1] == eng.n_params


def test_pack_arena_for_bf16_and_fp32_operands():
    cfg = resnet.cifar10_config(8)
    cls = _engines()[(2, False)]
    e8, e4 = _plan(cls, cfg, 8), _plan(cls, cfg, 4)
    assert all(s.pack == 1 for s in e8.segments if s.Cout) and all(s.pack == 3 for s in e4.segments if s.Cout)
    for c8, c4 in zip(e8.convs, e4.convs):
        k = c8.k * c8.k * c8.cin
        # [K / E rounded up to an even chunk count][N][E] elements
        assert c8.fwd_pack % 128 == 0 and c4.fwd_pack % 128 == 0
    # conv0: K = 9 * 8 = 72 -> 9 chunks of 8 (padded to 10) or 18 chunks of 4
    nxt8 = e8.convs[1].fwd_pack - e8.convs[0].fwd_pack
    nxt4 = e4.convs[1].fwd_pack - e4.convs[0].fwd_pack
    assert nxt8 == 10 * 16 * 8 and nxt4 == 18 * 16 * 4
    assert e4.n_pack < e8.n_pack * 1.2 and e4.n_pack >= e8.n_pack * 0.9       # same element count up to padding


def test_bottleneck_engine_limits():
    cls = _engines()[(2, True)]
    with pytest.raises(NotImplementedError):
        _plan(cls, resnet.cifar10_config(8))                                              # not a bottleneck config
    wide = resnet.ResNetConfig(bottleneck=True, num_filters=32, block_sizes=(1, 1, 1), final_size=512)
    with pytest.raises(NotImplementedError):
        _plan(cls, wide)                                                                  # 512 channels
    with pytest.raises(NotImplementedError):
        _plan(_engines()[(2, False)], resnet.cifar10_config(11, bottleneck=True))


def test_config_plumbing_of_block_type_and_dtype():
    with pytest.raises(ValueError):
        resnet.cifar10_config(32, bottleneck=True)                    # 9n + 2
    assert resnet.cifar10_config(29, bottleneck=True).block_sizes == (3, 3, 3)
    with pytest.raises(ValueError):
        RunConfig(target_model="cifar10", resnet_size=32, resnet_bottleneck=True).validate()
    cfg = parse_args(["4", "--model", "cifar10", "--resnet-bottleneck", "--resnet-size", "20", "--dtype", "tf32",
                      "--resnet-version", "1"])
    assert cfg.resnet_bottleneck and cfg.resnet_size == 20 and cfg.dtype == "tf32" and cfg.resnet_version == 1
    from distributedtf_b200 import launch
    kw = launch.model_kwargs(cfg)
    assert kw["resnet_bottleneck"] is True and kw["dtype"] == "tf32" and kw["resnet_version"] == 1
    from distributedtf_b200.models import resnet_run
    ns = resnet_run.define_cifar_flags().parse(["--resnet_size", "29", "--resnet_bottleneck", "--dtype", "tf32"])
    assert ns.resnet_bottleneck is True and ns.resnet_size == "29" and ns.dtype == "tf32"
