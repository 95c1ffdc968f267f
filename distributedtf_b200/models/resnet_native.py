"""The B200 training engine for the CIFAR ResNet (v2, building blocks; v1 and bottleneck blocks: sibling modules).

Everything a step does runs in hand-written sm_100a kernels (``csrc/``): the input
preprocessing, every convolution forward / dgrad / wgrad as a tcgen05 implicit GEMM with
BatchNorm+ReLU applied in the operand gather and BN statistics / residual add in the
epilogue, the classifier head (final BN + ReLU + pool + dense + softmax-CE, fwd+bwd in one
launch), the BN-backward apply, and a single multi-tensor optimizer launch that also
re-emits the bf16 GEMM-ready weights.  The launches of a step (93 for ResNet-32) are captured once per
batch size in a CUDA graph; hyper-parameters, LR schedule and the step counter live in
device memory, so explore / exploit never force a re-capture.

Member state is one flat device arena::

    [ hparams | step | params fp32 | slot0 | slot1 | BN moving stats ]   <- exploit copies this
    [ bf16 packed weights (fwd + dgrad layouts) ]                        <- derived
    [ grads fp32 | BN batch statistics | loss, correct ]                 <- zeroed every step

Parity: the architecture and numerics follow reference ``resnet/resnet_model.py`` (block
wiring :127-212, stage layout :323-359, head :537-552) and ``resnet_run_loop.py:176-354``
(loss, optimizer, BN update ops); verified against the torch oracle in ``models/resnet.py``,
which itself reproduces the reference's golden fixtures.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple

import torch

from ..ops import native, reference
from ..utils.streams import new_stream
from . import resnet
from .native_base import NativeEngineBase, _align

IMG_C = 8          # image channels padded 3 -> 8 (one 16-byte chunk)
MAX_BATCH = 256    # explore keeps batch_size in [65, 256]


@dataclass
class ConvL:
    name: str
    cin: int          # storage Cin (conv0: 8)
    cout: int
    k: int
    stride: int
    hin: int
    hout: int
    w_off: int = 0            # element offset in params
    fwd_pack: int = 0         # element offset in pack arena
    dgrad_pack: int = -1
    seg: int = 0


@dataclass
class BnL:
    name: str
    c: int
    h: int                    # spatial size of the normalised tensor
    p_off: int = 0            # offset of (beta, gamma) in params / grads
    stat_off: int = 0         # offset of (sum, sumsq) in the stats region
    run_off: int = 0          # offset of (mean, var) in the running region


@dataclass
class Block:
    bn1: BnL
    conv1: ConvL
    bn2: BnL
    conv2: ConvL
    proj: Optional[ConvL]
    cin: int
    cout: int
    hin: int
    hout: int
    stride: int


class NativeResNetEngine(NativeEngineBase):
    version = 2
    act_dtype = torch.bfloat16
    E = 8

    def __init__(self, cfg: resnet.ResNetConfig, hparams: Dict[str, Any], device, *, seed: int = 0,
                 dtype: str = "bf16", max_batch: int = MAX_BATCH, use_graphs: bool = True,
                 arena_alloc=None, dp_group=None, all_reduce_alg=None, loss_scale: Optional[float] = None):
        self._check_config(cfg)
        if dtype not in ("bf16", "fp16", "tf32", "fp32"):
            raise NotImplementedError(f"native engine: dtype {dtype!r}")
        super().__init__(device, max_batch=max_batch, use_graphs=use_graphs, arena_alloc=arena_alloc,
                         dp_group=dp_group, seed=seed, all_reduce_alg=all_reduce_alg)
        # dtype "fp16" = the reference's mixed-precision mode (fp32 master weights, static loss scale 128,
        # resnet_model.py:439-474): on B200 the 16-bit operand type is bf16 (same tensor-core rate, fp32's exponent
        # range), the loss scaling is applied exactly as specified.
        # dtype "fp32" / "tf32" = the reference's default dtype (flags/_performance.py:107): activations, gradients
        # and packed weights are stored in fp32 and the convolutions run as kind::tf32 tcgen05 MMAs (10-bit mantissa
        # products, fp32 accumulation) on the generic implicit-GEMM kernels (csrc/conv_sm100.cu, E = 4).
        self.compute_dtype = "tf32" if dtype in ("tf32", "fp32") else "bf16"
        self.act_dtype = torch.float32 if self.compute_dtype == "tf32" else torch.bfloat16
        self.E = 4 if self.compute_dtype == "tf32" else 8          # elements per 16-byte operand chunk
        self.loss_scale = float(loss_scale) if loss_scale else (128.0 if dtype == "fp16" else 1.0)
        self.cfg = cfg
        self.image_size, self.image_channels = cfg.image_size, cfg.in_channels
        self._build_plan()
        self._alloc()
        self._init_params(hparams.get("initializer", "None"), seed)
        self.set_hparams(hparams, reset_slots=True)

    def _check_config(self, cfg) -> None:
        if cfg.resnet_version != self.version or cfg.bottleneck or cfg.first_pool_size:
            raise NotImplementedError("building-block engine: CIFAR configuration (no bottleneck, no initial "
                                      "max-pool); bottleneck blocks: resnet_native_bottleneck.py")

    def lr_schedule(self, hp):
        return reference.cifar_lr_schedule(hp)

    # ------------------------------------------------------------------ plan
    def _build_plan(self) -> None:
        cfg = self.cfg
        self.convs: List[ConvL] = []
        self.bns: List[BnL] = []
        self.blocks: List[Block] = []
        size = cfg.image_size

        def conv(cin, cout, k, s, hin):
            c = ConvL(f"conv{len(self.convs)}", cin, cout, k, s, hin, hin // s)
            self.convs.append(c)
            return c

        def bn(c, h):
            b = BnL(f"bn{len(self.bns)}", c, h)
            self.bns.append(b)
            return b

        self.conv0 = conv(IMG_C, cfg.num_filters, cfg.kernel_size, cfg.conv_stride, size)
        h = size // cfg.conv_stride
        cin = cfg.num_filters
        for i, nblocks in enumerate(cfg.block_sizes):
            filters = cfg.num_filters * (2 ** i)
            for b in range(nblocks):
                stride = cfg.block_strides[i] if b == 0 else 1
                bn1 = bn(cin, h)
                proj = conv(cin, filters, 1, stride, h) if b == 0 else None
                c1 = conv(cin, filters, 3, stride, h)
                bn2 = bn(filters, h // stride)
                c2 = conv(filters, filters, 3, 1, h // stride)
                self.blocks.append(Block(bn1, c1, bn2, c2, proj, cin, filters, h, h // stride, stride))
                cin = filters
                h //= stride
        self.bn_final = bn(cin, h)
        self.final_c, self.final_h = cin, h
        # ---- flat parameter layout: conv kernels | (beta, gamma) per BN | dense W | dense b
        off = 0
        pack = 0
        self.segments: List[native.OptSegment] = []
        for c in self.convs:
            c.w_off = off
            n = c.cout * c.k * c.k * c.cin
            E = self.E
            kch = c.k * c.k * c.cin // E
            c.fwd_pack = pack
            pack += _align(((kch + 1) & ~1) * c.cout * E, 128)
            if c is not self.conv0:
                kchd = c.k * c.k * c.cout // E
                c.dgrad_pack = pack
                pack += _align(((kchd + 1) & ~1) * c.cin * E, 128)
            c.seg = len(self.segments)
            self.segments.append(native.OptSegment(off, n, 1, 1 if E == 8 else 3, c.cout, c.k * c.k, c.cin,
                                                   c.fwd_pack, c.dgrad_pack))
            off += n
        self.n_conv_params = off
        stat = run = 0
        for b in self.bns:
            b.p_off, b.stat_off, b.run_off = off, stat, run
            self.segments.append(native.OptSegment(off, 2 * b.c, 0, 0, 0, 0, 0, -1, -1))
            off += 2 * b.c
            stat += 2 * b.c
            run += 2 * b.c
        self.dense_w_off = off
        self.segments.append(native.OptSegment(off, cfg.num_classes * self.final_c, 0, 0, 0, 0, 0, -1, -1))
        off += cfg.num_classes * self.final_c
        self.dense_b_off = off
        self.segments.append(native.OptSegment(off, cfg.num_classes, 0, 0, 0, 0, 0, -1, -1))
        off += cfg.num_classes
        self.n_params = off
        self.n_stats, self.n_running, self.n_pack = stat, run, pack
        self._finish_segments()

    # ------------------------------------------------------------------ memory
    def _alloc(self) -> None:
        self._alloc_arena()
        dev = self.device
        # static IO + activations, sized for the largest batch so pointers never move
        mb = self.max_batch
        cfg = self.cfg
        S = cfg.image_size
        self.image = torch.zeros(mb, S, S, IMG_C, dtype=self.act_dtype, device=dev)

        def act(h, c):
            return torch.empty(mb, h, h, c, dtype=self.act_dtype, device=dev)

        self.r0 = act(self.conv0.hout, self.conv0.cout)
        self.acts = []
        for b in self.blocks:
            self.acts.append({"c1": act(b.hout, b.cout), "out": act(b.hout, b.cout),
                              "sc": act(b.hout, b.cout) if b.proj else None})
        big = max(b.hin * b.hin * b.cin for b in self.blocks)
        self.gbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]
        self.tbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]
        self.logits = torch.zeros(mb, cfg.num_classes, dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ params
    def conv_weight(self, c: ConvL) -> torch.Tensor:
        return self.params[c.w_off: c.w_off + c.cout * c.k * c.k * c.cin].view(c.cout, c.k, c.k, c.cin)

    def bn_view(self, b: BnL, base: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        return base[b.p_off: b.p_off + b.c], base[b.p_off + b.c: b.p_off + 2 * b.c]      # (beta, gamma)

    def _init_params(self, initializer: str, seed: int) -> None:
        """Initial weights are drawn on the device (``dtf_init_params``: Philox keyed by the member's seed) for the
        uniform / truncated-normal initialisers; ``orthogonal`` needs a QR factorisation and comes from the host
        restatement (``models/resnet.py init_params``), as does everything when ``DTF_HOST_INIT=1``."""
        import os
        if initializer == "orthogonal" or os.environ.get("DTF_HOST_INIT") == "1":
            self.load_named_params(resnet.init_params(self.cfg, initializer, seed))
            return
        segs = []
        for c in self.convs:
            cin_real = self.cfg.in_channels if c is self.conv0 else c.cin
            fan_in, fan_out = c.k * c.k * cin_real, c.k * c.k * c.cout
            n = c.cout * c.k * c.k * c.cin
            if initializer == "glorot_normal":
                kind, scale = native.INIT_TRUNC_NORMAL, math.sqrt(2.0 / (fan_in + fan_out))
            elif initializer == "he_init":
                kind, scale = native.INIT_TRUNC_NORMAL, math.sqrt(2.0 / fan_in)
            else:                                   # layer default: glorot_uniform
                kind, scale = native.INIT_UNIFORM, math.sqrt(6.0 / (fan_in + fan_out))
            pad = (c.cin, cin_real) if cin_real != c.cin else (0, 0)
            segs.append(native.InitSegment(c.w_off, n, kind, scale, pad[0], pad[1]))
        for b in self.bns:
            segs.append(native.InitSegment(b.p_off, b.c, native.INIT_CONST, 0.0, 0, 0))            # beta
            segs.append(native.InitSegment(b.p_off + b.c, b.c, native.INIT_CONST, 1.0, 0, 0))      # gamma
        nd = self.cfg.num_classes * self.final_c                                                   # dense: glorot_uniform
        segs.append(native.InitSegment(self.dense_w_off, nd, native.INIT_UNIFORM,
                                       math.sqrt(6.0 / (self.final_c + self.cfg.num_classes)), 0, 0))
        segs.append(native.InitSegment(self.dense_b_off, self.cfg.num_classes, native.INIT_CONST, 0.0, 0, 0))
        with torch.cuda.device(self.device):
            native.init_params(self.params, segs, 0x9E3779B97F4A7C15 * (seed + 1))
            for b in self.bns:
                self.running[b.run_off: b.run_off + b.c].zero_()
                self.running[b.run_off + b.c: b.run_off + 2 * b.c].fill_(1.0)
            self.repack()

    @torch.no_grad()
    def load_named_params(self, named: Dict[str, torch.Tensor]) -> None:
        """Adopt parameters in the oracle's naming/layout (``models/resnet.py``)."""
        dev = self.device
        for i, c in enumerate(self.convs):
            w = named[f"conv{i}/kernel"].to(dev, torch.float32)
            if c is self.conv0:
                full = torch.zeros(c.cout, c.k, c.k, c.cin, device=dev)
                full[..., : w.shape[-1]] = w
                w = full
            self.conv_weight(c).copy_(w)
        for i, b in enumerate(self.bns):
            beta, gamma = self.bn_view(b, self.params)
            beta.copy_(named[f"bn{i}/beta"].to(dev))
            gamma.copy_(named[f"bn{i}/gamma"].to(dev))
            self.running[b.run_off: b.run_off + b.c].copy_(named[f"bn{i}/moving_mean"].to(dev))
            self.running[b.run_off + b.c: b.run_off + 2 * b.c].copy_(named[f"bn{i}/moving_variance"].to(dev))
        n = self.cfg.num_classes * self.final_c
        self.params[self.dense_w_off: self.dense_w_off + n].copy_(named["dense/kernel"].to(dev).reshape(-1))
        self.params[self.dense_b_off: self.dense_b_off + self.cfg.num_classes].copy_(named["dense/bias"].to(dev))
        self.repack()

    def export_named(self, base: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """Parameters (or, with ``base=self.grads``, gradients) in the oracle's naming."""
        base = self.params if base is None else base
        out = {}
        for i, c in enumerate(self.convs):
            w = base[c.w_off: c.w_off + c.cout * c.k * c.k * c.cin].view(c.cout, c.k, c.k, c.cin)
            out[f"conv{i}/kernel"] = (w[..., : self.cfg.in_channels] if c is self.conv0 else w).clone()
        for i, b in enumerate(self.bns):
            beta, gamma = self.bn_view(b, base)
            out[f"bn{i}/beta"], out[f"bn{i}/gamma"] = beta.clone(), gamma.clone()
            if base is self.params:
                out[f"bn{i}/moving_mean"] = self.running[b.run_off: b.run_off + b.c].clone()
                out[f"bn{i}/moving_variance"] = self.running[b.run_off + b.c: b.run_off + 2 * b.c].clone()
        n = self.cfg.num_classes * self.final_c
        out["dense/kernel"] = base[self.dense_w_off: self.dense_w_off + n].view(self.cfg.num_classes, -1).clone()
        out["dense/bias"] = base[self.dense_b_off: self.dense_b_off + self.cfg.num_classes].clone()
        return out

    # ------------------------------------------------------------------ kernels of one step
    def _bn(self, b: BnL, nb: int, training: bool) -> native.BnRef:
        beta, gamma = self.bn_view(b, self.params)
        return native.bn_ref(b.c, nb * b.h * b.h, stats=self.stats[b.stat_off: b.stat_off + 2 * b.c],
                             gamma=gamma, beta=beta, run_mean=self.running[b.run_off: b.run_off + b.c],
                             run_var=self.running[b.run_off + b.c: b.run_off + 2 * b.c],
                             use_running=not training)

    def _stats_out(self, b: Optional[BnL]):
        return None if b is None else self.stats[b.stat_off: b.stat_off + 2 * b.c]

    def _bwd_stats(self, b: BnL) -> torch.Tensor:
        return self.grads[b.p_off: b.p_off + 2 * b.c]

    def _wpk(self, c: ConvL, dgrad: bool = False) -> torch.Tensor:
        return self.pack[(c.dgrad_pack if dgrad else c.fwd_pack):]

    def _fwd_conv(self, c: ConvL, src, out, nb, *, bn: Optional[BnL], training, update, res=None,
                  stats_for: Optional[BnL] = None) -> None:
        g = native.gather(src, c.hin, c.hin, c.cin, KH=c.k, KW=c.k, stride=c.stride, pad=(c.k - 1) // 2,
                          mode=native.PRO_BN_RELU if bn else native.PRO_NONE,
                          bn=self._bn(bn, nb, training) if bn else None, update_running=update and training)
        native.conv_gemm(g, self._wpk(c), out, NB=nb, Ho=c.hout, Wo=c.hout, N=c.cout,
                         epi_mode=native.EPI_RESIDUAL if res is not None else native.EPI_STORE, epi_res=res,
                         epi_stats_out=self._stats_out(stats_for) if training else None)
        self._count += 1

    # -- independent kernels (projection shortcut, strided-conv weight gradient) run on a side stream: every
    #    kernel here is a latency chain (launch -> BN constants -> first tile -> tail), so a second chain in
    #    flight hides a good part of it.  Inside CUDA-graph capture the fork/join become graph branches.
    def _fork(self):
        if self._side is None:
            self._side = new_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._side.wait_event(ev)
        return torch.cuda.stream(self._side)

    def _join(self) -> None:
        ev = torch.cuda.Event()
        ev.record(self._side)
        torch.cuda.current_stream().wait_event(ev)

    def _forward(self, nb: int, training: bool) -> torch.Tensor:
        """Runs the trunk; returns the tensor feeding the final BN."""
        self._fwd_conv(self.conv0, self.image, self.r0, nb, bn=None, training=training, update=False,
                       stats_for=self.blocks[0].bn1)
        x = self.r0
        for i, (b, a) in enumerate(zip(self.blocks, self.acts)):
            nxt = self.blocks[i + 1].bn1 if i + 1 < len(self.blocks) else self.bn_final
            if b.proj:
                with self._fork():
                    self._fwd_conv(b.proj, x, a["sc"], nb, bn=b.bn1, training=training, update=False)
            self._fwd_conv(b.conv1, x, a["c1"], nb, bn=b.bn1, training=training, update=True, stats_for=b.bn2)
            if b.proj:
                self._join()
            self._fwd_conv(b.conv2, a["c1"], a["out"], nb, bn=b.bn2, training=training, update=True,
                           res=a["sc"] if b.proj else x, stats_for=nxt)
            x = a["out"]
        return x

    def _head(self, x, nb: int, training: bool, t_out=None, loss=None, correct=None) -> None:
        b = self.bn_final
        n = self.cfg.num_classes * self.final_c
        h = native.Head()
        h.x, h.labels = x.data_ptr(), self.labels.data_ptr()
        h.w = self.params[self.dense_w_off:].data_ptr()
        h.bias = self.params[self.dense_b_off:].data_ptr()
        h.bn = self._bn(b, nb, training)
        h.loss_out = (loss if loss is not None else self.loss_cell).data_ptr()
        h.correct_out = (correct if correct is not None else self.correct_cell).data_ptr()
        h.logits_out = self.logits.data_ptr()
        h.dw = self.grads[self.dense_w_off:].data_ptr()
        h.db = self.grads[self.dense_b_off:].data_ptr()
        h.t_out = t_out.data_ptr() if t_out is not None else 0
        h.bwd_stats_out = self._bwd_stats(b).data_ptr()
        h.loss_scale = float(self.loss_scale)
        h.batch, h.HW, h.C, h.classes = nb, b.h * b.h, b.c, self.cfg.num_classes
        h.has_bn, h.training, h.update_running = 1, int(training), int(training)
        h.board_hits = 0 if training else self.board_slot_ptr     # evaluation hits go straight to the score board
        h.dtype = native.dtype_code(x)
        native.head(h)
        self._count += 1

    def _wgrad(self, c: ConvL, x_src, x_bn: Optional[BnL], gy_src, nb, *, gy_bn: Optional[BnL] = None,
               gy_x=None) -> None:
        gx = native.gather(x_src, c.hin, c.hin, c.cin, KH=c.k, KW=c.k, stride=c.stride, pad=(c.k - 1) // 2,
                           mode=native.PRO_BN_RELU if x_bn else native.PRO_NONE,
                           bn=self._bn(x_bn, nb, True) if x_bn else None)
        gg = native.gather(gy_src, c.hout, c.hout, c.cout,
                           mode=native.PRO_BN_BWD if gy_bn else native.PRO_NONE,
                           bn=self._bn(gy_bn, nb, True) if gy_bn else None, src2=gy_x,
                           bwd_stats=self._bwd_stats(gy_bn) if gy_bn else None)
        dw = self.grads[c.w_off:]
        native.conv_wgrad(gx, gg, dw, NB=nb, Ho=c.hout, Wo=c.hout, Cout=c.cout)
        self._count += 1

    def _dgrad(self, c: ConvL, gy_src, nb, out, *, mask_x, mask_bn: BnL, gy_bn: Optional[BnL] = None, gy_x=None,
               accumulate: bool = False) -> None:
        g = native.gather(gy_src, c.hout, c.hout, c.cout, KH=c.k, KW=c.k, stride=c.stride, pad=(c.k - 1) // 2,
                          transposed=True, mode=native.PRO_BN_BWD if gy_bn else native.PRO_NONE,
                          bn=self._bn(gy_bn, nb, True) if gy_bn else None, src2=gy_x,
                          bwd_stats=self._bwd_stats(gy_bn) if gy_bn else None)
        native.conv_gemm(g, self._wpk(c, dgrad=True), out, NB=nb, Ho=c.hin, Wo=c.hin, N=c.cin,
                         epi_mode=native.EPI_RELU_MASK, epi_x=mask_x, epi_bn=self._bn(mask_bn, nb, True),
                         epi_stats_out=self._bwd_stats(mask_bn), epi_accumulate=accumulate)
        self._count += 1

    def _bwd_conv(self, c: ConvL, x_src, x_bn: Optional[BnL], gy_src, nb, out, *, mask_x, mask_bn: BnL,
                  gy_bn: Optional[BnL] = None, gy_x=None) -> None:
        """Weight gradient + input gradient of one conv.  3x3 / stride-1 layers go out as ONE launch
        (dgrad CTAs and wgrad CTAs share the grid); the others as two."""
        if not (c.k == 3 and c.stride == 1 and c.cin == c.cout):
            with self._fork():
                self._wgrad(c, x_src, x_bn, gy_src, nb, gy_bn=gy_bn, gy_x=gy_x)
            self._dgrad(c, gy_src, nb, out, mask_x=mask_x, mask_bn=mask_bn, gy_bn=gy_bn, gy_x=gy_x)
            self._join()
            return
        pro = native.PRO_BN_BWD if gy_bn else native.PRO_NONE
        bn_gy = self._bn(gy_bn, nb, True) if gy_bn else None
        bwd = self._bwd_stats(gy_bn) if gy_bn else None
        gd = native.gather(gy_src, c.hout, c.hout, c.cout, KH=3, KW=3, stride=1, pad=1, transposed=True, mode=pro,
                           bn=bn_gy, src2=gy_x, bwd_stats=bwd)
        dgrad = native.conv_gemm_params(gd, self._wpk(c, dgrad=True), out, NB=nb, Ho=c.hin, Wo=c.hin, N=c.cin,
                                        epi_mode=native.EPI_RELU_MASK, epi_x=mask_x,
                                        epi_bn=self._bn(mask_bn, nb, True), epi_stats_out=self._bwd_stats(mask_bn))
        gx = native.gather(x_src, c.hin, c.hin, c.cin, KH=3, KW=3, stride=1, pad=1,
                           mode=native.PRO_BN_RELU if x_bn else native.PRO_NONE,
                           bn=self._bn(x_bn, nb, True) if x_bn else None)
        gg = native.gather(gy_src, c.hout, c.hout, c.cout, mode=pro, bn=bn_gy, src2=gy_x, bwd_stats=bwd)
        native.conv_bwd(dgrad, gx, gg, self.grads[c.w_off:], NB=nb, Ho=c.hout, Wo=c.hout, Cout=c.cout)
        self._count += 1

    def _backward(self, x_last, nb: int) -> None:
        t_f = self.tbuf[0]
        self._head(x_last, nb, True, t_out=t_f)
        bf = self.bn_final
        g_out = self.gbuf[0]
        native.bn_bwd_apply(t_f, x_last, None, g_out, self._bn(bf, nb, True), self._bwd_stats(bf),
                            nb * bf.h * bf.h, bf.c)
        self._count += 1
        cur = 0
        for i in range(len(self.blocks) - 1, -1, -1):
            b, a = self.blocks[i], self.acts[i]
            x_in = self.acts[i - 1]["out"] if i > 0 else self.r0
            t2, t1 = self.tbuf[0], self.tbuf[1]
            if b.proj:
                with self._fork():
                    self._wgrad(b.proj, x_in, b.bn1, g_out, nb)
            self._bwd_conv(b.conv2, a["c1"], b.bn2, g_out, nb, t2, mask_x=a["c1"], mask_bn=b.bn2)
            if b.proj:
                self._join()
            self._bwd_conv(b.conv1, x_in, b.bn1, t2, nb, t1, mask_x=x_in, mask_bn=b.bn1, gy_bn=b.bn2, gy_x=a["c1"])
            if b.proj:
                self._dgrad(b.proj, g_out, nb, t1, mask_x=x_in, mask_bn=b.bn1, accumulate=True)
            nxt = self.gbuf[cur ^ 1]
            native.bn_bwd_apply(t1, x_in, None if b.proj else g_out, nxt, self._bn(b.bn1, nb, True),
                                self._bwd_stats(b.bn1), nb * b.hin * b.hin, b.cin)
            self._count += 1
            g_out = nxt
            cur ^= 1
        self._wgrad(self.conv0, self.image, None, g_out, nb)

    def _train_kernels(self, nb: int, pp_mode: int) -> None:
        self._count = 0
        self._begin_step()
        self.preprocess_batch(nb, pp_mode)
        self._count += 1
        x_last = self._forward(nb, True)
        self._backward(x_last, nb)
        self._apply_update()

    def _eval_kernels(self, nb: int) -> None:
        self._count = 0
        self.preprocess_batch(nb, 1)
        x_last = self._forward(nb, False)
        self._head(x_last, nb, False, loss=self.eval_cells[0:1], correct=self.eval_cells[1:2].view(torch.int32))

