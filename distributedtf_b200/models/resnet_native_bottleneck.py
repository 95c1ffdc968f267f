"""B200 training engines for ResNets built from **bottleneck** blocks (v2 pre-activation and v1).

Parity: reference ``resnet/resnet_model.py:215-264`` (``_bottleneck_block_v1``: 1x1 -> BN -> ReLU -> 3x3 (strided)
-> BN -> ReLU -> 1x1 (4 x filters) -> BN, add the shortcut, ReLU) and ``:267-320`` (``_bottleneck_block_v2``:
BN -> ReLU in front of each of the three convolutions, projection taken after the first BN + ReLU), selected by
``Model.__init__`` (``:362-437``) when ``bottleneck=True``; the final dense layer then sees ``4 x`` the channels.

The engines reuse every kernel of the building-block engines (``resnet_native.py``, ``resnet_native_v1.py``): a
block is a chain of three implicit-GEMM convolutions whose BatchNorm + ReLU ride in the next operand gather and
whose BN statistics are produced by the previous epilogue.  What the wider blocks needed from ``csrc/``:

* 1x1 convolutions with 128 / 256 input channels gather their rows in batches of eight 16-byte loads
  (``gather_chunks_wide`` in ``conv_common.cuh``) instead of one chunk at a time;
* the weight-gradient kernel splits more than 128 output channels over ``blockIdx.z`` (``conv_wgrad_kernel``), since
  one ``tcgen05.mma`` covers 128 of them (M = 128);
* the 3x3 convolution in the middle has as many outputs as inputs (16 / 32 / 64), so with stride 1 it runs on the
  halo-tile kernels and its backward is the fused dgrad + wgrad launch; with stride 2 the input gradient runs on
  the parity-class kernel.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops import native
from .resnet_native import IMG_C, BnL, ConvL, NativeResNetEngine
from .resnet_native_v1 import NativeResNetV1Engine


@dataclass
class BlockB:
    """One bottleneck block.  v2: ``bns[j]`` is the BatchNorm in FRONT of ``convs[j]``; v1: the one BEHIND it."""
    bns: List[BnL]
    convs: List[ConvL]
    proj: Optional[ConvL]
    bn_p: Optional[BnL]          # v1 only: BatchNorm of the projection shortcut
    cin: int
    mid: int                     # filters of the 1x1 reduce and the 3x3
    cout: int                    # 4 x mid
    hin: int
    hout: int
    stride: int


def _check_cfg(cfg) -> None:
    if not cfg.bottleneck or cfg.first_pool_size:
        raise NotImplementedError("bottleneck engine: bottleneck blocks, no initial max-pool")
    widest = cfg.num_filters * 2 ** (len(cfg.block_sizes) - 1) * 4
    if widest > 256 or cfg.num_filters % 16 != 0:
        raise NotImplementedError("bottleneck engine: up to 256 channels (num_filters 16, three stages), filters a "
                                  "multiple of 16")


class NativeResNetBottleneckEngine(NativeResNetEngine):
    """ResNet v2 with bottleneck blocks."""
    version = 2
    bottleneck = True

    def _check_config(self, cfg) -> None:
        if cfg.resnet_version != 2:
            raise NotImplementedError("v2 bottleneck engine")
        _check_cfg(cfg)

    # ------------------------------------------------------------------ plan (parameter order = models/resnet.py)
    def _build_plan(self) -> None:
        cfg = self.cfg
        self.convs: List[ConvL] = []
        self.bns: List[BnL] = []
        self.blocks: List[BlockB] = []
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
            f = cfg.num_filters * (2 ** i)
            for b in range(nblocks):
                s = cfg.block_strides[i] if b == 0 else 1
                bn1 = bn(cin, h)
                proj = conv(cin, 4 * f, 1, s, h) if b == 0 else None
                c1 = conv(cin, f, 1, 1, h)
                bn2 = bn(f, h)
                c2 = conv(f, f, 3, s, h)
                bn3 = bn(f, h // s)
                c3 = conv(f, 4 * f, 1, 1, h // s)
                self.blocks.append(BlockB([bn1, bn2, bn3], [c1, c2, c3], proj, None, cin, f, 4 * f, h, h // s, s))
                cin = 4 * f
                h //= s
        self.bn_final = bn(cin, h)
        self.final_c, self.final_h = cin, h
        NativeResNetV1Engine._layout_params(self)          # the flat layout is the same for every engine

    # ------------------------------------------------------------------ memory
    def _alloc(self) -> None:
        self._alloc_arena()
        dev, mb, cfg = self.device, self.max_batch, self.cfg
        S = cfg.image_size
        self.image = torch.zeros(mb, S, S, IMG_C, dtype=self.act_dtype, device=dev)

        def act(h, c):
            return torch.empty(mb, h, h, c, dtype=self.act_dtype, device=dev)

        self.r0 = act(self.conv0.hout, self.conv0.cout)
        self.acts = [{"m1": act(b.hin, b.mid), "m2": act(b.hout, b.mid), "out": act(b.hout, b.cout),
                      "sc": act(b.hout, b.cout) if b.proj else None} for b in self.blocks]
        big = max(max(b.hin * b.hin * max(b.cin, b.mid), b.hout * b.hout * b.cout) for b in self.blocks)
        self.gbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]
        self.tbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]
        self.logits = torch.zeros(mb, cfg.num_classes, dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ forward / backward
    def _forward(self, nb: int, training: bool) -> torch.Tensor:
        self._fwd_conv(self.conv0, self.image, self.r0, nb, bn=None, training=training, update=False,
                       stats_for=self.blocks[0].bns[0])
        x = self.r0
        for i, (b, a) in enumerate(zip(self.blocks, self.acts)):
            nxt = self.blocks[i + 1].bns[0] if i + 1 < len(self.blocks) else self.bn_final
            bn1, bn2, bn3 = b.bns
            c1, c2, c3 = b.convs
            if b.proj:
                with self._fork():
                    self._fwd_conv(b.proj, x, a["sc"], nb, bn=bn1, training=training, update=False)
            self._fwd_conv(c1, x, a["m1"], nb, bn=bn1, training=training, update=True, stats_for=bn2)
            self._fwd_conv(c2, a["m1"], a["m2"], nb, bn=bn2, training=training, update=True, stats_for=bn3)
            if b.proj:
                self._join()
            self._fwd_conv(c3, a["m2"], a["out"], nb, bn=bn3, training=training, update=True,
                           res=a["sc"] if b.proj else x, stats_for=nxt)
            x = a["out"]
        return x

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
            bn1, bn2, bn3 = b.bns
            c1, c2, c3 = b.convs
            x_in = self.acts[i - 1]["out"] if i > 0 else self.r0
            t3, t2 = self.tbuf[0], self.tbuf[1]
            t1 = self.tbuf[0]                      # t3 is dead once the 3x3's backward has been issued
            if b.proj:
                with self._fork():
                    self._wgrad(b.proj, x_in, bn1, g_out, nb)
            self._bwd_conv(c3, a["m2"], bn3, g_out, nb, t3, mask_x=a["m2"], mask_bn=bn3)
            if b.proj:
                self._join()
            self._bwd_conv(c2, a["m1"], bn2, t3, nb, t2, mask_x=a["m1"], mask_bn=bn2, gy_bn=bn3, gy_x=a["m2"])
            self._bwd_conv(c1, x_in, bn1, t2, nb, t1, mask_x=x_in, mask_bn=bn1, gy_bn=bn2, gy_x=a["m1"])
            if b.proj:
                self._dgrad(b.proj, g_out, nb, t1, mask_x=x_in, mask_bn=bn1, accumulate=True)
            nxt = self.gbuf[cur ^ 1]
            native.bn_bwd_apply(t1, x_in, None if b.proj else g_out, nxt, self._bn(bn1, nb, True),
                                self._bwd_stats(bn1), nb * b.hin * b.hin, b.cin)
            self._count += 1
            g_out = nxt
            cur ^= 1
        self._wgrad(self.conv0, self.image, None, g_out, nb)


class NativeResNetBottleneckV1Engine(NativeResNetV1Engine):
    """ResNet v1 with bottleneck blocks (post-activation; the block output is written by the tail kernel)."""
    version = 1
    bottleneck = True

    def _check_config(self, cfg) -> None:
        if cfg.resnet_version != 1:
            raise NotImplementedError("v1 bottleneck engine")
        _check_cfg(cfg)

    def _build_plan(self) -> None:
        cfg = self.cfg
        self.convs: List[ConvL] = []
        self.bns: List[BnL] = []
        self.blocks: List[BlockB] = []
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
        self.bn0 = bn(cfg.num_filters, h)
        cin = cfg.num_filters
        for i, nblocks in enumerate(cfg.block_sizes):
            f = cfg.num_filters * (2 ** i)
            for b in range(nblocks):
                s = cfg.block_strides[i] if b == 0 else 1
                proj = conv(cin, 4 * f, 1, s, h) if b == 0 else None
                bn_p = bn(4 * f, h // s) if b == 0 else None
                c1 = conv(cin, f, 1, 1, h)
                bn1 = bn(f, h)
                c2 = conv(f, f, 3, s, h)
                bn2 = bn(f, h // s)
                c3 = conv(f, 4 * f, 1, 1, h // s)
                bn3 = bn(4 * f, h // s)
                self.blocks.append(BlockB([bn1, bn2, bn3], [c1, c2, c3], proj, bn_p, cin, f, 4 * f, h, h // s, s))
                cin = 4 * f
                h //= s
        self.bn_final = None
        self.final_c, self.final_h = cin, h
        self._layout_params()

    def _alloc(self) -> None:
        self._alloc_arena()
        dev, mb, cfg = self.device, self.max_batch, self.cfg
        S = cfg.image_size
        self.image = torch.zeros(mb, S, S, IMG_C, dtype=self.act_dtype, device=dev)

        def act(h, c):
            return torch.empty(mb, h, h, c, dtype=self.act_dtype, device=dev)

        self.c0 = act(self.conv0.hout, self.conv0.cout)
        self.y0 = act(self.conv0.hout, self.conv0.cout)
        self.acts = [{"c1": act(b.hin, b.mid), "c2": act(b.hout, b.mid), "c3": act(b.hout, b.cout),
                      "y": act(b.hout, b.cout), "p": act(b.hout, b.cout) if b.proj else None} for b in self.blocks]
        big = max(max(b.hin * b.hin * max(b.cin, b.mid), b.hout * b.hout * b.cout) for b in self.blocks)
        self.gbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]
        self.tbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(3)]
        self.logits = torch.zeros(mb, cfg.num_classes, dtype=torch.float32, device=dev)

    def _forward(self, nb: int, training: bool) -> torch.Tensor:
        self._conv_raw(self.conv0, self.image, self.c0, nb, bn=None, training=training, stats_for=self.bn0)
        native.bn_act_apply(self.c0, self._bn(self.bn0, nb, training), self.y0, nb * self.bn0.h ** 2, self.bn0.c,
                            update_running=training)
        self._count += 1
        x = self.y0
        for b, a in zip(self.blocks, self.acts):
            bn1, bn2, bn3 = b.bns
            c1, c2, c3 = b.convs
            if b.proj:
                with self._fork():
                    self._conv_raw(b.proj, x, a["p"], nb, bn=None, training=training, stats_for=b.bn_p)
            self._conv_raw(c1, x, a["c1"], nb, bn=None, training=training, stats_for=bn1)
            self._conv_raw(c2, a["c1"], a["c2"], nb, bn=bn1, training=training, stats_for=bn2)
            self._conv_raw(c3, a["c2"], a["c3"], nb, bn=bn2, training=training, stats_for=bn3)
            if b.proj:
                self._join()
            native.bn_act_apply(a["c3"], self._bn(bn3, nb, training), a["y"], nb * b.hout ** 2, b.cout,
                                sc=a["p"] if b.proj else x,
                                bn_sc=self._bn(b.bn_p, nb, training) if b.proj else None, update_running=training)
            self._count += 1
            x = a["y"]
        return x

    def _backward(self, x_last, nb: int) -> None:
        dy = self.gbuf[0]
        self._head(x_last, nb, True, t_out=dy)
        cur = 0
        for i in range(len(self.blocks) - 1, -1, -1):
            b, a = self.blocks[i], self.acts[i]
            bn1, bn2, bn3 = b.bns
            c1, c2, c3 = b.convs
            y_in = self.acts[i - 1]["y"] if i > 0 else self.y0
            t, ta, tb = self.tbuf
            pix = nb * b.hout ** 2
            native.relu_bwd_stats(dy, a["y"], t, a["c3"], self._bn(bn3, nb, True), self._bwd_stats(bn3), pix, b.cout,
                                  x2=a["p"] if b.proj else None, bn2=self._bn(b.bn_p, nb, True) if b.proj else None,
                                  stats2=self._bwd_stats(b.bn_p) if b.proj else None)
            self._count += 1
            nxt = self.gbuf[cur ^ 1]
            if b.proj:
                with self._fork():
                    gx = native.gather(y_in, b.hin, b.hin, b.cin, KH=1, KW=1, stride=b.stride, pad=0)
                    gg = native.gather(t, b.hout, b.hout, b.cout, mode=native.PRO_BN_BWD, bn=self._bn(b.bn_p, nb, True),
                                       src2=a["p"], bwd_stats=self._bwd_stats(b.bn_p))
                    native.conv_wgrad(gx, gg, self.grads[b.proj.w_off:], NB=nb, Ho=b.hout, Wo=b.hout, Cout=b.cout)
                    self._count += 1
            self._bwd_v1(c3, a["c2"], bn2, t, a["c3"], bn3, nb, ta, mask_x=a["c2"], mask_bn=bn2)
            if b.proj:
                self._join()
            self._bwd_v1(c2, a["c1"], bn1, ta, a["c2"], bn2, nb, tb, mask_x=a["c1"], mask_bn=bn1)
            if b.proj:
                self._bwd_v1(c1, y_in, None, tb, a["c1"], bn1, nb, nxt)
                gd = native.gather(t, b.hout, b.hout, b.cout, KH=1, KW=1, stride=b.stride, pad=0, transposed=True,
                                   mode=native.PRO_BN_BWD, bn=self._bn(b.bn_p, nb, True), src2=a["p"],
                                   bwd_stats=self._bwd_stats(b.bn_p))
                native.conv_gemm(gd, self._wpk(b.proj, dgrad=True), nxt, NB=nb, Ho=b.hin, Wo=b.hin, N=b.cin,
                                 epi_mode=native.EPI_RESIDUAL, epi_res=nxt)
                self._count += 1
            else:
                self._bwd_v1(c1, y_in, None, tb, a["c1"], bn1, nb, nxt, res=t)
            dy = nxt
            cur ^= 1
        t0 = self.tbuf[0]
        native.relu_bwd_stats(dy, self.y0, t0, self.c0, self._bn(self.bn0, nb, True), self._bwd_stats(self.bn0),
                              nb * self.bn0.h ** 2, self.bn0.c)
        c0 = self.conv0
        gx = native.gather(self.image, c0.hin, c0.hin, c0.cin, KH=c0.k, KW=c0.k, stride=c0.stride, pad=(c0.k - 1) // 2)
        gg = native.gather(t0, c0.hout, c0.hout, c0.cout, mode=native.PRO_BN_BWD, bn=self._bn(self.bn0, nb, True),
                           src2=self.c0, bwd_stats=self._bwd_stats(self.bn0))
        native.conv_wgrad(gx, gg, self.grads[c0.w_off:], NB=nb, Ho=c0.hout, Wo=c0.hout, Cout=c0.cout)
        self._count += 2
