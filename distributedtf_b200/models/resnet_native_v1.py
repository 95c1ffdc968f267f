"""The B200 training engine for the CIFAR ResNet **v1** (post-activation building blocks).

Parity: reference ``resnet/resnet_model.py:127-168`` (``_building_block_v1``: conv -> BN -> ReLU -> conv -> BN,
add the shortcut, ReLU; a projection shortcut is a 1x1 conv followed by its own BN), ``:505-520`` (v1 applies
BN + ReLU right after the first convolution) and ``:537-539`` (no final BN: that belongs to v2).

Same kernels and the same fusion idea as the v2 engine (``resnet_native.py``): every convolution is a tcgen05
implicit GEMM that leaves its raw output plus the per-channel (sum, sum of squares) of the *next* BatchNorm in its
epilogue.  What differs is where the residual stream is materialised.  In v1 the block output
``y = relu(bn2(c2) + shortcut)`` is consumed by two convolutions and by the next shortcut, after a ReLU that sits
behind the add, so it is written once by a small tail kernel (``dtf_bn_act_apply``) and the next block gathers it
without a prologue; inside the block ``bn1 + ReLU`` still rides in the second convolution's operand gather.
Backward mirrors it: ``dtf_relu_bwd_stats`` turns ``dL/dy`` into ``t = dL/dy * (y > 0)`` and the BatchNorm-backward
sums, and the BN-backward combination itself is applied while the upstream kernels gather their operands
(``DTF_PRO_BN_BWD``), exactly as in the v2 blocks.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from ..ops import native
from .resnet_native import IMG_C, BnL, ConvL, NativeResNetEngine


@dataclass
class BlockV1:
    conv1: ConvL
    bn1: BnL
    conv2: ConvL
    bn2: BnL
    proj: Optional[ConvL]
    bn_p: Optional[BnL]
    cin: int
    cout: int
    hin: int
    hout: int
    stride: int


class NativeResNetV1Engine(NativeResNetEngine):
    version = 1

    # ------------------------------------------------------------------ plan (parameter order = models/resnet.py)
    def _build_plan(self) -> None:
        cfg = self.cfg
        self.convs: List[ConvL] = []
        self.bns: List[BnL] = []
        self.blocks: List[BlockV1] = []
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
            filters = cfg.num_filters * (2 ** i)
            for b in range(nblocks):
                stride = cfg.block_strides[i] if b == 0 else 1
                proj = conv(cin, filters, 1, stride, h) if b == 0 else None
                bn_p = bn(filters, h // stride) if b == 0 else None
                c1 = conv(cin, filters, 3, stride, h)
                bn1 = bn(filters, h // stride)
                c2 = conv(filters, filters, 3, 1, h // stride)
                bn2 = bn(filters, h // stride)
                self.blocks.append(BlockV1(c1, bn1, c2, bn2, proj, bn_p, cin, filters, h, h // stride, stride))
                cin = filters
                h //= stride
        self.bn_final = None
        self.final_c, self.final_h = cin, h
        self._layout_params()

    def _layout_params(self) -> None:
        """Flat parameter layout, identical in structure to the v2 engine's."""
        from .native_base import _align
        cfg = self.cfg
        off = pack = 0
        self.segments = []
        for c in self.convs:
            c.w_off = off
            n = c.cout * c.k * c.k * c.cin
            E = self.E                                      # 8 (bf16 operands) or 4 (fp32 operands, v2 engines only)
            kch = c.k * c.k * c.cin // E
            c.fwd_pack = pack
            pack += _align(((kch + 1) & ~1) * c.cout * E, 128)
            if c is not self.conv0:
                kchd = c.k * c.k * c.cout // E
                c.dgrad_pack = pack
                pack += _align(((kchd + 1) & ~1) * c.cin * E, 128)
            c.seg = len(self.segments)
            self.segments.append(native.OptSegment(off, n, 1, 1 if E == 8 else 3, c.cout, c.k * c.k, c.cin, c.fwd_pack,
                                                   c.dgrad_pack))
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
        dev, mb, cfg = self.device, self.max_batch, self.cfg
        S = cfg.image_size
        self.image = torch.zeros(mb, S, S, IMG_C, dtype=self.act_dtype, device=dev)

        def act(h, c):
            return torch.empty(mb, h, h, c, dtype=self.act_dtype, device=dev)

        self.c0 = act(self.conv0.hout, self.conv0.cout)        # raw conv0 output
        self.y0 = act(self.conv0.hout, self.conv0.cout)        # relu(bn0(c0)): input of the first block
        self.acts = [{"c1": act(b.hout, b.cout), "c2": act(b.hout, b.cout), "y": act(b.hout, b.cout),
                      "p": act(b.hout, b.cout) if b.proj else None} for b in self.blocks]
        big = max(b.hin * b.hin * b.cin for b in self.blocks)
        self.gbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]   # dL/dy, ping-pong
        self.tbuf = [torch.empty(mb * big, dtype=self.act_dtype, device=dev) for _ in range(2)]   # masked gradients
        self.logits = torch.zeros(mb, cfg.num_classes, dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ forward
    def _conv_raw(self, c: ConvL, src, out, nb, *, bn: Optional[BnL], training, stats_for: Optional[BnL]) -> None:
        """Convolution with an optional BN+ReLU prologue on its input; raw output + statistics of `stats_for`."""
        self._fwd_conv(c, src, out, nb, bn=bn, training=training, update=True, stats_for=stats_for)

    def _forward(self, nb: int, training: bool) -> torch.Tensor:
        self._conv_raw(self.conv0, self.image, self.c0, nb, bn=None, training=training, stats_for=self.bn0)
        native.bn_act_apply(self.c0, self._bn(self.bn0, nb, training), self.y0, nb * self.bn0.h ** 2, self.bn0.c,
                            update_running=training)
        self._count += 1
        x = self.y0
        for b, a in zip(self.blocks, self.acts):
            if b.proj:
                with self._fork():
                    self._conv_raw(b.proj, x, a["p"], nb, bn=None, training=training, stats_for=b.bn_p)
            self._conv_raw(b.conv1, x, a["c1"], nb, bn=None, training=training, stats_for=b.bn1)
            self._conv_raw(b.conv2, a["c1"], a["c2"], nb, bn=b.bn1, training=training, stats_for=b.bn2)
            if b.proj:
                self._join()
            native.bn_act_apply(a["c2"], self._bn(b.bn2, nb, training), a["y"], nb * b.hout ** 2, b.cout,
                                sc=a["p"] if b.proj else x,
                                bn_sc=self._bn(b.bn_p, nb, training) if b.proj else None, update_running=training)
            self._count += 1
            x = a["y"]
        return x

    def _head(self, x, nb: int, training: bool, t_out=None, loss=None, correct=None) -> None:
        h = native.Head()
        h.x, h.labels = x.data_ptr(), self.labels.data_ptr()
        h.w = self.params[self.dense_w_off:].data_ptr()
        h.bias = self.params[self.dense_b_off:].data_ptr()
        h.loss_out = (loss if loss is not None else self.loss_cell).data_ptr()
        h.correct_out = (correct if correct is not None else self.correct_cell).data_ptr()
        h.logits_out = self.logits.data_ptr()
        h.dw = self.grads[self.dense_w_off:].data_ptr()
        h.db = self.grads[self.dense_b_off:].data_ptr()
        h.t_out = t_out.data_ptr() if t_out is not None else 0
        h.bwd_stats_out = 0
        h.loss_scale = float(self.loss_scale)
        h.batch, h.HW, h.C, h.classes = nb, self.final_h ** 2, self.final_c, self.cfg.num_classes
        h.has_bn, h.training, h.update_running = 0, int(training), 0        # v1: no BN in front of the pool
        h.board_hits = 0 if training else self.board_slot_ptr
        h.dtype = native.dtype_code(x)
        native.head(h)
        self._count += 1

    # ------------------------------------------------------------------ backward
    def _bwd_v1(self, c: ConvL, x_src, x_bn: Optional[BnL], t_src, t_x, t_bn: BnL, nb, out, *, mask_x=None,
                mask_bn: Optional[BnL] = None, res=None) -> None:
        """Weight + input gradient of one conv whose output gradient is BN-backward(t_src; t_x, t_bn).
        Input-gradient epilogue: ReLU mask + statistics of `mask_bn` (inside a block), or plain / residual store
        (block input: the mask of the previous tail is applied by that tail's backward kernel)."""
        pro, bn_t, st = native.PRO_BN_BWD, self._bn(t_bn, nb, True), self._bwd_stats(t_bn)
        pad = (c.k - 1) // 2
        gd = native.gather(t_src, c.hout, c.hout, c.cout, KH=c.k, KW=c.k, stride=c.stride, pad=pad, transposed=True,
                           mode=pro, bn=bn_t, src2=t_x, bwd_stats=st)
        if mask_bn is not None:
            epi = dict(epi_mode=native.EPI_RELU_MASK, epi_x=mask_x, epi_bn=self._bn(mask_bn, nb, True),
                       epi_stats_out=self._bwd_stats(mask_bn))
        elif res is not None:
            epi = dict(epi_mode=native.EPI_RESIDUAL, epi_res=res)
        else:
            epi = dict(epi_mode=native.EPI_STORE)
        gx = native.gather(x_src, c.hin, c.hin, c.cin, KH=c.k, KW=c.k, stride=c.stride, pad=pad,
                           mode=native.PRO_BN_RELU if x_bn else native.PRO_NONE,
                           bn=self._bn(x_bn, nb, True) if x_bn else None)
        gg = native.gather(t_src, c.hout, c.hout, c.cout, mode=pro, bn=bn_t, src2=t_x, bwd_stats=st)
        dw = self.grads[c.w_off:]
        if c.k == 3 and c.stride == 1 and c.cin == c.cout:      # one fused launch (dgrad CTAs + wgrad CTAs)
            dgrad = native.conv_gemm_params(gd, self._wpk(c, dgrad=True), out, NB=nb, Ho=c.hin, Wo=c.hin, N=c.cin, **epi)
            native.conv_bwd(dgrad, gx, gg, dw, NB=nb, Ho=c.hout, Wo=c.hout, Cout=c.cout)
            self._count += 1
            return
        with self._fork():
            native.conv_wgrad(gx, gg, dw, NB=nb, Ho=c.hout, Wo=c.hout, Cout=c.cout)
        native.conv_gemm(gd, self._wpk(c, dgrad=True), out, NB=nb, Ho=c.hin, Wo=c.hin, N=c.cin, **epi)
        self._join()
        self._count += 2

    def _backward(self, x_last, nb: int) -> None:
        dy = self.gbuf[0]
        self._head(x_last, nb, True, t_out=dy)                 # dL/dy of the last block output (no BN, no mask)
        cur = 0
        for i in range(len(self.blocks) - 1, -1, -1):
            b, a = self.blocks[i], self.acts[i]
            y_in = self.acts[i - 1]["y"] if i > 0 else self.y0
            t, t1 = self.tbuf[0], self.tbuf[1]
            pix = nb * b.hout ** 2
            # tail: t = dy * (y > 0); BN-backward sums of bn2 (and of the projection's BN)
            native.relu_bwd_stats(dy, a["y"], t, a["c2"], self._bn(b.bn2, nb, True), self._bwd_stats(b.bn2), pix, b.cout,
                                  x2=a["p"] if b.proj else None, bn2=self._bn(b.bn_p, nb, True) if b.proj else None,
                                  stats2=self._bwd_stats(b.bn_p) if b.proj else None)
            self._count += 1
            nxt = self.gbuf[cur ^ 1]
            if b.proj:                                          # projection weight gradient on the side stream
                with self._fork():
                    gx = native.gather(y_in, b.hin, b.hin, b.cin, KH=1, KW=1, stride=b.stride, pad=0)
                    gg = native.gather(t, b.hout, b.hout, b.cout, mode=native.PRO_BN_BWD, bn=self._bn(b.bn_p, nb, True),
                                       src2=a["p"], bwd_stats=self._bwd_stats(b.bn_p))
                    native.conv_wgrad(gx, gg, self.grads[b.proj.w_off:], NB=nb, Ho=b.hout, Wo=b.hout, Cout=b.cout)
                    self._count += 1
            # conv2: output gradient = BN-backward(t; c2, bn2); input gradient masked by relu(bn1(c1)) -> t1
            self._bwd_v1(b.conv2, a["c1"], b.bn1, t, a["c2"], b.bn2, nb, t1, mask_x=a["c1"], mask_bn=b.bn1)
            if b.proj:
                self._join()
            # conv1: output gradient = BN-backward(t1; c1, bn1); input gradient + shortcut gradient -> dL/dy_in
            if b.proj:
                self._bwd_v1(b.conv1, y_in, None, t1, a["c1"], b.bn1, nb, nxt)
                gd = native.gather(t, b.hout, b.hout, b.cout, KH=1, KW=1, stride=b.stride, pad=0, transposed=True,
                                   mode=native.PRO_BN_BWD, bn=self._bn(b.bn_p, nb, True), src2=a["p"],
                                   bwd_stats=self._bwd_stats(b.bn_p))
                native.conv_gemm(gd, self._wpk(b.proj, dgrad=True), nxt, NB=nb, Ho=b.hin, Wo=b.hin, N=b.cin,
                                 epi_mode=native.EPI_RESIDUAL, epi_res=nxt)
                self._count += 1
            else:
                self._bwd_v1(b.conv1, y_in, None, t1, a["c1"], b.bn1, nb, nxt, res=t)
            dy = nxt
            cur ^= 1
        # first layer: y0 = relu(bn0(c0))
        t0 = self.tbuf[0]
        native.relu_bwd_stats(dy, self.y0, t0, self.c0, self._bn(self.bn0, nb, True), self._bwd_stats(self.bn0),
                              nb * self.bn0.h ** 2, self.bn0.c)
        gx = native.gather(self.image, self.conv0.hin, self.conv0.hin, self.conv0.cin, KH=self.conv0.k, KW=self.conv0.k,
                           stride=self.conv0.stride, pad=(self.conv0.k - 1) // 2)
        gg = native.gather(t0, self.conv0.hout, self.conv0.hout, self.conv0.cout, mode=native.PRO_BN_BWD,
                           bn=self._bn(self.bn0, nb, True), src2=self.c0, bwd_stats=self._bwd_stats(self.bn0))
        native.conv_wgrad(gx, gg, self.grads[self.conv0.w_off:], NB=nb, Ho=self.conv0.hout, Wo=self.conv0.hout,
                          Cout=self.conv0.cout)
        self._count += 2
