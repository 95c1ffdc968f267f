"""The B200 training engine for the MNIST CNN.

Parity: reference ``mnist_model.py:62-126`` — conv 5x5 (1->32) + bias + ReLU, max-pool 2x2, conv 5x5
(32->64) + bias + ReLU, max-pool 2x2, dense 3136->1024 + ReLU, dropout 0.4 (training only), dense
1024->10, sparse softmax cross-entropy; six optimizers with a constant learning rate, no regulariser.

Kernels (``csrc/``): both convolutions forward / dgrad / wgrad on the generic tcgen05 implicit-GEMM
kernels (bias + ReLU fused into the epilogue), the 3136->1024 dense layer forward / dgrad / wgrad on the
tcgen05 GEMM (bias + ReLU + inverted dropout in the epilogue; the operands are consumed K-major or
MN-major straight from their row-major storage, so no transposed weight copy exists), max-pool forward
and backward (ReLU mask and bias gradient fused), the classifier head (dense 1024->10 + softmax-CE,
forward + backward), and the shared multi-tensor optimizer.  One CUDA graph per batch size.
"""
from __future__ import annotations

from typing import Any, Dict, Optional

import torch

from ..ops import native
from . import mnist_net
from .native_base import NativeEngineBase, _align

IMG_C = 8
DROPOUT_KEEP = 1.0 - mnist_net.DROPOUT_RATE


class NativeMnistEngine(NativeEngineBase):
    uses_regularizer = False
    train_pp_mode, eval_pp_mode = 0, 0        # pixels stay 0..255 floats, as in the reference

    def __init__(self, hparams: Dict[str, Any], device, *, seed: int = 0, dtype: str = "bf16",
                 max_batch: int = 256, use_graphs: bool = True, arena_alloc=None, dp_group=None,
                 all_reduce_alg=None):
        if dtype != "bf16":
            raise NotImplementedError("native engine computes in bf16 with fp32 master weights / accumulation")
        super().__init__(device, max_batch=max_batch, use_graphs=use_graphs, arena_alloc=arena_alloc,
                         dp_group=dp_group, seed=seed, all_reduce_alg=all_reduce_alg)
        self.image_size, self.image_channels = 28, 1
        self.dropout_keep = DROPOUT_KEEP
        self._build_plan()
        self._alloc_arena()
        self._alloc_buffers()
        self._init_params(hparams.get("initializer", "None"), seed)
        self.set_hparams(hparams, reset_slots=True)

    def _init_params(self, initializer: str, seed: int) -> None:
        """Weights drawn on the device (``dtf_init_params``, Philox keyed by the member's seed): the two convolutions
        and the 3136 -> 1024 layer take the member's initialiser, the logits layer the layer default, biases zero
        (``mnist_model.py:62-126``).  ``orthogonal`` (QR) and ``DTF_HOST_INIT=1`` use the host restatement."""
        import math
        import os
        if initializer == "orthogonal" or os.environ.get("DTF_HOST_INIT") == "1":
            self.load_named_params(mnist_net.init_params(initializer, seed))
            return

        def kernel(off, count, fan_in, fan_out, init, c_pad=0, c_real=0):
            if init == "glorot_normal":
                kind, scale = native.INIT_TRUNC_NORMAL, math.sqrt(2.0 / (fan_in + fan_out))
            elif init == "he_init":
                kind, scale = native.INIT_TRUNC_NORMAL, math.sqrt(2.0 / fan_in)
            else:
                kind, scale = native.INIT_UNIFORM, math.sqrt(6.0 / (fan_in + fan_out))
            return native.InitSegment(off, count, kind, scale, c_pad, c_real)

        zero = lambda off, n: native.InitSegment(off, n, native.INIT_CONST, 0.0, 0, 0)
        segs = [kernel(self.c0_w, 32 * 25 * IMG_C, 25 * 1, 25 * 32, initializer, IMG_C, 1), zero(self.c0_b, 32),
                kernel(self.c1_w, 64 * 25 * 32, 25 * 32, 25 * 64, initializer), zero(self.c1_b, 64),
                kernel(self.d0_w, 1024 * 3136, 3136, 1024, initializer), zero(self.d0_b, 1024),
                kernel(self.d1_w, 10 * 1024, 1024, 10, "None"), zero(self.d1_b, 10)]
        with torch.cuda.device(self.device):
            native.init_params(self.params, segs, 0x9E3779B97F4A7C15 * (seed + 1))
            self.repack()

    # ------------------------------------------------------------------ plan
    def _build_plan(self) -> None:
        S = native.OptSegment
        off = pack = 0
        self.segments = []

        def seg(count, pack_mode=0, cout=0, taps=0, cin=0, fwd=-1, dgrad=-1):
            nonlocal off
            self.segments.append(S(off, count, 0, pack_mode, cout, taps, cin, fwd, dgrad))
            o = off
            off += count
            return o

        def pack_room(chunks, n):
            nonlocal pack
            o = pack
            pack += _align(((chunks + 1) & ~1) * n * 8, 128)
            return o

        self.c0_pack = pack_room(25 * IMG_C // 8, 32)
        self.c0_w = seg(32 * 25 * IMG_C, 1, 32, 25, IMG_C, self.c0_pack, -1)
        self.c0_b = seg(32)
        self.c1_pack = pack_room(25 * 32 // 8, 64)
        self.c1_dpack = pack_room(25 * 64 // 8, 32)
        self.c1_w = seg(64 * 25 * 32, 1, 64, 25, 32, self.c1_pack, self.c1_dpack)
        self.c1_b = seg(64)
        self.d0_pack = pack
        pack += _align(1024 * 3136, 128)
        self.d0_w = seg(1024 * 3136, 2, 0, 0, 0, self.d0_pack, -1)
        self.d0_b = seg(1024)
        self.d1_w = seg(10 * 1024)
        self.d1_b = seg(10)
        self.n_params, self.n_pack, self.n_stats, self.n_running = off, pack, 0, 0
        self._finish_segments()

    def _alloc_buffers(self) -> None:
        dev, mb = self.device, self.max_batch
        bf = lambda *shape: torch.empty(*shape, dtype=torch.bfloat16, device=dev)
        self.image = torch.zeros(mb, 28, 28, IMG_C, dtype=torch.bfloat16, device=dev)
        self.c0, self.p0 = bf(mb, 28, 28, 32), bf(mb, 14, 14, 32)
        self.c1, self.p1 = bf(mb, 14, 14, 64), bf(mb, 7, 7, 64)
        self.h = bf(mb, 1024)
        self.t_h, self.g_z = bf(mb, 1024), bf(mb, 1024)
        self.g_p1, self.g_c1 = bf(mb, 7, 7, 64), bf(mb, 14, 14, 64)
        self.g_p0, self.g_c0 = bf(mb, 14, 14, 32), bf(mb, 28, 28, 32)
        self.logits = torch.zeros(mb, 10, dtype=torch.float32, device=dev)
        # split-K workspace of the dense-layer GEMMs (self-cleaning: zero between launches)
        self.gemm_ws = torch.zeros(mb * 3136, dtype=torch.float32, device=dev)
        self.gemm_ctr = torch.zeros(64, dtype=torch.int32, device=dev)

    # ------------------------------------------------------------------ params
    def _view(self, base, off, *shape):
        n = 1
        for s in shape:
            n *= s
        return base[off: off + n].view(*shape)

    @torch.no_grad()
    def load_named_params(self, named: Dict[str, torch.Tensor]) -> None:
        dev = self.device
        w0 = torch.zeros(32, 5, 5, IMG_C, device=dev)
        w0[..., :1] = named["conv0/kernel"].to(dev)
        self._view(self.params, self.c0_w, 32, 5, 5, IMG_C).copy_(w0)
        self._view(self.params, self.c0_b, 32).copy_(named["conv0/bias"].to(dev))
        self._view(self.params, self.c1_w, 64, 5, 5, 32).copy_(named["conv1/kernel"].to(dev))
        self._view(self.params, self.c1_b, 64).copy_(named["conv1/bias"].to(dev))
        self._view(self.params, self.d0_w, 1024, 3136).copy_(named["dense0/kernel"].to(dev))
        self._view(self.params, self.d0_b, 1024).copy_(named["dense0/bias"].to(dev))
        self._view(self.params, self.d1_w, 10, 1024).copy_(named["dense1/kernel"].to(dev))
        self._view(self.params, self.d1_b, 10).copy_(named["dense1/bias"].to(dev))
        self.repack()

    def export_named(self, base: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        b = self.params if base is None else base
        return {"conv0/kernel": self._view(b, self.c0_w, 32, 5, 5, IMG_C)[..., :1].clone(),
                "conv0/bias": self._view(b, self.c0_b, 32).clone(),
                "conv1/kernel": self._view(b, self.c1_w, 64, 5, 5, 32).clone(),
                "conv1/bias": self._view(b, self.c1_b, 64).clone(),
                "dense0/kernel": self._view(b, self.d0_w, 1024, 3136).clone(),
                "dense0/bias": self._view(b, self.d0_b, 1024).clone(),
                "dense1/kernel": self._view(b, self.d1_w, 10, 1024).clone(),
                "dense1/bias": self._view(b, self.d1_b, 10).clone()}

    # ------------------------------------------------------------------ kernels
    def _forward(self, nb: int, training: bool) -> None:
        n = native
        g0 = n.gather(self.image, 28, 28, IMG_C, KH=5, KW=5, stride=1, pad=2)
        n.conv_gemm(g0, self.pack[self.c0_pack:], self.c0, NB=nb, Ho=28, Wo=28, N=32, epi_mode=n.EPI_BIAS_RELU,
                    epi_bias=self.params[self.c0_b:])
        n.maxpool_fwd(self.c0, self.p0, nb, 28, 28, 32)
        g1 = n.gather(self.p0, 14, 14, 32, KH=5, KW=5, stride=1, pad=2)
        n.conv_gemm(g1, self.pack[self.c1_pack:], self.c1, NB=nb, Ho=14, Wo=14, N=64, epi_mode=n.EPI_BIAS_RELU,
                    epi_bias=self.params[self.c1_b:])
        n.maxpool_fwd(self.c1, self.p1, nb, 14, 14, 64)
        n.gemm(self.p1, self.pack[self.d0_pack:], self.h, M=nb, N=1024, K=3136, lda=3136, ldb=3136, ldo=1024,
               bias=self.params[self.d0_b:], relu=True, dropout_keep=self.dropout_keep if training else 1.0,
               seed=self._seed, step=self.step_dev, ws=self.gemm_ws, tile_counters=self.gemm_ctr)
        self._count += 5

    def _head(self, nb: int, training: bool, loss=None, correct=None) -> None:
        h = native.Head()
        h.x, h.labels = self.h.data_ptr(), self.labels.data_ptr()
        h.w, h.bias = self.params[self.d1_w:].data_ptr(), self.params[self.d1_b:].data_ptr()
        h.bn = native.BnRef()
        h.loss_out = (loss if loss is not None else self.loss_cell).data_ptr()
        h.correct_out = (correct if correct is not None else self.correct_cell).data_ptr()
        h.logits_out = self.logits.data_ptr()
        h.dw, h.db = self.grads[self.d1_w:].data_ptr(), self.grads[self.d1_b:].data_ptr()
        h.t_out, h.bwd_stats_out = self.t_h.data_ptr(), 0
        h.loss_scale = 1.0
        h.batch, h.HW, h.C, h.classes = nb, 1, 1024, 10
        h.has_bn, h.training, h.update_running = 0, int(training), 0
        h.board_hits = 0 if training else self.board_slot_ptr     # evaluation hits go straight to the score board
        native.head(h)
        self._count += 1

    def _backward(self, nb: int) -> None:
        n = native
        # dense 1024: ReLU + dropout backward (+ bias gradient), then wgrad and dgrad GEMMs
        n.act_mask(self.t_h, self.h, self.g_z, self.grads[self.d0_b:], 1.0 / self.dropout_keep, nb, 1024)
        n.gemm(self.g_z, self.p1, self._view(self.grads, self.d0_w, 1024, 3136), M=1024, N=3136, K=nb, lda=1024,
               ldb=3136, ldo=3136, a_mn_major=True, b_mn_major=True)
        n.gemm(self.g_z, self.pack[self.d0_pack:], self.g_p1, M=nb, N=3136, K=1024, lda=1024, ldb=3136, ldo=3136,
               b_mn_major=True, ws=self.gemm_ws, tile_counters=self.gemm_ctr)
        # conv 2
        n.maxpool_bwd(self.c1, self.p1, self.g_p1, self.g_c1, self.grads[self.c1_b:], nb, 14, 14, 64)
        n.conv_wgrad(n.gather(self.p0, 14, 14, 32, KH=5, KW=5, stride=1, pad=2), n.gather(self.g_c1, 14, 14, 64),
                     self.grads[self.c1_w:], NB=nb, Ho=14, Wo=14, Cout=64)
        gd = n.gather(self.g_c1, 14, 14, 64, KH=5, KW=5, stride=1, pad=2, transposed=True)
        n.conv_gemm(gd, self.pack[self.c1_dpack:], self.g_p0, NB=nb, Ho=14, Wo=14, N=32)
        # conv 1
        n.maxpool_bwd(self.c0, self.p0, self.g_p0, self.g_c0, self.grads[self.c0_b:], nb, 28, 28, 32)
        n.conv_wgrad(n.gather(self.image, 28, 28, IMG_C, KH=5, KW=5, stride=1, pad=2), n.gather(self.g_c0, 28, 28, 32),
                     self.grads[self.c0_w:], NB=nb, Ho=28, Wo=28, Cout=32)
        self._count += 8

    def _train_kernels(self, nb: int, pp_mode: int) -> None:
        self._count = 0
        self._begin_step()
        self.preprocess_batch(nb, 0)
        self._count += 1
        self._forward(nb, True)
        self._head(nb, True)
        self._backward(nb)
        self._apply_update()

    def _eval_kernels(self, nb: int) -> None:
        self._count = 0
        self.preprocess_batch(nb, 0)
        self._forward(nb, False)
        self._head(nb, False, loss=self.eval_cells[0:1], correct=self.eval_cells[1:2].view(torch.int32))
