"""ctypes bindings to ``libdtf_b200.so`` (C ABI in ``include/dtf_b200.h``).

The structures below mirror the header field for field.  Wrappers take torch CUDA
tensors, pass raw device pointers and the *current* torch CUDA stream, and raise on any
non-zero return code — there is no eager fallback: on a GPU box a missing library is an
error (``require()``), so tests can never silently pass on a library path.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Any, Optional, Sequence

import torch

LIB_PATH = os.environ.get("DTF_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "lib", "libdtf_b200.so")
_lib: Optional[C.CDLL] = None

OK = 0
ERRORS = {-1: "DTF_ERR_SHAPE", -2: "DTF_ERR_SMEM", -3: "DTF_ERR_CUDA", -4: "DTF_ERR_ARG",
          -5: "DTF_ERR_LIMIT", -6: "DTF_ERR_IO"}

PRO_NONE, PRO_BN_RELU, PRO_BN_BWD = 0, 1, 3
EPI_STORE, EPI_RESIDUAL, EPI_BIAS_RELU, EPI_RELU_MASK, EPI_BIAS = 0, 1, 2, 3, 4
OPT_CHUNK = 1024
MAX_LR_BOUNDARIES = 100
BN_EPS, BN_DECAY = 1e-5, 0.997


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.isfile(LIB_PATH)


_lib_lock = threading.Lock()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        with _lib_lock:                      # published only after its signatures are declared
            if _lib is None:
                if not available():
                    raise NativeError(f"{LIB_PATH} is missing: run `python build_native.py` "
                                      "(or __graft_entry__.build()) before using the native engine")
                l = C.CDLL(os.path.abspath(LIB_PATH))
                _declare(l)
                if os.environ.get("DTF_CRASH_BT"):
                    l.dtf_install_crash_handler()
                _lib = l
    return _lib


def require() -> C.CDLL:
    """Loud failure when the extension is absent on a machine that has a GPU."""
    return lib()


def check(rc: int, what: str) -> None:
    if rc != OK:
        raise NativeError(f"{what} failed: {ERRORS.get(rc, rc)}")


def stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


# --------------------------------------------------------------------------- structs
class BnRef(C.Structure):
    _fields_ = [("stats", C.c_void_p), ("gamma", C.c_void_p), ("beta", C.c_void_p),
                ("run_mean", C.c_void_p), ("run_var", C.c_void_p),
                ("inv_count", C.c_float), ("eps", C.c_float), ("decay", C.c_float),
                ("channels", C.c_int), ("use_running", C.c_int)]


class Gather(C.Structure):
    _fields_ = [("src", C.c_void_p), ("src2", C.c_void_p), ("bwd_stats", C.c_void_p), ("bn", BnRef),
                ("mode", C.c_int), ("update_running", C.c_int),
                ("Hs", C.c_int), ("Ws", C.c_int), ("C", C.c_int),
                ("KH", C.c_int), ("KW", C.c_int), ("stride", C.c_int), ("pad", C.c_int),
                ("transposed", C.c_int), ("dtype", C.c_int), ("cpt_shift", C.c_int), ("kchunks_real", C.c_int)]


class ConvGemm(C.Structure):
    _fields_ = [("a", Gather), ("wpk", C.c_void_p), ("out", C.c_void_p), ("epi_res", C.c_void_p),
                ("epi_bias", C.c_void_p), ("epi_x", C.c_void_p), ("epi_bn", BnRef),
                ("epi_stats_out", C.c_void_p), ("epi_mode", C.c_int), ("epi_accumulate", C.c_int),
                ("NB", C.c_int), ("Ho", C.c_int), ("Wo", C.c_int), ("N", C.c_int),
                ("kchunks", C.c_int), ("kb_chunks", C.c_int), ("smem_b_off", C.c_int), ("smem_a_off", C.c_int)]


class ConvWgrad(C.Structure):
    _fields_ = [("x", Gather), ("gy", Gather), ("dw", C.c_void_p),
                ("NB", C.c_int), ("Ho", C.c_int), ("Wo", C.c_int), ("Cout", C.c_int),
                ("kchunks", C.c_int), ("kchunks_real", C.c_int), ("chunks_per_cta", C.c_int),
                ("smem_a_off", C.c_int)]


class Head(C.Structure):
    _fields_ = [("x", C.c_void_p), ("labels", C.c_void_p), ("w", C.c_void_p), ("bias", C.c_void_p),
                ("bn", BnRef), ("loss_out", C.c_void_p), ("correct_out", C.c_void_p),
                ("logits_out", C.c_void_p), ("dw", C.c_void_p), ("db", C.c_void_p), ("t_out", C.c_void_p),
                ("bwd_stats_out", C.c_void_p), ("loss_scale", C.c_float),
                ("batch", C.c_int), ("HW", C.c_int), ("C", C.c_int), ("classes", C.c_int),
                ("has_bn", C.c_int), ("training", C.c_int), ("update_running", C.c_int),
                ("board_hits", C.c_void_p), ("dtype", C.c_int)]


class Hparams(C.Structure):
    _fields_ = [("optimizer", C.c_int), ("regularizer", C.c_int), ("momentum", C.c_float),
                ("grad_decay", C.c_float), ("weight_decay", C.c_float), ("inv_loss_scale", C.c_float),
                ("n_boundaries", C.c_int), ("batch_size", C.c_int),
                ("boundaries", C.c_longlong * MAX_LR_BOUNDARIES),
                ("lr_values", C.c_float * (MAX_LR_BOUNDARIES + 1)), ("reserved", C.c_float * 3)]


class OptSegment(C.Structure):
    _fields_ = [("offset", C.c_int), ("count", C.c_int), ("regularized", C.c_int), ("pack", C.c_int),
                ("Cout", C.c_int), ("taps", C.c_int), ("Cin", C.c_int),
                ("fwd_pack_off", C.c_int), ("dgrad_pack_off", C.c_int), ("reserved", C.c_int * 3)]


class InitSegment(C.Structure):
    _fields_ = [("offset", C.c_int), ("count", C.c_int), ("kind", C.c_int), ("scale", C.c_float),
                ("c_pad", C.c_int), ("c_real", C.c_int), ("reserved", C.c_int * 2)]


INIT_CONST, INIT_UNIFORM, INIT_TRUNC_NORMAL = 0, 1, 2


def _declare(l: C.CDLL) -> None:
    vp, i, ll, f = C.c_void_p, C.c_int, C.c_longlong, C.c_float
    l.dtf_conv_gemm.argtypes = [C.POINTER(ConvGemm), vp]
    l.dtf_conv_wgrad.argtypes = [C.POINTER(ConvWgrad), vp]
    l.dtf_conv_bwd.argtypes = [C.POINTER(ConvGemm), C.POINTER(ConvWgrad), vp]
    l.dtf_conv_bwd.restype = i
    l.dtf_bn_bwd_apply.argtypes = [vp, vp, vp, vp, C.POINTER(BnRef), vp, ll, i, vp]
    l.dtf_bn_bwd_apply_dt.argtypes = [vp, vp, vp, vp, C.POINTER(BnRef), vp, ll, i, i, vp]
    l.dtf_bn_bwd_apply_dt.restype = i
    l.dtf_head.argtypes = [C.POINTER(Head), vp]
    l.dtf_bn_act_apply.argtypes = [vp, C.POINTER(BnRef), vp, C.POINTER(BnRef), vp, ll, i, i, vp]
    l.dtf_bn_act_apply.restype = i
    l.dtf_relu_bwd_stats.argtypes = [vp, vp, vp, vp, C.POINTER(BnRef), vp, vp, C.POINTER(BnRef), vp, ll, i, vp]
    l.dtf_relu_bwd_stats.restype = i
    l.dtf_bn_act_apply_dt.argtypes = [vp, C.POINTER(BnRef), vp, C.POINTER(BnRef), vp, ll, i, i, i, vp]
    l.dtf_bn_act_apply_dt.restype = i
    l.dtf_relu_bwd_stats_dt.argtypes = [vp, vp, vp, vp, C.POINTER(BnRef), vp, vp, C.POINTER(BnRef), vp, ll, i, i, vp]
    l.dtf_relu_bwd_stats_dt.restype = i
    l.dtf_pack_weights.argtypes = [vp, vp, vp, vp, i, vp]
    l.dtf_optimizer_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, i, vp, vp, f, vp]
    l.dtf_step_begin.argtypes = [vp, ll, vp, i, vp]
    l.dtf_fill.argtypes = [vp, ll, f, vp]
    l.dtf_init_params.argtypes = [vp, vp, i, i, C.c_ulonglong, vp]
    l.dtf_init_params.restype = i
    l.dtf_preprocess.argtypes = [vp, vp, vp, vp, i, i, i, i, i, C.c_ulonglong, C.c_ulonglong, vp, vp]
    l.dtf_preprocess_indexed.argtypes = [vp, vp, vp, vp, vp, vp, i, i, i, i, i, C.c_ulonglong, C.c_ulonglong, vp, vp]
    l.dtf_gemm.argtypes = [C.POINTER(Gemm), vp]
    l.dtf_maxpool_fwd.argtypes = [vp, vp, i, i, i, i, vp]
    l.dtf_maxpool_bwd.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, vp]
    l.dtf_act_mask.argtypes = [vp, vp, vp, vp, f, ll, i, vp]
    l.dtf_set_sm_limit.argtypes = [i]
    l.dtf_cifar_decode.argtypes = [vp, ll, vp, vp, i]
    l.dtf_read_file.argtypes = [C.c_char_p, ll, vp, ll]
    l.dtf_read_file.restype = ll
    l.dtf_idx_header.argtypes = [vp, ll, vp, vp, i]
    l.dtf_io_create.argtypes = [i]
    l.dtf_io_create.restype = vp
    l.dtf_io_submit.argtypes = [vp, C.c_char_p, vp, C.c_ulonglong, vp, i, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p]
    l.dtf_io_submit.restype = ll
    l.dtf_io_wait.argtypes = [vp, ll]
    l.dtf_io_wait_all.argtypes = [vp]
    l.dtf_io_pending.argtypes = [vp]
    l.dtf_io_bytes_written.argtypes = [vp]
    l.dtf_io_bytes_written.restype = ll
    l.dtf_io_destroy.argtypes = [vp]
    l.dtf_io_destroy.restype = None
    l.dtf_sm_count.argtypes = []
    for name in ("dtf_gemm", "dtf_maxpool_fwd", "dtf_maxpool_bwd", "dtf_act_mask"):
        getattr(l, name).restype = i
    for name in ("dtf_conv_gemm", "dtf_conv_wgrad", "dtf_bn_bwd_apply", "dtf_head", "dtf_pack_weights",
                 "dtf_optimizer_step", "dtf_step_begin", "dtf_fill", "dtf_preprocess"):
        getattr(l, name).restype = i
    _declare_optional(l)


class Gemm(C.Structure):
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("out", C.c_void_p), ("bias", C.c_void_p), ("step", C.c_void_p),
                ("lda", C.c_longlong), ("ldb", C.c_longlong), ("ldo", C.c_longlong),
                ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
                ("a_mn_major", C.c_int), ("b_mn_major", C.c_int), ("out_fp32", C.c_int), ("relu", C.c_int),
                ("dropout_keep", C.c_float), ("seed", C.c_uint),
                ("ws", C.c_void_p), ("tile_counters", C.c_void_p), ("max_tiles", C.c_int), ("k_split", C.c_int)]


def gemm(a, b, out, *, M: int, N: int, K: int, lda: int, ldb: int, ldo: int, a_mn_major=False, b_mn_major=False,
         bias=None, relu=False, dropout_keep: float = 1.0, seed: int = 0, step=None, ws=None, tile_counters=None) -> None:
    """out[M][N] = A x B^T (see include/dtf_b200.h: DtfGemm); ``out`` fp32 or bf16.  ``ws`` (fp32, >= M*N, zero)
    and ``tile_counters`` (int32, zero) enable split-K for shapes with few output tiles."""
    use_ws = ws is not None and tile_counters is not None and ws.numel() >= M * N
    p = Gemm(ptr(a), ptr(b), ptr(out), ptr(bias), ptr(step), lda, ldb, ldo, M, N, K, int(a_mn_major),
             int(b_mn_major), int(out.dtype == torch.float32), int(relu), dropout_keep, seed & 0xFFFFFFFF,
             ptr(ws) if use_ws else None, ptr(tile_counters) if use_ws else None,
             int(tile_counters.numel()) if use_ws else 0, 1)
    check(lib().dtf_gemm(C.byref(p), stream_ptr()), "dtf_gemm")


def maxpool_fwd(x, y, NB, H, W, Cc) -> None:
    check(lib().dtf_maxpool_fwd(ptr(x), ptr(y), NB, H, W, Cc, stream_ptr()), "dtf_maxpool_fwd")


def maxpool_bwd(x, y, gy, gx, db, NB, H, W, Cc) -> None:
    check(lib().dtf_maxpool_bwd(ptr(x), ptr(y), ptr(gy), ptr(gx), ptr(db), NB, H, W, Cc, stream_ptr()), "dtf_maxpool_bwd")


def act_mask(g_in, act, g_out, db, scale: float, rows: int, Cc: int) -> None:
    check(lib().dtf_act_mask(ptr(g_in), ptr(act), ptr(g_out), ptr(db), scale, rows, Cc, stream_ptr()), "dtf_act_mask")


def _declare_optional(l: C.CDLL) -> None:
    """Entry points added by later source files (exploit, all-reduce, score board, host IO)."""
    from . import native_ext
    native_ext.declare(l)


# --------------------------------------------------------------------------- builders
def bn_ref(channels: int, count: int, *, stats=None, gamma=None, beta=None, run_mean=None, run_var=None,
           use_running: bool = False, eps: float = BN_EPS, decay: float = BN_DECAY) -> BnRef:
    r = BnRef(ptr(stats), ptr(gamma), ptr(beta), ptr(run_mean), ptr(run_var),
              1.0 / float(count), eps, decay, channels, int(use_running))
    # the struct holds raw device pointers: keep the tensors alive as long as it lives, so that a temporary passed
    # here cannot be recycled by the caching allocator before the launch that reads it has been enqueued
    r._keep = (stats, gamma, beta, run_mean, run_var)
    return r


BF16, F32 = 0, 1               # DTF_BF16 / DTF_F32: element type of the tensors of a launch
PP_F32 = F32 << 8              # DTF_PP_F32: preprocess writes fp32


def dtype_code(t: Optional[torch.Tensor]) -> int:
    return F32 if t is not None and t.dtype == torch.float32 else BF16


def gather(src: torch.Tensor, Hs: int, Ws: int, Cc: int, *, KH=1, KW=1, stride=1, pad=0, transposed=False,
           mode=PRO_NONE, bn: Optional[BnRef] = None, src2=None, bwd_stats=None, update_running=False) -> Gather:
    """The element type of the launch follows the source tensor: bf16, or fp32 (tf32 tensor-core math)."""
    g = Gather(ptr(src), ptr(src2), ptr(bwd_stats), bn if bn is not None else BnRef(), mode,
               int(update_running), Hs, Ws, Cc, KH, KW, stride, pad, int(transposed), dtype_code(src), 0, 0)
    g._keep = (src, src2, bwd_stats, getattr(bn, "_keep", None))       # see bn_ref
    return g


def conv_gemm(a: Gather, wpk: torch.Tensor, out: torch.Tensor, *, NB: int, Ho: int, Wo: int, N: int,
              epi_mode=EPI_STORE, epi_res=None, epi_bias=None, epi_x=None, epi_bn: Optional[BnRef] = None,
              epi_stats_out=None, epi_accumulate=False) -> None:
    p = ConvGemm(a, ptr(wpk), ptr(out), ptr(epi_res), ptr(epi_bias), ptr(epi_x),
                 epi_bn if epi_bn is not None else BnRef(), ptr(epi_stats_out), epi_mode,
                 int(epi_accumulate), NB, Ho, Wo, N, 0, 0, 0, 0)
    check(lib().dtf_conv_gemm(C.byref(p), stream_ptr()), "dtf_conv_gemm")


def conv_gemm_params(a: Gather, wpk, out, *, NB, Ho, Wo, N, epi_mode=EPI_STORE, epi_res=None, epi_bias=None, epi_x=None,
                     epi_bn=None, epi_stats_out=None, epi_accumulate=False) -> ConvGemm:
    return ConvGemm(a, ptr(wpk), ptr(out), ptr(epi_res), ptr(epi_bias), ptr(epi_x),
                    epi_bn if epi_bn is not None else BnRef(), ptr(epi_stats_out), epi_mode,
                    int(epi_accumulate), NB, Ho, Wo, N, 0, 0, 0, 0)


def conv_bwd(dgrad: ConvGemm, x: Gather, gy: Gather, dw: torch.Tensor, *, NB: int, Ho: int, Wo: int, Cout: int) -> None:
    """dgrad + wgrad of one convolution (one fused launch for 3x3 / stride-1 layers)."""
    w = ConvWgrad(x, gy, ptr(dw), NB, Ho, Wo, Cout, 0, 0, 0, 0)
    check(lib().dtf_conv_bwd(C.byref(dgrad), C.byref(w), stream_ptr()), "dtf_conv_bwd")


def conv_wgrad(x: Gather, gy: Gather, dw: torch.Tensor, *, NB: int, Ho: int, Wo: int, Cout: int) -> None:
    p = ConvWgrad(x, gy, ptr(dw), NB, Ho, Wo, Cout, 0, 0, 0, 0)
    check(lib().dtf_conv_wgrad(C.byref(p), stream_ptr()), "dtf_conv_wgrad")


def bn_bwd_apply(t, x, res, out, bn: BnRef, bwd_stats, pixels: int, channels: int) -> None:
    check(lib().dtf_bn_bwd_apply_dt(ptr(t), ptr(x), ptr(res), ptr(out), C.byref(bn), ptr(bwd_stats),
                                    pixels, channels, dtype_code(out), stream_ptr()), "dtf_bn_bwd_apply")


def bn_act_apply(x, bn: BnRef, out, pixels: int, channels: int, *, sc=None, bn_sc: Optional[BnRef] = None,
                 update_running: bool = False) -> None:
    """ResNet v1 block tail: out = relu(bn(x) + shortcut); shortcut = none | sc | bn_sc(sc)."""
    check(lib().dtf_bn_act_apply_dt(ptr(x), C.byref(bn), ptr(sc), C.byref(bn_sc) if bn_sc is not None else None,
                                    ptr(out), pixels, channels, int(update_running), dtype_code(out), stream_ptr()),
          "dtf_bn_act_apply")


def relu_bwd_stats(dy, y, t_out, x, bn: BnRef, stats, pixels: int, channels: int, *, x2=None,
                   bn2: Optional[BnRef] = None, stats2=None) -> None:
    """Backward of the v1 block tail: t = dy * (y > 0) plus the BN-backward sums against x (and x2)."""
    check(lib().dtf_relu_bwd_stats_dt(ptr(dy), ptr(y), ptr(t_out), ptr(x), C.byref(bn), ptr(stats), ptr(x2),
                                      C.byref(bn2) if bn2 is not None else None, ptr(stats2), pixels, channels,
                                      dtype_code(t_out), stream_ptr()), "dtf_relu_bwd_stats")


def head(p: Head) -> None:
    check(lib().dtf_head(C.byref(p), stream_ptr()), "dtf_head")


def pack_weights(params, pack, segs_dev, chunks_dev, n_chunks: int) -> None:
    check(lib().dtf_pack_weights(ptr(params), ptr(pack), ptr(segs_dev), ptr(chunks_dev), n_chunks,
                                 stream_ptr()), "dtf_pack_weights")


def optimizer_step(params, grads, slot0, slot1, pack, segs_dev, chunks_dev, n_chunks: int, hp_dev, step_dev,
                   grad_scale: float = 1.0) -> None:
    check(lib().dtf_optimizer_step(ptr(params), ptr(grads), ptr(slot0), ptr(slot1), ptr(pack), ptr(segs_dev),
                                   ptr(chunks_dev), n_chunks, ptr(hp_dev), ptr(step_dev), grad_scale,
                                   stream_ptr()), "dtf_optimizer_step")


def step_begin(zero_region: torch.Tensor, step_dev: torch.Tensor, bump: bool = True) -> None:
    n = zero_region.numel() * zero_region.element_size() // 4
    check(lib().dtf_step_begin(ptr(zero_region), n, ptr(step_dev), int(bump), stream_ptr()), "dtf_step_begin")


def set_sm_limit(n: int) -> int:
    """SMs the persistent kernels occupy from now on (0 = all); baked into CUDA graphs captured afterwards."""
    return int(lib().dtf_set_sm_limit(int(n)))


def sm_count() -> int:
    return int(lib().dtf_sm_count())


def fill(t: torch.Tensor, value: float) -> None:
    check(lib().dtf_fill(ptr(t), t.numel(), float(value), stream_ptr()), "dtf_fill")


def init_params(params: torch.Tensor, segments, seed: int) -> None:
    """Device-side initialisation of the flat parameter arena: `segments` = InitSegment list (Philox, keyed by seed)."""
    arr = (InitSegment * len(segments))(*segments)
    dev = struct_to_tensor(arr, params.device)
    check(lib().dtf_init_params(ptr(params), ptr(dev), len(segments), max(s.count for s in segments),
                                seed & (2 ** 64 - 1), stream_ptr()), "dtf_init_params")
    del dev


def preprocess(images_u8: torch.Tensor, indices: Optional[torch.Tensor], out: torch.Tensor, *, mode: int,
               crop_flip: Optional[torch.Tensor] = None, seed: int = 0, counter: int = 0,
               n: Optional[int] = None, labels_all: Optional[torch.Tensor] = None,
               labels_out: Optional[torch.Tensor] = None, step: Optional[torch.Tensor] = None) -> None:
    _, H, W, Cc = images_u8.shape
    count = int(n if n is not None else (indices.numel() if indices is not None else images_u8.shape[0]))
    if out.dtype == torch.float32:
        mode |= PP_F32
    if labels_all is not None:          # batch gathered inside the kernel (images and labels)
        check(lib().dtf_preprocess_indexed(ptr(images_u8), ptr(indices), ptr(labels_all), ptr(labels_out), ptr(out),
                                           ptr(crop_flip), count, H, W, Cc, mode, seed & (2 ** 64 - 1),
                                           counter & (2 ** 64 - 1), ptr(step), stream_ptr()), "dtf_preprocess_indexed")
        return
    check(lib().dtf_preprocess(ptr(images_u8), ptr(indices), ptr(out), ptr(crop_flip), count, H, W, Cc, mode,
                               seed & (2 ** 64 - 1), counter & (2 ** 64 - 1), ptr(step), stream_ptr()), "dtf_preprocess")


def struct_to_tensor(s: C.Structure, device) -> torch.Tensor:
    """Copy a ctypes struct (or array of structs) into a device uint8 tensor."""
    raw = bytes(memoryview(s).cast("B"))
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device)


def pack_weight_host(w_ohwi: torch.Tensor, transpose_for_dgrad: bool = False, dtype=torch.bfloat16) -> torch.Tensor:
    """Host restatement of the packing: OHWI fp32 -> bf16 [K/8 (even)][N][8], or fp32 [K/4 (even)][N][4] for the
    tf32 kernels (oracle + tests)."""
    cout, kh, kw, cin = w_ohwi.shape
    if dtype == torch.float32:
        taps = kh * kw
        w = w_ohwi.reshape(cout, taps, cin)
        mat = w.permute(2, 1, 0).reshape(cin, taps * cout) if transpose_for_dgrad else w.reshape(cout, taps * cin)
        n, k = mat.shape
        kch_pad = (((k + 3) // 4) + 1) & ~1
        buf = torch.zeros(kch_pad * 4, n, dtype=torch.float32, device=w_ohwi.device)
        buf[:k] = mat.t()
        return buf.reshape(kch_pad, 4, n).permute(0, 2, 1).contiguous()
    taps = kh * kw
    w = w_ohwi.reshape(cout, taps, cin)
    if transpose_for_dgrad:
        mat = w.permute(2, 1, 0).reshape(cin, taps * cout)       # [N=ci][K = tap*Cout + co]
    else:
        mat = w.reshape(cout, taps * cin)                        # [N=co][K = tap*Cin + ci]
    n, k = mat.shape
    kch = (k + 7) // 8
    kch_pad = (kch + 1) & ~1
    buf = torch.zeros(kch_pad * 8, n, dtype=torch.float32, device=w_ohwi.device)
    buf[:k] = mat.t()
    return buf.reshape(kch_pad, 8, n).permute(0, 2, 1).contiguous().to(torch.bfloat16)
