#!/usr/bin/env python
"""Which shared-memory word does tcgen05.mma read for A[m][k]?  (bench/csrc/probe_umma.cu)  Prints, for a few
descriptor settings, the byte offset the hardware used for selected (m, k) -- the evidence behind the operand
layouts of the tf32 kernels."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtf_b200.ops import native as nat
import build_native
lib = ctypes.CDLL(build_native.build_bench())
lib.dtf_probe_umma.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_uint, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                               ctypes.c_void_p]
out = torch.zeros(128, 16, device="cuda")


def probe(kind, a_mn, lbo, sbo, layout=0):
    parts = []
    for p in (0, 1):
        out.zero_()
        rc = lib.dtf_probe_umma(kind, a_mn, lbo, sbo, layout, p, out.data_ptr(), nat.stream_ptr().value)
        torch.cuda.synchronize()
        assert rc == 0, rc
        parts.append(out.clone().round().long().cpu())
    if kind == 1:
        idx = parts[0] + (parts[1] << 10)      # 32-bit word index
        return idx * 4
    idx = parts[0] + (parts[1] << 7)           # 16-bit half-word index
    return idx * 2


CASES = [(1, 0, 2048, 128, 0, "tf32 K-major, no swizzle, LBO 2048 SBO 128 (forward path)"),
         (0, 1, 128, 2048, 0, "bf16 MN-major, no swizzle, LBO 128 SBO 2048 (bf16 wgrad path)"),
         (1, 1, 128, 2048, 0, "tf32 MN-major, no swizzle, LBO 128 SBO 2048"),
         (1, 1, 16384, 1024, 2, "tf32 MN-major, 128B swizzle, LBO 16384 SBO 1024"),
         (1, 1, 16384, 1024, 1, "tf32 MN-major, 128B swizzle with 32-byte base, LBO 16384 SBO 1024"),
         (1, 1, 16384, 512, 1, "tf32 MN-major, 128B swizzle with 32-byte base, LBO 16384 SBO 512"),
         (1, 1, 8192, 512, 4, "tf32 MN-major, 64B swizzle, LBO 8192 SBO 512"),
         (1, 1, 4096, 256, 6, "tf32 MN-major, 32B swizzle, LBO 4096 SBO 256"),
         (0, 1, 16384, 1024, 2, "bf16 MN-major, 128B swizzle, LBO 16384 SBO 1024")]
for kind, a_mn, lbo, sbo, layout, label in CASES:
    K = 8 if kind == 1 else 16
    off = probe(kind, a_mn, lbo, sbo, layout)
    print("==", label)
    for m in (0, 1, 2, 3, 4, 5, 7, 8, 9, 16, 31, 32, 33, 64, 127):
        print(f"  m={m:3d}: byte offsets for k=0..{K-1}:", off[m, :K].tolist())
