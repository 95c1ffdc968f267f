import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtf_b200.ops import native as nat
import build_native
lib = ctypes.CDLL(build_native.build_bench())     # micro-benchmark kernels live in their own library (bench/csrc)
lib.dtf_bench_umma.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_void_p]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
names = {0: "none", 6: "sw32", 4: "sw64", 2: "sw128"}
for mn in (0, 1):
    for layout in (0, 6, 4, 2):
        for n in (16, 32, 64, 144, 256):
            for shift in (0, 48 if layout == 0 else 128 * 3):
                if mn and layout != 0:
                    continue
                iters = 64
                rc = lib.dtf_bench_umma(layout, n, iters, shift, mn, out.data_ptr(), nat.stream_ptr().value)
                torch.cuda.synchronize()
                issue, total = out.tolist()
                print(f"mn_major={mn} layout={names[layout]:6s} N={n:3d} a_shift={shift:4d}: issue {issue/iters:7.1f} cyc/mma, "
                      f"complete {total/iters:7.1f} cyc/mma")
