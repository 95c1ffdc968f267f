"""In-kernel timeline of one conv3x3 launch (library built with -DDTF_TIMING)."""
import ctypes, sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtf_b200.ops import native as nat
NB, H, Cc = 128, int(sys.argv[1]) if len(sys.argv) > 1 else 16, int(sys.argv[2]) if len(sys.argv) > 2 else 32
x = torch.randn(NB, H, H, Cc, device="cuda").bfloat16()
w = (torch.randn(Cc, 3, 3, Cc, device="cuda") * 0.05)
res = torch.randn(NB, H, H, Cc, device="cuda").bfloat16()
xf = x.float().reshape(-1, Cc)
stats = torch.cat([xf.sum(0), (xf * xf).sum(0)]).contiguous()
gamma, beta = torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
rm, rv = torch.zeros(Cc, device="cuda"), torch.ones(Cc, device="cuda")
bn = nat.bn_ref(Cc, NB * H * H, stats=stats, gamma=gamma, beta=beta, run_mean=rm, run_var=rv)
wpk = nat.pack_weight_host(w)
out = torch.empty_like(x)
so = torch.zeros(2 * Cc, device="cuda")
g = nat.gather(x, H, H, Cc, KH=3, KW=3, stride=1, pad=1, mode=nat.PRO_BN_RELU, bn=bn)
lib = nat.lib()
lib.dtf_debug_timeline.argtypes = [ctypes.c_void_p]
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    nat.conv_gemm(g, wpk, out, NB=NB, Ho=H, Wo=H, N=Cc, epi_mode=nat.EPI_RESIDUAL, epi_res=res, epi_stats_out=so)
    e1.record()
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 128)()
    rc = lib.dtf_debug_timeline(buf)
    t = list(buf)
    t0 = t[0]
    names = {0: "setup done (pre-sync)", 1: "after sync", 2: "weights landed", 3: "role done (tid0)", 4: "final sync"}
    for i in range(4): names[8 + i] = f"mma: full[{i}] ready"; names[12 + i] = f"mma: tile {i} issued"
    for i in range(8): names[16 + i] = f"gather: batch {i} start"; names[24 + i] = f"gather: batch {i} arrived"
    for i in range(8): names[64 + i] = f"gather: batch {i} empty ok, loads next"; names[72 + i] = f"gather: batch {i} stored"; names[80 + i] = f"gather: batch {i} fenced"
    for i in range(8): names[88 + i] = f"MMA {i}: acc free"; names[96 + i] = f"MMA {i}: operands ready"; names[104 + i] = f"MMA {i}: issued+committed"; names[112 + i] = f"epi {i}: waiting for acc"
    for i in range(8): names[32 + i] = f"epi: acc_full tile {i}"; names[40 + i] = f"epi: tile {i} done"
    print(f"--- rep {rep}: event time {e0.elapsed_time(e1)*1000:.1f} us, rc={rc}")
    for k, v in sorted(((k, v) for k, v in enumerate(t) if v and k in names), key=lambda kv: kv[1]):
        print(f"  {(v - t0)/1000:8.2f} us  {names[k]}")
