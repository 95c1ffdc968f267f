"""In-kernel timeline of one halo wgrad launch (library built with DTF_TIMING=1; use DTF_LIB to point at it)."""
import ctypes, sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributedtf_b200.ops import native as nat
NB, H, Cc = 128, int(sys.argv[1]) if len(sys.argv) > 1 else 32, int(sys.argv[2]) if len(sys.argv) > 2 else 16
dev = "cuda"
x, gy, c_out = [torch.randn(NB, H, H, Cc, device=dev).bfloat16() for _ in range(3)]
def stats_of(t):
    f = t.float().reshape(-1, Cc); return torch.cat([f.sum(0), (f * f).sum(0)]).contiguous()
st_x, st_o = stats_of(x), stats_of(c_out)
ones, zeros, bwd = torch.ones(Cc, device=dev), torch.zeros(Cc, device=dev), torch.zeros(2 * Cc, device=dev)
bn_x = nat.bn_ref(Cc, NB * H * H, stats=st_x, gamma=ones, beta=zeros)
bn_o = nat.bn_ref(Cc, NB * H * H, stats=st_o, gamma=ones, beta=zeros)
dw = torch.zeros(Cc, 3, 3, Cc, device=dev)
gx = nat.gather(x, H, H, Cc, KH=3, KW=3, stride=1, pad=1, mode=nat.PRO_BN_RELU, bn=bn_x)
gg = nat.gather(gy, H, H, Cc, mode=nat.PRO_BN_BWD, bn=bn_o, src2=c_out, bwd_stats=bwd)
lib = nat.lib()
lib.dtf_debug_timeline.argtypes = [ctypes.c_void_p]
names = {0: "setup done (pre-sync)", 1: "after sync", 4: "final sync", 32: "flush: acc_full", 33: "flush done"}
for i in range(4):
    names[8 + i] = f"mma: full[{i}] ready"; names[12 + i] = f"mma: tile {i} issued+committed"
    names[16 + 4 * i] = f"gather {i}: stage free, start loads"; names[17 + 4 * i] = f"gather {i}: rows stored"
    names[18 + 4 * i] = f"gather {i}: barrier passed"; names[19 + 4 * i] = f"gather {i}: expanded + arrived"
for i in range(4):
    names[64 + i] = f"expander {i}: halo ready"; names[72 + i] = f"expander {i}: copies issued"
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    nat.conv_wgrad(gx, gg, dw, NB=NB, Ho=H, Wo=H, Cout=Cc)
    e1.record()
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 128)()
    lib.dtf_debug_timeline(buf)
    t = list(buf); t0 = t[0]
    print(f"--- rep {rep}: event time {e0.elapsed_time(e1)*1000:.1f} us")
    for k, v in sorted(((k, v) for k, v in enumerate(t) if v and k in names), key=lambda kv: kv[1]):
        print(f"  {(v - t0)/1000:8.2f} us  {names[k]}")
