"""Step-by-step comparison of the native backward pass with autograd on the oracle (debug aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from distributedtf_b200.models import resnet
from distributedtf_b200.models.resnet_native import NativeResNetEngine
from distributedtf_b200.ops import native
from distributedtf_b200.utils.data import cifar_preprocess_torch

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
size, nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8, 32
cfg = resnet.cifar10_config(size)
hp = {"opt_case": {"optimizer": "gd", "lr": 0.0}, "decay_steps": 0, "decay_rate": 0.1, "weight_decay": 0.0,
      "regularizer": "None", "initializer": "he_init", "batch_size": nb}
eng = NativeResNetEngine(cfg, hp, "cuda", seed=5, use_graphs=False)
named = eng.export_named()
g = torch.Generator().manual_seed(1)
imgs = torch.randint(0, 256, (nb, 32, 32, 3), dtype=torch.uint8, generator=g).cuda()
labels = torch.randint(0, 10, (nb,), generator=g).cuda()

x = cifar_preprocess_torch(imgs, False).bfloat16().float().permute(0, 3, 1, 2)
params = {k: v.clone().requires_grad_("moving" not in k) for k, v in named.items()}
taps = {}
logits = resnet.forward(cfg, params, x, True, taps, emulate_bf16=True)
for t in taps.values():
    t.retain_grad()
F.cross_entropy(logits, labels).backward()


def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-20)).item()


def nhwc(t):
    return t.permute(0, 2, 3, 1).contiguous()


eng.load_batch(imgs, labels)
eng._count = 0
native.step_begin(eng.zero_region, eng.step_dev, bump=True)
native.preprocess(eng.in_u8, None, eng.image, mode=1, n=nb)
x_last = eng._forward(nb, True)
torch.cuda.synchronize()
print("fwd image", rel(eng.image[:nb, ..., :3], nhwc(x)))
print("fwd conv0", rel(eng.r0[:nb], nhwc(taps["conv0"])))
for i, a in enumerate(eng.acts):
    print(f"fwd block{i} out", rel(a["out"][:nb], nhwc(taps[f"add{i}"])))

# backward, step by step (mirrors NativeResNetEngine._backward)
t_f = eng.tbuf[0]
eng._head(x_last, nb, True, t_out=t_f)
bfin = eng.bn_final
nbn = len(eng.bns)
shape = lambda h, c: (nb, h, h, c)
view = lambda buf, h, c: buf[: nb * h * h * c].view(nb, h, h, c)
print("t_final", rel(view(t_f, bfin.h, bfin.c), nhwc(taps[f"bn{nbn-1}"].grad)))
g_out = eng.gbuf[0]
native.bn_bwd_apply(t_f, x_last, None, g_out, eng._bn(bfin, nb, True), eng._bwd_stats(bfin), nb * bfin.h * bfin.h, bfin.c)
nblk = len(eng.blocks)
print("g_out(last)", rel(view(g_out, bfin.h, bfin.c), nhwc(taps[f"add{nblk-1}"].grad)))
cur = 0
bn_index = {id(b): i for i, b in enumerate(eng.bns)}
conv_index = {id(c): i for i, c in enumerate(eng.convs)}
for i in range(nblk - 1, -1, -1):
    b, a = eng.blocks[i], eng.acts[i]
    x_in = eng.acts[i - 1]["out"] if i > 0 else eng.r0
    t2, t1 = eng.tbuf[0], eng.tbuf[1]
    eng._wgrad(b.conv2, a["c1"], b.bn2, g_out, nb)
    eng._dgrad(b.conv2, g_out, nb, t2, mask_x=a["c1"], mask_bn=b.bn2)
    torch.cuda.synchronize()
    print(f"block{i} t2", rel(view(t2, b.hout, b.cout), nhwc(taps[f"bn{bn_index[id(b.bn2)]}"].grad)))
    eng._wgrad(b.conv1, x_in, b.bn1, t2, nb, gy_bn=b.bn2, gy_x=a["c1"])
    if b.proj:
        eng._wgrad(b.proj, x_in, b.bn1, g_out, nb)
    eng._dgrad(b.conv1, t2, nb, t1, mask_x=x_in, mask_bn=b.bn1, gy_bn=b.bn2, gy_x=a["c1"])
    torch.cuda.synchronize()
    if b.proj:
        eng._dgrad(b.proj, g_out, nb, t1, mask_x=x_in, mask_bn=b.bn1, accumulate=True)
    torch.cuda.synchronize()
    print(f"block{i} t1", rel(view(t1, b.hin, b.cin), nhwc(taps[f"bn{bn_index[id(b.bn1)]}"].grad)))
    nxt = eng.gbuf[cur ^ 1]
    native.bn_bwd_apply(t1, x_in, None if b.proj else g_out, nxt, eng._bn(b.bn1, nb, True), eng._bwd_stats(b.bn1),
                        nb * b.hin * b.hin, b.cin)
    g_out = nxt
    cur ^= 1
    torch.cuda.synchronize()
    ref = taps[f"add{i-1}"].grad if i > 0 else taps["conv0"].grad
    print(f"block{i} g_prev", rel(view(g_out, b.hin, b.cin), nhwc(ref)))
eng._wgrad(eng.conv0, eng.image, None, g_out, nb)
torch.cuda.synchronize()
grads = eng.export_named(eng.grads)
print({k: round(rel(grads[k], params[k].grad), 4) for k in grads if params[k].grad is not None})
