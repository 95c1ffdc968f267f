"""Dense GEMM, pooling and the native MNIST engine vs PyTorch references."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from distributedtf_b200.ops import native
    native.require()
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return native


def rnd(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", generator=g) * scale


def close(got, want, rtol, atol, what=""):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    bad = (err > atol + rtol * want.abs()).sum().item()
    assert bad == 0, f"{what}: {bad}/{got.numel()} off, max err {err.max().item():.4g} (ref max {want.abs().max().item():.4g})"


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (65, 1024, 3136), (256, 1024, 3136), (100, 200, 72), (7, 16, 8)])
def test_gemm_k_major(nat, M, N, K):
    a, b = rnd(M, K, seed=1).bfloat16(), rnd(N, K, scale=1 / math.sqrt(K), seed=2).bfloat16()
    bias = rnd(N, scale=0.1, seed=3)
    want = a.float() @ b.float().t()
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    nat.gemm(a, b, out, M=M, N=N, K=K, lda=K, ldb=K, ldo=N)
    close(out, want, 1e-2, 1e-2, "plain")
    nat.gemm(a, b, out, M=M, N=N, K=K, lda=K, ldb=K, ldo=N, bias=bias, relu=True)
    close(out, torch.relu(want + bias), 1e-2, 1e-2, "bias+relu")
    out32 = torch.zeros(M, N, device="cuda")
    nat.gemm(a, b, out32, M=M, N=N, K=K, lda=K, ldb=K, ldo=N)
    close(out32, want, 1e-4, 1e-3, "fp32 out")


@pytest.mark.parametrize("batch", [65, 128, 256])
def test_gemm_dense_backward_shapes(nat, batch):
    """dW = gz^T x flat (both operands MN-major, K = batch) and g_flat = gz x W (B MN-major)."""
    gz = rnd(batch, 1024, scale=0.05, seed=1).bfloat16()
    flat = rnd(batch, 3136, seed=2).bfloat16()
    w = rnd(1024, 3136, scale=0.02, seed=3).bfloat16()
    dw = torch.full((1024, 3136), float("nan"), device="cuda")
    nat.gemm(gz, flat, dw, M=1024, N=3136, K=batch, lda=1024, ldb=3136, ldo=3136, a_mn_major=True, b_mn_major=True)
    close(dw, gz.float().t() @ flat.float(), 1e-3, 2e-3, "dense wgrad")
    gflat = torch.full((batch, 3136), float("nan"), device="cuda", dtype=torch.bfloat16)
    nat.gemm(gz, w, gflat, M=batch, N=3136, K=1024, lda=1024, ldb=3136, ldo=3136, b_mn_major=True)
    close(gflat, gz.float() @ w.float(), 1e-2, 1e-3, "dense dgrad")


@pytest.mark.parametrize("batch", [65, 128, 256])
def test_gemm_split_k_matches_and_cleans_up(nat, batch):
    """Dense-layer shapes with few output tiles run split-K (partial tiles summed in an fp32 workspace, the last
    CTA of a tile applies the epilogue): same results as the single-pass kernel, twice in a row, and the
    workspace / counters are zero again afterwards."""
    x = rnd(batch, 3136, seed=1).bfloat16()
    w = rnd(1024, 3136, scale=0.02, seed=2).bfloat16()
    bias = rnd(1024, scale=0.1, seed=3)
    gz = rnd(batch, 1024, scale=0.05, seed=4).bfloat16()
    ws = torch.zeros(batch * 3136, device="cuda")
    ctr = torch.zeros(64, dtype=torch.int32, device="cuda")
    want_f = torch.relu(x.float() @ w.float().t() + bias)
    want_d = gz.float() @ w.float()
    for rep in range(2):
        h = torch.full((batch, 1024), float("nan"), device="cuda", dtype=torch.bfloat16)
        nat.gemm(x, w, h, M=batch, N=1024, K=3136, lda=3136, ldb=3136, ldo=1024, bias=bias, relu=True,
                 ws=ws, tile_counters=ctr)
        close(h, want_f, 1e-2, 2e-2, f"split-K forward rep {rep}")
        g = torch.full((batch, 3136), float("nan"), device="cuda", dtype=torch.bfloat16)
        nat.gemm(gz, w, g, M=batch, N=3136, K=1024, lda=1024, ldb=3136, ldo=3136, b_mn_major=True,
                 ws=ws, tile_counters=ctr)
        close(g, want_d, 1e-2, 2e-3, f"split-K dgrad rep {rep}")
        torch.cuda.synchronize()
        assert float(ws.abs().max()) == 0.0 and int(ctr.abs().max()) == 0


def test_gemm_dropout_statistics(nat):
    M, N, K = 256, 1024, 64
    a, b = torch.ones(M, K, device="cuda").bfloat16(), torch.ones(N, K, device="cuda").bfloat16() / K
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    step = torch.tensor([3], dtype=torch.int64, device="cuda")
    nat.gemm(a, b, out, M=M, N=N, K=K, lda=K, ldb=K, ldo=N, dropout_keep=0.6, seed=11, step=step)
    kept = (out > 0).float().mean().item()
    assert abs(kept - 0.6) < 0.01
    close(out[out > 0], torch.full_like(out[out > 0], 1 / 0.6), 1e-2, 1e-2, "inverted scaling")
    out2 = torch.empty_like(out)
    nat.gemm(a, b, out2, M=M, N=N, K=K, lda=K, ldb=K, ldo=N, dropout_keep=0.6, seed=11, step=step)
    assert torch.equal(out, out2)                                # same (seed, step) -> same mask
    step.fill_(4)
    nat.gemm(a, b, out2, M=M, N=N, K=K, lda=K, ldb=K, ldo=N, dropout_keep=0.6, seed=11, step=step)
    assert not torch.equal(out, out2)


def test_maxpool_and_act_mask(nat):
    NB, H, Cc = 5, 14, 64
    x = torch.relu(rnd(NB, H, H, Cc, seed=1)).bfloat16()
    y = torch.empty(NB, H // 2, H // 2, Cc, device="cuda", dtype=torch.bfloat16)
    nat.maxpool_fwd(x, y, NB, H, H, Cc)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, 2, 2)
    assert torch.equal(y.float(), yr.detach().permute(0, 2, 3, 1))
    gy = rnd(NB, H // 2, H // 2, Cc, seed=2).bfloat16()
    yr.backward(gy.float().permute(0, 3, 1, 2))
    want = xr.grad.permute(0, 2, 3, 1) * (x.float() > 0)
    gx = torch.full_like(x, float("nan"))
    db = torch.zeros(Cc, device="cuda")
    nat.maxpool_bwd(x, y, gy, gx, db, NB, H, H, Cc)
    # ties (several zeros in a window) may route differently, but those positions are masked by x > 0
    close(gx, want, 0, 0, "maxpool bwd")
    close(db, want.reshape(-1, Cc).sum(0), 1e-3, 1e-2, "bias grad")
    act = torch.relu(rnd(77, 1024, seed=3)).bfloat16()
    g = rnd(77, 1024, seed=4).bfloat16()
    out = torch.empty_like(g)
    db2 = torch.zeros(1024, device="cuda")
    nat.act_mask(g, act, out, db2, 1 / 0.6, 77, 1024)
    ref = g.float() * (act.float() > 0) / 0.6
    close(out, ref, 1e-2, 1e-3, "act mask")
    close(db2, ref.sum(0), 1e-2, 2e-2, "act mask db")


def _hp(opt="Momentum", lr=0.01, **extra):
    return {"opt_case": {"optimizer": opt, "lr": lr, **extra}, "decay_steps": 0, "decay_rate": 1.0, "weight_decay": 0.0,
            "regularizer": "None", "initializer": "glorot_normal", "batch_size": 96}


def test_mnist_device_initialisers(nat):
    """Weights of the MNIST member are drawn on the device: distribution per layer, zero biases, zero padding
    channels, a function of the seed."""
    from distributedtf_b200.models.mnist_native import NativeMnistEngine
    a = NativeMnistEngine(_hp(), "cuda", seed=4, use_graphs=False)
    b = NativeMnistEngine(_hp(), "cuda", seed=4, use_graphs=False)
    c = NativeMnistEngine(_hp(), "cuda", seed=5, use_graphs=False)
    assert torch.equal(a.params, b.params) and not torch.equal(a.params, c.params)
    n = a.export_named()
    assert all((n[k] == 0).all() for k in n if k.endswith("/bias"))
    assert tuple(n["conv0/kernel"].shape) == (32, 5, 5, 1)
    assert (a._view(a.params, a.c0_w, 32, 5, 5, 8)[..., 1:] == 0).all()
    std0 = math.sqrt(2.0 / (3136 + 1024))                                   # glorot_normal, cut at two sigma
    d0 = n["dense0/kernel"]
    assert d0.abs().max().item() <= 2 * std0 * 1.0001 and abs(d0.std().item() - 0.8796 * std0) < 0.01 * std0
    lim = math.sqrt(6.0 / (1024 + 10))                                      # logits layer: glorot_uniform
    d1 = n["dense1/kernel"]
    assert d1.abs().max().item() <= lim and abs(d1.std().item() - lim / math.sqrt(3)) < 0.03 * lim


def test_native_mnist_step_matches_oracle(nat):
    from distributedtf_b200.models import mnist_net
    from distributedtf_b200.models.mnist_native import NativeMnistEngine
    nb = 96
    eng = NativeMnistEngine(_hp(momentum=0.9), "cuda", seed=4, use_graphs=False)
    eng.dropout_keep = 1.0                       # the oracle cannot reproduce the counter-based mask
    named = {k: v.clone() for k, v in eng.export_named().items()}
    g = torch.Generator().manual_seed(1)
    imgs = (torch.rand(nb, 28, 28, 1, generator=g) < 0.2).to(torch.uint8) * torch.randint(1, 255, (nb, 28, 28, 1), generator=g, dtype=torch.uint8)
    imgs, labels = imgs.cuda(), torch.randint(0, 10, (nb,), generator=g).cuda()
    loss = eng.train_step_u8(imgs, labels, augment=True)
    torch.cuda.synchronize()
    params = {k: v.clone().requires_grad_(True) for k, v in named.items()}
    x = imgs.float().permute(0, 3, 1, 2)
    logits = mnist_net.forward({k: (v + (v.bfloat16().float() - v).detach()) if "kernel" in k and k != "dense1/kernel" else v
                                for k, v in params.items()}, x, False)
    ref_loss = F.cross_entropy(logits, labels)
    ref_loss.backward()
    close(eng.logits[:nb], logits.detach(), 3e-2, 3e-2 * logits.abs().max().item(), "logits")
    close(loss, ref_loss.detach().reshape(1), 3e-2, 3e-2, "loss")
    grads = eng.export_named(eng.grads)
    for k, p in params.items():
        err = ((grads[k] - p.grad).norm() / (p.grad.norm() + 1e-12)).item()
        assert err < 0.06, f"gradient {k}: relative error {err:.4f}"
    after = eng.export_named()
    for k in named:                               # Momentum, first step: w -= lr * g (constant lr, no schedule)
        close(after[k], named[k] - 0.01 * grads[k], 1e-4, 1e-7, f"updated {k}")
    assert eng.step == 1


def test_native_mnist_trains_and_evaluates(nat):
    from distributedtf_b200.models.mnist_native import NativeMnistEngine
    eng = NativeMnistEngine(_hp("Adam", 1e-3), "cuda", seed=2, use_graphs=True)
    g = torch.Generator().manual_seed(3)
    labels = torch.randint(0, 10, (128,), generator=g)
    imgs = torch.zeros(128, 28, 28, 1, dtype=torch.uint8)
    for i, l in enumerate(labels.tolist()):          # a learnable toy set: a bright bar whose row encodes the label
        imgs[i, 2 * l + 3: 2 * l + 6, 4:24, 0] = 200
    imgs, labels = imgs.cuda(), labels.cuda()
    losses = [eng.train_step_u8(imgs, labels).clone() for _ in range(60)]
    torch.cuda.synchronize()
    losses = [l.item() for l in losses]
    assert all(math.isfinite(l) for l in losses) and losses[-1] < 0.3 * losses[0], (losses[0], losses[-1])
    hits = eng.eval_batch_u8(imgs, labels).item()
    assert hits > 100, hits
    assert eng.step == 60 and ("train", 128) in eng._graphs
