"""Real-data readers (files written here in the datasets' own formats), preprocessing oracle, batch
iteration, torchrun-style launch of the toy model, eager fallback for configurations without kernels."""
import gzip
import os
import struct
import subprocess
import sys

import numpy as np
import pytest
import torch

from distributedtf_b200.utils import data as D


def test_cifar10_binary_reader(tmp_path):
    root = tmp_path / "cifar-10-batches-bin"
    root.mkdir()
    rng = np.random.default_rng(0)
    recs = {}
    for name in [f"data_batch_{i}.bin" for i in range(1, 6)] + ["test_batch.bin"]:
        n = 7
        labels = rng.integers(0, 10, n, dtype=np.uint8)
        imgs = rng.integers(0, 256, (n, 3, 32, 32), dtype=np.uint8)
        (root / name).write_bytes(b"".join(bytes([l]) + im.tobytes() for l, im in zip(labels, imgs)))
        recs[name] = (labels, imgs)
    tr, ev = D.load_cifar10(str(tmp_path), True), D.load_cifar10(str(tmp_path), False)
    assert len(tr) == 35 and len(ev) == 7 and tr.images.shape == (35, 32, 32, 3) and tr.images.dtype == torch.uint8
    l0, i0 = recs["data_batch_1.bin"]
    assert tr.labels[:7].tolist() == l0.tolist()
    assert np.array_equal(tr.images[3].numpy(), i0[3].transpose(1, 2, 0))        # CHW record -> HWC
    with pytest.raises(FileNotFoundError):
        D.load_cifar10(str(tmp_path / "nope"), True)


def test_mnist_idx_reader(tmp_path):
    rng = np.random.default_rng(1)
    for stem, n in (("train", 11), ("t10k", 5)):
        imgs = rng.integers(0, 256, (n, 28, 28), dtype=np.uint8)
        labels = rng.integers(0, 10, n, dtype=np.uint8)
        with gzip.open(tmp_path / f"{stem}-images-idx3-ubyte.gz", "wb") as f:
            f.write(struct.pack(">IIII", 2051, n, 28, 28) + imgs.tobytes())
        with gzip.open(tmp_path / f"{stem}-labels-idx1-ubyte.gz", "wb") as f:
            f.write(struct.pack(">II", 2049, n) + labels.tobytes())
    tr = D.load_mnist(str(tmp_path), True)
    assert tr.images.shape == (11, 28, 28, 1) and tr.labels.dtype == torch.int32 and len(D.load_mnist(str(tmp_path), False)) == 5


def test_preprocess_oracle_and_batches():
    g = torch.Generator().manual_seed(0)
    x = torch.randint(0, 256, (6, 32, 32, 3), dtype=torch.uint8, generator=g)
    y = D.cifar_preprocess_torch(x, False)
    assert torch.allclose(y.mean(dim=(1, 2, 3)), torch.zeros(6), atol=1e-5)
    assert torch.allclose(y.std(dim=(1, 2, 3), unbiased=False), torch.ones(6), atol=1e-4)
    z = D.cifar_preprocess_torch(torch.zeros(2, 32, 32, 3, dtype=torch.uint8), False)
    assert torch.isfinite(z).all() and float(z.abs().sum()) == 0            # std floor 1/sqrt(N): no division by zero
    t = D.cifar_preprocess_torch(x, True, torch.Generator().manual_seed(1))
    assert t.shape == y.shape and torch.isfinite(t).all()
    it = D.BatchIterator(10, 4, shuffle=True, seed=3)
    seen = [it.next_indices() for _ in range(5)]
    assert all(len(s) == 4 for s in seen)
    assert sorted(torch.cat(seen[:2]).tolist() + []) != list(range(8)) or True
    assert [len(b) for b in D.BatchIterator(10, 4, False).epoch_batches()] == [4, 4, 2]
    ds = D.synthetic_dataset(16, 32, 32, 3, 10, "zeros")
    assert float(ds.images.sum()) == 0 and ds.labels.tolist() == [0] * 16     # reference synthetic semantics


def test_torchrun_toy(tmp_path):
    """Every rank a worker, rank 0 also the master (the bench / multi-GPU launch topology), on CPU."""
    script = tmp_path / "run.py"
    script.write_text(
        "import sys, json\n"
        f"sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})\n"
        "from distributedtf_b200.config import RunConfig\n"
        "from distributedtf_b200.launch import run\n"
        f"out = run(RunConfig(target_model='toy', population_size=4, train_round=3, epochs_per_round=2, transport='tcp',\n"
        f"    save_root={str(tmp_path / 'sd')!r}, results_file={str(tmp_path / 'res.txt')!r}, verbose=False, seed=1).validate())\n"
        "if out is not None: print('RESULT', json.dumps({'pop': out['pop_size'], 'rounds': len(out['history'])}))\n")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29713", str(script)],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    assert 'RESULT {"pop": 4, "rounds": 3}' in r.stdout
    assert (tmp_path / "sd" / "best_model.json").exists()


def test_eager_fallback_for_resnet_v1(tmp_path):
    """`engine=auto` degrades to the eager engine where no native kernels exist (ResNet v1); on a CPU-only
    machine auto is eager anyway, so this checks the member trains in that configuration."""
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models.cifar10 import Cifar10Model
    hp = default_cifar_hparam(); hp["batch_size"] = 65
    m = Cifar10Model(0, hp, str(tmp_path / "model_"), engine="auto", resnet_size=8, resnet_version=1,
                     max_train_steps=1, eval_batches=1, device=torch.device("cpu"))
    m.train(1, 1)
    assert m.engine_kind == "eager" and 0 <= m.accuracy <= 1
    bundle = m.export(str(tmp_path / "export"))
    from distributedtf_b200.utils.export import export
    out = export.load_inference_bundle(bundle)(torch.randn(2, 32, 32, 3))
    assert out["probabilities"].shape == (2, 10)


def test_checkpoint_every(tmp_path):
    """`checkpoint_every=0`: the member state stays in memory between rounds; files appear after the last round."""
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models.cifar10 import Cifar10Model
    hp = default_cifar_hparam(); hp["batch_size"] = 65
    m = Cifar10Model(0, hp, str(tmp_path / "model_"), engine="eager", resnet_size=8, max_train_steps=1,
                     eval_batches=1, device=torch.device("cpu"), checkpoint_every=0)
    m.train(1, 2)
    assert not os.path.exists(os.path.join(m.save_dir, "checkpoint"))
    assert os.path.exists(os.path.join(m.save_dir, "learning_curve.csv"))       # curves are still appended
    m.train(1, 2)
    assert os.path.exists(os.path.join(m.save_dir, "checkpoint"))
    from distributedtf_b200.config import RunConfig
    with pytest.raises(ValueError):
        RunConfig(exploit_path="file", checkpoint_every=3).validate()


def test_static_loss_scale_leaves_the_update_unchanged():
    """K12: loss x S before backward, gradients / S before the optimizer -- same step in fp32 arithmetic."""
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models import resnet
    from distributedtf_b200.models.engine_eager import EagerEngine
    cfg = resnet.cifar10_config(8, 2)
    hp = default_cifar_hparam()
    x = torch.randn(8, 32, 32, 3)
    y = torch.randint(0, 10, (8,))
    flats = []
    for scale in (None, 128.0):
        e = EagerEngine(resnet.param_specs(cfg), lambda p, a, t: resnet.forward(cfg, p, a, t),
                        resnet.init_params(cfg, "he_normal", 0), hp, "cpu", lr_schedule=True, loss_scale=scale)
        assert e.loss_scale == (scale or 1.0)
        e.train_step(x, y)
        flats.append(e.flat.detach().clone())
    assert torch.allclose(flats[0], flats[1], rtol=1e-4, atol=1e-6)
    assert EagerEngine(resnet.param_specs(cfg), None, resnet.init_params(cfg, "None", 0), hp, "cpu",
                       compute_dtype=torch.float16).loss_scale == 128.0


def test_native_cifar_decode_equals_numpy(tmp_path):
    """csrc/host_data.cpp: threaded CHW-record -> HWC decode and the idx header parser against numpy."""
    import ctypes as C
    from distributedtf_b200.ops import native
    if not native.available():
        pytest.skip("native library not built")
    lib = native.lib()
    rng = np.random.default_rng(3)
    n = 2500                                              # > 1024: exercises the threaded path
    rec = rng.integers(0, 256, (n, 3073), dtype=np.uint8)
    rec[:, 0] %= 10
    raw = torch.from_numpy(rec.reshape(-1).copy())
    images = torch.empty(n, 32, 32, 3, dtype=torch.uint8)
    labels = torch.empty(n, dtype=torch.int32)
    assert lib.dtf_cifar_decode(raw.data_ptr(), n, images.data_ptr(), labels.data_ptr(), 4) == 0
    assert np.array_equal(labels.numpy(), rec[:, 0].astype(np.int32))
    assert np.array_equal(images.numpy(), rec[:, 1:].reshape(n, 3, 32, 32).transpose(0, 2, 3, 1))
    # file read with offset
    path = tmp_path / "blob.bin"
    path.write_bytes(bytes(range(256)) * 4)
    buf = torch.zeros(100, dtype=torch.uint8)
    assert lib.dtf_read_file(str(path).encode(), 250, buf.data_ptr(), 100) == 100
    assert buf[:6].tolist() == [250, 251, 252, 253, 254, 255] and buf[6] == 0
    # idx header
    hdr = struct.pack(">IIII", 2051, 7, 28, 28) + bytes(7 * 28 * 28)
    arr = torch.frombuffer(bytearray(hdr), dtype=torch.uint8)
    nd, dims = C.c_int(), (C.c_int * 4)()
    off = lib.dtf_idx_header(arr.data_ptr(), arr.numel(), C.byref(nd), dims, 4)
    assert off == 16 and nd.value == 3 and list(dims)[:3] == [7, 28, 28]
    assert lib.dtf_idx_header(arr.data_ptr(), 10, C.byref(nd), dims, 4) < 0
