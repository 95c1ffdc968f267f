"""The allocation-free exploit entry point, called through the C ABI.

Collected last on purpose (file name): a run that stops at the first failure has then already
exercised everything else.  The check is against the regular entry point on identical inputs —
the copy is a byte copy, so the tolerance is zero — and it is made twice in a row because the
library-owned scratch is shared between calls and keeps whatever the previous call left in it.

Why the entry point exists: a worker that adopts a winner's state while other members train on
the same GPU must not allocate (cudaMalloc would synchronise the device and stall them, and an
embedding application may own all device memory), yet the regular call needs 4 bytes of zeroed
device scratch from its caller.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from distributedtf_b200.ops import native
    l = native.require()
    vp, ll, ull = C.c_void_p, C.c_longlong, C.c_ulonglong
    l.dtf_exploit_copy.argtypes = [vp, vp, ll, ll, vp, vp, ll, vp, ull, vp, C.c_int, vp]
    l.dtf_exploit_copy.restype = C.c_int
    l.dtf_exploit_copy_noalloc.argtypes = [vp, vp, ll, ll, vp, vp, ll, vp, ull, vp]
    l.dtf_exploit_copy_noalloc.restype = C.c_int
    l.dtf_noalloc_last_error.argtypes = []
    l.dtf_noalloc_last_error.restype = C.c_char_p
    return native, l


def _hp(native, lr):
    h = native.Hparams()
    h.optimizer, h.regularizer, h.momentum, h.weight_decay, h.inv_loss_scale = 1, 2, 0.9, 2e-4, 1.0
    h.n_boundaries, h.batch_size = 1, 128
    h.boundaries[0] = 12345
    h.lr_values[0], h.lr_values[1] = lr, lr * 0.1
    return h


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    native, l = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    import gc
    gc.collect()                                             # members of earlier tests: let them go now, not mid-test
    torch.cuda.synchronize(dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    nbytes, hp_off = 5 * 1024 * 1024 + 4096, 1024
    for call in range(2):                                    # back to back: the scratch carries over
        g = torch.Generator(device=dev).manual_seed(call)
        src = torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device=dev, generator=g)
        hp = _hp(native, 0.1 * (call + 1))
        epoch = 10 + call
        # regular entry point: the caller supplies flag + zeroed arrival counter
        want = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        cells_a = torch.ones(32, dtype=torch.float32, device=dev)
        ctrl_a = torch.zeros(4, dtype=torch.int64, device=dev)
        rc = l.dtf_exploit_copy(want.data_ptr(), src.data_ptr(), nbytes, hp_off, C.byref(hp), cells_a.data_ptr(),
                                cells_a.numel() * 4, ctrl_a.data_ptr(), epoch, ctrl_a.data_ptr() + 8, 0, stream)
        assert rc == 0
        # allocation-free entry point: no scratch argument at all
        got = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        cells_b = torch.ones(32, dtype=torch.float32, device=dev)
        flag_b = torch.zeros(1, dtype=torch.int64, device=dev)
        before = torch.cuda.memory_allocated(dev)
        rc = l.dtf_exploit_copy_noalloc(got.data_ptr(), src.data_ptr(), nbytes, hp_off, C.byref(hp),
                                        cells_b.data_ptr(), cells_b.numel() * 4, flag_b.data_ptr(), epoch, stream)
        assert rc == 0, l.dtf_noalloc_last_error().decode()
        assert torch.cuda.memory_allocated(dev) == before     # nothing went through torch's allocator either
        torch.cuda.synchronize(dev)
        assert torch.equal(got, want), f"call {call}: outputs differ"
        assert torch.equal(cells_b, cells_a) and float(cells_b.abs().sum()) == 0.0
        assert int(flag_b.item()) == epoch == int(ctrl_a[0].item())
        n_hp = C.sizeof(native.Hparams)
        assert torch.equal(got[:hp_off], src[:hp_off]) and torch.equal(got[hp_off + n_hp:], src[hp_off + n_hp:])


def test_noalloc_rejects_oversized_and_misaligned_requests():
    native, l = _lib()
    torch.cuda.set_device(0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    rc = l.dtf_exploit_copy_noalloc(buf.data_ptr(), buf.data_ptr(), (1 << 30) + 16, 0, None, None, 0, None, 0, stream)
    assert rc == -5                                           # DTF_ERR_LIMIT
    msg = l.dtf_noalloc_last_error().decode()
    assert "exceeds the supported maximum" in msg and str(1 << 30) in msg
    rc = l.dtf_exploit_copy_noalloc(buf.data_ptr(), buf.data_ptr(), 100, 0, None, None, 0, None, 0, stream)
    assert rc == -4 and "multiples of 16" in l.dtf_noalloc_last_error().decode()
