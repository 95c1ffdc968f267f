"""Exploit copy, score board, fused all-reduce+update: single-GPU semantics and 2-GPU peer paths."""
import ctypes as C
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from distributedtf_b200.ops import native
    native.require()
    return native


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _hp_struct(nat, lr=0.25, opt=1):
    h = nat.Hparams()
    h.optimizer, h.regularizer, h.momentum, h.grad_decay = opt, 2, 0.9, 0.8
    h.weight_decay, h.inv_loss_scale, h.n_boundaries, h.batch_size = 1e-3, 1.0, 1, 96
    h.boundaries[0] = 1000
    h.lr_values[0], h.lr_values[1] = lr, lr / 2
    return h


def test_exploit_copy_local(nat):
    """state copy + hyper-parameter block rewrite + round-cell reset + completion epoch, one launch."""
    from distributedtf_b200.parallel import exploit
    nbytes = 3 * 1024 * 1024 + 256
    src = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device="cuda")
    dst = torch.zeros_like(src)
    hp_off = 512
    hp = _hp_struct(nat, lr=0.125)
    cells = torch.ones(64, dtype=torch.float32, device="cuda")
    ctrl = torch.zeros(32, dtype=torch.int64, device="cuda")
    exploit.copy_kernel(dst.data_ptr(), src.data_ptr(), nbytes, hp_offset=hp_off, new_hp=hp, zero=cells,
                        done_flag_ptr=ctrl.data_ptr(), epoch=7, arrive_ptr=ctrl.data_ptr() + 8)
    torch.cuda.synchronize()
    n_hp = C.sizeof(nat.Hparams)
    assert torch.equal(dst[:hp_off], src[:hp_off]) and torch.equal(dst[hp_off + n_hp:], src[hp_off + n_hp:])
    want = torch.frombuffer(bytearray(bytes(memoryview(hp).cast("B"))), dtype=torch.uint8).cuda()
    assert torch.equal(dst[hp_off: hp_off + n_hp], want)
    assert torch.all(cells == 0)
    assert ctrl[0].item() == 7 and (ctrl[1].item() & 0xFFFFFFFF) == 0
    # plain copy (publish) leaves everything identical
    exploit.copy_kernel(dst.data_ptr(), src.data_ptr(), nbytes)
    torch.cuda.synchronize()
    assert torch.equal(dst, src)
    with pytest.raises(nat.NativeError):
        exploit.copy_kernel(dst.data_ptr(), src.data_ptr(), nbytes - 3)


def test_scoreboard_local(nat):
    from distributedtf_b200.parallel import scoreboard
    board = scoreboard.ScoreBoard(8, torch.device("cuda", 0))
    for member, hits in [(0, 10), (3, 73), (5, 100)]:
        cell = torch.tensor([hits], dtype=torch.int32, device="cuda")
        scoreboard.publish(board.ptr, member, cell, 100, epoch=4)
        torch.cuda.synchronize()
        assert cell.item() == 0                      # counter reset for the next round
    got = board.read(4, members=[0, 3, 5], timeout=5)
    assert got == {0: pytest.approx(0.10), 3: pytest.approx(0.73), 5: pytest.approx(1.0)}
    with pytest.raises(TimeoutError):
        board.read(5, members=[0], timeout=0.2)


def test_member_exploit_p2p_same_process(nat, tmp_path):
    """Two native CIFAR members in one process: SET + EXPLORE => loser == winner's snapshot, bit-identical
    (weights, slots, moving statistics, step), with the loser's explored hyper-parameters on the device."""
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200.models.cifar10 import Cifar10Model
    base = str(tmp_path / "model_")
    kw = dict(engine="native", resnet_size=8, max_train_steps=3, eval_batches=1)
    hp_w = default_cifar_hparam(); hp_w["batch_size"] = 72
    hp_l = default_cifar_hparam(); hp_l["batch_size"] = 80; hp_l["opt_case"] = {"optimizer": "Adam", "lr": 1e-3}
    winner, loser = Cifar10Model(0, hp_w, base, **kw), Cifar10Model(1, hp_l, base, **kw)
    winner.train(1, 1); loser.train(1, 1)
    assert winner.engine.step == 3 and loser.engine.step == 3
    snap = winner.engine.state.clone()
    winner.train(1, 1)                                  # winner already trains the next round ...
    import copy
    loser.set_values([1, winner.accuracy, copy.deepcopy(hp_w),
                      {"path": "p2p", "src_id": 0, "src_rank": 1, "round": 0, "explore_follows": True}])
    loser.need_explore = True
    loser.seed(3)
    loser.perturb_hparams()                             # ... while the loser pulls round 0's snapshot
    torch.cuda.synchronize()
    eng = loser.engine
    off, n_hp = eng.hp_state_offset, C.sizeof(nat.Hparams)
    assert torch.equal(eng.state[:off], snap[:off]) and torch.equal(eng.state[off + n_hp:], snap[off + n_hp:])
    assert eng.step == 3 and winner.engine.step == 6
    assert loser.hparams["opt_case"]["optimizer"] == "Momentum"          # adopted winner's optimizer type
    assert loser.hparams["opt_case"]["lr"] != hp_w["opt_case"]["lr"]      # ... and explored around its values
    dev_hp = nat.Hparams.from_buffer_copy(bytes(eng.hp_dev.cpu().numpy()))
    assert dev_hp.optimizer == 1 and dev_hp.batch_size == loser.hparams["batch_size"]
    want = eng.hparams_struct(loser.hparams)
    assert abs(dev_hp.lr_values[0] - want.lr_values[0]) < 1e-9 and abs(dev_hp.momentum - want.momentum) < 1e-9
    # packed weights were refreshed from the adopted parameters: training continues finite
    loser.train(1, 1)
    assert eng.step == 6 and loser.accuracy == loser.accuracy


# --------------------------------------------------------------------------- 2 GPUs
def _two_rank(rank, world, port, fn_name, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        out = globals()[fn_name](rank, world)
        q.put((rank, "ok", out))
    except Exception as e:  # surfaced in the parent
        import traceback
        q.put((rank, "error", traceback.format_exc()))
    finally:
        dist.barrier()
        dist.destroy_process_group()


def _run2(fn_name, world=2):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_two_rank, args=(r, world, port, fn_name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(60)
    for rank, status, out in res:
        assert status == "ok", f"rank {rank}:\n{out}"
    return {rank: out for rank, _, out in res}


def _peer_exploit_and_board(rank, world):
    import torch.distributed as dist
    from distributedtf_b200.ops import native as nat
    from distributedtf_b200.parallel import exploit, scoreboard, symm
    dev = torch.device("cuda", rank)
    nbytes = 8 * 1024 * 1024
    mb = exploit.MemberBuffer(nbytes, dev)
    state = mb.alloc_state(nbytes)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    state.copy_(torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=dev, generator=g))
    symm.DIRECTORY.register_local(rank, mb.buf, rank)
    board = scoreboard.ScoreBoard(4, dev) if rank == 0 else None
    handles = [None, None]
    dist.all_gather_object(handles, board.handle() if rank == 0 else None)
    symm.exchange_directory({rank: mb.buf})
    exploit.publish_buffer(mb, 0)                     # snapshot + release of snapshot_epoch[0] = 1
    out = {}
    if rank == 1:                                     # no barrier: the pull acquires the winner's flag on the device
        hp = nat.Hparams(); hp.optimizer = 4; hp.batch_size = 77
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        exploit.pull_buffer(mb, symm.DIRECTORY.pointer(0), 0, hp_offset=0, new_hp=hp)
        e1.record()
        torch.cuda.synchronize()
        out["ms"] = e0.elapsed_time(e1)
        out["state"] = state.cpu()
        out["flag"] = int(mb.control_tensor().view(torch.int64)[0].item())
        # score board lives on rank 0; publish through the peer mapping
        bptr = C.c_void_p()
        nat.check(nat.lib().dtf_ipc_open_handle(handles[0], C.byref(bptr)), "open board")
        cell = torch.tensor([41], dtype=torch.int32, device=dev)
        scoreboard.publish(int(bptr.value), 2, cell, 50, epoch=9)
        # member 3: hits arrive one by one from an evaluation kernel's reductions (dtf_head, board_hits) ...
        from distributedtf_b200.models import resnet
        from distributedtf_b200.models.resnet_native import NativeResNetEngine
        from distributedtf_b200.hpspace import default_cifar_hparam
        eng = NativeResNetEngine(resnet.cifar10_config(8), default_cifar_hparam(), dev, seed=3, use_graphs=False)
        eng.board_slot_ptr = scoreboard.slot_ptr(int(bptr.value), 3)
        g2 = torch.Generator().manual_seed(5)
        imgs = torch.randint(0, 256, (96, 32, 32, 3), dtype=torch.uint8, generator=g2).to(dev)
        labels = torch.randint(0, 10, (96,), generator=g2).to(dev)
        scoreboard.reset(int(bptr.value), 3)
        hits = int(eng.eval_batch_u8(imgs[:64], labels[:64]).item()) + int(eng.eval_batch_u8(imgs[64:], labels[64:]).item())
        scoreboard.release(int(bptr.value), 3, 9, loss=eng.loss_cell)
        # ... and member 1 is marked dropped because its loss is not finite
        scoreboard.reset(int(bptr.value), 1)
        scoreboard.release(int(bptr.value), 1, 9, loss=torch.full((1,), float("nan"), device=dev))
        torch.cuda.synchronize()
        out["hits"] = hits
        view = scoreboard.ScoreBoard.view(int(bptr.value), 4, dev)     # a non-owner can read the board as well
        out["view"] = view.read(9, members=[2, 3], timeout=10)
    else:
        out["state"] = state.cpu()
    dist.barrier()
    if rank == 0:
        out["scores"] = board.read(9, members=[1, 2, 3], timeout=10)
    return out


@pytest.mark.multigpu
def test_peer_exploit_and_scoreboard_two_gpus(nat):
    res = _run2("_peer_exploit_and_board")
    n_hp = C.sizeof(nat.Hparams)
    assert torch.equal(res[1]["state"][n_hp:], res[0]["state"][n_hp:])         # bit-exact over NVLink
    got_hp = nat.Hparams.from_buffer_copy(bytes(res[1]["state"][:n_hp].numpy()))
    assert got_hp.optimizer == 4 and got_hp.batch_size == 77
    assert res[1]["flag"] == 1
    sc = res[0]["scores"]
    assert sc[2] == pytest.approx(0.82) and sc[3] == pytest.approx(res[1]["hits"] / 96) and sc[1] != sc[1]
    assert res[1]["view"] == {2: pytest.approx(0.82), 3: pytest.approx(res[1]["hits"] / 96)}
    print(f"8 MiB peer pull: {res[1]['ms']*1000:.1f} us")


def _dp_step(rank, world):
    """DP-2 member: two replicas, each half of the batch.  Step 1 (eager launches) is checked against
    the TF-1.x Momentum rule applied to the MEAN of the two replicas' gradients; steps 2-4 replay the
    captured CUDA graph and must keep the replicas bit-identical."""
    import torch.distributed as dist
    from distributedtf_b200.models import resnet
    from distributedtf_b200.models.resnet_native import NativeResNetEngine
    dev = torch.device("cuda", rank)
    cfg = resnet.cifar10_config(8)
    hp = {"opt_case": {"optimizer": "Momentum", "lr": 0.1, "momentum": 0.9}, "decay_steps": 20, "decay_rate": 0.1,
          "weight_decay": 2e-4, "regularizer": "l2_regularizer", "initializer": "he_init", "batch_size": 64}
    eng = NativeResNetEngine(cfg, hp, dev, seed=5, use_graphs=True, dp_group=dist.group.WORLD)
    start = eng.params[: eng.n_params].clone()
    g = torch.Generator().manual_seed(1)
    imgs = torch.randint(0, 256, (64, 32, 32, 3), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (64,), generator=g)
    per = 64 // world
    lo, hi = rank * per, rank * per + per
    losses = [eng.train_step_u8(imgs[lo:hi].to(dev), labels[lo:hi].to(dev), augment=False).clone()]
    torch.cuda.synchronize()
    dist.barrier()
    mine = eng.grads[: eng.n_params].clone()
    both = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(both, mine)
    gmean = sum(both) / world
    reg = torch.zeros_like(start)
    reg[: eng.n_conv_params] = 2e-4 * start[: eng.n_conv_params]
    want = start - 0.1 * 64 / 128 * (gmean + reg)          # first Momentum step: accumulator = gradient
    err = (eng.params[: eng.n_params] - want).abs().max().item()
    for _ in range(3):
        losses.append(eng.train_step_u8(imgs[lo:hi].to(dev), labels[lo:hi].to(dev), augment=False).clone())
    torch.cuda.synchronize()
    return {"params": eng.params[: eng.n_params].cpu(), "step": eng.step, "loss": [l.item() for l in losses],
            "err": err, "grads_differ": not torch.equal(both[0], both[1]), "ref_start": start.cpu()}


@pytest.mark.multigpu
def test_dp2_allreduce_update_two_gpus(nat):
    res = _run2("_dp_step")
    assert res[0]["step"] == 4 and res[1]["step"] == 4
    # replicas stay bit-identical: every rank sums the same values in the same order
    assert torch.equal(res[0]["params"], res[1]["params"])
    assert not torch.equal(res[0]["params"], res[0]["ref_start"])
    assert all(l == l for l in res[0]["loss"])
    assert res[0]["grads_differ"]                      # the replicas really saw different shards
    assert res[0]["err"] < 1e-6 and res[1]["err"] < 1e-6, (res[0]["err"], res[1]["err"])


@pytest.mark.multigpu
def test_dp4_allreduce_update_four_gpus(nat):
    """Same check with the member sharded over four GPUs (one-shot all-reduce over three peers)."""
    if torch.cuda.device_count() < 4:
        pytest.skip("needs 4 GPUs")
    res = _run2("_dp_step", world=4)
    for r in range(1, 4):
        assert torch.equal(res[0]["params"], res[r]["params"])
    assert all(res[r]["step"] == 4 and res[r]["err"] < 1e-6 for r in range(4)), [res[r]["err"] for r in range(4)]
    assert res[0]["grads_differ"]


def _dp_step_counter_jump(rank, world):
    """The member's step counter is part of the exploited / restored state: it jumps when a data-parallel
    member adopts a winner that ran a different batch size, or resumes from a checkpoint.  The flag handshake
    counts exchanges with its own counter, so both directions (higher and lower step) keep working."""
    import torch.distributed as dist
    from distributedtf_b200.models import resnet
    from distributedtf_b200.models.resnet_native import NativeResNetEngine
    dev = torch.device("cuda", rank)
    hp = {"opt_case": {"optimizer": "Momentum", "lr": 0.05, "momentum": 0.9}, "decay_steps": 20, "decay_rate": 0.1,
          "weight_decay": 2e-4, "regularizer": "l2_regularizer", "initializer": "he_init", "batch_size": 32}
    eng = NativeResNetEngine(resnet.cifar10_config(8), hp, dev, seed=5, use_graphs=True, dp_group=dist.group.WORLD)
    g = torch.Generator().manual_seed(1)
    imgs = torch.randint(0, 256, (32, 32, 32, 3), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (32,), generator=g)
    per = 32 // world
    x, y = imgs[rank * per:(rank + 1) * per].to(dev), labels[rank * per:(rank + 1) * per].to(dev)
    steps = []
    for jump in (None, 5000, 3):          # adopt a member that is far ahead, then one that is behind
        if jump is not None:
            torch.cuda.synchronize()
            dist.barrier()
            eng.step_dev.fill_(jump)
        for _ in range(3):
            eng.train_step_u8(x, y, augment=False)
        torch.cuda.synchronize()
        steps.append(eng.step)
    return {"steps": steps, "params": eng.params[: eng.n_params].cpu()}


@pytest.mark.multigpu
def test_dp2_survives_step_counter_jumps(nat):
    res = _run2("_dp_step_counter_jump")
    assert res[0]["steps"] == [3, 5003, 6] and res[1]["steps"] == [3, 5003, 6]
    assert torch.equal(res[0]["params"], res[1]["params"])
    assert torch.isfinite(res[0]["params"]).all()


def _spin(lib, ns):
    if ns > 0:
        lib.dtf_debug_spin(C.c_ulonglong(int(ns)), C.c_void_p(torch.cuda.current_stream().cuda_stream))


def _publish_pull_stress(rank, world):
    """Flag protocol of the exploit transfer under randomised delays: in every round the winners overwrite their
    state with a fresh pattern and publish it, the losers pull it, each side behind a spin kernel of random length
    (0-300 us), with NO synchronisation between publish and pull inside a round.  The loser must end every round
    bit-identical to the winner's pattern of that round."""
    import random
    import torch.distributed as dist
    import build_native
    from distributedtf_b200.parallel import exploit, symm
    lib = C.CDLL(build_native.build_bench())
    lib.dtf_debug_spin.argtypes = [C.c_ulonglong, C.c_void_p]
    dev = torch.device("cuda", rank)
    nbytes = 1 << 20
    mb = exploit.MemberBuffer(nbytes, dev)
    state = mb.alloc_state(nbytes).view(torch.int32)
    symm.DIRECTORY.register_local(rank, mb.buf, rank)
    symm.exchange_directory({rank: mb.buf})
    half = world // 2
    is_winner = rank < half
    partner = (rank + half) % world
    peer_base = symm.DIRECTORY.pointer(partner) if not is_winner else 0
    rounds = int(os.environ.get("DTF_STRESS_ROUNDS", "1000"))
    rng = random.Random(1234 + rank)
    bad = 0
    pulled_early = 0
    for r in range(rounds):
        if is_winner:
            state.fill_(r * 65537 + rank)                 # this round's "weights"
            _spin(lib, rng.choice([0, 0, 20e3, 100e3, 300e3]))
            exploit.publish_buffer(mb, r)
        else:
            _spin(lib, rng.choice([0, 0, 20e3, 100e3, 300e3]))
            exploit.pull_buffer(mb, peer_base, r)
            torch.cuda.synchronize()
            want = r * 65537 + partner
            if not bool((state == want).all()):
                bad += 1
        # buffer reuse (parity) is protected by the round structure of PBT, here by a host barrier per round
        torch.cuda.synchronize()
        dist.barrier()
    flag = int(mb.control_tensor().view(torch.int64)[0].item())
    return {"bad": bad, "rounds": rounds, "done_flag": flag, "winner": is_winner}


@pytest.mark.multigpu
def test_exploit_flag_protocol_under_random_delays(nat):
    """1000 rounds, random spin kernels before publish and before pull, bit-identity every time."""
    world = 8 if torch.cuda.device_count() >= 8 else 2
    res = _run2("_publish_pull_stress", world=world)
    for r, out in res.items():
        assert out["bad"] == 0, (r, out)
        if not out["winner"]:
            assert out["done_flag"] == out["rounds"]        # the last pull released its completion epoch
