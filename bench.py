#!/usr/bin/env python
"""Headline benchmark: CIFAR-10 ResNet-32 population-based training throughput.

    python bench.py --gpus N --steps K --warmup W [--impl native|restated|reference]

Metric (BASELINE.json): *CIFAR-10 ResNet PBT images/sec (box, max over ranks)*.  One
population member per GPU (population = N, weak scaling), ResNet-32 v2 on 32x32x3 images,
batch 128, Momentum-SGD with l2 regulariser (the hand-tuned point of reference
``cifar10_model.py:15-22``), synthetic uint8 data of CIFAR-10's shape, random-init weights.
Every timed step is a complete training step: input gather + augmentation, forward, backward,
BatchNorm statistic updates and the optimizer update.  One PBT exploit + explore event
(rank by loss -> winner->loser state transfer -> perturb hyper-parameters) happens in the
middle of the timed region when N > 1.

For N > 1 the driver launches this file under ``torch.distributed.run``; ranks talk NCCL.
The kernel-only number (`value`) draws batches from a device-resident 50 000-image dataset
(153.6 MB > L2); the end-to-end number (`e2e`) additionally copies every batch from pinned host
memory and reads the loss back every step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "CIFAR-10 ResNet-32 PBT images/sec (box, max over ranks)"
BATCH = 128
DATASET_IMAGES = 50000


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", choices=["native", "restated", "restated-graph", "reference"], default="native",
                   help="native = this framework; restated = the same model / update rules in torch eager on "
                        "cuDNN/cuBLAS (launch-bound); restated-graph = that step captured in a CUDA graph (what a "
                        "TF-1 graph executor amounts to; the strong library baseline); reference = unavailable here")
    p.add_argument("--batch", type=int, default=BATCH)
    p.add_argument("--resnet-size", type=int, default=32)
    p.add_argument("--resnet-version", type=int, default=2, choices=[1, 2])
    p.add_argument("--dtype", default="bf16", choices=["bf16", "tf32"],
                   help="native arm: bf16 operands (headline) or fp32 tensors with tf32 tensor-core math")
    p.add_argument("--bottleneck", action="store_true", help="bottleneck blocks (resnet-size 9n + 2); not the headline")
    p.add_argument("--dp-size", type=int, default=1, choices=[1, 2, 4],
                   help="GPUs per member (BASELINE config 4: population N/D x data-parallel D)")
    p.add_argument("--mode", choices=["step", "pbt"], default="step",
                   help="step = bare training steps (+ one PBT event); pbt = a whole PBT run through launch.run "
                        "(master, workers, control plane, evaluation, score board, exploit, explore, checkpoints)")
    p.add_argument("--model", choices=["cifar10", "mnist"], default="cifar10", help="--mode pbt: member model")
    p.add_argument("--rounds", type=int, default=6, help="--mode pbt: timed rounds (one more runs first as warm-up)")
    p.add_argument("--steps-per-round", type=int, default=100, help="--mode pbt: training steps per member per round")
    p.add_argument("--eval-batches", type=int, default=8, help="--mode pbt: evaluation batches per member per round")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-graphs", action="store_true")
    return p.parse_args()


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line).
    NVML in a background thread (20 ms period); falls back to an `nvidia-smi -lms` child process."""
    REASONS = {0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x8: "hw_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(visible.split(",")[self.gpu]) if visible and visible.split(",")[0].isdigit() else self.gpu
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._nvml = (pynvml, h)
            self._thread = threading.Thread(target=self._poll, daemon=True)
            self._thread.start()
            return
        except Exception:
            self._nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        pynvml, h = self._nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                mask = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self._nvml is not None:
            self._stop.set()
            self._thread.join(1.0)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# --------------------------------------------------------------------------- whole PBT runs
def pbt_mode(args) -> int:
    """BASELINE configs 2-4 through the product: `launch.spawn_tcp` starts one worker process per member GPU group
    (N / dp-size members), the master ranks from the device score board, losers pull winners over NVLink, every
    member evaluates each round and writes its checkpoint at the end.  One warm-up round (graph capture, peer
    mappings) runs before the timed rounds.  `value` is device-timed: images trained in the timed rounds divided by
    the slowest worker's train + eval + exploit device time; `wall` is the master's wall clock for the same rounds
    (control plane, host bookkeeping and CSV / checkpoint hand-over included)."""
    import tempfile
    import torch
    from distributedtf_b200 import launch
    from distributedtf_b200.config import RunConfig
    from distributedtf_b200.hpspace import default_cifar_hparam
    D = args.dp_size
    members = max(args.gpus // D, 1)
    tmp = tempfile.mkdtemp(prefix="dtf_pbt_bench_")
    hp = default_cifar_hparam()
    hp["batch_size"] = args.batch
    if args.model == "mnist":
        hp = {"opt_case": {"optimizer": "Momentum", "lr": 1e-4, "momentum": 0.9}, "batch_size": args.batch,
              "initializer": "glorot_normal", "decay_steps": 0, "decay_rate": 1.0, "weight_decay": 0.0,
              "regularizer": "None"}
    import json as _json
    hp_file = os.path.join(tmp, "initial_hp.json")
    with open(hp_file, "w") as f:
        _json.dump([{"model_id": i, "accuracy": 0.0, "hparams": hp} for i in range(members)], f)
    cfg = RunConfig(target_model=args.model, population_size=members, train_round=args.rounds + 1,
                    epochs_per_round=1, workers=members, transport="tcp" if args.gpus > 1 else "local", master_port=0,
                    dp_size=D, exploit_path="p2p", engine="native", resnet_size=args.resnet_size,
                    max_train_steps=args.steps_per_round, eval_batches=args.eval_batches, checkpoint_every=0,
                    save_root=os.path.join(tmp, "savedata"), results_file=os.path.join(tmp, "res.txt"),
                    initial_hp=hp_file, verbose=False, seed=1, hooks=[], recv_timeout=600.0).validate()
    timed = {}
    orig_train = launch.PBTCluster.train

    def train_with_warmup(self, round_num):
        """Round 0 untimed (graph capture, peer mappings, first checkpoints of nothing), then the timed rounds."""
        orig_train(self, 1)
        before = self.profiling_info()
        t0 = time.time()
        out = orig_train(self, round_num - 1)
        timed["wall"] = time.time() - t0
        after = self.profiling_info()
        timed["phases"] = {k: after[k] - before.get(k, 0.0) for k in after if k.startswith("device_") or k in
                           ("train", "explore", "exploit", "gather_wait")}
        timed["workers"] = len(self.worker_ranks)
        timed["history"] = len(self.history)
        timed["fallbacks"] = self.board_fallbacks
        return out
    launch.PBTCluster.train = train_with_warmup
    sampler = ClockSampler(0)
    sampler.start()
    try:
        out = launch.run(cfg)
    finally:
        launch.PBTCluster.train = orig_train
    clocks = sampler.stop()
    ph = timed["phases"]
    # images actually trained: explore perturbs batch sizes, so members leave the nominal batch after round 1
    # (the master's profile holds the mean over the workers it talks to)
    nominal = members * args.rounds * args.steps_per_round * args.batch
    images = ph.pop("device_images", 0.0) * timed["workers"] or nominal
    ph.pop("device_images_max", None)
    dev_s = ph.get("device_train_max", 0.0) + ph.get("device_eval_max", 0.0) + ph.get("device_exploit_max", 0.0)
    res = {"metric": METRIC if args.model == "cifar10" else "MNIST deep CNN PBT images/sec (box, max over ranks)",
           "value": images / max(dev_s, 1e-9), "unit": "images/s", "n_gpus": args.gpus, "steps": args.rounds * args.steps_per_round,
           "warmup": args.steps_per_round, "ms_per_step": dev_s / (args.rounds * args.steps_per_round) * 1e3,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "impl": "native", "mode": "pbt",
           "data": "synthetic uint8 images of the dataset's shape (random), random labels, random-init weights",
           "config": {"model": f"CIFAR-10 ResNet-{args.resnet_size} v2" if args.model == "cifar10" else "MNIST deep CNN",
                      "global_batch": args.batch * members, "per_gpu_batch": args.batch // D, "seq_len": None,
                      "parallelism": f"population {members} x {D} GPU, master + {members} workers over TCP, "
                                     f"p2p exploit, device score board",
                      "rounds": args.rounds, "steps_per_round": args.steps_per_round, "eval_batches": args.eval_batches,
                      "images_trained": images, "images_at_nominal_batch": nominal,
                      "what_is_timed": "train + eval + exploit device time of the slowest worker over the timed rounds"},
           "wall": {"seconds": timed["wall"], "images_per_s": images / max(timed["wall"], 1e-9)},
           "phases_s": ph, "exploit_events": timed["history"], "board_fallbacks": timed["fallbacks"],
           "pop_size_end": out["pop_size"], "clocks": clocks}
    print(json.dumps(res))
    return 0


# --------------------------------------------------------------------------- main
def main():
    args = parse()
    if args.mode == "pbt" and args.impl == "native":
        return pbt_mode(args)
    if args.impl == "reference":
        # The reference cannot be installed or run here: it has no packaging files at all and needs
        # TensorFlow 1.x + mpi4py + hyperopt (none in the image, no network).  See DESIGN.md.
        if os.environ.get("RANK", "0") == "0":           # one line per job, also under torchrun
            print(json.dumps({"impl": "reference",
                              "unavailable": "reference has no setup.py/pyproject (pip install fails) and needs "
                                             "TensorFlow 1.x (tf.contrib), mpi4py, hyperopt, mpirun: none installable offline"}))
        return 0

    import torch
    import torch.distributed as dist
    from distributedtf_b200.hpspace import default_cifar_hparam
    from distributedtf_b200 import explore

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    hp = default_cifar_hparam()
    hp["batch_size"] = args.batch
    D = args.dp_size
    if world % D or args.batch % D:
        raise SystemExit("--dp-size must divide both the GPU count and the batch")
    members, member_id, dp_rank = world // D, rank // D, rank % D
    B = args.batch // D                      # images per replica per step (global batch per member stays args.batch)
    dp_group = None
    if D > 1:
        for m in range(members):
            grp = dist.new_group(list(range(m * D, (m + 1) * D)))
            if m == member_id:
                dp_group = grp

    # ---- member under test ----
    from distributedtf_b200.models import resnet
    cfg = resnet.cifar10_config(args.resnet_size, args.resnet_version, bottleneck=args.bottleneck)
    symm_buf = None
    if args.impl == "native":
        from distributedtf_b200.models.resnet_native import NativeResNetEngine
        if args.bottleneck or args.resnet_version == 1:
            from distributedtf_b200.models import resnet_native_bottleneck as _rb, resnet_native_v1 as _r1
            NativeResNetEngine = {(2, True): _rb.NativeResNetBottleneckEngine, (1, True): _rb.NativeResNetBottleneckV1Engine,
                                  (1, False): _r1.NativeResNetV1Engine}[(args.resnet_version, args.bottleneck)]
        from distributedtf_b200.parallel import exploit, symm

        def alloc(n):
            nonlocal symm_buf
            symm_buf = exploit.MemberBuffer(n, dev)
            symm.DIRECTORY.register_local(rank, symm_buf.buf, rank)
            return symm_buf.alloc_state(n)

        eng = NativeResNetEngine(cfg, hp, dev, seed=1234 + member_id, use_graphs=not args.no_graphs,
                                 arena_alloc=alloc, dp_group=dp_group, dtype=args.dtype)
        if world > 1:
            symm.exchange_directory({rank: symm_buf.buf})
            for r in range(world):                       # map every peer's buffer now (cudaIpcOpenMemHandle is
                symm.DIRECTORY.pointer(r)                # milliseconds): PBT does this once after ADD_GRAPHS
    else:
        from distributedtf_b200.models.engine_eager import EagerEngine
        from distributedtf_b200.utils.data import cifar_preprocess_torch
        init = resnet.init_params(cfg, hp["initializer"], 1234 + member_id)
        eng = EagerEngine(resnet.param_specs(cfg), lambda p, x, t: resnet.forward(cfg, p, x, t), init, hp, dev,
                          compute_dtype=torch.bfloat16, lr_schedule=True, dp_group=dp_group)   # NCCL all-reduce for D > 1

    # ---- synthetic CIFAR-shaped data: resident uint8 dataset larger than L2 + pinned host copy ----
    g = torch.Generator().manual_seed(99 + rank)
    host_images = torch.randint(0, 256, (DATASET_IMAGES, 32, 32, 3), dtype=torch.uint8, generator=g).pin_memory()
    host_labels = torch.randint(0, 10, (DATASET_IMAGES,), dtype=torch.int32, generator=g).pin_memory()
    dev_images, dev_labels = host_images.to(dev), host_labels.to(dev)
    perm = torch.randperm(DATASET_IMAGES, generator=g)
    nbatches = DATASET_IMAGES // B
    idx_dev = perm[: nbatches * B].view(nbatches, B).to(dev)
    idx_host = perm[: nbatches * B].view(nbatches, B)

    def preprocess_on_device(x_u8):
        """cifar_preprocess_torch with the random draw on the device (no host round trip, graph-capturable)."""
        x = x_u8.float()
        n, h, w, c = x.shape
        padded = torch.zeros(n, h + 8, w + 8, c, device=dev)
        padded[:, 4:4 + h, 4:4 + w] = x
        oy = torch.randint(0, 9, (n,), device=dev)
        ox = torch.randint(0, 9, (n,), device=dev)
        flip = torch.rand(n, device=dev) < 0.5
        rows = oy[:, None] + torch.arange(h, device=dev)[None]
        cols = ox[:, None] + torch.arange(w, device=dev)[None]
        cols = torch.where(flip[:, None], cols.flip(1), cols)
        x = padded[torch.arange(n, device=dev)[:, None, None], rows[:, :, None], cols[:, None, :]]
        mean = x.mean(dim=(1, 2, 3), keepdim=True)
        std = x.std(dim=(1, 2, 3), keepdim=True, unbiased=False)
        return (x - mean) / torch.clamp(std, min=1.0 / float(h * w * c) ** 0.5)

    graphed = None
    if args.impl == "restated-graph":
        if D > 1:
            raise SystemExit("restated-graph covers one GPU per member; use --impl restated for the NCCL data-parallel arm")

        class GraphedStep:
            """Whole step (augmentation, forward, backward, update) of the torch model in ONE CUDA graph:
            static input buffers, autograd captured, cuDNN / cuBLAS kernels underneath."""

            def __init__(self):
                self.x = torch.zeros(B, 32, 32, 3, dtype=torch.uint8, device=dev)
                self.y = torch.zeros(B, dtype=torch.int64, device=dev)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    # a fresh leaf whose gradient-accumulator node is born on this (non-default) stream: autograd
                    # synchronises backward with the stream that node was created on, and the legacy default
                    # stream must not be touched during capture
                    eng.flat = eng.flat.detach().clone().requires_grad_(True)
                    eng._refresh_views()
                    for _ in range(3):
                        eng.train_step(preprocess_on_device(self.x), self.y)
                torch.cuda.current_stream().wait_stream(side)
                torch.cuda.synchronize()
                self.graph = torch.cuda.CUDAGraph()
                eng.flat.grad = None
                with torch.cuda.graph(self.graph):
                    self.loss = eng.train_step(preprocess_on_device(self.x), self.y)

            def __call__(self, idx):
                torch.index_select(dev_images, 0, idx, out=self.x)
                self.y.copy_(dev_labels[idx])
                self.graph.replay()
                return self.loss

            def step_from(self, images_u8, labels):
                self.x.copy_(images_u8, non_blocking=True)
                self.y.copy_(labels, non_blocking=True)
                self.graph.replay()
                return self.loss
        graphed = GraphedStep()

    def step_resident(i):
        idx = idx_dev[i % nbatches]
        if args.impl == "native":       # the step's first kernel gathers the batch from the resident dataset
            return eng.train_step_indexed(dev_images, dev_labels, idx, augment=True)
        if graphed is not None:
            return graphed(idx)
        x = cifar_preprocess_torch(dev_images[idx], True)
        return eng.train_step(x, dev_labels[idx])

    # staging for the end-to-end path: batch assembled in pinned memory, copied H2D every step
    stage_img = torch.empty(B, 32, 32, 3, dtype=torch.uint8).pin_memory()
    stage_lab = torch.empty(B, dtype=torch.int32).pin_memory()
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    d_img = torch.empty(B, 32, 32, 3, dtype=torch.uint8, device=dev)
    d_lab = torch.empty(B, dtype=torch.int32, device=dev)
    h2d_bytes = stage_img.numel() + stage_lab.numel() * 4
    d2h_bytes = 4

    def step_e2e(i):
        ih = idx_host[i % nbatches]
        torch.index_select(host_images, 0, ih, out=stage_img)      # host-side batch assembly (data loader)
        torch.index_select(host_labels, 0, ih, out=stage_lab)
        d_img.copy_(stage_img, non_blocking=True)
        d_lab.copy_(stage_lab, non_blocking=True)
        if args.impl == "native":
            loss = eng.train_step_u8(d_img, d_lab, augment=True)
        elif graphed is not None:
            loss = graphed.step_from(d_img, d_lab)
        else:
            loss = eng.train_step(cifar_preprocess_torch(d_img, True), d_lab)
        loss_host.copy_(loss.reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()                   # the user reads the loss every step
        return float(loss_host[0])

    # ---- PBT event through the product's data plane (no NCCL, no barrier, no host synchronisation of the
    #      training stream): every member publishes its round snapshot (device epoch flag), evaluates one batch
    #      whose hits land on the master's score board straight from the head kernel, and releases the round's
    #      epoch; the board is the only barrier.  Every rank reads the same board and derives the same
    #      truncation-selection pairing (a replicated controller: the ranking of parallel/scoreboard + master.py
    #      without the TCP hop); the loser's pull kernel acquires the winner's snapshot flag on the device, writes
    #      the explored hyper-parameters and resets its round cells in the same launch (exploit.pull_buffer).
    exploit_events = []
    pbt_round = [0]
    board = board_view = None
    board_ptr = 0
    eval_idx = idx_dev[nbatches - 1]
    if args.impl == "native":
        from distributedtf_b200.parallel import scoreboard as sb
        from distributedtf_b200.master import exploit_pairs
        info = [None]
        if rank == 0:
            board = sb.ScoreBoard(members, dev)
            info = [board.handle()]
        if world > 1:
            dist.broadcast_object_list(info, src=0)
        if rank == 0:
            board_ptr, board_view = board.ptr, board
        else:
            import ctypes as C
            from distributedtf_b200.ops import native as _nat
            out = C.c_void_p()
            _nat.check(_nat.lib().dtf_ipc_open_handle(info[0], C.byref(out)), "open score board")
            board_ptr = int(out.value)
            board_view = sb.ScoreBoard.view(board_ptr, members, dev)
        if dp_rank == 0:
            eng.board_slot_ptr = sb.slot_ptr(board_ptr, member_id)

    def pbt_begin_native():
        """Device half of the event: nothing here waits for the host or for another rank."""
        rnd = pbt_round[0]
        exploit.publish_buffer(symm_buf, rnd)                           # snapshot + release of its epoch flag
        if dp_rank == 0:
            sb.reset(board_ptr, member_id)
            eng.eval_batch_indexed(dev_images, dev_labels, eval_idx)    # hits -> board slot (head kernel, peer red)
            sb.release(board_ptr, member_id, rnd + 1, loss=eng.loss_cell)

    def pbt_finish_native():
        """Controller half: the board is the barrier.  It runs PBT_LOOKAHEAD training steps after the device half,
        so the GPU works on while this host thread waits for the slowest member's score (a loser's look-ahead
        steps are overwritten by the winner's state, which is what exploit means anyway)."""
        rnd = pbt_round[0]
        scores = board_view.read(rnd + 1, members=range(members), timeout=60.0, poll=0.0)
        order = sorted(range(members), key=lambda m: (scores[m], m))    # ascending, ties by id: same on every rank
        for top, bottom in exploit_pairs(order):
            winner, loser = order[top], order[bottom]
            if winner == loser or member_id != loser:
                continue
            import random as _random
            new_hp = dict(hp, opt_case=dict(hp["opt_case"]))
            explore.perturb_hparams(new_hp, _random.Random(1000 * rnd + loser))   # same draw on every replica
            new_hp["batch_size"] = args.batch                           # keep the benchmark's batch fixed
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            exploit.pull_buffer(symm_buf, symm.DIRECTORY.pointer(winner * D + dp_rank), rnd,
                                hp_offset=eng.hp_state_offset, new_hp=eng.hparams_struct(new_hp), zero=eng.round_cells)
            eng.adopt_hparams(new_hp)
            eng.after_state_replaced()
            e1.record()
            exploit_events.append((e0, e1))                             # read after the timed region
        pbt_round[0] += 1

    def pbt_event_restated():
        """Library arm: scores over NCCL all_gather, host ranking, NCCL send/recv of weights + slots."""
        if members == 1:
            return
        loss = eng.last_loss.reshape(1).float().clone()
        allv = [torch.zeros_like(loss) for _ in range(world)]
        dist.all_gather(allv, loss)
        scores = [-float(allv[m * D].item()) for m in range(members)]
        order = sorted(range(members), key=lambda m: scores[m])
        n = -(-members // 4)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for winner, loser in [(order[members - n + i], order[i]) for i in range(n)]:
            if winner == loser:
                continue
            flat = eng.flat.detach()
            if member_id == winner:       # replica r of the winner feeds replica r of the loser
                for t in (flat, eng.slot0, eng.slot1):
                    dist.send(t, dst=loser * D + dp_rank)
            elif member_id == loser:
                for t in (flat, eng.slot0, eng.slot1):
                    dist.recv(t, src=winner * D + dp_rank)
        e1.record()
        exploit_events.append((e0, e1))
        pbt_round[0] += 1

    if args.impl == "native":
        pbt_begin, pbt_finish = pbt_begin_native, pbt_finish_native
    else:
        pbt_begin, pbt_finish = (lambda: None), pbt_event_restated
    PBT_LOOKAHEAD = 2

    # ---- warm-up (includes one PBT event: eval graph capture, peer mappings and -- in the library arm -- NCCL's
    #      lazy point-to-point connection set-up all happen here, not in the timed region) ----
    for i in range(max(args.warmup, 3)):
        step_resident(i)
    if args.impl != "native" and world > 1:
        # library arms: open every NCCL point-to-point connection now (the timed event may pair other ranks than
        # the warm-up event did, and NCCL sets a connection up lazily on first use: ~300-450 ms)
        for peer in range(world):
            if peer == rank:
                continue
            t = torch.zeros(1, device=dev)
            if rank < peer:
                dist.send(t, dst=peer); dist.recv(t, src=peer)
            else:
                dist.recv(t, src=peer); dist.send(t, dst=peer)
        torch.cuda.synchronize()
    pbt_begin()
    pbt_finish()
    exploit_events.clear()
    barrier()

    def timed(fn, steps, with_pbt):
        sampler = ClockSampler(local)
        sampler.start()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for i in range(steps):
            fn(1000 + i)
            if with_pbt and i == steps // 2:
                pbt_begin()
            if with_pbt and i == min(steps // 2 + PBT_LOOKAHEAD, steps - 1):
                pbt_finish()
        e1.record()
        barrier()
        wall = time.time() - t0
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), wall, sampler.stop()

    dp_before = eng.dp.comm_raw() if getattr(eng, "dp", None) is not None else None
    dev_ms, wall, clocks = timed(step_resident, args.steps, True)
    dp_after = eng.dp.comm_raw() if dp_before is not None else None
    value = members * args.batch * args.steps / (dev_ms / 1e3)
    exploit_ms = None
    ems = torch.tensor([max([a.elapsed_time(b) for a, b in exploit_events], default=-1.0)], device=dev)
    if world > 1:
        dist.all_reduce(ems, op=dist.ReduceOp.MAX)
    if float(ems.item()) >= 0:
        exploit_ms = float(ems.item())

    e2e = None
    if not args.no_e2e:
        for i in range(3):
            step_e2e(i)
        e2e_ms, _, _ = timed(step_e2e, args.steps, False)
        e2e = {"value": members * args.batch * args.steps / (e2e_ms / 1e3), "unit": "images/s",
               "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
               "ms_per_step": e2e_ms / args.steps}

    launches = getattr(eng, "launches_per_step", 0) * args.steps
    dp_all = None
    if dp_before is not None:
        mine = eng.dp.comm_stats_between(dp_before, dp_after)
        dp_all = [None] * world
        dist.all_gather_object(dp_all, mine)
    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.dtype if args.impl == "native" else "bf16",
            "data": "synthetic uint8 CIFAR-10-shaped images (random), random labels, random-init weights",
            "impl": args.impl,
            "config": {"model": f"CIFAR-10 ResNet-{args.resnet_size} v{args.resnet_version} (resnet_model, "
                                + ("bottleneck blocks, 9n+2)" if args.bottleneck else "6n+2)"),
                       "global_batch": args.batch * members, "per_gpu_batch": B, "seq_len": None,
                       "parallelism": f"population {members} x {D} GPU" + (f" (pbt{members})" if D == 1 else f" (pbt{members} x dp{D}, fused all-reduce+update)"),
                       "optimizer": "Momentum 0.9 + l2 2e-4",
                       "l2": "inputs larger than L2: each step gathers a fresh batch from a 153.6 MB resident dataset",
                       "pbt_event_in_timed_region": True,
                       "pbt_event": ("snapshot publish + one eval batch scored on the device board + board read + "
                                     "truncation selection + device-ordered P2P pull with explore write"
                                     if args.impl == "native" else "NCCL all_gather of scores + NCCL send/recv exploit"),
                       "cuda_graphs": not args.no_graphs},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "launches_per_step": getattr(eng, "launches_per_step", 0), "exploit_ms": exploit_ms,
            "state_bytes": getattr(eng, "state_bytes", lambda: None)(), "wall_s": wall,
        }
        if getattr(eng, "dp", None) is not None:
            # device-timestamped accounting from inside the fused all-reduce + update kernel (CTA 0, %globaltimer):
            # time blocked on the peers' ready flags = exposed communication
            # -- over the device-timed region only (the e2e half adds host jitter between the replicas' launches);
            # the wait also contains the time this replica's backward finished ahead of the slowest one
            st = max(dp_all, key=lambda d: d["wait_us_per_call"])
            out["dp_comm"] = dict(st, per_rank_wait_us=[round(d["wait_us_per_call"], 2) for d in dp_all],
                                  per_rank_load_us=[round(d["load_us_per_call"], 2) for d in dp_all],
                                  exposed_fraction_of_step=st["wait_us_per_call"] / 1e3 / (dev_ms / args.steps),
                                  load_fraction_of_step=st["load_us_per_call"] / 1e3 / (dev_ms / args.steps))
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
