#!/usr/bin/env python
"""The reference's own benchmark protocol (test_runner.sh:5-24, main_manager.py:32-61): PBT wall time for
`mpirun -n N` (1 master + N-1 workers) x pop_size, 20 rounds x 1 epoch, MNIST model with 10 train steps per
epoch (mnist_model.py:162-165), synthetic data.  Prints the reference's result line
    n = {}, pop_size = {}, time = {}s
for every cell, then one JSON line with the grid.

Arms:  --arm native     native MNIST engine, fused device exploit, persistent members
       --arm native-mf  same, `--members-in-flight K` members of a worker trained at once
       --arm restated   torch-eager members, checkpoint-file exploit (what the reference's TF graph amounts to)
All workers of a run share the visible GPUs round-robin (the reference: one GPU per pod)."""
import argparse
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedtf_b200.config import RunConfig  # noqa: E402
from distributedtf_b200.launch import run_local  # noqa: E402


def one(arm, n, pop, rounds, k, ckpt_every=1):
    d = tempfile.mkdtemp(prefix="dtf_sweep_")
    cfg = RunConfig(target_model="mnist", population_size=pop, train_round=rounds, epochs_per_round=1,
                    workers=n - 1, transport="local", max_train_steps=10, eval_batches=4,
                    engine="eager" if arm == "restated" else "native",
                    exploit_path="file" if arm == "restated" else "p2p",
                    members_in_flight=k if arm == "native-mf" else 1,
                    checkpoint_every=1 if arm == "restated" else ckpt_every,
                    save_root=d + "/savedata", results_file=d + "/test_results.txt", verbose=False, seed=7).validate()
    t0 = time.time()
    out = run_local(cfg)
    return time.time() - t0, out["pop_size"]


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["native", "native-mf", "restated"], default="native")
    ap.add_argument("--n", type=int, nargs="+", default=[2, 3, 4, 5])
    ap.add_argument("--pop", type=int, nargs="+", default=[10, 20, 30, 40, 50])
    ap.add_argument("--rounds", type=int, default=20)
    ap.add_argument("--members-in-flight", type=int, default=4)
    ap.add_argument("--checkpoint-every", type=int, default=1, help="native arms: rounds between checkpoints (0 = last only)")
    a = ap.parse_args()
    import faulthandler
    faulthandler.enable()
    if os.environ.get("DTF_DUMP_AFTER"):             # debugging aid: all thread stacks if the run stalls
        faulthandler.dump_traceback_later(float(os.environ["DTF_DUMP_AFTER"]), exit=True)
    grid = []
    one(a.arm, 2, 2, 1, a.members_in_flight)          # warm-up: library load, kernel attributes
    for n in a.n:
        for pop in a.pop:
            secs, alive = one(a.arm, n, pop, a.rounds, a.members_in_flight, a.checkpoint_every)
            print("n = {}, pop_size = {}, time = {}s".format(n, pop, secs), flush=True)
            grid.append({"n": n, "pop_size": pop, "time_s": round(secs, 3), "alive_at_end": alive})
    print(json.dumps({"bench": "pbt_sweep", "arm": a.arm, "rounds": a.rounds, "model": "mnist (10 train steps/epoch)",
                      "members_in_flight": a.members_in_flight if a.arm == "native-mf" else 1,
                      "checkpoint_every": 1 if a.arm == "restated" else a.checkpoint_every, "grid": grid}))
