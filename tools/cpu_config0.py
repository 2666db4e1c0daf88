#!/usr/bin/env python
"""BASELINE.json configs[0], literally: the reference CPU backend (oracle/_ref, the reference's own
sources built with g++) on `scripts/cpu_benchmark.py 2000 1920 0 0` semantics — 2000 worlds x
1920 steps after 5 warm-up steps, sim_flags 0, seed 0, 2 hiders + 2 seekers, x / y actions
randint(-5, 5) every step (scripts/cpu_benchmark.py:21-35, 105-143; the two lines that need the
renderer are dropped, sim only) — with every host core and with ONE core, plus the C++ driver's
variant `src/headless.cpp CPU 2000 1920 --rand-actions` (3 hiders + 2 seekers, seed 5, x / y / r in
[0, 10], random grab / lock bits, src/headless.cpp:24-103).  TEST / MEASUREMENT INFRASTRUCTURE.

  python tools/cpu_config0.py [--one-core-steps 240] [--out profiles/r06_config0.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refsim


def run_benchmark_py(worlds, steps, threads):
    ref = refsim.RefSim(worlds, sim_flags=0, seed=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2,
                        num_threads=threads)
    ref.init()
    secs = ref.bench(5, steps, 0)
    rew = float(np.abs(ref.tensors()["reward"]).sum())
    ref.close()
    return {"worlds": worlds, "steps": steps, "seconds": secs, "env_steps_per_s": worlds * steps / secs,
            "threads": threads if threads > 0 else len(os.sched_getaffinity(0)), "reward_abs_sum": rew}


def run_headless(worlds, steps, threads):
    ref = refsim.RefSim(worlds, sim_flags=0, seed=5, min_hiders=3, max_hiders=3, min_seekers=2, max_seekers=2,
                        num_threads=threads)
    ref.init()
    act = ref.tensors()["action"]
    rng = np.random.RandomState(5)
    t0 = time.perf_counter()
    for _ in range(steps):
        a = np.empty(act.shape, dtype=np.int32)
        a[:, 0:3] = rng.randint(0, 11, size=(act.shape[0], 3))
        a[:, 3:5] = rng.randint(0, 2, size=(act.shape[0], 2))
        act[:] = a
        ref.step()
    secs = time.perf_counter() - t0
    ref.close()
    return {"worlds": worlds, "steps": steps, "seconds": secs, "env_steps_per_s": worlds * steps / secs,
            "threads": threads if threads > 0 else len(os.sched_getaffinity(0))}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=2000)
    ap.add_argument("--steps", type=int, default=1920)
    ap.add_argument("--one-core-steps", type=int, default=240)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    out = {"host_cores": len(os.sched_getaffinity(0)),
           "build": "oracle/_ref: reference sources, g++ -O2 -march=x86-64-v3 -ffp-contract=off"}
    out["cpu_benchmark_py_all_cores"] = run_benchmark_py(a.worlds, a.steps, 0)
    print(json.dumps(out["cpu_benchmark_py_all_cores"]), flush=True)
    out["cpu_benchmark_py_one_core"] = run_benchmark_py(a.worlds, a.one_core_steps, 1)
    print(json.dumps(out["cpu_benchmark_py_one_core"]), flush=True)
    out["headless_cpp_rand_actions_all_cores"] = run_headless(a.worlds, a.steps, 0)
    print(json.dumps(out["headless_cpp_rand_actions_all_cores"]), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)
