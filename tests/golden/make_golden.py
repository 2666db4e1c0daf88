#!/usr/bin/env python
"""Generates the committed golden fixtures from the reference itself.

Runs the UNMODIFIED reference CPU backend (oracle/_ref/libhs_ref.so, built from the sources
under /root/reference by oracle/Makefile) single-threaded on fixed seeds and a fixed action
tape, and stores per-world internal state + every exported tensor after init and after
selected steps.  The fixtures travel with the repo; the GPU box has no /root/reference.

    python tests/golden/make_golden.py [case ...]   # rewrites tests/golden/<case>.npz (default: all)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refsim  # noqa: E402

# (name, worlds, seed, sim_flags, (min,max) hiders, (min,max) seekers, steps, snapshots, grab/lock)
CASES = [
    ("bench_seed0", 4, 0, 0, (2, 2), (2, 2), 12, (0, 1, 6, 12), False),
    ("ragged_seed5_grablock", 4, 5, 0, (1, 3), (1, 3), 110, (0, 100, 110), True),
    ("train_flags_seed5", 3, 5, 1 | 4 | 8, (3, 3), (3, 3), 104, (0, 104), True),
    ("debug_levels", 8, 2, 0, (1, 2), (1, 2), 90, (0, 3, 45, 90), True),
]

# triggerReset(world, level) issued before step 1 (src/level_gen.cpp:336-526 fixtures)
LEVELS = {"debug_levels": [2, 3, 4, 5, 6, 7, 8, 1]}

EXPORT_KEYS = ["prep_counter", "action", "self_obs", "self_type", "self_mask", "agent_obs", "box_obs",
               "ramp_obs", "agent_vis", "box_vis", "ramp_vis", "lidar", "seed", "reward", "done",
               "global_pos", "episode_result"]


def action_tape(name, steps, rows, zero_vel, grab_lock):
    rng = np.random.RandomState(sum(map(ord, name)))
    a = np.zeros((steps, rows, 5), dtype=np.int32)
    a[:, :, 0:3] = rng.randint(0, 5 if zero_vel else 11, size=(steps, rows, 3))
    if grab_lock:
        a[:, :, 3] = rng.rand(steps, rows) < 0.3
        a[:, :, 4] = rng.rand(steps, rows) < 0.15
    return a


def snapshot(ref, N):
    worlds = np.stack([ref.dump_world(w) for w in range(N)])
    rt = ref.tensors()
    out = {"worlds": worlds}
    for k in EXPORT_KEYS:
        out["ex_" + k] = np.array(rt[k])
    return out


def main():
    only = set(sys.argv[1:])
    for name, N, seed, flags, hs, sk, steps, snaps, grab_lock in CASES:
        if only and name not in only:
            continue
        ref = refsim.RefSim(N, sim_flags=flags, seed=seed, min_hiders=hs[0], max_hiders=hs[1],
                            min_seekers=sk[0], max_seekers=sk[1], num_threads=1)
        ref.init()
        A = hs[1] + sk[1]
        tape = action_tape(name, steps, N * A, bool(flags & 8), grab_lock)
        data = {"config": np.array([N, seed, flags, hs[0], hs[1], sk[0], sk[1], steps], dtype=np.int64),
                "snaps": np.array(snaps, dtype=np.int64), "tape": tape.astype(np.int8)}
        levels = LEVELS.get(name)
        if levels is not None:
            data["levels"] = np.array(levels, dtype=np.int64)
        rt = ref.tensors()
        for s in range(steps + 1):
            if s in snaps:
                for k, v in snapshot(ref, N).items():
                    data[f"s{s}_{k}"] = v
            if s < steps:
                rt["action"][:] = tape[s]
                if s == 1 and levels is not None:
                    for w, lvl in enumerate(levels):
                        ref.trigger_reset(w, lvl)
                ref.step()
        ref.close()
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **data)
        print(name, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
