"""Whole-shard diff of the CUDA path against the reference CPU backend (oracle/_ref): both sides
run the same configuration and action tape; at the listed steps every world's compact record
(tests/compact.py: body state, bookkeeping, BVH, candidate pairs, last-substep contacts, lidar
hit bodies) and every exported tensor are compared.  TEST INFRASTRUCTURE ONLY.

  python tests/full_parity.py [--worlds 16000 --steps 245 --seed 0 --check 0,1,96,200,239,240,245]
"""
import argparse
import json
import sys
import time

import numpy as np

POS_TOL = 1e-4      # body state / BVH bounds / contact rows (the north star's single-step float bar)
OBS_TOL = 5e-4      # observations, lidar depth


def benchmark_tape(rows, step, seed):
    """scripts/benchmark.py:82-84: x, y <- randint(-5, 5) every step; r/g/l stay consumed."""
    rng = np.random.RandomState((seed * 1000003 + step) & 0x7FFFFFFF)
    a = np.zeros((rows, 5), dtype=np.int32)
    a[:, 0:2] = rng.randint(-5, 5, size=(rows, 2))
    a[:, 2] = 2
    return a


def grablock_tape(rows, step, seed):
    rng = np.random.RandomState((seed * 1000003 + step) & 0x7FFFFFFF)
    a = np.zeros((rows, 5), dtype=np.int32)
    a[:, 0:3] = rng.randint(0, 11, size=(rows, 3))
    a[:, 3] = rng.rand(rows) < 0.3
    a[:, 4] = rng.rand(rows) < 0.15
    return a


def run(N, steps, seed, check, hs=(2, 2), sk=(2, 2), flags=0, tape=benchmark_tape, chunk=4000,
        threads=0, log=print):
    import torch
    import gpu_hideseek as gh
    import refsim
    from compact import (ref_step_probe, ref_dump_compact, hs_dump_compact, hs_set_debug, compare_compact)
    from parity import compare_exports

    A = hs[1] + sk[1]
    ref = refsim.RefSim(N, sim_flags=flags, seed=seed, min_hiders=hs[0], max_hiders=hs[1],
                        min_seekers=sk[0], max_seekers=sk[1], num_threads=threads)
    ref.init()
    sim = gh.HideAndSeekSimulator(num_worlds=N, sim_flags=flags, rand_seed=seed, min_hiders=hs[0],
                                  max_hiders=hs[1], min_seekers=sk[0], max_seekers=sk[1])
    hs_set_debug(sim, 1)
    sim.init()
    rt = ref.tensors()
    act = sim.action_tensor().to_torch()
    check = sorted(set(check))
    summary = {"worlds": N, "steps": steps, "seed": seed, "agents": A, "checked_steps": [], "errors": []}
    agg = {}

    def diff(step, probes, prev_active):
        t0 = time.time()
        errs, st_all = [], {}
        for w0 in range(0, N, chunk):
            n = min(chunk, N - w0)
            r = ref_dump_compact(ref, w0, n, threads)
            h = hs_dump_compact(sim, w0, n)
            e, st = compare_compact(r, h, tol=POS_TOL, exact=False, probes=probes, w0=w0,
                                    check_scores=step > 0)
            errs += e
            for k, v in st.items():
                if k.startswith("max_"):
                    st_all[k] = max(st_all.get(k, 0), v)
                elif k.startswith("mean_"):
                    st_all[k] = st_all.get(k, 0.0) + v * n / N
                else:
                    st_all[k] = st_all.get(k, 0) + v
            walls = (h["obj"] == 3) & (np.arange(52)[None, :] < h["num_bodies"][:, None])
            st_all["max_walls"] = max(st_all.get("max_walls", 0), int(walls.sum(axis=1).max()))
            st_all["worlds_ge30_walls"] = st_all.get("worlds_ge30_walls", 0) + int((walls.sum(axis=1) >= 30).sum())
            st_all["max_bodies"] = max(st_all.get("max_bodies", 0), int(h["num_bodies"].max()))
            st_all["worlds_with_joints"] = st_all.get("worlds_with_joints", 0) + int((h["joint_other"] >= 0).any(axis=1).sum())
            st_all["locked_bodies"] = st_all.get("locked_bodies", 0) + int(((h["response"] == 2) & (h["owner"] != 3) & (np.arange(52)[None, :] < h["num_bodies"][:, None])).sum())
        # reward / done / episode result are first written by a step (src/sim.cpp:805-893): after
        # init() the reference rows hold uninitialised memory
        skip = ("reward", "done", "episode_result") if step == 0 else ()
        e, maxd = compare_exports(sim, rt, hs[0] + sk[0], tol=OBS_TOL, prev_active=prev_active, skip=skip)
        errs += e
        st_all["max_export_diff"] = maxd
        for k, v in st_all.items():
            if k.startswith("max_"):
                agg[k] = max(agg.get(k, 0), v)
        summary["checked_steps"].append({"step": step, "stats": st_all, "errors": errs})
        summary["errors"] += [f"step {step}: {x}" for x in errs]
        log(f"step {step}: {len(errs)} mismatching fields, {time.time() - t0:.1f} s, "
            + ", ".join(f"{k}={v:.3g}" if isinstance(v, float) else f"{k}={v}" for k, v in sorted(st_all.items())))

    if 0 in check:
        diff(0, False, None)
    t_ref = t_gpu = 0.0
    for step in range(1, steps + 1):
        a = tape(N * A, step, seed)
        rt["action"][:] = a
        act.copy_(torch.from_numpy(a).to(act.device))
        prev_active = rt["self_mask"].reshape(-1) == 1.0
        t0 = time.time()
        if step in check:
            ref_step_probe(ref, threads)
        else:
            ref.step()
        t_ref += time.time() - t0
        t0 = time.time()
        sim.step()
        t_gpu += time.time() - t0
        if step in check:
            diff(step, True, prev_active)
    summary["max"] = agg
    summary["overflow"] = sim.overflow_count()
    summary["seconds_reference"] = t_ref
    summary["seconds_cuda"] = t_gpu
    ref.close()
    return summary


if __name__ == "__main__":
    import os
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=16000)
    ap.add_argument("--steps", type=int, default=245)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--check", default="0,1,96,200,239,240,245")
    ap.add_argument("--ragged", action="store_true", help="0-3 hiders / seekers per world, grab + lock actions")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    kw = dict(hs=(0, 3), sk=(0, 3), tape=grablock_tape) if a.ragged else {}
    s = run(a.worlds, a.steps, a.seed, [int(x) for x in a.check.split(",")], **kw)
    print(json.dumps({k: v for k, v in s.items() if k != "checked_steps"}, indent=1))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(s, f, indent=1)
    sys.exit(1 if s["errors"] else 0)
