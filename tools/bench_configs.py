#!/usr/bin/env python
"""Throughput of the BASELINE.json configurations that are not the bench.py line (they are
parity-test cases; these numbers fill BASELINE.md §4).  One GPU, device-timed with CUDA events.

    python tools/bench_configs.py [--worlds 16000] > gpurun_out/configs.json
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))
import gpu_hideseek as gh  # noqa: E402


def timed(sim, steps, warmup, before_step):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.current_stream().cuda_stream
    for i in range(warmup):
        before_step(i)
        sim.step_async(stream)
    torch.cuda.synchronize()
    ev0.record()
    for i in range(steps):
        before_step(warmup + i)
        sim.step_async(stream)
    ev1.record()
    torch.cuda.synchronize()
    return ev0.elapsed_time(ev1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=16000)
    args = ap.parse_args()
    N = args.worlds
    dev = torch.device("cuda", 0)
    out = []

    def run(name, steps, kw, make_hook):
        sim = gh.HideAndSeekSimulator(num_worlds=N, **kw)
        sim.init()
        A = sim.agents_per_world
        hook = make_hook(sim, A)
        ms = timed(sim, steps, 5, hook)
        st = sim.world_stats()
        out.append({"config": name, "worlds": N, "agents_per_world": A, "steps": steps,
                    "ms_per_step": ms / steps, "env_steps_per_s": N * steps / (ms / 1000.0),
                    "mean_bodies": st["bodies"], "mean_candidates": st["candidates"],
                    "mean_contacts": st["contacts"], "overflow": sim.overflow_count()})
        del sim

    base = dict(rand_seed=0, sim_flags=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2)

    def moves(sim, A):
        act = sim.action_tensor().to_torch()
        return lambda i: act[:, 0:2].copy_(torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev))

    # config 2 at the reference's full step count (scripts/benchmark.py 16000 1920)
    run("benchmark.py 16000x1920, random move actions", 1920, base, moves)

    # config 4: reset-heavy (scripts/benchmark.py:77-80 reset_chance mechanism)
    def resets(p):
        def make(sim, A):
            mv = moves(sim, A)
            rst = sim.reset_tensor().to_torch()
            def hook(i):
                mv(i)
                rst.copy_((torch.rand(N, 1, device=dev) < p).int())
            return hook
        return make
    run("reset-heavy: 25% of worlds reset every step", 240, base, resets(0.25))

    def every4(sim, A):
        mv = moves(sim, A)
        rst = sim.reset_tensor().to_torch()
        def hook(i):
            mv(i)
            if i % 4 == 0:
                rst.fill_(1)
        return hook
    run("reset-heavy: every world reset every 4 steps", 240, base, every4)

    # config 5: jax_train.py flags, 3+3 agents, seed 5, random policy over all 5 action columns
    jaxkw = dict(rand_seed=5, sim_flags=1 | 4 | 8, min_hiders=3, max_hiders=3, min_seekers=3, max_seekers=3)

    def policy(sim, A):
        act = sim.action_tensor().to_torch()
        def hook(i):
            act[:, 0:3].copy_(torch.randint(0, 5, (N * A, 3), dtype=torch.int32, device=dev))
            act[:, 3:5].copy_(torch.randint(0, 2, (N * A, 2), dtype=torch.int32, device=dev))
        return hook
    run("jax_train.py flags (fixed world, flip teams, zero agent vel), 3+3 agents, random policy", 480, jaxkw, policy)

    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
