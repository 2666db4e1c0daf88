#!/usr/bin/env python
"""Agent camera views (enable_batch_renderer=True): cost of k_agent_views next to the step, at
scripts/benchmark.py's settings (64x64 views, 2 hiders + 2 seekers).  One GPU, CUDA events.

    python tools/bench_views.py [--worlds 16000] [--size 64] [--data-dir DIR] > gpurun_out/views.json

Roofline: the kernel must write 8 bytes per pixel (f32 depth + RGBA8) and read one world's
body records per view; those algorithmic bytes over the measured time are `achieved_gbs`.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))
import gpu_hideseek as gh  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=16000)
    ap.add_argument("--size", type=int, default=64)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--data-dir", default=None)
    args = ap.parse_args()
    N, S = args.worlds, args.size
    dev = torch.device("cuda", 0)
    sim = gh.HideAndSeekSimulator(num_worlds=N, enable_batch_renderer=True, batch_render_width=S,
                                  batch_render_height=S, render_data_dir=args.data_dir)
    sim.init()
    A = sim.agents_per_world
    act = sim.action_tensor().to_torch()
    stream = torch.cuda.current_stream().cuda_stream
    lib = gh.load_library()
    import ctypes as C
    lib.hs_render_views.argtypes = [C.c_void_p, C.c_void_p]
    lib.hs_step.argtypes = [C.c_void_p, C.c_void_p]
    h = C.c_void_p(sim.handle())

    def move():
        act[:, 0:2].copy_(torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev))

    for _ in range(20):          # let the agents spread out of the spawn poses
        move()
        assert lib.hs_step(h, C.c_void_p(stream)) == 0
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_step = t_view = 0.0
    for _ in range(args.steps):
        move()
        e[0].record()
        assert lib.hs_step(h, C.c_void_p(stream)) == 0
        e[1].record()
        assert lib.hs_render_views(h, C.c_void_p(stream)) == 0
        e[2].record()
        torch.cuda.synchronize()
        t_step += e[0].elapsed_time(e[1])
        t_view += e[1].elapsed_time(e[2])
    ms_step, ms_view = t_step / args.steps, t_view / args.steps
    st = sim.world_stats()
    depth = sim.depth_tensor().to_torch()
    rays = N * A * S * S
    alg_bytes = rays * 8 + N * A * st["bodies"] * (32 + 16 + 4)
    print(json.dumps({
        "kernel": "k_agent_views", "worlds": N, "agents_per_world": A, "view": [S, S],
        "meshes": "reference OBJ" if args.data_dir else "built-in",
        "ms_step_without_views": ms_step, "ms_views": ms_view,
        "env_steps_per_s_with_views": N / ((ms_step + ms_view) / 1000.0),
        "rays_per_s": rays / (ms_view / 1000.0),
        "alg_bytes_per_launch": alg_bytes, "achieved_gbs": alg_bytes / (ms_view / 1000.0) / 1e9,
        "hit_fraction": float((depth > 0).float().mean()), "mean_bodies": st["bodies"],
    }))


if __name__ == "__main__":
    main()
