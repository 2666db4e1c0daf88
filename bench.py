#!/usr/bin/env python
"""Benchmark: env-steps/s (worlds x steps) of the hide-and-seek batch simulator.

Workload = BASELINE.json configs[1]: scripts/benchmark.py semantics — 16000 worlds
per GPU, 2 hiders + 2 seekers, sim_flags 0, seed 0, renderer off (graphics path
unavailable on this box), every step x,y actions <- randint(-5,5) on the device,
then sim.step().  Episodes auto-reset every 240 steps.

  python bench.py [--gpus N --steps K --warmup W]        our CUDA path
  python bench.py --impl reference [...]                 the reference's own CPU backend
                                                         (oracle/_ref) on the host cores

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))

WORLDS_PER_GPU = 16000
AGENTS = 4
METRIC = "env steps/sec (worlds x steps) at 16K worlds/GPU, random actions"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=480)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--worlds", type=int, default=WORLDS_PER_GPU, help="worlds per GPU")
    ap.add_argument("--profile-steps", type=int, default=24)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--views", action="store_true",
                    help="also measure the agent camera views (enable_batch_renderer=True, 64x64) and add a "
                         "`views` object; off by default: the headline configuration has the renderer off")
    return ap.parse_args()


_JSON_OUT = None


def guard_stdout():
    """Keep stdout for the ONE JSON line: libraries that print to fd 1 (NCCL's version banner,
    the reference backend's log lines) are sent to stderr instead."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi SM clock / throttle-reason samples taken DURING the timed region."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.max_mhz = gpu, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=3)
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---- algorithmic bytes (DESIGN.md "Kernels and rooflines") ---------------------------
def step_alg_bytes(A, E):
    """bytes one env-step must move if everything derived stays on chip."""
    return A * (20 + 4 + 1260) + (4 + 136 + 8) + E * (2 * 52 + 24)


def kernel_alg_bytes(name, A, E, C, K):
    """Per-world ALGORITHMIC bytes of one launch of `name`, by SURVEY.md §8(d)'s rule: only
    persistent state counts — per live body 52 B mutable (pose 28 + velocity 24) and 24 B
    immutable (scale 12, object id / response / owner 12), per agent the action row and the
    exported rows — each touched once per launch.  Everything derived (BVH nodes and leaf
    records, candidate pairs, contact rows, previous / pre-solve scratch, forces, hull tables)
    counts 0: an ideal kernel holds it on chip."""
    pose, vel = 28, 24
    table = {
        "k_move": A * (20 + pose),                       # action row + agent pose in (force is derived)
        "k_leaf_refit": E * (pose + 12 + 12 + 4),        # pose, linear velocity, scale, object id in
        "k_bvh_rebuild": 0,
        "k_action": A * (20 + 20),                       # action row in, consumed action row out
        "k_broadphase": E * 4,                           # response / entity id word in
        "k_substep": E * (pose + vel + 4 + pose),        # pose + velocity + meta in, pose out
        "k_narrowphase": E * (pose + 12 + 4),            # pose, scale, object id in (contacts are derived)
        "k_solve": E * (pose + 4 + pose + vel),          # pose + meta in, pose + velocity out
        "k_solve_global": E * (pose + 4 + pose + vel),
        "k_solve_positions": E * (pose + 4 + pose),
        "k_set_velocities": E * (pose + vel),
        "k_solve_velocities": E * (vel + 4 + vel),
        "k_zero_agent_vel": A * 2 * vel,
        "k_rewards": A * (pose + 8),                     # agent pose in, reward + done out
        "k_reset": 4,
        "k_reset_worlds": 0,
        "k_observations": A * (13 + 70 + 153 + 28) * 4 + E * (pose + vel),
        "k_rays": A * (pose + 120 + 64),                 # agent pose in, 30 lidar floats + 16 visibility flags out
        "k_debug_positions": 136 + 17 * 8,
    }
    return float(table.get(name, 0))


def aggregate_value(n_per_rank, world, steps, ms_total):
    """Whole-job env-steps/s: every rank steps n_per_rank worlds (weak scaling, no data-path
    collective); ms_total is the max over ranks of the device-timed region."""
    return n_per_rank * world * steps / (ms_total / 1000.0)


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    `ncu --set full` capture (profiles/traffic.json, written by tools/ncu_digest.py --traffic)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t["kernels"].get(kernel), t.get("source")
    except Exception:
        return None, None


def kernel_rooflines(kernels, peak):
    """Adds the measured-traffic view to every kernel entry: `traffic` (ncu DRAM bytes per launch),
    `frac` = algorithmic bytes / time / peak, `dram_frac` = traffic / time / peak.  Returns the
    kernels whose algorithmic figure exceeds the DRAM bytes ncu saw inside the launch."""
    over = []
    for k in kernels:
        t, _ = measured_traffic(k["kernel"])
        k["traffic"] = t
        k["frac"] = k["achieved_gbs"] / peak
        k["dram_frac"] = (t / 1e9) / (k["avg_ms"] / 1e3) / peak if t and k["avg_ms"] > 0 else None
        if t and k["alg_bytes_per_launch"] > t * 1.02:
            over.append(f"{k['kernel']}: algorithmic {k['alg_bytes_per_launch']:.3g} B > DRAM bytes inside the launch "
                        f"{t:.3g} B (stores still in the write-back L2 when the kernel ends)")
    return over


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def cpu_reference_run(worlds, warmup, steps, threads):
    """Times the UNMODIFIED reference CPU backend (oracle/_ref) on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refsim
    ref = refsim.RefSim(worlds, sim_flags=0, seed=0, min_hiders=2, max_hiders=2, min_seekers=2,
                        max_seekers=2, num_threads=threads)
    ref.init()
    secs = ref.bench(warmup, steps, 0)
    cores = threads if threads > 0 else len(os.sched_getaffinity(0))
    ref.close()
    return worlds * steps / secs, secs, cores


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refsim
    if not refsim.available():
        # oracle/_ref is a build product that travels with the repo snapshot; without it (and
        # without /root/reference to build it from) there is nothing to time
        emit({"impl": "reference", "unavailable": "oracle/_ref/libhs_ref.so is not built on this machine"})
        return
    # the full 16000-world shard when the step budget allows, else fewer worlds (the CPU
    # backend runs one world per job, so throughput is flat in world count once worlds >> cores)
    worlds = args.worlds
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    t0 = time.time()
    # keep the whole run within a few minutes: scale the per-call step count down
    per_step_budget = 120.0 / max(1, steps + warmup)
    probe_value, probe_secs, cores = cpu_reference_run(worlds, 1, 3, 0)
    est_step = probe_secs / 3
    if est_step > per_step_budget:
        worlds = max(64, int(worlds * per_step_budget / est_step))
    value, secs, cores = cpu_reference_run(worlds, warmup, steps, 0)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1000.0 * secs / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "scripts/benchmark.py 16000 worlds x steps, random move actions "
                               "(reference CPU backend, bounded sample)",
                   "worlds_sampled": worlds, "agents_per_world": AGENTS},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": f"{worlds} worlds x {steps} steps, all host threads, "
                                   "oracle/_ref (reference sources built with g++ -O2)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.time() - t0,
    }
    emit(line)


def measure_views(args, gh, torch, dev, local, peak):
    """k_agent_views at scripts/benchmark.py's renderer settings (64x64 per agent), device-timed,
    next to its CPU restatement (oracle/views_oracle.c, the cpu_baseline leg) on a few views."""
    import ctypes as C
    import time
    N, A, S, launches = args.worlds, AGENTS, 64, 10
    sim = gh.HideAndSeekSimulator(gpu_id=local, num_worlds=N, sim_flags=0, rand_seed=0, min_hiders=2, max_hiders=2,
                                  min_seekers=2, max_seekers=2, enable_batch_renderer=True,
                                  batch_render_width=S, batch_render_height=S)
    sim.init()
    lib = gh.load_library()
    lib.hs_render_views.argtypes = [C.c_void_p, C.c_void_p]
    lib.hs_step.argtypes = [C.c_void_p, C.c_void_p]
    h = C.c_void_p(sim.handle())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    act = sim.action_tensor().to_torch()
    for _ in range(20):                     # let the agents leave their spawn poses
        act[:, 0:2] = torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev)
        assert lib.hs_step(h, stream) == 0
    for _ in range(3):
        assert lib.hs_render_views(h, stream) == 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(launches):
        assert lib.hs_render_views(h, stream) == 0
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / launches
    E = sim.world_stats()["bodies"]
    rays = N * A * S * S
    alg = rays * 8 + N * A * E * 52          # f32 depth + RGBA8 written per pixel, body records read per view
    out = {"kernel": "k_agent_views", "view": [S, S], "views_per_launch": N * A, "launches": launches,
           "ms_per_launch": ms, "rays_per_s": rays / (ms / 1e3), "meshes": "built-in",
           "roofline": {"bound": "hbm", "achieved": alg / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg / (ms / 1e3) / 1e9 / peak, "alg_bytes_per_launch": alg,
                        "traffic": measured_traffic("k_agent_views")[0],
                        "traffic_source": measured_traffic("k_agent_views")[1]}}
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import views_util as vu
        meshes = vu.Meshes(gh.view_mesh_table())
        depth = sim.depth_tensor().to_torch()
        t0, done, maxd = time.perf_counter(), 0, 0.0
        for w in range(2):
            dump = sim.dump_world(w)
            for k in range(A):
                rd, _ = vu.render_world_view(meshes, dump, k, S, S)
                maxd = max(maxd, float(abs(depth[w * A + k, :, :, 0].cpu().numpy() - rd).max()))
                done += 1
        secs = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": done * S * S / secs, "unit": "rays/s", "cores": 1, "kind": "port",
                               "sample": f"{done} views of {S}x{S} ({secs:.1f} s), oracle/views_oracle.c: brute force "
                                         f"over every triangle, no acceleration structure; max |depth diff| to the "
                                         f"kernel on them {maxd:.2g}"}
    except Exception as e:      # the oracle is a checker, never a dependency of the product arm
        out["cpu_baseline"] = {"value": None, "unit": "rays/s", "cores": 0, "kind": "port", "sample": f"unavailable: {e}"}
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    import gpu_hideseek as gh

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    N, A = args.worlds, AGENTS
    sim = gh.HideAndSeekSimulator(gpu_id=local, num_worlds=N, sim_flags=0, rand_seed=0, min_hiders=2,
                                  max_hiders=2, min_seekers=2, max_seekers=2, world_base=rank * N)
    sim.init()
    actions = sim.action_tensor().to_torch()
    reward = sim.reward_tensor().to_torch()
    torch.manual_seed(rank)
    stream = torch.cuda.current_stream().cuda_stream

    def one_step():
        # scripts/benchmark.py:82-88
        actions[:, 0:2] = torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev)
        sim.step_async(stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        one_step()
    barrier()

    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        one_step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    value = aggregate_value(N, world, args.steps, ms_total)

    # ---- end to end through the C ABI with HOST buffers (pinned), copies timed ------------
    R = N * A
    h_act = torch.zeros((R, 5), dtype=torch.int32).pin_memory()
    h_act[:, 2] = 5
    h_rew = torch.zeros((R,), dtype=torch.float32).pin_memory()
    h_done = torch.zeros((R,), dtype=torch.int32).pin_memory()
    e2e_steps = max(16, min(args.steps, 128))
    gen = torch.Generator().manual_seed(1 + rank)
    tape = torch.randint(-5, 5, (8, R, 2), dtype=torch.int32, generator=gen)
    for i in range(3):
        h_act[:, 0:2] = tape[i % 8]
        sim.step_host(h_act.data_ptr(), h_rew.data_ptr(), h_done.data_ptr(), stream)
        torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        h_act[:, 0:2] = tape[i % 8]          # the caller's per-step host write
        sim.step_host(h_act.data_ptr(), h_rew.data_ptr(), h_done.data_ptr(), stream)
        torch.cuda.synchronize()             # result (reward, done) is on the host
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = N * world * e2e_steps / float(e2e_s.item())
    checksum = float(h_rew.sum())

    # ---- end to end with EVERY output tensor copied to the host (pinned), PCIe-bound ------------
    out_names = ("prep_counter", "self_data", "self_type", "self_mask", "lidar", "agent_data", "box_data",
                 "ramp_data", "visible_agents_mask", "visible_boxes_mask", "visible_ramps_mask", "reward", "done")
    outs = [getattr(sim, n + "_tensor")().to_torch() for n in out_names]
    h_outs = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in outs]
    d_act = sim.action_tensor().to_torch()
    obs_bytes = sum(t.numel() * t.element_size() for t in outs)
    obs_steps = 24
    for i in range(obs_steps + 2):
        if i == 2:
            barrier()
            t0 = time.perf_counter()
        h_act[:, 0:2] = tape[i % 8]
        d_act.copy_(h_act, non_blocking=True)
        sim.step_async(stream)
        for src, dst in zip(outs, h_outs):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
    obs_s = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(obs_s, op=dist.ReduceOp.MAX)
    e2e_obs_value = N * world * obs_steps / float(obs_s.item())

    # ---- one FULL EPISODE, device-timed: 240 steps from wherever the run stands, so the window
    # always holds one synchronized reset wave and the whole contact build-up (the `--steps` region
    # above starts at a fresh init(): with few steps it only sees the cheap prep phase) -----------
    _ = float(reward.abs().sum())            # load the reduction kernel outside the timed window
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    # reward checksum from two steps 119 apart: at most one of them lies in the 96-step
    # preparation phase (rewards 0), whatever phase the episode starts in
    rew_sum = torch.zeros((), device=dev, dtype=torch.float64)
    for i in range(240):
        one_step()
        if i in (120, 239):
            rew_sum += reward.abs().sum()
    f1.record()
    barrier()
    fms = torch.tensor([f0.elapsed_time(f1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(fms, op=dist.ReduceOp.MAX)
    full_ms = float(fms.item())
    full_value = aggregate_value(N, world, 240, full_ms)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-kernel device times (CUDA events between launches, same stream): `profile_steps`
    # un-graphed steps spread evenly over one more episode, world statistics sampled with them ----
    totals, counts = {}, {}
    every = max(1, 240 // max(1, args.profile_steps))
    stat_sum, samples = {"bodies": 0.0, "candidates": 0.0, "contacts": 0.0}, 0
    for i in range(240):
        actions[:, 0:2] = torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev)
        if i % every == every // 2 and samples < args.profile_steps:
            for name, kms in sim.step_profiled(stream):
                totals[name] = totals.get(name, 0.0) + kms
                counts[name] = counts.get(name, 0) + 1
            st = sim.world_stats()
            for k in stat_sum:
                stat_sum[k] += st[k]
            samples += 1
        else:
            sim.step_async(stream)
    torch.cuda.synchronize()
    E, C, K = (stat_sum[k] / max(1, samples) for k in ("bodies", "candidates", "contacts"))
    peak, peak_src = measured_peak_gbs()
    step_ms_prof = sum(totals.values()) / max(1, samples)
    kernels = []
    for name in sorted(totals, key=lambda n: -totals[n]):
        avg_ms = totals[name] / counts[name]
        algb = kernel_alg_bytes(name, A, E, C, K) * N
        kernels.append({"kernel": name, "launches_per_step": counts[name] / max(1, samples),
                        "avg_ms": avg_ms, "share": totals[name] / sum(totals.values()),
                        "alg_bytes_per_launch": algb,
                        "achieved_gbs": (algb / 1e9) / (avg_ms / 1e3) if avg_ms > 0 else 0.0})
    violations = kernel_rooflines(kernels, peak)
    top = kernels[0]
    traffic, traffic_src = measured_traffic(top["kernel"])
    per_gpu = full_value / world
    roofline = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved_gbs"], "peak": peak,
                "unit": "GB/s", "frac": top["achieved_gbs"] / peak, "traffic": traffic,
                "dram_frac": top["dram_frac"], "traffic_source": traffic_src,
                "peak_source": peak_src, "avg_launch_ms": top["avg_ms"],
                "share_of_step": top["share"],
                "alg_bytes_per_launch": top["alg_bytes_per_launch"],
                "alg_gt_traffic": violations,
                "whole_step": {"alg_bytes_per_env_step": step_alg_bytes(A, E),
                               "achieved": per_gpu * step_alg_bytes(A, E) / 1e9,
                               "frac": per_gpu * step_alg_bytes(A, E) / 1e9 / peak,
                               "from": "full_episode"},
                "how": f"CUDA events after every launch of {samples} un-graphed steps spread over one "
                       f"episode (sum {step_ms_prof:.3f} ms/step); algorithmic bytes by SURVEY.md §8(d) "
                       "(derived data counts 0)"}
    full_episode = {"steps": 240, "value": full_value, "unit": UNIT, "ms_per_step": full_ms / 240,
                    "mean_bodies": E, "mean_candidates": C, "mean_contacts": K,
                    "reward_checksum": float(rew_sum.item()),
                    "how": "240 device-timed steps after the --steps region: one whole episode incl. its reset wave"}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        # same shard size; step count chosen for ~12 s of CPU work from a 2-step probe
        cw = N
        try:
            _, probe_secs, _ = cpu_reference_run(cw, 1, 2, 0)
            cs = int(min(96, max(8, 12.0 / max(probe_secs / 2, 1e-3))))
            v, secs, cores = cpu_reference_run(cw, 2, cs, 0)
            cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference",
                            "sample": f"{cw} worlds x {cs} steps ({secs:.1f} s), all host threads, "
                                      "oracle/_ref = reference CPU backend built with g++ -O2"}
        except Exception as e:      # the oracle is a checker, never a dependency of the product arm
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                            "sample": f"unavailable: {e}"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "scripts/benchmark.py semantics: 16000 worlds/GPU x steps, 2 hiders + 2 "
                               "seekers, random move actions every step, renderer off",
                   "worlds_per_gpu": N, "agents_per_world": A, "sharding": f"worlds x{world}",
                   "l2": "per-GPU state (BVH 208 MB + body records 186 MB + contacts 164 MB + exports 81 MB) exceeds the 126 MB L2",
                   "mean_bodies": E, "mean_candidates": C, "mean_contacts": K,
                   "note": "`value` times exactly --steps steps from a fresh init(); `full_episode` is the "
                           "same metric over one whole 240-step episode (reset wave included)"},
        # bytes are whole-job totals like `value` (every rank copies its own shard's tensors)
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": R * 5 * 4 * world,
                "d2h_bytes_per_step": R * 8 * world, "steps": e2e_steps,
                "how": "hs_step_host: pinned host actions -> H2D, step graph, reward+done D2H, "
                       "stream sync every step; observations stay on the device (a GPU-resident policy "
                       "reads them through DLPack)", "reward_checksum": checksum},
        "e2e_obs_to_host": {"value": e2e_obs_value, "unit": UNIT, "h2d_bytes_per_step": R * 5 * 4 * world,
                            "d2h_bytes_per_step": obs_bytes * world, "steps": obs_steps,
                            "how": "pinned host actions -> H2D, step graph, all 13 output tensors D2H into pinned "
                                   "host buffers, sync every step (PCIe-bound)"},
        "full_episode": full_episode,
        "gpu_launches": sim.last_step_launches() * args.steps * world,     # whole job
        "gpu_launches_per_rank": sim.last_step_launches() * args.steps,
        "launches_per_step": sim.last_step_launches(),
        "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "overflow": sim.overflow_count(),
    }
    if cpu_baseline:
        line["cpu_baseline"] = cpu_baseline
    if args.views and world == 1:
        del sim
        line["views"] = measure_views(args, gh, torch, dev, local, measured_peak_gbs()[0])
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    guard_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
