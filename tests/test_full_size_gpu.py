"""Size-independent properties at BASELINE.json's full shard size (16000 worlds, 2+2 agents),
where the CPU oracle is too slow to diff every world:

* slicing invariance — the step graph cut into 1 slice on 1 branch and into 3 slices on 3
  parallel branches must produce BIT-identical exports (worlds are independent);
* sharding invariance — a shard created with world_base = 8000 reproduces worlds
  [8000, 16000) of the full shard bit for bit (the RNG is keyed by GLOBAL world index,
  src/sim.cpp:105-114), which is what makes N-GPU runs equal to one big run;
* episode bookkeeping — Seed = (episode, global world), done fires on step 239 for every row
  and the prep counter restarts (src/sim.cpp:172-200, 763-893).
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N, A = 16000, 4
EXPORTS = ("self_data", "agent_data", "box_data", "ramp_data", "visible_agents_mask",
           "visible_boxes_mask", "visible_ramps_mask", "lidar", "reward", "done", "prep_counter",
           "seed", "global_positions", "self_type", "self_mask")


def _make(num_worlds, world_base=0, slices=None, branches=None):
    import gpu_hideseek as gh
    env = {}
    if slices is not None:
        env["HIDESEEK_STEP_SLICES"] = str(slices)
    if branches is not None:
        env["HIDESEEK_STEP_BRANCHES"] = str(branches)
    os.environ.update(env)
    try:
        sim = gh.HideAndSeekSimulator(num_worlds=num_worlds, rand_seed=0, min_hiders=2, max_hiders=2,
                                      min_seekers=2, max_seekers=2, world_base=world_base)
    finally:
        for k in env:
            del os.environ[k]
    sim.init()
    return sim


def _tape(steps, rows):
    gen = torch.Generator(device="cpu").manual_seed(99)
    out = []
    for _ in range(steps):
        a = torch.zeros((rows, 5), dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 11, (rows, 3), generator=gen, dtype=torch.int32)
        a[:, 3] = (torch.rand(rows, generator=gen) < 0.2).int()
        a[:, 4] = (torch.rand(rows, generator=gen) < 0.1).int()
        out.append(a)
    return out


def _exports(sim):
    return {n: getattr(sim, n + "_tensor")().to_torch().clone().cpu().numpy() for n in EXPORTS}


def _bits(a):
    return a.view(np.uint32) if a.dtype.itemsize == 4 else a


def test_slicing_and_sharding_invariance():
    steps = 30
    tape = _tape(steps, N * A)
    one = _make(N, slices=1, branches=1)
    three = _make(N, slices=3, branches=3)
    upper = _make(N // 2, world_base=N // 2)          # second half as its own shard
    acts = [s.action_tensor().to_torch() for s in (one, three, upper)]
    for a in tape:
        acts[0].copy_(a.to(acts[0].device))
        acts[1].copy_(a.to(acts[1].device))
        acts[2].copy_(a[(N // 2) * A:].to(acts[2].device))
        one.step(); three.step(); upper.step()
    e1, e3, eu = _exports(one), _exports(three), _exports(upper)
    for n in EXPORTS:
        assert np.array_equal(_bits(e1[n]), _bits(e3[n])), f"slicing changed {n}"
        per_world = e1[n].shape[0] // N
        assert np.array_equal(_bits(e1[n][(N // 2) * per_world:]), _bits(eu[n])), f"sharding changed {n}"
    # Seed = {episode index, GLOBAL world index} per agent row (src/level_gen.cpp:23)
    seed = eu["seed"].reshape(N // 2, A, 2)
    assert np.array_equal(seed[:, :, 1], np.repeat(np.arange(N // 2, N)[:, None], A, 1))
    assert (seed[:, :, 0] == 0).all()
    assert one.overflow_count() == 0 and three.overflow_count() == 0


def test_episode_rollover_full_size():
    sim = _make(N)
    act = sim.action_tensor().to_torch()
    done = sim.done_tensor().to_torch()
    prep = sim.prep_counter_tensor().to_torch()
    seed = sim.seed_tensor().to_torch()
    gen = torch.Generator(device="cuda").manual_seed(5)
    done_steps = []
    for step in range(242):
        act[:, 0:2] = torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=act.device, generator=gen)
        sim.step()
        d = int(done.sum().item())
        assert d in (0, N * A), (step, d)       # worlds roll over in lockstep
        if d:
            done_steps.append(step)
    assert done_steps == [239], done_steps      # episodeLen = 240 (src/sim.cpp:14-17)
    # the auto-reset ran: episode index 1 everywhere, prep counter in lockstep again
    assert int((seed.view(N, A, 2)[:, :, 0] == 1).sum().item()) == N * A
    assert int(prep.min().item()) == int(prep.max().item())
    lidar = sim.lidar_tensor().to_torch()
    assert torch.isfinite(lidar).all() and float(lidar.min().item()) >= 0.0
    assert sim.overflow_count() == 0
