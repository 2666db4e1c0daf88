"""The BVH update at the head of a step (setupBroadphaseTasks at the start of the reference's Step
graph, src/sim.cpp:1307-1313) is skipped by default: the previous step (or init / a checkpoint
load) ended with the same update on the same poses and velocities, so it cannot change a leaf box
or grow a node.  HIDESEEK_HEAD_REFIT=1 restores the launch; the two must agree bit for bit on the
whole world record — BVH nodes and leaves, candidate pairs, contacts, poses — over an episode, its
reset wave, forced resets and a checkpoint round trip.  Same harness for the other launch the step
graph drops: k_substep of substeps 2-4 (done at the tail of k_solve)."""
import os

import numpy as np
import pytest
import torch

from compact import hs_dump_compact, hs_set_debug

pytestmark = pytest.mark.gpu


KNOBS = ("HIDESEEK_HEAD_REFIT", "HIDESEEK_FUSED_SUBSTEP", "HIDESEEK_SOLVE_ORDER", "HIDESEEK_STAGE_CAP")


def _run(env, N=384, steps=250):
    import gpu_hideseek as gh
    for k in KNOBS:
        os.environ.pop(k, None)
    os.environ.update({k: v for k, v in env.items() if not k.startswith("_")})
    try:
        sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=11, min_hiders=1, max_hiders=3,
                                      min_seekers=1, max_seekers=3)
        A = 6
        sim.init()
        hs_set_debug(sim, 1)
        launches = None
        act = sim.action_tensor().to_torch()
        gen = torch.Generator(device="cpu").manual_seed(5)
        dumps = []
        for step in range(steps):
            a = torch.zeros((N * A, 5), dtype=torch.int32)
            a[:, 0:3] = torch.randint(0, 11, (N * A, 3), generator=gen, dtype=torch.int32)
            a[:, 3] = (torch.rand(N * A, generator=gen) < 0.3).int()
            a[:, 4] = (torch.rand(N * A, generator=gen) < 0.15).int()
            act.copy_(a.to(act.device))
            if step == 60:
                sim.trigger_reset(7, 1)
            if step == 100:
                for w in (3, 200):
                    sim.save_checkpoint(w)
            if step == 130:
                for w in (3, 200):
                    sim.load_checkpoint(w)
            if env.get("_PROFILED_EVERY") and step % int(env["_PROFILED_EVERY"]) == 2:
                sim.step_profiled()              # un-graphed launches of the same step, whole shard as one range
            else:
                sim.step()
            if launches is None:
                launches = sim.last_step_launches()
            if step % 20 == 0 or step in (61, 131, 239, 240, 241):
                dumps.append((step, hs_dump_compact(sim).tobytes()))
        assert sim.overflow_count() == 0
        return launches, dumps
    finally:
        for k in KNOBS:
            os.environ.pop(k, None)


def _same(with_launch, without):
    assert len(with_launch) == len(without)
    for (s1, a), (s0, b) in zip(with_launch, without):
        assert s1 == s0
        if a != b:
            ra = np.frombuffer(a, dtype=np.uint8).reshape(384, -1)
            rb = np.frombuffer(b, dtype=np.uint8).reshape(384, -1)
            bad = np.nonzero((ra != rb).any(axis=1))[0]
            raise AssertionError(f"step {s1}: {len(bad)} worlds differ, first {bad[:5]}")


def test_head_of_step_bvh_update_changes_nothing():
    l1, with_refit = _run({"HIDESEEK_HEAD_REFIT": "1"})
    l0, without = _run({})
    assert l1 > l0, (l1, l0)                     # the default really launches less
    _same(with_refit, without)


def test_substep_at_the_tail_of_the_solver_is_the_substep_kernel():
    """Substeps 2-4 of a step are integrated on the staged bodies at the tail of k_solve instead of
    by a k_substep launch (HIDESEEK_FUSED_SUBSTEP=0 launches it): same arithmetic on the same values,
    so the whole world record must be bit-identical."""
    l1, launched = _run({"HIDESEEK_FUSED_SUBSTEP": "0"})
    l0, fused = _run({})
    assert l1 > l0 and (l1 - l0) % 3 == 0, (l1, l0)          # three launches per world slice
    _same(launched, fused)


def test_solver_world_order_is_only_a_schedule():
    """k_solve hands its worlds out grouped by the rounds their contact sweeps took in the previous
    substep (HIDESEEK_SOLVE_ORDER=0: index order).  Worlds are independent, so which warp solves a
    world must not show in any bit of the result."""
    l1, plain = _run({"HIDESEEK_SOLVE_ORDER": "0"})
    l0, ordered = _run({})
    assert l1 == l0
    _same(plain, ordered)


def test_global_memory_fallback_of_the_staged_solver():
    """A world whose movable bodies do not fit the solver's shared-memory body area is solved (and,
    for substeps 2-4, integrated) out of global memory inside the same kernel.  Generated levels
    always fit, so the path is forced here: with room for 10 bodies most worlds fall back, a few
    stay staged, and every world record must still match the default build's bit for bit."""
    l1, forced = _run({"HIDESEEK_STAGE_CAP": "10"})
    l0, default = _run({})
    assert l1 == l0
    _same(forced, default)


def test_profiled_steps_interleave_with_graph_steps():
    """hs_step_profiled (what bench.py times kernels with) runs the step as plain launches over the
    whole shard, the step graph as world slices; both use the solver's world-order buffers, which are
    cleared whenever the view changes.  Mixing the two must not change a bit."""
    _, mixed = _run({"_PROFILED_EVERY": "5"})
    _, graph = _run({})
    _same(mixed, graph)
