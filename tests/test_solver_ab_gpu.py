"""The staged lane-pair solver (k_solve: level schedule from the broadphase, bodies in shared
memory via one bulk-async copy per world, two lanes per contact) against the global-memory
solver it replaced (k_solve_global, HIDESEEK_SOLVER=global: own level schedule, one lane per
contact): same Gauss-Seidel order, same expressions, so every body state, contact multiplier and
export must be BIT-identical.  4096 worlds x 60 steps with grab / lock actions (joints, locked
boxes), 1-3 agents per team."""
import os

import numpy as np
import pytest
import torch

from compact import hs_dump_compact

pytestmark = pytest.mark.gpu

NAMES = ("self_data", "agent_data", "box_data", "ramp_data", "lidar", "reward", "visible_boxes_mask")


def _run(solver, N, steps, flags=0):
    import gpu_hideseek as gh
    os.environ["HIDESEEK_SOLVER"] = solver
    try:
        sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=7, sim_flags=flags, min_hiders=1, max_hiders=3,
                                      min_seekers=1, max_seekers=3)
    finally:
        del os.environ["HIDESEEK_SOLVER"]
    sim.init()
    act = sim.action_tensor().to_torch()
    gen = torch.Generator(device="cpu").manual_seed(21)
    for _ in range(steps):
        a = torch.zeros(act.shape, dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 11, (act.shape[0], 3), generator=gen, dtype=torch.int32)
        a[:, 3] = (torch.rand(act.shape[0], generator=gen) < 0.3).int()
        a[:, 4] = (torch.rand(act.shape[0], generator=gen) < 0.15).int()
        act.copy_(a.to(act.device))
        sim.step()
    ex = [getattr(sim, n + "_tensor")().to_torch().clone().cpu().numpy() for n in NAMES]
    return ex, hs_dump_compact(sim), sim.overflow_count()


@pytest.mark.parametrize("flags", [0, 8])
def test_staged_solver_equals_global_solver(flags):
    N, steps = 4096, 60
    ea, ca, oa = _run("global", N, steps, flags)
    eb, cb, ob = _run("staged", N, steps, flags)
    assert oa == 0 and ob == 0
    for n, a, b in zip(NAMES, ea, eb):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), n
    for f in ("pose", "vel", "contact_pts", "contact_normal", "contact_lambda", "leaf_aabb"):
        assert np.array_equal(ca[f].view(np.uint32), cb[f].view(np.uint32)), f
    for f in ("num_cand", "num_contacts", "cand", "contact_ref", "contact_alt", "contact_npts", "joint_other", "response"):
        assert np.array_equal(ca[f], cb[f]), f
    assert int((ca["joint_other"] >= 0).sum()) > 0          # grabs happened
    assert int(((ca["response"] == 2) & (ca["owner"] != 3)).sum()) > 0      # locks happened
