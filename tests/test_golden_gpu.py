"""CUDA path vs the committed golden fixtures (generated from the reference CPU backend by
tests/golden/make_golden.py).  Nothing here touches oracle/ or /root/reference, so this is
the parity check that still runs where neither exists."""
import numpy as np
import pytest
import torch

from golden_util import Golden, case_names
from parity import compare_world_static, compare_bvh, compare_exports

pytestmark = pytest.mark.gpu

POS_TOL = 1e-4      # abs; positions / quaternions / velocities / lidar depth / observations


@pytest.mark.parametrize("name", case_names())
def test_cuda_matches_golden(name):
    import gpu_hideseek as gh
    g = Golden(name)
    sim = gh.HideAndSeekSimulator(num_worlds=g.N, rand_seed=g.seed, **g.kwargs())
    sim.init()
    act = sim.action_tensor().to_torch()
    maxd = 0.0
    for s in range(g.steps + 1):
        if s in g.snaps:
            gw = g.worlds(s)
            for w in range(g.N):
                hw = sim.dump_world(w)
                errs = compare_world_static(gw[w], hw, exact=(s == 0), tol=POS_TOL, check_scores=(s > 0))
                errs += compare_bvh(gw[w], hw, exact=(s == 0), tol=POS_TOL)
                assert not errs, (name, s, w, errs[:4])
            skip = ("reward", "done", "episode_result") if s == 0 else ()
            errs, d = compare_exports(sim, g.exports(s), g.min_h + g.min_s, tol=5e-4, skip=skip)
            assert not errs, (name, s, errs)
            maxd = max(maxd, d)
        if s < g.steps:
            act.copy_(torch.from_numpy(g.tape[s]).to(act.device))
            if s == 1 and g.levels is not None:
                for w, lvl in enumerate(g.levels):
                    sim.trigger_reset(w, lvl)
            sim.step()
    assert sim.overflow_count() == 0
    print(f"{name}: max |export diff| vs golden = {maxd:g}")
