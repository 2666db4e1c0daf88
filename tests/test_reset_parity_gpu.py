"""Level generation / reset / BVH build: bit-exact vs the reference CPU backend."""
import numpy as np
import pytest

import refsim
from parity import compare_world_static, compare_bvh, compare_exports

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,flags,hs,sk", [(0, 0, (2, 2), (2, 2)), (5, 0, (1, 3), (1, 3)),
                                              (7, 4, (1, 3), (1, 2))])
def test_init_bit_exact(seed, flags, hs, sk):
    import gpu_hideseek as gh
    N = 64
    ref = refsim.RefSim(N, sim_flags=flags, seed=seed, min_hiders=hs[0], max_hiders=hs[1],
                        min_seekers=sk[0], max_seekers=sk[1], num_threads=1)
    ref.init()
    sim = gh.HideAndSeekSimulator(num_worlds=N, sim_flags=flags, rand_seed=seed, min_hiders=hs[0],
                                  max_hiders=hs[1], min_seekers=sk[0], max_seekers=sk[1])
    sim.init()
    bad = []
    for w in range(N):
        rw, hw = ref.dump_world(w), sim.dump_world(w)
        errs = compare_world_static(rw, hw, exact=True, check_scores=False) + compare_bvh(rw, hw, exact=True)
        if errs:
            bad.append((w, errs[:4]))
    assert not bad, bad[:3]
    errs, maxd = compare_exports(sim, ref.tensors(), hs[0] + sk[0], tol=2e-6,
                                 skip=("reward", "done", "episode_result"))
    assert not errs, errs
    assert sim.overflow_count() == 0
