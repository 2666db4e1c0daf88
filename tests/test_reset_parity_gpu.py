"""Level generation / reset / BVH build: bit-exact vs the reference CPU backend."""
import numpy as np
import pytest

import refsim
from parity import compare_world_static, compare_bvh, f32bits

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
        errs = compare_world_static(rw, hw, exact=True) + compare_bvh(rw, hw, exact=True)
        if errs:
            bad.append((w, errs[:4]))
    assert not bad, bad[:3]
    rt = ref.tensors()
    for name, key in (("seed", "seed"), ("self_type", "self_type"), ("self_mask", "self_mask"),
                      ("action", "action"), ("lidar", "lidar"), ("self_data", "self_obs"),
                      ("visible_boxes_mask", "box_vis"), ("visible_ramps_mask", "ramp_vis"),
                      ("visible_agents_mask", "agent_vis"), ("prep_counter", "prep_counter"),
                      ("global_positions", "global_pos")):
        mine = getattr(sim, name + "_tensor")().to_torch().cpu().numpy()
        theirs = rt[key]
        if name in ("self_data",):
            np.testing.assert_allclose(mine, theirs, rtol=0, atol=2e-6, err_msg=name)
        elif mine.dtype == np.float32:
            assert np.array_equal(f32bits(mine), f32bits(theirs)), (name, np.abs(mine - theirs).max())
        else:
            assert np.array_equal(mine, theirs.reshape(mine.shape)), name
    assert sim.overflow_count() == 0
