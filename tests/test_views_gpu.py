"""Agent camera views (depth_tensor / rgb_tensor with enable_batch_renderer=True): the CUDA
ray caster k_agent_views against the brute-force CPU restatement oracle/views_oracle.c on the
same world state.  Parity is UNPINNED for this component (the reference's ray caster is
device-only code that cannot run here) — see the oracle's header.

Bars: depth within 1e-4 * max(1, depth), colour bytes identical, on all but <= 0.2 % of a
view's pixels (silhouette pixels, where an ulp of difference in the transformed ray flips
hit / miss or the instance hit)."""
import numpy as np
import pytest
import torch

import views_util as vu

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not vu.available(), reason="oracle/_ref/libviews_oracle.so not built")]

DEPTH_TOL = 1e-4
FLIP_FRAC = 2e-3


def make(N, data_dir=None, size=64, **kw):
    import gpu_hideseek as gh
    sim = gh.HideAndSeekSimulator(num_worlds=N, enable_batch_renderer=True, batch_render_width=size,
                                  batch_render_height=size, render_data_dir=data_dir, **kw)
    sim.init()
    return sim, gh.view_mesh_table(data_dir), (gh.view_tex_table(data_dir) if data_dir else None)


def drive(sim, steps, seed=0, A=4):
    act = sim.action_tensor().to_torch()
    g = torch.Generator(device="cpu").manual_seed(seed)
    for _ in range(steps):
        a = torch.zeros(act.shape, dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 11, (act.shape[0], 3), generator=g, dtype=torch.int32)
        act.copy_(a.to(act.device))
        sim.step()


def check_views(sim, table, worlds, A, size=64, height=None, tex_table=None, data_dir=None):
    height = height or size
    meshes = vu.Meshes(table, tex_table, data_dir)
    rgb_tol = 2 if tex_table is not None else 0       # textured materials go through the texture unit's filter
    depth = sim.depth_tensor().to_torch().cpu().numpy()
    rgb = sim.rgb_tensor().to_torch().cpu().numpy()
    assert depth.shape == (sim.num_worlds * A, height, size, 1) and depth.dtype == np.float32
    assert rgb.shape == (sim.num_worlds * A, height, size, 4) and rgb.dtype == np.uint8
    worst, maxd, covered, active = 0.0, 0.0, 0.0, 0
    for w in worlds:
        dump = sim.dump_world(w)
        for k in range(A):
            rd, rc = vu.render_world_view(meshes, dump, k, size, height)
            ok, frac, d = vu.compare(depth[w * A + k, :, :, 0], rgb[w * A + k], rd, rc, DEPTH_TOL, FLIP_FRAC, rgb_tol)
            assert ok, (w, k, frac, d)
            worst, maxd = max(worst, frac), max(maxd, d)
            if dump["iface_slot"][k] >= 0:
                covered += (rd > 0).mean()
                active += 1
    assert covered / active > 0.3        # the views do look at geometry
    return worst, maxd


@pytest.mark.parametrize("data_dir", [None, "ref"])
def test_views_match_oracle(data_dir):
    if data_dir == "ref":
        if not vu.have_ref_data():
            pytest.skip("reference render meshes not present")
        data_dir = vu.REF_DATA
    N, A = 48, 4
    sim, table, tex = make(N, data_dir)
    check_views(sim, table, range(0, N, 12), A, tex_table=tex, data_dir=data_dir)           # right after init
    drive(sim, 12)
    worst, maxd = check_views(sim, table, range(0, N, 6), A, tex_table=tex, data_dir=data_dir)
    if tex is not None:
        # the textures do show: the ground is no longer one flat colour
        rgb = sim.rgb_tensor().to_torch()[:, 40:, :, :3].reshape(-1, 3)
        assert len(torch.unique(rgb, dim=0)) > 20
    print(f"views vs oracle: worst flipped-pixel fraction {worst:.2e}, max |depth diff| {maxd:.2e}")


def test_inactive_slots_and_rectangular_views():
    import gpu_hideseek as gh
    N, A = 32, 6
    sim = gh.HideAndSeekSimulator(num_worlds=N, min_hiders=1, max_hiders=3, min_seekers=1, max_seekers=3,
                                  enable_batch_renderer=True, batch_render_width=44, batch_render_height=26)
    sim.init()
    drive(sim, 3, A=A)
    mask = sim.self_mask_tensor().to_torch().cpu().numpy().reshape(N * A)
    depth = sim.depth_tensor().to_torch().cpu().numpy()
    rgb = sim.rgb_tensor().to_torch().cpu().numpy()
    assert (mask == 0).any() and (mask == 1).any()
    assert (depth[mask == 0] == 0).all() and (rgb[mask == 0] == 0).all()
    assert (rgb[mask == 1][..., 3] == 255).all()
    assert (depth[mask == 1].reshape(int(mask.sum()), -1).max(1) > 0).all()
    # 44 x 26: neither a multiple of the 8 x 4 tile, so edge tiles are partly outside the image
    check_views(sim, gh.view_mesh_table(), [0, 7, 31], A, size=44, height=26)


def test_views_need_the_flag():
    import gpu_hideseek as gh
    import os
    os.environ["HIDESEEK_THROW"] = "1"
    try:
        sim = gh.HideAndSeekSimulator(num_worlds=4)
        sim.init()
        with pytest.raises(RuntimeError):
            sim.depth_tensor()
        with pytest.raises(RuntimeError):
            sim.rgb_tensor()
    finally:
        os.environ.pop("HIDESEEK_THROW", None)


def test_views_sharding_invariance_full_shard():
    """Size-independent property at a large shard: the image of global world g does not depend
    on which shard / slot holds it (and rendering does not disturb the simulation state)."""
    import gpu_hideseek as gh
    N, A, S = 4096, 4, 32
    big = gh.HideAndSeekSimulator(num_worlds=N, enable_batch_renderer=True, batch_render_width=S, batch_render_height=S)
    big.init()
    base = 3000
    small = gh.HideAndSeekSimulator(num_worlds=64, world_base=base, enable_batch_renderer=True,
                                    batch_render_width=S, batch_render_height=S)
    small.init()
    plain = gh.HideAndSeekSimulator(num_worlds=64, world_base=base)
    plain.init()
    for _ in range(5):
        big.step(); small.step(); plain.step()
    db = big.depth_tensor().to_torch()[base * A:(base + 64) * A]
    ds = small.depth_tensor().to_torch()
    assert torch.equal(db, ds)
    assert torch.equal(big.rgb_tensor().to_torch()[base * A:(base + 64) * A], small.rgb_tensor().to_torch())
    assert torch.equal(small.self_data_tensor().to_torch(), plain.self_data_tensor().to_torch())
    assert (ds.flatten(1).amax(1) > 0).all()


def test_benchmark_script_call_sequence():
    """scripts/benchmark.py of the reference, call for call (constructor keywords :20-35, tensors
    :38-47, loop :64-88), at a reduced world count: it must run unmodified against this module."""
    import gpu_hideseek
    num_worlds, num_steps, gpu_id = 1024, 12, 0
    sim = gpu_hideseek.HideAndSeekSimulator(
        exec_mode=gpu_hideseek.madrona.ExecMode.CUDA, gpu_id=gpu_id, num_worlds=num_worlds, sim_flags=0,
        rand_seed=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2, num_pbt_policies=1,
        enable_batch_renderer=True, batch_render_width=64, batch_render_height=64)
    sim.init()
    actions = sim.action_tensor().to_torch()
    resets = sim.reset_tensor().to_torch()
    assert actions.shape == (num_worlds * 4, 5) and resets.shape == (num_worlds, 1)
    rgb_observations = sim.rgb_tensor().to_torch()
    assert rgb_observations.shape == (num_worlds * 4, 64, 64, 4) and rgb_observations.dtype == torch.uint8
    first = rgb_observations.clone()
    for i in range(5 + num_steps):
        sim.step()
        actions[..., 0] = torch.randint_like(actions[..., 0], -5, 5)
        actions[..., 1] = torch.randint_like(actions[..., 1], -5, 5)
    assert not torch.equal(first, rgb_observations)            # the zero-copy view follows the simulation
    assert (rgb_observations[..., 3] == 255).all()
    del sim
