"""Checkpoint save / load vs the reference CPU backend: the Checkpoint
tensor keeps the reference's byte layout (src/sim.hpp:283-313), loading regenerates the
level bit-exactly from the saved episode key and restores dynamic state + grab joints
(src/sim.cpp:956-1137), and the trajectories stay matched afterwards."""
import numpy as np
import pytest
import torch

import refsim
from parity import compare_world_static, compare_bvh, compare_exports

pytestmark = pytest.mark.gpu

TOL = 1e-4

ckpt_body = [("pos", "f4", 3), ("rot", "f4", 4), ("vel", "f4", 6)]
ckpt_agent = np.dtype(ckpt_body + [("grab_idx", "i4"), ("r1", "f4", 3), ("r2", "f4", 3),
                                   ("attach1", "f4", 4), ("attach2", "f4", 4), ("sep", "f4")])
ckpt_obj = np.dtype(ckpt_body + [("team", "u4"), ("locked", "u1"), ("pad", "u1", 3)])
ckpt_dt = np.dtype([("key", "u4", 2), ("scores", "i4", 2), ("step", "i4"), ("agents", ckpt_agent, 6),
                    ("boxes", ckpt_obj, 9), ("ramps", ckpt_obj, 2), ("num_hiders", "i4"),
                    ("num_seekers", "i4"), ("num_boxes", "i4"), ("num_ramps", "i4")])
assert ckpt_dt.itemsize == 1392


def _compare_ckpt(r, h):
    for f in ("key", "scores", "step", "num_hiders", "num_seekers", "num_boxes", "num_ramps"):
        assert np.array_equal(r[f], h[f]), f
    na = int(r["num_hiders"] + r["num_seekers"])
    assert np.array_equal(r["agents"]["grab_idx"][:na], h["agents"]["grab_idx"][:na])
    for f in ("pos", "rot", "vel", "r1", "r2", "attach1", "attach2", "sep"):
        assert np.abs(r["agents"][f][:na] - h["agents"][f][:na]).max() <= TOL, f
    for grp, n in (("boxes", int(r["num_boxes"])), ("ramps", int(r["num_ramps"]))):
        assert np.array_equal(r[grp]["team"][:n], h[grp]["team"][:n])
        assert np.array_equal(r[grp]["locked"][:n], h[grp]["locked"][:n])
        for f in ("pos", "rot", "vel"):
            assert np.abs(r[grp][f][:n] - h[grp][f][:n]).max() <= TOL, (grp, f)


def test_save_load_roundtrip_matches_reference():
    import gpu_hideseek as gh
    N, seed, hs, sk = 8, 5, (1, 3), (1, 3)
    A = hs[1] + sk[1]
    ref = refsim.RefSim(N, seed=seed, min_hiders=hs[0], max_hiders=hs[1], min_seekers=sk[0],
                        max_seekers=sk[1], num_threads=1)
    sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=seed, min_hiders=hs[0], max_hiders=hs[1],
                                  min_seekers=sk[0], max_seekers=sk[1])
    ref.init()
    sim.init()
    rt = ref.tensors()
    act = sim.action_tensor().to_torch()
    rng = np.random.RandomState(99)

    def run(steps):
        for _ in range(steps):
            a = np.zeros((N * A, 5), dtype=np.int32)
            a[:, 0:3] = rng.randint(0, 11, size=(N * A, 3))
            a[:, 3] = rng.rand(N * A) < 0.3
            a[:, 4] = rng.rand(N * A) < 0.15
            rt["action"][:] = a
            act.copy_(torch.from_numpy(a).to(act.device))
            ref.step()
            sim.step()

    def check(tag):
        for w in range(N):
            rw, hw = ref.dump_world(w), sim.dump_world(w)
            errs = compare_world_static(rw, hw, exact=False, tol=TOL) + compare_bvh(rw, hw, exact=False, tol=TOL)
            assert not errs, (tag, w, errs[:4])
        errs, _ = compare_exports(sim, rt, hs[0] + sk[0], tol=5e-4)
        assert not errs, (tag, errs)

    run(120)                       # past the prep phase, grabs and locks in place
    saved = [1, 4, 6]
    for w in saved:
        ref.l.hsref_save_checkpoint(ref.h, w)
        sim.save_checkpoint(w)
    rck = ref._view("Checkpoint", "u1", (N, 1392)).view(ckpt_dt).reshape(N)
    hck = sim.ckpt_tensor().to_torch().cpu().numpy().view(ckpt_dt).reshape(N)
    for w in saved:
        _compare_ckpt(rck[w], hck[w])
    assert any(int(x) >= 0 for w in saved for x in rck[w]["agents"]["grab_idx"][:6]), "no grab in fixture"

    run(15)                        # diverge from the saved state
    for w in saved:
        ref.l.hsref_load_checkpoint(ref.h, w)
        sim.load_checkpoint(w)
    check("after load")
    run(20)
    check("after load + 20 steps")
    assert sim.overflow_count() == 0


def test_load_refuses_out_of_range_agent_counts():
    """The checkpoint tensor is user-writable: a record whose agent counts exceed the Manager's
    team bounds (the reference only asserts) must not index past the world's agent rows — the
    world keeps its state and the refusal is counted."""
    import gpu_hideseek as gh
    from compact import hs_dump_compact
    N = 8
    sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=1, min_hiders=1, max_hiders=2, min_seekers=1, max_seekers=2)
    sim.init()
    for _ in range(5):
        sim.step()
    for w in range(N):
        sim.save_checkpoint(w)
    blob = sim.ckpt_tensor().to_torch()
    host = blob.cpu().numpy().copy().view(ckpt_dt).reshape(N)
    host["num_hiders"][2] = 3              # above max_hiders = 2
    host["num_seekers"][5] = -1
    host["num_hiders"][6] = 1000000
    blob.copy_(torch.from_numpy(host.view(np.uint8).reshape(N, 1392)).to(blob.device))
    for _ in range(3):
        sim.step()
    before = hs_dump_compact(sim)
    sim.ckpt_ctrl_tensor().to_torch().fill_(1)
    sim.load_checkpoints()
    after = hs_dump_compact(sim)
    assert sim.overflow_count() == 3
    for w in range(N):
        same = before[w]["pose"].tobytes() == after[w]["pose"].tobytes() and \
            int(before[w]["cur_episode_step"]) == int(after[w]["cur_episode_step"])
        assert same == (w in (2, 5, 6)), w          # refused worlds keep their state, the others rewind
    sim.step()                                      # and the simulator keeps running
    assert np.isfinite(sim.self_data_tensor().to_torch().cpu().numpy()).all()
