"""The probing step of the oracle harness (hsref_step_probe) must be the reference's step: a
simulator advanced with it ends bit-identical to one advanced with Manager::step(), and what it
captures (candidate pairs, last-substep contacts, lidar hit bodies) is self-consistent."""
import numpy as np
import pytest

import refsim
from compact import ref_step_probe, ref_dump_compact, compare_compact


@pytest.mark.skipif(not refsim.available(), reason="oracle/_ref not built")
def test_probe_step_equals_plain_step():
    N, A, steps = 12, 4, 12
    a = refsim.RefSim(N, seed=0, num_threads=1)
    b = refsim.RefSim(N, seed=0, num_threads=1)
    a.init(); b.init()
    ta, tb = a.tensors(), b.tensors()
    rng = np.random.RandomState(3)
    for s in range(steps):
        act = np.zeros((N * A, 5), dtype=np.int32)
        act[:, 0:3] = rng.randint(0, 11, size=(N * A, 3))
        act[:, 3] = rng.rand(N * A) < 0.3
        ta["action"][:] = act
        tb["action"][:] = act
        a.step()
        ref_step_probe(b, 3 if s % 2 else 1)
        for k in ("reward", "done", "lidar", "self_obs", "box_obs", "agent_vis"):
            assert np.array_equal(ta[k].view(np.uint32), tb[k].view(np.uint32)), (s, k)
    ca, cb = ref_dump_compact(a), ref_dump_compact(b)
    errs, _ = compare_compact(ca, cb, exact=True, probes=False)
    assert not errs, errs
    assert (ca["num_cand"] == -1).all()          # never probed
    ok = cb["num_cand"] >= 0
    assert ok.sum() >= N - 1
    c = cb[ok]
    assert (c["num_cand"] > 0).all() and (c["num_contacts"] > 0).all()
    assert (c["num_contacts"] <= c["num_cand"]).all()
    for w in range(len(c)):
        nb, nc, nk = int(c["num_bodies"][w]), int(c["num_cand"][w]), int(c["num_contacts"][w])
        assert (c["cand"][w][:nc] < nb).all()
        assert (c["contact_ref"][w][:nk] < nb).all() and (c["contact_alt"][w][:nk] < nb).all()
        assert ((c["contact_npts"][w][:nk] >= 1) & (c["contact_npts"][w][:nk] <= 4)).all()
        # a contact pair is one of the candidate pairs
        pairs = {tuple(sorted(p)) for p in c["cand"][w][:nc].tolist()}
        for k in range(nk):
            assert tuple(sorted((int(c["contact_ref"][w][k]), int(c["contact_alt"][w][k])))) in pairs
    # lidar: a ray reports a body exactly when the exported depth is non-zero
    lid = tb["lidar"].reshape(N, A, 30)[ok]
    hit = c["lidar_hit"][:, :A, :]
    assert (hit > -2).all()
    assert np.array_equal(hit >= 0, lid != 0.0)
