"""Bulk parity support: the compact per-world record both sides can dump for a whole shard in
one call (hs_compact_world in include/hideseek/debug_dump.h, hsref_compact in
oracle/ref_dump.h) and a vectorised comparison of two arrays of them.

TEST INFRASTRUCTURE ONLY.
"""
import ctypes as C

import numpy as np

B, NODES, CAND, CON, AG = 52, 69, 256, 96, 6

compact_dt = np.dtype([
    ("cur_episode_step", "i4"), ("cur_world_episode", "i4"), ("rng_key", "u4", 2), ("rng_count", "u4"),
    ("num_hiders", "i4"), ("num_seekers", "i4"), ("num_boxes", "i4"), ("num_ramps", "i4"),
    ("hider_team_reward", "f4"), ("seekers_first", "i4"), ("running_scores", "i4", 2),
    ("num_bodies", "i4"), ("num_dyn", "i4"), ("bvh_num_leaves", "i4"), ("bvh_num_nodes", "i4"),
    ("num_cand", "i4"), ("num_contacts", "i4"),
    ("obj", "u1", B), ("response", "u1", B), ("owner", "u1", B), ("leaf", "u1", B), ("eid_rank", "u1", B),
    ("joint_other", "i1", AG), ("pad0", "i1", 2), ("joint_sep", "f4", AG),
    ("pose", "f4", (B, 7)), ("scale", "f4", (B, 3)), ("vel", "f4", (B, 6)), ("force", "f4", (B, 6)),
    ("node_bounds", "f4", (NODES, 6, 4)), ("node_children", "i4", (NODES, 4)), ("node_parent", "i4", NODES),
    ("leaf_aabb", "f4", (B, 6)), ("leaf_parent", "u4", B),
    ("cand", "u1", (CAND, 2)),
    ("contact_ref", "u1", CON), ("contact_alt", "u1", CON), ("contact_npts", "u1", CON),
    ("contact_pts", "f4", (CON, 4, 4)), ("contact_normal", "f4", (CON, 3)), ("contact_lambda", "f4", CON),
    ("lidar_hit", "i1", (AG, 30))], align=True)


def ref_step_probe(ref, num_threads=0):
    l = ref.l
    l.hsref_step_probe.argtypes = [C.c_void_p, C.c_int]
    rc = l.hsref_step_probe(ref.h, num_threads)
    assert rc == 0, f"hsref_step_probe: {rc}"


def ref_dump_compact(ref, w0=0, count=None, num_threads=0):
    l = ref.l
    assert l.hsref_compact_bytes() == compact_dt.itemsize, (l.hsref_compact_bytes(), compact_dt.itemsize)
    count = ref.N - w0 if count is None else count
    out = np.zeros(count, dtype=compact_dt)
    l.hsref_dump_compact.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
    rc = l.hsref_dump_compact(ref.h, w0, count, out.ctypes.data, num_threads)
    assert rc == 0
    return out


def hs_set_debug(sim, flags):
    lib = sim._clib
    lib.hs_set_debug.argtypes = [C.c_void_p, C.c_uint]
    assert lib.hs_set_debug(sim._h, flags) == 0, lib.hs_last_error()


def hs_dump_compact(sim, w0=0, count=None):
    lib = sim._clib
    assert lib.hs_compact_dump_bytes() == compact_dt.itemsize, (lib.hs_compact_dump_bytes(), compact_dt.itemsize)
    count = sim.num_worlds - w0 if count is None else count
    out = np.zeros(count, dtype=compact_dt)
    lib.hs_dump_compact.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64]
    rc = lib.hs_dump_compact(sim._h, w0, count, out.ctypes.data, out.nbytes)
    assert rc == 0, lib.hs_last_error()
    return out


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _rows(mask):
    """world indices with any True along the remaining axes"""
    return np.nonzero(mask.reshape(mask.shape[0], -1).any(axis=1))[0]


def compare_compact(r, h, tol=1e-4, exact=False, check_scores=True, probes=True, w0=0):
    """Vectorised diff of reference records `r` against product records `h`.

    Discrete state must be equal; float state bit-equal when `exact` else within `tol` (abs).
    Returns (errors, stats): errors is a list of strings naming the field and the first worlds
    that differ, stats carries max |diff| per float field and the probe counters."""
    errs, stats = [], {}
    n = len(r)

    def bad(name, mask):
        idx = _rows(mask)
        if len(idx):
            errs.append(f"{name}: {len(idx)} worlds differ, first {[int(i) + w0 for i in idx[:5]]}")

    scalars = ["cur_episode_step", "cur_world_episode", "rng_count", "num_hiders", "num_seekers", "num_boxes",
               "num_ramps", "seekers_first", "num_bodies", "num_dyn", "bvh_num_leaves", "bvh_num_nodes"]
    for f in scalars:
        bad(f, r[f] != h[f])
    bad("rng_key", r["rng_key"] != h["rng_key"])
    bad("hider_team_reward", r["hider_team_reward"] != h["hider_team_reward"])
    if check_scores:
        bad("running_scores", r["running_scores"] != h["running_scores"])
    same_nb = r["num_bodies"] == h["num_bodies"]
    live = (np.arange(B)[None, :] < r["num_bodies"][:, None]) & same_nb[:, None]
    for f in ("obj", "response", "owner", "leaf", "eid_rank"):
        bad(f, (r[f] != h[f]) & live)
    bad("joint_other", r["joint_other"] != h["joint_other"])
    for f in ("pose", "scale", "vel", "force"):
        a, b = r[f], h[f]
        m = live[:, :, None]
        if exact:
            bad(f + " (bits)", (_bits(a) != _bits(b)) & m)
        d = np.where(m, np.abs(a.astype(np.float64) - b), 0.0)
        stats["max_" + f] = float(d.max()) if n else 0.0
        if not exact:
            bad(f"{f} (> {tol})", ~(d <= tol))
    a, b = r["joint_sep"], h["joint_sep"]
    d = np.abs(a.astype(np.float64) - b)
    stats["max_joint_sep"] = float(d.max()) if n else 0.0
    bad("joint_sep", ~(d <= (0.0 if exact else tol)))

    # BVH: topology exact; bounds of live child slots exact or within tol; leaf boxes likewise
    same_bvh = (r["bvh_num_nodes"] == h["bvh_num_nodes"]) & (r["bvh_num_leaves"] == h["bvh_num_leaves"])
    # only nodes reachable from the root are defined (num_nodes is an allocation bound)
    nodelive = np.zeros((n, NODES), dtype=bool)
    if n:
        nodelive[:, 0] = True
    ch = r["node_children"]
    for _ in range(8):
        valid = nodelive[:, :, None] & (ch >= 0) & (ch < NODES)
        wi = np.nonzero(valid)[0]
        nodelive[wi, ch[valid]] = True
    nodelive &= same_bvh[:, None]
    bad("node_children", (r["node_children"] != h["node_children"]) & nodelive[:, :, None])
    bad("node_parent", (r["node_parent"] != h["node_parent"]) & nodelive)
    child_live = (r["node_children"] != -1) & nodelive[:, :, None]          # [n, NODES, 4]
    with np.errstate(invalid="ignore"):                # inf - inf of equal infinite bounds (the ground plane)
        db = np.where(child_live[:, :, None, :], np.abs(r["node_bounds"].astype(np.float64) - h["node_bounds"]), 0.0)
    db = np.nan_to_num(db, nan=0.0, posinf=np.inf)       # inf - inf of equal infinite bounds
    eqb = (_bits(r["node_bounds"]) == _bits(h["node_bounds"])) | ~child_live[:, :, None, :]
    stats["max_node_bounds"] = float(np.where(eqb, 0.0, db).max()) if n else 0.0
    if exact:
        bad("node_bounds (bits)", ~eqb)
    else:
        bad(f"node_bounds (> {tol})", ~eqb & ~(db <= tol))
    leaflive = (np.arange(B)[None, :] < r["bvh_num_leaves"][:, None]) & same_bvh[:, None]
    bad("leaf_parent", (r["leaf_parent"] != h["leaf_parent"]) & leaflive)
    eql = (_bits(r["leaf_aabb"]) == _bits(h["leaf_aabb"])) | ~leaflive[:, :, None]
    dl = np.where(eql, 0.0, np.abs(r["leaf_aabb"].astype(np.float64) - h["leaf_aabb"]))
    stats["max_leaf_aabb"] = float(dl.max()) if n else 0.0
    if exact:
        bad("leaf_aabb (bits)", ~eql)
    else:
        bad(f"leaf_aabb (> {tol})", ~(dl <= tol))

    if probes:
        ok = r["num_cand"] >= 0                       # worlds the reference-side discovery pass spoiled
        stats["probe_worlds"] = int(ok.sum())
        bad("num_cand", (r["num_cand"] != h["num_cand"]) & ok)
        same_c = ok & (r["num_cand"] == h["num_cand"])
        cl = (np.arange(CAND)[None, :] < np.minimum(r["num_cand"], CAND)[:, None]) & same_c[:, None]
        bad("candidate pairs (order-exact)", (r["cand"] != h["cand"]) & cl[:, :, None])
        bad("num_contacts", (r["num_contacts"] != h["num_contacts"]) & ok)
        same_k = ok & (r["num_contacts"] == h["num_contacts"])
        kl = (np.arange(CON)[None, :] < np.minimum(r["num_contacts"], CON)[:, None]) & same_k[:, None]
        for f in ("contact_ref", "contact_alt", "contact_npts"):
            bad(f, (r[f] != h[f]) & kl)
        pl = kl[:, :, None] & (np.arange(4)[None, None, :] < r["contact_npts"][:, :, None])
        for f, m in (("contact_pts", pl[:, :, :, None]), ("contact_normal", kl[:, :, None]), ("contact_lambda", kl)):
            d = np.where(m, np.abs(r[f].astype(np.float64) - h[f]), 0.0)
            stats["max_" + f] = float(d.max()) if n else 0.0
            if exact:
                bad(f + " (bits)", (_bits(r[f]) != _bits(h[f])) & m)
            else:
                bad(f"{f} (> {tol})", ~(d <= tol))
        bad("lidar hit body", (r["lidar_hit"] != h["lidar_hit"]) & ok[:, None, None])
        stats["max_cand"] = int(max(r["num_cand"].max(), h["num_cand"].max())) if n else 0
        stats["max_contacts"] = int(max(r["num_contacts"].max(), h["num_contacts"].max())) if n else 0
        stats["mean_cand"] = float(h["num_cand"].mean()) if n else 0.0
        stats["mean_contacts"] = float(h["num_contacts"].mean()) if n else 0.0
        stats["lidar_rays"] = int((h["lidar_hit"] > -2).sum())
        stats["lidar_hits"] = int((h["lidar_hit"] >= 0).sum())
    return errs, stats
