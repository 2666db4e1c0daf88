"""Shared comparison helpers: reference dump (oracle/_ref) vs hs_dump_world."""
import numpy as np


def f32bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def ref_bodies(rw):
    return rw["bodies"][: rw["num_bodies"]]


def compare_world_static(rw, hw, exact=True, tol=0.0, check_scores=True):
    """Compare body tables + bookkeeping. Returns list of mismatch strings."""
    errs = []
    nb = int(rw["num_bodies"])
    if nb != int(hw["num_bodies"]):
        return [f"num_bodies ref {nb} hs {int(hw['num_bodies'])}"]
    rb, hb = rw["bodies"][:nb], hw["bodies"][:nb]
    for f, g in (("obj_id", "obj"), ("response", "response"), ("owner", "owner"), ("leaf_id", "leaf")):
        if not np.array_equal(rb[f], hb[g]):
            errs.append(f"{f} mismatch {rb[f].tolist()} vs {hb[g].tolist()}")
    # entity ids: only the relative order inside the world matters
    if not np.array_equal(np.argsort(rb["entity_id"], kind="stable"), np.argsort(hb["eid"], kind="stable")):
        errs.append(f"entity id order mismatch {rb['entity_id'].tolist()} vs {hb['eid'].tolist()}")
    for f in ("pos", "rot", "scale", "vel", "ext_force", "ext_torque"):
        if exact:
            if not np.array_equal(f32bits(rb[f]), f32bits(hb[f])):
                d = np.abs(rb[f] - hb[f]).max()
                errs.append(f"{f} not bit-exact (max abs diff {d})")
        else:
            d = np.abs(rb[f].astype(np.float64) - hb[f]).max() if nb else 0.0
            if d > tol:
                errs.append(f"{f} max abs diff {d} > {tol}")
    for f, g in (("cur_episode_step", "cur_episode_step"), ("num_hiders", "num_hiders"),
                 ("num_seekers", "num_seekers"), ("num_boxes", "num_boxes"), ("num_ramps", "num_ramps"),
                 ("seekers_first", "seekers_first"), ("rng_count", "rng_count"),
                 ("cur_world_episode", "cur_world_episode")):
        if int(rw[f]) != int(hw[g]):
            errs.append(f"{f}: ref {int(rw[f])} hs {int(hw[g])}")
    if not np.array_equal(rw["rng_key"], hw["rng_key"]):
        errs.append("episode rng key mismatch")
    # EpisodeStats is uninitialised memory in the reference until the first step of an episode
    if check_scores and not np.array_equal(rw["running_scores"], hw["running_scores"]):
        errs.append("running scores mismatch")
    if float(rw["hider_team_reward"]) != float(hw["hider_team_reward"]):
        errs.append("hider_team_reward mismatch")
    return errs


def compare_bvh(rw, hw, exact=True, tol=0.0):
    errs = []
    nl = int(rw["bvh_num_leaves"])
    if nl != int(hw["bvh_num_leaves"]) or int(rw["bvh_num_nodes"]) != int(hw["bvh_num_nodes"]):
        return [f"bvh sizes ref ({nl},{int(rw['bvh_num_nodes'])}) hs ({int(hw['bvh_num_leaves'])},{int(hw['bvh_num_nodes'])})"]
    if not np.array_equal(rw["leaf_parents"][:nl], hw["leaf_parents"][:nl]):
        errs.append("leaf parents mismatch")
    # walk the tree from the root comparing structure and bounds
    stack = [0]
    seen = 0
    while stack:
        n = stack.pop()
        seen += 1
        rn, hn = rw["nodes"][n], hw["nodes"][n]
        if not np.array_equal(rn["children"], hn["children"]) or int(rn["parent"]) != int(hn["parent"]):
            errs.append(f"node {n} topology mismatch {rn['children']} vs {hn['children']}")
            break
        for i in range(4):
            c = int(rn["children"][i])
            if c == -1:
                continue
            rbnd = np.asarray(rn["bounds"]).reshape(6, 4)[:, i]
            hbnd = np.asarray(hn["bounds"]).reshape(6, 4)[:, i]
            if exact:
                if not np.array_equal(f32bits(rbnd), f32bits(hbnd)):
                    errs.append(f"node {n} child {i} bounds differ {rbnd} vs {hbnd}")
            elif np.abs(rbnd.astype(np.float64) - hbnd).max() > tol:
                errs.append(f"node {n} child {i} bounds differ beyond tol")
            if c >= 0:
                stack.append(c)
    la_r, la_h = rw["leaf_aabbs"][:nl], hw["leaf_aabbs"][:nl]
    if exact:
        if not np.array_equal(f32bits(la_r), f32bits(la_h)):
            errs.append("leaf aabbs not bit-exact")
    elif np.abs(la_r.astype(np.float64) - la_h).max() > tol:
        errs.append("leaf aabbs differ beyond tol")
    return errs


# (binding getter, reference key, float-exactness class)
EXPORTS = [("seed", "seed", "int"), ("self_type", "self_type", "int"), ("self_mask", "self_mask", "mask"),
           ("action", "action", "int"), ("prep_counter", "prep_counter", "int"),
           ("reward", "reward", "exactf"), ("done", "done", "int"),
           ("lidar", "lidar", "float"), ("self_data", "self_obs", "float"),
           ("agent_data", "agent_obs", "float"), ("box_data", "box_obs", "float"),
           ("ramp_data", "ramp_obs", "float"),
           ("visible_agents_mask", "agent_vis", "exactf"), ("visible_boxes_mask", "box_vis", "exactf"),
           ("visible_ramps_mask", "ramp_vis", "exactf"),
           ("episode_result", "episode_result", "world"), ("global_positions", "global_pos", "globalpos")]


def compare_exports(sim, rt, min_agents, tol, skip=(), prev_active=None):
    """Compare every exported tensor.  Rows of inactive agent interfaces hold
    uninitialised memory in the reference and are skipped.
    Reward / done are written BEFORE the reset stage of a step (src/sim.cpp:805-842 runs ahead
    of resetSystem), so a row whose agent only appeared in this step's reset still holds
    uninitialised reward / done: pass the activity mask from before the step as `prev_active`.
    Returns (errors, max float difference)."""
    errs, maxd = [], 0.0
    active = rt["self_mask"].reshape(-1) == 1.0
    written = active if prev_active is None else (active & prev_active)
    for name, key, kind in EXPORTS:
        if name in skip:
            continue
        mine = getattr(sim, name + "_tensor")().to_torch().cpu().numpy()
        theirs = np.asarray(rt[key]).reshape(mine.shape)
        if kind == "mask":
            if not np.array_equal(mine, theirs):
                errs.append(name)
            continue
        if kind == "globalpos":
            k = 11 + min_agents
            mine, theirs = mine[:, :k], theirs[:, :k]
        elif kind != "world":
            rows = written if name in ("reward", "done") else active
            mine, theirs = mine[rows], theirs[rows]
        if kind == "int":
            if not np.array_equal(mine, theirs):
                errs.append(name)
        else:
            d = float(np.abs(mine.astype(np.float64) - theirs).max()) if mine.size else 0.0
            maxd = max(maxd, d)
            lim = 0.0 if kind == "exactf" else tol
            if not (d <= lim):
                errs.append(f"{name}: max abs diff {d}")
    return errs, maxd
