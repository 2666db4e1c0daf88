"""numpy mirror of include/hideseek/debug_dump.h (hs_world_dump)."""
import numpy as np

body_dt = np.dtype([("eid", "i4"), ("obj", "i4"), ("response", "i4"), ("owner", "i4"),
                    ("leaf", "i4"), ("pad", "i4", 3), ("pos", "f4", 3), ("rot", "f4", 4),
                    ("scale", "f4", 3), ("vel", "f4", 6), ("ext_force", "f4", 3),
                    ("ext_torque", "f4", 3)])
joint_dt = np.dtype([("active", "i4"), ("other_slot", "i4"), ("eid", "i4"), ("seq", "i4"),
                     ("attach1", "f4", 4), ("attach2", "f4", 4), ("separation", "f4"),
                     ("r1", "f4", 3), ("r2", "f4", 3)])
node_dt = np.dtype([("bounds", "f4", (6, 4)), ("children", "i4", 4), ("parent", "i4")])
world_dt = np.dtype([
    ("cur_episode_step", "i4"), ("cur_world_episode", "i4"), ("episode_key", "u4", 2),
    ("rng_key", "u4", 2), ("rng_count", "u4"),
    ("num_hiders", "i4"), ("num_seekers", "i4"), ("num_agents", "i4"),
    ("num_dyn", "i4"), ("num_bodies", "i4"), ("num_boxes", "i4"), ("num_ramps", "i4"),
    ("hider_team_reward", "f4"), ("seekers_first", "i4"), ("running_scores", "i4", 2),
    ("num_cand", "i4"),
    ("box_slot", "i4", 9), ("ramp_slot", "i4", 2), ("hider_slot", "i4", 3), ("seeker_slot", "i4", 3),
    ("iface_slot", "i4", 6), ("iface_type", "i4", 6),
    ("bodies", body_dt, 52), ("joints", joint_dt, 6),
    ("bvh_num_nodes", "i4"), ("bvh_num_leaves", "i4"), ("bvh_force_rebuild", "i4"),
    ("nodes", node_dt, 69), ("leaf_aabbs", "f4", (52, 6)), ("leaf_parents", "u4", 52),
    ("leaf_slot", "i4", 52),
    ("cand_a", "i4", 256), ("cand_b", "i4", 256), ("contact_meta", "i4", 256),
    ("contacts", "f4", (256, 20))])

# mirror of hs::ViewMeshTable (csrc/hs_views.cuh)
view_box_dt = np.dtype([("lo", "f4", 3), ("first", "i4"), ("hi", "f4", 3), ("count", "i4")])
view_mesh_dt = np.dtype([
    ("obj_first_sub", "i4", 8), ("obj_num_sub", "i4", 8), ("obj_aabb", "f4", (8, 6)),
    ("sub_color", "f4", (8, 3, 3)), ("sub", view_box_dt, 16), ("cluster", view_box_dt, 200),
    ("num_sub", "i4"), ("num_clusters", "i4"), ("num_tris", "i4"), ("num_small_tris", "i4"),
    ("fov_scale", "f4"), ("tri", "f4", (1552, 9))])
# mirror of hs::ViewTexTable
view_tex_dt = np.dtype([("tri_uv", "f4", (1552, 6)), ("sub_tex", "i4", (8, 3)), ("tex", "u8", 3)])
