"""ctypes driver for the reference oracle (oracle/_ref/libhs_ref.so).

TEST INFRASTRUCTURE ONLY.  The library is the UNMODIFIED reference CPU backend
(src/mgr.cpp, src/sim.cpp, Madrona physics) compiled by oracle/Makefile; this
module only loads it, steps it and reads tensors / per-world dumps.
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libhs_ref.so")

# ExportID (reference src/sim.hpp:45-68)
EXPORT = dict(Reset=0, PrepCounter=1, Action=2, SelfObs=3, SelfType=4, SelfMask=5,
              AgentObsData=6, BoxObsData=7, RampObsData=8, AgentVisMasks=9,
              BoxVisMasks=10, RampVisMasks=11, Lidar=12, Seed=13, Reward=14,
              Done=15, GlobalDebugPositions=16, AgentPolicy=17, EpisodeResult=18,
              CheckpointControl=19, Checkpoint=20)

MAX_BODIES, MAX_NODES, MAX_JOINTS, MAX_AGENTS = 64, 128, 16, 6

body_dt = np.dtype([
    ("entity_id", "i4"), ("archetype", "i4"), ("row", "i4"), ("obj_id", "i4"),
    ("response", "i4"), ("owner", "i4"), ("leaf_id", "i4"), ("grab_joint", "i4"),
    ("pos", "f4", 3), ("rot", "f4", 4), ("scale", "f4", 3), ("vel", "f4", 6),
    ("ext_force", "f4", 3), ("ext_torque", "f4", 3)])
joint_dt = np.dtype([
    ("entity_id", "i4"), ("e1", "i4"), ("e2", "i4"), ("attach1", "f4", 4),
    ("attach2", "f4", 4), ("separation", "f4"), ("r1", "f4", 3), ("r2", "f4", 3)])
node_dt = np.dtype([("bounds", "f4", (6, 4)), ("children", "i4", 4), ("parent", "i4")])
world_dt = np.dtype([
    ("cur_episode_step", "i4"), ("cur_world_episode", "i4"),
    ("episode_key", "u4", 2), ("rng_key", "u4", 2), ("rng_count", "u4"),
    ("num_hiders", "i4"), ("num_seekers", "i4"), ("num_active_agents", "i4"),
    ("num_obstacles", "i4"), ("num_boxes", "i4"), ("num_ramps", "i4"),
    ("hider_team_reward", "f4"), ("seekers_first", "i4"), ("running_scores", "i4", 2),
    ("obstacles", "i4", MAX_BODIES), ("boxes", "i4", 9), ("ramps", "i4", 2),
    ("hiders", "i4", 3), ("seekers", "i4", 3), ("agent_iface_sim", "i4", MAX_AGENTS),
    ("box_sizes", "f4", (9, 3)),
    ("num_bodies", "i4"), ("bodies", body_dt, MAX_BODIES),
    ("num_joints", "i4"), ("joints", joint_dt, MAX_JOINTS),
    ("bvh_num_leaves", "i4"), ("bvh_num_nodes", "i4"), ("bvh_force_rebuild", "i4"),
    ("nodes", node_dt, MAX_NODES),
    ("leaf_aabbs", "f4", (MAX_BODIES, 6)), ("leaf_parents", "u4", MAX_BODIES),
    ("leaf_entities", "i4", MAX_BODIES), ("leaf_obj_ids", "i4", MAX_BODIES)])
object_dt = np.dtype([
    ("prim_type", "i4"), ("num_verts", "i4"), ("num_faces", "i4"), ("num_hedges", "i4"),
    ("prim_aabb", "f4", 6), ("obj_aabb", "f4", 6), ("inv_mass", "f4"),
    ("inv_inertia", "f4", 3), ("mu_s", "f4"), ("mu_d", "f4"),
    ("verts", "f4", (8, 3)), ("planes", "f4", (6, 4)), ("face_base_hedge", "u4", 6),
    ("hedges", "u4", (24, 3))])

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(LIB_PATH)
        l.hsref_create.restype = C.c_void_p
        l.hsref_create.argtypes = [C.c_int, C.c_uint32, C.c_uint32] + [C.c_int] * 5
        for n in ("hsref_destroy", "hsref_init", "hsref_step"):
            getattr(l, n).argtypes = [C.c_void_p]
            getattr(l, n).restype = None
        l.hsref_tensor.restype = C.c_void_p
        l.hsref_tensor.argtypes = [C.c_void_p, C.c_int]
        l.hsref_trigger_reset.argtypes = [C.c_void_p, C.c_int, C.c_int]
        l.hsref_set_action.argtypes = [C.c_void_p] + [C.c_int] * 6
        l.hsref_dump_world.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        l.hsref_dump_objects.argtypes = [C.c_void_p, C.c_void_p]
        l.hsref_bench.restype = C.c_double
        l.hsref_bench.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint32]
        l.hsref_save_checkpoint.argtypes = [C.c_void_p, C.c_int]
        l.hsref_load_checkpoint.argtypes = [C.c_void_p, C.c_int]
        assert l.hsref_world_dump_bytes() == world_dt.itemsize, \
            (l.hsref_world_dump_bytes(), world_dt.itemsize)
        assert l.hsref_object_dump_bytes() == object_dt.itemsize
        _lib = l
    return _lib


class RefSim:
    """The reference `Manager` on its CPU backend (src/mgr.hpp:14-99)."""

    def __init__(self, num_worlds, sim_flags=0, seed=0, min_hiders=2, max_hiders=2,
                 min_seekers=2, max_seekers=2, num_threads=1):
        self.l = lib()
        self.N = num_worlds
        self.A = max_hiders + max_seekers
        self.h = self.l.hsref_create(num_worlds, sim_flags, seed, min_hiders, max_hiders,
                                     min_seekers, max_seekers, num_threads)

    def close(self):
        if self.h:
            self.l.hsref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def init(self):
        self.l.hsref_init(self.h)

    def step(self):
        self.l.hsref_step(self.h)

    def _view(self, name, dtype, shape):
        p = self.l.hsref_tensor(self.h, EXPORT[name])
        n = int(np.prod(shape))
        buf = (C.c_byte * (n * np.dtype(dtype).itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def tensors(self):
        N, A = self.N, self.A
        return dict(
            reset=self._view("Reset", "i4", (N, 1)),
            prep_counter=self._view("PrepCounter", "i4", (N * A, 1)),
            action=self._view("Action", "i4", (N * A, 5)),
            self_obs=self._view("SelfObs", "f4", (N * A, 13)),
            self_type=self._view("SelfType", "i4", (N * A, 1)),
            self_mask=self._view("SelfMask", "f4", (N * A, 1)),
            agent_obs=self._view("AgentObsData", "f4", (N * A, 5, 14)),
            box_obs=self._view("BoxObsData", "f4", (N * A, 9, 17)),
            ramp_obs=self._view("RampObsData", "f4", (N * A, 2, 14)),
            agent_vis=self._view("AgentVisMasks", "f4", (N * A, 5, 1)),
            box_vis=self._view("BoxVisMasks", "f4", (N * A, 9, 1)),
            ramp_vis=self._view("RampVisMasks", "f4", (N * A, 2, 1)),
            lidar=self._view("Lidar", "f4", (N * A, 30)),
            seed=self._view("Seed", "i4", (N * A, 2)),
            reward=self._view("Reward", "f4", (N * A, 1)),
            done=self._view("Done", "i4", (N * A, 1)),
            global_pos=self._view("GlobalDebugPositions", "f4", (N, 17, 2)),
            policy=self._view("AgentPolicy", "i4", (N * A, 1)),
            episode_result=self._view("EpisodeResult", "f4", (N, 2)),
        )

    def trigger_reset(self, world, level):
        self.l.hsref_trigger_reset(self.h, world, level)

    def dump_world(self, w):
        out = np.zeros(1, dtype=world_dt)
        self.l.hsref_dump_world(self.h, w, out.ctypes.data)
        return out[0]

    def dump_objects(self):
        out = np.zeros(8, dtype=object_dt)
        rc = self.l.hsref_dump_objects(self.h, out.ctypes.data)
        assert rc == 0, rc
        return out

    def bench(self, warmup, steps, action_seed=0):
        return self.l.hsref_bench(self.h, warmup, steps, action_seed)
