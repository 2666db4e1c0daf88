"""Driver for the agent-view oracle (oracle/views_oracle.c -> oracle/_ref/libviews_oracle.so).

TEST INFRASTRUCTURE ONLY.  The oracle is a brute-force CPU restatement of Madrona's software
ray caster (parity unpinned: the reference's own is device-only code); the product's
render-mesh table is flattened into the per-object triangle lists it takes.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "_ref", "libviews_oracle.so")
REF_DATA = os.path.join(ROOT, "oracle", "_ref", "data")

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def have_ref_data():
    return os.path.exists(os.path.join(REF_DATA, "agent_render.obj"))


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.hs_oracle_render_view.argtypes = [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 6 + \
            [C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return _lib


class Meshes:
    """Per-object flat triangle lists out of a ViewMeshTable record."""

    def __init__(self, table):
        tris, sub_of, first, num = [], [], np.zeros(8, np.int32), np.zeros(8, np.int32)
        for o in range(8):
            first[o] = len(tris)
            for s in range(int(table["obj_num_sub"][o])):
                sb = table["sub"][int(table["obj_first_sub"][o]) + s]
                for c in range(int(sb["count"])):
                    cb = table["cluster"][int(sb["first"]) + c]
                    for t in range(int(cb["count"])):
                        tris.append(table["tri"][int(cb["first"]) + t])
                        sub_of.append(s)
            num[o] = len(tris) - first[o]
        self.tris = np.ascontiguousarray(np.array(tris, dtype=np.float32).reshape(-1, 9))
        self.tri_sub = np.array(sub_of, dtype=np.int32)
        self.first, self.num = first, num
        self.color = np.ascontiguousarray(table["sub_color"], dtype=np.float32)


def render(meshes, inst_obj, inst_pos, inst_rot, inst_scale, cam_pos, cam_rot, width=64, height=64, fov=100.0):
    inst_obj = np.ascontiguousarray(inst_obj, dtype=np.int32)
    inst_pos = np.ascontiguousarray(inst_pos, dtype=np.float32)
    inst_rot = np.ascontiguousarray(inst_rot, dtype=np.float32)
    inst_scale = np.ascontiguousarray(inst_scale, dtype=np.float32)
    cam_pos = np.ascontiguousarray(cam_pos, dtype=np.float32)
    cam_rot = np.ascontiguousarray(cam_rot, dtype=np.float32)
    depth = np.zeros((height, width), dtype=np.float32)
    rgb = np.zeros((height, width, 4), dtype=np.uint8)
    p = lambda a: a.ctypes.data
    rc = lib().hs_oracle_render_view(p(meshes.tris), p(meshes.tri_sub), p(meshes.first), p(meshes.num),
                                     p(meshes.color), len(inst_obj), p(inst_obj), p(inst_pos), p(inst_rot),
                                     p(inst_scale), p(cam_pos), p(cam_rot), fov, width, height, p(depth), p(rgb))
    assert rc == 0
    return depth, rgb


def render_world_view(meshes, dump, iface, width=64, height=64):
    """Oracle image of agent interface `iface` of a world given the product's hs_world_dump."""
    n = int(dump["num_bodies"])
    b = dump["bodies"][:n]
    slot = int(dump["iface_slot"][iface])
    if slot < 0:
        return np.zeros((height, width), np.float32), np.zeros((height, width, 4), np.uint8)
    return render(meshes, b["obj"], b["pos"], b["rot"], b["scale"], b["pos"][slot], b["rot"][slot], width, height)


def compare(depth, rgb, ref_depth, ref_rgb, tol=1e-4, max_flip_frac=2e-3):
    """Depth within tol*max(1,depth) and identical colour, except for at most max_flip_frac of
    the pixels (silhouette pixels where a 1-ulp difference flips hit/miss or the hit instance)."""
    d = np.abs(depth - ref_depth)
    bad = (d > tol * np.maximum(1.0, np.abs(ref_depth))) | (rgb != ref_rgb).any(-1)
    frac = bad.mean()
    ok = ~bad
    return frac <= max_flip_frac, frac, float(d[ok].max()) if ok.any() else 0.0
