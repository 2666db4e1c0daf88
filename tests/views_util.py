"""Driver for the agent-view oracle (oracle/views_oracle.c -> oracle/_ref/libviews_oracle.so).

TEST INFRASTRUCTURE ONLY.  The oracle's per-instance ray / triangle stage is the REFERENCE's own
MeshBVH::traceRay (oracle/_ref/libhs_ref.so, hsref_mesh_trace: M/include/madrona/mesh_bvh.inl
compiled unchanged) whenever that library is built — `Meshes.bind()` installs it; camera rays,
the instance loop and shading stay a restatement (views_oracle.c header).  Without libhs_ref the
oracle falls back to its brute-force restatement.  The product's render-mesh table is flattened
into the per-object triangle lists both take.
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
        _lib.hs_oracle_set_uv.argtypes = [C.c_void_p, C.c_void_p]
        _lib.hs_oracle_set_texture.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int]
    return _lib


TEXTURE_FILES = ["green_grid.png", "smile.png", "red_smile.png"]      # src/mgr.cpp:651-658


class Meshes:
    """Per-object flat triangle lists out of a ViewMeshTable record (+ texture coordinates and
    images when the ViewTexTable of a data directory is given)."""

    def __init__(self, table, tex_table=None, data_dir=None):
        tris, sub_of, first, num = [], [], np.zeros(8, np.int32), np.zeros(8, np.int32)
        uvs = []
        for o in range(8):
            first[o] = len(tris)
            for s in range(int(table["obj_num_sub"][o])):
                sb = table["sub"][int(table["obj_first_sub"][o]) + s]
                for c in range(int(sb["count"])):
                    cb = table["cluster"][int(sb["first"]) + c]
                    for t in range(int(cb["count"])):
                        tris.append(table["tri"][int(cb["first"]) + t])
                        sub_of.append(s)
                        if tex_table is not None:
                            uvs.append(tex_table["tri_uv"][int(cb["first"]) + t])
            num[o] = len(tris) - first[o]
        self.tris = np.ascontiguousarray(np.array(tris, dtype=np.float32).reshape(-1, 9))
        self.tri_sub = np.array(sub_of, dtype=np.int32)
        self.first, self.num = first, num
        self.color = np.ascontiguousarray(table["sub_color"], dtype=np.float32)
        self.uv = self.sub_tex = None
        self.images = []
        if tex_table is not None:
            from PIL import Image
            self.uv = np.ascontiguousarray(np.array(uvs, dtype=np.float32).reshape(-1, 6))
            self.sub_tex = np.ascontiguousarray(tex_table["sub_tex"], dtype=np.int32)
            for f in TEXTURE_FILES:
                self.images.append(np.ascontiguousarray(np.array(Image.open(os.path.join(data_dir, f)).convert("RGBA"), dtype=np.uint8)))

    def _ref_meshes(self):
        """One reference MeshBVH per object (built once), or False when libhs_ref is absent."""
        if getattr(self, "_ref", None) is None:
            import refsim
            if not refsim.available() or os.environ.get("HIDESEEK_VIEWS_ORACLE") == "restated":
                self._ref = False
            else:
                rl = refsim.lib()
                rl.hsref_mesh_create.restype = C.c_void_p
                rl.hsref_mesh_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
                handles = (C.c_void_p * 8)()
                for o in range(8):
                    n, f = int(self.num[o]), int(self.first[o])
                    if n == 0:
                        continue
                    tri = np.ascontiguousarray(self.tris[f:f + n])
                    uv = None if self.uv is None else np.ascontiguousarray(self.uv[f:f + n])
                    handles[o] = rl.hsref_mesh_create(tri.ctypes.data, None if uv is None else uv.ctypes.data, n)
                self._ref = (handles, C.cast(rl.hsref_mesh_trace, C.c_void_p))
        return self._ref

    def pinned(self):
        """True when the ray / triangle stage of renders with this mesh set is reference code."""
        return bool(self._ref_meshes())

    def bind(self):
        """Point the oracle's tracer and texturing state at this mesh set (or clear them)."""
        l = lib()
        l.hs_oracle_set_tracer.argtypes = [C.c_void_p, C.c_void_p]
        ref = self._ref_meshes()
        if ref:
            l.hs_oracle_set_tracer(ref[1], C.cast(ref[0], C.c_void_p))
        else:
            l.hs_oracle_set_tracer(None, None)
        if self.uv is None:
            l.hs_oracle_set_uv(None, None)
            return
        l.hs_oracle_set_uv(self.uv.ctypes.data, self.sub_tex.ctypes.data)
        for i, im in enumerate(self.images):
            l.hs_oracle_set_texture(i, im.ctypes.data, im.shape[1], im.shape[0])


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
    meshes.bind()
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


def compare(depth, rgb, ref_depth, ref_rgb, tol=1e-4, max_flip_frac=2e-3, rgb_tol=0):
    """Depth within tol*max(1,depth) and colour bytes within rgb_tol (0 without textures: the
    hardware bilinear filter is only restated to its documented 8 fractional weight bits),
    except for at most max_flip_frac of the pixels (silhouette pixels where a 1-ulp difference
    flips hit/miss or the hit instance)."""
    d = np.abs(depth - ref_depth)
    bad = (d > tol * np.maximum(1.0, np.abs(ref_depth))) | (np.abs(rgb.astype(np.int16) - ref_rgb.astype(np.int16)) > rgb_tol).any(-1)
    frac = bad.mean()
    ok = ~bad
    return frac <= max_flip_frac, frac, float(d[ok].max()) if ok.any() else 0.0
