"""Row T3 of SURVEY.md §8: the collision asset table.  hs_create derives it from the five
`*_collision.obj` files (OBJ import with first-use vertex numbering, half-edge mesh, Newell
planes, AABBs, tetrahedron mass properties, 8 approximate Jacobi sweeps) or from the built-in
hull descriptions; both must equal the reference's ObjectManager — what every reference world
reads at run time (oracle hsref_dump_objects) — BIT FOR BIT, field by field."""
import ctypes as C
import os

import numpy as np
import pytest

import refsim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "oracle", "_ref", "data")

hull_dt = np.dtype([("num_verts", "i4"), ("num_faces", "i4"), ("num_hedges", "i4"), ("pad", "i4"),
                    ("verts", "f4", (8, 3)), ("planes", "f4", (6, 4)), ("hedge_next", "u1", 24),
                    ("hedge_root", "u1", 24), ("hedge_face", "u1", 24), ("face_base", "u1", 6), ("pad2", "u1", 2)])
obj_dt = np.dtype([("prim_type", "i4"), ("aabb", "f4", 6), ("inv_mass", "f4"), ("inv_inertia", "f4", 3),
                   ("mu_s", "f4"), ("mu_d", "f4")])
table_dt = np.dtype([("obj", obj_dt, 8), ("hull", hull_dt, 8)])


def _table(data_dir):
    import gpu_hideseek as gh
    lib = gh.load_library()
    assert lib.hs_asset_table_bytes() == table_dt.itemsize, (lib.hs_asset_table_bytes(), table_dt.itemsize)
    out = np.zeros(1, dtype=table_dt)
    lib.hs_dump_assets.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
    rc = lib.hs_dump_assets((data_dir or "").encode(), out.ctypes.data, table_dt.itemsize)
    assert rc == 0, lib.hs_last_error()
    return out[0]


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _compare(T, ref):
    for o in range(8):
        r, m, h = ref[o], T["obj"][o], T["hull"][o]
        assert int(m["prim_type"]) == int(r["prim_type"]), o
        assert np.array_equal(_bits(m["aabb"]), _bits(r["prim_aabb"])), (o, "prim aabb", m["aabb"], r["prim_aabb"])
        assert np.array_equal(_bits(m["aabb"]), _bits(r["obj_aabb"])), (o, "obj aabb")
        for f, g in (("inv_mass", "inv_mass"), ("mu_s", "mu_s"), ("mu_d", "mu_d")):
            assert _bits(m[f]) == _bits(r[g]), (o, f, float(m[f]), float(r[g]))
        assert np.array_equal(_bits(m["inv_inertia"]), _bits(r["inv_inertia"])), \
            (o, "inverse inertia", m["inv_inertia"], r["inv_inertia"], _bits(m["inv_inertia"]), _bits(r["inv_inertia"]))
        if int(r["prim_type"]) != 2:
            continue
        nv, nf, nh = int(r["num_verts"]), int(r["num_faces"]), int(r["num_hedges"])
        assert (int(h["num_verts"]), int(h["num_faces"]), int(h["num_hedges"])) == (nv, nf, nh), o
        assert np.array_equal(_bits(h["verts"][:nv]), _bits(r["verts"][:nv])), (o, "vertices")
        assert np.array_equal(_bits(h["planes"][:nf]), _bits(r["planes"][:nf])), (o, "face planes", h["planes"][:nf], r["planes"][:nf])
        assert np.array_equal(h["face_base"][:nf], r["face_base_hedge"][:nf]), (o, "face base half-edges")
        assert np.array_equal(h["hedge_next"][:nh], r["hedges"][:nh, 0]), (o, "half-edge next")
        assert np.array_equal(h["hedge_root"][:nh], r["hedges"][:nh, 1]), (o, "half-edge root vertex")
        assert np.array_equal(h["hedge_face"][:nh], r["hedges"][:nh, 2]), (o, "half-edge face")


@pytest.fixture(scope="module")
def ref_objects():
    if not refsim.available():
        pytest.skip("oracle/_ref not built")
    ref = refsim.RefSim(1, seed=0, num_threads=1)
    objs = ref.dump_objects()
    ref.close()
    return objs


def test_table_from_collision_objs_equals_reference(ref_objects):
    if not os.path.exists(os.path.join(DATA, "cube_collision.obj")):
        pytest.skip("reference collision hulls not present")
    _compare(_table(DATA), ref_objects)


def test_builtin_table_equals_reference(ref_objects):
    _compare(_table(None), ref_objects)


def test_builtin_table_equals_obj_table():
    if not os.path.exists(os.path.join(DATA, "cube_collision.obj")):
        pytest.skip("reference collision hulls not present")
    a, b = _table(DATA), _table(None)
    assert a.tobytes() == b.tobytes()


def test_missing_hull_file_is_an_error(tmp_path):
    import gpu_hideseek as gh
    lib = gh.load_library()
    out = np.zeros(1, dtype=table_dt)
    lib.hs_dump_assets.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
    assert lib.hs_dump_assets(str(tmp_path).encode(), out.ctypes.data, table_dt.itemsize) != 0
    assert b"cube_collision.obj" in lib.hs_last_error()
