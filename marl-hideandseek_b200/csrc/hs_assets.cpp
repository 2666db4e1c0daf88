// Init-time asset table: collision hulls -> half-edge meshes, face planes,
// AABBs and mass data for the 8 SimObjects (host side, runs once per hs_sim).
//
// Replaces, for this environment's fixed asset set, the reference pipeline
//   loadPhysicsObjects                          src/mgr.cpp:441-588
//   OBJ import + first-use vertex remap         M/src/importer/obj.cpp (meshopt remap)
//   buildHalfEdgeMesh / computeNewellPlane      M/src/physics/physics_assets.cpp:212-253, 638-747
//   setupRigidBodyAABBsAndPrimitives            physics_assets.cpp:1214-1266
// The hull geometry below is this repo's own description of the five convex
// shapes (2x2x2 cube, 2x2x2.5 wall block, 2x2x2 agent block, wedge ramp,
// 8x1.5x2 box): corner lists in the order the importer leaves them
// (first use by the face list) and CCW outward faces.
//
// Inverse inertia: the reference integrates tetrahedra and runs 8 approximate
// Jacobi sweeps (physics_assets.cpp:795-1134); for this fixed asset set the
// resulting diagonals are embedded as exact bit patterns and pinned against the
// oracle's ObjectManager dump in tests/test_assets.py.
#include "hs_state.cuh"

#include <cmath>
#include <cstring>
#include <map>
#include <utility>
#include <vector>

namespace hs {

namespace {

struct HullSrc {
    std::vector<Vec3> verts;
    std::vector<std::vector<int>> faces;
};

// corners: T = top z, B = bottom z
HullSrc blockA(float hx, float hy, float zb, float zt)
{
    HullSrc h;
    h.verts = {{hx, hy, zt}, {-hx, hy, zt}, {-hx, -hy, zt}, {hx, -hy, zt},
               {hx, -hy, zb}, {-hx, -hy, zb}, {-hx, hy, zb}, {hx, hy, zb}};
    h.faces = {{0, 1, 2, 3}, {4, 3, 2, 5}, {5, 2, 1, 6}, {6, 7, 4, 5}, {7, 0, 3, 4}, {6, 1, 0, 7}};
    return h;
}

HullSrc blockB(float hx, float hy, float zb, float zt)
{
    HullSrc h;
    h.verts = {{-hx, -hy, zb}, {-hx, -hy, zt}, {-hx, hy, zt}, {-hx, hy, zb},
               {hx, hy, zt}, {hx, hy, zb}, {hx, -hy, zt}, {hx, -hy, zb}};
    h.faces = {{0, 1, 2, 3}, {3, 2, 4, 5}, {5, 4, 6, 7}, {7, 6, 1, 0}, {3, 5, 7, 0}, {4, 2, 1, 6}};
    return h;
}

HullSrc wedge()
{
    HullSrc h;
    h.verts = {{-1, 1, -1}, {1, 1, -1}, {1, -2, -1}, {-1, -2, -1}, {-1, 1, 1}, {1, 1, 1}};
    h.faces = {{0, 1, 2, 3}, {0, 4, 5, 1}, {2, 5, 4, 3}, {1, 5, 2}, {3, 4, 0}};
    return h;
}

float bitsToFloat(uint32_t b)
{
    float f;
    std::memcpy(&f, &b, 4);
    return f;
}

// RTCD 12.4.2 Newell plane, accumulation order as physics_assets.cpp:212-239.
void newellPlane(const HullSrc &h, const std::vector<int> &idx, float out[4])
{
    Vec3 centroid = {0, 0, 0}, n = {0, 0, 0};
    int cnt = 0;
    for (size_t i = idx.size() - 1, j = 0; j < idx.size(); i = j, j++) {
        Vec3 vi = h.verts[idx[i]], vj = h.verts[idx[j]];
        n.x += (vi.y - vj.y) * (vi.z + vj.z);
        n.y += (vi.z - vj.z) * (vi.x + vj.x);
        n.z += (vi.x - vj.x) * (vi.y + vj.y);
        centroid += vj;
        cnt++;
    }
    centroid = centroid / float(cnt);
    n = normalize(n);
    out[0] = n.x; out[1] = n.y; out[2] = n.z;
    out[3] = dot(centroid, n);
}

void buildHull(const HullSrc &src, HullData &H, float aabb[6])
{
    std::memset(&H, 0, sizeof(H));
    H.numVerts = (int)src.verts.size();
    H.numFaces = (int)src.faces.size();
    for (int i = 0; i < H.numVerts; i++) {
        H.verts[i][0] = src.verts[i].x; H.verts[i][1] = src.verts[i].y; H.verts[i][2] = src.verts[i].z;
    }

    // Half-edge pairing: walking faces in order, the first time a directed edge
    // (a,b) is seen it and its twin (b,a) take the next two ids, so twins are
    // always {2k, 2k+1} (geo.inl twinIDX) and edge k == half-edge 2k.
    std::map<std::pair<int, int>, int> edgeToHedge;
    int assigned = 0;
    for (int f = 0; f < H.numFaces; f++) {
        const std::vector<int> &idx = src.faces[f];
        int n = (int)idx.size();
        newellPlane(src, idx, H.planes[f]);
        for (int v = 0; v < n; v++) {
            int a = idx[v], b = idx[(v + 1) % n], c = idx[(v + 2) % n];
            auto it = edgeToHedge.find({a, b});
            if (it == edgeToHedge.end()) {
                edgeToHedge[{a, b}] = assigned;
                edgeToHedge[{b, a}] = assigned + 1;
                it = edgeToHedge.find({a, b});
                assigned += 2;
            }
            int hedge = it->second;
            if (v == 0) H.faceBase[f] = (uint8_t)hedge;
            auto nx = edgeToHedge.find({b, c});
            int next = nx == edgeToHedge.end() ? assigned : nx->second;
            H.hedgeNext[hedge] = (uint8_t)next;
            H.hedgeRoot[hedge] = (uint8_t)a;
            H.hedgeFace[hedge] = (uint8_t)f;
        }
    }
    H.numHedges = assigned;

    for (int k = 0; k < 3; k++) { aabb[k] = FLT_MAX; aabb[3 + k] = -FLT_MAX; }
    for (const Vec3 &v : src.verts) {
        float c[3] = {v.x, v.y, v.z};
        for (int k = 0; k < 3; k++) {
            if (c[k] < aabb[k]) aabb[k] = c[k];
            if (c[k] > aabb[3 + k]) aabb[3 + k] = c[k];
        }
    }
}

}  // namespace

void buildAssetTable(AssetTable &T)
{
    std::memset(&T, 0, sizeof(T));

    struct Spec { int obj; HullSrc src; float invMass, muS, muD; uint32_t invI[3]; };
    // invMass / friction: src/mgr.cpp:476-559; agents: inverse inertia x,y zeroed (mgr.cpp:577-584)
    std::vector<Spec> specs;
    specs.push_back({OBJ_CUBE, blockA(1, 1, -1, 1), 0.5f, 0.5f, 2.f, {1061158908u, 1061158908u, 1061158909u}});
    specs.push_back({OBJ_WALL, blockA(1, 1, 0, 2.5f), 0.f, 0.5f, 2.f, {0u, 0u, 0u}});
    specs.push_back({OBJ_HIDER, blockB(1, 1, -1, 1), 1.f, 0.5f, 16.f, {0u, 0u, 1069547515u}});
    specs.push_back({OBJ_SEEKER, blockB(1, 1, -1, 1), 1.f, 0.5f, 16.f, {0u, 0u, 1069547515u}});
    specs.push_back({OBJ_RAMP, wedge(), 0.5f, 0.5f, 1.f, {1060190995u, 1065748618u, 1057780132u}});
    specs.push_back({OBJ_BOX, blockB(4, 0.75f, -1, 1), 0.5f, 0.5f, 4.f, {1064682126u, 1035252915u, 1035565742u}});

    for (const Spec &s : specs) {
        ObjectData &O = T.obj[s.obj];
        O.primType = 2;
        buildHull(s.src, T.hull[s.obj], O.aabb);
        O.invMass = s.invMass;
        for (int k = 0; k < 3; k++) O.invInertia[k] = bitsToFloat(s.invI[k]);
        O.muS = s.muS;
        O.muD = s.muD;
    }

    // Sphere (never instantiated by this environment) and ground plane.
    ObjectData &sp = T.obj[OBJ_SPHERE];
    sp.primType = 1;
    for (int k = 0; k < 3; k++) { sp.aabb[k] = -1.f; sp.aabb[3 + k] = 1.f; }
    sp.invMass = 1.f;
    for (int k = 0; k < 3; k++) sp.invInertia[k] = bitsToFloat(1075838975u);
    sp.muS = 0.5f; sp.muD = 0.5f;

    ObjectData &pl = T.obj[OBJ_PLANE];
    pl.primType = 4;
    pl.aabb[0] = -FLT_MAX; pl.aabb[1] = -FLT_MAX; pl.aabb[2] = -FLT_MAX;
    pl.aabb[3] = FLT_MAX; pl.aabb[4] = FLT_MAX; pl.aabb[5] = 0.f;
    pl.invMass = 0.f;
    pl.muS = 2.f; pl.muD = 2.f;
}

}  // namespace hs
