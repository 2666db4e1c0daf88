// Init-time asset table: collision hulls -> half-edge meshes, face planes, AABBs, mass and
// diagonalised inertia for the 8 SimObjects (host side, runs once per hs_sim).
//
// The same pipeline as the reference's, restated:
//   loadPhysicsObjects (which OBJ per object, masses, friction, agent hack)   src/mgr.cpp:441-588
//   OBJ import: corners un-indexed, vertices re-indexed by first use           M/src/importer/obj.cpp:302-460
//   buildHalfEdgeMesh / computeNewellPlane                 M/src/physics/physics_assets.cpp:212-253, 638-747
//   computeMassProperties (tetrahedron covariance sums)    physics_assets.cpp:956-1134
//   diagonalizeInertiaTensor (8 approximate Jacobi sweeps, McAdams et al.)     physics_assets.cpp:795-954
//   toMassData, setupRigidBodyAABBsAndPrimitives           physics_assets.cpp:1136-1266
// Geometry source: the five `*_collision.obj` files of the reference's data directory when one is
// given (HIDESEEK_DATA_DIR, or hs_dump_assets' argument), else this repo's own description of the
// same five convex shapes (2x2x2 cube, 2x2x2.5 wall block, 2x2x2 agent block, wedge ramp, 8x1.5x2
// box: corner lists in the order the importer leaves them and CCW outward faces).  Either way every
// derived number — planes, half-edges, AABBs, inverse inertia — is COMPUTED here in the reference's
// operation order; tests/test_assets_cpu.py compares the whole table bit for bit with the
// reference's ObjectManager (oracle hsref_dump_objects), for both sources.
#include "hs_state.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
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

// ---- OBJ import (positions + polygon faces) ------------------------------------------------------
// M/src/importer/obj.cpp: every face corner becomes a vertex ("un-indexing"), then identical
// vertices are merged and numbered in order of first use (meshopt_generateVertexRemapMulti).
// The collision files carry positions only, so a vertex is its position.
bool loadHullObj(const std::string &path, HullSrc &out)
{
    FILE *f = fopen(path.c_str(), "r");
    if (!f) return false;
    std::vector<Vec3> pos;
    std::vector<std::vector<int>> faces;
    char line[512];
    while (fgets(line, sizeof(line), f)) {
        if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
            char *p = line + 2;
            Vec3 v;
            v.x = strtof(p, &p); v.y = strtof(p, &p); v.z = strtof(p, &p);
            pos.push_back(v);
        } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
            std::vector<int> idx;
            char *p = line + 2;
            while (*p) {
                while (*p == ' ' || *p == '\t') p++;
                if (*p == 0 || *p == '\n' || *p == '\r') break;
                long i = strtol(p, &p, 10);
                while (*p && *p != ' ' && *p != '\t' && *p != '\n' && *p != '\r') p++;    // skip /vt/vn
                if (i < 0) i = (long)pos.size() + 1 + i;
                if (i < 1 || i > (long)pos.size()) { fclose(f); return false; }
                idx.push_back(int(i - 1));
            }
            if (idx.size() >= 3) faces.push_back(idx);
        }
    }
    fclose(f);
    out.verts.clear();
    out.faces.clear();
    std::vector<int> remap(pos.size(), -1);
    for (const auto &fc : faces) {
        std::vector<int> nf;
        for (int i : fc) {
            if (remap[i] < 0) {
                // merge bit-identical positions (what the remap's hash compares)
                int found = -1;
                for (size_t k = 0; k < out.verts.size(); k++)
                    if (std::memcmp(&out.verts[k], &pos[i], sizeof(Vec3)) == 0) { found = int(k); break; }
                if (found < 0) { found = int(out.verts.size()); out.verts.push_back(pos[i]); }
                remap[i] = found;
            }
            nf.push_back(remap[i]);
        }
        out.faces.push_back(nf);
    }
    return (int)out.verts.size() <= kMaxHullVerts && (int)out.faces.size() <= kMaxHullFaces && !out.faces.empty();
}

// ---- mass properties (physics_assets.cpp:956-1134) -------------------------------------------------
struct Sym3 { float d[3]; float o[3]; };      // diag {11, 22, 33}, off {12, 13, 23}

// McAdams et al. 2011, Algorithm 2 (physics_assets.cpp:808-834); rsqrtApprox: math.inl:14-24
float rsqrtApprox(float x)
{
    uint32_t xb;
    std::memcpy(&xb, &x, 4);
    uint32_t yb = 0x5F1FFFF9u - (xb >> 1);
    float y;
    std::memcpy(&y, &yb, 4);
    return y * (0.703952253f * (2.38924456f - x * y * y));
}

void approxGivens(const Sym3 &m, float *chOut, float *shOut)
{
    const float gamma = 5.82842712474619f, c_star = 0.9238795325112867f, s_star = 0.3826834323650898f;
    float a11 = m.d[0], a12 = m.o[0], a22 = m.d[1];
    float ch = 2.f * (a11 - a22);
    float sh = a12;
    float sh2 = sh * sh;
    if (sh2 < 1e-20f) { *chOut = 1.f; *shOut = 0.f; return; }
    float ch2 = ch * ch;
    bool b = (gamma * sh2) < ch2;
    float omega = rsqrtApprox(ch2 + sh2);
    *chOut = b ? (omega * ch) : c_star;
    *shOut = b ? (omega * sh) : s_star;
}

Sym3 jacobiConjugate(const Sym3 &m, float ch, float sh)
{
    float ch2 = ch * ch, sh2 = sh * sh;
    float q_scale = ch2 + sh2;
    float q11 = (ch2 - sh2) / q_scale;
    float q12 = (-2.f * sh * ch) / q_scale;
    float q21 = (2.f * sh * ch) / q_scale;
    float q22 = (ch2 - sh2) / q_scale;
    float m11 = m.d[0], m22 = m.d[1], m33 = m.d[2], m12 = m.o[0], m13 = m.o[1], m23 = m.o[2];
    float m11q11_m12q21 = m11 * q11 + m12 * q21;
    float m11q12_m12q22 = m11 * q12 + m12 * q22;
    float m12q11_m22q21 = m12 * q11 + m22 * q21;
    float m12q12_m22q22 = m12 * q12 + m22 * q22;
    Sym3 r;
    r.d[0] = q11 * m11q11_m12q21 + q21 * m12q11_m22q21;
    r.d[1] = q12 * m11q12_m12q22 + q22 * m12q12_m22q22;
    r.d[2] = m33;
    r.o[0] = q12 * m11q11_m12q21 + q22 * m12q11_m22q21;
    r.o[1] = m13 * q11 + m23 * q21;
    r.o[2] = m13 * q12 + m23 * q22;
    return r;
}

void diagonalizeInertia(const Sym3 &m, float outDiag[3])
{
    Sym3 cur = m;
    Quat acc = {1.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < 8; i++) {
        float ch1, sh1, ch2, sh2, ch3, sh3;
        approxGivens(cur, &ch1, &sh1);
        cur = jacobiConjugate(cur, ch1, sh1);
        std::swap(cur.d[1], cur.d[2]);
        std::swap(cur.o[0], cur.o[1]);
        approxGivens(cur, &ch2, &sh2);
        cur = jacobiConjugate(cur, ch2, sh2);
        std::swap(cur.d[0], cur.d[2]);
        std::swap(cur.o[0], cur.o[2]);
        approxGivens(cur, &ch3, &sh3);
        cur = jacobiConjugate(cur, ch3, sh3);
        Sym3 t = {{cur.d[2], cur.d[0], cur.d[1]}, {cur.o[1], cur.o[2], cur.o[0]}};
        cur = t;
        acc = Quat{ch1, 0, 0, sh1} * Quat{ch2, 0, sh2, 0} * Quat{ch3, sh3, 0, 0} * acc;
    }
    Quat fin = qnormalize(acc);
    Mat3 q = mat3FromQuat(fin);
    float m11 = m.d[0], m22 = m.d[1], m33 = m.d[2], m12 = m.o[0], m13 = m.o[1], m23 = m.o[2];
    float q11 = q.c0.x, q21 = q.c0.y, q31 = q.c0.z;
    float q12 = q.c1.x, q22 = q.c1.y, q32 = q.c1.z;
    float q13 = q.c2.x, q23 = q.c2.y, q33 = q.c2.z;
    outDiag[0] = q11 * (m11 * q11 + m12 * q21 + m13 * q31) + q21 * (m12 * q11 + m22 * q21 + m23 * q31) +
                 q31 * (m13 * q11 + m23 * q21 + m33 * q31);
    outDiag[1] = q12 * (m11 * q12 + m12 * q22 + m13 * q32) + q22 * (m12 * q12 + m22 * q22 + m23 * q32) +
                 q32 * (m13 * q12 + m23 * q22 + m33 * q32);
    outDiag[2] = q13 * (m11 * q13 + m12 * q23 + m13 * q33) + q23 * (m12 * q13 + m22 * q23 + m23 * q33) +
                 q33 * (m13 * q13 + m23 * q23 + m33 * q33);
}

struct MassAcc {
    Sym3 C = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    float m = 0.f;
    Vec3 x = {0.f, 0.f, 0.f};
};

// one tetrahedron (origin, v1, v2, v3): covariance about the origin, mass, centre of mass
void accumulateTet(MassAcc &A, Vec3 e1, Vec3 e2, Vec3 e3)
{
    const float xd = 1.f / 60.f, xo = 1.f / 120.f;       // C_canonical
    float det = e1.x * (e2.y * e3.z - e3.y * e2.z) - e1.y * (e2.x * e3.z - e3.x * e2.z) + e1.z * (e2.x * e3.y - e3.x * e2.y);
    // Symmetric3x3::AXAT(A, X), A's columns = e1 e2 e3 (math.inl:1279-1315)
    float a11 = e1.x, a21 = e1.y, a31 = e1.z, a12 = e2.x, a22 = e2.y, a32 = e2.z, a13 = e3.x, a23 = e3.y, a33 = e3.z;
    float x11 = xd, x22 = xd, x33 = xd, x12 = xo, x13 = xo, x23 = xo;
    float r1a = a11 * x11 + a12 * x12 + a13 * x13;
    float r1b = a11 * x12 + a12 * x22 + a13 * x23;
    float r1c = a11 * x13 + a12 * x23 + a13 * x33;
    float r2a = a21 * x11 + a22 * x12 + a23 * x13;
    float r2b = a21 * x12 + a22 * x22 + a23 * x23;
    float r2c = a21 * x13 + a22 * x23 + a23 * x33;
    Sym3 C;
    C.d[0] = a11 * r1a + a12 * r1b + a13 * r1c;
    C.d[1] = a21 * r2a + a22 * r2b + a23 * r2c;
    C.d[2] = a31 * (a31 * x11 + a32 * x12 + a33 * x13) + a32 * (a31 * x12 + a32 * x22 + a33 * x23) +
             a33 * (a31 * x13 + a32 * x23 + a33 * x33);
    C.o[0] = a21 * r1a + a22 * r1b + a23 * r1c;
    C.o[1] = a31 * r1a + a32 * r1b + a33 * r1c;
    C.o[2] = a31 * r2a + a32 * r2b + a33 * r2c;
    for (int k = 0; k < 3; k++) { C.d[k] *= det; C.o[k] *= det; }
    float volume = 1.f / 6.f * det;
    float m = volume * 1.f;
    Vec3 x = 0.25f * e1 + 0.25f * e2 + 0.25f * e3;
    float old = A.m;
    A.m += m;
    A.x = (x * m + A.x * old) / A.m;
    for (int k = 0; k < 3; k++) { A.C.d[k] += C.d[k]; A.C.o[k] += C.o[k]; }
}

// centre the covariance, turn it into the inertia tensor of unit total mass, diagonalise
void finishMass(MassAcc &A, float inertiaDiag[3])
{
    Vec3 x = A.x, dx = -A.x;
    Sym3 a, b;
    a.d[0] = 2.f * (x.x * dx.x); a.d[1] = 2.f * (x.y * dx.y); a.d[2] = 2.f * (x.z * dx.z);
    a.o[0] = x.x * dx.y + x.y * dx.x; a.o[1] = x.x * dx.z + x.z * dx.x; a.o[2] = x.y * dx.z + x.z * dx.y;
    b.d[0] = dx.x * dx.x; b.d[1] = dx.y * dx.y; b.d[2] = dx.z * dx.z;
    b.o[0] = dx.x * dx.y; b.o[1] = dx.x * dx.z; b.o[2] = dx.y * dx.z;
    Sym3 C;
    for (int k = 0; k < 3; k++) {
        C.d[k] = A.C.d[k] + (a.d[k] + b.d[k]) * A.m;
        C.o[k] = A.C.o[k] + (a.o[k] + b.o[k]) * A.m;
    }
    float tr = C.d[0] + C.d[1] + C.d[2];
    Sym3 I;
    for (int k = 0; k < 3; k++) { I.d[k] = tr - C.d[k]; I.o[k] = 0.f - C.o[k]; }
    float invMass = 1.f / A.m;
    for (int k = 0; k < 3; k++) { I.d[k] *= invMass; I.o[k] *= invMass; }
    diagonalizeInertia(I, inertiaDiag);
}

void hullInverseInertia(const HullData &H, float invMass, float out[3])
{
    MassAcc A;
    for (int f = 0; f < H.numFaces; f++) {
        int root = H.faceBase[f];
        Vec3 v1 = {H.verts[H.hedgeRoot[root]][0], H.verts[H.hedgeRoot[root]][1], H.verts[H.hedgeRoot[root]][2]};
        int cur = H.hedgeNext[root];
        while (true) {
            int next = H.hedgeNext[cur];
            if (next == root) break;
            Vec3 v2 = {H.verts[H.hedgeRoot[cur]][0], H.verts[H.hedgeRoot[cur]][1], H.verts[H.hedgeRoot[cur]][2]};
            Vec3 v3 = {H.verts[H.hedgeRoot[next]][0], H.verts[H.hedgeRoot[next]][1], H.verts[H.hedgeRoot[next]][2]};
            accumulateTet(A, v1, v2, v3);
            cur = next;
        }
    }
    float diag[3];
    finishMass(A, diag);
    for (int k = 0; k < 3; k++) out[k] = invMass / diag[k];
}

}  // namespace

const char *buildAssetTable(AssetTable &T, const char *dataDir)
{
    std::memset(&T, 0, sizeof(T));

    // which collision hull, inverse mass and friction per object: src/mgr.cpp:457-559
    struct Spec { int obj; const char *file; HullSrc builtin; float invMass, muS, muD; bool agent; };
    std::vector<Spec> specs;
    specs.push_back({OBJ_CUBE, "cube_collision.obj", blockA(1, 1, -1, 1), 0.5f, 0.5f, 2.f, false});
    specs.push_back({OBJ_WALL, "wall_collision.obj", blockA(1, 1, 0, 2.5f), 0.f, 0.5f, 2.f, false});
    specs.push_back({OBJ_HIDER, "agent_collision.obj", blockB(1, 1, -1, 1), 1.f, 0.5f, 16.f, true});
    specs.push_back({OBJ_SEEKER, "agent_collision.obj", blockB(1, 1, -1, 1), 1.f, 0.5f, 16.f, true});
    specs.push_back({OBJ_RAMP, "ramp_collision.obj", wedge(), 0.5f, 0.5f, 1.f, false});
    specs.push_back({OBJ_BOX, "elongated_collision.obj", blockB(4, 0.75f, -1, 1), 0.5f, 0.5f, 4.f, false});

    static thread_local std::string err;
    for (Spec &s : specs) {
        HullSrc src = s.builtin;
        if (dataDir && dataDir[0]) {
            std::string path = std::string(dataDir) + "/" + s.file;
            if (!loadHullObj(path, src)) {
                err = "cannot read a convex hull of at most 8 vertices / 6 faces from " + path;
                return err.c_str();
            }
        }
        ObjectData &O = T.obj[s.obj];
        O.primType = 2;
        buildHull(src, T.hull[s.obj], O.aabb);
        O.invMass = s.invMass;
        hullInverseInertia(T.hull[s.obj], s.invMass, O.invInertia);
        if (s.agent) { O.invInertia[0] = 0.f; O.invInertia[1] = 0.f; }      // the HACK of src/mgr.cpp:577-584
        O.muS = s.muS;
        O.muD = s.muD;
    }

    // Sphere (never instantiated by this environment): covariance diag r^2/5 of a unit-mass ball
    // (physics_assets.cpp:1003-1016), through the same centring / diagonalisation.
    ObjectData &sp = T.obj[OBJ_SPHERE];
    sp.primType = 1;
    for (int k = 0; k < 3; k++) { sp.aabb[k] = -1.f; sp.aabb[3 + k] = 1.f; }
    sp.invMass = 1.f;
    {
        MassAcc A;
        A.m += 1.f;
        float v = 1.f / 5.f * 1.f * 1.f;
        for (int k = 0; k < 3; k++) A.C.d[k] += v;
        float diag[3];
        finishMass(A, diag);
        for (int k = 0; k < 3; k++) sp.invInertia[k] = sp.invMass / diag[k];
    }
    sp.muS = 0.5f; sp.muD = 0.5f;

    // ground plane: infinite inertia (physics_assets.cpp:1017-1026) -> inverse inertia 0 / inf
    ObjectData &pl = T.obj[OBJ_PLANE];
    pl.primType = 4;
    pl.aabb[0] = -FLT_MAX; pl.aabb[1] = -FLT_MAX; pl.aabb[2] = -FLT_MAX;
    pl.aabb[3] = FLT_MAX; pl.aabb[4] = FLT_MAX; pl.aabb[5] = 0.f;
    pl.invMass = 0.f;
    pl.muS = 2.f; pl.muD = 2.f;
    return "";
}

}  // namespace hs
