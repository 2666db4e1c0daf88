// Narrowphase: SAT between convex hulls / hull vs ground plane, face clipping,
// manifold reduction.  One contact slot per candidate (slot index == candidate
// index) so the solver sees contacts in the reference's creation order.
//
// Behavioural contract (CPU path of M/src/physics/narrowphase.cpp):
//   makeHullState 151-222 · getHullDistanceFromPlane 296-337 · queryFaceDirections 339-365
//   isMinkowskiFace / edgeDistance / queryEdgeDirections 367-567 · findIncidentFace 569-615
//   clipPolygon 617-652 · doSAT 659-739 · doSATPlane 741-769
//   buildFaceContactManifold 771-879 · createFaceContact 881-989
//   createFacePlaneContact 991-1034 · shortestSegmentBetween / createEdgeContact 1037-1138
//   narrowphaseDispatch (HullHull, HullPlane) 1278-1325, 1459-1511 · generateContacts 1516-1680
//   runNarrowphase 1682-1907
#pragma once

#include "hs_bvh.cuh"

namespace hs {

struct WorldHull {
    Vec3 v[kMaxHullVerts];
    Plane p[kMaxHullFaces];
    Vec3 center;
    const HullData *h;
};

__device__ inline void makeWorldHull(WorldHull &W, const HullData &H, Vec3 t, Quat rot, Vec3 scale)
{
    W.h = &H;
    Mat3 r = mat3FromQuat(rot);
    Mat3 vt = {r.c0 * scale.x, r.c1 * scale.y, r.c2 * scale.z};
    Mat3 nt = {r.c0 * (1.f / scale.x), r.c1 * (1.f / scale.y), r.c2 * (1.f / scale.z)};
    Vec3 center = {0.f, 0.f, 0.f};
    for (int i = 0; i < H.numVerts; i++) {
        Vec3 wp = vt * v3(H.verts[i][0], H.verts[i][1], H.verts[i][2]) + t;
        W.v[i] = wp;
        center += wp;
    }
    W.center = center / float(H.numVerts);
    for (int i = 0; i < H.numFaces; i++) {
        Vec3 n = {H.planes[i][0], H.planes[i][1], H.planes[i][2]};
        Vec3 origin = vt * (n * H.planes[i][3]) + t;
        Vec3 tn = normalize(nt * n);
        W.p[i] = {tn, dot(tn, origin)};
    }
}

__device__ __forceinline__ float hullDistanceFromPlane(const Plane &pl, const WorldHull &h)
{
    float m = FLT_MAX;
    for (int i = 0; i < h.h->numVerts; i++) {
        float d = dot(h.v[i], pl.n);
        if (d < m) m = d;
    }
    return m - pl.d;
}

struct FaceQuery { float sep; int face; Plane plane; };

__device__ inline FaceQuery queryFaces(const WorldHull &a, const WorldHull &b)
{
    FaceQuery q;
    q.sep = -FLT_MAX; q.face = -1; q.plane = {{0.f, 0.f, 0.f}, 0.f};
    for (int f = 0; f < a.h->numFaces; f++) {
        Plane pl = a.p[f];
        float d = hullDistanceFromPlane(pl, b);
        if (d > q.sep) {
            q.sep = d; q.face = f; q.plane = pl;
            if (q.sep > 0.f) break;
        }
    }
    return q;
}

__device__ __forceinline__ bool isMinkowskiFace(Vec3 a, Vec3 b, Vec3 c, Vec3 d)
{
    Vec3 bxa = cross(b, a);
    Vec3 dxc = cross(d, c);
    float cba = dot(c, bxa), dba = dot(d, bxa), adc = dot(a, dxc), bdc = dot(b, dxc);
    return cba * dba < 0.0f && adc * bdc < 0.0f && cba * bdc > 0.0f;
}

struct EdgeQuery { float sep; Vec3 normal; int edgeA, edgeB; };

__device__ inline EdgeQuery queryEdges(const WorldHull &a, const WorldHull &b)
{
    EdgeQuery q;
    q.sep = -FLT_MAX; q.normal = {0.f, 0.f, 0.f}; q.edgeA = 0; q.edgeB = 0;
    const HullData &A = *a.h, &B = *b.h;
    const int na = A.numHedges / 2, nb = B.numHedges / 2;
    for (int ea = 0; ea < na; ea++) {
        int ha = ea * 2;
        Vec3 aN1 = a.p[A.hedgeFace[ha]].n, aN2 = a.p[A.hedgeFace[ha + 1]].n;
        Vec3 a1 = a.v[A.hedgeRoot[ha]], a2 = a.v[A.hedgeRoot[A.hedgeNext[ha]]];
        Vec3 dirA = a2 - a1;
        for (int eb = 0; eb < nb; eb++) {
            int hb = eb * 2;
            Vec3 bN1 = b.p[B.hedgeFace[hb]].n, bN2 = b.p[B.hedgeFace[hb + 1]].n;
            float sep;
            Vec3 normal = {0.f, 0.f, 0.f};
            if (isMinkowskiFace(aN1, aN2, -bN1, -bN2)) {
                Vec3 b1 = b.v[B.hedgeRoot[hb]], b2 = b.v[B.hedgeRoot[B.hedgeNext[hb]]];
                Vec3 dirB = b2 - b1;
                Vec3 c = cross(dirA, dirB);
                float len2 = length2(c);
                if (len2 == 0.f) {
                    sep = -FLT_MAX;
                } else {
                    float inv = 1.f / sqrtf(len2);
                    normal = c * inv;
                    if (dot(normal, a1 - a.center) < 0.0f) normal = -normal;
                    sep = dot(normal, b1 - a1);
                }
            } else {
                sep = -FLT_MAX;
            }
            if (sep > q.sep) {
                q.sep = sep; q.normal = normal; q.edgeA = ha; q.edgeB = hb;
                if (q.sep > 0.f) return q;
            }
        }
    }
    return q;
}

__device__ __forceinline__ int findIncidentFace(const WorldHull &h, Vec3 refNormal)
{
    float m = FLT_MAX;
    int face = -1;
    for (int f = 0; f < h.h->numFaces; f++) {
        float d = dot(h.p[f].n, refNormal);
        if (d < m) { m = d; face = f; }
    }
    return face;
}

__device__ __forceinline__ float distFromPlane(const Plane &pl, Vec3 a) { return dot(a, pl.n) - pl.d; }

__device__ __forceinline__ Vec3 planeIntersection(const Plane &pl, Vec3 p1, Vec3 p2)
{
    float dist = distFromPlane(pl, p1);
    return p1 + (p2 - p1) * (-dist / dot(pl.n, p2 - p1));
}

constexpr int kClipCap = 12;

__device__ inline int clipPolygon(Vec3 *dst, Plane pl, const Vec3 *in, int n)
{
    if (n <= 0) return 0;
    int out = 0;
    Vec3 v1 = in[n - 1];
    float d1 = distFromPlane(pl, v1);
    for (int i = 0; i < n; ++i) {
        Vec3 v2 = in[i];
        float d2 = distFromPlane(pl, v2);
        if (d1 <= 0.0f && d2 <= 0.0f) {
            if (out < kClipCap) dst[out++] = v2;
        } else if (d1 <= 0.0f && d2 > 0.0f) {
            if (out < kClipCap) dst[out++] = planeIntersection(pl, v1, v2);
        } else if (d2 <= 0.0f && d1 > 0.0f) {
            if (out < kClipCap) dst[out++] = planeIntersection(pl, v1, v2);
            if (out < kClipCap) dst[out++] = v2;
        }
        v1 = v2;
        d1 = d2;
    }
    return out;
}

struct Manifold {
    Vec3 pts[4];
    float depth[4];
    int n;
    Vec3 normal;
};

// buildFaceContactManifold (narrowphase.cpp:771-879).  Note the reference stores
// the winding sign in a `bool`, so its clockwise flip never triggers; kept as is.
__device__ inline Manifold buildManifold(Vec3 normal, const Vec3 *c, const float *dep, int n)
{
    Manifold m;
    if (n <= 4) {
        m.n = n;
        for (int i = 0; i < n; i++) { m.pts[i] = c[i]; m.depth[i] = dep[i]; }
    } else {
        m.n = 4;
        m.pts[0] = c[0]; m.depth[0] = dep[0];
        m.pts[1] = m.pts[2] = m.pts[3] = {0.f, 0.f, 0.f};
        m.depth[1] = m.depth[2] = m.depth[3] = 0.f;

        float maxDistSq = 0.f;
        for (int i = 1; i < n; i++) {
            float d = length2(m.pts[0] - c[i]);
            if (d > maxDistSq) { maxDistSq = d; m.pts[1] = c[i]; m.depth[1] = dep[i]; }
        }
        Vec3 ba = m.pts[1] - m.pts[0];

        float maxArea = 0.0f;
        for (int i = 1; i < n; i++) {
            Vec3 bc = c[i] - m.pts[1];
            float area = fabsf(dot(normal, cross(ba, bc)));
            if (area > maxArea) { maxArea = area; m.pts[2] = c[i]; m.depth[2] = dep[i]; }
        }

        Vec3 cb = m.pts[2] - m.pts[1];
        Vec3 ac = m.pts[0] - m.pts[2];
        float mostNeg = 0.f;
        for (int i = 1; i < n; i++) {
            Vec3 aq = m.pts[0] - c[i];
            Vec3 qc = c[i] - m.pts[2];
            float abq = dot(normal, cross(ba, aq));
            float bcq = dot(normal, cross(cb, qc));
            float caq = dot(normal, cross(aq, ac));
            float mn = fminf(abq, fminf(bcq, caq));
            if (mn < mostNeg) { mostNeg = mn; m.pts[3] = c[i]; m.depth[3] = dep[i]; }
        }
        if (maxDistSq == 0.f || maxArea == 0.f || mostNeg == 0.f) {
            m.n = 0;
            m.normal = {0.f, 0.f, 0.f};
            return m;
        }
    }
    const Quat id = {1.f, 0.f, 0.f, 0.f};
    const Vec3 zero = {0.f, 0.f, 0.f};
    for (int i = 0; i < m.n; i++) m.pts[i] = rotate(id, m.pts[i]) + zero;
    m.normal = rotate(id, normal);
    return m;
}

__device__ inline Manifold faceContact(Plane ref, int refFace, int incFace,
                                       const WorldHull &R, const WorldHull &O)
{
    Vec3 bufA[kClipCap], bufB[kClipCap];
    Vec3 *in = bufA, *dst = bufB;
    int n = 0;
    {
        const HullData &H = *O.h;
        int start = H.faceBase[incFace], he = start;
        do {
            if (n < kClipCap) in[n++] = O.v[H.hedgeRoot[he]];
            he = H.hedgeNext[he];
        } while (he != start);
    }
    {
        const HullData &H = *R.h;
        int start = H.faceBase[refFace], he = start;
        Vec3 cur = R.v[H.hedgeRoot[he]];
        do {
            he = H.hedgeNext[he];
            Vec3 next = R.v[H.hedgeRoot[he]];
            Vec3 edge = next - cur;
            Vec3 pn = cross(edge, ref.n);
            float d = dot(pn, cur);
            cur = next;
            n = clipPolygon(dst, Plane{pn, d}, in, n);
            Vec3 *t = dst; dst = in; in = t;
        } while (he != start);
    }
    float depth[kClipCap];
    int below = 0;
    for (int i = 0; i < n; ++i) {
        Vec3 v = in[i];
        float d = distFromPlane(ref, v);
        if (d <= 0.0f) {
            in[below] = v - d * ref.n;
            depth[below] = -d;
            below++;
        }
    }
    return buildManifold(ref.n, in, depth, below);
}

__device__ inline Manifold facePlaneContact(Plane pl, int incFace, const WorldHull &Hh)
{
    Vec3 pts[kClipCap];
    float depth[kClipCap];
    int n = 0;
    const HullData &H = *Hh.h;
    int start = H.faceBase[incFace], he = start;
    do {
        Vec3 v = Hh.v[H.hedgeRoot[he]];
        he = H.hedgeNext[he];
        float d = distFromPlane(pl, v);
        if (d <= 0.0f && n < kClipCap) {
            pts[n] = v - d * pl.n;
            depth[n] = -d;
            n++;
        }
    } while (he != start);
    return buildManifold(pl.n, pts, depth, n);
}

__device__ inline Manifold edgeContact(const EdgeQuery &q, const WorldHull &a, const WorldHull &b)
{
    const HullData &A = *a.h, &B = *b.h;
    Vec3 a1 = a.v[A.hedgeRoot[q.edgeA]], a2 = a.v[A.hedgeRoot[A.hedgeNext[q.edgeA]]];
    Vec3 b1 = b.v[B.hedgeRoot[q.edgeB]], b2 = b.v[B.hedgeRoot[B.hedgeNext[q.edgeB]]];
    // shortestSegmentBetween (narrowphase.cpp:1037-1068); only the point on A is used
    Vec3 v1 = a2 - a1, v2 = b2 - b1, v21 = b1 - a1;
    float d22 = dot(v2, v2), d11 = dot(v1, v1), d21 = dot(v2, v1);
    float d211 = dot(v21, v1), d212 = dot(v21, v2);
    float denom = d21 * d21 - d22 * d11;
    float s;
    if (fabsf(denom) < 0.00001f) s = 0.0f;
    else s = (d212 * d21 - d22 * d211) / denom;
    s = fmaxf(fminf(s, 1.0f), 0.0f);
    Vec3 contact = a1 + s * v1;

    const Quat id = {1.f, 0.f, 0.f, 0.f};
    const Vec3 zero = {0.f, 0.f, 0.f};
    Manifold m;
    m.pts[0] = rotate(id, contact) + zero;
    m.depth[0] = -q.sep;
    for (int i = 1; i < 4; i++) { m.pts[i] = zero; m.depth[i] = 0.f; }
    m.n = 1;
    m.normal = rotate(id, q.normal);
    return m;
}

// runNarrowphase for one candidate; returns false when no contact is produced.
__device__ inline bool narrowphasePair(const DeviceState &S, int w, int a, int b,
                                       int *refOut, int *altOut, Manifold *out)
{
    uint32_t ma = META(S, a, w), mb = META(S, b, w);
    int objA = metaObj(ma), objB = metaObj(mb);
    int typeA = c_assets.obj[objA].primType, typeB = c_assets.obj[objB].primType;
    if (typeA > typeB) {
        int t = a; a = b; b = t;
        t = objA; objA = objB; objB = t;
        t = typeA; typeA = typeB; typeB = t;
    }
    Vec3 aPos = loadPos(S, a, w), bPos = loadPos(S, b, w);
    Quat aRot = loadRot(S, a, w), bRot = loadRot(S, b, w);
    Vec3 aScale = loadScale(S, a, w), bScale = loadScale(S, b, w);

    AABB aBox = aabbApplyTRS(objAABB(objA), aPos, aRot, aScale);
    AABB bBox = aabbApplyTRS(objAABB(objB), bPos, bRot, bScale);
    if (!aabbIntersects(aBox, bBox)) return false;

    if (typeA == 2 && typeB == 4) {
        WorldHull ha;
        makeWorldHull(ha, c_assets.hull[objA], aPos, aRot, aScale);
        Vec3 n = rotate(bRot, v3(0, 0, 1));
        Plane pl = {n, dot(n, bPos)};
        if (hullDistanceFromPlane(pl, ha) > 0.0f) return false;
        int inc = findIncidentFace(ha, pl.n);
        Manifold m = facePlaneContact(pl, inc, ha);
        if (m.n <= 0) return false;
        *refOut = b; *altOut = a; *out = m;
        return true;
    }
    if (typeA != 2 || typeB != 2) return false;   // sphere / plane-plane pairs never occur here

    WorldHull ha, hb;
    makeWorldHull(ha, c_assets.hull[objA], aPos, aRot, aScale);
    makeWorldHull(hb, c_assets.hull[objB], bPos, bRot, bScale);

    FaceQuery fa = queryFaces(ha, hb);
    if (fa.sep > 0.0f) return false;
    FaceQuery fb = queryFaces(hb, ha);
    if (fb.sep > 0.0f) return false;
    EdgeQuery eq = queryEdges(ha, hb);
    if (eq.sep > 0.0f) return false;

    bool faceA = fa.sep > eq.sep, faceB = fb.sep > eq.sep;
    if (faceA || faceB) {
        bool aIsRef = fa.sep >= fb.sep;
        Plane ref = aIsRef ? fa.plane : fb.plane;
        int refFace = aIsRef ? fa.face : fb.face;
        const WorldHull &R = aIsRef ? ha : hb;
        const WorldHull &O = aIsRef ? hb : ha;
        int inc = findIncidentFace(O, ref.n);
        Manifold m = faceContact(ref, refFace, inc, R, O);
        if (m.n <= 0) return false;
        *refOut = aIsRef ? a : b; *altOut = aIsRef ? b : a; *out = m;
        return true;
    }
    *refOut = a; *altOut = b;
    *out = edgeContact(eq, ha, hb);
    return true;
}

// one warp per world, lanes stride over the world's candidates
__global__ void k_narrowphase(DeviceState S)
{
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= S.p.numWorlds) return;
    const int nc = WI(S, WF_NUM_CAND, w);
    const uint16_t *cand = S.cand + size_t(w) * kMaxCand;
    for (int c = lane; c < nc; c += 32) {
        uint16_t pr = cand[c];
        int ref, alt;
        Manifold m;
        uint32_t meta = 0;
        if (narrowphasePair(S, w, pr & 0xFF, pr >> 8, &ref, &alt, &m)) {
            float *o = S.contact + (size_t(w) * kMaxCand + c) * kContactWords;
            for (int i = 0; i < 4; i++) {
                bool v = i < m.n;
                o[i * 4 + 0] = v ? m.pts[i].x : 0.f;
                o[i * 4 + 1] = v ? m.pts[i].y : 0.f;
                o[i * 4 + 2] = v ? m.pts[i].z : 0.f;
                o[i * 4 + 3] = v ? m.depth[i] : 0.f;
            }
            o[16] = m.normal.x; o[17] = m.normal.y; o[18] = m.normal.z;
            o[19] = 0.f;
            meta = uint32_t(ref) | (uint32_t(alt) << 8) | (uint32_t(m.n) << 16);
        }
        S.cmeta[size_t(w) * kMaxCand + c] = meta;
    }
}

}  // namespace hs
