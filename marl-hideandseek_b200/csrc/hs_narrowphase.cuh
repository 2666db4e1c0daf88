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

// A hull in world space.  The vertex and plane arrays are indexed by data-dependent ids (half-edge
// roots, face ids), so a per-thread copy lives in local memory: 1 KB of stack per thread for a
// pair, whose footprint over a CTA wave exceeds the L1 (11 % of the local loads missed to L2, and
// every extra lane made it worse).  The arrays therefore live in a shared-memory slot of 51 words
// per thread (odd stride: lanes reading the same element of their own slots hit 32 banks), and
// WorldHull is a view of one slot.
struct HullSlot {
    Vec3 v[kMaxHullVerts];
    Plane p[kMaxHullFaces];
    Vec3 center;
};
static_assert(sizeof(HullSlot) == 51 * 4, "HullSlot stride must stay odd in words");

struct WorldHull {
    const Vec3 *v;
    const Plane *p;
    Vec3 center;
    const HullData *h;
};

__device__ __forceinline__ WorldHull viewHull(const HullSlot &slot, const HullData &H)
{
    return {slot.v, slot.p, slot.center, &H};
}

__device__ inline void makeWorldHull(HullSlot &W, const HullData &H, Vec3 t, Quat rot, Vec3 scale)
{
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

// Edges [ea0, ea1) of `a` against every edge of `b`.  The reference's running maximum takes a
// later pair only when it is strictly larger, so the query over all edges equals the left-biased
// combination of the queries over consecutive ranges (combineEdgeQueries); a positive separation
// ends the reference's loop, and then only its sign is used (doSAT, narrowphase.cpp:659-739).
__device__ inline EdgeQuery queryEdgeRange(const WorldHull &a, const WorldHull &b, int ea0, int ea1)
{
    EdgeQuery q;
    q.sep = -FLT_MAX; q.normal = {0.f, 0.f, 0.f}; q.edgeA = 0; q.edgeB = 0;
    const HullData &A = *a.h, &B = *b.h;
    const int nb = B.numHedges / 2;
    static_assert(kMaxHullHedges / 2 <= 32, "one mark bit per B edge");
    for (int ea = ea0; ea < ea1; ea++) {
        const int ha = ea * 2;
        const Vec3 aN1 = a.p[A.hedgeFace[ha]].n, aN2 = a.p[A.hedgeFace[ha + 1]].n;
        const Vec3 a1 = a.v[A.hedgeRoot[ha]], a2 = a.v[A.hedgeRoot[A.hedgeNext[ha]]];
        const Vec3 dirA = a2 - a1;
        // Two phases per row.  Only ~1 edge pair in 12 forms a face of the Minkowski difference, but
        // with 32 different pairs of hulls in a warp some lane's pair does at almost every step, so a
        // single loop drags the whole warp through the separation arithmetic (cross, sqrt, division)
        // at one or two active lanes each time.  Phase 1 only marks the row's candidates
        // (branch-free), phase 2 walks each lane's marks in order: the lanes are in the arithmetic
        // together.  Pairs that are no Minkowski face (or are parallel) never raise the maximum, so
        // skipping them and visiting the others in the same order gives the same result.
        // (Marking all rows first and walking all marks at the end measured slightly slower.)
        uint32_t marks = 0u;
        for (int eb = 0; eb < nb; eb++) {
            const int hb = eb * 2;
            const Vec3 bN1 = b.p[B.hedgeFace[hb]].n, bN2 = b.p[B.hedgeFace[hb + 1]].n;
            if (isMinkowskiFace(aN1, aN2, -bN1, -bN2)) marks |= 1u << eb;
        }
        for (; marks; marks &= marks - 1u) {
            const int hb = (__ffs(int(marks)) - 1) * 2;
            const Vec3 b1 = b.v[B.hedgeRoot[hb]], b2 = b.v[B.hedgeRoot[B.hedgeNext[hb]]];
            const Vec3 dirB = b2 - b1;
            const Vec3 c = cross(dirA, dirB);
            const float len2 = length2(c);
            if (len2 == 0.f) continue;
            const float inv = 1.f / sqrtf(len2);
            Vec3 normal = c * inv;
            if (dot(normal, a1 - a.center) < 0.0f) normal = -normal;
            const float sep = dot(normal, b1 - a1);
            if (sep > q.sep) {
                q.sep = sep; q.normal = normal; q.edgeA = ha; q.edgeB = hb;
            }
        }
        // the reference leaves at the first positive separation; only its sign is used then, and the
        // running maximum stays positive, so the exit can wait for the end of the row
        if (q.sep > 0.f) return q;
    }
    return q;
}

__device__ inline EdgeQuery queryEdges(const WorldHull &a, const WorldHull &b)
{
    return queryEdgeRange(a, b, 0, a.h->numHedges / 2);
}

__device__ __forceinline__ EdgeQuery combineEdgeQueries(const EdgeQuery &first, const EdgeQuery &second)
{
    return second.sep > first.sep ? second : first;
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

// createFacePlaneContact for a face of any size (the reference's form: arrays indexed as they fill)
__device__ __noinline__ Manifold facePlaneContactGeneric(Plane pl, int incFace, const WorldHull &Hh)
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

// The same for faces of at most four vertices — every face of this environment's hulls — without
// arrays that fill at a running index: those live in local memory, and with 2 x 97 KB of shared
// memory per SM there is little L1 left to keep them in.  Candidates stay in registers and are
// packed into the manifold's slots by selects, in the same order, through the same operations.
__device__ inline Manifold facePlaneContact(Plane pl, int incFace, const WorldHull &Hh)
{
    const HullData &H = *Hh.h;
    const int start = H.faceBase[incFace];
    int he = start;
    bool more = true;
    Vec3 pv[4];
    float pd[4];
    bool ok[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        ok[k] = false;
        pv[k] = {0.f, 0.f, 0.f};
        pd[k] = 0.f;
        if (more) {
            const Vec3 v = Hh.v[H.hedgeRoot[he]];
            he = H.hedgeNext[he];
            const float d = distFromPlane(pl, v);
            if (d <= 0.0f) {
                pv[k] = v - d * pl.n;
                pd[k] = -d;
                ok[k] = true;
            }
            more = he != start;
        }
    }
    if (more) return facePlaneContactGeneric(pl, incFace, Hh);       // a face with more than four vertices
    const int c1 = ok[0], c2 = c1 + ok[1], c3 = c2 + ok[2];
    Manifold m;
    m.n = c3 + ok[3];
    const Quat id = {1.f, 0.f, 0.f, 0.f};
    const Vec3 zero = {0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const bool s0 = ok[0] && j == 0, s1 = ok[1] && c1 == j, s2 = ok[2] && c2 == j, s3 = ok[3] && c3 == j;
        const Vec3 p = s0 ? pv[0] : s1 ? pv[1] : s2 ? pv[2] : pv[3];
        m.depth[j] = s0 ? pd[0] : s1 ? pd[1] : s2 ? pd[2] : pd[3];
        m.pts[j] = j < m.n ? rotate(id, p) + zero : p;          // buildManifold's tail, see there
        (void)s3;
    }
    m.normal = rotate(id, pl.n);
    return m;
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

// ---- kernel ------------------------------------------------------------------------------------
// One CTA serves kNpWorlds consecutive worlds.  A world has only ~27 candidate pairs of two very
// different costs (hull vs ground plane: one hull transform; hull vs hull: two transforms + SAT +
// clipping), so a warp per world would run at a handful of active lanes.  Instead:
//   pass 1 (8 lanes per world): load both poses, world-AABB reject (narrowphase.cpp:1560-1577),
//           classify the survivors and append them to two CTA-wide shared lists (plane / hull);
//   pass 2: the whole CTA drains the plane list, then the hull list — lanes are full and every
//           warp runs one code path at a time.
// Each pair still writes the contact slot of its own candidate index, so the solver sees
// contacts in the reference's creation order no matter which lane produced them.
#ifndef HS_NP_WORLDS
#define HS_NP_WORLDS 32
#endif
#ifndef HS_NP_THREADS
#define HS_NP_THREADS 384
#endif
#ifndef HS_NP_MINB
#define HS_NP_MINB 2
#endif
constexpr int kNpWorlds = HS_NP_WORLDS;
constexpr int kNpThreads = HS_NP_THREADS;
constexpr int kNpListCap = kNpWorlds * kMaxCand;
static_assert(kNpWorlds <= 32 && kNpThreads % kNpWorlds == 0 && kNpThreads % 32 == 0, "narrowphase CTA shape");

struct NpSmem {
    AssetTable assets;
    uint32_t list[kNpListCap];      // plane pairs grow from the front, hull pairs from the back
    int numPlane, numHull;
    HullSlot hull[kNpThreads];      // one world-space hull per thread (see HullSlot)
};

__device__ __forceinline__ AABB objAABBOf(const AssetTable &AT, int obj)
{
    const float *a = AT.obj[obj].aabb;
    return {{a[0], a[1], a[2]}, {a[3], a[4], a[5]}};
}

__device__ __forceinline__ void writeContact(const DeviceState &S, int w, int c, int ref, int alt, const Manifold &m)
{
    float4 *o = reinterpret_cast<float4 *>(S.contact + (size_t(w) * kMaxCand + c) * kContactWords);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        bool v = i < m.n;
        o[i] = v ? make_float4(m.pts[i].x, m.pts[i].y, m.pts[i].z, m.depth[i]) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    o[4] = make_float4(m.normal.x, m.normal.y, m.normal.z, 0.f);
    S.cmeta[size_t(w) * kMaxCand + c] = uint32_t(ref) | (uint32_t(alt) << 8) | (uint32_t(m.n) << 16);
}

__global__ void __launch_bounds__(kNpThreads, HS_NP_MINB) k_narrowphase(DeviceState S)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    NpSmem &sm = *reinterpret_cast<NpSmem *>(smemRaw);
    const int tid = threadIdx.x;
    {
        const uint32_t *ca = reinterpret_cast<const uint32_t *>(&c_assets);
        uint32_t *sa = reinterpret_cast<uint32_t *>(&sm.assets);
        for (int i = tid; i < int(sizeof(AssetTable) / 4); i += kNpThreads) sa[i] = ca[i];
        if (tid == 0) { sm.numPlane = 0; sm.numHull = 0; }
    }
    __syncthreads();
    const AssetTable &AT = sm.assets;
    const int w0 = blockIdx.x * kNpWorlds;

    // ---- pass 1: AABB reject + classification, 8 lanes per world ---------------------------------
    {
        constexpr int kLanes = kNpThreads / kNpWorlds;
        const int wi = tid / kLanes, l = tid % kLanes;
        const int w = w0 + wi;
        const int nc = w < S.p.numWorlds ? WI(S, WF_NUM_CAND, w) : 0;
        const uint16_t *cand = S.cand + size_t(w) * kMaxCand;
        for (int c = l; c < nc; c += kLanes) {
            uint16_t pr = cand[c];
            int a = pr & 0xFF, b = pr >> 8;
            int objA = metaObj(META(S, a, w)), objB = metaObj(META(S, b, w));
            int typeA = AT.obj[objA].primType, typeB = AT.obj[objB].primType;
            if (typeA > typeB) {
                int t = a; a = b; b = t;
                t = objA; objA = objB; objB = t;
                t = typeA; typeA = typeB; typeB = t;
            }
            const Pose pa = loadPose(S, REC_POSE, a, w), pb = loadPose(S, REC_POSE, b, w);
            AABB aBox = aabbApplyTRS(objAABBOf(AT, objA), pa.x, pa.q, loadScale(S, a, w));
            AABB bBox = aabbApplyTRS(objAABBOf(AT, objB), pb.x, pb.q, loadScale(S, b, w));
            bool plane = typeA == 2 && typeB == 4, hull = typeA == 2 && typeB == 2;
            if (!aabbIntersects(aBox, bBox) || !(plane || hull)) {
                S.cmeta[size_t(w) * kMaxCand + c] = 0;
                continue;
            }
            uint32_t rec = uint32_t(wi) | (uint32_t(c) << 5) | (uint32_t(a) << 13) | (uint32_t(b) << 21);
            if (plane) sm.list[atomicAdd(&sm.numPlane, 1)] = rec;
            else sm.list[kNpListCap - 1 - atomicAdd(&sm.numHull, 1)] = rec;
        }
    }
    __syncthreads();

    // ---- pass 2a: hull vs ground plane -------------------------------------------------------------
    const int numPlane = sm.numPlane, numHull = sm.numHull;
    for (int i = tid; i < numPlane; i += kNpThreads) {
        const uint32_t rec = sm.list[i];
        const int w = w0 + int(rec & 31), c = int(rec >> 5) & 0xFF, a = int(rec >> 13) & 0xFF, b = int(rec >> 21) & 0xFF;
        const Pose pa = loadPose(S, REC_POSE, a, w), pb = loadPose(S, REC_POSE, b, w);
        const HullData &HA = AT.hull[metaObj(META(S, a, w))];
        makeWorldHull(sm.hull[tid], HA, pa.x, pa.q, loadScale(S, a, w));
        const WorldHull ha = viewHull(sm.hull[tid], HA);
        Vec3 n = rotate(pb.q, v3(0, 0, 1));
        Plane pl = {n, dot(n, pb.x)};
        bool has = false;
        Manifold m;
        if (!(hullDistanceFromPlane(pl, ha) > 0.0f)) {
            int inc = findIncidentFace(ha, pl.n);
            m = facePlaneContact(pl, inc, ha);
            has = m.n > 0;
        }
        if (has) writeContact(S, w, c, b, a, m);
        else S.cmeta[size_t(w) * kMaxCand + c] = 0;
    }

    // ---- pass 2b: hull vs hull, two lanes per pair --------------------------------------------------
    // The SAT of one pair is a long serial chain (two face queries, then up to 12 x 12 edge pairs)
    // and a launch holds only ~5 such pairs per world, so the pass is bound by that chain, not by
    // issue slots.  Lanes 2k / 2k+1 share pair k: lane 0 runs queryFaces(A, B) and the first half
    // of A's edges, lane 1 queryFaces(B, A) and the second half; the halves are exchanged with
    // full-mask shuffles between warp-uniform phases and combined in the reference's order
    // (combineEdgeQueries), so the contact is bit-identical to the one-lane pass.  Lane 0 then
    // builds and writes the manifold.  Each lane also transforms only one of the two hulls.
    const int hullLanes = numHull * 2;
    for (int it = tid; (it & ~31) < hullLanes; it += kNpThreads) {
        const bool valid = it < hullLanes;
        const int half = it & 1;
        const uint32_t rec = sm.list[kNpListCap - 1 - (valid ? it >> 1 : 0)];
        const int w = w0 + int(rec & 31), c = int(rec >> 5) & 0xFF, a = int(rec >> 13) & 0xFF, b = int(rec >> 21) & 0xFF;
        // each lane transforms ONE of the two hulls into its slot; the pair reads both
        const HullData &HA = AT.hull[metaObj(META(S, a, w))], &HB = AT.hull[metaObj(META(S, b, w))];
        __syncwarp();                                       // the slots' previous readers are done
        {
            const int body = half ? b : a;
            const Pose ps = loadPose(S, REC_POSE, body, w);
            makeWorldHull(sm.hull[tid], half ? HB : HA, ps.x, ps.q, loadScale(S, body, w));
        }
        __syncwarp();
        const WorldHull ha = viewHull(sm.hull[tid & ~1], HA), hb = viewHull(sm.hull[tid | 1], HB);

        // one call site: `half ? queryFaces(hb, ha) : queryFaces(ha, hb)` would be two inlined copies that
        // the two lanes of every pair run one after the other
        const WorldHull fx = half ? hb : ha, fy = half ? ha : hb;
        FaceQuery mine = queryFaces(fx, fy);
        __syncwarp();
        FaceQuery other;
        other.sep = __shfl_xor_sync(0xFFFFFFFFu, mine.sep, 1);
        other.face = __shfl_xor_sync(0xFFFFFFFFu, mine.face, 1);
        other.plane.n.x = __shfl_xor_sync(0xFFFFFFFFu, mine.plane.n.x, 1);
        other.plane.n.y = __shfl_xor_sync(0xFFFFFFFFu, mine.plane.n.y, 1);
        other.plane.n.z = __shfl_xor_sync(0xFFFFFFFFu, mine.plane.n.z, 1);
        other.plane.d = __shfl_xor_sync(0xFFFFFFFFu, mine.plane.d, 1);
        const FaceQuery fa = half ? other : mine, fb = half ? mine : other;
        bool separated = fa.sep > 0.0f || fb.sep > 0.0f;

        EdgeQuery em;
        em.sep = -FLT_MAX; em.normal = {0.f, 0.f, 0.f}; em.edgeA = 0; em.edgeB = 0;
        if (!separated) {
            const int na = ha.h->numHedges / 2, mid = na / 2;
            em = queryEdgeRange(ha, hb, half ? mid : 0, half ? na : mid);
        }
        __syncwarp();
        EdgeQuery eo;
        eo.sep = __shfl_xor_sync(0xFFFFFFFFu, em.sep, 1);
        eo.normal.x = __shfl_xor_sync(0xFFFFFFFFu, em.normal.x, 1);
        eo.normal.y = __shfl_xor_sync(0xFFFFFFFFu, em.normal.y, 1);
        eo.normal.z = __shfl_xor_sync(0xFFFFFFFFu, em.normal.z, 1);
        eo.edgeA = __shfl_xor_sync(0xFFFFFFFFu, em.edgeA, 1);
        eo.edgeB = __shfl_xor_sync(0xFFFFFFFFu, em.edgeB, 1);
        if (half || !valid) continue;
        const EdgeQuery eq = combineEdgeQueries(em, eo);

        bool has = false;
        int ref = a, alt = b;
        Manifold m;
        do {
            if (separated || eq.sep > 0.0f) break;
            bool faceA = fa.sep > eq.sep, faceB = fb.sep > eq.sep;
            if (faceA || faceB) {
                bool aIsRef = fa.sep >= fb.sep;
                Plane rp = aIsRef ? fa.plane : fb.plane;
                int refFace = aIsRef ? fa.face : fb.face;
                const WorldHull &R = aIsRef ? ha : hb;
                const WorldHull &O = aIsRef ? hb : ha;
                int inc = findIncidentFace(O, rp.n);
                m = faceContact(rp, refFace, inc, R, O);
                if (m.n <= 0) break;
                ref = aIsRef ? a : b; alt = aIsRef ? b : a;
                has = true;
            } else {
                m = edgeContact(eq, ha, hb);
                has = true;
            }
        } while (false);
        if (has) writeContact(S, w, c, ref, alt, m);
        else S.cmeta[size_t(w) * kMaxCand + c] = 0;
    }
}

}  // namespace hs
