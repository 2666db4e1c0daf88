// Ray tracing and overlap queries against the per-world BVH.
//
// Behavioural contract:
//   BVH::traceRay / traceRayIntoLeaf / traceRayIntoConvexPolyhedron / traceRayIntoPlane
//                                              M/src/physics/broadphase.cpp:658-890
//   BVH::findIntersecting (DFS, LIFO stack, children 0..3)
//                                              M/include/madrona/broadphase.inl:22-59
//   findIntersectingEntry (candidate emission)  broadphase.cpp:930-993
// `blob` may point at global or shared memory (generic addressing).
#pragma once

#include "hs_bvh.cuh"

namespace hs {

// Ray vs one leaf, in the leaf's object space.  Returns true and the hit t.
__device__ inline bool traceLeaf(const uint32_t *blob, int leaf, Vec3 o, Vec3 d,
                                 float t_min, float t_max, float *hit_t)
{
    const float *L = leafF(blob, leaf);
    int obj = int(blob[kBvhLeafOff + leaf * kLeafWords + kLeafInfo] >> 8) & 0xFF;
    Vec3 pos = {L[kLeafPos], L[kLeafPos + 1], L[kLeafPos + 2]};
    Quat rot = {L[kLeafRot], L[kLeafRot + 1], L[kLeafRot + 2], L[kLeafRot + 3]};
    Vec3 scale = {L[kLeafScale], L[kLeafScale + 1], L[kLeafScale + 2]};

    Quat toLocal = qinv(rot);
    Vec3 oo = rotate(toLocal, o - pos);
    oo.x /= scale.x; oo.y /= scale.y; oo.z /= scale.z;
    Vec3 od = rotate(toLocal, d);
    od.x /= scale.x; od.y /= scale.y; od.z /= scale.z;
    Vec3 inv = {1.f / od.x, 1.f / od.y, 1.f / od.z};

    const ObjectData &O = c_assets.obj[obj];
    Vec3 lo = {O.aabb[0], O.aabb[1], O.aabb[2]}, hi = {O.aabb[3], O.aabb[4], O.aabb[5]};
    if (!aabbRayIntersects(lo, hi, oo, inv, 0.f, t_max)) return false;

    if (O.primType == 4) {
        // plane z = 0, normal +z
        float denom = od.z;
        if (denom == 0.f) return false;
        float t = -oo.z / denom;
        if (t < t_min || t > t_max) return false;
        *hit_t = t;
        return true;
    }

    // RTCD 5.3.8 clipped against outward-facing hull planes
    const HullData &H = c_assets.hull[obj];
    float tfirst = t_min, tlast = t_max;
    bool entered = false;
    for (int f = 0; f < H.numFaces; f++) {
        Vec3 n = {H.planes[f][0], H.planes[f][1], H.planes[f][2]};
        float pd = H.planes[f][3];
        float denom = dot(n, od);
        float negDist = pd - dot(n, oo);
        if (denom == 0.0f) {
            if (negDist < 0.0f) return false;
        } else {
            float t = negDist / denom;
            if (denom < 0.0f) {
                if (t >= tfirst) { tfirst = t; entered = true; }
            } else {
                if (t <= tlast) tlast = t;
            }
            if (tfirst > tlast) return false;
        }
    }
    if (!entered) return false;     // only back faces hit -> miss
    *hit_t = tfirst;
    return true;
}

// Returns the body slot of the closest hit (or -1) and writes *out_t.
__device__ inline int traceRay(const uint32_t *blob, Vec3 o, Vec3 d, float t_max, float *out_t)
{
    Vec3 inv = {1.f / d.x, 1.f / d.y, 1.f / d.z};
    int32_t stack[32];
    stack[0] = 0;
    int sp = 1;
    int hitLeaf = -1;

    while (sp > 0) {
        int node = stack[--sp];
        const float *nf = nodeF(blob, node);
        const int32_t *ni = nodeI(blob, node);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            int32_t child = ni[kNodeChild + i];
            if (child == kSentinel) continue;
            Vec3 lo = {nf[0 + i], nf[4 + i], nf[8 + i]};
            Vec3 hi = {nf[12 + i], nf[16 + i], nf[20 + i]};
            if (!aabbRayIntersects(lo, hi, o, inv, 0.f, t_max)) continue;
            if (child < 0) {
                int leaf = child & 0x7FFFFFFF;
                float t;
                if (traceLeaf(blob, leaf, o, d, 0.f, t_max, &t)) {
                    t_max = t;
                    hitLeaf = leaf;
                }
            } else if (sp < 32) {
                stack[sp++] = child;
            }
        }
    }
    if (hitLeaf < 0) return -1;
    *out_t = t_max;
    return int(blob[kBvhLeafOff + hitLeaf * kLeafWords + kLeafInfo] & 0xFF);
}

// DFS overlap query; fn(slot) is called for every overlapping leaf in the
// reference's visiting order.
template <typename Fn>
__device__ __forceinline__ void forEachOverlap(const uint32_t *blob, const AABB &q, Fn &&fn)
{
    int32_t stack[32];
    stack[0] = 0;
    int sp = 1;
    while (sp > 0) {
        int node = stack[--sp];
        const float *nf = nodeF(blob, node);
        const int32_t *ni = nodeI(blob, node);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            int32_t child = ni[kNodeChild + i];
            if (child == kSentinel) continue;
            AABB c = {{nf[0 + i], nf[4 + i], nf[8 + i]}, {nf[12 + i], nf[16 + i], nf[20 + i]}};
            if (!aabbOverlaps(q, c)) continue;
            if (child < 0) {
                fn(int(blob[kBvhLeafOff + (child & 0x7FFFFFFF) * kLeafWords + kLeafInfo] & 0xFF));
            } else if (sp < 32) {
                stack[sp++] = child;
            }
        }
    }
}

// findIntersectingEntry: one warp per world, lanes own bodies {lane, lane+32}.
// Pass 1 counts each body's emitted pairs, a warp scan turns counts into
// offsets in body-row order, pass 2 re-walks and writes — so the candidate list
// is in exactly the order the sequential reference produces it.
__global__ void k_broadphase(DeviceState S)
{
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= S.p.numWorlds) return;

    const uint32_t *blob = BVH(S, w);
    const int nb = WI(S, WF_NUM_BODIES, w);
    const int N = S.p.numWorlds;

    int counts[2] = {0, 0};
    uint32_t metas[2];
    AABB boxes[2];
    for (int r = 0; r < 2; r++) {
        int s = lane + 32 * r;
        if (s >= nb) continue;
        uint32_t m = S.meta[size_t(s) * N + w];
        metas[r] = m;
        const float *L = leafF(blob, metaLeaf(m));
        boxes[r] = {{L[0], L[1], L[2]}, {L[3], L[4], L[5]}};
        int eid = metaEid(m);
        bool aStatic = metaResp(m) == RESP_STATIC;
        int c = 0;
        forEachOverlap(blob, boxes[r], [&](int other) {
            uint32_t mo = S.meta[size_t(other) * N + w];
            if (eid < metaEid(mo) && !(aStatic && metaResp(mo) == RESP_STATIC)) c++;
        });
        counts[r] = c;
    }

    int offs[2];
    int base = 0;
    for (int r = 0; r < 2; r++) {
        int v = counts[r];
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += t;
        }
        offs[r] = base + incl - v;
        base += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }

    int total = base;
    if (total > kMaxCand) {
        if (lane == 0) atomicAdd(S.overflow, 1u);
        total = kMaxCand;
    }
    if (lane == 0) WI(S, WF_NUM_CAND, w) = total;

    uint16_t *out = S.cand + size_t(w) * kMaxCand;
    for (int r = 0; r < 2; r++) {
        int s = lane + 32 * r;
        if (s >= nb || counts[r] == 0) continue;
        uint32_t m = metas[r];
        int eid = metaEid(m);
        bool aStatic = metaResp(m) == RESP_STATIC;
        int o = offs[r];
        forEachOverlap(blob, boxes[r], [&](int other) {
            uint32_t mo = S.meta[size_t(other) * N + w];
            if (eid < metaEid(mo) && !(aStatic && metaResp(mo) == RESP_STATIC)) {
                if (o < kMaxCand) out[o] = uint16_t(s | (other << 8));
                o++;
            }
        });
    }
}

}  // namespace hs
