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

// findIntersectingEntry: one warp per world, lanes own bodies {lane, lane+32}.  The world's BVH
// nodes, leaf->slot map and body meta words are staged in shared memory (one coalesced copy)
// because every lane walks the same small tree.  Pass 1 counts each body's emitted pairs, a
// warp scan turns counts into offsets in body-row order, pass 2 re-walks and writes — so the
// candidate list is in exactly the order the sequential reference produces it.
constexpr int kBroadWarps = 4;

struct alignas(16) BroadWarpSmem {
    uint32_t nodes[kBvhNodeOff + kMaxNodes * kNodeWords];    // header + nodes, blob layout
    uint32_t meta[kMaxBodies];
    uint8_t leafSlot[kMaxBodies + 4];
};

template <typename Fn>
__device__ __forceinline__ void forEachOverlapS(const BroadWarpSmem &ws, const AABB &q, Fn &&fn)
{
    uint8_t stack[24];
    stack[0] = 0;
    int sp = 1;
    while (sp > 0) {
        int node = stack[--sp];
        const float4 *nf = reinterpret_cast<const float4 *>(ws.nodes + kBvhNodeOff + node * kNodeWords);
        const float4 lox = nf[0], loy = nf[1], loz = nf[2], hix = nf[3], hiy = nf[4], hiz = nf[5];
        const int4 ch = *reinterpret_cast<const int4 *>(ws.nodes + kBvhNodeOff + node * kNodeWords + kNodeChild);
#define HS_OVL(C, LX, LY, LZ, HX, HY, HZ)                                                     \
        if (C != kSentinel && q.lo.x < HX && LX < q.hi.x && q.lo.y < HY && LY < q.hi.y &&     \
            q.lo.z < HZ && LZ < q.hi.z) {                                                      \
            if (C < 0) fn(int(ws.leafSlot[C & 0x7FFFFFFF]));                                   \
            else if (sp < 24) stack[sp++] = (uint8_t)C;                                        \
        }
        HS_OVL(ch.x, lox.x, loy.x, loz.x, hix.x, hiy.x, hiz.x)
        HS_OVL(ch.y, lox.y, loy.y, loz.y, hix.y, hiy.y, hiz.y)
        HS_OVL(ch.z, lox.z, loy.z, loz.z, hix.z, hiy.z, hiz.z)
        HS_OVL(ch.w, lox.w, loy.w, loz.w, hix.w, hiy.w, hiz.w)
#undef HS_OVL
    }
}

__global__ void __launch_bounds__(kBroadWarps * 32) k_broadphase(DeviceState S)
{
    __shared__ __align__(16) BroadWarpSmem sm[kBroadWarps];
    const int lane = threadIdx.x & 31;
    const int w = blockIdx.x * kBroadWarps + (threadIdx.x >> 5);
    if (w >= S.p.numWorlds) return;
    BroadWarpSmem &ws = sm[threadIdx.x >> 5];

    const uint32_t *blob = BVH(S, w);
    const int nb = WI(S, WF_NUM_BODIES, w);
    {
        const int numNodes = min((int)blob[0], kMaxNodes), numLeaves = min((int)blob[1], kMaxBodies);
        const uint4 *src = reinterpret_cast<const uint4 *>(blob);
        uint4 *dst = reinterpret_cast<uint4 *>(ws.nodes);
        const int nvec = (kBvhNodeOff + numNodes * kNodeWords) / 4;
        for (int i = lane; i < nvec; i += 32) dst[i] = src[i];
        for (int l = lane; l < numLeaves; l += 32)
            ws.leafSlot[l] = (uint8_t)(blob[kBvhLeafOff + l * kLeafWords + kLeafInfo] & 0xFF);
        for (int s = lane; s < nb; s += 32) ws.meta[s] = META(S, s, w);
    }
    __syncwarp();

    int counts[2] = {0, 0};
    uint32_t metas[2];
    AABB boxes[2];
    for (int r = 0; r < 2; r++) {
        int s = lane + 32 * r;
        if (s >= nb) continue;
        uint32_t m = ws.meta[s];
        metas[r] = m;
        const float *L = leafF(blob, metaLeaf(m));
        boxes[r] = {{L[0], L[1], L[2]}, {L[3], L[4], L[5]}};
        int eid = metaEid(m);
        bool aStatic = metaResp(m) == RESP_STATIC;
        int c = 0;
        forEachOverlapS(ws, boxes[r], [&](int other) {
            uint32_t mo = ws.meta[other];
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
        forEachOverlapS(ws, boxes[r], [&](int other) {
            uint32_t mo = ws.meta[other];
            if (eid < metaEid(mo) && !(aStatic && metaResp(mo) == RESP_STATIC)) {
                if (o < kMaxCand) out[o] = uint16_t(s | (other << 8));
                o++;
            }
        });
    }
}

}  // namespace hs
