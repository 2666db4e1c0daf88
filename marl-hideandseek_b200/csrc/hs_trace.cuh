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
//
// The reference walks the BVH once per body and emits a pair for every leaf whose stored node
// slot overlaps the body's leaf box.  Node slots nest (a slot encloses everything below it, both
// after a rebuild and under grow-only refit), so "the walk reaches leaf l" is equivalent to "the
// query overlaps l's own slot in its parent node", and a pruned walk reports leaves in the same
// relative order as the full walk.  So instead of 29 divergent tree walks per world the warp
// stages, in visiting order (dfsLeaf, written by the rebuild), each leaf's slot bounds, body slot
// and meta word in shared memory, and every lane runs ONE uniform loop over that list —
// broadcast reads, no divergence — collecting its hits in a 64-bit mask.  A warp scan of the
// popcounts gives each body its offset in body-row order and the set bits are written out in
// visiting order: the candidate list is exactly the sequential reference's.
constexpr int kBroadWarps = 4;

constexpr int kStageMax = 18;        // bodies the solver stages per world: 9 boxes + 2 ramps + plane + 6 agents

struct alignas(16) BroadWarpSmem {
    float bounds[6][kDfsStride];     // slot bounds of the r-th visited leaf
    uint32_t meta[kDfsStride];       // meta word of the body in that leaf
    uint8_t slot[kDfsStride];
    // per-pair solver info (see the tail of k_broadphase)
    uint16_t cand[kMaxCand];         // the candidate pairs as written out
    uint8_t binfo[kDfsStride];       // per body slot: obj | carrying << 4 | static << 5
};

__global__ void __launch_bounds__(kBroadWarps * 32) k_broadphase(DeviceState S)
{
    __shared__ BroadWarpSmem sm[kBroadWarps];
    const int lane = threadIdx.x & 31;
    const int w = blockIdx.x * kBroadWarps + (threadIdx.x >> 5);
    if (w >= S.p.numWorlds) return;
    BroadWarpSmem &ws = sm[threadIdx.x >> 5];

    const uint32_t *blob = BVH(S, w);
    const int nb = WI(S, WF_NUM_BODIES, w);
    const int numLeaves = min((int)blob[1], kMaxBodies);
    const uint8_t *dfs = S.dfsLeaf + size_t(w) * kDfsStride;
    for (int r = lane; r < numLeaves; r += 32) {
        const int leaf = dfs[r];
        const uint32_t lp = blob[kBvhLeafOff + leaf * kLeafWords + kLeafParent];
        const float *nf = nodeF(blob, int(lp >> 2));
        const int sub = int(lp & 3);
#pragma unroll
        for (int k = 0; k < 6; k++) ws.bounds[k][r] = nf[k * 4 + sub];
        const int slot = int(blob[kBvhLeafOff + leaf * kLeafWords + kLeafInfo] & 0xFF);
        ws.slot[r] = (uint8_t)slot;
        ws.meta[r] = META(S, slot, w);
    }
    __syncwarp();

    unsigned long long hits[2] = {0ull, 0ull};
    int counts[2] = {0, 0};
    for (int rd = 0; rd < 2; rd++) {
        const int s = lane + 32 * rd;
        if (s >= nb) continue;
        const uint32_t m = META(S, s, w);
        const float *L = leafF(blob, metaLeaf(m));
        const float qlx = L[0], qly = L[1], qlz = L[2], qhx = L[3], qhy = L[4], qhz = L[5];
        const int eid = metaEid(m);
        const bool aStatic = metaResp(m) == RESP_STATIC;
        unsigned long long h = 0ull;
        for (int r = 0; r < numLeaves; r++) {
            // AABB::overlaps(query, slot) — strict
            const bool ov = qlx < ws.bounds[3][r] && ws.bounds[0][r] < qhx &&
                            qly < ws.bounds[4][r] && ws.bounds[1][r] < qhy &&
                            qlz < ws.bounds[5][r] && ws.bounds[2][r] < qhz;
            const uint32_t mo = ws.meta[r];
            if (ov && eid < metaEid(mo) && !(aStatic && metaResp(mo) == RESP_STATIC)) h |= 1ull << r;
        }
        hits[rd] = h;
        counts[rd] = __popcll(h);
    }

    int offs[2];
    int base = 0;
    for (int rd = 0; rd < 2; rd++) {
        int v = counts[rd];
        int incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= d) incl += t;
        }
        offs[rd] = base + incl - v;
        base += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }

    int total = base;
    if (total > kMaxCand) {
        if (lane == 0) atomicAdd(S.overflow, 1u);
        total = kMaxCand;
    }
    if (lane == 0) WI(S, WF_NUM_CAND, w) = total;

    uint16_t *out = S.cand + size_t(w) * kMaxCand;
    for (int rd = 0; rd < 2; rd++) {
        const int s = lane + 32 * rd;
        unsigned long long h = hits[rd];
        int o = offs[rd];
        while (h) {
            const int r = __ffsll((long long)h) - 1;
            h &= h - 1ull;
            if (o < kMaxCand) {
                out[o] = uint16_t(s | (int(ws.slot[r]) << 8));
                ws.cand[o] = uint16_t(s | (int(ws.slot[r]) << 8));
            }
            o++;
        }
    }

    // ---- what the solver needs to know about this step's pairs (hs_solver.cuh) ----------------------
    // Which bodies are Static and which rotations are identity is fixed for the four substeps of a
    // step, so it is decided here once: a body is "carrying" (the Gauss-Seidel sweep may change it,
    // so contacts sharing it are ordered) unless it is Static with identity rotation — walls, the
    // ground plane — which the reference never moves and whose quaternion it re-normalises to the
    // same bits.  Every candidate pair gets an entry with both bodies' slot, object id and flags.
    // The solver keeps the bodies it may write in shared memory: they are the slots from the first
    // non-wall, non-plane or carrying body to the end of the world (walls come first in a generated
    // level), at most kStageMax of them; a world with a longer range is solved out of global memory.
    int firstStaged = nb;
    for (int rd = 0; rd < 2; rd++) {
        const int s = lane + 32 * rd;
        if (s >= nb) continue;
        const uint32_t m = META(S, s, w);
        const bool isStatic = metaResp(m) == RESP_STATIC;
        bool carrying = !isStatic;
        if (isStatic) {
            const Quat q = loadRot(S, s, w);
            carrying = !(q.w == 1.f && q.x == 0.f && q.y == 0.f && q.z == 0.f);
        }
        const int obj = metaObj(m);
        ws.binfo[s] = uint8_t(obj | (carrying ? 16 : 0) | (isStatic ? 32 : 0));
        if (carrying || (obj != OBJ_WALL && obj != OBJ_PLANE)) firstStaged = min(firstStaged, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) firstStaged = min(firstStaged, __shfl_xor_sync(0xFFFFFFFFu, firstStaged, o));
    if (lane == 0) {
        WI(S, WF_STAGE_FIRST, w) = firstStaged;
        WI(S, WF_STAGE_COUNT, w) = nb - firstStaged > kStageMax ? -1 : nb - firstStaged;
    }
    __syncwarp();
    SchedEntry *sched = S.sched + size_t(w) * kMaxCand;
    for (int c = lane; c < total; c += 32) {
        const int a = ws.cand[c] & 0xFF, b = ws.cand[c] >> 8;
        const int ia = ws.binfo[a], ib = ws.binfo[b];
        SchedEntry e;
        e.x = uint32_t(c) | (uint32_t(a) << 16) | (uint32_t(b) << 24);
        e.y = uint32_t(ia & 15) | (uint32_t(ib & 15) << 4) | (uint32_t((ia >> 4) & 1) << 8) | (uint32_t((ib >> 4) & 1) << 9) |
              (uint32_t((ia >> 5) & 1) << 10) | (uint32_t((ib >> 5) & 1) << 11);
        sched[c] = e;
    }
}

}  // namespace hs
