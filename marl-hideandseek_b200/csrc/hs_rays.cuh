// Fused ray stage: 30-ray lidar + visibility masks for every agent of a world in ONE kernel.
//
// Behavioural contract (unchanged from hs_obs.cuh / hs_trace.cuh):
//   computeVisibilitySystem (CPU branch, incl. hiderTeamReward side effect)  src/sim.cpp:567-605, 663-709
//   lidarSystem                                                             src/sim.cpp:712-759
//   BVH::traceRay / traceRayIntoLeaf / ...                                  M/src/physics/broadphase.cpp:658-890
//
// Mapping (B200): one CTA per world.  The world's BVH tile (header + nodes + leaves, the
// contiguous world-major blob of hs_state.cuh) and the hull/object table are staged in shared
// memory once; every ray of the world — A*30 lidar rays plus the visibility rays that pass the
// view-cone test, compacted into a shared queue — is then one lane's work item, so warps are
// full regardless of how many targets an agent has.
//
// Traversal is warp-convergent: every lane walks its own ray through the 4-wide tree (all four
// child slabs of a node are tested in straight-line code from float4 shared loads), but a lane
// that reaches a leaf parks until a warp vote finds enough parked lanes, and the (expensive,
// formerly 4x-inlined) ray-vs-hull test then runs once for all of them.  Per-lane visiting order
// — pop LIFO, children 0..3, leaves tested at the moment they are met with the current t_max —
// is exactly the reference's, so hit entities are identical even for exact ties.
#pragma once

#include "hs_trace.cuh"

namespace hs {

#ifndef HS_VOTE_NUM
#define HS_VOTE_NUM 1
#define HS_VOTE_DEN 1
#endif
#ifndef HS_RAY_THREADS
#define HS_RAY_THREADS 160
#endif
constexpr int kRayThreads = HS_RAY_THREADS;
#ifndef HS_RAY_MINB
#define HS_RAY_MINB 8
#endif
constexpr int kMaxVisRays = kMaxAgents * (kMaxBoxes + kMaxRamps + kMaxAgents - 1);   // 96
constexpr int kMaxRays = kMaxAgents * 30 + kMaxVisRays;                               // 276 per world
constexpr int kRayQueueA = 1024, kRayQueueB = 384;
static_assert(kMaxRays <= 512 && kMaxBodies <= 64, "queue entries pack a ray in 9 bits and a leaf in 6");

__constant__ float c_lidarDir[30][2];     // (cos, sin) of theta_i, hs_det_sincosf on the host

// kFullBvh = false (the flat nearest-first query, which never touches the inner nodes) keeps only
// the leaf records: 4.2 KB instead of the 13 KB blob, which is what lets an SM hold 8 CTAs.
// blob() is the address the blob WOULD start at, so kBvhLeafOff-based accessors work unchanged.
template <bool kFullBvh>
struct RaySmemT {
    uint32_t bvh[kFullBvh ? kBvhWords : kMaxBodies * kLeafWords];
    __device__ __forceinline__ uint32_t *blob() { return bvh - (kFullBvh ? 0 : kBvhLeafOff); }
    AssetTable assets;
    float agentPos[kMaxAgents][3];
    float agentRot[kMaxAgents][4];
    int agentSlot[kMaxAgents];
    int agentType[kMaxAgents];
    int agentSkipLeaf[kMaxAgents];        // the agent's own leaf when its rays provably miss it, else -1
    // visibility ray queue
    float visDir[kMaxVisRays][3];
    uint16_t visInfo[kMaxVisRays];        // agent k | target slot << 3 | seesHiderCandidate << 11
    float *visOut[kMaxVisRays];
    int numVis;
    unsigned long long leafBar;           // mbarrier of the leaf-tile bulk copy
    // queued query (k_rays, nearest-first mode): per-ray state + the two CTA-wide work lists
    unsigned long long rayKey[kMaxRays];  // closest hit so far: float bits of t << 32 | (0x7FFFFFFF - leaf); low word ~0 = none
    float rayD[kMaxRays][3], rayInv[kMaxRays][3];
    uint16_t qa[kRayQueueA];              // (ray | leaf << 9): slab-entered leaves still to be re-checked
    uint16_t qb[kRayQueueB];              // (ray | leaf << 9): leaves whose hull has to be intersected
    int qaCount, qbCount;
};

// Ray vs one leaf in object space; operand order of hs_trace.cuh::traceLeaf, tables in shared.
__device__ __forceinline__ bool traceLeafS(const uint32_t *blob, const AssetTable &AT, int leaf,
                                           Vec3 o, Vec3 d, float t_max, float *hit_t)
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

    const ObjectData &O = AT.obj[obj];
    Vec3 lo = {O.aabb[0], O.aabb[1], O.aabb[2]}, hi = {O.aabb[3], O.aabb[4], O.aabb[5]};
    if (!aabbRayIntersects(lo, hi, oo, inv, 0.f, t_max)) return false;

    if (O.primType == 4) {
        float denom = od.z;
        if (denom == 0.f) return false;
        float t = -oo.z / denom;
        if (t < 0.f || t > t_max) return false;
        *hit_t = t;
        return true;
    }

    const HullData &H = AT.hull[obj];
    float tfirst = 0.f, tlast = t_max;
    bool entered = false;
    const int nf = H.numFaces;
    for (int f = 0; f < nf; f++) {
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
    if (!entered) return false;
    *hit_t = tfirst;
    return true;
}

// True when `o` lies strictly inside the convex hull of `leaf` (every face plane at least 1e-4
// away in object space).  traceLeafS then returns false for EVERY direction from `o`: no entering
// plane has t >= 0, so the "only back faces" rule rejects the leaf.  Every ray an agent casts
// starts at its own centre, so its own leaf — always the nearest slab entry (t = 0) — can be left
// out of the query without changing any result.
__device__ inline bool originInsideLeaf(const uint32_t *blob, const AssetTable &AT, int leaf, Vec3 o)
{
    const float *L = leafF(blob, leaf);
    int obj = int(blob[kBvhLeafOff + leaf * kLeafWords + kLeafInfo] >> 8) & 0xFF;
    if (AT.obj[obj].primType != 2) return false;
    Vec3 pos = {L[kLeafPos], L[kLeafPos + 1], L[kLeafPos + 2]};
    Quat rot = {L[kLeafRot], L[kLeafRot + 1], L[kLeafRot + 2], L[kLeafRot + 3]};
    Vec3 scale = {L[kLeafScale], L[kLeafScale + 1], L[kLeafScale + 2]};
    Vec3 oo = rotate(qinv(rot), o - pos);
    oo.x /= scale.x; oo.y /= scale.y; oo.z /= scale.z;
    const HullData &H = AT.hull[obj];
    for (int f = 0; f < H.numFaces; f++) {
        Vec3 n = {H.planes[f][0], H.planes[f][1], H.planes[f][2]};
        if (!(H.planes[f][3] - dot(n, oo) > 1e-4f)) return false;
    }
    return true;
}

// Warp-convergent closest-hit query.  Must be called by all 32 lanes of a warp; lanes without
// a ray pass active = false.  Returns the body slot of the closest hit or -1.
__device__ inline int traceRayWarp(const uint32_t *blob, const AssetTable &AT, bool active,
                                   Vec3 o, Vec3 d, float t_max, float *out_t)
{
    const Vec3 inv = {1.f / d.x, 1.f / d.y, 1.f / d.z};
    uint8_t stack[24];
    int sp = 0;
    int cur = -1;             // node whose children are being scanned, -1: pop next
    int ci = 0;               // next child index of cur
    int hitBits = 0;          // children of cur whose slab test passes without the t_max clamp
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;    // slab entry distances of cur's children
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    int pending = -1;         // leaf waiting for the shared leaf phase
    int hitLeaf = -1;
    bool done = !active;
    if (active) { stack[0] = 0; sp = 1; }

    while (true) {
        if (!done && pending < 0) {
            if (cur < 0) {
                if (sp == 0) {
                    done = true;
                } else {
                    cur = stack[--sp];
                    ci = 0;
                    const float4 *nf = reinterpret_cast<const float4 *>(blob + kBvhNodeOff + cur * kNodeWords);
                    const float4 lox = nf[0], loy = nf[1], loz = nf[2], hix = nf[3], hiy = nf[4], hiz = nf[5];
                    const int4 ch = *reinterpret_cast<const int4 *>(blob + kBvhNodeOff + cur * kNodeWords + kNodeChild);
                    c0 = ch.x; c1 = ch.y; c2 = ch.z; c3 = ch.w;
                    hitBits = 0;
#define HS_SLAB(i, LX, LY, LZ, HX, HY, HZ, A, C)                                              \
                    {                                                                          \
                        float tlx = inv.x * (LX - o.x), tly = inv.y * (LY - o.y), tlz = inv.z * (LZ - o.z); \
                        float tux = inv.x * (HX - o.x), tuy = inv.y * (HY - o.y), tuz = inv.z * (HZ - o.z); \
                        float mnx = fminf(tlx, tux), mny = fminf(tly, tuy), mnz = fminf(tlz, tuz);        \
                        float mxx = fmaxf(tlx, tux), mxy = fmaxf(tly, tuy), mxz = fmaxf(tlz, tuz);        \
                        A = fmaxf(mnx, fmaxf(mny, fmaxf(mnz, 0.f)));                           \
                        float b = fminf(mxx, fminf(mxy, fminf(mxz, INFINITY)));                \
                        if (C != kSentinel && A <= b) hitBits |= 1 << i;                       \
                    }
                    HS_SLAB(0, lox.x, loy.x, loz.x, hix.x, hiy.x, hiz.x, a0, c0)
                    HS_SLAB(1, lox.y, loy.y, loz.y, hix.y, hiy.y, hiz.y, a1, c1)
                    HS_SLAB(2, lox.z, loy.z, loz.z, hix.z, hiy.z, hiz.z, a2, c2)
                    HS_SLAB(3, lox.w, loy.w, loz.w, hix.w, hiy.w, hiz.w, a3, c3)
#undef HS_SLAB
                }
            }
            if (cur >= 0) {
                // scan children in order; stop at the first leaf that needs testing
                while (ci < 4) {
                    int i = ci++;
                    if (!((hitBits >> i) & 1)) continue;
                    float a = i == 0 ? a0 : (i == 1 ? a1 : (i == 2 ? a2 : a3));
                    if (!(a <= t_max)) continue;
                    int child = i == 0 ? c0 : (i == 1 ? c1 : (i == 2 ? c2 : c3));
                    if (child < 0) { pending = child & 0x7FFFFFFF; break; }
                    if (sp < 24) stack[sp++] = (uint8_t)child;
                }
                if (ci >= 4 && pending < 0) cur = -1;
            }
        }
        const unsigned act = __ballot_sync(0xFFFFFFFFu, !done);
        if (act == 0) break;
        const unsigned pend = __ballot_sync(0xFFFFFFFFu, pending >= 0);
        if (pend != 0 && (pend == act || __popc(pend) * HS_VOTE_DEN >= __popc(act) * HS_VOTE_NUM)) {
            if (pending >= 0) {
                float t;
                if (traceLeafS(blob, AT, pending, o, d, t_max, &t)) { t_max = t; hitLeaf = pending; }
                pending = -1;
                if (ci >= 4) cur = -1;
            }
        }
    }
    if (hitLeaf < 0) return -1;
    *out_t = t_max;
    return int(blob[kBvhLeafOff + hitLeaf * kLeafWords + kLeafInfo] & 0xFF);
}

// Default closest-hit query: no hierarchy walk at all.  A world has at most 52 leaves (~29 live),
// so every lane slab-tests ALL leaf boxes in one uniform loop — the whole warp reads the same
// leaf record (a shared-memory broadcast) and no lane diverges — remembering the set of boxes
// its ray enters (one 64-bit mask) and the nearest of them.  The nearest box is tested first
// (all lanes together); the few remaining candidates are re-checked against the now short
// t_max and tested with the same parked-lane vote as the reference-order walk.
// The closest hit is a property of the ray and the leaves, not of the visiting order or of the
// hierarchy; the only order-dependent case is an exact floating-point tie between two different
// leaves, which changes neither a lidar depth nor — short of two bodies sharing their entry face
// at the hit point — a visibility bit (tests/test_ray_order_gpu.py compares both modes bit-wise).
__device__ inline int traceRayFlat(const uint32_t *blob, const AssetTable &AT, bool active,
                                   Vec3 o, Vec3 d, float t_max, float *out_t, int skipLeaf = -1)
{
    const Vec3 inv = {1.f / d.x, 1.f / d.y, 1.f / d.z};
    const int numLeaves = min((int)blob[1], kMaxBodies);
    unsigned long long mask = 0ull;
    float bestA = INFINITY;
    int bestL = -1;
#define HS_LEAF_SLAB(L, A, OK)                                                                 \
    {                                                                                          \
        float tlx = inv.x * (L[0] - o.x), tly = inv.y * (L[1] - o.y), tlz = inv.z * (L[2] - o.z); \
        float tux = inv.x * (L[3] - o.x), tuy = inv.y * (L[4] - o.y), tuz = inv.z * (L[5] - o.z); \
        float mnx = fminf(tlx, tux), mny = fminf(tly, tuy), mnz = fminf(tlz, tuz);            \
        float mxx = fmaxf(tlx, tux), mxy = fmaxf(tly, tuy), mxz = fmaxf(tlz, tuz);            \
        A = fmaxf(mnx, fmaxf(mny, fmaxf(mnz, 0.f)));                                           \
        OK = A <= fminf(mxx, fminf(mxy, fminf(mxz, t_max)));                                   \
    }
    for (int l = 0; l < numLeaves; l++) {
        const float *L = leafF(blob, l);
        float a;
        bool ok;
        HS_LEAF_SLAB(L, a, ok)
        if (ok && active && l != skipLeaf) {
            mask |= 1ull << l;
            if (a < bestA) { bestA = a; bestL = l; }
        }
    }
    int hitLeaf = -1;
    if (bestL >= 0) {
        float t;
        if (traceLeafS(blob, AT, bestL, o, d, t_max, &t)) { t_max = t; hitLeaf = bestL; }
        mask &= ~(1ull << bestL);
    }
    int pending = -1;
    bool done = mask == 0ull;
    while (true) {
        if (!done && pending < 0) {
            while (mask) {
                const int l = __ffsll((long long)mask) - 1;
                mask &= mask - 1ull;
                const float *L = leafF(blob, l);
                float a;
                bool ok;
                HS_LEAF_SLAB(L, a, ok)
                if (ok) { pending = l; break; }
            }
            if (pending < 0) done = true;
        }
        const unsigned act = __ballot_sync(0xFFFFFFFFu, !done);
        if (act == 0) break;
        const unsigned pend = __ballot_sync(0xFFFFFFFFu, pending >= 0);
        if (pend == act) {
            if (pending >= 0) {
                float t;
                if (traceLeafS(blob, AT, pending, o, d, t_max, &t)) { t_max = t; hitLeaf = pending; }
                pending = -1;
            }
        }
    }
#undef HS_LEAF_SLAB
    if (hitLeaf < 0) return -1;
    *out_t = t_max;
    return int(blob[kBvhLeafOff + hitLeaf * kLeafWords + kLeafInfo] & 0xFF);
}

// ---- queued flat query (what k_rays runs in nearest-first mode) -------------------------------------
// traceRayFlat above leaves every ray's follow-up work — the leaves its ray enters besides the
// nearest one — to the ray's own lane: a few rays per warp have long candidate lists, so the
// re-checks ran at ~6 and the hull tests at ~2.6 of 32 lanes and cost more than the first tests.
// The closest hit is a MINIMUM over leaves (a leaf's hit distance does not depend on the bound it
// is tested with; the bound only prunes), so the follow-up work of all rays of a world is put on
// two CTA-wide lists and done at full lanes: list A = (ray, leaf) slab re-checks against the
// ray's best hit so far, list B = the survivors' hull tests; results merge with atomicMin on a
// packed (t, leaf) key per ray.  A full list falls back to doing the item in place.
__device__ __forceinline__ unsigned long long rayKeyPack(float t, int leaf)
{
    return ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)(0x7FFFFFFFu - (unsigned)leaf);
}
__device__ __forceinline__ unsigned long long rayKeyNone(float tmax)
{
    return ((unsigned long long)__float_as_uint(tmax) << 32) | 0xFFFFFFFFull;
}
__device__ __forceinline__ float rayKeyT(unsigned long long k) { return __uint_as_float((unsigned)(k >> 32)); }

__device__ __forceinline__ bool leafSlab(const float *L, Vec3 o, Vec3 inv, float t_max, float *entry)
{
    float tlx = inv.x * (L[0] - o.x), tly = inv.y * (L[1] - o.y), tlz = inv.z * (L[2] - o.z);
    float tux = inv.x * (L[3] - o.x), tuy = inv.y * (L[4] - o.y), tuz = inv.z * (L[5] - o.z);
    float mnx = fminf(tlx, tux), mny = fminf(tly, tuy), mnz = fminf(tlz, tuz);
    float mxx = fmaxf(tlx, tux), mxy = fmaxf(tly, tuy), mxz = fmaxf(tlz, tuz);
    const float a = fmaxf(mnx, fmaxf(mny, fmaxf(mnz, 0.f)));
    *entry = a;
    return a <= fminf(mxx, fminf(mxy, fminf(mxz, t_max)));
}

// hull test of (ray r, leaf) merged into the ray's key; `o` = the ray's origin
template <class SM>
__device__ __forceinline__ void rayTestLeaf(SM &sm, int r, int leaf, Vec3 o)
{
    const Vec3 d = {sm.rayD[r][0], sm.rayD[r][1], sm.rayD[r][2]};
    const float best = rayKeyT(*(volatile unsigned long long *)&sm.rayKey[r]);
    float t;
    if (traceLeafS(sm.blob(), sm.assets, leaf, o, d, best, &t)) atomicMin(&sm.rayKey[r], rayKeyPack(t, leaf));
}

// slab re-check of (ray r, leaf) against the ray's best hit so far; survivors go to list B
template <class SM>
__device__ __forceinline__ void rayRecheckLeaf(SM &sm, int r, int leaf, Vec3 o)
{
    const Vec3 inv = {sm.rayInv[r][0], sm.rayInv[r][1], sm.rayInv[r][2]};
    const float best = rayKeyT(*(volatile unsigned long long *)&sm.rayKey[r]);
    float a;
    if (!leafSlab(leafF(sm.blob(), leaf), o, inv, best, &a)) return;
    const int i = atomicAdd(&sm.qbCount, 1);
    if (i < kRayQueueB) sm.qb[i] = uint16_t(r | (leaf << 9));
    else rayTestLeaf(sm, r, leaf, o);
}

// kNearestFirst = false walks the BVH in the reference's visiting order (ties resolved
// identically); true is the flat nearest-first query and the default (hs_sim.cu: HIDESEEK_RAY_ORDER=reference selects the other at run time).
template <bool kNearestFirst>
__device__ __forceinline__ int traceRaySel(const uint32_t *blob, const AssetTable &AT, bool active,
                                           Vec3 o, Vec3 d, float t_max, float *out_t, int skipLeaf = -1)
{
    if (kNearestFirst) return traceRayFlat(blob, AT, active, o, d, t_max, out_t, skipLeaf);
    return traceRayWarp(blob, AT, active, o, d, t_max, out_t);
}

template <bool kNearestFirst>
__global__ void __launch_bounds__(kRayThreads, HS_RAY_MINB) k_rays(DeviceState S, float cosThreshold)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    using RaySmem = RaySmemT<!kNearestFirst>;
    RaySmem &sm = *reinterpret_cast<RaySmem *>(smemRaw);
    const int w = blockIdx.x;
    const int tid = threadIdx.x;
    const int A = S.p.agentsPerWorld;

    // ---- stage the world's BVH tile + the asset table ---------------------------------------
    const uint32_t *g = BVH(S, w);
    const int numNodes = min((int)g[0], kMaxNodes), numLeaves = min((int)g[1], kMaxBodies);
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(g);
        uint4 *dst = reinterpret_cast<uint4 *>(sm.blob());
        // the flat query never touches the header or the inner nodes (and has no room for them)
        const int nodeVec = kNearestFirst ? 0 : (kBvhNodeOff + numNodes * kNodeWords) / 4;
        for (int i = tid; i < nodeVec; i += kRayThreads) dst[i] = src[i];
        // the world's leaf records are one contiguous, 16-byte aligned run of numLeaves * 80 bytes:
        // ONE bulk-async copy (TMA engine), completion counted on an mbarrier; it overlaps the
        // staging of the asset table and the agent poses below
        static_assert((kBvhLeafOff * 4) % 16 == 0 && (kLeafWords * 4) % 16 == 0 && (kBvhWords * 4) % 16 == 0, "bulk copy alignment");
        if (tid == 0) {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sm.leafBar);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            if (numLeaves > 0) {
                const uint32_t bytes = uint32_t(numLeaves) * kLeafWords * 4u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"((uint32_t)__cvta_generic_to_shared(sm.blob() + kBvhLeafOff)), "l"(g + kBvhLeafOff), "r"(bytes), "r"(bar)
                             : "memory");
            }
        }
        const uint32_t *ca = reinterpret_cast<const uint32_t *>(&c_assets);
        uint32_t *sa = reinterpret_cast<uint32_t *>(&sm.assets);
        for (int i = tid; i < int(sizeof(AssetTable) / 4); i += kRayThreads) sa[i] = ca[i];
    }
    if (tid < A) {
        int slot = WB(S, WB_IFACE_SLOT + tid, w);
        sm.agentSlot[tid] = slot == 0xFF ? -1 : slot;
        sm.agentType[tid] = WB(S, WB_IFACE_TYPE + tid, w);
        if (slot != 0xFF) {
            Vec3 p = loadPos(S, slot, w);
            Quat q = loadRot(S, slot, w);
            sm.agentPos[tid][0] = p.x; sm.agentPos[tid][1] = p.y; sm.agentPos[tid][2] = p.z;
            sm.agentRot[tid][0] = q.w; sm.agentRot[tid][1] = q.x; sm.agentRot[tid][2] = q.y; sm.agentRot[tid][3] = q.z;
        }
    }
    if (tid == 0) sm.numVis = 0;
    __syncthreads();
    if (numLeaves > 0) {                     // leaf tile landed?
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sm.leafBar);
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
    }
    if (tid < A) {
        int skip = -1;
        if (sm.agentSlot[tid] >= 0) {
            const int leaf = metaLeaf(META(S, sm.agentSlot[tid], w));
            Vec3 p = {sm.agentPos[tid][0], sm.agentPos[tid][1], sm.agentPos[tid][2]};
            if (leaf < numLeaves && originInsideLeaf(sm.blob(), sm.assets, leaf, p)) skip = leaf;
        }
        sm.agentSkipLeaf[tid] = skip;
    }

    // ---- visibility: cone test per (agent, target); survivors join the ray queue --------------
    constexpr int kTargets = kMaxBoxes + kMaxRamps + kMaxAgents;      // 17 lanes of the old kernel
    for (int e = tid; e < A * kTargets; e += kRayThreads) {
        const int k = e / kTargets, lane = e % kTargets;
        if (sm.agentSlot[k] < 0) continue;
        const size_t row = size_t(w) * A + k;
        float *out = nullptr;
        int other = -1;
        bool hiderTarget = false;
        if (lane < kMaxBoxes) {
            out = S.exBoxVis + row * kMaxBoxes + lane;
            if (lane < WI(S, WF_NUM_BOXES, w)) other = WB(S, WB_BOX_SLOT + lane, w);
        } else if (lane < kMaxBoxes + kMaxRamps) {
            int i = lane - kMaxBoxes;
            out = S.exRampVis + row * kMaxRamps + i;
            if (i < WI(S, WF_NUM_RAMPS, w)) other = WB(S, WB_RAMP_SLOT + i, w);
        } else {
            int j = lane - kMaxBoxes - kMaxRamps;
            int numAgents = WI(S, WF_NUM_AGENTS, w);
            if (j == k && j < numAgents) continue;
            int outIdx = j < numAgents ? otherAgentIndex(k, j) : j - 1;
            if (outIdx < 0 || outIdx >= kMaxAgents - 1) continue;
            out = S.exAgentVis + row * (kMaxAgents - 1) + outIdx;
            if (j < numAgents) {
                other = WB(S, WB_IFACE_SLOT + j, w);
                hiderTarget = sm.agentType[k] == AGENT_SEEKER && WB(S, WB_IFACE_TYPE + j, w) == AGENT_HIDER;
            }
        }
        if (other < 0) { *out = 0.f; continue; }
        Vec3 apos = {sm.agentPos[k][0], sm.agentPos[k][1], sm.agentPos[k][2]};
        Quat arot = {sm.agentRot[k][0], sm.agentRot[k][1], sm.agentRot[k][2], sm.agentRot[k][3]};
        Vec3 afwd = rotate(arot, v3(0, 1, 0));
        Vec3 to = loadPos(S, other, w) - apos;
        float cosA = dot(normalize(to), afwd);
        if (cosA < cosThreshold) { *out = 0.f; continue; }
        int q = atomicAdd(&sm.numVis, 1);
        sm.visDir[q][0] = to.x; sm.visDir[q][1] = to.y; sm.visDir[q][2] = to.z;
        sm.visInfo[q] = uint16_t(k | (other << 3) | (hiderTarget ? 1 << 11 : 0));
        sm.visOut[q] = out;
    }
    __syncthreads();

    // ---- all rays of the world: [0, A*30) lidar, then the visibility queue ------------------
    const int numLidar = A * 30;
    const int numRays = numLidar + sm.numVis;
    const int rounded = (numRays + 31) & ~31;
    if (kNearestFirst) {
        if (tid == 0) { sm.qaCount = 0; sm.qbCount = 0; }
        __syncthreads();
        // ---- stage 1, lane = ray: slab-test every leaf, intersect the nearest-entry leaf's hull,
        //      queue the other entered leaves ----------------------------------------------------------
        for (int r = tid; r < numRays; r += kRayThreads) {
            Vec3 d = {1.f, 0.f, 0.f};
            float tmax = 0.f;
            int k;
            bool active = true;
            if (r < numLidar) {
                k = r / 30;
                const int i = r - k * 30;
                if (sm.agentSlot[k] < 0) {
                    active = false;
                } else {
                    Quat rot = {sm.agentRot[k][0], sm.agentRot[k][1], sm.agentRot[k][2], sm.agentRot[k][3]};
                    Vec3 fwd = rotate(rot, v3(0, 1, 0));
                    Vec3 right = rotate(rot, v3(1, 0, 0));
                    float x = c_lidarDir[i][0], y = c_lidarDir[i][1];
                    d = normalize(x * right + y * fwd);
                    tmax = 200.f;
                }
            } else {
                const int q = r - numLidar;
                k = sm.visInfo[q] & 7;
                d = {sm.visDir[q][0], sm.visDir[q][1], sm.visDir[q][2]};
                tmax = 1.f;
            }
            unsigned long long key = rayKeyNone(tmax);
            if (active) {
                const Vec3 o = {sm.agentPos[k][0], sm.agentPos[k][1], sm.agentPos[k][2]};
                const Vec3 inv = {1.f / d.x, 1.f / d.y, 1.f / d.z};
                const int skipLeaf = sm.agentSkipLeaf[k];
                unsigned long long mask = 0ull;
                float bestA = INFINITY;
                int bestL = -1;
                for (int l = 0; l < numLeaves; l++) {
                    float a;
                    const bool ok = leafSlab(leafF(sm.blob(), l), o, inv, tmax, &a);
                    if (ok && l != skipLeaf) {
                        mask |= 1ull << l;
                        if (a < bestA) { bestA = a; bestL = l; }
                    }
                }
                if (bestL >= 0) {
                    float t;
                    if (traceLeafS(sm.blob(), sm.assets, bestL, o, d, tmax, &t)) key = rayKeyPack(t, bestL);
                    mask &= ~(1ull << bestL);
                }
                sm.rayD[r][0] = d.x; sm.rayD[r][1] = d.y; sm.rayD[r][2] = d.z;
                sm.rayInv[r][0] = inv.x; sm.rayInv[r][1] = inv.y; sm.rayInv[r][2] = inv.z;
                sm.rayKey[r] = key;
                if (mask) {
                    int pos = atomicAdd(&sm.qaCount, __popcll(mask));
                    for (; mask; mask &= mask - 1ull, pos++) {
                        const int l = __ffsll((long long)mask) - 1;
                        if (pos < kRayQueueA) sm.qa[pos] = uint16_t(r | (l << 9));
                        else rayRecheckLeaf(sm, r, l, o);              // list full: in place
                    }
                }
            } else {
                sm.rayKey[r] = key;
            }
        }
        __syncthreads();
        // ---- stage 2, lane = queued (ray, leaf): slab re-check against the ray's best hit ----------
        const int nA = min(sm.qaCount, kRayQueueA);
        for (int i = tid; i < nA; i += kRayThreads) {
            const int e = sm.qa[i], r = e & 511, l = e >> 9;
            const int k = r < numLidar ? r / 30 : (sm.visInfo[r - numLidar] & 7);
            rayRecheckLeaf(sm, r, l, {sm.agentPos[k][0], sm.agentPos[k][1], sm.agentPos[k][2]});
        }
        __syncthreads();
        // ---- stage 3, lane = surviving (ray, leaf): hull intersection, merged by atomicMin ----------
        const int nB = min(sm.qbCount, kRayQueueB);
        for (int i = tid; i < nB; i += kRayThreads) {
            const int e = sm.qb[i], r = e & 511, l = e >> 9;
            const int k = r < numLidar ? r / 30 : (sm.visInfo[r - numLidar] & 7);
            rayTestLeaf(sm, r, l, {sm.agentPos[k][0], sm.agentPos[k][1], sm.agentPos[k][2]});
        }
        __syncthreads();
        // ---- results ------------------------------------------------------------------------------------
        for (int r = tid; r < numRays; r += kRayThreads) {
            const unsigned long long key = sm.rayKey[r];
            const unsigned low = (unsigned)key;
            const int hit = low == 0xFFFFFFFFu ? -1
                          : int(sm.blob()[kBvhLeafOff + int(0x7FFFFFFFu - low) * kLeafWords + kLeafInfo] & 0xFF);
            const float t = rayKeyT(key);
            if (r < numLidar) {
                const int k = r / 30;
                if (sm.agentSlot[k] < 0) continue;
                S.exLidar[(size_t(w) * A + k) * 30 + (r - k * 30)] = hit < 0 ? 0.f : t;
                if (S.p.debugFlags & 1u) S.dbgLidarHit[(size_t(w) * A + k) * 30 + (r - k * 30)] = int8_t(hit);
            } else {
                const int q = r - numLidar;
                const int info = sm.visInfo[q];
                const bool vis = hit == ((info >> 3) & 0xFF);
                *sm.visOut[q] = vis ? 1.f : 0.f;
                // CPU-backend side effect (src/sim.cpp:700-705)
                if (vis && (info >> 11)) WI(S, WF_HIDER_REWARD, w) = __float_as_int(-1.f);
            }
        }
        return;
    }
    for (int r = tid; r < rounded; r += kRayThreads) {
        bool active = r < numRays;
        Vec3 o = {0.f, 0.f, 0.f}, d = {1.f, 0.f, 0.f};
        float tmax = 0.f;
        int k = 0, q = 0;
        const bool isLidar = r < numLidar;
        if (active) {
            if (isLidar) {
                k = r / 30;
                const int i = r - k * 30;
                if (sm.agentSlot[k] < 0) {
                    active = false;
                } else {
                    Quat rot = {sm.agentRot[k][0], sm.agentRot[k][1], sm.agentRot[k][2], sm.agentRot[k][3]};
                    Vec3 fwd = rotate(rot, v3(0, 1, 0));
                    Vec3 right = rotate(rot, v3(1, 0, 0));
                    float x = c_lidarDir[i][0], y = c_lidarDir[i][1];
                    d = normalize(x * right + y * fwd);
                    tmax = 200.f;
                }
            } else {
                q = r - numLidar;
                k = sm.visInfo[q] & 7;
                d = {sm.visDir[q][0], sm.visDir[q][1], sm.visDir[q][2]};
                tmax = 1.f;
            }
            o = {sm.agentPos[k][0], sm.agentPos[k][1], sm.agentPos[k][2]};
        }
        float t = 0.f;
        int hit = traceRaySel<kNearestFirst>(sm.blob(), sm.assets, active, o, d, tmax, &t, active ? sm.agentSkipLeaf[k] : -1);
        if (!active) continue;
        if (isLidar) {
            S.exLidar[(size_t(w) * A + k) * 30 + (r - k * 30)] = hit < 0 ? 0.f : t;
            if (S.p.debugFlags & 1u) S.dbgLidarHit[(size_t(w) * A + k) * 30 + (r - k * 30)] = int8_t(hit);
        } else {
            const int info = sm.visInfo[q];
            const bool vis = hit == ((info >> 3) & 0xFF);
            *sm.visOut[q] = vis ? 1.f : 0.f;
            // CPU-backend side effect (src/sim.cpp:700-705)
            if (vis && (info >> 11)) WI(S, WF_HIDER_REWARD, w) = __float_as_int(-1.f);
        }
    }
}

// rewardsVisSystem + outputRewardsDonesSystem + updateEpisodeResultsSystem (src/sim.cpp:763-893).
// A world has at most 3 x 3 (seeker, hider) occlusion rays, so three worlds share a warp (lanes
// 0-8, 9-17, 18-26); the reference's early `break` only skips work, "any seeker sees any hider"
// is order-free.  Lanes < A of each group then write the per-agent reward / done rows and the
// group's first lane the per-world episode result.
constexpr int kRewardWarps = 4;
constexpr int kRewardWorldsPerWarp = 3;
constexpr int kRewardLanes = 9;

template <bool kNearestFirst>
__global__ void __launch_bounds__(kRewardWarps * 32) k_rewards(DeviceState S, float cosThreshold)
{
    __shared__ AssetTable s_assets;
    {
        const uint32_t *ca = reinterpret_cast<const uint32_t *>(&c_assets);
        uint32_t *sa = reinterpret_cast<uint32_t *>(&s_assets);
        for (int i = threadIdx.x; i < int(sizeof(AssetTable) / 4); i += blockDim.x) sa[i] = ca[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int grp = lane / kRewardLanes, lg = lane - grp * kRewardLanes;
    const int warp = blockIdx.x * kRewardWarps + (threadIdx.x >> 5);
    const int w = warp * kRewardWorldsPerWarp + grp;
    const bool valid = grp < kRewardWorldsPerWarp && w < S.p.numWorlds;
    const int wc = valid ? w : 0;                  // lanes without a world still join the warp votes
    const unsigned gmask = valid ? (((1u << kRewardLanes) - 1u) << (grp * kRewardLanes)) : 0u;
    const int A = S.p.agentsPerWorld;
    const int curStep = WI(S, WF_EPISODE_STEP, wc);
    float teamReward = __int_as_float(WI(S, WF_HIDER_REWARD, wc));
    const int numHiders = WI(S, WF_NUM_HIDERS, wc), numSeekers = WI(S, WF_NUM_SEEKERS, wc);

    bool active = valid && lg < numHiders * numSeekers;
    int hslot = -1, skip = -1;
    Vec3 spos = {0.f, 0.f, 0.f}, toHider = {1.f, 0.f, 0.f};
    if (active) {
        int si = lg / numHiders, hi = lg - si * numHiders;
        int sslot = WB(S, WB_SEEKER_SLOT + si, w);
        skip = metaLeaf(META(S, sslot, w));
        hslot = WB(S, WB_HIDER_SLOT + hi, w);
        const Pose sp = loadPose(S, REC_POSE, sslot, w);
        spos = sp.x;
        Vec3 sfwd = rotate(sp.q, v3(0, 1, 0));
        toHider = loadPos(S, hslot, w) - spos;
        float cosA = dot(normalize(toHider), sfwd);
        if (cosA < cosThreshold) active = false;
        else if (!originInsideLeaf(BVH(S, w), s_assets, skip, spos)) skip = -1;
    }
    float t;
    int hit = traceRaySel<kNearestFirst>(BVH(S, wc), s_assets, active, spos, toHider, 1.f, &t, skip);
    const unsigned seen = __ballot_sync(0xFFFFFFFFu, active && hit == hslot);
    if (!valid) return;
    if (seen & gmask) teamReward = -1.f;

    if (lg < A) {
        const int k = lg;
        int slot = WB(S, WB_IFACE_SLOT + k, w);
        if (slot != 0xFF) {
            size_t row = size_t(w) * A + k;
            if (curStep == 0) S.exDone[row] = 0;
            if (curStep < kNumPrepSteps - 1) {
                S.exReward[row] = 0.f;
            } else {
                if (curStep == kEpisodeLen - 1) S.exDone[row] = 1;
                float r = teamReward;
                if (WB(S, WB_IFACE_TYPE + k, w) == AGENT_SEEKER) r *= -1.f;
                Vec3 pos = loadPos(S, slot, w);
                if (fabsf(pos.x) >= 18.f || fabsf(pos.y) >= 18.f) r -= 10.f;
                S.exReward[row] = r;
            }
        }
    }
    if (lg == 0) {
        WI(S, WF_HIDER_REWARD, w) = __float_as_int(teamReward);
        float *res = S.exEpisodeResult + size_t(w) * 2;
        int s0 = WI(S, WF_SCORE0, w), s1 = WI(S, WF_SCORE1, w);
        if (curStep == 0) { res[0] = 0.f; res[1] = 0.f; s0 = 0; s1 = 0; }
        if (curStep >= kNumPrepSteps) {
            bool seekersFirst = WI(S, WF_SEEKERS_FIRST, w) != 0;
            bool hidden = teamReward == 1.f;
            if (hidden == seekersFirst) s1 += 1; else s0 += 1;
        }
        if (curStep == kEpisodeLen - 1) {
            if (s0 > s1) { res[0] = 1.f; res[1] = 0.f; }
            else if (s0 < s1) { res[0] = 0.f; res[1] = 1.f; }
            else { res[0] = 0.5f; res[1] = 0.5f; }
        }
        WI(S, WF_SCORE0, w) = s0;
        WI(S, WF_SCORE1, w) = s1;
    }
}

// actionSystem (src/sim.cpp:270-370).  Lock and grab mutate shared body state and the entity-id
// pool, so the agents of a world are processed in interface order by ONE lane; but the ray both
// actions cast (same origin, direction and reach) only depends on the geometry, which no action
// changes, so the rays of all agents are traced first, one lane per agent, 4 worlds per warp.
constexpr int kActLanes = 8, kActWarps = 4;
static_assert(kMaxAgents <= kActLanes, "one lane per agent interface");

template <bool kNearestFirst>
__global__ void __launch_bounds__(kActWarps * 32) k_action(DeviceState S)
{
    __shared__ AssetTable s_assets;
    __shared__ int s_hit[kActWarps * 32];
    __shared__ float s_t[kActWarps * 32];
    {
        const uint32_t *ca = reinterpret_cast<const uint32_t *>(&c_assets);
        uint32_t *sa = reinterpret_cast<uint32_t *>(&s_assets);
        for (int i = threadIdx.x; i < int(sizeof(AssetTable) / 4); i += blockDim.x) sa[i] = ca[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int grp = lane / kActLanes, lg = lane % kActLanes;
    const int w = (blockIdx.x * kActWarps + (threadIdx.x >> 5)) * (32 / kActLanes) + grp;
    const bool valid = w < S.p.numWorlds;
    const int wc = valid ? w : 0;                  // every lane joins the warp votes of the query
    const int A = S.p.agentsPerWorld;
    const int nDyn = WI(S, WF_NUM_DYN, wc);

    bool need = false;
    int skip = -1;
    Vec3 ro = {0.f, 0.f, 0.f}, rd = {0.f, 1.f, 0.f};
    if (valid && lg < A) {
        const int slot = WB(S, WB_IFACE_SLOT + lg, w);
        if (slot != 0xFF && !agentFrozen(S, w, WB(S, WB_IFACE_TYPE + lg, w))) {
            const int32_t *act = S.exAction + (size_t(w) * A + lg) * 5;
            const bool lock = act[4] == 1;
            const bool grab = act[3] == 1 && JMETA(S, slot - nDyn, w) == 0xFFFFFFFFu;   // else: release
            if (lock || grab) {
                need = true;
                const Pose p = loadPose(S, REC_POSE, slot, w);
                ro = p.x + 0.5f * v3(0, 0, 1);
                rd = rotate(p.q, v3(0, 1, 0));
                skip = metaLeaf(META(S, slot, w));
                if (!originInsideLeaf(BVH(S, w), s_assets, skip, ro)) skip = -1;
            }
        }
    }
    float t = 0.f;
    int hit = -1;
    if (__any_sync(0xFFFFFFFFu, need))       // warp-uniform: no lock / grab in these 4 worlds, no query
        hit = traceRaySel<kNearestFirst>(BVH(S, wc), s_assets, need, ro, rd, 2.5f, &t, skip);
    s_hit[threadIdx.x] = need ? hit : -1;
    s_t[threadIdx.x] = t;
    __syncwarp();
    if (!valid || lg != 0) return;

    const int base = threadIdx.x;          // lane 0 of the group: agent k's ray is at base + k
    for (int k = 0; k < A; k++) {
        int slot = WB(S, WB_IFACE_SLOT + k, w);
        if (slot == 0xFF) continue;
        int type = WB(S, WB_IFACE_TYPE + k, w);
        if (agentFrozen(S, w, type)) continue;
        int32_t *act = S.exAction + (size_t(w) * A + k) * 5;
        const int hitK = s_hit[base + k];
        const float tK = s_t[base + k];

        if (act[4] == 1 && hitK >= 0) {
            uint32_t m = META(S, hitK, w);
            int owner = metaOwner(m), resp = metaResp(m);
            if (resp == RESP_STATIC) {
                if ((type == AGENT_SEEKER && owner == OWNER_SEEKER) ||
                    (type == AGENT_HIDER && owner == OWNER_HIDER)) {
                    resp = RESP_DYNAMIC; owner = OWNER_NONE;
                }
            } else if (owner == OWNER_NONE) {
                resp = RESP_STATIC;
                owner = type == AGENT_HIDER ? OWNER_HIDER : OWNER_SEEKER;
            }
            META(S, hitK, w) = packMeta(metaObj(m), resp, owner, metaEid(m), metaLeaf(m));
        }

        if (act[3] == 1) {
            int ak = slot - nDyn;       // DynAgent row
            uint32_t jm = JMETA(S, ak, w);
            if (jm != 0xFFFFFFFFu) {
                IdPool ids(S, w);
                ids.release((jm >> 8) & 0xFF);
                ids.commit();
                JMETA(S, ak, w) = 0xFFFFFFFFu;
            } else if (hitK >= 0) {
                uint32_t m = META(S, hitK, w);
                if (metaOwner(m) == OWNER_NONE && metaResp(m) == RESP_DYNAMIC) {
                    const Pose p = loadPose(S, REC_POSE, slot, w);
                    Vec3 rok = p.x + 0.5f * v3(0, 0, 1);
                    Vec3 rdk = rotate(p.q, v3(0, 1, 0));
                    Vec3 opos = loadPos(S, hitK, w);
                    Quat orot = loadRot(S, hitK, w);
                    Vec3 r1 = 1.25f * v3(0, 1, 0) + 0.5f * v3(0, 0, 1);
                    Vec3 hitPos = rok + rdk * tK;
                    Vec3 r2 = rotate(qinv(orot), hitPos - opos);
                    Quat a2 = qnormalize(qinv(orot) * p.q);
                    IdPool ids(S, w);
                    int eid = ids.acquire();
                    ids.commit();
                    int seq = WI(S, WF_JOINT_SEQ, w);
                    WI(S, WF_JOINT_SEQ, w) = seq + 1;
                    JMETA(S, ak, w) = (uint32_t(seq & 0xFFFF) << 16) | (uint32_t(eid) << 8) | uint32_t(hitK);
                    JF(S, JF_A1W, ak, w) = 1.f; JF(S, JF_A1X, ak, w) = 0.f;
                    JF(S, JF_A1Y, ak, w) = 0.f; JF(S, JF_A1Z, ak, w) = 0.f;
                    JF(S, JF_A2W, ak, w) = a2.w; JF(S, JF_A2X, ak, w) = a2.x;
                    JF(S, JF_A2Y, ak, w) = a2.y; JF(S, JF_A2Z, ak, w) = a2.z;
                    JF(S, JF_SEP, ak, w) = tK - 1.25f;
                    JF(S, JF_R1X, ak, w) = r1.x; JF(S, JF_R1Y, ak, w) = r1.y; JF(S, JF_R1Z, ak, w) = r1.z;
                    JF(S, JF_R2X, ak, w) = r2.x; JF(S, JF_R2Y, ak, w) = r2.y; JF(S, JF_R2Z, ak, w) = r2.z;
                }
            }
        }

        act[0] = 2; act[1] = 2; act[2] = 2; act[3] = 0; act[4] = 0;
    }
}

}  // namespace hs
