// Per-world 4-wide BVH: leaf update, grow-only refit, top-down rebuild.
//
// Behavioural contract (reference CPU backend):
//   BVH::updateLeafPosition / expandAABBWithMotion   M/src/physics/broadphase.cpp:440-485
//   BVH::refitLeaf (grow-only min/max, stop when not expanded)   broadphase.cpp:550-647
//   BVH::rebuild (two-level midpoint split, LIFO stack)          broadphase.cpp:47-285
//   BVH::updateTree (rebuild only when forced by reset)           broadphase.cpp:649-656
// Node / leaf records live in the world-major blob described in hs_state.cuh.
#pragma once

#include "hs_state.cuh"
#include "hs_levelgen.cuh"   // c_assets

namespace hs {

__device__ __forceinline__ float *nodeF(uint32_t *blob, int n)
{
    return reinterpret_cast<float *>(blob + kBvhNodeOff + n * kNodeWords);
}
__device__ __forceinline__ int32_t *nodeI(uint32_t *blob, int n)
{
    return reinterpret_cast<int32_t *>(blob + kBvhNodeOff + n * kNodeWords);
}
__device__ __forceinline__ float *leafF(uint32_t *blob, int l)
{
    return reinterpret_cast<float *>(blob + kBvhLeafOff + l * kLeafWords);
}
__device__ __forceinline__ const float *nodeF(const uint32_t *blob, int n)
{
    return reinterpret_cast<const float *>(blob + kBvhNodeOff + n * kNodeWords);
}
__device__ __forceinline__ const int32_t *nodeI(const uint32_t *blob, int n)
{
    return reinterpret_cast<const int32_t *>(blob + kBvhNodeOff + n * kNodeWords);
}
__device__ __forceinline__ const float *leafF(const uint32_t *blob, int l)
{
    return reinterpret_cast<const float *>(blob + kBvhLeafOff + l * kLeafWords);
}

__device__ __forceinline__ float atomicMinF(float *addr, float v)
{
    if (!signbit(v)) return __int_as_float(atomicMin((int *)addr, __float_as_int(v)));
    return __uint_as_float(atomicMax((unsigned int *)addr, __float_as_uint(v)));
}
__device__ __forceinline__ float atomicMaxF(float *addr, float v)
{
    if (!signbit(v)) return __int_as_float(atomicMax((int *)addr, __float_as_int(v)));
    return __uint_as_float(atomicMin((unsigned int *)addr, __float_as_uint(v)));
}

// updateLeafPositionsEntry + refitEntry fused: one thread per (slot, world).  The thread that
// recomputes a body's (velocity-expanded) leaf box also grows the node slots above it, so the
// box never makes a round trip through HBM.  min/max are exact and commutative and a node slot
// always encloses its subtree, so the result does not depend on the order leaves are processed
// in (sequential on the CPU).  Worlds whose tree is about to be rebuilt (force_rebuild, set by a
// reset) only refresh their leaf records: the rebuild that follows recomputes every node and the
// reference's refit after a rebuild finds nothing to grow.
// onlyIfReset: the third BVH update of a step (after resetSystem) sees unchanged bodies in every
// world that did not reset, so it is skipped there.
__global__ void k_leaf_refit(DeviceState S, bool onlyIfReset)
{
    const int N = S.p.numWorlds;
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)kMaxBodies * N) return;
    int s = int(idx % kMaxBodies), w = int(idx / kMaxBodies);
    if (s >= WI(S, WF_NUM_BODIES, w)) return;
    if (onlyIfReset && WI(S, WF_DID_RESET, w) == 0) return;

    uint32_t m = META(S, s, w);
    const Pose pose = loadPose(S, REC_POSE, s, w);
    Vec3 pos = pose.x;
    Quat rot = pose.q;
    Vec3 scale = loadScale(S, s, w);
    float4 vel4 = recPtr(S, REC_VEL, s, w)[0];

    AABB box = aabbApplyTRS(objAABB(metaObj(m)), pos, rot, scale);

    // physics.cpp:104-111: velocity expansion 2*dt, accel expansion 100*dt*dt
    const float vexp = 2.f * kDeltaT;
    const float aexp = 100.f * kDeltaT * kDeltaT;
    float a[6] = {box.lo.x, box.lo.y, box.lo.z, box.hi.x, box.hi.y, box.hi.z};
    float v[3] = {vel4.x, vel4.y, vel4.z};
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float pd = vexp * v[i];
        float mn = pd - aexp, mx = pd + aexp;
        if (mn < 0.f) a[i] += mn;
        if (mx > 0.f) a[3 + i] += mx;
    }

    uint32_t *blob = BVH(S, w);
    float4 *L4 = reinterpret_cast<float4 *>(blob + kBvhLeafOff + metaLeaf(m) * kLeafWords);
    const float4 keep = L4[1];                       // .z = parent word, .w = slot | obj << 8
    L4[0] = make_float4(a[0], a[1], a[2], a[3]);
    L4[1] = make_float4(a[4], a[5], keep.z, keep.w);
    L4[2] = make_float4(pos.x, pos.y, pos.z, rot.w);
    L4[3] = make_float4(rot.x, rot.y, rot.z, scale.x);
    L4[4] = make_float4(scale.y, scale.z, 0.f, 0.f);
    if (blob[2] != 0) return;                        // rebuild pending

    const uint32_t lp = __float_as_uint(keep.z);
    int node = int(lp >> 2), sub = int(lp & 3);
    bool expanded = false;
    {
        float *nf = nodeF(blob, node);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float old = nf[k * 4 + sub];
            if (a[k] < old) { nf[k * 4 + sub] = a[k]; expanded = true; }
            float oldh = nf[(3 + k) * 4 + sub];
            if (a[3 + k] > oldh) { nf[(3 + k) * 4 + sub] = a[3 + k]; expanded = true; }
        }
    }
    if (!expanded) return;

    int child = node;
    node = nodeI(blob, node)[kNodeParent];
    while (node != kSentinel) {
        const int32_t *ni = nodeI(blob, node);
        int off = 0;
#pragma unroll
        for (int j = 3; j >= 0; j--) if (ni[kNodeChild + j] == child) off = j;
        float *nf = nodeF(blob, node);
        bool ex = false;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float old = atomicMinF(&nf[k * 4 + off], a[k]);
            ex |= a[k] < old;
            float oldh = atomicMaxF(&nf[(3 + k) * 4 + off], a[3 + k]);
            ex |= a[3 + k] > oldh;
        }
        if (!ex) break;
        child = node;
        node = ni[kNodeParent];
    }
}

// BVH::rebuild of one world (no-op unless force_rebuild is set).
__device__ inline void rebuildWorldBvh(const DeviceState &S, int w)
{
    uint32_t *blob = BVH(S, w);
    if (blob[2] == 0) return;
    blob[2] = 0;

    const int numLeaves = (int)blob[1];
    int tmp = numLeaves - 1;
    int third = (tmp + 2) / 3;
    blob[0] = uint32_t((third > 1 ? third : 1) + numLeaves);   // numInternalNodes

    uint8_t sorted[kMaxBodies];
    for (int i = 0; i < numLeaves; i++) sorted[i] = (uint8_t)i;

    struct Entry { int nodeID, parentID, offset, numObjs; };
    Entry stack[64];
    stack[0] = {kSentinel, kSentinel, 0, numLeaves};
    int stackSize = 1;
    int curNode = 0;

    auto center = [&](int leaf) {
        const float *L = leafF(blob, leaf);
        Vec3 lo = {L[0], L[1], L[2]}, hi = {L[3], L[4], L[5]};
        return (lo + hi) * 0.5f;     // (pMin + pMax) / 2.f
    };

    auto midpointSplit = [&](int base, int n) {
        Vec3 cmin = {FLT_MAX, FLT_MAX, FLT_MAX}, cmax = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = 0; i < n; i++) {
            Vec3 c = center(sorted[base + i]);
            cmin = vmin(cmin, c);
            cmax = vmax(cmax, c);
        }
        Vec3 d = cmax - cmin;
        int axis = (d.x > d.y && d.x > d.z) ? 0 : ((d.y > d.x && d.y > d.z) ? 1 : 2);
        float splitVal = 0.5f * (comp(cmin, axis) + comp(cmax, axis));
        int start = 0, end = n;
        while (start < end) {
            while (start < end && comp(center(sorted[base + start]), axis) < splitVal) ++start;
            while (start < end && comp(center(sorted[base + end - 1]), axis) >= splitVal) --end;
            if (start < end) {
                uint8_t t = sorted[base + start];
                sorted[base + start] = sorted[base + end - 1];
                sorted[base + end - 1] = t;
                ++start; --end;
            }
        }
        return (start > 0 && start < n) ? start : n / 2;
    };

    while (stackSize > 0) {
        Entry &e = stack[stackSize - 1];
        int nodeId;
        if (e.numObjs <= 4) {
            nodeId = curNode++;
            if (nodeId >= kMaxNodes) { atomicAdd(S.overflow, 1u); return; }
            float *nf = nodeF(blob, nodeId);
            int32_t *ni = nodeI(blob, nodeId);
            ni[kNodeParent] = e.parentID;
            for (int i = 0; i < 4; i++) {
                if (i < e.numObjs) {
                    int leaf = sorted[e.offset + i];
                    const float *L = leafF(blob, leaf);
                    blob[kBvhLeafOff + leaf * kLeafWords + kLeafParent] = (uint32_t(nodeId) << 2) | uint32_t(i);
                    ni[kNodeChild + i] = int32_t(0x80000000u | uint32_t(leaf));
                    for (int k = 0; k < 6; k++) nf[k * 4 + i] = L[k];
                } else {
                    ni[kNodeChild + i] = kSentinel;
                    for (int k = 0; k < 3; k++) { nf[k * 4 + i] = FLT_MAX; nf[(3 + k) * 4 + i] = -FLT_MAX; }
                }
            }
        } else if (e.nodeID == kSentinel) {
            nodeId = curNode++;
            if (nodeId >= kMaxNodes) { atomicAdd(S.overflow, 1u); return; }
            e.nodeID = nodeId;
            int32_t *ni = nodeI(blob, nodeId);
            for (int i = 0; i < 4; i++) ni[kNodeChild + i] = kSentinel;
            ni[kNodeParent] = e.parentID;

            int second = midpointSplit(e.offset, e.numObjs);
            int h1 = second, h2 = e.numObjs - second;
            int first = midpointSplit(e.offset, h1);
            int thirdS = midpointSplit(e.offset + second, h2);
            int off = e.offset, self = e.nodeID;
            stack[stackSize++] = {-1, self, off + h1 + thirdS, h2 - thirdS};
            stack[stackSize++] = {-1, self, off + h1, thirdS};
            stack[stackSize++] = {-1, self, off + first, h1 - first};
            stack[stackSize++] = {-1, self, off, first};
            continue;
        } else {
            nodeId = e.nodeID;
        }

        stackSize -= 1;
        const int32_t *ni = nodeI(blob, nodeId);
        int parent = ni[kNodeParent];
        if (parent == -1) continue;

        const float *nf = nodeF(blob, nodeId);
        float c[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = 0; i < 4; i++) {
            if (ni[kNodeChild + i] == kSentinel) break;
            for (int k = 0; k < 3; k++) {
                c[k] = fminf(c[k], nf[k * 4 + i]);
                c[3 + k] = fmaxf(c[3 + k], nf[(3 + k) * 4 + i]);
            }
        }
        int32_t *pi = nodeI(blob, parent);
        float *pf = nodeF(blob, parent);
        int off = 0;
        while (pi[kNodeChild + off] != kSentinel) off++;
        pi[kNodeChild + off] = nodeId;
        for (int k = 0; k < 6; k++) pf[k * 4 + off] = c[k];
    }

    // Leaves in the order BVH::findIntersecting meets them (LIFO stack, children 0..3, leaves
    // reported as soon as their node is popped — M/include/madrona/broadphase.inl:22-59).  A pruned
    // walk reports a subset of the leaves in this same relative order, which is all k_broadphase
    // needs to emit candidate pairs in the reference's order without walking the tree.
    {
        uint8_t *dfs = S.dfsLeaf + size_t(w) * kDfsStride;
        int32_t st[32];
        st[0] = 0;
        int sp = 1, n = 0;
        while (sp > 0) {
            const int32_t *ni = nodeI(blob, st[--sp]);
            for (int i = 0; i < 4; i++) {
                int32_t child = ni[kNodeChild + i];
                if (child == kSentinel) continue;
                if (child < 0) { if (n < kMaxBodies) dfs[n++] = uint8_t(child & 0x7FFFFFFF); }
                else if (sp < 32) st[sp++] = child;
            }
        }
    }
}

// One thread per world; worlds without a pending rebuild return at once.
__global__ void k_bvh_rebuild(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    rebuildWorldBvh(S, w);
}

}  // namespace hs
