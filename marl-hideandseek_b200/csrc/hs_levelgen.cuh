// Episode reset: episode bookkeeping, entity-id recycling, procedural walls,
// rejection-sampled placement of boxes / ramps / agents.
//
// Behavioural contract (bit-exact vs the reference CPU backend):
//   resetSystem / resetEnvironment / initEpisodeRNG   src/sim.cpp:105-200
//   makeWalls / applyWallOperation / findAnotherWall / addDoor / populateStaticGeometry
//                                                     src/geo_gen.cpp:139-505
//   generateTrainingEnvironment / makeAgent / generateEnvironment
//                                                     src/level_gen.cpp:12-334
//   entity-id reuse order                             M/include/madrona/impl/id_map_impl.inl:69-225
//
// One thread per world: the generator is a serial, branchy, RNG-driven program;
// in the benchmark it runs on 1 step in 240.  All per-world scratch lives in
// registers / local memory and the finished bodies are scattered once into the
// SoA state.
#pragma once

#include "hs_state.cuh"
#include "hs_rng.cuh"
#include <hideseek/detmath.h>

namespace hs {

__constant__ AssetTable c_assets;

__device__ __forceinline__ AABB objAABB(int obj)
{
    const float *a = c_assets.obj[obj].aabb;
    return {{a[0], a[1], a[2]}, {a[3], a[4], a[5]}};
}

// Quat::angleAxis (math.inl:728-739) with the deterministic sin/cos.
__device__ __forceinline__ Quat angleAxisZ(float angle)
{
    float s, c;
    hs_det_sincosf(angle / 2.f, &s, &c);
    return {c, 0.f * s, 0.f * s, 1.f * s};
}

__device__ __forceinline__ Quat angleAxis(float angle, Vec3 n)
{
    float s, c;
    hs_det_sincosf(angle / 2.f, &s, &c);
    return {c, n.x * s, n.y * s, n.z * s};
}

struct WallSet {
    static constexpr int kCap = 36;
    float p1x[kCap], p1y[kCap], p2x[kCap], p2y[kCap];
    uint8_t horiz[kCap], vert[kCap];
    int n, nH, nV;
};

__device__ __forceinline__ bool wallIsHorizontal(float p1y, float p2y)
{
    return fabsf(p1y - p2y) < 0.000001f;
}

__device__ __forceinline__ void wallSort(float &p1x, float &p1y, float &p2x, float &p2y)
{
    if (p1x > p2x || p1y > p2y) {
        float tx = p1x, ty = p1y;
        p1x = p2x; p1y = p2y; p2x = tx; p2y = ty;
    }
}

__device__ inline int wallAdd(WallSet &W, float ax, float ay, float bx, float by)
{
    wallSort(ax, ay, bx, by);
    int idx = W.n;
    if (idx >= WallSet::kCap) return idx - 1;   // reference asserts here (geo_gen.cpp:24)
    if (wallIsHorizontal(ay, by)) W.horiz[W.nH++] = (uint8_t)idx;
    else W.vert[W.nV++] = (uint8_t)idx;
    W.p1x[idx] = ax; W.p1y[idx] = ay; W.p2x[idx] = bx; W.p2y[idx] = by;
    W.n = idx + 1;
    return idx;
}

__device__ __forceinline__ float wallLength(const WallSet &W, int i)
{
    return wallIsHorizontal(W.p1y[i], W.p2y[i]) ? (W.p2x[i] - W.p1x[i]) : (W.p2y[i] - W.p1y[i]);
}

// geo_gen.cpp:177-270.  `axis` selects which coordinate runs along the walls
// of the list (horizontal list: along = x, across = y).
__device__ inline int findPartnerWall(const WallSet &W, const uint8_t *list, int size,
                                      int chosenIdx, bool horizontal, Rng &rng)
{
    const float *a1 = horizontal ? W.p1x : W.p1y;   // along-axis start
    const float *a2 = horizontal ? W.p2x : W.p2y;   // along-axis end
    const float *c1 = horizontal ? W.p1y : W.p1x;   // across-axis coordinate
    const float min_len = horizontal ? 0.3f : 0.5f;

    int chosen = list[chosenIdx];
    int start = chosenIdx + 1 + rng.sampleI32(0, size - 1);
    for (int i = 0; i < size - 1; ++i) {
        int cur = (start + i) % size;
        if (cur == chosenIdx) cur = (cur + 1) % size;
        int other = list[cur];
        if (!(a1[chosen] >= a2[other] || a2[chosen] <= a1[other]) &&
            wallLength(W, chosen) >= min_len && wallLength(W, other) >= min_len) {
            float high = a2[other] < a2[chosen] ? a2[other] : a2[chosen];
            float low = a1[chosen] < a1[other] ? a1[other] : a1[chosen];
            bool works = true;
            for (int j = 0; j < size; ++j) {
                if (j == cur) continue;
                int wj = list[j];
                float lo_m = low - 0.1f, hi_p = high + 0.1f;
                float in_low = a1[wj] < lo_m ? lo_m : a1[wj];
                float in_high = hi_p < a2[wj] ? hi_p : a2[wj];
                if (in_low < in_high) {
                    float c = c1[wj];
                    float cmin = c1[other] < c1[chosen] ? c1[other] : c1[chosen];
                    float cmax = c1[chosen] < c1[other] ? c1[other] : c1[chosen];
                    if (c > cmin && c < cmax) { works = false; break; }
                }
            }
            if (works) return cur;
        }
    }
    return -1;
}

// geo_gen.cpp:275-307
__device__ inline void wallAddDoor(WallSet &W, int wi, float doorSize, Rng &rng)
{
    if (wallIsHorizontal(W.p1y[wi], W.p2y[wi])) {
        float low = W.p1x[wi] + doorSize;
        float high = W.p2x[wi] - doorSize;
        float rat = 0.3f + rng.sampleUniform() * 0.4f;
        float x = low + rat * (high - low);
        float oldP2x = W.p2x[wi];
        W.p2x[wi] = x - doorSize * 0.5f;
        wallSort(W.p1x[wi], W.p1y[wi], W.p2x[wi], W.p2y[wi]);
        wallAdd(W, x + doorSize * 0.5f, W.p1y[wi], oldP2x, W.p1y[wi]);
    } else {
        float low = W.p1y[wi] + doorSize;
        float high = W.p2y[wi] - doorSize;
        float rat = 0.3f + rng.sampleUniform() * 0.4f;
        float y = low + rat * (high - low);
        float oldP2y = W.p2y[wi];
        W.p2y[wi] = y - doorSize * 0.5f;
        wallSort(W.p1x[wi], W.p1y[wi], W.p2x[wi], W.p2y[wi]);
        wallAdd(W, W.p1x[wi], y + doorSize * 0.5f, W.p1x[wi], oldP2y);
    }
}

// geo_gen.cpp:309-427, case WallConnectAndAddDoor
__device__ inline void wallConnect(WallSet &W, Rng &rng)
{
    bool isH = rng.sampleI32(0, 2) != 0;
    const uint8_t *list = isH ? W.horiz : W.vert;
    int size = isH ? W.nH : W.nV;
    int wallIdx = rng.sampleI32(0, size);
    int otherIdx;
    int counter = 0;
    while ((otherIdx = findPartnerWall(W, list, size, wallIdx, isH, rng)) == -1) {
        isH = rng.sampleI32(0, 2) != 0;
        list = isH ? W.horiz : W.vert;
        size = isH ? W.nH : W.nV;
        wallIdx = rng.sampleI32(0, size);
        if (counter++ > 4) return;
    }

    int first = list[wallIdx], second = list[otherIdx];
    if (isH) {
        float high = W.p2x[second] < W.p2x[first] ? W.p2x[second] : W.p2x[first];
        float low = W.p1x[first] < W.p1x[second] ? W.p1x[second] : W.p1x[first];
        if (W.p1y[first] > W.p1y[second]) { int t = first; first = second; second = t; }
        float rat = 0.4f + rng.sampleUniform() * 0.2f;
        float x = low + rat * (high - low);
        int newIdx = wallAdd(W, x, W.p1y[first], x, W.p1y[second]);
        float firstOld = W.p2x[first], secondOld = W.p2x[second];
        W.p2x[first] = x;
        wallSort(W.p1x[first], W.p1y[first], W.p2x[first], W.p2y[first]);
        W.p2x[second] = x;
        wallSort(W.p1x[second], W.p1y[second], W.p2x[second], W.p2y[second]);
        float fy = W.p1y[first], sy = W.p1y[second];
        wallAdd(W, x, fy, firstOld, fy);
        wallAdd(W, x, sy, secondOld, sy);
        wallAddDoor(W, newIdx, 0.1f, rng);
    } else {
        float high = W.p2y[second] < W.p2y[first] ? W.p2y[second] : W.p2y[first];
        float low = W.p1y[first] < W.p1y[second] ? W.p1y[second] : W.p1y[first];
        if (W.p1x[first] > W.p1x[second]) { int t = first; first = second; second = t; }
        float rat = 0.4f + rng.sampleUniform() * 0.2f;
        float y = low + rat * (high - low);
        int newIdx = wallAdd(W, W.p1x[first], y, W.p1x[second], y);
        float firstOld = W.p2y[first], secondOld = W.p2y[second];
        W.p2y[first] = y;
        wallSort(W.p1x[first], W.p1y[first], W.p2x[first], W.p2y[first]);
        W.p2y[second] = y;
        wallSort(W.p1x[second], W.p1y[second], W.p2x[second], W.p2y[second]);
        float fx = W.p1x[first], sx = W.p1x[second];
        wallAdd(W, fx, y, fx, firstOld);
        wallAdd(W, sx, y, sx, secondOld);
        wallAddDoor(W, newIdx, 0.1f, rng);
    }
}

// geo_gen.cpp:429-465
__device__ inline void makeWalls(WallSet &W, Rng &rng)
{
    W.n = W.nH = W.nV = 0;
    wallAdd(W, 0.f, 0.f, 1.f, 0.f);
    wallAdd(W, 0.f, 0.f, 0.f, 1.f);
    wallAdd(W, 0.f, 1.f, 1.f, 1.f);
    wallAdd(W, 1.f, 1.f, 1.f, 0.f);

    int counts[2];
    counts[0] = 1 + rng.sampleI32(0, 6);      // WallConnectAndAddDoor
    counts[1] = 4 + rng.sampleI32(0, 7 - 4);  // WallAddDoor
    int ops[2] = {0, 1};
    int nSel = 2;

    do {
        int opIdx = rng.sampleI32(0, nSel);
        int op = ops[opIdx];
        --counts[op];
        if (counts[op] == 0) {
            --nSel;
            ops[opIdx] = ops[nSel];
        }
        if (op == 0) {
            wallConnect(W, rng);
        } else {
            float doorSize = 0.1f * 2.0f;
            int wi = rng.sampleI32(0, W.n);
            if (wallLength(W, wi) > 3.0f * doorSize) wallAddDoor(W, wi, doorSize, rng);
        }
    } while (counts[0] != 0 || counts[1] != 0);
}

// ---- entity-id cache emulation (id_map_impl.inl acquireID/releaseID) ---------
struct IdPool {
    const DeviceState &S;
    int w;
    int n, fresh;
    __device__ IdPool(const DeviceState &S_, int w_) : S(S_), w(w_)
    {
        n = WI(S, WF_ID_STACK_N, w);
        fresh = WI(S, WF_ID_FRESH, w);
    }
    __device__ int acquire()
    {
        if (n > 0) { --n; return WB(S, WB_ID_STACK + n, w); }
        return fresh++;
    }
    __device__ void release(int id)
    {
        if (n < kIdPool) { WB(S, WB_ID_STACK + n, w) = (uint8_t)id; ++n; }
    }
    __device__ void commit()
    {
        WI(S, WF_ID_STACK_N, w) = n;
        WI(S, WF_ID_FRESH, w) = fresh;
    }
};

struct NewBody {
    Vec3 pos; Quat rot; Vec3 scale;
    int obj, resp, owner;
};

__device__ inline void writeBody(const DeviceState &S, int w, int slot, const NewBody &b, int eid, int leaf)
{
    storeV3(S, BF_PX, slot, w, b.pos);
    storeQ(S, BF_QW, slot, w, b.rot);
    storeV3(S, BF_SX, slot, w, b.scale);
    for (int f = BF_VX; f <= BF_TZ; f++) BF(S, f, slot, w) = 0.f;
    META(S, slot, w) = packMeta(b.obj, b.resp, b.owner, eid, leaf);
    uint32_t *blob = BVH(S, w);
    blob[kBvhLeafOff + leaf * kLeafWords + kLeafInfo] = uint32_t(slot) | (uint32_t(b.obj) << 8);
}

// resetEnvironment (src/sim.cpp:116-159): teardown + id release order.
__device__ inline void teardownWorld(const DeviceState &S, int w, IdPool &ids)
{
    int nDyn = WI(S, WF_NUM_DYN, w);
    int nAgents = WI(S, WF_NUM_AGENTS, w);
    // obstacles[] == DynamicObject rows in creation order
    for (int s = 0; s < nDyn; s++) ids.release(metaEid(META(S, s, w)));
    // hiders then seekers, each preceded by its grab joint entity
    int nh = WI(S, WF_NUM_HIDERS, w), ns = WI(S, WF_NUM_SEEKERS, w);
    for (int t = 0; t < 2; t++) {
        int cnt = t == 0 ? nh : ns;
        int base = t == 0 ? WB_HIDER_SLOT : WB_SEEKER_SLOT;
        for (int i = 0; i < cnt; i++) {
            int slot = WB(S, base + i, w);
            int k = slot - nDyn;
            uint32_t jm = JMETA(S, k, w);
            if (jm != 0xFFFFFFFFu) {
                ids.release((jm >> 8) & 0xFF);
                JMETA(S, k, w) = 0xFFFFFFFFu;
            }
            ids.release(metaEid(META(S, slot, w)));
        }
    }
    (void)nAgents;
    WI(S, WF_NUM_BODIES, w) = 0;
    WI(S, WF_NUM_DYN, w) = 0;
    WI(S, WF_NUM_AGENTS, w) = 0;
    WI(S, WF_NUM_HIDERS, w) = 0;
    WI(S, WF_NUM_SEEKERS, w) = 0;
    WI(S, WF_NUM_BOXES, w) = 0;
    WI(S, WF_NUM_RAMPS, w) = 0;
}

// generateTrainingEnvironment (src/level_gen.cpp:79-308)
__device__ inline void generateTrainingWorld(const DeviceState &S, int w, Rng &episodeRng,
                                             RandKey levelKey, int numHiders, int numSeekers,
                                             IdPool &ids)
{
    Rng rng{levelKey, 0};

    int totalBoxes = rng.sampleI32(3, 10);
    int numElongated = rng.sampleI32(3, totalBoxes);
    int numCubes = totalBoxes - numElongated;

    WallSet W;
    makeWalls(W, rng);

    // world AABBs of everything placed so far that later placements must avoid
    // (walls, boxes, ramps — never agents, level_gen.cpp:104-121 + makeAgent)
    AABB placed[kMaxBodies];
    int numPlaced = 0;
    int slot = 0;      // DynamicObject rows
    int leaf = 0;      // creation order
    const Quat identity = angleAxisZ(0.f);     // == angleAxis(0, {1,0,0})

    // populateStaticGeometry (geo_gen.cpp:467-505): scale to [-18,18]^2
    for (int i = 0; i < W.n; i++) {
        float p1x = -18.f + 36.f * W.p1x[i], p1y = -18.f + 36.f * W.p1y[i];
        float p2x = -18.f + 36.f * W.p2x[i], p2y = -18.f + 36.f * W.p2y[i];
        NewBody b;
        b.pos = {0.5f * (p1x + p2x), 0.5f * (p1y + p2y), 0.f};
        if (wallIsHorizontal(p1y, p2y)) b.scale = {p2x - b.pos.x, 0.2f, 1.0f};
        else b.scale = {0.2f, p2y - b.pos.y, 1.0f};
        b.rot = identity;
        b.obj = OBJ_WALL; b.resp = RESP_STATIC; b.owner = OWNER_UNOWNABLE;
        placed[numPlaced++] = aabbApplyTRS(objAABB(OBJ_WALL), b.pos, b.rot, b.scale);
        writeBody(S, w, slot, b, ids.acquire(), leaf);
        slot++; leaf++;
    }

    // any-overlap is order-free, so the scan has no early exit and no per-component branches: the
    // loads of several boxes are in flight together (this loop is the critical path of a reset)
    auto overlapsPlaced = [&](const AABB &a) {
        bool any = false;
#pragma unroll 4
        for (int i = 0; i < numPlaced; i++) {
            const AABB b = placed[i];
            any |= (a.lo.x < b.hi.x) & (b.lo.x < a.hi.x) & (a.lo.y < b.hi.y) & (b.lo.y < a.hi.y) &
                   (a.lo.z < b.hi.z) & (b.lo.z < a.hi.z);
        }
        return any;
    };

    // shared rejection-sampling loop (level_gen.cpp:123-229, 266-292)
    auto samplePose = [&](int obj, NewBody &b, AABB &box) {
        int rejections = 0;
        while (true) {
            float x = -18.f + rng.sampleUniform() * 36.f;
            float y = -18.f + rng.sampleUniform() * 36.f;
            b.pos = {x, y, 1.0f};
            b.rot = angleAxisZ(rng.sampleUniform() * 3.14159265358979323846f);
            b.scale = {1.f, 1.f, 1.f};
            box = aabbApplyTRS(objAABB(obj), b.pos, b.rot, b.scale);
            if (!overlapsPlaced(box) || rejections == 20) return;
            rejections++;
        }
    };

    int nBoxes = 0;
    for (int i = 0; i < totalBoxes; i++) {
        int obj = i < numElongated ? OBJ_BOX : OBJ_CUBE;
        NewBody b; AABB box;
        samplePose(obj, b, box);
        b.obj = obj; b.resp = RESP_DYNAMIC; b.owner = OWNER_NONE;
        placed[numPlaced++] = box;
        writeBody(S, w, slot, b, ids.acquire(), leaf);
        WB(S, WB_BOX_SLOT + nBoxes, w) = (uint8_t)slot;
        nBoxes++; slot++; leaf++;
    }
    for (int i = 0; i < kMaxRamps; i++) {
        NewBody b; AABB box;
        samplePose(OBJ_RAMP, b, box);
        b.obj = OBJ_RAMP; b.resp = RESP_DYNAMIC; b.owner = OWNER_NONE;
        placed[numPlaced++] = box;
        writeBody(S, w, slot, b, ids.acquire(), leaf);
        WB(S, WB_RAMP_SLOT + i, w) = (uint8_t)slot;
        slot++; leaf++;
    }

    // team order uses the EPISODE rng and is always drawn (level_gen.cpp:232-238)
    bool seekersFirst = episodeRng.sampleI32(0, 2) == 1;
    if ((S.p.simFlags & FLAG_RANDOM_FLIP_TEAMS) == 0) seekersFirst = false;
    WI(S, WF_SEEKERS_FIRST, w) = seekersFirst ? 1 : 0;

    const int numDyn = slot + 1;     // plane row closes the DynamicObject table
    const int A = S.p.agentsPerWorld;
    int nAgents = 0, nh = 0, nsk = 0;
    for (int team = 0; team < 2; team++) {
        bool teamIsSeeker = (team == 0) == seekersFirst;
        int cnt = teamIsSeeker ? numSeekers : numHiders;
        int obj = teamIsSeeker ? OBJ_SEEKER : OBJ_HIDER;
        for (int i = 0; i < cnt; i++) {
            NewBody b; AABB box;
            samplePose(obj, b, box);
            b.obj = obj; b.resp = RESP_DYNAMIC; b.owner = OWNER_UNOWNABLE;
            // makeAgent (level_gen.cpp:12-66)
            int k = nAgents++;
            int aslot = numDyn + k;
            writeBody(S, w, aslot, b, ids.acquire(), leaf);
            leaf++;
            JMETA(S, k, w) = 0xFFFFFFFFu;
            WB(S, WB_IFACE_SLOT + k, w) = (uint8_t)aslot;
            WB(S, WB_IFACE_TYPE + k, w) = teamIsSeeker ? AGENT_SEEKER : AGENT_HIDER;
            if (teamIsSeeker) WB(S, WB_SEEKER_SLOT + nsk++, w) = (uint8_t)aslot;
            else WB(S, WB_HIDER_SLOT + nh++, w) = (uint8_t)aslot;
            size_t row = size_t(w) * A + k;
            S.exSelfType[row] = teamIsSeeker ? AGENT_SEEKER : AGENT_HIDER;
            S.exSelfMask[row] = 1.f;
            S.exSeed[row * 2 + 0] = WI(S, WF_EP_KEY_A, w);
            S.exSeed[row * 2 + 1] = WI(S, WF_EP_KEY_B, w);
            int32_t *act = S.exAction + row * 5;
            act[0] = 2; act[1] = 2; act[2] = 2; act[3] = 0; act[4] = 0;
        }
    }

    // ground plane is created last but sits in the DynamicObject table (level_gen.cpp:294)
    {
        NewBody b;
        b.pos = {0.f, 0.f, 0.f}; b.rot = identity; b.scale = {1.f, 1.f, 1.f};
        b.obj = OBJ_PLANE; b.resp = RESP_STATIC; b.owner = OWNER_UNOWNABLE;
        writeBody(S, w, slot, b, ids.acquire(), leaf);
        slot++; leaf++;
    }

    WI(S, WF_NUM_DYN, w) = numDyn;
    WI(S, WF_NUM_BODIES, w) = numDyn + nAgents;
    WI(S, WF_NUM_AGENTS, w) = nAgents;
    WI(S, WF_NUM_HIDERS, w) = nh;
    WI(S, WF_NUM_SEEKERS, w) = nsk;
    WI(S, WF_NUM_BOXES, w) = nBoxes;
    WI(S, WF_NUM_RAMPS, w) = kMaxRamps;
}

// generateDebugEnvironment (src/level_gen.cpp:336-526): the fixed physics fixtures behind
// triggerReset(world, 2..8).  Every level lists its DynamicObject rows first and its agents
// after them, so slots and BVH leaves are both handed out in creation order.  Boxes / ramps are
// NOT registered with the observation lists (numActiveBoxes / numActiveRamps stay 0).
// Level 5 never records its plane in obstacles[] (level_gen.cpp:396-405), so in the reference
// that plane survives the world's next reset as an orphan row; here it is torn down with the
// rest of the world, i.e. parity for a level-5 world holds until its next reset.
struct DebugWorld {
    const DeviceState &S;
    int w;
    IdPool &ids;
    int slot = 0, leaf = 0, nAgents = 0, nh = 0, nsk = 0;

    __device__ int object(Vec3 pos, Quat rot, int obj, int resp = RESP_DYNAMIC, int owner = OWNER_NONE,
                          Vec3 scale = {1.f, 1.f, 1.f})
    {
        NewBody b{pos, rot, scale, obj, resp, owner};
        writeBody(S, w, slot, b, ids.acquire(), leaf++);
        return slot++;
    }
    __device__ int plane(Vec3 pos, Quat rot) { return object(pos, rot, OBJ_PLANE, RESP_STATIC, OWNER_UNOWNABLE); }
    // makeAgent (level_gen.cpp:12-66); only called once every object row exists
    __device__ void agent(Vec3 pos, Quat rot, bool seeker)
    {
        const int A = S.p.agentsPerWorld;
        if (nAgents >= A) return;
        NewBody b{pos, rot, {1.f, 1.f, 1.f}, seeker ? OBJ_SEEKER : OBJ_HIDER, RESP_DYNAMIC, OWNER_UNOWNABLE};
        int k = nAgents++;
        int aslot = slot + k;
        writeBody(S, w, aslot, b, ids.acquire(), leaf++);
        JMETA(S, k, w) = 0xFFFFFFFFu;
        WB(S, WB_IFACE_SLOT + k, w) = (uint8_t)aslot;
        WB(S, WB_IFACE_TYPE + k, w) = seeker ? AGENT_SEEKER : AGENT_HIDER;
        if (seeker) WB(S, WB_SEEKER_SLOT + nsk++, w) = (uint8_t)aslot;
        else WB(S, WB_HIDER_SLOT + nh++, w) = (uint8_t)aslot;
        size_t row = size_t(w) * A + k;
        S.exSelfType[row] = seeker ? AGENT_SEEKER : AGENT_HIDER;
        S.exSelfMask[row] = 1.f;
        S.exSeed[row * 2 + 0] = WI(S, WF_EP_KEY_A, w);
        S.exSeed[row * 2 + 1] = WI(S, WF_EP_KEY_B, w);
        int32_t *act = S.exAction + row * 5;
        act[0] = 2; act[1] = 2; act[2] = 2; act[3] = 0; act[4] = 0;
    }
};

__device__ inline void generateDebugWorld(const DeviceState &S, int w, int level, IdPool &ids)
{
    constexpr float kPi = 3.14159265358979323846264338327950288f;
    constexpr float kPiD2 = kPi / 2.f;
    constexpr float kDeg = kPi / 180.f;            // toRadians (math.inl:7-11)
    const Vec3 X = {1.f, 0.f, 0.f}, Y = {0.f, 1.f, 0.f}, Z = {0.f, 0.f, 1.f};
    const Quat flat = angleAxis(0.f, X);
    DebugWorld D{S, w, ids};
    switch (level) {
    case 2: {   // cube balanced on a corner; atanf(1/sqrtf(2)) as glibc rounds it (0x3f1d9014)
        Quat r = qnormalize(qnormalize(angleAxis(__int_as_float(0x3f1d9014), Y) * angleAxis(kDeg * 45.f, X)));
        D.object({0.f, 0.f, 5.f}, r, OBJ_CUBE);
        D.plane({0.f, 0.f, 0.f}, flat);
    } break;
    case 3:
        D.object({0.f, 0.f, 5.f}, angleAxis(0.f, Z), OBJ_CUBE);
        D.plane({0.f, 0.f, 0.f}, flat);
        break;
    case 4:
        D.object(v3(0.f, 0.f, 5.f) + v3(0.f, 0.f, 5.f), qnormalize(angleAxis(kDeg * 45.f, Y)), OBJ_BOX);
        D.plane({0.f, 0.f, 0.f}, flat);
        break;
    case 5:
        D.plane({0.f, 0.f, 0.f}, flat);
        D.agent({0.f, 0.f, 1.f}, Quat{1.f, 0.f, 0.f, 0.f}, false);
        break;
    case 6:
        D.plane({0.f, 0.f, 0.f}, flat);
        D.object({0.f, 0.f, 0.f}, flat, OBJ_WALL, RESP_STATIC, OWNER_UNOWNABLE, {10.f, 0.2f, 1.f});
        D.object({0.f, -5.f, 1.f}, flat, OBJ_CUBE);
        D.agent({-15.f, -15.f, 1.5f}, angleAxis(kDeg * -45.f, Z), false);
        D.agent({-15.f, -10.f, 1.5f}, angleAxis(kDeg * 45.f, Z), true);
        break;
    case 7: {
        Quat r = qnormalize(angleAxis(kDeg * 45.f, Y) * angleAxis(kDeg * 40.f, X));
        D.object({0.f, 0.f, 5.f}, r, OBJ_CUBE);
        D.object(v3(0.f, 0.f, 5.f) + v3(0.f, 0.f, 5.f), r, OBJ_CUBE);
        D.plane({0.f, 0.f, 0.f}, flat);
        D.plane({-20.f, 0.f, 0.f}, angleAxis(kPiD2, Y));
        D.plane({20.f, 0.f, 0.f}, angleAxis(-kPiD2, Y));
    } break;
    case 8: {
        Quat r = qnormalize(angleAxis(kDeg * 25.f, Y) * angleAxis(kDeg * 90.f, Z) * angleAxis(kDeg * 45.f, X));
        int ramp = D.object({0.f, 0.f, 10.f}, r, OBJ_RAMP);
        BF(S, BF_VZ, ramp, w) = -30.f;
        D.object({-0.5f, -0.5f, 1.f}, qnormalize(angleAxis(kDeg * -90.f, X) * angleAxis(kPi, Y)), OBJ_RAMP,
                 RESP_STATIC, OWNER_NONE);
        D.plane({0.f, 0.f, 0.f}, flat);
        D.plane({-20.f, 0.f, 0.f}, angleAxis(kPiD2, Y));
        D.plane({20.f, 0.f, 0.f}, angleAxis(-kPiD2, Y));
    } break;
    default: break;     // unknown level ids generate an empty world, as in the reference
    }
    WI(S, WF_NUM_DYN, w) = D.slot;
    WI(S, WF_NUM_BODIES, w) = D.slot + D.nAgents;
    WI(S, WF_NUM_AGENTS, w) = D.nAgents;
    WI(S, WF_NUM_HIDERS, w) = D.nh;
    WI(S, WF_NUM_SEEKERS, w) = D.nsk;
    WI(S, WF_NUM_BOXES, w) = 0;
    WI(S, WF_NUM_RAMPS, w) = 0;
}

// resetEnvironment + generateEnvironment for one world.  `ckpt == nullptr`: the episode
// reset of resetSystem (new episode key, agent counts drawn from the episode RNG).
// `ckpt != nullptr`: the regeneration half of loadCheckpointSystem (src/sim.cpp:966-985):
// the episode key comes from the checkpoint, the episode counter is not advanced, the two
// agent-count draws are burned and the saved counts are used instead.
struct CkptRegen { RandKey key; int numHiders, numSeekers; };

__device__ inline void regenerateWorld(const DeviceState &S, int w, const CkptRegen *ckpt, int level = 1)
{
    IdPool ids(S, w);
    WI(S, WF_EPISODE_STEP, w) = 0;
    uint32_t *blob = BVH(S, w);
    blob[2] = 1;        // PhysicsSystem::reset -> rebuildOnUpdate + clearLeaves
    blob[1] = 0;
    teardownWorld(S, w, ids);

    RandKey epKey;
    if (ckpt) {
        epKey = ckpt->key;
    } else {
        // initEpisodeRNG (src/sim.cpp:105-114)
        uint32_t ep = (uint32_t)WI(S, WF_WORLD_EPISODE, w);
        WI(S, WF_WORLD_EPISODE, w) = int32_t(ep + 1);
        epKey = RandKey{ep, uint32_t(S.p.worldBase + w)};
    }
    WI(S, WF_EP_KEY_A, w) = (int32_t)epKey.a;
    WI(S, WF_EP_KEY_B, w) = (int32_t)epKey.b;
    Rng rng{splitKey(RandKey{S.p.initKeyA, S.p.initKeyB}, epKey.a, epKey.b), 0};

    int numHiders = rng.sampleI32(S.p.minHiders, S.p.maxHiders + 1);
    int numSeekers = rng.sampleI32(S.p.minSeekers, S.p.maxSeekers + 1);
    if (ckpt) { numHiders = ckpt->numHiders; numSeekers = ckpt->numSeekers; }
    RandKey lvlKey = rng.advance();        // levelGenRandKey always draws
    if (S.p.simFlags & FLAG_FIXED_WORLD) lvlKey = {0, 0};

    if (level == 1) generateTrainingWorld(S, w, rng, lvlKey, numHiders, numSeekers, ids);
    else generateDebugWorld(S, w, level, ids);

    // generateEnvironment tail (level_gen.cpp:326-333): park unused interfaces
    int nAgents = WI(S, WF_NUM_AGENTS, w);
    for (int k = nAgents; k < S.p.agentsPerWorld; k++) {
        WB(S, WB_IFACE_SLOT + k, w) = 0xFF;
        S.exSelfMask[size_t(w) * S.p.agentsPerWorld + k] = 0.f;
    }

    blob[1] = (uint32_t)WI(S, WF_NUM_BODIES, w);
    WI(S, WF_RNG_A, w) = (int32_t)rng.k.a;
    WI(S, WF_RNG_B, w) = (int32_t)rng.k.b;
    WI(S, WF_RNG_COUNT, w) = (int32_t)rng.count;
    WI(S, WF_DID_RESET, w) = 1;
    ids.commit();
}

// resetSystem (src/sim.cpp:172-200) in two launches.  k_reset: one thread per world decides
// (explicit trigger or episode end), advances the step counter of the worlds that carry on and
// appends the others to a compact list.  k_reset_worlds: one thread per LISTED world runs the
// serial level-generation program, so when only some worlds reset (training with staggered
// episodes, the reset-heavy configuration) the warps doing the long program are full instead
// of carrying 32 lanes for a few resetting worlds.  Worlds are independent, so the list order
// (atomic, arbitrary) does not matter.
__global__ void k_reset(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;

    int level = S.exReset[w];
    int step = WI(S, WF_EPISODE_STEP, w);
    if ((S.p.simFlags & FLAG_IGNORE_EPISODE_LEN) == 0 && step == kEpisodeLen - 1) level = 1;

    if (level != 0) {
        S.exReset[w] = 0;
        S.resetList[atomicAdd(S.resetCount, 1)] = uint32_t(w) | (uint32_t(level) << 24);
    } else {
        WI(S, WF_EPISODE_STEP, w) = step + 1;
        WI(S, WF_DID_RESET, w) = 0;
    }
    WI(S, WF_HIDER_REWARD, w) = __float_as_int(1.f);
}

// Launched with kResetSpread threads per world of the slice.  Level generation is a long,
// branchy, per-world serial program: 32 worlds in one warp follow 32 different paths and the
// warp executes their union (measured: 4.6 of 32 lanes active, ~7x the single-world
// instruction count).  When few worlds reset, each gets a warp of its own (lane 0 works), which
// removes that serialisation; when most worlds reset the dense mapping issues fewer
// instructions in total and the GPU is full anyway.
constexpr int kResetSpread = 32;
static_assert(kResetSpread >= 4, "the launch covers N * kResetSpread / 4 threads");

__global__ void k_reset_worlds(DeviceState S)
{
    const int count = *S.resetCount;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    int i;
    if (count * 4 <= S.p.numWorlds) {
        if (t % kResetSpread) return;
        i = t / kResetSpread;
    } else {
        i = t;
    }
    if (i >= count) return;
    const uint32_t e = S.resetList[i];
    regenerateWorld(S, int(e & 0xFFFFFFu), nullptr, int(e >> 24));
}

}  // namespace hs
