// Device state layout for one shard of worlds (one hs_sim == one GPU).
//
// * Rigid-body state is structure-of-arrays ACROSS worlds, world index fastest:
//   field f of body slot s of world w lives at  body[(f*HS_MAX_BODIES + s)*N + w],
//   so thread-per-(slot,world) and thread-per-world kernels read coalesced lines.
// * Slot order inside a world equals the reference's ECS iteration order
//   (SURVEY.md App. B-1,2): DynamicObject rows in creation order, then DynAgent
//   rows in creation order.  Candidate / contact order follows from it.
// * The per-world BVH (nodes + leaf records) is a contiguous world-major blob so
//   that ray / overlap kernels can stage one world's working set into shared
//   memory with a single bulk copy.
// * Exported tensors keep the reference's row-major AoS layouts (src/sim.hpp:149-243)
//   because they are the ABI.
#pragma once

#include <cstdint>
#include "hs_math.cuh"

namespace hs {

constexpr int kMaxBodies = 52;     // src/sim.cpp:1356-1357 (9+2+6+35)
constexpr int kMaxAgents = 6;      // src/sim.hpp:41
constexpr int kMaxBoxes = 9;
constexpr int kMaxRamps = 2;
constexpr int kMaxWalls = 33;      // src/geo_gen.cpp:145,430-433
constexpr int kMaxNodes = 69;      // BVH::numInternalNodes(52), broadphase.cpp:38-45
constexpr int kMaxCand = 128;     // candidate pairs per world; largest seen over every tested configuration: 86 (overflow is counted)
constexpr int kIdPool = 64;        // ids_per_cache_ (id_map.hpp)
constexpr int kCheckpointBytes = 1392;   // sizeof(Checkpoint), src/sim.hpp:283-313

constexpr float kDeltaT = 1.f / 30.f;       // src/sim.cpp:14
constexpr int kSubsteps = 4;
constexpr int kNumPrepSteps = 96;
constexpr int kEpisodeLen = 240;

// SimObject (src/sim.hpp:78-88)
enum : int { OBJ_SPHERE = 0, OBJ_PLANE, OBJ_CUBE, OBJ_WALL, OBJ_HIDER, OBJ_SEEKER, OBJ_RAMP, OBJ_BOX, OBJ_COUNT };
// ResponseType (M/include/madrona/physics.hpp:27-31)
enum : int { RESP_DYNAMIC = 0, RESP_KINEMATIC = 1, RESP_STATIC = 2 };
// OwnerTeam (src/sim.hpp:127-132)
enum : int { OWNER_NONE = 0, OWNER_SEEKER, OWNER_HIDER, OWNER_UNOWNABLE };
// AgentType (src/sim.hpp:138-141)
enum : int { AGENT_SEEKER = 0, AGENT_HIDER = 1 };
// SimFlags (src/sim_flags.hpp:7-13)
enum : uint32_t { FLAG_FIXED_WORLD = 1, FLAG_IGNORE_EPISODE_LEN = 2, FLAG_RANDOM_FLIP_TEAMS = 4, FLAG_ZERO_AGENT_VEL = 8 };

// ---- per-body float fields -------------------------------------------------
enum BodyField : int {
    BF_PX, BF_PY, BF_PZ,
    BF_QW, BF_QX, BF_QY, BF_QZ,
    BF_SX, BF_SY, BF_SZ,
    BF_VX, BF_VY, BF_VZ,
    BF_WX, BF_WY, BF_WZ,
    BF_FX, BF_FY, BF_FZ,          // ExternalForce
    BF_TX, BF_TY, BF_TZ,          // ExternalTorque
    // XPBD per-substep scratch (M/src/physics/xpbd.cpp:25-38)
    BF_PREV_PX, BF_PREV_PY, BF_PREV_PZ,
    BF_PREV_QW, BF_PREV_QX, BF_PREV_QY, BF_PREV_QZ,
    BF_PRE_PX, BF_PRE_PY, BF_PRE_PZ,
    BF_PRE_QW, BF_PRE_QX, BF_PRE_QY, BF_PRE_QZ,
    BF_PRE_VX, BF_PRE_VY, BF_PRE_VZ,
    BF_PRE_WX, BF_PRE_WY, BF_PRE_WZ,
    BF_COUNT
};

// ---- body records -------------------------------------------------------------------------------
// In HBM a body is seven 32-byte records, world-major, in two arrays:
//   body [world][slot][4][8 floats]   what the XPBD solver reads at the start of a substep:
//       R0 pose (pos3, quat4)   R1 previous pose (substep start)   R2 pre-solve pose
//       R3 pre-solve velocity
//   bodyB[world][slot][3][8 floats]   R4 velocity (lin3, ang3)   R5 scale3   R6 force3, torque3
// One record = one 32-byte sector = two 128-bit accesses, so the pair-centric stages
// (narrowphase, rays) move whole sectors of useful bytes per body.  The solver records of
// consecutive slots are contiguous (128 B per slot, no holes), so the solver stages the movable
// tail of a world — boxes, ramps, ground plane, agents: the slots after the walls — into shared
// memory with ONE bulk-async copy (cp.async.bulk, hs_solver.cuh).  BodyField indices address
// components through recOf / compOf, so scalar code is layout-agnostic.
constexpr int kRecWords = 8, kBodyRecs = 7, kSolverRecs = 4, kAuxRecs = kBodyRecs - kSolverRecs;
constexpr int kBodyWords = kRecWords * kSolverRecs, kBodyAuxWords = kRecWords * kAuxRecs;
enum BodyRec : int { REC_POSE = 0, REC_PREV, REC_PRE, REC_PREVEL, REC_VEL, REC_SCALE, REC_FORCE };
HS_HD constexpr int recOf(int f) { return f < BF_SX ? REC_POSE : f < BF_VX ? REC_SCALE : f < BF_FX ? REC_VEL : f < BF_PREV_PX ? REC_FORCE : f < BF_PRE_PX ? REC_PREV : f < BF_PRE_VX ? REC_PRE : REC_PREVEL; }
HS_HD constexpr int compOf(int f) { return f < BF_SX ? f : f < BF_VX ? f - BF_SX : f < BF_FX ? f - BF_VX : f < BF_PREV_PX ? f - BF_FX : f < BF_PRE_PX ? f - BF_PREV_PX : f < BF_PRE_VX ? f - BF_PRE_PX : f - BF_PRE_VX; }

// body meta word: obj | resp<<4 | owner<<6 | entity_id<<8 | leaf<<16
HS_HD uint32_t packMeta(int obj, int resp, int owner, int eid, int leaf)
{
    return uint32_t(obj) | (uint32_t(resp) << 4) | (uint32_t(owner) << 6) |
           (uint32_t(eid) << 8) | (uint32_t(leaf) << 16);
}
HS_HD int metaObj(uint32_t m) { return m & 0xF; }
HS_HD int metaResp(uint32_t m) { return (m >> 4) & 3; }
HS_HD int metaOwner(uint32_t m) { return (m >> 6) & 3; }
HS_HD int metaEid(uint32_t m) { return (m >> 8) & 0xFF; }
HS_HD int metaLeaf(uint32_t m) { return (m >> 16) & 0xFF; }

// ---- per-world int fields (SoA [field][world]) ------------------------------
enum WorldField : int {
    WF_NUM_BODIES, WF_NUM_DYN,          // total bodies; DynamicObject rows (slots [0,numDyn))
    WF_NUM_AGENTS, WF_NUM_HIDERS, WF_NUM_SEEKERS,
    WF_NUM_BOXES, WF_NUM_RAMPS,
    WF_EPISODE_STEP, WF_WORLD_EPISODE,
    WF_EP_KEY_A, WF_EP_KEY_B,           // curEpisodeRNDCounter (exported as Seed)
    WF_RNG_A, WF_RNG_B, WF_RNG_COUNT,   // episode RNG (src/sim.hpp:334)
    WF_HIDER_REWARD,                    // float bits of Sim::hiderTeamReward
    WF_SEEKERS_FIRST,
    WF_SCORE0, WF_SCORE1,               // EpisodeStats::runningScores
    WF_ID_STACK_N, WF_ID_FRESH,         // entity-id cache emulation
    WF_JOINT_SEQ,                       // joint creation counter
    WF_NUM_CAND,
    WF_DID_RESET,                       // set by the reset stage for this step
    WF_NUM_CONTACTS,                    // contacts of the last substep (statistics)
    WF_STAGE_FIRST, WF_STAGE_COUNT,     // solver staging range [first, first + count) of this step; count < 0: too large
    WF_COUNT
};

// per-world small byte tables (SoA [index][world])
enum WorldByte : int {
    WB_BOX_SLOT = 0,                         // 9
    WB_RAMP_SLOT = WB_BOX_SLOT + kMaxBoxes,  // 2
    WB_HIDER_SLOT = WB_RAMP_SLOT + kMaxRamps,   // 3
    WB_SEEKER_SLOT = WB_HIDER_SLOT + 3,         // 3
    WB_IFACE_SLOT = WB_SEEKER_SLOT + 3,         // 6 : body slot of interface i, 0xFF = none
    WB_IFACE_TYPE = WB_IFACE_SLOT + kMaxAgents, // 6
    WB_ID_STACK = WB_IFACE_TYPE + kMaxAgents,   // 64
    WB_COUNT = WB_ID_STACK + kIdPool
};

// ---- grab joints: one optional Fixed joint per agent k (DynAgent row k) -------
enum JointField : int {
    JF_A1W, JF_A1X, JF_A1Y, JF_A1Z,
    JF_A2W, JF_A2X, JF_A2Y, JF_A2Z,
    JF_SEP,
    JF_R1X, JF_R1Y, JF_R1Z,
    JF_R2X, JF_R2Y, JF_R2Z,
    JF_COUNT
};
// jmeta[k][w]: seq<<16 | eid<<8 | other_slot ; 0xFFFFFFFF = no joint

// ---- BVH blob (world-major, 32-bit words) -------------------------------------
constexpr int kNodeWords = 32;       // bounds[24] children[4] parent pad[3]
constexpr int kLeafWords = 20;       // aabb[6] parent slot|obj<<8 pos[3] rot[4] scale[3] pad[2]
constexpr int kBvhHdrWords = 4;      // num_nodes, num_leaves, force_rebuild, pad
constexpr int kBvhNodeOff = kBvhHdrWords;
constexpr int kBvhLeafOff = kBvhNodeOff + kMaxNodes * kNodeWords;
constexpr int kBvhWords = kBvhLeafOff + kMaxBodies * kLeafWords;   // 3252 words = 13008 B
constexpr int kNodeChild = 24, kNodeParent = 28;
constexpr int kLeafParent = 6, kLeafInfo = 7, kLeafPos = 8, kLeafRot = 11, kLeafScale = 15;
constexpr int32_t kSentinel = -1;    // 0xFFFFFFFF
constexpr int kDfsStride = 64;       // bytes per world of the leaf visiting-order table

// ---- contacts (world-major) ----------------------------------------------------
constexpr int kContactWords = 20;    // points[4]{x,y,z,depth} normal[3] lambdaN
// cmeta[w][c]: ref | alt<<8 | npts<<16 ; npts==0 -> no contact for candidate c

// ---- hull / object tables (constant memory) ---------------------------------------
constexpr int kMaxHullVerts = 8, kMaxHullFaces = 6, kMaxHullHedges = 24;
struct HullData {
    int numVerts, numFaces, numHedges, pad;
    float verts[kMaxHullVerts][3];
    float planes[kMaxHullFaces][4];
    uint8_t hedgeNext[kMaxHullHedges];
    uint8_t hedgeRoot[kMaxHullHedges];
    uint8_t hedgeFace[kMaxHullHedges];
    uint8_t faceBase[kMaxHullFaces];
    uint8_t pad2[2];
};
struct ObjectData {
    int primType;            // 1 sphere, 2 hull, 4 plane (CollisionPrimitive::Type)
    float aabb[6];           // object == primitive AABB (one primitive per object)
    float invMass;
    float invInertia[3];
    float muS, muD;
};
struct AssetTable {
    ObjectData obj[OBJ_COUNT];
    HullData hull[OBJ_COUNT];    // indexed by object id (agents share geometry, duplicated)
};

// what the solver needs to know about one candidate pair (written once per step by k_broadphase,
// hs_trace.cuh; the solver adds the contact's dependency level and point count per substep)
//   x: candidate index | level << 8 | slot a << 16 | slot b << 24
//   y: obj a | obj b << 4 | carrying a << 8 | carrying b << 9 | static a << 10 | static b << 11
struct alignas(8) SchedEntry { uint32_t x, y; };

constexpr int kSolveOrderBufs = 4, kSolveOrderBuckets = 8;     // DeviceState::solveOrder / solveCnt

struct SimParams {
    int numWorlds;           // worlds this launch covers (the whole shard, or one slice of it)
    int stride;              // worlds on the shard = SoA row stride of every per-world array
    int worldBase;           // global index of world 0 of this shard (RNG keying)
    int agentsPerWorld;      // A = maxHiders + maxSeekers
    uint32_t simFlags;
    uint32_t initKeyA, initKeyB;     // rand::initKey(seed)
    int minHiders, maxHiders, minSeekers, maxSeekers;
    uint32_t debugFlags;             // hs_set_debug: bit 0 = record lidar hit bodies
};

struct DeviceState {
    SimParams p;
    float *body;             // [N][kMaxBodies][kSolverRecs][8]  pose, previous / pre-solve pose, pre-solve velocity
    float *bodyB;            // [N][kMaxBodies][kAuxRecs][8]     velocity; scale; force + torque
    uint32_t *meta;          // [N][kMaxBodies]
    int32_t *wi;             // [WF_COUNT][N]
    uint8_t *wb;             // [WB_COUNT][N]
    float *joint;            // [JF_COUNT][kMaxAgents][N]
    uint32_t *jmeta;         // [kMaxAgents][N]
    uint32_t *bvh;           // [N][kBvhWords]
    uint8_t *dfsLeaf;        // [N][kDfsStride] leaf ids in BVH visiting order (set by the rebuild)
    uint32_t *resetList;     // [N] worlds to regenerate this step: world | level << 24 (compacted)
    int *resetCount;         // [N] element 0 of a slice's view counts that slice's list
    // k_solve's world order (a performance hint, any permutation is valid): 4 rotating buffers of
    // 8 buckets keyed by the rounds a world's contact sweep took in the previous substep
    uint32_t *solveOrder;    // [4 * 8][N] worlds of a bucket (slice-relative), in arrival order
    int *solveCnt;           // [N + 64] elements 0..31 of a slice's view: worlds per (buffer, bucket)
    uint16_t *cand;          // [N][kMaxCand]  a | b<<8
    float *contact;          // [N][kMaxCand][kContactWords]
    uint32_t *cmeta;         // [N][kMaxCand]
    SchedEntry *sched;       // [N][kMaxCand] per-pair solver info of this step's candidates (hs_solver.cuh)
    uint32_t *overflow;      // [1] candidate-capacity overflow counter
    int8_t *dbgLidarHit;     // [N*A][30] body slot every lidar ray hit (debugFlags bit 0), -1 miss

    // exported tensors (reference ExportID order, src/sim.hpp:45-68)
    int32_t *exReset;        // [N,1]
    int32_t *exPrepCounter;  // [N*A,1]
    int32_t *exAction;       // [N*A,5]
    float *exSelfObs;        // [N*A,13]
    int32_t *exSelfType;     // [N*A,1]
    float *exSelfMask;       // [N*A,1]
    float *exAgentObs;       // [N*A,5,14]
    float *exBoxObs;         // [N*A,9,17]
    float *exRampObs;        // [N*A,2,14]
    float *exAgentVis;       // [N*A,5]
    float *exBoxVis;         // [N*A,9]
    float *exRampVis;        // [N*A,2]
    float *exLidar;          // [N*A,30]
    int32_t *exSeed;         // [N*A,2]
    float *exReward;         // [N*A,1]
    int32_t *exDone;         // [N*A,1]
    float *exGlobalPos;      // [N,17,2]
    int32_t *exPolicy;       // [N*A,1]
    float *exEpisodeResult;  // [N,2]
    int32_t *exCkptCtrl;     // [N,1]
    uint8_t *exCkpt;         // [N,sizeof(Checkpoint)]
};

// A slice [begin, begin + count) of the shard as its own DeviceState: SoA arrays keep the shard
// stride and are offset by `begin` worlds, world-major arrays are offset by whole rows.  Worlds
// never interact, so a step can run as several independent slices (hs_sim.cu: graph branches).
inline DeviceState sliceOf(const DeviceState &S, int begin, int count)
{
    DeviceState V = S;
    const size_t b = size_t(begin), A = size_t(S.p.agentsPerWorld);
    V.p.numWorlds = count;
    V.p.worldBase = S.p.worldBase + begin;
    V.body += b * kMaxBodies * kBodyWords; V.bodyB += b * kMaxBodies * kBodyAuxWords; V.meta += b * kMaxBodies; V.wi += b; V.wb += b; V.joint += b; V.jmeta += b;
    V.bvh += b * kBvhWords;
    V.dfsLeaf += b * kDfsStride;
    V.resetList += b; V.resetCount += b;
    V.solveOrder += b; V.solveCnt += b;
    V.cand += b * kMaxCand;
    V.contact += b * kMaxCand * kContactWords;
    V.cmeta += b * kMaxCand;
    V.sched += b * kMaxCand;
    V.dbgLidarHit += b * A * 30;
    V.exReset += b; V.exPrepCounter += b * A; V.exAction += b * A * 5; V.exSelfObs += b * A * 13;
    V.exSelfType += b * A; V.exSelfMask += b * A; V.exAgentObs += b * A * 5 * 14;
    V.exBoxObs += b * A * 9 * 17; V.exRampObs += b * A * 2 * 14; V.exAgentVis += b * A * 5;
    V.exBoxVis += b * A * 9; V.exRampVis += b * A * 2; V.exLidar += b * A * 30; V.exSeed += b * A * 2;
    V.exReward += b * A; V.exDone += b * A; V.exGlobalPos += b * 17 * 2; V.exPolicy += b * A;
    V.exEpisodeResult += b * 2; V.exCkptCtrl += b; V.exCkpt += b * kCheckpointBytes;
    return V;
}

#if defined(__CUDACC__)
__device__ __forceinline__ float *recBase(const DeviceState &S, int rec, int s, int w)
{
    return rec < kSolverRecs ? S.body + ((size_t(w) * kMaxBodies + s) * kSolverRecs + rec) * kRecWords
                             : S.bodyB + ((size_t(w) * kMaxBodies + s) * kAuxRecs + (rec - kSolverRecs)) * kRecWords;
}
__device__ __forceinline__ float &BF(const DeviceState &S, int f, int s, int w)
{
    return recBase(S, recOf(f), s, w)[compOf(f)];
}
__device__ __forceinline__ uint32_t &META(const DeviceState &S, int s, int w)
{
    return S.meta[size_t(w) * kMaxBodies + s];
}
__device__ __forceinline__ int32_t &WI(const DeviceState &S, int f, int w)
{
    return S.wi[size_t(f) * S.p.stride + w];
}
__device__ __forceinline__ uint8_t &WB(const DeviceState &S, int f, int w)
{
    return S.wb[size_t(f) * S.p.stride + w];
}
__device__ __forceinline__ float &JF(const DeviceState &S, int f, int k, int w)
{
    return S.joint[(size_t(f) * kMaxAgents + k) * S.p.stride + w];
}
__device__ __forceinline__ uint32_t &JMETA(const DeviceState &S, int k, int w)
{
    return S.jmeta[size_t(k) * S.p.stride + w];
}
__device__ __forceinline__ uint32_t *BVH(const DeviceState &S, int w)
{
    return S.bvh + size_t(w) * kBvhWords;
}

// ---- record-granular (128-bit) access --------------------------------------------------------------
__device__ __forceinline__ float4 *recPtr(const DeviceState &S, int rec, int s, int w)
{
    return reinterpret_cast<float4 *>(recBase(S, rec, s, w));
}
struct Pose { Vec3 x; Quat q; };
struct Vel6 { Vec3 v; Vec3 w; };
// pose-shaped records: REC_POSE, REC_PREV, REC_PRE
__device__ __forceinline__ Pose loadPose(const DeviceState &S, int rec, int s, int w)
{
    const float4 *p = recPtr(S, rec, s, w);
    float4 a = p[0], b = p[1];
    return {{a.x, a.y, a.z}, {a.w, b.x, b.y, b.z}};
}
__device__ __forceinline__ void storePose(const DeviceState &S, int rec, int s, int w, Vec3 x, Quat q)
{
    float4 *p = recPtr(S, rec, s, w);
    p[0] = make_float4(x.x, x.y, x.z, q.w);
    p[1] = make_float4(q.x, q.y, q.z, 0.f);
}
// six-vector records: REC_VEL (linear, angular), REC_FORCE (force, torque), REC_PREVEL
__device__ __forceinline__ Vel6 loadVel6(const DeviceState &S, int rec, int s, int w)
{
    const float4 *p = recPtr(S, rec, s, w);
    float4 a = p[0], b = p[1];
    return {{a.x, a.y, a.z}, {a.w, b.x, b.y}};
}
__device__ __forceinline__ void storeVel6(const DeviceState &S, int rec, int s, int w, Vec3 v, Vec3 o)
{
    float4 *p = recPtr(S, rec, s, w);
    p[0] = make_float4(v.x, v.y, v.z, o.x);
    p[1] = make_float4(o.y, o.z, 0.f, 0.f);
}

__device__ __forceinline__ Vec3 loadPos(const DeviceState &S, int s, int w)
{
    float4 a = recPtr(S, REC_POSE, s, w)[0];
    return {a.x, a.y, a.z};
}
__device__ __forceinline__ Quat loadRot(const DeviceState &S, int s, int w)
{
    return loadPose(S, REC_POSE, s, w).q;
}
__device__ __forceinline__ Vec3 loadScale(const DeviceState &S, int s, int w)
{
    float4 a = recPtr(S, REC_SCALE, s, w)[0];
    return {a.x, a.y, a.z};
}
__device__ __forceinline__ Vec3 loadV3(const DeviceState &S, int f0, int s, int w)
{
    return {BF(S, f0, s, w), BF(S, f0 + 1, s, w), BF(S, f0 + 2, s, w)};
}
__device__ __forceinline__ Quat loadQ(const DeviceState &S, int f0, int s, int w)
{
    return {BF(S, f0, s, w), BF(S, f0 + 1, s, w), BF(S, f0 + 2, s, w), BF(S, f0 + 3, s, w)};
}
__device__ __forceinline__ void storeV3(const DeviceState &S, int f0, int s, int w, Vec3 v)
{
    BF(S, f0, s, w) = v.x; BF(S, f0 + 1, s, w) = v.y; BF(S, f0 + 2, s, w) = v.z;
}
__device__ __forceinline__ void storeQ(const DeviceState &S, int f0, int s, int w, Quat q)
{
    BF(S, f0, s, w) = q.w; BF(S, f0 + 1, s, w) = q.x; BF(S, f0 + 2, s, w) = q.y; BF(S, f0 + 3, s, w) = q.z;
}
#endif

}  // namespace hs
