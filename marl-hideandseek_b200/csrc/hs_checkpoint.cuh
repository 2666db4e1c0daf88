// Checkpoint save / load: the SaveCheckpoints and LoadCheckpoints task graphs.
//
// Behavioural contract:
//   saveCheckpointSystem   src/sim.cpp:1040-1137
//   loadCheckpointSystem   src/sim.cpp:956-1038 (regenerate the level from the saved episode
//                          key, then overwrite dynamic state and recreate grab joints)
//   Checkpoint byte layout src/sim.hpp:283-313 (1392 bytes, checked below)
// One thread per world; both graphs touch only worlds whose CheckpointControl.trigger is set.
#pragma once

#include "hs_levelgen.cuh"

namespace hs {

struct CkptBody { float pos[3]; float rot[4]; float vel[6]; };
struct CkptObject : CkptBody { uint32_t team; uint8_t isLocked; uint8_t pad[3]; };
struct CkptAgent : CkptBody {
    int32_t grabIdx;
    float grabR1[3];
    float grabR2[3];
    float attachRot1[4];
    float attachRot2[4];
    float separation;
};
struct CkptData {
    uint32_t episodeKey[2];
    int32_t runningScores[2];
    int32_t episodeStep;
    CkptAgent agents[kMaxAgents];
    CkptObject boxes[kMaxBoxes];
    CkptObject ramps[kMaxRamps];
    int32_t numHiders, numSeekers, numBoxes, numRamps;
};
static_assert(sizeof(CkptAgent) == 116 && sizeof(CkptObject) == 60, "Checkpoint member layout");
static_assert(sizeof(CkptData) == kCheckpointBytes, "Checkpoint layout (src/sim.hpp:283-313)");

__device__ inline void ckptSaveBody(const DeviceState &S, int w, int slot, CkptBody &b)
{
    for (int k = 0; k < 3; k++) b.pos[k] = BF(S, BF_PX + k, slot, w);
    for (int k = 0; k < 4; k++) b.rot[k] = BF(S, BF_QW + k, slot, w);
    for (int k = 0; k < 6; k++) b.vel[k] = BF(S, BF_VX + k, slot, w);
}

__device__ inline void ckptLoadBody(const DeviceState &S, int w, int slot, const CkptBody &b)
{
    for (int k = 0; k < 3; k++) BF(S, BF_PX + k, slot, w) = b.pos[k];
    for (int k = 0; k < 4; k++) BF(S, BF_QW + k, slot, w) = b.rot[k];
    for (int k = 0; k < 6; k++) BF(S, BF_VX + k, slot, w) = b.vel[k];
}

__global__ void k_save_checkpoint(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    if (!S.exCkptCtrl[w]) return;
    S.exCkptCtrl[w] = 0;

    CkptData &c = *reinterpret_cast<CkptData *>(S.exCkpt + size_t(w) * kCheckpointBytes);
    c.episodeKey[0] = (uint32_t)WI(S, WF_EP_KEY_A, w);
    c.episodeKey[1] = (uint32_t)WI(S, WF_EP_KEY_B, w);
    c.runningScores[0] = WI(S, WF_SCORE0, w);
    c.runningScores[1] = WI(S, WF_SCORE1, w);
    c.episodeStep = WI(S, WF_EPISODE_STEP, w);

    const int nDyn = WI(S, WF_NUM_DYN, w);
    const int nBoxes = WI(S, WF_NUM_BOXES, w), nRamps = WI(S, WF_NUM_RAMPS, w);
    const int nh = WI(S, WF_NUM_HIDERS, w), ns = WI(S, WF_NUM_SEEKERS, w);
    int idx = 0;
    for (int t = 0; t < 2; t++) {
        int cnt = t == 0 ? nh : ns;
        int base = t == 0 ? WB_HIDER_SLOT : WB_SEEKER_SLOT;
        for (int i = 0; i < cnt; i++) {
            int slot = WB(S, base + i, w);
            CkptAgent &a = c.agents[idx++];
            ckptSaveBody(S, w, slot, a);
            a.grabIdx = -1;
            for (int k = 0; k < 3; k++) { a.grabR1[k] = 0.f; a.grabR2[k] = 0.f; }
            for (int k = 0; k < 4; k++) { a.attachRot1[k] = 0.f; a.attachRot2[k] = 0.f; }
            a.separation = 0.f;
            int ak = slot - nDyn;
            uint32_t jm = JMETA(S, ak, w);
            if (jm != 0xFFFFFFFFu) {
                int other = int(jm & 0xFF);
                for (int k = 0; k < 3; k++) {
                    a.grabR1[k] = JF(S, JF_R1X + k, ak, w);
                    a.grabR2[k] = JF(S, JF_R2X + k, ak, w);
                }
                for (int k = 0; k < 4; k++) {
                    a.attachRot1[k] = JF(S, JF_A1W + k, ak, w);
                    a.attachRot2[k] = JF(S, JF_A2W + k, ak, w);
                }
                a.separation = JF(S, JF_SEP, ak, w);
                for (int j = 0; j < nBoxes; j++) if (WB(S, WB_BOX_SLOT + j, w) == other) a.grabIdx = j;
                for (int j = 0; j < nRamps; j++) if (WB(S, WB_RAMP_SLOT + j, w) == other) a.grabIdx = j + nBoxes;
            }
        }
    }
    c.numHiders = nh;
    c.numSeekers = ns;

    auto saveObj = [&](int slot, CkptObject &o) {
        ckptSaveBody(S, w, slot, o);
        uint32_t m = META(S, slot, w);
        o.team = (uint32_t)metaOwner(m);
        o.isLocked = metaResp(m) == RESP_STATIC ? 1 : 0;
    };
    c.numBoxes = nBoxes;
    for (int i = 0; i < nBoxes; i++) saveObj(WB(S, WB_BOX_SLOT + i, w), c.boxes[i]);
    c.numRamps = nRamps;
    for (int i = 0; i < nRamps; i++) saveObj(WB(S, WB_RAMP_SLOT + i, w), c.ramps[i]);
}

__global__ void k_load_checkpoint(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    WI(S, WF_DID_RESET, w) = 0;
    if (!S.exCkptCtrl[w]) return;
    S.exCkptCtrl[w] = 1;       // the reference leaves the trigger set (src/sim.cpp:963)

    const CkptData &c = *reinterpret_cast<const CkptData *>(S.exCkpt + size_t(w) * kCheckpointBytes);
    // The checkpoint tensor is user-writable.  The reference only asserts on the agent counts
    // (src/sim.cpp:1034-1036); counts outside this Manager's team bounds would index past the
    // world's agent rows, so such a record is refused: the world keeps its state and the refusal
    // is counted in the overflow counter (hs_overflow_count).
    if (c.numHiders < 0 || c.numHiders > S.p.maxHiders || c.numSeekers < 0 || c.numSeekers > S.p.maxSeekers ||
        c.numHiders + c.numSeekers > S.p.agentsPerWorld) {
        atomicAdd(S.overflow, 1u);
        return;
    }
    CkptRegen regen{{c.episodeKey[0], c.episodeKey[1]}, c.numHiders, c.numSeekers};
    regenerateWorld(S, w, &regen);
    WI(S, WF_SCORE0, w) = c.runningScores[0];
    WI(S, WF_SCORE1, w) = c.runningScores[1];
    WI(S, WF_EPISODE_STEP, w) = c.episodeStep;

    auto loadObj = [&](int slot, const CkptObject &o) {
        ckptLoadBody(S, w, slot, o);
        uint32_t m = META(S, slot, w);
        META(S, slot, w) = packMeta(metaObj(m), o.isLocked ? RESP_STATIC : RESP_DYNAMIC, int(o.team & 3),
                                    metaEid(m), metaLeaf(m));
    };
    const int nBoxes = WI(S, WF_NUM_BOXES, w), nRamps = WI(S, WF_NUM_RAMPS, w);
    for (int i = 0; i < c.numBoxes && i < nBoxes; i++) loadObj(WB(S, WB_BOX_SLOT + i, w), c.boxes[i]);
    for (int i = 0; i < c.numRamps && i < nRamps; i++) loadObj(WB(S, WB_RAMP_SLOT + i, w), c.ramps[i]);

    const int nDyn = WI(S, WF_NUM_DYN, w);
    const int nh = WI(S, WF_NUM_HIDERS, w), ns = WI(S, WF_NUM_SEEKERS, w);
    int idx = 0;
    for (int t = 0; t < 2; t++) {
        int cnt = t == 0 ? nh : ns;
        int base = t == 0 ? WB_HIDER_SLOT : WB_SEEKER_SLOT;
        for (int i = 0; i < cnt; i++) {
            int slot = WB(S, base + i, w);
            const CkptAgent &a = c.agents[idx++];
            ckptLoadBody(S, w, slot, a);
            int ak = slot - nDyn;
            if (a.grabIdx < 0 || a.grabIdx >= nBoxes + nRamps) { JMETA(S, ak, w) = 0xFFFFFFFFu; continue; }
            int other = a.grabIdx < nBoxes ? WB(S, WB_BOX_SLOT + a.grabIdx, w)
                                           : WB(S, WB_RAMP_SLOT + (a.grabIdx - nBoxes), w);
            // PhysicsSystem::makeFixedJoint (M/src/physics/physics.cpp:255-279)
            IdPool ids(S, w);
            int eid = ids.acquire();
            ids.commit();
            int seq = WI(S, WF_JOINT_SEQ, w);
            WI(S, WF_JOINT_SEQ, w) = seq + 1;
            JMETA(S, ak, w) = (uint32_t(seq & 0xFFFF) << 16) | (uint32_t(eid) << 8) | uint32_t(other);
            for (int k = 0; k < 4; k++) {
                JF(S, JF_A1W + k, ak, w) = a.attachRot1[k];
                JF(S, JF_A2W + k, ak, w) = a.attachRot2[k];
            }
            JF(S, JF_SEP, ak, w) = a.separation;
            for (int k = 0; k < 3; k++) {
                JF(S, JF_R1X + k, ak, w) = a.grabR1[k];
                JF(S, JF_R2X + k, ak, w) = a.grabR2[k];
            }
        }
    }
}

}  // namespace hs
