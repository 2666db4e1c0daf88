// Agent-facing stages: action application (move / lock / grab), rewards and
// dones, episode results.
//
// Behavioural contract:
//   movementSystem / instantMovementSystem / agentZeroVelSystem   src/sim.cpp:202-268
//   actionSystem (lock ray, grab ray, joint create/destroy)        src/sim.cpp:270-370
//   rewardsVisSystem / outputRewardsDonesSystem / updateEpisodeResultsSystem
//                                                                  src/sim.cpp:763-893
#pragma once

#include "hs_trace.cuh"

namespace hs {

__device__ __forceinline__ bool agentFrozen(const DeviceState &S, int w, int type)
{
    return type == AGENT_SEEKER && WI(S, WF_EPISODE_STEP, w) < kNumPrepSteps - 1;
}

// one thread per agent-interface row
__global__ void k_move(DeviceState S)
{
    const int A = S.p.agentsPerWorld;
    int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= S.p.numWorlds * A) return;
    int w = row / A, k = row % A;
    int slot = WB(S, WB_IFACE_SLOT + k, w);
    if (slot == 0xFF) return;
    if (agentFrozen(S, w, WB(S, WB_IFACE_TYPE + k, w))) return;

    const int32_t *act = S.exAction + size_t(row) * 5;
    float fx, fy, tz;
    if (S.p.simFlags & FLAG_ZERO_AGENT_VEL) {
        fx = 400.f * float(act[0] - 2);
        fy = 400.f * float(act[1] - 2);
        tz = 120.f * float(act[2] - 2);
    } else {
        fx = 12.f * float(act[0] - 5);
        fy = 12.f * float(act[1] - 5);
        tz = 3.f * float(act[2] - 5);
    }
    Quat rot = loadRot(S, slot, w);
    storeV3(S, BF_FX, slot, w, rotate(rot, {fx, fy, 0.f}));
    storeV3(S, BF_TX, slot, w, {0.f, 0.f, tz});
}

// one thread per DynAgent body (flag-gated)
__global__ void k_zero_agent_vel(DeviceState S)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= S.p.numWorlds * kMaxAgents) return;
    int w = idx % S.p.numWorlds, k = idx / S.p.numWorlds;
    if (k >= WI(S, WF_NUM_AGENTS, w)) return;
    int slot = WI(S, WF_NUM_DYN, w) + k;
    BF(S, BF_VX, slot, w) = 0.f;
    BF(S, BF_VY, slot, w) = 0.f;
    BF(S, BF_VZ, slot, w) = fminf(BF(S, BF_VZ, slot, w), 0.f);
    BF(S, BF_WX, slot, w) = 0.f;
    BF(S, BF_WY, slot, w) = 0.f;
    BF(S, BF_WZ, slot, w) = 0.f;
}

// actionSystem: one thread per world, agents in interface order (lock / grab
// mutate shared body state and the entity-id pool, so order matters).
__global__ void k_action(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    const int A = S.p.agentsPerWorld;
    const uint32_t *blob = BVH(S, w);
    const int nDyn = WI(S, WF_NUM_DYN, w);

    for (int k = 0; k < A; k++) {
        int slot = WB(S, WB_IFACE_SLOT + k, w);
        if (slot == 0xFF) continue;
        int type = WB(S, WB_IFACE_TYPE + k, w);
        if (agentFrozen(S, w, type)) continue;
        int32_t *act = S.exAction + (size_t(w) * A + k) * 5;

        if (act[4] == 1) {
            Vec3 pos = loadPos(S, slot, w);
            Quat rot = loadRot(S, slot, w);
            float t;
            int hit = traceRay(blob, pos + 0.5f * v3(0, 0, 1), rotate(rot, v3(0, 1, 0)), 2.5f, &t);
            if (hit >= 0) {
                uint32_t m = META(S, hit, w);
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
                META(S, hit, w) = packMeta(metaObj(m), resp, owner, metaEid(m), metaLeaf(m));
            }
        }

        if (act[3] == 1) {
            int ak = slot - nDyn;       // DynAgent row
            uint32_t jm = JMETA(S, ak, w);
            if (jm != 0xFFFFFFFFu) {
                IdPool ids(S, w);
                ids.release((jm >> 8) & 0xFF);
                ids.commit();
                JMETA(S, ak, w) = 0xFFFFFFFFu;
            } else {
                Vec3 pos = loadPos(S, slot, w);
                Quat rot = loadRot(S, slot, w);
                Vec3 ro = pos + 0.5f * v3(0, 0, 1);
                Vec3 rd = rotate(rot, v3(0, 1, 0));
                float t;
                int hit = traceRay(blob, ro, rd, 2.5f, &t);
                if (hit >= 0) {
                    uint32_t m = META(S, hit, w);
                    if (metaOwner(m) == OWNER_NONE && metaResp(m) == RESP_DYNAMIC) {
                        Vec3 opos = loadPos(S, hit, w);
                        Quat orot = loadRot(S, hit, w);
                        Vec3 r1 = 1.25f * v3(0, 1, 0) + 0.5f * v3(0, 0, 1);
                        Vec3 hitPos = ro + rd * t;
                        Vec3 r2 = rotate(qinv(orot), hitPos - opos);
                        Quat a2 = qnormalize(qinv(orot) * rot);
                        IdPool ids(S, w);
                        int eid = ids.acquire();
                        ids.commit();
                        int seq = WI(S, WF_JOINT_SEQ, w);
                        WI(S, WF_JOINT_SEQ, w) = seq + 1;
                        JMETA(S, ak, w) = (uint32_t(seq & 0xFFFF) << 16) | (uint32_t(eid) << 8) | uint32_t(hit);
                        JF(S, JF_A1W, ak, w) = 1.f; JF(S, JF_A1X, ak, w) = 0.f;
                        JF(S, JF_A1Y, ak, w) = 0.f; JF(S, JF_A1Z, ak, w) = 0.f;
                        JF(S, JF_A2W, ak, w) = a2.w; JF(S, JF_A2X, ak, w) = a2.x;
                        JF(S, JF_A2Y, ak, w) = a2.y; JF(S, JF_A2Z, ak, w) = a2.z;
                        JF(S, JF_SEP, ak, w) = t - 1.25f;
                        JF(S, JF_R1X, ak, w) = r1.x; JF(S, JF_R1Y, ak, w) = r1.y; JF(S, JF_R1Z, ak, w) = r1.z;
                        JF(S, JF_R2X, ak, w) = r2.x; JF(S, JF_R2Y, ak, w) = r2.y; JF(S, JF_R2Z, ak, w) = r2.z;
                    }
                }
            }
        }

        act[0] = 2; act[1] = 2; act[2] = 2; act[3] = 0; act[4] = 0;
    }
}

}  // namespace hs
