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

// actionSystem lives in hs_rays.cuh (it casts rays through the shared query).

}  // namespace hs
