// Observation stages: relative observations, visibility masks, 30-ray lidar,
// global debug positions.  One warp per agent-interface row; lanes own
// entities (obs / visibility) or rays (lidar).
//
// Behavioural contract:
//   collectObservationsSystem / computeRelativePosVelObs / quatToEuler /
//   computeLockObservation                         src/sim.cpp:372-565
//   computeVisibilitySystem (CPU branch, incl. the hiderTeamReward side effect)
//                                                  src/sim.cpp:567-605, 663-709
//   lidarSystem                                    src/sim.cpp:712-759
//   globalPositionsDebugSystem                     src/sim.cpp:895-941
#pragma once

#include "hs_trace.cuh"

namespace hs {

__device__ __forceinline__ Vec3 quatToEuler(Quat q)
{
    float sinr_cosp = 2.f * (q.w * q.x + q.y * q.z);
    float cosr_cosp = 1.f - 2.f * (q.x * q.x + q.y * q.y);
    float roll = atan2f(sinr_cosp, cosr_cosp);
    float sinp = 2.f * (q.w * q.y - q.z * q.x);
    float pitch = fabsf(sinp) >= 1.f ? copysignf(3.14159265358979323846f / 2.f, sinp) : asinf(sinp);
    float siny_cosp = 2.f * (q.w * q.z + q.x * q.y);
    float cosy_cosp = 1.f - 2.f * (q.y * q.y + q.z * q.z);
    float yaw = atan2f(siny_cosp, cosy_cosp);
    return {roll, pitch, yaw};
}

// writes 12 floats: pos, euler, linear vel, angular vel
__device__ __forceinline__ void writeRelObs(float *out, Vec3 origin, Quat toFrame, Vec3 fv, Vec3 fw,
                                            Vec3 x, Quat q, Vec3 v, Vec3 wv)
{
    x = rotate(toFrame, x - origin);
    q = qnormalize(toFrame * q);
    v = v - fv;
    wv = wv - fw;
    Vec3 e = quatToEuler(q);
    Vec3 lv = rotate(toFrame, v), av = rotate(toFrame, wv);
    out[0] = x.x; out[1] = x.y; out[2] = x.z;
    out[3] = e.x; out[4] = e.y; out[5] = e.z;
    out[6] = lv.x; out[7] = lv.y; out[8] = lv.z;
    out[9] = av.x; out[10] = av.y; out[11] = av.z;
}

// index of interface `other` among "other agents" of interface `self`, following
// the reference's enumeration (skip self, pad to maxAgents-1 with empty slots).
__device__ __forceinline__ int otherAgentIndex(int self, int other) { return other < self ? other : other - 1; }

__global__ void k_observations(DeviceState S)
{
    const int A = S.p.agentsPerWorld;
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= S.p.numWorlds * A) return;
    const int w = row / A, k = row % A;
    const int slot = WB(S, WB_IFACE_SLOT + k, w);
    if (slot == 0xFF) return;

    const int curStep = WI(S, WF_EPISODE_STEP, w);
    const int nDyn = WI(S, WF_NUM_DYN, w);
    Vec3 apos = loadPos(S, slot, w);
    Quat arot = loadRot(S, slot, w);
    Vec3 avel = loadV3(S, BF_VX, slot, w), aang = loadV3(S, BF_WX, slot, w);
    Quat toFrame = qinv(arot);

    if (lane == 0) {
        if (curStep <= kNumPrepSteps) S.exPrepCounter[row] = kNumPrepSteps - curStep;
        float *o = S.exSelfObs + size_t(row) * 13;
        Vec3 e = quatToEuler(arot);
        Vec3 lv = rotate(toFrame, avel), av = rotate(toFrame, aang);
        o[0] = apos.x; o[1] = apos.y; o[2] = apos.z;
        o[3] = e.x; o[4] = e.y; o[5] = e.z;
        o[6] = lv.x; o[7] = lv.y; o[8] = lv.z;
        o[9] = av.x; o[10] = av.y; o[11] = av.z;
        o[12] = JMETA(S, slot - nDyn, w) != 0xFFFFFFFFu ? 1.f : 0.f;
    } else if (lane <= kMaxBoxes) {
        int i = lane - 1;
        float *o = S.exBoxObs + (size_t(row) * kMaxBoxes + i) * 17;
        if (i >= WI(S, WF_NUM_BOXES, w)) {
            for (int c = 0; c < 17; c++) o[c] = 0.f;
        } else {
            int b = WB(S, WB_BOX_SLOT + i, w);
            uint32_t m = META(S, b, w);
            writeRelObs(o, apos, toFrame, avel, aang, loadPos(S, b, w), loadRot(S, b, w),
                        loadV3(S, BF_VX, b, w), loadV3(S, BF_WX, b, w));
            bool elongated = metaObj(m) == OBJ_BOX;     // boxSizes, level_gen.cpp:150,191
            o[12] = elongated ? 8.f : 2.f; o[13] = elongated ? 1.5f : 2.f; o[14] = 2.f;
            bool locked = metaResp(m) == RESP_STATIC;
            bool hider = metaOwner(m) == OWNER_HIDER;
            o[15] = locked && hider ? 1.f : 0.f;
            o[16] = locked && !hider ? 1.f : 0.f;
        }
    } else if (lane <= kMaxBoxes + kMaxRamps) {
        int i = lane - 1 - kMaxBoxes;
        float *o = S.exRampObs + (size_t(row) * kMaxRamps + i) * 14;
        if (i >= WI(S, WF_NUM_RAMPS, w)) {
            for (int c = 0; c < 14; c++) o[c] = 0.f;
        } else {
            int b = WB(S, WB_RAMP_SLOT + i, w);
            uint32_t m = META(S, b, w);
            writeRelObs(o, apos, toFrame, avel, aang, loadPos(S, b, w), loadRot(S, b, w),
                        loadV3(S, BF_VX, b, w), loadV3(S, BF_WX, b, w));
            bool locked = metaResp(m) == RESP_STATIC;
            bool hider = metaOwner(m) == OWNER_HIDER;
            o[12] = locked && hider ? 1.f : 0.f;
            o[13] = locked && !hider ? 1.f : 0.f;
        }
    } else if (lane <= kMaxBoxes + kMaxRamps + kMaxAgents) {
        int j = lane - 1 - kMaxBoxes - kMaxRamps;     // interface index 0..5
        int numAgents = WI(S, WF_NUM_AGENTS, w);
        if (j == k && j < numAgents) return;
        // enumeration of src/sim.cpp:527-564: active others first (skipping self),
        // then one empty entry per inactive index
        int outIdx = j < numAgents ? otherAgentIndex(k, j) : j - 1;
        if (outIdx < 0 || outIdx >= kMaxAgents - 1) return;
        float *o = S.exAgentObs + (size_t(row) * (kMaxAgents - 1) + outIdx) * 14;
        if (j >= numAgents) {
            for (int c = 0; c < 14; c++) o[c] = 0.f;
        } else {
            int b = WB(S, WB_IFACE_SLOT + j, w);
            writeRelObs(o, apos, toFrame, avel, aang, loadPos(S, b, w), loadRot(S, b, w),
                        loadV3(S, BF_VX, b, w), loadV3(S, BF_WX, b, w));
            o[12] = WB(S, WB_IFACE_TYPE + j, w) == AGENT_HIDER ? 1.f : 0.f;
            o[13] = JMETA(S, b - nDyn, w) != 0xFFFFFFFFu ? 1.f : 0.f;
        }
    }
}

__global__ void k_visibility(DeviceState S, float cosThreshold)
{
    const int A = S.p.agentsPerWorld;
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= S.p.numWorlds * A) return;
    const int w = row / A, k = row % A;
    const int slot = WB(S, WB_IFACE_SLOT + k, w);
    if (slot == 0xFF) return;

    const uint32_t *blob = BVH(S, w);
    Vec3 apos = loadPos(S, slot, w);
    Vec3 afwd = rotate(loadRot(S, slot, w), v3(0, 1, 0));

    auto visible = [&](int other) {
        Vec3 to = loadPos(S, other, w) - apos;
        float cosA = dot(normalize(to), afwd);
        if (cosA < cosThreshold) return 0.f;
        float t;
        return traceRay(blob, apos, to, 1.f, &t) == other ? 1.f : 0.f;
    };

    if (lane < kMaxBoxes) {
        float v = 0.f;
        if (lane < WI(S, WF_NUM_BOXES, w)) v = visible(WB(S, WB_BOX_SLOT + lane, w));
        S.exBoxVis[size_t(row) * kMaxBoxes + lane] = v;
    } else if (lane < kMaxBoxes + kMaxRamps) {
        int i = lane - kMaxBoxes;
        float v = 0.f;
        if (i < WI(S, WF_NUM_RAMPS, w)) v = visible(WB(S, WB_RAMP_SLOT + i, w));
        S.exRampVis[size_t(row) * kMaxRamps + i] = v;
    } else if (lane < kMaxBoxes + kMaxRamps + kMaxAgents) {
        int j = lane - kMaxBoxes - kMaxRamps;
        int numAgents = WI(S, WF_NUM_AGENTS, w);
        if (j == k && j < numAgents) return;
        int outIdx = j < numAgents ? otherAgentIndex(k, j) : j - 1;
        if (outIdx < 0 || outIdx >= kMaxAgents - 1) return;
        float v = 0.f;
        if (j < numAgents) {
            v = visible(WB(S, WB_IFACE_SLOT + j, w));
            // CPU-backend side effect (src/sim.cpp:700-705)
            if (v == 1.f && WB(S, WB_IFACE_TYPE + k, w) == AGENT_SEEKER &&
                WB(S, WB_IFACE_TYPE + j, w) == AGENT_HIDER) {
                WI(S, WF_HIDER_REWARD, w) = __float_as_int(-1.f);
            }
        }
        S.exAgentVis[size_t(row) * (kMaxAgents - 1) + outIdx] = v;
    }
}

__global__ void k_lidar(DeviceState S)
{
    const int A = S.p.agentsPerWorld;
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= S.p.numWorlds * A) return;
    const int w = row / A, k = row % A;
    const int slot = WB(S, WB_IFACE_SLOT + k, w);
    if (slot == 0xFF || lane >= 30) return;

    const float pi = 3.14159265358979323846f;
    Vec3 pos = loadPos(S, slot, w);
    Quat rot = loadRot(S, slot, w);
    Vec3 fwd = rotate(rot, v3(0, 1, 0));
    Vec3 right = rotate(rot, v3(1, 0, 0));

    float theta = 2.f * pi * (float(lane) / float(30)) + pi / 2.f;
    float x, y;
    hs_det_sincosf(theta, &y, &x);
    Vec3 dir = normalize(x * right + y * fwd);
    float t;
    int hit = traceRay(BVH(S, w), pos, dir, 200.f, &t);
    S.exLidar[size_t(row) * 30 + lane] = hit < 0 ? 0.f : t;
}

__global__ void k_debug_positions(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    float *o = S.exGlobalPos + size_t(w) * 17 * 2;
    int nb = WI(S, WF_NUM_BOXES, w), nr = WI(S, WF_NUM_RAMPS, w);
    for (int i = 0; i < kMaxBoxes; i++) {
        float x = 0.f, y = 0.f;
        if (i < nb) { int s = WB(S, WB_BOX_SLOT + i, w); x = BF(S, BF_PX, s, w); y = BF(S, BF_PY, s, w); }
        o[i * 2] = x; o[i * 2 + 1] = y;
    }
    for (int i = 0; i < kMaxRamps; i++) {
        float x = 0.f, y = 0.f;
        if (i < nr) { int s = WB(S, WB_RAMP_SLOT + i, w); x = BF(S, BF_PX, s, w); y = BF(S, BF_PY, s, w); }
        o[(kMaxBoxes + i) * 2] = x; o[(kMaxBoxes + i) * 2 + 1] = y;
    }
    float *ao = o + (kMaxBoxes + kMaxRamps) * 2;
    int off = 0;
    int nh = WI(S, WF_NUM_HIDERS, w), ns = WI(S, WF_NUM_SEEKERS, w);
    for (int i = 0; i < nh; i++, off++) {
        int s = WB(S, WB_HIDER_SLOT + i, w);
        ao[off * 2] = BF(S, BF_PX, s, w); ao[off * 2 + 1] = BF(S, BF_PY, s, w);
    }
    for (int i = 0; i < ns; i++, off++) {
        int s = WB(S, WB_SEEKER_SLOT + i, w);
        ao[off * 2] = BF(S, BF_PX, s, w); ao[off * 2 + 1] = BF(S, BF_PY, s, w);
    }
    // src/sim.cpp:937-939 increments twice per iteration: every other trailing slot is cleared
    for (; off < kMaxAgents; off += 2) { ao[off * 2] = 0.f; ao[off * 2 + 1] = 0.f; }
}

}  // namespace hs
