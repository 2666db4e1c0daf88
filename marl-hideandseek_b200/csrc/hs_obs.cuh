// Observation stages: relative observations and global debug positions (visibility masks and
// the 30-ray lidar live in hs_rays.cuh).
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

// collectObservationsSystem.  One thread per output record, records grouped BY KIND across the
// whole slice — [self rows][box records][ramp records][other-agent records] — so a warp runs a
// single code path (the euler-angle math is ~200 instructions and the four kinds differ).  A CTA
// owns kObsRecs consecutive records of ONE kind, i.e. a contiguous range of the export tensor:
// threads format their record into shared memory and the CTA then streams the range out with
// consecutive lanes writing consecutive words (the observation tensors are the biggest HBM
// writer of the step; a thread writing its own 17 floats costs 17 scattered sectors per warp
// store).  Records of inactive agent rows are not written, as in the reference.
constexpr int kObsRecs = 252;            // divisible by 9 (boxes), 6 (agents), 2 (ramps): CTAs start on a row
constexpr int kObsThreads = 256;
constexpr int kObsMaxLen = 17;

struct ObsFrame { Vec3 pos; Quat toFrame; Vec3 vel, ang; };

__device__ __forceinline__ ObsFrame obsFrame(const DeviceState &S, int slot, int w)
{
    const Pose p = loadPose(S, REC_POSE, slot, w);
    const Vel6 v = loadVel6(S, REC_VEL, slot, w);
    return {p.x, qinv(p.q), v.v, v.w};
}

__device__ __forceinline__ void writeRelObsOf(float *o, const ObsFrame &F, const DeviceState &S, int b, int w)
{
    const Pose p = loadPose(S, REC_POSE, b, w);
    const Vel6 v = loadVel6(S, REC_VEL, b, w);
    writeRelObs(o, F.pos, F.toFrame, F.vel, F.ang, p.x, p.q, v.v, v.w);
}

__host__ __device__ __forceinline__ long long obsBlocks(long long recs) { return (recs + kObsRecs - 1) / kObsRecs; }

__global__ void __launch_bounds__(kObsThreads) k_observations(DeviceState S)
{
    __shared__ float s_out[kObsRecs * kObsMaxLen];
    __shared__ uint8_t s_valid[kObsRecs];
    const int A = S.p.agentsPerWorld;
    const long long R = (long long)S.p.numWorlds * A;
    const int tid = threadIdx.x;
    long long blk = blockIdx.x;
    const long long nbSelf = obsBlocks(R), nbBox = obsBlocks(R * kMaxBoxes), nbRamp = obsBlocks(R * kMaxRamps);

    if (tid < kObsRecs) s_valid[tid] = 0;
    __syncthreads();

    float *dst = nullptr;          // first word of this CTA's output range
    int len = 0, numOut = 0;       // floats per record, records in the range
    if (blk < nbSelf) {
        // ---- self observation + prep counter ---------------------------------------------------
        const long long base = blk * kObsRecs;
        len = 13; numOut = int(min((long long)kObsRecs, R - base));
        dst = S.exSelfObs + base * 13;
        const long long idx = base + tid;
        if (tid < kObsRecs && idx < R) {
            const int row = int(idx), w = row / A, k = row % A;
            const int slot = WB(S, WB_IFACE_SLOT + k, w);
            if (slot != 0xFF) {
                const int curStep = WI(S, WF_EPISODE_STEP, w);
                if (curStep <= kNumPrepSteps) S.exPrepCounter[row] = kNumPrepSteps - curStep;
                const Pose p = loadPose(S, REC_POSE, slot, w);
                const Vel6 v = loadVel6(S, REC_VEL, slot, w);
                const Quat toFrame = qinv(p.q);
                float *o = s_out + tid * 13;
                Vec3 e = quatToEuler(p.q);
                Vec3 lv = rotate(toFrame, v.v), av = rotate(toFrame, v.w);
                o[0] = p.x.x; o[1] = p.x.y; o[2] = p.x.z;
                o[3] = e.x; o[4] = e.y; o[5] = e.z;
                o[6] = lv.x; o[7] = lv.y; o[8] = lv.z;
                o[9] = av.x; o[10] = av.y; o[11] = av.z;
                o[12] = JMETA(S, slot - WI(S, WF_NUM_DYN, w), w) != 0xFFFFFFFFu ? 1.f : 0.f;
                s_valid[tid] = 1;
            }
        }
    } else if ((blk -= nbSelf) < nbBox) {
        // ---- boxes ------------------------------------------------------------------------------
        const long long base = blk * kObsRecs, total = R * kMaxBoxes;
        len = 17; numOut = int(min((long long)kObsRecs, total - base));
        dst = S.exBoxObs + base * 17;
        const long long idx = base + tid;
        if (tid < kObsRecs && idx < total) {
            const int row = int(idx / kMaxBoxes), i = int(idx % kMaxBoxes), w = row / A, k = row % A;
            const int slot = WB(S, WB_IFACE_SLOT + k, w);
            if (slot != 0xFF) {
                float *o = s_out + tid * 17;
                if (i >= WI(S, WF_NUM_BOXES, w)) {
                    for (int c = 0; c < 17; c++) o[c] = 0.f;
                } else {
                    const int b = WB(S, WB_BOX_SLOT + i, w);
                    const uint32_t m = META(S, b, w);
                    writeRelObsOf(o, obsFrame(S, slot, w), S, b, w);
                    bool elongated = metaObj(m) == OBJ_BOX;     // boxSizes, level_gen.cpp:150,191
                    o[12] = elongated ? 8.f : 2.f; o[13] = elongated ? 1.5f : 2.f; o[14] = 2.f;
                    bool locked = metaResp(m) == RESP_STATIC;
                    bool hider = metaOwner(m) == OWNER_HIDER;
                    o[15] = locked && hider ? 1.f : 0.f;
                    o[16] = locked && !hider ? 1.f : 0.f;
                }
                s_valid[tid] = 1;
            }
        }
    } else if ((blk -= nbBox) < nbRamp) {
        // ---- ramps ------------------------------------------------------------------------------
        const long long base = blk * kObsRecs, total = R * kMaxRamps;
        len = 14; numOut = int(min((long long)kObsRecs, total - base));
        dst = S.exRampObs + base * 14;
        const long long idx = base + tid;
        if (tid < kObsRecs && idx < total) {
            const int row = int(idx / kMaxRamps), i = int(idx % kMaxRamps), w = row / A, k = row % A;
            const int slot = WB(S, WB_IFACE_SLOT + k, w);
            if (slot != 0xFF) {
                float *o = s_out + tid * 14;
                if (i >= WI(S, WF_NUM_RAMPS, w)) {
                    for (int c = 0; c < 14; c++) o[c] = 0.f;
                } else {
                    const int b = WB(S, WB_RAMP_SLOT + i, w);
                    const uint32_t m = META(S, b, w);
                    writeRelObsOf(o, obsFrame(S, slot, w), S, b, w);
                    bool locked = metaResp(m) == RESP_STATIC;
                    bool hider = metaOwner(m) == OWNER_HIDER;
                    o[12] = locked && hider ? 1.f : 0.f;
                    o[13] = locked && !hider ? 1.f : 0.f;
                }
                s_valid[tid] = 1;
            }
        }
    } else {
        // ---- other agents: interface index j = 0..5 (src/sim.cpp:527-564: active others first,
        // skipping self, then one empty entry per inactive index); 6 threads produce the 5 output
        // records of a row ---------------------------------------------------------------------
        blk -= nbRamp;
        const long long base = blk * kObsRecs, total = R * kMaxAgents;
        constexpr int kRowsPerCta = kObsRecs / kMaxAgents, kOutPerRow = kMaxAgents - 1;
        const long long rowBase = base / kMaxAgents;
        len = 14; numOut = int(min((long long)kRowsPerCta, R - rowBase)) * kOutPerRow;
        dst = S.exAgentObs + rowBase * kOutPerRow * 14;
        const long long idx = base + tid;
        if (tid < kObsRecs && idx < total) {
            const int lrow = tid / kMaxAgents, j = tid % kMaxAgents;
            const int row = int(rowBase) + lrow, w = row / A, k = row % A;
            const int slot = WB(S, WB_IFACE_SLOT + k, w);
            const int numAgents = WI(S, WF_NUM_AGENTS, w);
            const int outIdx = j < numAgents ? otherAgentIndex(k, j) : j - 1;
            if (slot != 0xFF && !(j == k && j < numAgents) && outIdx >= 0 && outIdx < kOutPerRow) {
                const int pos = lrow * kOutPerRow + outIdx;
                float *o = s_out + pos * 14;
                if (j >= numAgents) {
                    for (int c = 0; c < 14; c++) o[c] = 0.f;
                } else {
                    const int b = WB(S, WB_IFACE_SLOT + j, w);
                    writeRelObsOf(o, obsFrame(S, slot, w), S, b, w);
                    o[12] = WB(S, WB_IFACE_TYPE + j, w) == AGENT_HIDER ? 1.f : 0.f;
                    o[13] = JMETA(S, b - WI(S, WF_NUM_DYN, w), w) != 0xFFFFFFFFu ? 1.f : 0.f;
                }
                s_valid[pos] = 1;
            }
        }
    }
    __syncthreads();
    // ---- stream the range out: consecutive lanes, consecutive words ---------------------------
    const int words = numOut * len;
    for (int e = tid; e < words; e += kObsThreads)
        if (s_valid[e / len]) dst[e] = s_out[e];
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
