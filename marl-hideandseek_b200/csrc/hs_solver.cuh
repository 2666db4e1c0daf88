// XPBD rigid-body solver: substep integration, sequential position solve over
// contacts then joints, velocity derivation, restitution + dynamic friction.
//
// Behavioural contract (M/src/physics/xpbd.cpp):
//   substepRigidBodies 100-185 · computePositionalLambda / applyPositionalUpdate 206-285
//   computeAngularUpdate / applyAngularUpdate 287-319 · handleContactConstraint 321-403
//   getLocalSpaceContacts / getAvgContact / handleContact 405-550
//   applyJointOrientationConstraint / handleJointConstraint (Fixed) 552-718
//   solvePositions 720-736 · setVelocities 738-779
//   applyFrictionVelocityUpdate / applyRestitutionVelocityUpdate /
//   solveVelocitiesForContact / solveVelocities 781-1052
// Gauss-Seidel order = contact creation order, then joints in creation order
// (SURVEY.md App. B-6), so the position/velocity solves run one thread per world.
#pragma once

#include "hs_narrowphase.cuh"

namespace hs {

__device__ __forceinline__ float substepH() { return kDeltaT / float(kSubsteps); }

// one thread per (slot, world)
__global__ void k_substep(DeviceState S)
{
    const int N = S.p.numWorlds;
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)kMaxBodies * N) return;
    int s = int(idx % kMaxBodies), w = int(idx / kMaxBodies);
    if (s >= WI(S, WF_NUM_BODIES, w)) return;

    uint32_t m = META(S, s, w);
    Pose cur = loadPose(S, REC_POSE, s, w);
    Vec3 x = cur.x;
    Quat q = cur.q;
    storePose(S, REC_PREV, s, w, x, q);

    if (metaResp(m) == RESP_STATIC) {
        storePose(S, REC_PRE, s, w, x, q);
        storeVel6(S, REC_PREVEL, s, w, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f});
        return;
    }

    Vel6 vel = loadVel6(S, REC_VEL, s, w);
    Vel6 ft = loadVel6(S, REC_FORCE, s, w);
    Vec3 v = vel.v;
    Vec3 omega = vel.w;
    const ObjectData &O = c_assets.obj[metaObj(m)];
    float inv_m = O.invMass;
    Vec3 inv_I = {O.invInertia[0], O.invInertia[1], O.invInertia[2]};
    const float h = substepH();

    if (metaResp(m) == RESP_DYNAMIC) {
        const Vec3 g = -9.8f * v3(0, 0, 1);
        v += h * g;
    }
    v += (h * inv_m) * ft.v;
    x += h * v;

    Vec3 I = {inv_I.x == 0 ? 0.0f : 1.0f / inv_I.x,
              inv_I.y == 0 ? 0.0f : 1.0f / inv_I.y,
              inv_I.z == 0 ? 0.0f : 1.0f / inv_I.z};
    Quat toLocal = qinv(q);
    Vec3 tauLocal = rotate(toLocal, ft.w);
    Vec3 omegaLocal = rotate(toLocal, omega);
    Vec3 IOmega = mulDiag(I, omegaLocal);
    omegaLocal += h * mulDiag(inv_I, tauLocal - cross(omegaLocal, IOmega));
    omega = rotate(q, omegaLocal);

    Quat applyOmega = fromAngularVec((0.5f * h) * omega);
    q = q + applyOmega * q;
    q = qnormalize(q);

    storePose(S, REC_POSE, s, w, x, q);
    storePose(S, REC_PRE, s, w, x, q);
    storeVel6(S, REC_PREVEL, s, w, v, omega);
}

struct MassData { float inv_m; Vec3 inv_I; float muS, muD; };

__device__ __forceinline__ MassData massOf(uint32_t meta)
{
    const ObjectData &O = c_assets.obj[metaObj(meta)];
    MassData d = {O.invMass, {O.invInertia[0], O.invInertia[1], O.invInertia[2]}, O.muS, O.muD};
    if (metaResp(meta) == RESP_STATIC) { d.inv_m = 0.f; d.inv_I = {0.f, 0.f, 0.f}; }
    return d;
}

__device__ __forceinline__ float positionalLambda(Vec3 ta1, Vec3 ta2, Vec3 ra1, Vec3 ra2,
                                                  float inv_m1, float inv_m2, float c, float alpha)
{
    float w1 = inv_m1 + dot(ta1, ra1);
    float w2 = inv_m2 + dot(ta2, ra2);
    return -c / (w1 + w2 + alpha);
}

__device__ __forceinline__ void applyPositional(Vec3 &x1, Vec3 &x2, Quat &q1, Quat &q2,
                                                Vec3 ra1, Vec3 ra2, float inv_m1, float inv_m2,
                                                Vec3 n, float lambda)
{
    x1 += (lambda * inv_m1) * n;
    x2 -= (lambda * inv_m2) * n;
    float half = 0.5f * lambda;
    Vec3 u1 = rotate(q1, half * ra1);
    Vec3 u2 = rotate(q2, half * ra2);
    q1 = q1 + fromAngularVec(u1) * q1;
    q2 = q2 - fromAngularVec(u2) * q2;
    q1 = qnormalize(q1);
    q2 = qnormalize(q2);
}

__device__ __forceinline__ float applyPositionalFull(Vec3 &x1, Vec3 &x2, Quat &q1, Quat &q2,
                                                     Vec3 r1, Vec3 r2, const MassData &m1, const MassData &m2,
                                                     Vec3 n, float c, float alpha)
{
    Vec3 n1 = rotate(qinv(q1), n), n2 = rotate(qinv(q2), n);
    Vec3 ta1 = cross(r1, n1), ta2 = cross(r2, n2);
    Vec3 ra1 = mulDiag(m1.inv_I, ta1), ra2 = mulDiag(m2.inv_I, ta2);
    float lambda = positionalLambda(ta1, ta2, ra1, ra2, m1.inv_m, m2.inv_m, c, alpha);
    applyPositional(x1, x2, q1, q2, ra1, ra2, m1.inv_m, m2.inv_m, n, lambda);
    return lambda;
}

struct ContactView {
    Vec3 pts[4];
    float depth[4];
    int n;
    Vec3 normal;
    float lambdaN;
};

__device__ __forceinline__ ContactView loadContact(const float *c, uint32_t meta)
{
    // contact rows are 80 bytes, 16-byte aligned (hs_narrowphase.cuh::writeContact)
    const float4 *c4 = reinterpret_cast<const float4 *>(c);
    ContactView v;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float4 p = c4[i];
        v.pts[i] = {p.x, p.y, p.z};
        v.depth[i] = p.w;
    }
    float4 n = c4[4];
    v.n = int(meta >> 16) & 0xFF;
    v.normal = {n.x, n.y, n.z};
    v.lambdaN = n.w;
    return v;
}

// getAvgContact: returns true when the contact has zero total penetration
__device__ __forceinline__ bool avgContact(const ContactView &c, Vec3 *avg, float *pen)
{
    float maxPen = -FLT_MAX, sum = 0.f;
    for (int i = 0; i < c.n; i++) {
        if (c.depth[i] > maxPen) maxPen = c.depth[i];
        sum += c.depth[i];
    }
    if (sum == 0.f) return true;
    Vec3 a = {0.f, 0.f, 0.f};
    for (int i = 0; i < c.n; i++) a += (c.depth[i] / sum) * c.pts[i];
    *avg = a;
    *pen = maxPen;
    return false;
}

__device__ __forceinline__ void localContacts(Vec3 px1, Quat pq1, Vec3 px2, Quat pq2,
                                              Vec3 contact1, float pen, Vec3 normal, Vec3 *r1, Vec3 *r2)
{
    Vec3 contact2 = contact1 - normal * pen;
    *r1 = rotate(qinv(pq1), contact1 - px1);
    *r2 = rotate(qinv(pq2), contact2 - px2);
}

__device__ inline void handleContact(const DeviceState &S, int w, uint32_t meta, float *cdata,
                                     bool write1, bool write2)
{
    const int b1 = meta & 0xFF, b2 = (meta >> 8) & 0xFF;
    ContactView c = loadContact(cdata, meta);
    MassData m1 = massOf(META(S, b1, w)), m2 = massOf(META(S, b2, w));
    float avgMuS = 0.5f * (m1.muS + m2.muS);

    Vec3 avg; float pen;
    if (avgContact(c, &avg, &pen)) return;

    const Pose cur1 = loadPose(S, REC_POSE, b1, w), cur2 = loadPose(S, REC_POSE, b2, w);
    const Pose pre1 = loadPose(S, REC_PRE, b1, w), pre2 = loadPose(S, REC_PRE, b2, w);
    const Pose prev1 = loadPose(S, REC_PREV, b1, w), prev2 = loadPose(S, REC_PREV, b2, w);
    Vec3 x1 = cur1.x, x2 = cur2.x;
    Quat q1 = cur1.q, q2 = cur2.q;
    Vec3 r1, r2;
    localContacts(pre1.x, pre1.q, pre2.x, pre2.q, avg, pen, c.normal, &r1, &r2);

    // handleContactConstraint
    Vec3 n = c.normal;
    Vec3 p1 = rotate(q1, r1) + x1, p2 = rotate(q2, r2) + x2;
    float d = dot(p1 - p2, n);
    float lambda_n = 0.f;
    if (d > 0) {
        lambda_n = applyPositionalFull(x1, x2, q1, q2, r1, r2, m1, m2, n, d, 0.f);

        Vec3 p1_hat = rotate(prev1.q, r1) + prev1.x;
        Vec3 p2_hat = rotate(prev2.q, r2) + prev2.x;
        p1 = rotate(q1, r1) + x1;
        p2 = rotate(q2, r2) + x2;
        Vec3 dp = (p1 - p1_hat) - (p2 - p2_hat);
        Vec3 dpt = dp - dot(dp, n) * n;
        float tmag = length(dpt);
        if (tmag > 0.f) {
            Vec3 t = dpt / tmag;
            Vec3 t1 = rotate(qinv(q1), t), t2 = rotate(qinv(q2), t);
            Vec3 fa1 = cross(r1, t1), fa2 = cross(r2, t2);
            Vec3 fr1 = mulDiag(m1.inv_I, fa1), fr2 = mulDiag(m2.inv_I, fa2);
            float lambda_t = positionalLambda(fa1, fa2, fr1, fr2, m1.inv_m, m2.inv_m, tmag, 0.f);
            if (lambda_t > lambda_n * avgMuS) {
                applyPositional(x1, x2, q1, q2, fr1, fr2, m1.inv_m, m2.inv_m, t, lambda_t);
            }
        }
    }
    cdata[19] = lambda_n;

    if (write1) storePose(S, REC_POSE, b1, w, x1, q1);
    if (write2) storePose(S, REC_POSE, b2, w, x2, q2);
}

__device__ inline void handleFixedJoint(const DeviceState &S, int w, int k, int b1, int b2)
{
    MassData m1 = massOf(META(S, b1, w)), m2 = massOf(META(S, b2, w));
    const Pose jp1 = loadPose(S, REC_POSE, b1, w), jp2 = loadPose(S, REC_POSE, b2, w);
    Vec3 x1 = jp1.x, x2 = jp2.x;
    Quat q1 = jp1.q, q2 = jp2.q;
    Quat a1 = {JF(S, JF_A1W, k, w), JF(S, JF_A1X, k, w), JF(S, JF_A1Y, k, w), JF(S, JF_A1Z, k, w)};
    Quat a2 = {JF(S, JF_A2W, k, w), JF(S, JF_A2X, k, w), JF(S, JF_A2Y, k, w), JF(S, JF_A2Z, k, w)};
    Vec3 jr1 = {JF(S, JF_R1X, k, w), JF(S, JF_R1Y, k, w), JF(S, JF_R1Z, k, w)};
    Vec3 jr2 = {JF(S, JF_R2X, k, w), JF(S, JF_R2Y, k, w), JF(S, JF_R2Z, k, w)};
    float sep = JF(S, JF_SEP, k, w);

    // applyJointOrientationConstraint
    {
        Quat o1 = qnormalize(q1 * a1), o2 = qnormalize(q2 * a2);
        Quat diff = o1 * qinv(o2);
        Vec3 dq = 2.f * v3(diff.x, diff.y, diff.z);
        float mag = length(dq);
        if (mag > 0) {
            dq = dq / mag;
            Vec3 l1 = rotate(qinv(q1), dq), l2 = rotate(qinv(q2), dq);
            Vec3 ra1 = mulDiag(m1.inv_I, l1), ra2 = mulDiag(m2.inv_I, l2);
            float w1 = dot(l1, ra1), w2 = dot(l2, ra2);
            float dl = -mag / (w1 + w2 + 0.f);
            float half = 0.5f * dl;
            Quat u1 = fromAngularVec(rotate(q1, half * ra1));
            Quat u2 = fromAngularVec(rotate(q2, half * ra2));
            q1 = qnormalize(q1 + u1 * q1);
            q2 = qnormalize(q2 - u2 * q2);
        }
    }

    Vec3 r1w = rotate(q1, jr1) + x1, r2w = rotate(q2, jr2) + x2;
    Vec3 dr = r2w - r1w;
    Quat axes = qnormalize(q1 * a1);
    Vec3 ax = rotate(axes, v3(0, 1, 0));
    Vec3 bx = rotate(axes, v3(1, 0, 0));
    Vec3 cx = cross(ax, bx);
    Vec3 corr = {0.f, 0.f, 0.f};
    corr -= (dot(dr, ax) - sep) * ax;
    corr -= dot(dr, bx) * bx;
    corr -= dot(dr, cx) * cx;

    float cmag = length(corr);
    if (cmag > 0.f) {
        corr = corr / cmag;
        applyPositionalFull(x1, x2, q1, q2, jr1, jr2, m1, m2, corr, cmag, 0.f);
    }
    storePose(S, REC_POSE, b1, w, x1, q1);
    storePose(S, REC_POSE, b2, w, x2, q2);
}

// ---- ordered Gauss-Seidel without one-thread-per-world serialisation -----------------------------
// The reference solves contacts strictly in creation order (xpbd.cpp:724-735).  Contact i only
// depends on an earlier contact if they share a body whose state the pass can change, so each
// world's contacts are assigned a level = 1 + the level of the latest earlier contact sharing
// such a body, and a group of kSolveLanes lanes sweeps the levels, one contact per lane.
// Contacts inside a level touch disjoint mutable bodies: every read sees exactly the value the
// sequential order would produce, so results are bit-identical to the serial solver.  Bodies
// the pass cannot change carry no dependency and are never written: position pass — Static
// bodies with identity rotation (walls, ground plane; a rotated Static body still has its
// quaternion re-normalised by the reference); velocity pass — every Static body (inverse mass
// and inertia are zero).  4 worlds share a warp, so lanes stay busy although one world has
// only a handful of contacts per level.
constexpr int kSolveLanes = 8;
constexpr int kMaxLevels = 126;    // deeper dependency chains fall back to the serial order
constexpr int kSolveGroups = 16;
constexpr int kSolveThreads = kSolveGroups * kSolveLanes;

struct SolveGroupSmem {
    uint32_t cm[kMaxCand];       // compacted contact meta (creation order) | candidate slot << 24
    uint8_t clvl[kMaxCand];
    uint8_t order[kMaxCand];     // contact indices sorted by (level, creation order)
    uint16_t lvlStart[kMaxLevels + 2];
    uint8_t blvl[kMaxBodies + 4];
};

// Compacts the world's contacts (candidate order), assigns levels; returns K and the top level.
template <bool kVelocityPass>
__device__ __forceinline__ int buildLevels(const DeviceState &S, int w, SolveGroupSmem &g, int lg, unsigned gmask,
                                           int gshift, int *maxLvlOut)
{
    const int nc = WI(S, WF_NUM_CAND, w);
    const int nb = WI(S, WF_NUM_BODIES, w);
    const uint32_t *cmeta = S.cmeta + size_t(w) * kMaxCand;
    int K = 0;
    for (int c0 = 0; c0 < nc; c0 += kSolveLanes) {
        const int c = c0 + lg;
        uint32_t m = c < nc ? cmeta[c] : 0u;
        const bool has = (m >> 16) != 0;
        const unsigned bal = (__ballot_sync(gmask, has) >> gshift) & ((1u << kSolveLanes) - 1u);
        if (has) {
            int k = K + __popc(bal & ((1u << lg) - 1u));
            g.cm[k] = (m & 0x00FFFFFFu) | (uint32_t(c) << 24);
        }
        K += __popc(bal);
    }
    for (int s = lg; s < nb; s += kSolveLanes) {
        uint32_t m = META(S, s, w);
        bool fixedBody = metaResp(m) == RESP_STATIC;
        if (!kVelocityPass && fixedBody) {
            Quat q = loadRot(S, s, w);
            fixedBody = q.w == 1.f && q.x == 0.f && q.y == 0.f && q.z == 0.f;
        }
        g.blvl[s] = fixedBody ? 0xFF : 0;
    }
    __syncwarp(gmask);
    for (int i = lg; i < kMaxLevels + 2; i += kSolveLanes) g.lvlStart[i] = 0;
    __syncwarp(gmask);
    int maxLvl = 0;
    if (lg == 0) {
        for (int k = 0; k < K; k++) {
            uint32_t cm = g.cm[k];
            int b1 = cm & 0xFF, b2 = (cm >> 8) & 0xFF;
            int l1 = g.blvl[b1], l2 = g.blvl[b2];
            int lv = min(max(l1 == 0xFF ? 0 : l1, l2 == 0xFF ? 0 : l2), 0xFE);
            g.clvl[k] = (uint8_t)lv;
            if (lv < kMaxLevels) g.lvlStart[lv + 1] += 1;      // histogram, turned into offsets below
            if (l1 != 0xFF) g.blvl[b1] = (uint8_t)(lv + 1);
            if (l2 != 0xFF) g.blvl[b2] = (uint8_t)(lv + 1);
            maxLvl = max(maxLvl, lv);
        }
        if (maxLvl < kMaxLevels)
            for (int l = 1; l <= maxLvl + 1; l++) g.lvlStart[l] += g.lvlStart[l - 1];
    }
    maxLvl = __shfl_sync(gmask, maxLvl, gshift);
    __syncwarp(gmask);
    if (maxLvl >= kMaxLevels) {        // pathological chain: signal the serial fallback
        *maxLvlOut = -1;
        return K;
    }
    // stable counting sort by level: lanes stay dense in every sweep
    for (int k = lg; k < K; k += kSolveLanes) {
        const int lv = g.clvl[k];
        int rank = g.lvlStart[lv];
        for (int j = 0; j < k; j++) rank += g.clvl[j] == lv;
        g.order[rank] = (uint8_t)k;
    }
    __syncwarp(gmask);
    *maxLvlOut = maxLvl;
    return K;
}

// joints in creation order (at most one Fixed joint per agent; rare), after every contact
__device__ inline void solveJoints(const DeviceState &S, int w, int lg)
{
    if (lg != 0) return;
    const int na = WI(S, WF_NUM_AGENTS, w);
    const int nDyn = WI(S, WF_NUM_DYN, w);
    int lastSeq = -1;
    for (int it = 0; it < na; it++) {
        int best = -1, bestSeq = 0x7FFFFFFF;
        for (int k = 0; k < na; k++) {
            uint32_t jm = JMETA(S, k, w);
            if (jm == 0xFFFFFFFFu) continue;
            int seq = int(jm >> 16);
            if (seq > lastSeq && seq < bestSeq) { bestSeq = seq; best = k; }
        }
        if (best < 0) break;
        lastSeq = bestSeq;
        handleFixedJoint(S, w, best, nDyn + best, int(JMETA(S, best, w) & 0xFF));
    }
}

// solvePositions
__global__ void __launch_bounds__(kSolveThreads) k_solve_positions(DeviceState S)
{
    __shared__ SolveGroupSmem sg[kSolveGroups];
    const int lane = threadIdx.x & 31;
    const int lg = lane % kSolveLanes;
    const int gshift = lane - lg;
    const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
    const int w = blockIdx.x * kSolveGroups + threadIdx.x / kSolveLanes;
    if (w >= S.p.numWorlds) return;
    SolveGroupSmem &g = sg[threadIdx.x / kSolveLanes];

    int maxLvl;
    const int K = buildLevels<false>(S, w, g, lg, gmask, gshift, &maxLvl);
    float *contacts = S.contact + size_t(w) * kMaxCand * kContactWords;
    for (int lv = 0; lv <= maxLvl; lv++) {
        const int end = K ? g.lvlStart[lv + 1] : 0;
        for (int i = g.lvlStart[lv] + lg; i < end; i += kSolveLanes) {
            const int k = g.order[i];
            uint32_t cm = g.cm[k];
            handleContact(S, w, cm, contacts + size_t(cm >> 24) * kContactWords,
                          g.blvl[cm & 0xFF] != 0xFF, g.blvl[(cm >> 8) & 0xFF] != 0xFF);
        }
        __syncwarp(gmask);
    }
    if (maxLvl < 0 && lg == 0)
        for (int k = 0; k < K; k++) handleContact(S, w, g.cm[k], contacts + size_t(g.cm[k] >> 24) * kContactWords, true, true);
    __syncwarp(gmask);

    solveJoints(S, w, lg);
}

// setVelocities (xpbd.cpp:738-779) for one body
__device__ __forceinline__ void setBodyVelocity(const DeviceState &S, int s, int w)
{
    const float h = substepH();
    const Pose cur = loadPose(S, REC_POSE, s, w), prev = loadPose(S, REC_PREV, s, w);
    Vec3 x = cur.x, xp = prev.x;
    Quat q = cur.q, qp = prev.q;
    Quat dq;
    if (q.w != qp.w || q.x != qp.x || q.y != qp.y || q.z != qp.z) dq = q * qinv(qp);
    else dq = {1.f, 0.f, 0.f, 0.f};
    Vec3 newOmega = (2.f / h) * v3(dq.x, dq.y, dq.z);
    storeVel6(S, REC_VEL, s, w, (x - xp) / h, dq.w > 0.f ? newOmega : -newOmega);
}

// setVelocities: one thread per (slot, world)
__global__ void k_set_velocities(DeviceState S)
{
    const int N = S.p.numWorlds;
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)kMaxBodies * N) return;
    int s = int(idx % kMaxBodies), w = int(idx / kMaxBodies);
    if (s >= WI(S, WF_NUM_BODIES, w)) return;
    setBodyVelocity(S, s, w);
}

__device__ __forceinline__ Vec3 relVel(Vec3 v1, Vec3 v2, Vec3 o1, Vec3 o2, Vec3 d1, Vec3 d2)
{
    return (v1 + cross(o1, d1)) - (v2 + cross(o2, d2));
}

__device__ inline void solveContactVelocity(const DeviceState &S, int w, uint32_t meta, const float *cdata,
                                            bool write1, bool write2)
{
    const int b1 = meta & 0xFF, b2 = (meta >> 8) & 0xFF;
    ContactView c = loadContact(cdata, meta);
    const float lambdaN = c.lambdaN;
    MassData m1 = massOf(META(S, b1, w)), m2 = massOf(META(S, b2, w));
    const float h = substepH();
    const float gMag = sqrtf(0.f * 0.f + 0.f * 0.f + 9.8f * 9.8f);
    const float restThreshold = 2.f * gMag * h;

    const Vel6 vel1 = loadVel6(S, REC_VEL, b1, w), vel2 = loadVel6(S, REC_VEL, b2, w);
    const Vel6 pv1 = loadVel6(S, REC_PREVEL, b1, w), pv2 = loadVel6(S, REC_PREVEL, b2, w);
    const Pose pre1 = loadPose(S, REC_PRE, b1, w), pre2 = loadPose(S, REC_PRE, b2, w);
    Vec3 v1 = vel1.v, o1 = vel1.w;
    Vec3 v2 = vel2.v, o2 = vel2.w;
    Quat q1 = loadRot(S, b1, w), q2 = loadRot(S, b2, w);
    Vec3 px1 = pre1.x, px2 = pre2.x;
    Quat pq1 = pre1.q, pq2 = pre2.q;
    float muD = 0.5f * (m1.muD + m2.muD);
    Vec3 n = c.normal;

    {
        Vec3 avg; float pen;
        if (avgContact(c, &avg, &pen)) return;
        Vec3 r1, r2;
        localContacts(px1, pq1, px2, pq2, avg, pen, n, &r1, &r2);
        Vec3 r1pre = rotate(pq1, r1), r2pre = rotate(pq2, r2);
        Vec3 vbar = relVel(pv1.v, pv2.v, pv1.w, pv2.w, r1pre, r2pre);
        float vnBar = dot(n, vbar);
        Vec3 r1w = rotate(q1, r1), r2w = rotate(q2, r2);
        Vec3 ta1 = cross(r1, rotate(qinv(q1), n)), ta2 = cross(r2, rotate(qinv(q2), n));

        // applyRestitutionVelocityUpdate
        Vec3 v = relVel(v1, v2, o1, o2, r1w, r2w);
        float vn = dot(n, v);
        float e = 0.3f;
        if (fabsf(vnBar) <= restThreshold) e = 0.f;
        float mag = fminf(-e * vnBar, 0) - vn;
        Vec3 ra1 = mulDiag(m1.inv_I, ta1), ra2 = mulDiag(m2.inv_I, ta2);
        float w1 = m1.inv_m + dot(ta1, ra1), w2 = m2.inv_m + dot(ta2, ra2);
        float invScale = 1.f / (w1 + w2);
        float imp = mag * invScale;
        if (imp != 0.f) {
            v1 += n * imp * m1.inv_m;
            v2 -= n * imp * m2.inv_m;
            o1 += rotate(q1, imp * ra1);
            o2 -= rotate(q2, imp * ra2);
        }
    }

    float penSum = 0.f;
    for (int i = 0; i < c.n; i++) penSum += c.depth[i];

    for (int i = 0; i < c.n; i++) {
        Vec3 r1, r2;
        localContacts(px1, pq1, px2, pq2, c.pts[i], c.depth[i], n, &r1, &r2);
        Vec3 r1w = rotate(q1, r1), r2w = rotate(q2, r2);
        float lambda = lambdaN * (c.depth[i] / penSum);

        // applyFrictionVelocityUpdate
        Vec3 v = relVel(v1, v2, o1, o2, r1w, r2w);
        float vn = dot(n, v);
        Vec3 vt = v - n * vn;
        float vtLen = length(vt);
        if (vtLen == 0.f) continue;
        Vec3 dw = vt / vtLen;
        Vec3 d1 = rotate(qinv(q1), dw), d2 = rotate(qinv(q2), dw);
        Vec3 fa1 = cross(r1, d1), fa2 = cross(r2, d2);
        Vec3 fr1 = mulDiag(m1.inv_I, fa1), fr2 = mulDiag(m2.inv_I, fa2);
        float w1 = m1.inv_m + dot(fa1, fr1), w2 = m2.inv_m + dot(fa2, fr2);
        float invScale = 1.f / (w1 + w2);
        float fmag = muD * fabsf(lambda) * invScale / h;
        float corrected = -fminf(fmag, vtLen);
        float imp = corrected * invScale;
        if (imp == 0.f) continue;
        v1 += dw * imp * m1.inv_m;
        v2 -= dw * imp * m2.inv_m;
        o1 += rotate(q1, imp * fr1);
        o2 -= rotate(q2, imp * fr2);
    }

    if (write1) storeVel6(S, REC_VEL, b1, w, v1, o1);
    if (write2) storeVel6(S, REC_VEL, b2, w, v2, o2);
}

// solveVelocities
__global__ void __launch_bounds__(kSolveThreads) k_solve_velocities(DeviceState S)
{
    __shared__ SolveGroupSmem sg[kSolveGroups];
    const int lane = threadIdx.x & 31;
    const int lg = lane % kSolveLanes;
    const int gshift = lane - lg;
    const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
    const int w = blockIdx.x * kSolveGroups + threadIdx.x / kSolveLanes;
    if (w >= S.p.numWorlds) return;
    SolveGroupSmem &g = sg[threadIdx.x / kSolveLanes];

    int maxLvl;
    const int K = buildLevels<true>(S, w, g, lg, gmask, gshift, &maxLvl);
    const float *contacts = S.contact + size_t(w) * kMaxCand * kContactWords;
    for (int lv = 0; lv <= maxLvl; lv++) {
        const int end = K ? g.lvlStart[lv + 1] : 0;
        for (int i = g.lvlStart[lv] + lg; i < end; i += kSolveLanes) {
            const int k = g.order[i];
            uint32_t cm = g.cm[k];
            solveContactVelocity(S, w, cm, contacts + size_t(cm >> 24) * kContactWords,
                                 g.blvl[cm & 0xFF] != 0xFF, g.blvl[(cm >> 8) & 0xFF] != 0xFF);
        }
        __syncwarp(gmask);
    }
    if (maxLvl < 0 && lg == 0)
        for (int k = 0; k < K; k++) solveContactVelocity(S, w, g.cm[k], contacts + size_t(g.cm[k] >> 24) * kContactWords, true, true);
    if (lg == 0) WI(S, WF_NUM_CONTACTS, w) = K;
}

// solvePositions + setVelocities + solveVelocities of one substep in ONE launch: the three stages
// of a world are consecutive and touch only that world, so the same 8-lane group runs them back
// to back.  The level schedule is built once, with the position-pass rule; it is also a valid
// schedule for the velocity pass, whose conflicts (contacts sharing a non-Static body) are a
// subset of the position pass's, so both passes reproduce the serial creation-order sweep.
__global__ void __launch_bounds__(kSolveThreads) k_solve(DeviceState S)
{
    __shared__ SolveGroupSmem sg[kSolveGroups];
    __shared__ unsigned long long s_static[kSolveGroups];
    const int lane = threadIdx.x & 31;
    const int lg = lane % kSolveLanes;
    const int gshift = lane - lg;
    const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
    const int grp = threadIdx.x / kSolveLanes;
    const int w = blockIdx.x * kSolveGroups + grp;
    if (w >= S.p.numWorlds) return;
    SolveGroupSmem &g = sg[grp];
    const int nb = WI(S, WF_NUM_BODIES, w);

    // Static bodies (never written by the velocity pass)
    {
        unsigned long long mine = 0ull;
        for (int s = lg; s < nb; s += kSolveLanes)
            if (metaResp(META(S, s, w)) == RESP_STATIC) mine |= 1ull << s;
#pragma unroll
        for (int o = kSolveLanes / 2; o > 0; o >>= 1) mine |= __shfl_xor_sync(gmask, mine, o);
        if (lg == 0) s_static[grp] = mine;
    }

    int maxLvl;
    const int K = buildLevels<false>(S, w, g, lg, gmask, gshift, &maxLvl);
    float *contacts = S.contact + size_t(w) * kMaxCand * kContactWords;
    // ---- positions ----------------------------------------------------------------------------
    for (int lv = 0; lv <= maxLvl; lv++) {
        const int end = K ? g.lvlStart[lv + 1] : 0;
        for (int i = g.lvlStart[lv] + lg; i < end; i += kSolveLanes) {
            const int k = g.order[i];
            uint32_t cm = g.cm[k];
            handleContact(S, w, cm, contacts + size_t(cm >> 24) * kContactWords,
                          g.blvl[cm & 0xFF] != 0xFF, g.blvl[(cm >> 8) & 0xFF] != 0xFF);
        }
        __syncwarp(gmask);
    }
    if (maxLvl < 0 && lg == 0)
        for (int k = 0; k < K; k++) handleContact(S, w, g.cm[k], contacts + size_t(g.cm[k] >> 24) * kContactWords, true, true);
    __syncwarp(gmask);
    solveJoints(S, w, lg);
    __syncwarp(gmask);
    // ---- velocities from the position change ----------------------------------------------------
    for (int s = lg; s < nb; s += kSolveLanes) setBodyVelocity(S, s, w);
    __syncwarp(gmask);
    // ---- velocity pass ----------------------------------------------------------------------------
    const unsigned long long st = s_static[grp];
    for (int lv = 0; lv <= maxLvl; lv++) {
        const int end = K ? g.lvlStart[lv + 1] : 0;
        for (int i = g.lvlStart[lv] + lg; i < end; i += kSolveLanes) {
            const int k = g.order[i];
            uint32_t cm = g.cm[k];
            solveContactVelocity(S, w, cm, contacts + size_t(cm >> 24) * kContactWords,
                                 !((st >> (cm & 0xFF)) & 1ull), !((st >> ((cm >> 8) & 0xFF)) & 1ull));
        }
        __syncwarp(gmask);
    }
    if (maxLvl < 0 && lg == 0)
        for (int k = 0; k < K; k++) solveContactVelocity(S, w, g.cm[k], contacts + size_t(g.cm[k] >> 24) * kContactWords, true, true);
    if (lg == 0) WI(S, WF_NUM_CONTACTS, w) = K;
}

}  // namespace hs
