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

// substepRigidBodies for one non-static body (xpbd.cpp:100-185): pose and velocity after the
// explicit integration step.  `vel` is the body's velocity record, `ft` its force / torque record.
__device__ __forceinline__ void integrateBody(uint32_t m, Vec3 &x, Quat &q, const Vel6 &vel, const Vel6 &ft, Vec3 &v, Vec3 &omega)
{
    v = vel.v;
    omega = vel.w;
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
}

// one body of one world, state in HBM
__device__ __forceinline__ void substepBody(const DeviceState &S, int s, int w)
{
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

    Vec3 v, omega;
    integrateBody(m, x, q, loadVel6(S, REC_VEL, s, w), loadVel6(S, REC_FORCE, s, w), v, omega);

    storePose(S, REC_POSE, s, w, x, q);
    storePose(S, REC_PRE, s, w, x, q);
    storeVel6(S, REC_PREVEL, s, w, v, omega);
}

// one thread per (slot, world)
__global__ void k_substep(DeviceState S)
{
    const int N = S.p.numWorlds;
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)kMaxBodies * N) return;
    int s = int(idx % kMaxBodies), w = int(idx / kMaxBodies);
    if (s >= WI(S, WF_NUM_BODIES, w)) return;
    substepBody(S, s, w);
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
    // fixed trip counts with a predicate (same operations in the same order): static indices keep the
    // contact in registers
    float maxPen = -FLT_MAX, sum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i < c.n) {
            if (c.depth[i] > maxPen) maxPen = c.depth[i];
            sum += c.depth[i];
        }
    }
    if (sum == 0.f) return true;
    Vec3 a = {0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (i < c.n) a += (c.depth[i] / sum) * c.pts[i];
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
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (i < c.n) penSum += c.depth[i];

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
__device__ __noinline__ void solveWorldGlobal(const DeviceState &S, int w, SolveGroupSmem &g, int lg, unsigned gmask, int gshift)
{
    const int nb = WI(S, WF_NUM_BODIES, w);

    // Static bodies (never written by the velocity pass)
    unsigned long long st = 0ull;
    {
        for (int s = lg; s < nb; s += kSolveLanes)
            if (metaResp(META(S, s, w)) == RESP_STATIC) st |= 1ull << s;
#pragma unroll
        for (int o = kSolveLanes / 2; o > 0; o >>= 1) st |= __shfl_xor_sync(gmask, st, o);
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

// HIDESEEK_SOLVER=global: every world through the global-memory path (A/B against k_solve below)
__global__ void __launch_bounds__(kSolveThreads) k_solve_global(DeviceState S)
{
    __shared__ SolveGroupSmem sg[kSolveGroups];
    const int lane = threadIdx.x & 31;
    const int lg = lane % kSolveLanes;
    const int gshift = lane - lg;
    const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
    const int grp = threadIdx.x / kSolveLanes;
    const int w = blockIdx.x * kSolveGroups + grp;
    if (w >= S.p.numWorlds) return;
    solveWorldGlobal(S, w, sg[grp], lg, gmask, gshift);
}

// =================================================================================================
// Staged solver: the default k_solve.
//
// One launch per substep runs solvePositions + setVelocities + solveVelocities of every world, like
// k_solve_global above, but
//   * which bodies are carrying / Static and the object ids of every candidate pair come from
//     k_broadphase (decided once per step, hs_trace.cuh); the group compacts the pairs that have a
//     contact in this substep into shared memory and levels them there;
//   * the bodies the sweep may write — the contiguous movable tail of the world, 160 B of solver
//     records per slot — are brought into shared memory with ONE bulk-async copy per world
//     (cp.async.bulk + mbarrier), the Gauss-Seidel sweeps read and write shared memory only, and
//     pose + velocity go back to HBM once at the end (walls are read straight from HBM: a Static
//     body with identity rotation is never written, its pre-solve / previous poses equal its pose
//     and its velocities are zero);
//   * a contact is solved by TWO lanes, one per body: every term of handleContact /
//     solveVelocitiesForContact that belongs to one body (lever arm, world point, generalised
//     inverse mass, the body's own update) is computed by that body's lane, the few scalars and
//     vectors that couple them cross with __shfl_xor.  Both lanes evaluate the reference's
//     expressions in the reference's operand order (a - b is formed as a + (-b), which is the same
//     float), so the result is bit-identical to the one-lane code above — tests/test_solver_ab_gpu.py
//     compares the two kernels bit for bit — at about half the instructions per contact.
// 8 lanes = 4 lane pairs per world, 4 worlds per warp, kStagedGroups (32) worlds per CTA, 2 CTAs per SM
// (measured: 8 / 16 / 32 worlds per CTA 1.592 / 1.587 / 1.577 ms per step).
// =================================================================================================
#ifndef HS_SOLVE_GROUPS
#define HS_SOLVE_GROUPS 32
#endif
#ifndef HS_SOLVE_MINB
#define HS_SOLVE_MINB 2
#endif
constexpr int kStagedGroups = HS_SOLVE_GROUPS;         // worlds per CTA of the staged solver
constexpr int kStagedThreads = kStagedGroups * kSolveLanes;
constexpr int kListCap = 48;          // contacts per world the staged path takes (more: global path)
constexpr int kStageRecs = 4;         // staged records per body: POSE, PREV (VEL once it is derived), PRE, PREVEL
enum StagedRec : int { SR_POSE = 0, SR_PREV = 1, SR_VEL = 1, SR_PRE = 2, SR_PREVEL = 3 };

// Per-world block of the staged solver's dynamic shared memory.  The body area holds `stageCap`
// bodies (9 boxes + 2 ramps + ground plane + the Manager's agent slots: hs_sim.cu sizes it), so the
// struct is a header in front of it.
struct alignas(16) StagedWorld {
    SchedEntry list[kListCap];                   // this substep's contacts in creation order
    unsigned long long bar;                      // mbarrier of the bulk copy
    uint8_t order[kListCap];                     // list indices sorted by (level, doubles before singles, creation order)
    uint8_t lvlStart[kListCap + 1];              // first position of level l (entry l + 1 = its end)
    uint8_t dblCount[kListCap];                  // contacts of level l whose two bodies are both carrying
    uint8_t sglCount[kListCap];                  // contacts of level l with one carrying body ("solo")
    uint8_t rank[kListCap];                      // rank of a contact inside its (level, kind)
    uint8_t blvl[kStageMax + 6];                 // next free level of staged body i
    float4 body[1][kStageRecs * 2];              // [stageCap]: records of slot first + i
};
__host__ __device__ constexpr int stagedWorldBytes(int stageCap)
{
    return (int(sizeof(StagedWorld)) + (stageCap - 1) * kStageRecs * 32 + 15) / 16 * 16;
}
constexpr int kMassBytes = OBJ_COUNT * 8 * 4;            // invMass, invInertia[3], muS, muD per object
__host__ __device__ constexpr int solveSmemBytes(int stageCap) { return kMassBytes + kStagedGroups * stagedWorldBytes(stageCap); }
static_assert(sizeof(SolveGroupSmem) <= sizeof(StagedWorld) + 11 * kStageRecs * 32, "the global path borrows the group's block");

__device__ __forceinline__ uint32_t smemAddr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr unsigned kFullMask = 0xFFFFFFFFu;

// Every shuffle of the staged solver runs in WARP-UNIFORM control flow with the full mask: a
// sub-warp mask that differs between the lanes of a warp (one mask per lane pair) makes each
// shuffle take the compiler's divergent MATCH / BRA.DIV path, which serialises the pairs and
// costs more than the arithmetic it connects.  So the code below is written as straight-line
// phases — per-lane work under `if (flag)` blocks that reconverge, then a full-mask exchange every
// lane takes part in — and idle lanes simply carry zeros through the exchanges.
__device__ __forceinline__ Vec3 shflXorV(Vec3 v)
{
    return {__shfl_xor_sync(kFullMask, v.x, 1), __shfl_xor_sync(kFullMask, v.y, 1), __shfl_xor_sync(kFullMask, v.z, 1)};
}
// v for the first body of a pair, -v for the second (x2 -= d is x2 + (-d) in IEEE arithmetic)
__device__ __forceinline__ Vec3 signedV(Vec3 v, uint32_t signBit)
{
    return {__uint_as_float(__float_as_uint(v.x) ^ signBit), __uint_as_float(__float_as_uint(v.y) ^ signBit),
            __uint_as_float(__float_as_uint(v.z) ^ signBit)};
}

__device__ __forceinline__ Pose poseOf(const float4 *rec)
{
    const float4 a = rec[0], b = rec[1];
    return {{a.x, a.y, a.z}, {a.w, b.x, b.y, b.z}};
}
__device__ __forceinline__ Vel6 vel6Of(const float4 *rec)
{
    const float4 a = rec[0], b = rec[1];
    return {{a.x, a.y, a.z}, {a.w, b.x, b.y}};
}
__device__ __forceinline__ void putPose(float4 *rec, Vec3 x, Quat q)
{
    rec[0] = make_float4(x.x, x.y, x.z, q.w);
    rec[1] = make_float4(q.x, q.y, q.z, 0.f);
}
__device__ __forceinline__ void putVel6(float4 *rec, Vec3 v, Vec3 o)
{
    rec[0] = make_float4(v.x, v.y, v.z, o.x);
    rec[1] = make_float4(o.y, o.z, 0.f, 0.f);
}

// one body's half of applyPositionalUpdate (xpbd.cpp:232-285)
__device__ __forceinline__ void applyPositionalSide(Vec3 &x, Quat &q, Vec3 ra, float inv_m, Vec3 n, float lambda,
                                                    uint32_t signBit)
{
    x = x + signedV((lambda * inv_m) * n, signBit);
    const float half = 0.5f * lambda;
    const Vec3 u = rotate(q, half * ra);
    const Quat dq = fromAngularVec(u) * q;
    q = q + Quat{__uint_as_float(__float_as_uint(dq.w) ^ signBit), __uint_as_float(__float_as_uint(dq.x) ^ signBit),
                 __uint_as_float(__float_as_uint(dq.y) ^ signBit), __uint_as_float(__float_as_uint(dq.z) ^ signBit)};
    q = qnormalize(q);
}

// What one lane of a pair needs to know about its contact and its body.
struct PairSide {
    int c, slot, npts;
    bool staged, carrying;
    float inv_m, muS, muD;          // muS / muD already averaged over the two bodies
    Vec3 inv_I;
    float4 *rec;                    // the body's solver records in shared memory (staged only)
};

__device__ __forceinline__ PairSide pairSide(const SchedEntry e, int side, StagedWorld &sw, int first,
                                             const float (*mass)[8])
{
    PairSide p;
    p.c = int(e.x & 0xFF);
    p.slot = int(e.x >> (16 + 8 * side)) & 0xFF;
    p.npts = int(e.y >> 12) & 7;
    const int obj = int(e.y >> (4 * side)) & 15, objRef = int(e.y) & 15, objAlt = int(e.y >> 4) & 15;
    p.carrying = (e.y >> (8 + side)) & 1;
    const bool isStatic = (e.y >> (10 + side)) & 1;
    p.staged = p.slot >= first;
    p.rec = sw.body[p.staged ? p.slot - first : 0];     // the body area extends past the declared [1]
    p.inv_m = isStatic ? 0.f : mass[obj][0];
    p.inv_I = isStatic ? Vec3{0.f, 0.f, 0.f} : Vec3{mass[obj][1], mass[obj][2], mass[obj][3]};
    p.muS = 0.5f * (mass[objRef][4] + mass[objAlt][4]);
    p.muD = 0.5f * (mass[objRef][5] + mass[objAlt][5]);
    return p;
}

// handleContact (xpbd.cpp:405-550 -> handleContactConstraint 321-403), one lane per body.
// Called by all 32 lanes; `mine` = this lane has a contact side to solve in this round.
// `solo`: the partner body is non-carrying (Static, identity rotation: never moved, zero inverse
// mass and inertia, previous = pre-solve = current pose), so this lane also evaluates the partner's
// contact point — by the very expressions the partner's lane would run — and the partner's other
// terms are exact zeros: generalised inverse mass 0 + dot(ta, 0 * ta), displacement p - p_hat of an
// unmoved point.
__device__ __forceinline__ void pairContactPositions(const DeviceState &S, int w, const PairSide &P, int side, bool mine,
                                                     bool solo, int slotO)
{
    const Vec3 zero = {0.f, 0.f, 0.f};
    float *row = S.contact + (size_t(w) * kMaxCand + P.c) * kContactWords;
    const uint32_t signBit = side ? 0x80000000u : 0u;
    bool act = mine;
    Vec3 n = zero, r = zero, x = zero, p = zero, pSolo = zero;
    Quat q = {1.f, 0.f, 0.f, 0.f};
    Pose prev = {zero, q};
    if (act) {
        const ContactView c = loadContact(row, uint32_t(P.npts) << 16);
        Vec3 avg; float pen;
        act = !avgContact(c, &avg, &pen);
        if (act) {
            Pose cur, pre;
            if (P.staged) { cur = poseOf(P.rec + 2 * SR_POSE); pre = poseOf(P.rec + 2 * SR_PRE); prev = poseOf(P.rec + 2 * SR_PREV); }
            else { cur = loadPose(S, REC_POSE, P.slot, w); pre = cur; prev = cur; }
            x = cur.x;
            q = cur.q;
            n = c.normal;
            // getLocalSpaceContacts: the second body's point lies `pen` below the first's along the normal
            const Vec3 contact = side ? avg - n * pen : avg;
            r = rotate(qinv(pre.q), contact - pre.x);
            p = rotate(q, r) + x;
            if (solo) {
                const Pose o = loadPose(S, REC_POSE, slotO, w);
                const Vec3 contactO = side ? avg : avg - n * pen;
                const Vec3 rO = rotate(qinv(o.q), contactO - o.x);
                pSolo = rotate(o.q, rO) + o.x;
            }
        }
    }
    Vec3 po = shflXorV(p);
    if (solo) po = pSolo;
    bool push = false;
    float d = 0.f, wgt = 0.f;
    Vec3 ra = zero;
    if (act) {
        d = dot(side ? po - p : p - po, n);
        push = d > 0;
        if (push) {
            const Vec3 nl = rotate(qinv(q), n);
            const Vec3 ta = cross(r, nl);
            ra = mulDiag(P.inv_I, ta);
            wgt = P.inv_m + dot(ta, ra);
        }
    }
    float wo = __shfl_xor_sync(kFullMask, wgt, 1);
    if (solo) wo = 0.f;
    float lambda_n = 0.f;
    Vec3 e = zero;
    if (push) {
        lambda_n = -d / ((side ? wo : wgt) + (side ? wgt : wo) + 0.f);
        applyPositionalSide(x, q, ra, P.inv_m, n, lambda_n, signBit);
        const Vec3 p_hat = rotate(prev.q, r) + prev.x;
        p = rotate(q, r) + x;
        e = p - p_hat;
    }
    Vec3 eo = shflXorV(e);
    if (solo) eo = zero;
    bool fric = false;
    float tmag = 0.f, wf = 0.f;
    Vec3 t = zero, fr = zero;
    if (push) {
        const Vec3 dp = side ? eo - e : e - eo;
        const Vec3 dpt = dp - dot(dp, n) * n;
        tmag = length(dpt);
        fric = tmag > 0.f;
        if (fric) {
            t = dpt / tmag;
            const Vec3 tl = rotate(qinv(q), t);
            const Vec3 fa = cross(r, tl);
            fr = mulDiag(P.inv_I, fa);
            wf = P.inv_m + dot(fa, fr);
        }
    }
    float wfo = __shfl_xor_sync(kFullMask, wf, 1);
    if (solo) wfo = 0.f;
    if (fric) {
        const float lambda_t = -tmag / ((side ? wfo : wf) + (side ? wf : wfo) + 0.f);
        if (lambda_t > lambda_n * P.muS) applyPositionalSide(x, q, fr, P.inv_m, t, lambda_t, signBit);
    }
    if (act) {
        if (side == 0 || solo) row[19] = lambda_n;
        if (P.carrying) putPose(P.rec + 2 * SR_POSE, x, q);
    }
}

// solveVelocitiesForContact (xpbd.cpp:781-1039), one lane per body; called by all 32 lanes.
// `solo`: the partner's velocities (current and pre-solve) are zero, so every term it would send is 0
__device__ __forceinline__ void pairContactVelocities(const DeviceState &S, int w, const PairSide &P, int side, bool mine,
                                                      bool solo)
{
    const Vec3 zero = {0.f, 0.f, 0.f};
    const float *row = S.contact + (size_t(w) * kMaxCand + P.c) * kContactWords;
    const uint32_t signBit = side ? 0x80000000u : 0u;
    const float h = substepH();
    const float gMag = sqrtf(0.f * 0.f + 0.f * 0.f + 9.8f * 9.8f);
    const float restThreshold = 2.f * gMag * h;

    bool act = mine;
    ContactView c;
    c.n = 0;
    Vec3 v = zero, o = zero, n = zero, ab = zero, a = zero, ra = zero;
    Pose pre = {zero, {1.f, 0.f, 0.f, 0.f}};
    Quat q = pre.q;
    float wgt = 0.f, lambdaN = 0.f;
    if (act) {
        c = loadContact(row, uint32_t(P.npts) << 16);
        lambdaN = c.lambdaN;
        Vec3 avg; float pen;
        act = !avgContact(c, &avg, &pen);
        if (act) {
            Vec3 pv = zero, po = zero;
            if (P.staged) {
                const Vel6 vel = vel6Of(P.rec + 2 * SR_VEL), pvel = vel6Of(P.rec + 2 * SR_PREVEL);
                v = vel.v; o = vel.w; pv = pvel.v; po = pvel.w;
                pre = poseOf(P.rec + 2 * SR_PRE);
                q = poseOf(P.rec + 2 * SR_POSE).q;
            } else {
                pre = loadPose(S, REC_POSE, P.slot, w);
                q = pre.q;
            }
            n = c.normal;
            const Vec3 contact = side ? avg - n * pen : avg;
            const Vec3 r = rotate(qinv(pre.q), contact - pre.x);
            const Vec3 rpre = rotate(pre.q, r);
            ab = pv + cross(po, rpre);
            const Vec3 rw = rotate(q, r);
            const Vec3 ta = cross(r, rotate(qinv(q), n));
            a = v + cross(o, rw);
            ra = mulDiag(P.inv_I, ta);
            wgt = P.inv_m + dot(ta, ra);
        }
    }
    {
        Vec3 abo = shflXorV(ab);
        Vec3 ao = shflXorV(a);
        float wo = __shfl_xor_sync(kFullMask, wgt, 1);
        if (solo) { abo = zero; ao = zero; wo = 0.f; }
        if (act) {
            // applyRestitutionVelocityUpdate
            const float vnBar = dot(n, side ? abo - ab : ab - abo);
            const float vn = dot(n, side ? ao - a : a - ao);
            float e = 0.3f;
            if (fabsf(vnBar) <= restThreshold) e = 0.f;
            const float mag = fminf(-e * vnBar, 0) - vn;
            const float invScale = 1.f / ((side ? wo : wgt) + (side ? wgt : wo));
            const float imp = mag * invScale;
            if (imp != 0.f) {
                v = v + signedV(n * imp * P.inv_m, signBit);
                o = o + signedV(rotate(q, imp * ra), signBit);
            }
        }
    }

    float penSum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (i < c.n) penSum += c.depth[i];

    // not unrolled: four copies of the friction update are 1100 more instructions in the solver's
    // hottest loop, and the instruction cache notices (1.457 -> 1.411 ms per step)
#pragma unroll 1
    for (int i = 0; i < 4; i++) {
        const bool pt = act && i < c.n;
        Vec3 r = zero, ai = zero;
        // point i comes from the contact row again (a 16-byte load that hits L1): indexing the register
        // copy with the loop counter would push the whole contact into local memory
        float4 p4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (pt) p4 = reinterpret_cast<const float4 *>(row)[i];
        const Vec3 cpt = {p4.x, p4.y, p4.z};
        const float cdepth = p4.w;
        if (pt) {
            const Vec3 contact = side ? cpt - n * cdepth : cpt;
            r = rotate(qinv(pre.q), contact - pre.x);
            const Vec3 rw = rotate(q, r);
            ai = v + cross(o, rw);
        }
        Vec3 aio = shflXorV(ai);
        if (solo) aio = zero;
        // applyFrictionVelocityUpdate
        bool go = false;
        float vtLen = 0.f, wi = 0.f;
        Vec3 dw = zero, fr = zero;
        if (pt) {
            const Vec3 rel = side ? aio - ai : ai - aio;
            const float vn = dot(n, rel);
            const Vec3 vt = rel - n * vn;
            vtLen = length(vt);
            go = vtLen != 0.f;
            if (go) {
                dw = vt / vtLen;
                const Vec3 dl = rotate(qinv(q), dw);
                const Vec3 fa = cross(r, dl);
                fr = mulDiag(P.inv_I, fa);
                wi = P.inv_m + dot(fa, fr);
            }
        }
        float wio = __shfl_xor_sync(kFullMask, wi, 1);
        if (solo) wio = 0.f;
        if (go) {
            const float lambda = lambdaN * (cdepth / penSum);
            const float invScale = 1.f / ((side ? wio : wi) + (side ? wi : wio));
            const float fmag = P.muD * fabsf(lambda) * invScale / h;
            const float corrected = -fminf(fmag, vtLen);
            const float imp = corrected * invScale;
            if (imp != 0.f) {
                v = v + signedV(dw * imp * P.inv_m, signBit);
                o = o + signedV(rotate(q, imp * fr), signBit);
            }
        }
    }

    if (act && P.carrying) putVel6(P.rec + 2 * SR_VEL, v, o);
}

// handleJointConstraint (Fixed) on the staged records; both bodies of a grab joint are movable.
__device__ inline void stagedFixedJoint(const DeviceState &S, int w, int k, float4 *r1, float4 *r2, const MassData &m1, const MassData &m2)
{
    const Pose jp1 = poseOf(r1 + 2 * SR_POSE), jp2 = poseOf(r2 + 2 * SR_POSE);
    Vec3 x1 = jp1.x, x2 = jp2.x;
    Quat q1 = jp1.q, q2 = jp2.q;
    Quat a1 = {JF(S, JF_A1W, k, w), JF(S, JF_A1X, k, w), JF(S, JF_A1Y, k, w), JF(S, JF_A1Z, k, w)};
    Quat a2 = {JF(S, JF_A2W, k, w), JF(S, JF_A2X, k, w), JF(S, JF_A2Y, k, w), JF(S, JF_A2Z, k, w)};
    Vec3 jr1 = {JF(S, JF_R1X, k, w), JF(S, JF_R1Y, k, w), JF(S, JF_R1Z, k, w)};
    Vec3 jr2 = {JF(S, JF_R2X, k, w), JF(S, JF_R2Y, k, w), JF(S, JF_R2Z, k, w)};
    float sep = JF(S, JF_SEP, k, w);
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
    putPose(r1 + 2 * SR_POSE, x1, q1);
    putPose(r2 + 2 * SR_POSE, x2, q2);
}

// Compacts the candidate pairs that have a contact in this substep into sw.list, in creation order
// and oriented (reference body first — the narrowphase may have swapped the pair).  Called by all
// 32 lanes (groups without work pass nc = 0).  Returns the number of contacts; more than kListCap
// means the list is incomplete.
__device__ __forceinline__ int fillContactList(const DeviceState &S, int w, StagedWorld &sw, int nc, int lg, int gshift)
{
    const SchedEntry *sched = S.sched + size_t(w) * kMaxCand;
    const uint32_t *cmeta = S.cmeta + size_t(w) * kMaxCand;
    int K = 0, pos = 0;
    while (__any_sync(kFullMask, pos < nc)) {
        const int i = pos + lg;
        SchedEntry e = {0u, 0u};
        uint32_t m = 0u;
        if (i < nc) { e = sched[i]; m = cmeta[i]; }
        const bool has = (m >> 16) != 0;
        const unsigned bal = (__ballot_sync(kFullMask, has) >> gshift) & ((1u << kSolveLanes) - 1u);
        const int k = K + __popc(bal & ((1u << lg) - 1u));
        if (has && k < kListCap) {
            const uint32_t a = (e.x >> 16) & 0xFF, b = e.x >> 24;
            if ((m & 0xFF) != a) {          // the narrowphase chose b as the reference body: swap the sides
                e.x = (e.x & 0xFFFFu) | (b << 16) | (a << 24);
                e.y = ((e.y >> 4) & 0xFu) | ((e.y & 0xFu) << 4) | (((e.y >> 9) & 1u) << 8) | (((e.y >> 8) & 1u) << 9) |
                      (((e.y >> 11) & 1u) << 10) | (((e.y >> 10) & 1u) << 11);
            }
            e.y = (e.y & 0xFFFu) | (((m >> 16) & 7u) << 12);
            sw.list[k] = e;
        }
        K += __popc(bal);
        pos += kSolveLanes;
    }
    return K;
}

// Dependency levels of the listed contacts (see the comment above buildLevels): level(c) = the
// next free level of its carrying bodies; a stable counting sort by level gives sw.order.  The scan
// is inherently serial (lane 0 of the group; K is ~17), the sort is lane-parallel.
__device__ __forceinline__ void levelContactList(StagedWorld &sw, int K, int first, int count, int lg)
{
    for (int i = lg; i < kListCap; i += kSolveLanes) { sw.dblCount[i] = 0; sw.sglCount[i] = 0; }
    for (int i = lg; i < count; i += kSolveLanes) sw.blvl[i] = 0;
    __syncwarp();
    if (lg == 0 && K > 0) {
        int maxLvl = 0;
        for (int k = 0; k < K; k++) {
            const SchedEntry e = sw.list[k];
            const bool ca = (e.y >> 8) & 1, cb = (e.y >> 9) & 1;
            const int ia = int((e.x >> 16) & 0xFF) - first, ib = int(e.x >> 24) - first;     // carrying => staged
            const int la = ca ? sw.blvl[ia] : 0, lb = cb ? sw.blvl[ib] : 0;
            const int lv = max(la, lb);                 // <= k < kListCap
            const bool solo = ca != cb;                 // the partner is never written: one lane does it all
            sw.list[k].x = (e.x & 0xFFFF00FFu) | (uint32_t(lv) << 8);
            sw.list[k].y = (e.y & 0x7FFFu) | (solo ? 0x8000u : 0u);
            sw.rank[k] = solo ? sw.sglCount[lv]++ : sw.dblCount[lv]++;
            if (ca) sw.blvl[ia] = uint8_t(lv + 1);
            if (cb) sw.blvl[ib] = uint8_t(lv + 1);
            maxLvl = max(maxLvl, lv);
        }
        int run = 0;
        for (int l = 0; l <= maxLvl; l++) { sw.lvlStart[l] = uint8_t(run); run += sw.dblCount[l] + sw.sglCount[l]; }
        sw.lvlStart[maxLvl + 1] = uint8_t(run);
    }
    __syncwarp();
    for (int k = lg; k < K; k += kSolveLanes) {
        const SchedEntry e = sw.list[k];
        const int lv = (e.x >> 8) & 0xFF;
        sw.order[sw.lvlStart[lv] + ((e.y & 0x8000u) ? sw.dblCount[lv] : 0) + sw.rank[k]] = uint8_t(k);
    }
    __syncwarp();
}

// One pass (positions or velocities) over the levelled contacts, one LEVEL at a time, levels in
// order; inside a level (disjoint carrying bodies, any order) a round hands the group's eight
// lanes out: a contact between two carrying bodies takes a lane pair (one lane per body), a "solo"
// contact — the partner is a wall, the ground plane or another Static body with identity rotation,
// ~95 % of all contacts — takes ONE lane, which supplies the partner's few terms itself.  Called
// by all 32 lanes; a round ends with a warp barrier, which orders the shared-memory body updates
// of one level before the next.
template <bool kVelocities>
__device__ __forceinline__ int sweepList(const DeviceState &S, int w, StagedWorld &sw, int K, int first, int lg,
                                         const float (*mass)[8])
{
    int i = 0, rounds = 0;
    while (__any_sync(kFullMask, i < K)) {
        const bool active = i < K;
        rounds += active;
        int nd = 0, ns = 0;
        if (active) {
            const int lv = int(sw.list[sw.order[i]].x >> 8) & 0xFF;
            const int dblEnd = sw.lvlStart[lv] + sw.dblCount[lv], lvlEnd = sw.lvlStart[lv + 1];
            nd = min(max(dblEnd - i, 0), kSolveLanes / 2);
            if (i + nd >= dblEnd) ns = min(lvlEnd - (i + nd), kSolveLanes - 2 * nd);
        }
        const bool solo = lg >= 2 * nd;
        const int j = solo ? i + nd + (lg - 2 * nd) : i + (lg >> 1);
        const bool mine = solo ? (lg - 2 * nd) < ns : true;
        const SchedEntry ej = mine ? sw.list[min(int(sw.order[min(j, kListCap - 1)]), kListCap - 1)] : SchedEntry{0u, 0u};
        // a solo lane works on the carrying body's side
        const int side = solo ? (((ej.y >> 8) & 1u) ? 0 : 1) : (lg & 1);
        const PairSide P = pairSide(ej, side, sw, first, mass);
        const int slotO = int(ej.x >> (16 + 8 * (1 - side))) & 0xFF;
        if (kVelocities) pairContactVelocities(S, w, P, side, mine, solo);
        else pairContactPositions(S, w, P, side, mine, solo, slotO);
        i += nd + ns;
        __syncwarp();
    }
    return rounds;          // rounds this group's world needed (the warp runs the maximum of its four)
}

__global__ void __launch_bounds__(kStagedThreads, HS_SOLVE_MINB) k_solve(const __grid_constant__ DeviceState S, int stageCap, int integrateNext, int orderIn)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    float (*mass)[8] = reinterpret_cast<float (*)[8]>(smemRaw);
    const int lane = threadIdx.x & 31;
    const int lg = lane % kSolveLanes;
    const int gshift = lane - lg;
    const int grp = threadIdx.x / kSolveLanes;
    const int wRaw = blockIdx.x * kStagedGroups + grp;
    const bool live = wRaw < S.p.numWorlds;
    // ---- which world this group solves ----------------------------------------------------------
    // A warp runs its four worlds' sweeps in lockstep, i.e. for as many rounds as the slowest of
    // them needs (3 or 4 for most worlds: 4.25 on average for four neighbours, 3.6 for four alike).
    // Worlds are independent, so ANY assignment of worlds to groups gives the same results; each
    // launch therefore files its worlds by the rounds they took (8 buckets, arrival order inside),
    // and the next substep's launch hands them out bucket by bucket, slowest first.  Four buffers
    // rotate with the substep index (orderIn): launch i reads buffer i, fills i + 1 and clears the
    // counters of i + 2.  Counters that do not add up to this view's world count (first launch, or
    // hs_sim.cu cleared them because the caller's view of the shard changed — step graph: world
    // slices, hs_step_profiled: one range) or orderIn < 0: identity order.
    int w = live ? wRaw : 0;                   // idle groups stay in the warp's control flow
    const int orderOut = orderIn >= 0 ? (orderIn + 1) % kSolveOrderBufs : -1;
    if (orderIn >= 0) {
        const int *cnt = S.solveCnt + orderIn * kSolveOrderBuckets;
        int total = 0, bucket = -1, base = 0;
#pragma unroll
        for (int b = kSolveOrderBuckets - 1; b >= 0; b--) {
            const int c = cnt[b];
            if (bucket < 0 && wRaw < total + c) { bucket = b; base = total; }
            total += c;
        }
        if (live && total == S.p.numWorlds && bucket >= 0)
            w = int(S.solveOrder[size_t(orderIn * kSolveOrderBuckets + bucket) * S.p.stride + (wRaw - base)]);
        if (blockIdx.x == 0 && threadIdx.x < kSolveOrderBuckets)
            S.solveCnt[((orderIn + 2) % kSolveOrderBufs) * kSolveOrderBuckets + threadIdx.x] = 0;
    }
    StagedWorld &sw = *reinterpret_cast<StagedWorld *>(smemRaw + kMassBytes + grp * stagedWorldBytes(stageCap));

    int first = 0, count = -1, nc = 0;
    if (live) {
        first = WI(S, WF_STAGE_FIRST, w);
        count = WI(S, WF_STAGE_COUNT, w);       // < 0: range longer than any Manager's body area
        nc = WI(S, WF_NUM_CAND, w);
    }
    const bool work = live && count >= 0 && count <= stageCap;
    if (!work) nc = 0;
    const uint32_t bar = smemAddr(&sw.bar);
    if (work && lg == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < OBJ_COUNT; i += kStagedThreads) {
        const ObjectData &O = c_assets.obj[i];
        mass[i][0] = O.invMass; mass[i][1] = O.invInertia[0]; mass[i][2] = O.invInertia[1];
        mass[i][3] = O.invInertia[2]; mass[i][4] = O.muS; mass[i][5] = O.muD;
    }
    __syncthreads();
    if (live && !work) {                      // staging range too long (never in generated levels)
        const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
        solveWorldGlobal(S, w, *reinterpret_cast<SolveGroupSmem *>(&sw), lg, gmask, gshift);
        if (integrateNext) {
            __syncwarp(gmask);
            for (int s = lg, nb = WI(S, WF_NUM_BODIES, w); s < nb; s += kSolveLanes) substepBody(S, s, w);
        }
    }
    __syncwarp();

    // ---- stage the movable tail of the world: one bulk-async copy ---------------------------------
    if (work && lg == 0 && count > 0) {
        static_assert(kStageRecs == kSolverRecs && REC_POSE == SR_POSE && REC_PREV == SR_PREV && REC_PRE == SR_PRE &&
                      REC_PREVEL == SR_PREVEL, "the staged block is a byte copy of the solver records");
        const uint32_t bytes = uint32_t(count) * kSolverRecs * kRecWords * 4u;
        const float *src = S.body + (size_t(w) * kMaxBodies + first) * kBodyWords;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smemAddr(&sw.body[0][0])), "l"(src), "r"(bytes), "r"(bar) : "memory");
    }
    // ... while it is in flight, gather and level this substep's contacts
    int K = fillContactList(S, w, sw, nc, lg, gshift);
    const bool tooMany = K > kListCap;              // never seen; the global path has no such limit
    const int total = K;
    if (tooMany) K = 0;
    levelContactList(sw, K, first, max(count, 0), lg);
    if (work && count > 0) {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
    }
    __syncwarp();
    if (tooMany) {
        const unsigned gmask = ((1u << kSolveLanes) - 1u) << gshift;
        solveWorldGlobal(S, w, *reinterpret_cast<SolveGroupSmem *>(&sw), lg, gmask, gshift);
        if (integrateNext) {
            __syncwarp(gmask);
            for (int s = lg, nb = WI(S, WF_NUM_BODIES, w); s < nb; s += kSolveLanes) substepBody(S, s, w);
        }
    }
    __syncwarp();
    const bool staged = work && !tooMany;

    // ---- positions ----------------------------------------------------------------------------
    const int rounds = sweepList<false>(S, w, sw, K, first, lg, mass);
    // joints in creation order (at most one Fixed joint per agent), after every contact
    if (staged && lg == 0) {
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
            const int b1 = nDyn + best, b2 = int(JMETA(S, best, w) & 0xFF);
            stagedFixedJoint(S, w, best, sw.body[b1 - first], sw.body[b2 - first], massOf(META(S, b1, w)), massOf(META(S, b2, w)));
        }
    }
    __syncwarp();
    // ---- velocities from the position change (setVelocities, xpbd.cpp:738-779) -----------------------
    for (int j = lg; staged && j < count; j += kSolveLanes) {
        const float h = substepH();
        const Pose cur = poseOf(sw.body[j] + 2 * SR_POSE), prev = poseOf(sw.body[j] + 2 * SR_PREV);
        Quat dq;
        if (cur.q.w != prev.q.w || cur.q.x != prev.q.x || cur.q.y != prev.q.y || cur.q.z != prev.q.z) dq = cur.q * qinv(prev.q);
        else dq = {1.f, 0.f, 0.f, 0.f};
        const Vec3 newOmega = (2.f / h) * v3(dq.x, dq.y, dq.z);
        // the previous pose is not needed past this point: the velocity takes its place
        putVel6(sw.body[j] + 2 * SR_VEL, (cur.x - prev.x) / h, dq.w > 0.f ? newOmega : -newOmega);
    }
    __syncwarp();
    // ---- velocity pass ----------------------------------------------------------------------------
    sweepList<true>(S, w, sw, K, first, lg, mass);
    // ---- pose + velocity back to HBM ----------------------------------------------------------------
    if (staged && !integrateNext) {
        for (int i = lg; i < count * 4; i += kSolveLanes) {
            const int j = i >> 2, q = i & 3;             // two float4 of POSE, then two of VEL
            float4 *dst = recPtr(S, q < 2 ? REC_POSE : REC_VEL, first + j, w);
            dst[q & 1] = sw.body[j][q < 2 ? 2 * SR_POSE + q : 2 * SR_VEL + (q - 2)];
        }
    }
    // ---- ... or, when another substep follows, its substepRigidBodies right here ------------------------
    // The next substep starts by integrating every body (k_substep).  For the staged bodies pose and
    // velocity are in shared memory at this point, so the integration runs on them and all five
    // records go out once.  Bodies outside the staged range are Static with identity rotation: no
    // solver writes them, so the PREV / PRE / PREVEL records the step's first k_substep gave them
    // are still what this substep's would write.
    if (staged && integrateNext) {
        __syncwarp();
        for (int j = lg; j < count; j += kSolveLanes) {
            const int s = first + j;
            const uint32_t m = META(S, s, w);
            const Pose cur = poseOf(sw.body[j] + 2 * SR_POSE);
            const Vel6 vel = vel6Of(sw.body[j] + 2 * SR_VEL);
            Vec3 x = cur.x;
            Quat q = cur.q;
            storePose(S, REC_PREV, s, w, x, q);
            storeVel6(S, REC_VEL, s, w, vel.v, vel.w);
            if (metaResp(m) == RESP_STATIC) {
                storePose(S, REC_POSE, s, w, x, q);
                storePose(S, REC_PRE, s, w, x, q);
                storeVel6(S, REC_PREVEL, s, w, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f});
            } else {
                Vec3 v, omega;
                integrateBody(m, x, q, vel, loadVel6(S, REC_FORCE, s, w), v, omega);
                storePose(S, REC_POSE, s, w, x, q);
                storePose(S, REC_PRE, s, w, x, q);
                storeVel6(S, REC_PREVEL, s, w, v, omega);
            }
        }
    }
    if (staged && lg == 0) WI(S, WF_NUM_CONTACTS, w) = total;
    if (orderOut >= 0 && live && lg == 0) {
        const int b = staged ? min(rounds, kSolveOrderBuckets - 1) : kSolveOrderBuckets - 1;
        const int pos = atomicAdd(&S.solveCnt[orderOut * kSolveOrderBuckets + b], 1);
        S.solveOrder[size_t(orderOut * kSolveOrderBuckets + b) * S.p.stride + pos] = uint32_t(w);
    }
}

}  // namespace hs
