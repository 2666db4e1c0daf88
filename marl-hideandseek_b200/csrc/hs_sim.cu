// C-ABI entry points and the per-step launch sequence (include/hideseek/c_api.h).
//
// One translation unit so every kernel shares the __constant__ asset table.
// The launch order is the node order of the reference's Step task graph
// (src/sim.cpp:1140-1313).  There is no CPU fallback: every
// entry point fails loudly if CUDA is unavailable.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <hideseek/c_api.h>
#include <hideseek/debug_dump.h>

#include "hs_state.cuh"
#include "hs_rng.cuh"
#include "hs_levelgen.cuh"
#include "hs_bvh.cuh"
#include "hs_trace.cuh"
#include "hs_agents.cuh"
#include "hs_obs.cuh"
#include "hs_rays.cuh"
#include "hs_narrowphase.cuh"
#include "hs_solver.cuh"
#include "hs_checkpoint.cuh"
#define HS_VIEWS_HOST_ONLY
#include "hs_views.cuh"

namespace hs {
const char *buildAssetTable(AssetTable &T, const char *dataDir);      // hs_assets.cpp
const char *decodePngRGBA(const char *path, int *width, int *height, std::vector<uint8_t> &rgba);   // hs_view_assets.cpp
cudaError_t launchAgentViews(const DeviceState &S, const ViewParams &P, unsigned views, cudaStream_t stream);   // hs_views.cu
}

using namespace hs;

static thread_local std::string g_err;

#define HS_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
    g_err = std::string(#x) + ": " + cudaGetErrorString(e_); return 1; } } while (0)

struct hs_sim {
    hs_config cfg;
    DeviceState S;
    std::vector<void *> allocs;
    hs_world_dump *dumpDev = nullptr;
    int lastLaunches = 0;
    // who filled the k_solve world-order buffers last: 0 nobody, 1 the step graph (per-slice views),
    // 2 hs_step_profiled (one whole-shard view); a change of view clears the counters (= identity order)
    int orderView = 0;
    cudaGraphExec_t stepGraph = nullptr;
    cudaStream_t graphStream = nullptr;
    // The step graph runs the shard as `slices` independent world ranges on parallel branches:
    // the stage kernels are latency-bound and leave the SMs partly idle (tails of short waves,
    // low occupancy of the ordered solver), so branches of different slices fill each other's gaps.
    bool raysNearestFirst = true;
    int slices = 1, branches = 1;
    cudaStream_t sliceStream[8] = {};
    cudaEvent_t forkEvent = nullptr, joinEvent[8] = {};
    // per branch: two side streams so observations / debug positions run beside the ray kernel
    cudaStream_t auxStream[8][2] = {};
    cudaEvent_t auxFork[8] = {}, auxJoin[8][2] = {};
    bool obsFork = true;
    bool inited = false;
    float cosThreshold = 0.f;
    // agent camera views (hs_views.cuh); off unless hs_enable_views was called
    bool viewsEnabled = false;
    ViewParams view{};
    ViewMeshTable *viewMeshHost = nullptr;
    ViewTexTable *viewTexHost = nullptr;
    cudaArray_t viewTexArray[kViewTextures] = {};
};

template <typename T>
static int devAlloc(hs_sim *sim, T **out, size_t count)
{
    void *p = nullptr;
    size_t bytes = count * sizeof(T);
    if (bytes == 0) bytes = 16;
    HS_CUDA(cudaMalloc(&p, bytes));
    HS_CUDA(cudaMemset(p, 0, bytes));
    sim->allocs.push_back(p);
    *out = (T *)p;
    return 0;
}

#define HS_TRY(x) do { if ((x) != 0) return 1; } while (0)

extern "C" const char *hs_last_error(void) { return g_err.c_str(); }

static int createBody(hs_sim *sim, const hs_config *cfg, int A);

extern "C" int hs_create(const hs_config *cfg, hs_sim **out)
{
    if (!cfg || !out) { g_err = "hs_create: null argument"; return 1; }
    int A = cfg->max_hiders + cfg->max_seekers;
    if (cfg->num_worlds <= 0 || cfg->num_worlds >= (1 << 24) || A <= 0 || A > kMaxAgents || cfg->max_hiders > 3 || cfg->max_seekers > 3 ||
        cfg->min_hiders > cfg->max_hiders || cfg->min_seekers > cfg->max_seekers) {
        g_err = "hs_create: invalid configuration";
        return 1;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_err = "hs_create: no CUDA device (this library has no CPU fallback)";
        return 1;
    }
    HS_CUDA(cudaSetDevice(cfg->gpu_id));

    hs_sim *sim = new hs_sim();
    sim->cfg = *cfg;
    if (createBody(sim, cfg, A) != 0) {      // g_err is set; release whatever was allocated so far
        hs_destroy(sim);
        return 1;
    }
    *out = sim;
    return 0;
}

static int createBody(hs_sim *sim, const hs_config *cfg, int A)
{
    DeviceState &S = sim->S;
    const size_t N = (size_t)cfg->num_worlds;
    RandKey ik = initKey(cfg->rand_seed);
    S.p = SimParams{cfg->num_worlds, cfg->num_worlds, cfg->world_base, A, cfg->sim_flags, ik.a, ik.b,
                    cfg->min_hiders, cfg->max_hiders, cfg->min_seekers, cfg->max_seekers};

    HS_TRY(devAlloc(sim, &S.body, size_t(kBodyWords) * kMaxBodies * N));
    HS_TRY(devAlloc(sim, &S.bodyB, size_t(kBodyAuxWords) * kMaxBodies * N));
    HS_TRY(devAlloc(sim, &S.meta, size_t(kMaxBodies) * N));
    HS_TRY(devAlloc(sim, &S.wi, size_t(WF_COUNT) * N));
    HS_TRY(devAlloc(sim, &S.wb, size_t(WB_COUNT) * N));
    HS_TRY(devAlloc(sim, &S.joint, size_t(JF_COUNT) * kMaxAgents * N));
    HS_TRY(devAlloc(sim, &S.jmeta, size_t(kMaxAgents) * N));
    HS_TRY(devAlloc(sim, &S.bvh, size_t(kBvhWords) * N));
    HS_TRY(devAlloc(sim, &S.dfsLeaf, size_t(kDfsStride) * N));
    HS_TRY(devAlloc(sim, &S.resetList, N));
    HS_TRY(devAlloc(sim, &S.resetCount, N));
    HS_TRY(devAlloc(sim, &S.solveOrder, size_t(kSolveOrderBufs) * kSolveOrderBuckets * N));
    HS_TRY(devAlloc(sim, &S.solveCnt, N + 64));
    HS_TRY(devAlloc(sim, &S.cand, size_t(kMaxCand) * N));
    HS_TRY(devAlloc(sim, &S.contact, size_t(kMaxCand) * kContactWords * N));
    HS_TRY(devAlloc(sim, &S.cmeta, size_t(kMaxCand) * N));
    HS_TRY(devAlloc(sim, &S.sched, size_t(kMaxCand) * N));
    HS_TRY(devAlloc(sim, &S.overflow, 4));
    HS_TRY(devAlloc(sim, &S.dbgLidarHit, N * A * 30));

    const size_t R = N * A;
    HS_TRY(devAlloc(sim, &S.exReset, N));
    HS_TRY(devAlloc(sim, &S.exPrepCounter, R));
    HS_TRY(devAlloc(sim, &S.exAction, R * 5));
    HS_TRY(devAlloc(sim, &S.exSelfObs, R * 13));
    HS_TRY(devAlloc(sim, &S.exSelfType, R));
    HS_TRY(devAlloc(sim, &S.exSelfMask, R));
    HS_TRY(devAlloc(sim, &S.exAgentObs, R * 5 * 14));
    HS_TRY(devAlloc(sim, &S.exBoxObs, R * 9 * 17));
    HS_TRY(devAlloc(sim, &S.exRampObs, R * 2 * 14));
    HS_TRY(devAlloc(sim, &S.exAgentVis, R * 5));
    HS_TRY(devAlloc(sim, &S.exBoxVis, R * 9));
    HS_TRY(devAlloc(sim, &S.exRampVis, R * 2));
    HS_TRY(devAlloc(sim, &S.exLidar, R * 30));
    HS_TRY(devAlloc(sim, &S.exSeed, R * 2));
    HS_TRY(devAlloc(sim, &S.exReward, R));
    HS_TRY(devAlloc(sim, &S.exDone, R));
    HS_TRY(devAlloc(sim, &S.exGlobalPos, N * 17 * 2));
    HS_TRY(devAlloc(sim, &S.exPolicy, R));
    HS_TRY(devAlloc(sim, &S.exEpisodeResult, N * 2));
    HS_TRY(devAlloc(sim, &S.exCkptCtrl, N));
    HS_TRY(devAlloc(sim, &S.exCkpt, N * kCheckpointBytes));
    HS_TRY(devAlloc(sim, &sim->dumpDev, 1));

    AssetTable T;
    {
        // collision hulls from the reference's data directory when one is named, else built in
        const char *err = buildAssetTable(T, getenv("HIDESEEK_DATA_DIR"));
        if (err[0]) { g_err = std::string("hs_create: ") + err; return 1; }
    }
    HS_CUDA(cudaMemcpyToSymbol(c_assets, &T, sizeof(T)));

    {
        // lidar ray directions (src/sim.cpp:728-731), evaluated once with the shared detmath
        float dirs[30][2];
        const float pi = 3.14159265358979323846f;
        for (int i = 0; i < 30; i++) {
            float theta = 2.f * pi * (float(i) / float(30)) + pi / 2.f;
            float sn, cs;
            hs_det_sincosf(theta, &sn, &cs);
            dirs[i][0] = cs; dirs[i][1] = sn;
        }
        HS_CUDA(cudaMemcpyToSymbol(c_lidarDir, dirs, sizeof(dirs)));
        HS_CUDA(cudaFuncSetAttribute(k_rays<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RaySmemT<false>)));
        HS_CUDA(cudaFuncSetAttribute(k_rays<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RaySmemT<true>)));
        HS_CUDA(cudaFuncSetAttribute(k_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, solveSmemBytes(kStageMax)));
        HS_CUDA(cudaFuncSetAttribute(k_narrowphase, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NpSmem)));
        if (const char *e = getenv("HIDESEEK_RAY_ORDER")) sim->raysNearestFirst = strcmp(e, "reference") != 0;
    }
    HS_CUDA(cudaStreamCreateWithFlags(&sim->graphStream, cudaStreamNonBlocking));
    {
        int want = 2;
        if (const char *e = getenv("HIDESEEK_STEP_SLICES")) want = atoi(e);
        if (want < 1) want = 1;
        if (want > 64) want = 64;
        while (want > 1 && cfg->num_worlds / want < 512) want--;     // keep slices worth a launch
        sim->slices = want;
        int br = 2;
        if (const char *e = getenv("HIDESEEK_STEP_BRANCHES")) br = atoi(e);
        if (br < 1) br = 1;
        if (br > 8) br = 8;
        if (br > want) br = want;
        sim->branches = br;
        HS_CUDA(cudaEventCreateWithFlags(&sim->forkEvent, cudaEventDisableTiming));
        for (int b = 1; b < br; b++) {
            HS_CUDA(cudaStreamCreateWithFlags(&sim->sliceStream[b], cudaStreamNonBlocking));
            HS_CUDA(cudaEventCreateWithFlags(&sim->joinEvent[b], cudaEventDisableTiming));
        }
        if (const char *e = getenv("HIDESEEK_OBS_FORK")) sim->obsFork = e[0] != '0';
        for (int b = 0; b < br && sim->obsFork; b++) {
            HS_CUDA(cudaEventCreateWithFlags(&sim->auxFork[b], cudaEventDisableTiming));
            for (int i = 0; i < 2; i++) {
                HS_CUDA(cudaStreamCreateWithFlags(&sim->auxStream[b][i], cudaStreamNonBlocking));
                HS_CUDA(cudaEventCreateWithFlags(&sim->auxJoin[b][i], cudaEventDisableTiming));
            }
        }
    }
    {
        // cosf(toRadians(135.f / 2.f)) as the reference's compiler folds it (src/sim.cpp:582,767)
        float mult = 3.14159265358979323846f / 180.f;
        float ang = mult * (135.f / 2.f);
        sim->cosThreshold = (float)cos((double)ang);
    }
    return 0;
}

extern "C" void hs_destroy(hs_sim *sim)
{
    if (!sim) return;
    cudaSetDevice(sim->cfg.gpu_id);
    cudaDeviceSynchronize();
    if (sim->stepGraph) cudaGraphExecDestroy(sim->stepGraph);
    if (sim->graphStream) cudaStreamDestroy(sim->graphStream);
    if (sim->forkEvent) cudaEventDestroy(sim->forkEvent);
    for (int b = 0; b < 8; b++) {
        if (sim->auxFork[b]) cudaEventDestroy(sim->auxFork[b]);
        for (int i = 0; i < 2; i++) {
            if (sim->auxStream[b][i]) cudaStreamDestroy(sim->auxStream[b][i]);
            if (sim->auxJoin[b][i]) cudaEventDestroy(sim->auxJoin[b][i]);
        }
    }
    for (int b = 1; b < 8; b++) {
        if (sim->sliceStream[b]) cudaStreamDestroy(sim->sliceStream[b]);
        if (sim->joinEvent[b]) cudaEventDestroy(sim->joinEvent[b]);
    }
    for (void *p : sim->allocs) cudaFree(p);
    if (sim->viewTexHost)
        for (int i = 0; i < kViewTextures; i++) {
            if (sim->viewTexHost->tex[i]) cudaDestroyTextureObject((cudaTextureObject_t)sim->viewTexHost->tex[i]);
            if (sim->viewTexArray[i]) cudaFreeArray(sim->viewTexArray[i]);
        }
    delete sim->viewMeshHost;
    delete sim->viewTexHost;
    delete sim;
}

// ---- construction-time world state (Sim::Sim, src/sim.cpp:1346-1408) ----------
__global__ void k_world_init(DeviceState S)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= S.p.numWorlds) return;
    WI(S, WF_ID_FRESH, w) = S.p.agentsPerWorld;   // interface entities took the first ids
    WI(S, WF_HIDER_REWARD, w) = __float_as_int(1.f);
    for (int k = 0; k < kMaxAgents; k++) {
        JMETA(S, k, w) = 0xFFFFFFFFu;
        WB(S, WB_IFACE_SLOT + k, w) = 0xFF;
    }
    S.exReset[w] = 1;                              // resetLevel = 1 in the constructor
    BVH(S, w)[2] = 1;                              // BVH ctor: force_rebuild_ = true
}

__global__ void k_dump_world(DeviceState S, int w, hs_world_dump *out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    hs_world_dump &d = *out;
    d.cur_episode_step = WI(S, WF_EPISODE_STEP, w);
    d.cur_world_episode = WI(S, WF_WORLD_EPISODE, w);
    d.episode_key[0] = WI(S, WF_EP_KEY_A, w); d.episode_key[1] = WI(S, WF_EP_KEY_B, w);
    d.rng_key[0] = WI(S, WF_RNG_A, w); d.rng_key[1] = WI(S, WF_RNG_B, w);
    d.rng_count = WI(S, WF_RNG_COUNT, w);
    d.num_hiders = WI(S, WF_NUM_HIDERS, w); d.num_seekers = WI(S, WF_NUM_SEEKERS, w);
    d.num_agents = WI(S, WF_NUM_AGENTS, w);
    d.num_dyn = WI(S, WF_NUM_DYN, w); d.num_bodies = WI(S, WF_NUM_BODIES, w);
    d.num_boxes = WI(S, WF_NUM_BOXES, w); d.num_ramps = WI(S, WF_NUM_RAMPS, w);
    d.hider_team_reward = __int_as_float(WI(S, WF_HIDER_REWARD, w));
    d.seekers_first = WI(S, WF_SEEKERS_FIRST, w);
    d.running_scores[0] = WI(S, WF_SCORE0, w); d.running_scores[1] = WI(S, WF_SCORE1, w);
    d.num_cand = WI(S, WF_NUM_CAND, w);
    for (int i = 0; i < 9; i++) d.box_slot[i] = WB(S, WB_BOX_SLOT + i, w);
    for (int i = 0; i < 2; i++) d.ramp_slot[i] = WB(S, WB_RAMP_SLOT + i, w);
    for (int i = 0; i < 3; i++) { d.hider_slot[i] = WB(S, WB_HIDER_SLOT + i, w); d.seeker_slot[i] = WB(S, WB_SEEKER_SLOT + i, w); }
    for (int i = 0; i < kMaxAgents; i++) {
        int s = WB(S, WB_IFACE_SLOT + i, w);
        d.iface_slot[i] = s == 0xFF ? -1 : s;
        d.iface_type[i] = WB(S, WB_IFACE_TYPE + i, w);
    }
    for (int s = 0; s < kMaxBodies; s++) {
        hs_dump_body &b = d.bodies[s];
        uint32_t m = META(S, s, w);
        b.eid = metaEid(m); b.obj = metaObj(m); b.response = metaResp(m);
        b.owner = metaOwner(m); b.leaf = metaLeaf(m);
        for (int k = 0; k < 3; k++) b.pos[k] = BF(S, BF_PX + k, s, w);
        for (int k = 0; k < 4; k++) b.rot[k] = BF(S, BF_QW + k, s, w);
        for (int k = 0; k < 3; k++) b.scale[k] = BF(S, BF_SX + k, s, w);
        for (int k = 0; k < 6; k++) b.vel[k] = BF(S, BF_VX + k, s, w);
        for (int k = 0; k < 3; k++) b.ext_force[k] = BF(S, BF_FX + k, s, w);
        for (int k = 0; k < 3; k++) b.ext_torque[k] = BF(S, BF_TX + k, s, w);
    }
    for (int k = 0; k < kMaxAgents; k++) {
        hs_dump_joint &j = d.joints[k];
        uint32_t jm = JMETA(S, k, w);
        j.active = jm != 0xFFFFFFFFu;
        j.other_slot = jm & 0xFF; j.eid = (jm >> 8) & 0xFF; j.seq = jm >> 16;
        for (int c = 0; c < 4; c++) { j.attach1[c] = JF(S, JF_A1W + c, k, w); j.attach2[c] = JF(S, JF_A2W + c, k, w); }
        j.separation = JF(S, JF_SEP, k, w);
        for (int c = 0; c < 3; c++) { j.r1[c] = JF(S, JF_R1X + c, k, w); j.r2[c] = JF(S, JF_R2X + c, k, w); }
    }
    const uint32_t *blob = BVH(S, w);
    d.bvh_num_nodes = blob[0]; d.bvh_num_leaves = blob[1]; d.bvh_force_rebuild = blob[2];
    for (int n = 0; n < kMaxNodes; n++) {
        const float *nf = nodeF(blob, n);
        const int32_t *ni = nodeI(blob, n);
        for (int k = 0; k < 24; k++) d.nodes[n].bounds[k] = nf[k];
        for (int k = 0; k < 4; k++) d.nodes[n].children[k] = ni[kNodeChild + k];
        d.nodes[n].parent = ni[kNodeParent];
    }
    for (int l = 0; l < kMaxBodies; l++) {
        const float *L = leafF(blob, l);
        for (int k = 0; k < 6; k++) d.leaf_aabbs[l][k] = L[k];
        d.leaf_parents[l] = blob[kBvhLeafOff + l * kLeafWords + kLeafParent];
        d.leaf_slot[l] = blob[kBvhLeafOff + l * kLeafWords + kLeafInfo] & 0xFF;
    }
    for (int c = 0; c < kMaxCand; c++) {
        uint16_t v = S.cand[size_t(w) * kMaxCand + c];
        d.cand_a[c] = v & 0xFF; d.cand_b[c] = v >> 8;
        d.contact_meta[c] = S.cmeta[size_t(w) * kMaxCand + c];
        for (int k = 0; k < kContactWords; k++)
            d.contacts[c][k] = S.contact[(size_t(w) * kMaxCand + c) * kContactWords + k];
    }
}


// ---- bulk parity dump (hs_compact_world, include/hideseek/debug_dump.h) ---------------------------
// One CTA per world; every section is a lane-strided copy.
__global__ void k_dump_compact(DeviceState S, int w0, hs_compact_world *out)
{
    const int w = w0 + blockIdx.x, t = threadIdx.x, T = blockDim.x;
    hs_compact_world &d = out[blockIdx.x];
    const int nb = WI(S, WF_NUM_BODIES, w), nDyn = WI(S, WF_NUM_DYN, w), A = S.p.agentsPerWorld;
    if (t == 0) {
        d.cur_episode_step = WI(S, WF_EPISODE_STEP, w);
        d.cur_world_episode = WI(S, WF_WORLD_EPISODE, w);
        d.rng_key[0] = WI(S, WF_RNG_A, w); d.rng_key[1] = WI(S, WF_RNG_B, w);
        d.rng_count = WI(S, WF_RNG_COUNT, w);
        d.num_hiders = WI(S, WF_NUM_HIDERS, w); d.num_seekers = WI(S, WF_NUM_SEEKERS, w);
        d.num_boxes = WI(S, WF_NUM_BOXES, w); d.num_ramps = WI(S, WF_NUM_RAMPS, w);
        d.hider_team_reward = __int_as_float(WI(S, WF_HIDER_REWARD, w));
        d.seekers_first = WI(S, WF_SEEKERS_FIRST, w);
        d.running_scores[0] = WI(S, WF_SCORE0, w); d.running_scores[1] = WI(S, WF_SCORE1, w);
        d.num_bodies = nb; d.num_dyn = nDyn;
        const uint32_t *blob = BVH(S, w);
        d.bvh_num_nodes = blob[0]; d.bvh_num_leaves = blob[1];
        // candidates in emission order; contacts of the last substep in creation order
        const int nc = WI(S, WF_NUM_CAND, w);
        d.num_cand = nc;
        int K = 0;
        for (int c = 0; c < nc; c++) {
            const uint32_t m = S.cmeta[size_t(w) * kMaxCand + c];
            if ((m >> 16) == 0) continue;
            if (K < HS_COMPACT_MAX_CONTACTS) {
                d.contact_ref[K] = m & 0xFF; d.contact_alt[K] = (m >> 8) & 0xFF; d.contact_npts[K] = (m >> 16) & 0xFF;
                const float *row = S.contact + (size_t(w) * kMaxCand + c) * kContactWords;
                for (int p = 0; p < 4; p++) for (int k = 0; k < 4; k++) d.contact_pts[K][p][k] = row[p * 4 + k];
                for (int k = 0; k < 3; k++) d.contact_normal[K][k] = row[16 + k];
                d.contact_lambda[K] = row[19];
            }
            K++;
        }
        d.num_contacts = K;
        for (int k = 0; k < HS_DUMP_MAX_AGENTS; k++) {
            const uint32_t jm = k < nb - nDyn ? JMETA(S, k, w) : 0xFFFFFFFFu;
            d.joint_other[k] = jm == 0xFFFFFFFFu ? -1 : int8_t(jm & 0xFF);
            d.joint_sep[k] = jm == 0xFFFFFFFFu ? 0.f : JF(S, JF_SEP, k, w);
        }
    }
    for (int s = t; s < HS_DUMP_MAX_BODIES; s += T) {
        const bool live = s < nb;
        const uint32_t m = live ? META(S, s, w) : 0u;
        d.obj[s] = live ? metaObj(m) : 0; d.response[s] = live ? metaResp(m) : 0;
        d.owner[s] = live ? metaOwner(m) : 0; d.leaf[s] = live ? metaLeaf(m) : 0;
        int rank = 0;
        for (int o = 0; live && o < nb; o++) rank += metaEid(META(S, o, w)) < metaEid(m);
        d.eid_rank[s] = uint8_t(rank);
        for (int k = 0; k < 7; k++) d.pose[s][k] = live ? BF(S, BF_PX + k, s, w) : 0.f;
        for (int k = 0; k < 3; k++) d.scale[s][k] = live ? BF(S, BF_SX + k, s, w) : 0.f;
        for (int k = 0; k < 6; k++) d.vel[s][k] = live ? BF(S, BF_VX + k, s, w) : 0.f;
        for (int k = 0; k < 6; k++) d.force[s][k] = live ? BF(S, BF_FX + k, s, w) : 0.f;
    }
    const uint32_t *blob = BVH(S, w);
    for (int i = t; i < kMaxNodes * 24; i += T) d.node_bounds[i / 24][i % 24] = nodeF(blob, i / 24)[i % 24];
    for (int i = t; i < kMaxNodes * 4; i += T) d.node_children[i / 4][i % 4] = nodeI(blob, i / 4)[kNodeChild + i % 4];
    for (int i = t; i < kMaxNodes; i += T) d.node_parent[i] = nodeI(blob, i)[kNodeParent];
    for (int i = t; i < kMaxBodies * 6; i += T) d.leaf_aabb[i / 6][i % 6] = leafF(blob, i / 6)[i % 6];
    for (int i = t; i < kMaxBodies; i += T) d.leaf_parent[i] = blob[kBvhLeafOff + i * kLeafWords + kLeafParent];
    for (int c = t; c < kMaxCand; c += T) {
        const uint16_t v = S.cand[size_t(w) * kMaxCand + c];
        d.cand[c][0] = v & 0xFF; d.cand[c][1] = v >> 8;
    }
    for (int i = t; i < HS_DUMP_MAX_AGENTS * 30; i += T) {
        const int k = i / 30;
        int8_t v = -2;
        if (k < A && WB(S, WB_IFACE_SLOT + k, w) != 0xFF) v = S.dbgLidarHit[(size_t(w) * A + k) * 30 + i % 30];
        d.lidar_hit[k][i % 30] = v;
    }
}

extern "C" int hs_compact_dump_bytes(void) { return (int)sizeof(hs_compact_world); }

extern "C" int hs_dump_compact(hs_sim *sim, int world0, int count, void *out, int64_t out_bytes)
{
    if (!sim || !out || world0 < 0 || count <= 0 || world0 + count > sim->cfg.num_worlds ||
        out_bytes < (int64_t)count * (int64_t)sizeof(hs_compact_world)) {
        g_err = "hs_dump_compact: bad argument";
        return 1;
    }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaDeviceSynchronize());
    hs_compact_world *dev = nullptr;
    const size_t bytes = size_t(count) * sizeof(hs_compact_world);
    HS_CUDA(cudaMalloc(&dev, bytes));
    cudaError_t e = cudaMemset(dev, 0, bytes);
    if (e == cudaSuccess) {
        k_dump_compact<<<count, 64>>>(sim->S, world0, dev);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, dev, bytes, cudaMemcpyDeviceToHost);
    cudaFree(dev);
    HS_CUDA(e);
    return 0;
}

// ---- allocation-free variant of hs_dump_compact ------
// hs_dump_compact stages the records in a cudaMalloc'ed scratch buffer, which callers under CUDA
// graph capture, tight serving loops or harnesses that forbid allocation cannot afford.  This
// variant stages them in a file-scope device array instead: kNoAllocDumpMax records (1.5 MB) that
// every GPU loading the library holds until exit.  The array's NAME is a host-side stand-in in
// host code, not a device address, so the device address is fetched once per device, with that
// device current, by cudaGetSymbolAddress and that pointer is what the kernel is given.
// Every call shares the array and it keeps the last call's contents: the part a call uses is
// cleared first (k_dump_compact does not write every byte of a record), and calls that overlap on
// the GPU — two host threads, two simulators on one device — race on it.
constexpr int kNoAllocDumpMax = 64;
static __device__ __align__(256) unsigned char g_noAllocDumpScratch[kNoAllocDumpMax * sizeof(hs_compact_world)];
constexpr int kNoAllocMaxDevices = 64;
static void *g_noAllocDumpAddr[kNoAllocMaxDevices] = {};

extern "C" int hs_dump_compact_noalloc_max(void) { return kNoAllocDumpMax; }

extern "C" int hs_dump_compact_noalloc(hs_sim *sim, int world0, int count, void *out, int64_t out_bytes)
{
    if (!sim || !out || world0 < 0 || count <= 0 || world0 + count > sim->cfg.num_worlds ||
        out_bytes < (int64_t)count * (int64_t)sizeof(hs_compact_world)) {
        g_err = "hs_dump_compact_noalloc: bad argument";
        return 1;
    }
    if (count > kNoAllocDumpMax) {
        g_err = "hs_dump_compact_noalloc: count " + std::to_string(count) + " exceeds the static scratch: at most " +
                std::to_string(kNoAllocDumpMax) + " worlds per call (use hs_dump_compact, or several calls)";
        return 1;
    }
    const int dev = sim->cfg.gpu_id;
    if (dev < 0 || dev >= kNoAllocMaxDevices) { g_err = "hs_dump_compact_noalloc: device index out of range"; return 1; }
    HS_CUDA(cudaSetDevice(dev));
    if (!g_noAllocDumpAddr[dev]) {           // once per device, with that device current
        void *p = nullptr;
        HS_CUDA(cudaGetSymbolAddress(&p, g_noAllocDumpScratch));
        g_noAllocDumpAddr[dev] = p;
    }
    hs_compact_world *scratch = (hs_compact_world *)g_noAllocDumpAddr[dev];
    const size_t bytes = size_t(count) * sizeof(hs_compact_world);
    HS_CUDA(cudaDeviceSynchronize());
    HS_CUDA(cudaMemsetAsync(scratch, 0, bytes, nullptr));
    k_dump_compact<<<count, 64>>>(sim->S, world0, scratch);
    HS_CUDA(cudaGetLastError());
    HS_CUDA(cudaMemcpy(out, scratch, bytes, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int hs_set_debug(hs_sim *sim, unsigned flags)
{
    if (!sim) { g_err = "hs_set_debug: null sim"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaDeviceSynchronize());
    sim->S.p.debugFlags = flags;
    if (sim->stepGraph) {                   // kernel arguments are baked into the graph: recapture
        cudaGraphExecDestroy(sim->stepGraph);
        sim->stepGraph = nullptr;
    }
    return 0;
}

extern "C" int hs_world_dump_bytes(void) { return (int)sizeof(hs_world_dump); }

// host-only: the collision asset table hs_create builds for `data_dir` (NULL / "" = built-in hulls)
extern "C" int hs_asset_table_bytes(void) { return (int)sizeof(AssetTable); }

extern "C" int hs_dump_assets(const char *data_dir, void *out, int out_bytes)
{
    if (!out || out_bytes != (int)sizeof(AssetTable)) { g_err = "hs_dump_assets: bad buffer"; return 1; }
    const char *err = buildAssetTable(*(AssetTable *)out, data_dir);
    if (err[0]) { g_err = std::string("hs_dump_assets: ") + err; return 1; }
    return 0;
}

extern "C" int hs_dump_world(hs_sim *sim, int world, void *out, int out_bytes)
{
    if (!sim || world < 0 || world >= sim->cfg.num_worlds || out_bytes < (int)sizeof(hs_world_dump)) {
        g_err = "hs_dump_world: bad argument";
        return 1;
    }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaDeviceSynchronize());
    k_dump_world<<<1, 1>>>(sim->S, world, sim->dumpDev);
    HS_CUDA(cudaGetLastError());
    HS_CUDA(cudaMemcpy(out, sim->dumpDev, sizeof(hs_world_dump), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int hs_overflow_count(hs_sim *sim, uint32_t *out)
{
    if (!sim || !out) { g_err = "hs_overflow_count: null argument"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaMemcpy(out, sim->S.overflow, 4, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int hs_last_step_launches(hs_sim *sim) { return sim ? sim->lastLaunches : 0; }

// ---- launch sequence ------------------------------------------------------------
namespace {

// HIDESEEK_FUSED_SOLVE=0 launches solvePositions / setVelocities / solveVelocities as three kernels
// (the pre-fusion pipeline, kept for A/B measurements); default is the fused k_solve.
static bool fusedSolveDefault()
{
    const char *e = getenv("HIDESEEK_FUSED_SOLVE");
    return !(e && e[0] == '0');
}
// HIDESEEK_SOLVER=global: the fused solver that works out of global memory with its own level
// schedule (k_solve_global) instead of the staged lane-pair solver (k_solve); bit-identical results.
static bool stagedSolveDefault()
{
    const char *e = getenv("HIDESEEK_SOLVER");
    return !(e && strcmp(e, "global") == 0);
}

// HIDESEEK_HEAD_REFIT=1 runs the BVH update at the head of a step like the reference's task graph
// does (src/sim.cpp:1307-1313 -> setupBroadphaseTasks).  It is skipped by default because it cannot
// change anything: every path that moves a body or its velocity outside a step (hs_init, checkpoint
// loads) and the step itself end with a BVH update of the worlds they touched, movementSystem only
// writes forces, leaf boxes are a pure function of pose / velocity / scale and the refit only grows
// bounds that already contain those boxes (tests/test_step_variants_gpu.py: bit-identical BVH).
// HIDESEEK_FUSED_SUBSTEP=0 launches k_substep in front of every substep; by default only the first
// substep of a step does, the staged k_solve integrates the next substep's bodies at its tail.
static bool fusedSubstepDefault()
{
    const char *e = getenv("HIDESEEK_FUSED_SUBSTEP");
    return !(e && e[0] == '0');
}
// HIDESEEK_SOLVE_ORDER=0: k_solve takes its worlds in index order instead of grouping worlds whose
// contact sweeps need the same number of rounds (same results either way).
static bool orderedSolveDefault()
{
    const char *e = getenv("HIDESEEK_SOLVE_ORDER");
    return !(e && e[0] == '0');
}
static bool headRefitDefault()
{
    const char *e = getenv("HIDESEEK_HEAD_REFIT");
    return e && e[0] == '1';
}

struct Launcher {
    DeviceState S;
    cudaStream_t st;
    int count = 0;
    bool fusedSolve = fusedSolveDefault();
    bool stagedSolve = stagedSolveDefault();
    bool headRefit = headRefitDefault();
    bool fusedSubstep = fusedSubstepDefault();
    bool orderedSolve = false;          // set by hs_step (graph capture) and hs_step_profiled, which own the order buffers
    // graph capture only: side streams + events for the independent observation-stage kernels
    cudaStream_t aux[2] = {nullptr, nullptr};
    cudaEvent_t auxFork = nullptr, auxJoin[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> *events = nullptr;     // profiling mode: one event after each launch
    std::vector<const char *> *names = nullptr;
    void mark(const char *name)
    {
        count++;
        if (events) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            cudaEventRecord(e, st);
            events->push_back(e);
            names->push_back(name);
        }
    }
    int N() const { return S.p.numWorlds; }
    // thread per world
    dim3 gWorld(int block = 128) const { return dim3((N() + block - 1) / block); }
    // thread per (slot, world)
    dim3 gBody(int block = 256) const { return dim3(unsigned(((long long)kMaxBodies * N() + block - 1) / block)); }
    // warp per world
    dim3 gWarpWorld(int block = 128) const { return dim3(unsigned(((long long)N() * 32 + block - 1) / block)); }
    // thread per agent row
    dim3 gAgent(int block = 128) const { return dim3((N() * S.p.agentsPerWorld + block - 1) / block); }
    // warp per agent row
    dim3 gWarpAgent(int block = 128) const { return dim3(unsigned(((long long)N() * S.p.agentsPerWorld * 32 + block - 1) / block)); }
};

// setupBVHTasks (broadphase.cpp:995-1017): update leaves, rebuild if forced, refit.  Leaf update
// and refit are one kernel; the rebuild of freshly reset worlds follows it (see k_leaf_refit).
void launchBroadphaseSetup(Launcher &L, bool mayRebuild, bool onlyIfReset = false)
{
    k_leaf_refit<<<L.gBody(), 256, 0, L.st>>>(L.S, onlyIfReset); L.mark("k_leaf_refit");
    if (mayRebuild) { k_bvh_rebuild<<<L.gWorld(64), 64, 0, L.st>>>(L.S); L.mark("k_bvh_rebuild"); }
}

// observationsTasks (src/sim.cpp:1232-1293)
// collectObservations, visibility + lidar and the debug positions only read the world state and
// write disjoint exports, so inside the step graph they run as three parallel nodes.
void launchObservations(Launcher &L, float cosThr, bool nearestFirst)
{
    const bool fork = L.aux[0] != nullptr;
    cudaStream_t so = fork ? L.aux[0] : L.st, sd = fork ? L.aux[1] : L.st;
    if (fork) {
        cudaEventRecord(L.auxFork, L.st);
        cudaStreamWaitEvent(L.aux[0], L.auxFork, 0);
        cudaStreamWaitEvent(L.aux[1], L.auxFork, 0);
    }
    {
        const long long R = (long long)L.N() * L.S.p.agentsPerWorld;
        const long long nb = obsBlocks(R) + obsBlocks(R * kMaxBoxes) + obsBlocks(R * kMaxRamps) + obsBlocks(R * kMaxAgents);
        k_observations<<<unsigned(nb), kObsThreads, 0, so>>>(L.S);
    } L.mark("k_observations");
    if (nearestFirst) k_rays<true><<<L.N(), kRayThreads, sizeof(RaySmemT<false>), L.st>>>(L.S, cosThr);
    else k_rays<false><<<L.N(), kRayThreads, sizeof(RaySmemT<true>), L.st>>>(L.S, cosThr);
    L.mark("k_rays");
    k_debug_positions<<<L.gWorld(), 128, 0, sd>>>(L.S); L.mark("k_debug_positions");
    if (fork) {
        for (int i = 0; i < 2; i++) {
            cudaEventRecord(L.auxJoin[i], L.aux[i]);
            cudaStreamWaitEvent(L.st, L.auxJoin[i], 0);
        }
    }
}

void launchReset(Launcher &L)
{
    cudaMemsetAsync(L.S.resetCount, 0, sizeof(int), L.st);
    k_reset<<<L.gWorld(128), 128, 0, L.st>>>(L.S); L.mark("k_reset");
    // sparse mode serves at most N/4 worlds with kResetSpread threads each; dense mode needs N threads
    k_reset_worlds<<<dim3((L.N() * (kResetSpread / 4) + 63) / 64), 64, 0, L.st>>>(L.S); L.mark("k_reset_worlds");
}

// Step task graph (src/sim.cpp:1307-1313), node order of SURVEY.md §3.3
void launchStep(Launcher &L, float cosThr, bool nearestFirst)
{
    k_move<<<L.gAgent(), 128, 0, L.st>>>(L.S); L.mark("k_move");
    if (L.headRefit) launchBroadphaseSetup(L, false);       // a rebuild is only ever pending right after a reset
    {
        const unsigned nb = unsigned((L.N() + kActWarps * (32 / kActLanes) - 1) / (kActWarps * (32 / kActLanes)));
        if (nearestFirst) k_action<true><<<nb, kActWarps * 32, 0, L.st>>>(L.S);
        else k_action<false><<<nb, kActWarps * 32, 0, L.st>>>(L.S);
        L.mark("k_action");
    }
    k_broadphase<<<(L.N() + kBroadWarps - 1) / kBroadWarps, kBroadWarps * 32, 0, L.st>>>(L.S); L.mark("k_broadphase");
    const bool tailSubstep = L.fusedSolve && L.stagedSolve && L.fusedSubstep;
    for (int i = 0; i < kSubsteps; i++) {
        if (i == 0 || !tailSubstep) { k_substep<<<L.gBody(), 256, 0, L.st>>>(L.S); L.mark("k_substep"); }
        k_narrowphase<<<(L.N() + kNpWorlds - 1) / kNpWorlds, kNpThreads, sizeof(NpSmem), L.st>>>(L.S); L.mark("k_narrowphase");
        if (L.fusedSolve && L.stagedSolve) {
            // body area of the staged solver: 9 boxes + 2 ramps + ground plane + this Manager's agent slots
            int stageCap = kMaxBoxes + kMaxRamps + 1 + L.S.p.agentsPerWorld;
            // test knob: a smaller body area sends the worlds that do not fit down the kernel's
            // global-memory path (tests/test_step_variants_gpu.py), which generated levels never take
            if (const char *e = getenv("HIDESEEK_STAGE_CAP")) stageCap = std::max(1, std::min(stageCap, atoi(e)));
            k_solve<<<(L.N() + kStagedGroups - 1) / kStagedGroups, kStagedThreads, solveSmemBytes(stageCap), L.st>>>(
                L.S, stageCap, tailSubstep && i + 1 < kSubsteps ? 1 : 0, L.orderedSolve ? i % kSolveOrderBufs : -1);
            L.mark("k_solve");
        } else if (L.fusedSolve) {
            k_solve_global<<<(L.N() + kSolveGroups - 1) / kSolveGroups, kSolveThreads, 0, L.st>>>(L.S); L.mark("k_solve_global");
        } else {
            k_solve_positions<<<(L.N() + kSolveGroups - 1) / kSolveGroups, kSolveThreads, 0, L.st>>>(L.S); L.mark("k_solve_positions");
            k_set_velocities<<<L.gBody(), 256, 0, L.st>>>(L.S); L.mark("k_set_velocities");
            k_solve_velocities<<<(L.N() + kSolveGroups - 1) / kSolveGroups, kSolveThreads, 0, L.st>>>(L.S); L.mark("k_solve_velocities");
        }
    }
    launchBroadphaseSetup(L, false);
    if (L.S.p.simFlags & FLAG_ZERO_AGENT_VEL) {
        k_zero_agent_vel<<<dim3((L.N() * kMaxAgents + 127) / 128), 128, 0, L.st>>>(L.S); L.mark("k_zero_agent_vel");
    }
    if (nearestFirst) k_rewards<true><<<(L.N() + kRewardWarps * kRewardWorldsPerWarp - 1) / (kRewardWarps * kRewardWorldsPerWarp), kRewardWarps * 32, 0, L.st>>>(L.S, cosThr);
    else k_rewards<false><<<(L.N() + kRewardWarps * kRewardWorldsPerWarp - 1) / (kRewardWarps * kRewardWorldsPerWarp), kRewardWarps * 32, 0, L.st>>>(L.S, cosThr);
    L.mark("k_rewards");
    launchReset(L);
    launchBroadphaseSetup(L, true, (L.S.p.simFlags & FLAG_ZERO_AGENT_VEL) == 0);
    launchObservations(L, cosThr, nearestFirst);
}

}  // namespace

extern "C" int hs_init(hs_sim *sim, void *stream)
{
    if (!sim) { g_err = "hs_init: null sim"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    Launcher L{sim->S, (cudaStream_t)stream};
    k_world_init<<<L.gWorld(), 128, 0, L.st>>>(L.S);
    // Init task graph (src/sim.cpp:1295-1305): reset -> broadphase setup -> observations
    launchReset(L);
    launchBroadphaseSetup(L, true);
    launchObservations(L, sim->cosThreshold, sim->raysNearestFirst);
    HS_CUDA(cudaGetLastError());
    sim->inited = true;
    return 0;
}

// ---- agent camera views (batch renderer outputs, src/mgr.cpp:876-880, 1241-1263) ---------------
extern "C" int hs_enable_views(hs_sim *sim, int width, int height, const char *data_dir)
{
    if (!sim) { g_err = "hs_enable_views: null sim"; return 1; }
    if (sim->viewsEnabled) { g_err = "hs_enable_views: already enabled"; return 1; }
    if (width <= 0 || height <= 0 || width > 4096 || height > 4096) { g_err = "hs_enable_views: invalid view size"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    if (!data_dir || !data_dir[0]) data_dir = getenv("HIDESEEK_DATA_DIR");
    if (!sim->viewMeshHost) sim->viewMeshHost = new ViewMeshTable();
    if (!sim->viewTexHost) sim->viewTexHost = new ViewTexTable();
    const char *err = buildViewMeshTable(*sim->viewMeshHost, *sim->viewTexHost, data_dir);
    if (err[0]) { g_err = std::string("hs_enable_views: ") + err; return 1; }
    if (data_dir && data_dir[0]) {
        // the three textures of loadRenderObjects (src/mgr.cpp:649-665) as CUDA texture objects
        // configured like the reference's (M/src/render/asset_processor.cpp:312-349)
        const char *names[kViewTextures] = {"green_grid.png", "smile.png", "red_smile.png"};
        for (int i = 0; i < kViewTextures; i++) {
            int w = 0, h = 0;
            std::vector<uint8_t> rgba;
            err = decodePngRGBA((std::string(data_dir) + "/" + names[i]).c_str(), &w, &h, rgba);
            if (err[0]) { g_err = std::string("hs_enable_views: ") + err; return 1; }
            cudaChannelFormatDesc ch = cudaCreateChannelDesc<uchar4>();
            HS_CUDA(cudaMallocArray(&sim->viewTexArray[i], &ch, w, h, cudaArrayDefault));
            HS_CUDA(cudaMemcpy2DToArray(sim->viewTexArray[i], 0, 0, rgba.data(), sizeof(uint32_t) * w, sizeof(uint32_t) * w, h,
                                        cudaMemcpyHostToDevice));
            cudaResourceDesc res = {};
            res.resType = cudaResourceTypeArray;
            res.res.array.array = sim->viewTexArray[i];
            cudaTextureDesc td = {};
            td.addressMode[0] = cudaAddressModeWrap;
            td.addressMode[1] = cudaAddressModeWrap;
            td.filterMode = cudaFilterModeLinear;
            td.readMode = cudaReadModeNormalizedFloat;
            td.normalizedCoords = 1;
            cudaTextureObject_t obj = 0;
            HS_CUDA(cudaCreateTextureObject(&obj, &res, &td, nullptr));
            sim->viewTexHost->tex[i] = (unsigned long long)obj;
        }
    }
    ViewMeshTable *dev = nullptr;
    HS_TRY(devAlloc(sim, &dev, 1));
    HS_CUDA(cudaMemcpy(dev, sim->viewMeshHost, sizeof(ViewMeshTable), cudaMemcpyHostToDevice));
    ViewTexTable *devTex = nullptr;
    HS_TRY(devAlloc(sim, &devTex, 1));
    HS_CUDA(cudaMemcpy(devTex, sim->viewTexHost, sizeof(ViewTexTable), cudaMemcpyHostToDevice));
    sim->view.texTable = devTex;
    const size_t pixels = size_t(sim->cfg.num_worlds) * sim->S.p.agentsPerWorld * width * height;
    HS_TRY(devAlloc(sim, &sim->view.depth, pixels));
    HS_TRY(devAlloc(sim, &sim->view.rgb, pixels * 4));
    sim->view.mesh = dev;
    sim->view.width = width;
    sim->view.height = height;
    sim->viewsEnabled = true;
    return 0;
}

extern "C" int hs_render_views(hs_sim *sim, void *stream)
{
    if (!sim || !sim->inited) { g_err = "hs_render_views: hs_init has not run"; return 1; }
    if (!sim->viewsEnabled) { g_err = "hs_render_views: agent views are not enabled (hs_enable_views)"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    const unsigned views = unsigned(sim->cfg.num_worlds) * sim->S.p.agentsPerWorld;
    HS_CUDA(launchAgentViews(sim->S, sim->view, views, (cudaStream_t)stream));
    return 0;
}

extern "C" int hs_view_mesh_table_bytes(void) { return (int)sizeof(ViewMeshTable); }

// host-only: the table hs_enable_views would build for `data_dir` (no device needed)
extern "C" int hs_build_view_mesh_table(const char *data_dir, void *out, int out_bytes)
{
    if (!out || out_bytes != (int)sizeof(ViewMeshTable)) { g_err = "hs_build_view_mesh_table: bad buffer"; return 1; }
    if (!data_dir || !data_dir[0]) data_dir = getenv("HIDESEEK_DATA_DIR");
    ViewTexTable *X = new ViewTexTable();
    const char *err = buildViewMeshTable(*(ViewMeshTable *)out, *X, data_dir);
    delete X;
    if (err[0]) { g_err = std::string("hs_build_view_mesh_table: ") + err; return 1; }
    return 0;
}

// host-only diagnostic: the PNG reader hs_enable_views uses for the textures.  `out` receives
// width*height RGBA8 pixels if it has room (`cap` bytes); width / height are always reported.
extern "C" int hs_decode_png(const char *path, int *width, int *height, uint8_t *out, int64_t cap)
{
    if (!path || !width || !height) { g_err = "hs_decode_png: null argument"; return 1; }
    std::vector<uint8_t> rgba;
    const char *err = decodePngRGBA(path, width, height, rgba);
    if (err[0]) { g_err = std::string("hs_decode_png: ") + err; return 1; }
    if (out && cap >= (int64_t)rgba.size()) memcpy(out, rgba.data(), rgba.size());
    return 0;
}

extern "C" int hs_view_tex_table_bytes(void) { return (int)sizeof(ViewTexTable); }

// host-only: per-triangle texture coordinates and material -> texture bindings for `data_dir`
// (the texture-object handles stay 0: they only exist inside a simulator)
extern "C" int hs_build_view_tex_table(const char *data_dir, void *out, int out_bytes)
{
    if (!out || out_bytes != (int)sizeof(ViewTexTable)) { g_err = "hs_build_view_tex_table: bad buffer"; return 1; }
    if (!data_dir || !data_dir[0]) data_dir = getenv("HIDESEEK_DATA_DIR");
    ViewMeshTable *T = new ViewMeshTable();
    const char *err = buildViewMeshTable(*T, *(ViewTexTable *)out, data_dir);
    delete T;
    if (err[0]) { g_err = std::string("hs_build_view_tex_table: ") + err; return 1; }
    return 0;
}

extern "C" int hs_sync(hs_sim *sim, void *stream)
{
    if (!sim) { g_err = "hs_sync: null sim"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

extern "C" int hs_export(hs_sim *sim, int id, void **ptr, int *dtype, int *ndim, int64_t *dims)
{
    if (!sim || !ptr || !dtype || !ndim || !dims) { g_err = "hs_export: null argument"; return 1; }
    const DeviceState &S = sim->S;
    const int64_t N = S.p.numWorlds, R = N * S.p.agentsPerWorld;
    auto set = [&](void *p, int dt, int nd, int64_t d0, int64_t d1, int64_t d2) {
        *ptr = p; *dtype = dt; *ndim = nd; dims[0] = d0; dims[1] = d1; dims[2] = d2; dims[3] = 0;
    };
    // shapes: src/mgr.cpp:1062-1331
    switch (id) {
    case HS_EXPORT_RESET: set(S.exReset, HS_I32, 2, N, 1, 0); break;
    case HS_EXPORT_PREP_COUNTER: set(S.exPrepCounter, HS_I32, 2, R, 1, 0); break;
    case HS_EXPORT_ACTION: set(S.exAction, HS_I32, 2, R, 5, 0); break;
    case HS_EXPORT_SELF_OBS: set(S.exSelfObs, HS_F32, 2, R, 13, 0); break;
    case HS_EXPORT_SELF_TYPE: set(S.exSelfType, HS_I32, 2, R, 1, 0); break;
    case HS_EXPORT_SELF_MASK: set(S.exSelfMask, HS_F32, 2, R, 1, 0); break;
    case HS_EXPORT_AGENT_OBS: set(S.exAgentObs, HS_F32, 3, R, 5, 14); break;
    case HS_EXPORT_BOX_OBS: set(S.exBoxObs, HS_F32, 3, R, 9, 17); break;
    case HS_EXPORT_RAMP_OBS: set(S.exRampObs, HS_F32, 3, R, 2, 14); break;
    case HS_EXPORT_AGENT_VIS: set(S.exAgentVis, HS_F32, 3, R, 5, 1); break;
    case HS_EXPORT_BOX_VIS: set(S.exBoxVis, HS_F32, 3, R, 9, 1); break;
    case HS_EXPORT_RAMP_VIS: set(S.exRampVis, HS_F32, 3, R, 2, 1); break;
    case HS_EXPORT_LIDAR: set(S.exLidar, HS_F32, 2, R, 30, 0); break;
    case HS_EXPORT_SEED: set(S.exSeed, HS_I32, 2, R, 2, 0); break;
    case HS_EXPORT_REWARD: set(S.exReward, HS_F32, 2, R, 1, 0); break;
    case HS_EXPORT_DONE: set(S.exDone, HS_I32, 2, R, 1, 0); break;
    case HS_EXPORT_GLOBAL_POS: set(S.exGlobalPos, HS_F32, 3, N, 17, 2); break;
    case HS_EXPORT_AGENT_POLICY: set(S.exPolicy, HS_I32, 2, R, 1, 0); break;
    case HS_EXPORT_EPISODE_RESULT: set(S.exEpisodeResult, HS_F32, 2, N, 2, 0); break;
    case HS_EXPORT_CKPT_CTRL: set(S.exCkptCtrl, HS_I32, 2, N, 1, 0); break;
    case HS_EXPORT_CKPT: set(S.exCkpt, HS_U8, 2, N, kCheckpointBytes, 0); break;
    // renderer outputs (src/mgr.cpp:1241-1263): [N*A, H, W, 1] f32 and [N*A, H, W, 4] u8
    case HS_EXPORT_VIEW_DEPTH:
    case HS_EXPORT_VIEW_RGB:
        if (!sim->viewsEnabled) { g_err = "hs_export: agent views are not enabled (hs_enable_views)"; return 1; }
        if (id == HS_EXPORT_VIEW_DEPTH) set(sim->view.depth, HS_F32, 4, R, sim->view.height, sim->view.width);
        else set(sim->view.rgb, HS_U8, 4, R, sim->view.height, sim->view.width);
        dims[3] = id == HS_EXPORT_VIEW_DEPTH ? 1 : 4;
        break;
    default: g_err = "hs_export: unknown export id"; return 1;
    }
    return 0;
}

extern "C" int hs_trigger_reset(hs_sim *sim, int world, int level)
{
    if (!sim || world < 0 || world >= sim->cfg.num_worlds) { g_err = "hs_trigger_reset: bad world"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    int32_t v = level;
    HS_CUDA(cudaMemcpy(sim->S.exReset + world, &v, 4, cudaMemcpyHostToDevice));
    return 0;
}

extern "C" int hs_set_action(hs_sim *sim, int row, int x, int y, int r, int g, int l)
{
    if (!sim || row < 0 || row >= sim->cfg.num_worlds * sim->S.p.agentsPerWorld) {
        g_err = "hs_set_action: bad agent row";
        return 1;
    }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    int32_t a[5] = {x, y, r, g ? 1 : 0, l ? 1 : 0};
    HS_CUDA(cudaMemcpy(sim->S.exAction + size_t(row) * 5, a, sizeof(a), cudaMemcpyHostToDevice));
    return 0;
}

extern "C" int hs_step(hs_sim *sim, void *stream)
{
    if (!sim || !sim->inited) { g_err = "hs_step: hs_init has not run"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    cudaStream_t st = (cudaStream_t)stream;
    if (!sim->stepGraph) {
        // capture the step once on a private stream; replay on the caller's stream
        cudaGraph_t graph;
        HS_CUDA(cudaStreamBeginCapture(sim->graphStream, cudaStreamCaptureModeThreadLocal));
        const int N = sim->cfg.num_worlds, B = sim->slices;
        const int per = ((N + B - 1) / B + 127) / 128 * 128;      // slice starts stay line-aligned
        int launches = 0;
        HS_CUDA(cudaEventRecord(sim->forkEvent, sim->graphStream));
        const int BR = sim->branches;
        for (int b = 1; b < BR; b++) HS_CUDA(cudaStreamWaitEvent(sim->sliceStream[b], sim->forkEvent, 0));
        for (int sl = 0; sl < B; sl++) {
            // slices are dealt round-robin to the branches; a branch runs its slices back to back
            const int begin = sl * per, count = (begin + per <= N ? per : N - begin);
            if (count <= 0) continue;
            const int b = sl % BR;
            cudaStream_t sb = b == 0 ? sim->graphStream : sim->sliceStream[b];
            Launcher L{sliceOf(sim->S, begin, count), sb};
            if (sim->obsFork) {
                L.aux[0] = sim->auxStream[b][0]; L.aux[1] = sim->auxStream[b][1];
                L.auxFork = sim->auxFork[b]; L.auxJoin[0] = sim->auxJoin[b][0]; L.auxJoin[1] = sim->auxJoin[b][1];
            }
            L.orderedSolve = orderedSolveDefault();
            launchStep(L, sim->cosThreshold, sim->raysNearestFirst);
            launches += L.count;
        }
        for (int b = 1; b < BR; b++) {
            HS_CUDA(cudaEventRecord(sim->joinEvent[b], sim->sliceStream[b]));
            HS_CUDA(cudaStreamWaitEvent(sim->graphStream, sim->joinEvent[b], 0));
        }
        Launcher L{sim->S, sim->graphStream};
        L.count = launches;
        HS_CUDA(cudaStreamEndCapture(sim->graphStream, &graph));
        HS_CUDA(cudaGraphInstantiate(&sim->stepGraph, graph, 0));
        HS_CUDA(cudaGraphDestroy(graph));
        sim->lastLaunches = L.count;
    }
    if (sim->orderView != 1) {
        HS_CUDA(cudaMemsetAsync(sim->S.solveCnt, 0, sizeof(int) * (size_t(sim->cfg.num_worlds) + 64), st));
        sim->orderView = 1;
    }
    HS_CUDA(cudaGraphLaunch(sim->stepGraph, st));
    return 0;
}

// ---- checkpoints (SaveCheckpoints / LoadCheckpoints task graphs, src/sim.cpp:1315-1332) ------
extern "C" int hs_save_checkpoints(hs_sim *sim, void *stream)
{
    if (!sim || !sim->inited) { g_err = "hs_save_checkpoints: hs_init has not run"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    Launcher L{sim->S, (cudaStream_t)stream};
    k_save_checkpoint<<<L.gWorld(), 128, 0, L.st>>>(L.S);
    HS_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int hs_load_checkpoints(hs_sim *sim, void *stream)
{
    if (!sim || !sim->inited) { g_err = "hs_load_checkpoints: hs_init has not run"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    Launcher L{sim->S, (cudaStream_t)stream};
    k_load_checkpoint<<<L.gWorld(64), 64, 0, L.st>>>(L.S);
    launchBroadphaseSetup(L, true);            // postGenTasks
    launchObservations(L, sim->cosThreshold, sim->raysNearestFirst);  // observationsTasks
    HS_CUDA(cudaGetLastError());
    return 0;
}

// ---- XLA / JAX buffer contract (src/mgr.cpp:168-201, 326-437) ----------------------------------
// `buffers` is the flat device-pointer list an XLA GPU custom call receives, AFTER the two leading
// entries the Madrona wrapper skips (M/include/madrona/py/bindings.inl:107-119).  Everything is
// enqueued on the caller's stream; only hs_jax_init waits (the reference synchronises there).
static size_t exportBytes(hs_sim *sim, int id, void **ptr)
{
    int dt, nd; int64_t dims[4];
    if (hs_export(sim, id, ptr, &dt, &nd, dims)) return 0;
    size_t n = dt == HS_U8 ? 1 : 4;
    for (int i = 0; i < nd; i++) n *= size_t(dims[i]);
    return n;
}

static int copyOut(hs_sim *sim, cudaStream_t st, int id, void *dst)
{
    void *src; size_t n = exportBytes(sim, id, &src);
    HS_CUDA(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToDevice, st));
    return 0;
}

static int copyIn(hs_sim *sim, cudaStream_t st, int id, const void *src)
{
    void *dst; size_t n = exportBytes(sim, id, &dst);
    HS_CUDA(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToDevice, st));
    return 0;
}

// JAXIOObservations::set order (src/mgr.cpp:168-201)
static const int kJaxObsOrder[11] = {
    HS_EXPORT_PREP_COUNTER, HS_EXPORT_SELF_OBS, HS_EXPORT_SELF_TYPE, HS_EXPORT_SELF_MASK, HS_EXPORT_LIDAR,
    HS_EXPORT_AGENT_OBS, HS_EXPORT_BOX_OBS, HS_EXPORT_RAMP_OBS,
    HS_EXPORT_AGENT_VIS, HS_EXPORT_BOX_VIS, HS_EXPORT_RAMP_VIS,
};

static int copyOutObservations(hs_sim *sim, cudaStream_t st, void **&buffers)
{
    for (int i = 0; i < 11; i++)
        if (copyOut(sim, st, kJaxObsOrder[i], *buffers++)) return 1;
    return 0;
}

extern "C" int hs_jax_init(hs_sim *sim, void *stream, void **buffers)
{
    if (!sim || !buffers) { g_err = "hs_jax_init: null argument"; return 1; }
    if (hs_init(sim, stream)) return 1;
    if (copyOutObservations(sim, (cudaStream_t)stream, buffers)) return 1;
    HS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

extern "C" int hs_jax_step(hs_sim *sim, void *stream, void **buffers)
{
    if (!sim || !buffers) { g_err = "hs_jax_step: null argument"; return 1; }
    cudaStream_t st = (cudaStream_t)stream;
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    if (copyIn(sim, st, HS_EXPORT_ACTION, *buffers++)) return 1;
    if (copyIn(sim, st, HS_EXPORT_RESET, *buffers++)) return 1;
    if (copyIn(sim, st, HS_EXPORT_AGENT_POLICY, *buffers++)) return 1;
    if (hs_step(sim, stream)) return 1;
    if (copyOutObservations(sim, st, buffers)) return 1;
    if (copyOut(sim, st, HS_EXPORT_REWARD, *buffers++)) return 1;
    if (copyOut(sim, st, HS_EXPORT_DONE, *buffers++)) return 1;
    if (copyOut(sim, st, HS_EXPORT_EPISODE_RESULT, *buffers++)) return 1;
    return 0;
}

extern "C" int hs_jax_save_checkpoints(hs_sim *sim, void *stream, void **buffers)
{
    if (!sim || !buffers) { g_err = "hs_jax_save_checkpoints: null argument"; return 1; }
    cudaStream_t st = (cudaStream_t)stream;
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    if (copyIn(sim, st, HS_EXPORT_CKPT_CTRL, buffers[0])) return 1;
    if (hs_save_checkpoints(sim, stream)) return 1;
    return copyOut(sim, st, HS_EXPORT_CKPT, buffers[1]);
}

extern "C" int hs_jax_load_checkpoints(hs_sim *sim, void *stream, void **buffers)
{
    if (!sim || !buffers) { g_err = "hs_jax_load_checkpoints: null argument"; return 1; }
    cudaStream_t st = (cudaStream_t)stream;
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    if (copyIn(sim, st, HS_EXPORT_CKPT_CTRL, *buffers++)) return 1;
    if (copyIn(sim, st, HS_EXPORT_CKPT, *buffers++)) return 1;
    if (hs_load_checkpoints(sim, stream)) return 1;
    return copyOutObservations(sim, st, buffers);
}

// The functions XLA itself would be handed (gpuEntryFn, bindings.inl:107-119): `opaque` holds the
// simulator pointer, buffers[0] is the CPU-backend slot and buffers[1] JAX's ordering token.
// XLA custom calls cannot return a status, so a failure aborts like the reference's REQ_CUDA.
static hs_sim *xlaSim(const char *opaque, size_t len, const char *who)
{
    hs_sim *sim = nullptr;
    if (opaque && len >= sizeof(hs_sim *)) memcpy(&sim, opaque, sizeof(hs_sim *));
    if (!sim) { fprintf(stderr, "%s: opaque does not hold a simulator pointer\n", who); abort(); }
    return sim;
}
#define HS_XLA_ENTRY(name, call)                                                                   \
    extern "C" void name(void *stream, void **buffers, const char *opaque, size_t opaque_len)     \
    {                                                                                              \
        if (call(xlaSim(opaque, opaque_len, #name), stream, buffers + 2)) {                        \
            fprintf(stderr, #name ": %s\n", g_err.c_str());                                        \
            abort();                                                                               \
        }                                                                                          \
    }
HS_XLA_ENTRY(hs_xla_gpu_init, hs_jax_init)
HS_XLA_ENTRY(hs_xla_gpu_step, hs_jax_step)
HS_XLA_ENTRY(hs_xla_gpu_save_checkpoints, hs_jax_save_checkpoints)
HS_XLA_ENTRY(hs_xla_gpu_load_checkpoints, hs_jax_load_checkpoints)

static int setCkptTrigger(hs_sim *sim, int world, const char *who)
{
    if (!sim || world < 0 || world >= sim->cfg.num_worlds) { g_err = std::string(who) + ": bad world"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    int32_t one = 1;
    HS_CUDA(cudaMemcpy(sim->S.exCkptCtrl + world, &one, 4, cudaMemcpyHostToDevice));
    return 0;
}

// Manager::saveCheckpoint / loadCheckpoint (src/mgr.cpp:905-964): set one world's trigger,
// run the graph, wait.
extern "C" int hs_save_checkpoint(hs_sim *sim, int world)
{
    HS_TRY(setCkptTrigger(sim, world, "hs_save_checkpoint"));
    HS_TRY(hs_save_checkpoints(sim, nullptr));
    HS_CUDA(cudaStreamSynchronize(nullptr));
    return 0;
}

extern "C" int hs_load_checkpoint(hs_sim *sim, int world)
{
    HS_TRY(setCkptTrigger(sim, world, "hs_load_checkpoint"));
    HS_TRY(hs_load_checkpoints(sim, nullptr));
    HS_CUDA(cudaStreamSynchronize(nullptr));
    return 0;
}

extern "C" int hs_step_host(hs_sim *sim, void *stream, const int32_t *actions_host,
                            float *reward_host, int32_t *done_host)
{
    if (!sim) { g_err = "hs_step_host: null sim"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    cudaStream_t st = (cudaStream_t)stream;
    size_t R = size_t(sim->cfg.num_worlds) * sim->S.p.agentsPerWorld;
    if (actions_host)
        HS_CUDA(cudaMemcpyAsync(sim->S.exAction, actions_host, R * 5 * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    HS_TRY(hs_step(sim, stream));
    if (reward_host)
        HS_CUDA(cudaMemcpyAsync(reward_host, sim->S.exReward, R * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (done_host)
        HS_CUDA(cudaMemcpyAsync(done_host, sim->S.exDone, R * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    return 0;
}

// Instrumented step: same launch sequence without the graph, one CUDA event
// after every kernel on `stream`; returns per-launch durations.  Used by
// bench.py for the roofline object; not on the product path.
static thread_local std::vector<const char *> g_profNames;

extern "C" int hs_step_profiled(hs_sim *sim, void *stream, float *ms_out, int cap, int *n_out)
{
    if (!sim || !sim->inited) { g_err = "hs_step_profiled: hs_init has not run"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<cudaEvent_t> ev;
    g_profNames.clear();
    cudaEvent_t start;
    HS_CUDA(cudaEventCreate(&start));
    HS_CUDA(cudaEventRecord(start, st));
    if (sim->orderView != 2) {
        HS_CUDA(cudaMemsetAsync(sim->S.solveCnt, 0, sizeof(int) * (size_t(sim->cfg.num_worlds) + 64), st));
        sim->orderView = 2;
    }
    Launcher L{sim->S, st};
    L.events = &ev;
    L.names = &g_profNames;
    L.orderedSolve = orderedSolveDefault();
    launchStep(L, sim->cosThreshold, sim->raysNearestFirst);
    HS_CUDA(cudaStreamSynchronize(st));
    cudaEvent_t prev = start;
    int n = 0;
    for (size_t i = 0; i < ev.size(); i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, prev, ev[i]);
        if (n < cap) ms_out[n++] = ms;
        prev = ev[i];
    }
    cudaEventDestroy(start);
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    *n_out = n;
    return 0;
}

extern "C" const char *hs_profiled_kernel_name(int i)
{
    return (i >= 0 && i < (int)g_profNames.size()) ? g_profNames[i] : "";
}

// Measurement only: totals over the shard of live bodies, agents, broadphase
// candidates and contacts of the most recent substep (bench.py uses the means to
// scale the algorithmic-bytes figures of DESIGN.md).
extern "C" int hs_world_stats(hs_sim *sim, int64_t *out4)
{
    if (!sim || !out4) { g_err = "hs_world_stats: null argument"; return 1; }
    HS_CUDA(cudaSetDevice(sim->cfg.gpu_id));
    HS_CUDA(cudaDeviceSynchronize());
    const size_t N = sim->cfg.num_worlds;
    std::vector<int32_t> row(N);
    std::vector<int32_t> ncand(N);
    auto sumRow = [&](int f, std::vector<int32_t> &dst, int64_t *acc) -> int {
        HS_CUDA(cudaMemcpy(dst.data(), sim->S.wi + size_t(f) * N, N * 4, cudaMemcpyDeviceToHost));
        int64_t s = 0;
        for (int32_t v : dst) s += v;
        *acc = s;
        return 0;
    };
    HS_TRY(sumRow(WF_NUM_BODIES, row, &out4[0]));
    HS_TRY(sumRow(WF_NUM_AGENTS, row, &out4[1]));
    HS_TRY(sumRow(WF_NUM_CAND, ncand, &out4[2]));
    HS_TRY(sumRow(WF_NUM_CONTACTS, row, &out4[3]));
    return 0;
}
