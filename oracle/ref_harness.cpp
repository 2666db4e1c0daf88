// TEST INFRASTRUCTURE ONLY — C-ABI driver around the UNMODIFIED reference.
//
// This translation unit textually includes the reference's src/mgr.cpp from
// where it lies under $(REF) (never copied into this repo) so that the harness
// can reach Manager::CPUImpl and, through it, each world's ECS context.  All
// simulation arithmetic executed through these entry points is the reference's
// own CPU backend; the harness only constructs, steps and reads.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load the resulting library.

#include <algorithm>
#include <array>
#include <atomic>
#include <charconv>
#include <chrono>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#include <sched.h>
#include <pthread.h>

// Headers that forward-declare nested types with an explicit access level
// must be seen before the access override below.
#include <madrona/memory.hpp>

// Read-only access to engine internals (BVH nodes, RNG counters).
#define private public
#define protected public
#include <mgr.cpp>
#undef private
#undef protected

#include "ref_dump.h"

// the reference's mesh ray tracing (header-only; M/include/madrona/mesh_bvh.hpp, mesh_bvh.inl)
#include <madrona/mesh_bvh.hpp>

// The contact solver state lives in an archetype column whose component type is private to
// M/src/physics/xpbd.cpp:9-11; component ids are tracked by type NAME, so declaring the same
// one-member struct under the same name lets the probe read the stored normal multiplier.
namespace madrona::phys::xpbd {
struct XPBDContactState {
    float lambdaN[4];
};
}

using namespace GPUHideSeek;
namespace bp = madrona::phys::broadphase;

namespace {

// what hsref_step_probe captures per world (the part of hsref_compact no export shows)
struct ProbeData {
    int32_t numCand, numContacts;
    uint8_t cand[HSREF_C_CAND][2];
    uint8_t ref[HSREF_C_CONTACTS], alt[HSREF_C_CONTACTS], npts[HSREF_C_CONTACTS];
    float pts[HSREF_C_CONTACTS][4][4];
    float normal[HSREF_C_CONTACTS][3];
    float lambda[HSREF_C_CONTACTS];
    int8_t lidarHit[HSREF_MAX_AGENTS][30];
};

struct Handle {
    Manager *mgr;
    Manager::CPUImpl *impl;
    int numWorlds;
    int agentsPerWorld;
    std::vector<ProbeData> probe;       // empty until the first hsref_step_probe
    int candClearNode = -1;             // Step-graph node that clears the candidate pairs
    int contactClearNode = -1;          // ... the contacts of the LAST substep
};

inline Manager::CPUImpl *cpuImpl(Manager *m)
{
    return static_cast<Manager::CPUImpl *>(m->impl_);
}

}

extern "C" {

// num_threads <= 0: use every core in the affinity mask (reference default).
void *hsref_create(int num_worlds, uint32_t sim_flags, uint32_t seed,
                   int min_hiders, int max_hiders,
                   int min_seekers, int max_seekers, int num_threads)
{
    cpu_set_t orig;
    pthread_getaffinity_np(pthread_self(), sizeof(orig), &orig);

    if (num_threads > 0) {
        cpu_set_t limited;
        CPU_ZERO(&limited);
        int taken = 0;
        for (int i = 0; i < CPU_SETSIZE && taken < num_threads; i++) {
            if (CPU_ISSET(i, &orig)) {
                CPU_SET(i, &limited);
                taken++;
            }
        }
        pthread_setaffinity_np(pthread_self(), sizeof(limited), &limited);
    }

    Manager *mgr = new Manager(Manager::Config {
        .execMode = madrona::ExecMode::CPU,
        .gpuID = 0,
        .numWorlds = (uint32_t)num_worlds,
        .simFlags = (SimFlags)sim_flags,
        .randSeed = seed,
        .minHiders = (uint32_t)min_hiders,
        .maxHiders = (uint32_t)max_hiders,
        .minSeekers = (uint32_t)min_seekers,
        .maxSeekers = (uint32_t)max_seekers,
        .numPBTPolicies = 0,
        .enableBatchRenderer = false,
    });

    pthread_setaffinity_np(pthread_self(), sizeof(orig), &orig);

    Handle *h = new Handle();
    h->mgr = mgr;
    h->impl = cpuImpl(mgr);
    h->numWorlds = num_worlds;
    h->agentsPerWorld = max_hiders + max_seekers;
    return h;
}

void hsref_destroy(void *hp)
{
    Handle *h = (Handle *)hp;
    delete h->mgr;
    delete h;
}

void hsref_init(void *hp) { ((Handle *)hp)->mgr->init(); }
void hsref_step(void *hp) { ((Handle *)hp)->mgr->step(); }

// Worker count the reference pool was created with (= CPUs in the affinity
// mask at construction, M/src/mw/cpu_exec.cpp:30-56,102).
int hsref_affinity_cpus()
{
    cpu_set_t cur;
    pthread_getaffinity_np(pthread_self(), sizeof(cur), &cur);
    return CPU_COUNT(&cur);
}

void *hsref_tensor(void *hp, int export_id)
{
    Handle *h = (Handle *)hp;
    return h->impl->cpuExec.getExported((uint32_t)export_id);
}

int hsref_checkpoint_bytes() { return (int)sizeof(Checkpoint); }

void hsref_trigger_reset(void *hp, int world, int level)
{
    ((Handle *)hp)->mgr->triggerReset(world, level);
}

void hsref_set_action(void *hp, int agent_row, int x, int y, int r, int g, int l)
{
    ((Handle *)hp)->mgr->setAction(agent_row, x, y, r, g != 0, l != 0);
}

void hsref_save_checkpoint(void *hp, int world) { ((Handle *)hp)->mgr->saveCheckpoint(world); }
void hsref_load_checkpoint(void *hp, int world) { ((Handle *)hp)->mgr->loadCheckpoint(world); }

static void fillBody(Engine &ctx, Entity e, int archetype, hsref_body *b)
{
    using namespace madrona;
    using namespace madrona::phys;
    using namespace madrona::base;

    Loc loc = ctx.loc(e);
    b->entity_id = e.id;
    b->archetype = archetype;
    b->row = loc.row;
    b->obj_id = ctx.get<ObjectID>(e).idx;
    b->response = (int32_t)ctx.get<ResponseType>(e);
    b->owner = (int32_t)ctx.get<OwnerTeam>(e);
    b->leaf_id = ctx.get<bp::LeafID>(e).id;
    b->grab_joint = -1;
    if (archetype == 1) {
        Entity j = ctx.get<GrabData>(e).constraintEntity;
        b->grab_joint = (j == Entity::none()) ? -1 : j.id;
    }
    Vector3 p = ctx.get<Position>(e);
    Quat q = ctx.get<Rotation>(e);
    Diag3x3 s = ctx.get<Scale>(e);
    Velocity v = ctx.get<Velocity>(e);
    Vector3 f = ctx.get<ExternalForce>(e);
    Vector3 t = ctx.get<ExternalTorque>(e);
    b->pos[0] = p.x; b->pos[1] = p.y; b->pos[2] = p.z;
    b->rot[0] = q.w; b->rot[1] = q.x; b->rot[2] = q.y; b->rot[3] = q.z;
    b->scale[0] = s.d0; b->scale[1] = s.d1; b->scale[2] = s.d2;
    b->vel[0] = v.linear.x; b->vel[1] = v.linear.y; b->vel[2] = v.linear.z;
    b->vel[3] = v.angular.x; b->vel[4] = v.angular.y; b->vel[5] = v.angular.z;
    b->ext_force[0] = f.x; b->ext_force[1] = f.y; b->ext_force[2] = f.z;
    b->ext_torque[0] = t.x; b->ext_torque[1] = t.y; b->ext_torque[2] = t.z;
}

int hsref_dump_world(void *hp, int world, hsref_world *out)
{
    using namespace madrona;
    using namespace madrona::phys;

    Handle *h = (Handle *)hp;
    Engine &ctx = h->impl->cpuExec.getWorldContext(world);
    Sim &sim = h->impl->cpuExec.getWorldData(world);

    memset(out, 0, sizeof(*out));

    out->cur_episode_step = (int32_t)sim.curEpisodeStep;
    out->cur_world_episode = (int32_t)sim.curWorldEpisode;
    out->episode_key[0] = sim.curEpisodeRNDCounter.a;
    out->episode_key[1] = sim.curEpisodeRNDCounter.b;
    out->rng_key[0] = sim.rng.k_.a;
    out->rng_key[1] = sim.rng.k_.b;
    out->rng_count = sim.rng.count_;
    out->num_hiders = sim.numHiders;
    out->num_seekers = sim.numSeekers;
    out->num_active_agents = (int32_t)sim.numActiveAgents;
    out->num_obstacles = sim.numObstacles;
    out->num_boxes = (int32_t)sim.numActiveBoxes;
    out->num_ramps = (int32_t)sim.numActiveRamps;
    out->hider_team_reward = sim.hiderTeamReward.load_relaxed();
    out->seekers_first = ctx.singleton<TeamState>().seekersFirst ? 1 : 0;
    out->running_scores[0] = ctx.singleton<EpisodeStats>().runningScores[0];
    out->running_scores[1] = ctx.singleton<EpisodeStats>().runningScores[1];

    for (int i = 0; i < sim.numObstacles && i < HSREF_MAX_BODIES; i++) {
        out->obstacles[i] = sim.obstacles[i].id;
    }
    for (int i = 0; i < 9; i++) {
        out->boxes[i] = i < sim.numActiveBoxes ? sim.boxes[i].id : -1;
        out->box_sizes[i][0] = sim.boxSizes[i].x;
        out->box_sizes[i][1] = sim.boxSizes[i].y;
        out->box_sizes[i][2] = sim.boxSizes[i].z;
    }
    for (int i = 0; i < 2; i++) {
        out->ramps[i] = i < sim.numActiveRamps ? sim.ramps[i].id : -1;
    }
    for (int i = 0; i < 3; i++) {
        out->hiders[i] = i < sim.numHiders ? sim.hiders[i].id : -1;
        out->seekers[i] = i < sim.numSeekers ? sim.seekers[i].id : -1;
    }
    for (int i = 0; i < HSREF_MAX_AGENTS; i++) {
        out->agent_iface_sim[i] = -1;
        if (i < sim.maxAgentsPerWorld) {
            Entity se = ctx.get<SimEntity>(sim.agentInterfaces[i]).e;
            out->agent_iface_sim[i] = (se == Entity::none()) ? -1 : se.id;
        }
    }

    // Bodies in ECS iteration order (src/core/state.cpp query order:
    // ascending archetype id => DynamicObject then DynAgent; rows in table
    // order, holes skipped).
    int nb = 0;
    {
        auto q = ctx.query<Entity, OwnerTeam, bp::LeafID>();
        ctx.iterateQuery(q, [&](Entity e, OwnerTeam &, bp::LeafID &) {
            if (nb >= HSREF_MAX_BODIES) return;
            int arch = ctx.getSafe<GrabData>(e).valid() ? 1 : 0;
            fillBody(ctx, e, arch, &out->bodies[nb++]);
        });
    }
    out->num_bodies = nb;

    int nj = 0;
    {
        auto q = ctx.query<Entity, JointConstraint>();
        ctx.iterateQuery(q, [&](Entity e, JointConstraint &j) {
            if (nj >= HSREF_MAX_JOINTS) return;
            hsref_joint &o = out->joints[nj++];
            o.entity_id = e.id;
            o.e1 = j.e1.id;
            o.e2 = j.e2.id;
            Quat a1 = j.fixed.attachRot1, a2 = j.fixed.attachRot2;
            o.attach1[0] = a1.w; o.attach1[1] = a1.x; o.attach1[2] = a1.y; o.attach1[3] = a1.z;
            o.attach2[0] = a2.w; o.attach2[1] = a2.x; o.attach2[2] = a2.y; o.attach2[3] = a2.z;
            o.separation = j.fixed.separation;
            o.r1[0] = j.r1.x; o.r1[1] = j.r1.y; o.r1[2] = j.r1.z;
            o.r2[0] = j.r2.x; o.r2[1] = j.r2.y; o.r2[2] = j.r2.z;
        });
    }
    out->num_joints = nj;

    bp::BVH &bvh = ctx.singleton<bp::BVH>();
    int nl = bvh.num_leaves_.load_relaxed();
    out->bvh_num_leaves = nl;
    out->bvh_num_nodes = (int32_t)bvh.num_nodes_;
    out->bvh_force_rebuild = bvh.force_rebuild_ ? 1 : 0;
    static_assert(sizeof(bp::BVH::Node) == sizeof(hsref_node));
    int nn = std::min<int>((int)bvh.num_nodes_, HSREF_MAX_NODES);
    memcpy(out->nodes, bvh.nodes_, sizeof(hsref_node) * nn);
    for (int i = 0; i < nl && i < HSREF_MAX_BODIES; i++) {
        math::AABB a = bvh.leaf_aabbs_[i];
        out->leaf_aabbs[i][0] = a.pMin.x; out->leaf_aabbs[i][1] = a.pMin.y;
        out->leaf_aabbs[i][2] = a.pMin.z; out->leaf_aabbs[i][3] = a.pMax.x;
        out->leaf_aabbs[i][4] = a.pMax.y; out->leaf_aabbs[i][5] = a.pMax.z;
        out->leaf_parents[i] = bvh.leaf_parents_[i];
        out->leaf_entities[i] = bvh.leaf_entities_[i].id;
        out->leaf_obj_ids[i] = bvh.leaf_obj_ids_[i].idx;
    }

    return 0;
}


// ---- bulk parity support -----------------------------------------------------------------------
namespace {

struct BodyIndex {
    int n = 0;
    madrona::Entity e[HSREF_MAX_BODIES];
    madrona::Loc loc[HSREF_MAX_BODIES];
    int arch[HSREF_MAX_BODIES];

    void build(Engine &ctx)
    {
        using namespace madrona;
        n = 0;
        auto q = ctx.query<Entity, OwnerTeam, bp::LeafID>();
        ctx.iterateQuery(q, [&](Entity ent, OwnerTeam &, bp::LeafID &) {
            if (n >= HSREF_MAX_BODIES) return;
            e[n] = ent;
            loc[n] = ctx.loc(ent);
            arch[n] = ctx.getSafe<GrabData>(ent).valid() ? 1 : 0;
            n++;
        });
    }
    int ofLoc(madrona::Loc l) const
    {
        for (int i = 0; i < n; i++)
            if (loc[i].archetype == l.archetype && loc[i].row == l.row) return i;
        return 255;
    }
    int ofEntity(madrona::Entity x) const
    {
        for (int i = 0; i < n; i++)
            if (e[i] == x) return i;
        return -1;
    }
};

void snapCandidates(Engine &ctx, ProbeData &P)
{
    using namespace madrona;
    using namespace madrona::phys;
    BodyIndex B;
    B.build(ctx);
    int n = 0;
    auto q = ctx.query<CandidateCollision>();
    ctx.iterateQuery(q, [&](CandidateCollision &c) {
        if (n < HSREF_C_CAND) {
            P.cand[n][0] = (uint8_t)B.ofLoc(c.a);
            P.cand[n][1] = (uint8_t)B.ofLoc(c.b);
        }
        n++;
    });
    P.numCand = n;
}

void snapContacts(Engine &ctx, ProbeData &P)
{
    using namespace madrona;
    using namespace madrona::phys;
    BodyIndex B;
    B.build(ctx);
    int n = 0;
    auto q = ctx.query<ContactConstraint, xpbd::XPBDContactState>();
    ctx.iterateQuery(q, [&](ContactConstraint &c, xpbd::XPBDContactState &st) {
        if (n < HSREF_C_CONTACTS) {
            P.ref[n] = (uint8_t)B.ofLoc(c.ref);
            P.alt[n] = (uint8_t)B.ofLoc(c.alt);
            P.npts[n] = (uint8_t)c.numPoints;
            for (int p = 0; p < 4; p++) {
                P.pts[n][p][0] = c.points[p].x; P.pts[n][p][1] = c.points[p].y;
                P.pts[n][p][2] = c.points[p].z; P.pts[n][p][3] = c.points[p].w;
            }
            P.normal[n][0] = c.normal.x; P.normal[n][1] = c.normal.y; P.normal[n][2] = c.normal.z;
            P.lambda[n] = st.lambdaN[0];
        }
        n++;
    });
    P.numContacts = n;
}

uint32_t countCandidates(Engine &ctx)
{
    auto q = ctx.query<madrona::phys::CandidateCollision>();
    uint32_t n = 0;
    ctx.iterateQuery(q, [&](madrona::phys::CandidateCollision &) { n++; });
    return n;
}

uint32_t countContacts(Engine &ctx)
{
    auto q = ctx.query<madrona::phys::ContactConstraint, madrona::phys::xpbd::XPBDContactState>();
    uint32_t n = 0;
    ctx.iterateQuery(q, [&](madrona::phys::ContactConstraint &, madrona::phys::xpbd::XPBDContactState &) { n++; });
    return n;
}

// The body every lidar ray hits: the ray set-up of lidarSystem (src/sim.cpp:712-759) handed to
// the reference's own BVH::traceRay on the state the step left behind (nothing moves between the
// lidar node and the end of the Step graph).
void snapLidarHits(Engine &ctx, Sim &sim, ProbeData &P)
{
    using namespace madrona;
    using namespace madrona::math;
    using namespace madrona::base;
    BodyIndex B;
    B.build(ctx);
    auto &bvh = ctx.singleton<bp::BVH>();
    for (int k = 0; k < HSREF_MAX_AGENTS; k++) {
        for (int i = 0; i < 30; i++) P.lidarHit[k][i] = -2;
        if (k >= sim.maxAgentsPerWorld) continue;
        Entity se = ctx.get<SimEntity>(sim.agentInterfaces[k]).e;
        if (se == Entity::none()) continue;
        Vector3 pos = ctx.get<Position>(se);
        Quat rot = ctx.get<Rotation>(se);
        Vector3 agent_fwd = rot.rotateVec(math::fwd);
        Vector3 right = rot.rotateVec(math::right);
        for (int32_t idx = 0; idx < 30; idx++) {
            float theta = 2.f * math::pi * (float(idx) / float(30)) + math::pi / 2.f;
            float x = cosf(theta);
            float y = sinf(theta);
            Vector3 ray_dir = (x * right + y * agent_fwd).normalize();
            float hit_t;
            Vector3 hit_normal;
            Entity hit = bvh.traceRay(pos, ray_dir, &hit_t, &hit_normal, 200.f);
            P.lidarHit[k][idx] = hit == Entity::none() ? -1 : (int8_t)B.ofEntity(hit);
        }
    }
}

// One world's Step graph, node by node (TaskGraph::run, M/src/core/taskgraph.cpp:164-170, is this
// loop without the probes).  discover: learn which nodes clear the candidates / contacts.
void runWorldProbed(Handle *h, int w, bool discover, std::vector<int> *contactClears)
{
    auto &exec = h->impl->cpuExec;
    const uint32_t tg = (uint32_t)TaskGraphID::Step;
    auto &jd = exec.job_datas_[(size_t)tg * h->numWorlds + w];
    Engine &ctx = exec.getWorldContext(w);
    madrona::TaskGraph &g = jd.taskgraph;
    ProbeData &P = h->probe[w];
    P.numCand = 0;
    P.numContacts = 0;
    int idx = 0;
    for (const auto &node : g.sorted_nodes_) {
        uint32_t candBefore = 0, conBefore = 0;
        if (discover) {
            candBefore = countCandidates(ctx);
            conBefore = countContacts(ctx);
            if (conBefore > 0) snapContacts(ctx, P);
            if (candBefore > 0) snapCandidates(ctx, P);
        } else {
            if (idx == h->contactClearNode) snapContacts(ctx, P);
            if (idx == h->candClearNode) snapCandidates(ctx, P);
        }
        node.fn((madrona::NodeBase *)(&g.node_datas_[node.dataIDX].userData[0]), jd.ctx, &g);
        if (discover) {
            if (candBefore > 0 && countCandidates(ctx) == 0) h->candClearNode = idx;
            if (conBefore > 0 && countContacts(ctx) == 0) contactClears->push_back(idx);
        }
        idx++;
    }
    snapLidarHits(ctx, exec.getWorldData(w), P);
}

}

// Manager::step() with probes: every world's Step graph is run node by node so that the
// candidate pairs and the contacts of the last substep can be copied out just before the
// reference clears them, and the lidar hit bodies are traced on the final state.  Worlds run on
// `num_threads` host threads (<= 0: every core in the affinity mask); they are independent.
// Returns 0, or -1 if the clearing nodes could not be identified.
int hsref_step_probe(void *hp, int num_threads)
{
    Handle *h = (Handle *)hp;
    const int N = h->numWorlds;
    if (h->probe.empty()) h->probe.resize(N);
    auto &exec = h->impl->cpuExec;
    const uint32_t tg = (uint32_t)TaskGraphID::Step;
    madrona::StateManager &sm = exec.job_datas_[(size_t)tg * N].taskgraph.stateManager();
    sm.copyInExportedColumns();

    // keep one instance of every query alive so worker threads never race to create them
    auto keepA = exec.getWorldContext(0).query<madrona::phys::CandidateCollision>();
    auto keepB = exec.getWorldContext(0).query<madrona::phys::ContactConstraint, madrona::phys::xpbd::XPBDContactState>();
    auto keepC = exec.getWorldContext(0).query<madrona::Entity, OwnerTeam, bp::LeafID>();

    int first = 0;
    while (first < N && (h->contactClearNode < 0 || h->candClearNode < 0)) {
        std::vector<int> clears;
        runWorldProbed(h, first, true, &clears);
        // four substeps (M/src/physics/xpbd.cpp:1080-1141) -> four clearing nodes, last one wins
        if (clears.size() == 4 && h->candClearNode >= 0) {
            h->contactClearNode = clears.back();
        } else {
            h->contactClearNode = -1;
            h->probe[first].numCand = -1;       // this world's probe data is not trustworthy
            h->probe[first].numContacts = -1;
        }
        first++;
    }
    if (h->contactClearNode < 0 || h->candClearNode < 0) {
        sm.copyOutExportedColumns();
        return -1;
    }

    int T = num_threads > 0 ? num_threads : hsref_affinity_cpus();
    if (T > N - first) T = N - first > 0 ? N - first : 1;
    std::atomic<int> next(first);
    auto worker = [&]() {
        for (;;) {
            int w = next.fetch_add(1);
            if (w >= N) break;
            runWorldProbed(h, w, false, nullptr);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < T; t++) pool.emplace_back(worker);
    worker();
    for (auto &t : pool) t.join();

    sm.copyOutExportedColumns();
    return 0;
}

int hsref_compact_bytes() { return (int)sizeof(hsref_compact); }

static void fillCompact(Handle *h, int world, hsref_compact *out)
{
    using namespace madrona;
    using namespace madrona::phys;
    using namespace madrona::base;

    Engine &ctx = h->impl->cpuExec.getWorldContext(world);
    Sim &sim = h->impl->cpuExec.getWorldData(world);
    memset(out, 0, sizeof(*out));
    out->cur_episode_step = (int32_t)sim.curEpisodeStep;
    out->cur_world_episode = (int32_t)sim.curWorldEpisode;
    out->rng_key[0] = sim.rng.k_.a;
    out->rng_key[1] = sim.rng.k_.b;
    out->rng_count = sim.rng.count_;
    out->num_hiders = sim.numHiders;
    out->num_seekers = sim.numSeekers;
    out->num_boxes = (int32_t)sim.numActiveBoxes;
    out->num_ramps = (int32_t)sim.numActiveRamps;
    out->hider_team_reward = sim.hiderTeamReward.load_relaxed();
    out->seekers_first = ctx.singleton<TeamState>().seekersFirst ? 1 : 0;
    out->running_scores[0] = ctx.singleton<EpisodeStats>().runningScores[0];
    out->running_scores[1] = ctx.singleton<EpisodeStats>().runningScores[1];

    BodyIndex B;
    B.build(ctx);
    const int nb = B.n < HSREF_C_BODIES ? B.n : HSREF_C_BODIES;
    out->num_bodies = B.n;
    int nDyn = 0;
    for (int i = 0; i < nb; i++) {
        Entity e = B.e[i];
        if (B.arch[i] == 0) nDyn++;
        out->obj[i] = (uint8_t)ctx.get<ObjectID>(e).idx;
        out->response[i] = (uint8_t)ctx.get<ResponseType>(e);
        out->owner[i] = (uint8_t)ctx.get<OwnerTeam>(e);
        out->leaf[i] = (uint8_t)ctx.get<bp::LeafID>(e).id;
        int rank = 0;
        for (int o = 0; o < nb; o++) rank += B.e[o].id < e.id;
        out->eid_rank[i] = (uint8_t)rank;
        Vector3 p = ctx.get<Position>(e);
        Quat q = ctx.get<Rotation>(e);
        Diag3x3 s = ctx.get<Scale>(e);
        Velocity v = ctx.get<Velocity>(e);
        Vector3 f = ctx.get<ExternalForce>(e);
        Vector3 t = ctx.get<ExternalTorque>(e);
        float pose[7] = {p.x, p.y, p.z, q.w, q.x, q.y, q.z};
        memcpy(out->pose[i], pose, sizeof(pose));
        out->scale[i][0] = s.d0; out->scale[i][1] = s.d1; out->scale[i][2] = s.d2;
        float vel[6] = {v.linear.x, v.linear.y, v.linear.z, v.angular.x, v.angular.y, v.angular.z};
        memcpy(out->vel[i], vel, sizeof(vel));
        float ft[6] = {f.x, f.y, f.z, t.x, t.y, t.z};
        memcpy(out->force[i], ft, sizeof(ft));
    }
    out->num_dyn = nDyn;
    for (int k = 0; k < HSREF_MAX_AGENTS; k++) out->joint_other[k] = -1;
    for (int i = 0, k = 0; i < nb; i++) {
        if (B.arch[i] != 1) continue;
        if (k < HSREF_MAX_AGENTS) {
            Entity j = ctx.get<GrabData>(B.e[i]).constraintEntity;
            if (j != Entity::none()) {
                JointConstraint &jc = ctx.get<JointConstraint>(j);
                out->joint_other[k] = (int8_t)B.ofEntity(jc.e2);
                out->joint_sep[k] = jc.fixed.separation;
            }
        }
        k++;
    }

    bp::BVH &bvh = ctx.singleton<bp::BVH>();
    int nl = bvh.num_leaves_.load_relaxed();
    out->bvh_num_leaves = nl;
    out->bvh_num_nodes = (int32_t)bvh.num_nodes_;
    int nn = std::min<int>((int)bvh.num_nodes_, HSREF_C_NODES);
    for (int n = 0; n < nn; n++) {
        const hsref_node *src = (const hsref_node *)&bvh.nodes_[n];
        memcpy(out->node_bounds[n], src->bounds, sizeof(src->bounds));
        memcpy(out->node_children[n], src->children, sizeof(src->children));
        out->node_parent[n] = src->parent;
    }
    for (int i = 0; i < nl && i < HSREF_C_BODIES; i++) {
        math::AABB a = bvh.leaf_aabbs_[i];
        float v[6] = {a.pMin.x, a.pMin.y, a.pMin.z, a.pMax.x, a.pMax.y, a.pMax.z};
        memcpy(out->leaf_aabb[i], v, sizeof(v));
        out->leaf_parent[i] = bvh.leaf_parents_[i];
    }

    out->num_cand = -1;
    out->num_contacts = -1;
    if (!h->probe.empty()) {
        const ProbeData &P = h->probe[world];
        out->num_cand = P.numCand;
        out->num_contacts = P.numContacts;
        memcpy(out->cand, P.cand, sizeof(P.cand));
        memcpy(out->contact_ref, P.ref, sizeof(P.ref));
        memcpy(out->contact_alt, P.alt, sizeof(P.alt));
        memcpy(out->contact_npts, P.npts, sizeof(P.npts));
        memcpy(out->contact_pts, P.pts, sizeof(P.pts));
        memcpy(out->contact_normal, P.normal, sizeof(P.normal));
        memcpy(out->contact_lambda, P.lambda, sizeof(P.lambda));
        memcpy(out->lidar_hit, P.lidarHit, sizeof(P.lidarHit));
    }
}

int hsref_dump_compact(void *hp, int world0, int count, hsref_compact *out, int num_threads)
{
    Handle *h = (Handle *)hp;
    if (world0 < 0 || count <= 0 || world0 + count > h->numWorlds) return -1;
    auto keep = h->impl->cpuExec.getWorldContext(0).query<madrona::Entity, OwnerTeam, bp::LeafID>();
    int T = num_threads > 0 ? num_threads : hsref_affinity_cpus();
    if (T > count) T = count;
    std::atomic<int> next(0);
    auto worker = [&]() {
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= count) break;
            fillCompact(h, world0 + i, &out[i]);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < T; t++) pool.emplace_back(worker);
    worker();
    for (auto &t : pool) t.join();
    return 0;
}


// ---- mesh ray tracing with the reference's own code (agent-view oracle) ------------------------------
// The ray / triangle stage of the agent views — QBVH node slab tests in quantised space, the leaf
// loop and Woop's watertight ray-triangle test with back-face culling — is the reference's
// MeshBVH::traceRay / traceRayLeaf / rayTriangleIntersection (M/include/madrona/mesh_bvh.inl:79-168,
// 319-378, 380-462), compiled here unchanged.  The reference BUILDS its MeshBVH with Embree
// (M/src/common/mesh_bvh_builder.cpp: <embree4/rtcore.h>, an external library that is not on this
// machine), so the tree handed to it is built here: consecutive triangles grouped 8 per leaf, four
// children per node, every node quantised by the reference's own QBVHNode::construct
// (mesh_bvh.hpp).  The topology decides the order triangles are visited in, not which triangle is
// hit first.
namespace {

struct RefMesh {
    std::vector<madrona::MeshBVH::BVHVertex> verts;
    std::vector<madrona::QBVHNode> nodes;
    madrona::MeshBVH bvh;
};

madrona::math::AABB triRangeAABB(const RefMesh &m, int first, int count)
{
    using madrona::math::AABB;
    AABB box = AABB::point(m.verts[first * 3].pos);
    for (int i = first * 3; i < (first + count) * 3; i++) box.expand(m.verts[i].pos);
    return box;
}

// builds the subtree over triangles [first, first + count) into node `self`; returns its bounds
madrona::math::AABB buildRefMeshNode(RefMesh &m, int self, int first, int count)
{
    using madrona::math::AABB;
    constexpr int kLeafTris = 8, kWidth = 4;
    AABB childBox[kWidth];
    int32_t childIdx[kWidth];
    uint8_t triSize[kWidth] = {0, 0, 0, 0};
    int n = 0;
    if (count <= kLeafTris * kWidth) {
        for (int off = 0; off < count; off += kLeafTris, n++) {
            int c = std::min(kLeafTris, count - off);
            childBox[n] = triRangeAABB(m, first + off, c);
            childIdx[n] = -(first + off) - 1;              // leaf: index of its first triangle
            triSize[n] = (uint8_t)c;
        }
    } else {
        int per = (count + kWidth - 1) / kWidth;
        for (int off = 0; off < count; off += per, n++) {
            int c = std::min(per, count - off);
            int child = (int)m.nodes.size();
            m.nodes.emplace_back();
            childBox[n] = buildRefMeshNode(m, child, first + off, c);
            childIdx[n] = child + 1;
        }
    }
    madrona::QBVHNode node = madrona::QBVHNode::construct((uint32_t)n, childBox, childIdx);
    for (int i = 0; i < kWidth; i++) node.triSize[i] = triSize[i];
    m.nodes[self] = node;
    AABB box = childBox[0];
    for (int i = 1; i < n; i++) box = AABB::merge(box, childBox[i]);
    return box;
}

}

// tris: [n][9] object-space corners; uvs: [n][6] per-corner texture coordinates or NULL
void *hsref_mesh_create(const float *tris, const float *uvs, int n)
{
    if (n <= 0) return nullptr;
    RefMesh *m = new RefMesh();
    m->verts.resize((size_t)n * 3);
    for (int t = 0; t < n; t++)
        for (int k = 0; k < 3; k++) {
            auto &v = m->verts[(size_t)t * 3 + k];
            v.pos = {tris[t * 9 + k * 3], tris[t * 9 + k * 3 + 1], tris[t * 9 + k * 3 + 2]};
            v.uv = uvs ? madrona::math::Vector2{uvs[t * 6 + k * 2], uvs[t * 6 + k * 2 + 1]} : madrona::math::Vector2{0.f, 0.f};
        }
    m->nodes.reserve((size_t)n + 8);
    m->nodes.emplace_back();
    madrona::math::AABB root = buildRefMeshNode(*m, 0, 0, n);
    m->bvh = {};
    m->bvh.nodes = m->nodes.data();
    m->bvh.leafMats = nullptr;
    m->bvh.vertices = m->verts.data();
    m->bvh.rootAABB = root;
    m->bvh.numNodes = (uint32_t)m->nodes.size();
    m->bvh.numLeaves = (uint32_t)n;
    m->bvh.numVerts = (uint32_t)m->verts.size();
    return m;
}

void hsref_mesh_destroy(void *mp) { delete (RefMesh *)mp; }

// closest front-facing hit of the ray o + t d, t in [0, t_max]: out = {t, normal xyz, uv}, *tri = triangle
int hsref_mesh_trace(void *mp, const float *o, const float *d, float t_max, float *out, int *tri)
{
    RefMesh *m = (RefMesh *)mp;
    madrona::MeshBVH::HitInfo hit;
    int32_t stack[64];
    int32_t stackSize = 0;
    bool ok = m->bvh.traceRay({o[0], o[1], o[2]}, {d[0], d[1], d[2]}, &hit, stack, stackSize, t_max);
    if (!ok) return 0;
    out[0] = hit.tHit;
    out[1] = hit.normal.x; out[2] = hit.normal.y; out[3] = hit.normal.z;
    out[4] = hit.uv.x; out[5] = hit.uv.y;
    *tri = (int)hit.leafMaterialIDX;
    return 1;
}

int hsref_world_dump_bytes() { return (int)sizeof(hsref_world); }
int hsref_object_dump_bytes() { return (int)sizeof(hsref_object); }

// Processed asset tables (ObjectManager) — what every world reads at run time.
int hsref_dump_objects(void *hp, hsref_object *out /* [HSREF_NUM_OBJECTS] */)
{
    using namespace madrona;
    using namespace madrona::phys;

    Handle *h = (Handle *)hp;
    ObjectManager &om = h->impl->physicsLoader.getObjectManager();

    for (int o = 0; o < HSREF_NUM_OBJECTS; o++) {
        hsref_object &d = out[o];
        memset(&d, 0, sizeof(d));
        uint32_t prim_idx = om.rigidBodyPrimitiveOffsets[o];
        if (om.rigidBodyPrimitiveCounts[o] != 1) return -1;
        const CollisionPrimitive &p = om.collisionPrimitives[prim_idx];
        d.prim_type = (int32_t)p.type;
        math::AABB pa = om.primitiveAABBs[prim_idx];
        math::AABB oa = om.rigidBodyAABBs[o];
        float pav[6] = {pa.pMin.x, pa.pMin.y, pa.pMin.z, pa.pMax.x, pa.pMax.y, pa.pMax.z};
        float oav[6] = {oa.pMin.x, oa.pMin.y, oa.pMin.z, oa.pMax.x, oa.pMax.y, oa.pMax.z};
        memcpy(d.prim_aabb, pav, sizeof(pav));
        memcpy(d.obj_aabb, oav, sizeof(oav));
        const RigidBodyMetadata &md = om.metadata[o];
        d.inv_mass = md.mass.invMass;
        d.inv_inertia[0] = md.mass.invInertiaTensor.x;
        d.inv_inertia[1] = md.mass.invInertiaTensor.y;
        d.inv_inertia[2] = md.mass.invInertiaTensor.z;
        d.mu_s = md.friction.muS;
        d.mu_d = md.friction.muD;
        if (p.type == CollisionPrimitive::Type::Hull) {
            const geo::HalfEdgeMesh &m = p.hull.halfEdgeMesh;
            if (m.numVertices > HSREF_MAX_HULL_VERTS ||
                m.numFaces > HSREF_MAX_HULL_FACES ||
                m.numHalfEdges > HSREF_MAX_HULL_HEDGES) return -2;
            d.num_verts = m.numVertices;
            d.num_faces = m.numFaces;
            d.num_hedges = m.numHalfEdges;
            for (uint32_t i = 0; i < m.numVertices; i++) {
                d.verts[i][0] = m.vertices[i].x;
                d.verts[i][1] = m.vertices[i].y;
                d.verts[i][2] = m.vertices[i].z;
            }
            for (uint32_t i = 0; i < m.numFaces; i++) {
                d.planes[i][0] = m.facePlanes[i].normal.x;
                d.planes[i][1] = m.facePlanes[i].normal.y;
                d.planes[i][2] = m.facePlanes[i].normal.z;
                d.planes[i][3] = m.facePlanes[i].d;
                d.face_base_hedge[i] = m.faceBaseHalfEdges[i];
            }
            for (uint32_t i = 0; i < m.numHalfEdges; i++) {
                d.hedges[i][0] = m.halfEdges[i].next;
                d.hedges[i][1] = m.halfEdges[i].rootVertex;
                d.hedges[i][2] = m.halfEdges[i].face;
            }
        }
    }
    return 0;
}

// Wall-clock timing of K reference steps with uniform random move actions,
// mirroring scripts/benchmark.py:74-88 (x,y <- randint(-5,5) every step).
double hsref_bench(void *hp, int warmup, int steps, uint32_t action_seed)
{
    Handle *h = (Handle *)hp;
    int32_t *actions = (int32_t *)hsref_tensor(hp, (int)ExportID::Action);
    int rows = h->numWorlds * h->agentsPerWorld;
    uint64_t s = 0x9E3779B97F4A7C15ull ^ action_seed;
    auto next = [&]() {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        return (uint32_t)(s >> 32);
    };
    auto fill = [&]() {
        for (int i = 0; i < rows; i++) {
            actions[i * 5 + 0] = (int32_t)(next() % 10) - 5;
            actions[i * 5 + 1] = (int32_t)(next() % 10) - 5;
        }
    };
    for (int i = 0; i < warmup; i++) { fill(); h->mgr->step(); }
    auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) { fill(); h->mgr->step(); }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}


// ---- known-answer entry points: the reference's own RNG / math routines, so that the
// tests can replay the reference's unit-test vectors (M/tests/rand.cpp:131-140,
// M/tests/math.cpp:23-47) against this build and against the CUDA library's host-compiled
// headers. ---------------------------------------------------------------------------------
void hsref_rand_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t idx_upper, uint32_t *out)
{
    madrona::RandKey k = madrona::rand::split_i(madrona::RandKey{a, b}, idx, idx_upper);
    out[0] = k.a; out[1] = k.b;
}

void hsref_rand_init_key(uint32_t seed, uint32_t *out)
{
    madrona::RandKey k = madrona::rand::initKey(seed);
    out[0] = k.a; out[1] = k.b;
}

uint32_t hsref_rand_bits32(uint32_t a, uint32_t b) { return madrona::rand::bits32(madrona::RandKey{a, b}); }

int32_t hsref_rand_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi)
{
    return madrona::rand::sampleI32(madrona::RandKey{a, b}, lo, hi);
}

float hsref_rand_sample_uniform(uint32_t a, uint32_t b)
{
    return madrona::rand::sampleUniform(madrona::RandKey{a, b});
}

// RNG(seed) stream: n draws of sampleUniform (even slots) / sampleI32(lo,hi) (odd slots)
void hsref_rng_stream(uint32_t seed, int n, int32_t lo, int32_t hi, float *uni, int32_t *ints)
{
    madrona::RNG rng(seed);
    for (int i = 0; i < n; i++) {
        uni[i] = rng.sampleUniform();
        ints[i] = rng.sampleI32(lo, hi);
    }
}

void hsref_quat_angle_axis(float angle, const float *axis, float *out)
{
    madrona::math::Quat q = madrona::math::Quat::angleAxis(angle, {axis[0], axis[1], axis[2]});
    out[0] = q.w; out[1] = q.x; out[2] = q.y; out[3] = q.z;
}

void hsref_quat_mul(const float *a, const float *b, float *out)
{
    madrona::math::Quat q = madrona::math::Quat{a[0], a[1], a[2], a[3]} *
                            madrona::math::Quat{b[0], b[1], b[2], b[3]};
    out[0] = q.w; out[1] = q.x; out[2] = q.y; out[3] = q.z;
}

void hsref_quat_rotate(const float *q, const float *v, float *out)
{
    madrona::math::Vector3 r = madrona::math::Quat{q[0], q[1], q[2], q[3]}.rotateVec({v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

}
