/* Flat per-world internal state returned by hs_dump_world (parity tests and
 * debugging only; not on the step path).  All members are 32-bit.
 * tests/hsdump.py mirrors this as a numpy dtype.
 */
#ifndef HIDESEEK_DEBUG_DUMP_H
#define HIDESEEK_DEBUG_DUMP_H

#include <stdint.h>

#define HS_DUMP_MAX_BODIES 52
#define HS_DUMP_MAX_NODES 69
#define HS_DUMP_MAX_AGENTS 6
#define HS_DUMP_MAX_CAND 256

typedef struct {
    int32_t eid, obj, response, owner, leaf, pad0, pad1, pad2;
    float pos[3];
    float rot[4];
    float scale[3];
    float vel[6];
    float ext_force[3];
    float ext_torque[3];
} hs_dump_body;

typedef struct {
    int32_t active, other_slot, eid, seq;
    float attach1[4], attach2[4];
    float separation;
    float r1[3], r2[3];
} hs_dump_joint;

typedef struct {
    float bounds[24];
    int32_t children[4];
    int32_t parent;
} hs_dump_node;

typedef struct {
    int32_t cur_episode_step, cur_world_episode;
    uint32_t episode_key[2];
    uint32_t rng_key[2];
    uint32_t rng_count;
    int32_t num_hiders, num_seekers, num_agents;
    int32_t num_dyn, num_bodies, num_boxes, num_ramps;
    float hider_team_reward;
    int32_t seekers_first;
    int32_t running_scores[2];
    int32_t num_cand;
    int32_t box_slot[9], ramp_slot[2], hider_slot[3], seeker_slot[3];
    int32_t iface_slot[HS_DUMP_MAX_AGENTS], iface_type[HS_DUMP_MAX_AGENTS];

    hs_dump_body bodies[HS_DUMP_MAX_BODIES];
    hs_dump_joint joints[HS_DUMP_MAX_AGENTS];

    int32_t bvh_num_nodes, bvh_num_leaves, bvh_force_rebuild;
    hs_dump_node nodes[HS_DUMP_MAX_NODES];
    float leaf_aabbs[HS_DUMP_MAX_BODIES][6];
    uint32_t leaf_parents[HS_DUMP_MAX_BODIES];
    int32_t leaf_slot[HS_DUMP_MAX_BODIES];

    int32_t cand_a[HS_DUMP_MAX_CAND], cand_b[HS_DUMP_MAX_CAND];
    int32_t contact_meta[HS_DUMP_MAX_CAND];      /* ref | alt<<8 | npts<<16 */
    float contacts[HS_DUMP_MAX_CAND][20];
} hs_world_dump;

/* Compact per-world record for BULK parity runs (hs_dump_compact fills `count` of them in one
 * call; the oracle harness fills the same layout from the reference, oracle/ref_dump.h).  Body
 * arrays are indexed by body slot = ECS iteration order (DynamicObject rows, then DynAgent rows).
 * The probe part holds what a step leaves behind that no export shows: the broadphase candidate
 * pairs in emission order, the contacts of the step's LAST substep in creation order (after the
 * solver, so lambda is the normal multiplier it stored) and the body each lidar ray hit. */
#define HS_COMPACT_MAX_CONTACTS 96

typedef struct {
    int32_t cur_episode_step, cur_world_episode;
    uint32_t rng_key[2];
    uint32_t rng_count;
    int32_t num_hiders, num_seekers, num_boxes, num_ramps;
    float hider_team_reward;
    int32_t seekers_first;
    int32_t running_scores[2];
    int32_t num_bodies, num_dyn;
    int32_t bvh_num_leaves, bvh_num_nodes;
    int32_t num_cand, num_contacts;              /* num_contacts may exceed the stored rows */
    uint8_t obj[HS_DUMP_MAX_BODIES], response[HS_DUMP_MAX_BODIES], owner[HS_DUMP_MAX_BODIES],
            leaf[HS_DUMP_MAX_BODIES];
    uint8_t eid_rank[HS_DUMP_MAX_BODIES];        /* rank of the body's entity id inside the world */
    int8_t joint_other[HS_DUMP_MAX_AGENTS];      /* DynAgent row k: slot of the grabbed body, -1 none */
    int8_t pad0[2];
    float joint_sep[HS_DUMP_MAX_AGENTS];
    float pose[HS_DUMP_MAX_BODIES][7];           /* pos3, quat wxyz */
    float scale[HS_DUMP_MAX_BODIES][3];
    float vel[HS_DUMP_MAX_BODIES][6];
    float force[HS_DUMP_MAX_BODIES][6];          /* ExternalForce, ExternalTorque */
    float node_bounds[HS_DUMP_MAX_NODES][24];
    int32_t node_children[HS_DUMP_MAX_NODES][4];
    int32_t node_parent[HS_DUMP_MAX_NODES];
    float leaf_aabb[HS_DUMP_MAX_BODIES][6];
    uint32_t leaf_parent[HS_DUMP_MAX_BODIES];
    uint8_t cand[HS_DUMP_MAX_CAND][2];           /* body slots a, b */
    uint8_t contact_ref[HS_COMPACT_MAX_CONTACTS], contact_alt[HS_COMPACT_MAX_CONTACTS],
            contact_npts[HS_COMPACT_MAX_CONTACTS];
    float contact_pts[HS_COMPACT_MAX_CONTACTS][4][4];
    float contact_normal[HS_COMPACT_MAX_CONTACTS][3];
    float contact_lambda[HS_COMPACT_MAX_CONTACTS];
    int8_t lidar_hit[HS_DUMP_MAX_AGENTS][30];    /* body slot, -1 miss, -2 inactive interface */
} hs_compact_world;

#endif
