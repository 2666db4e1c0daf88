/* TEST INFRASTRUCTURE ONLY.
 * Flat C structs the reference harness (oracle/ref_harness.cpp) fills when a
 * test asks the UNMODIFIED reference CPU backend for one world's internal
 * state.  tests/ mirror them as numpy dtypes (tests/refdump.py).  All fields
 * are 32-bit so the layout is packing-independent.
 */
#ifndef HS_REF_DUMP_H
#define HS_REF_DUMP_H

#include <stdint.h>

#define HSREF_MAX_BODIES 64
#define HSREF_MAX_NODES 128
#define HSREF_MAX_JOINTS 16
#define HSREF_MAX_AGENTS 6

typedef struct {
    int32_t entity_id;      /* Entity.id (global id-map index) */
    int32_t archetype;      /* 0 = DynamicObject row, 1 = DynAgent row */
    int32_t row;            /* row in its per-world table */
    int32_t obj_id;         /* SimObject */
    int32_t response;       /* ResponseType */
    int32_t owner;          /* OwnerTeam */
    int32_t leaf_id;
    int32_t grab_joint;     /* joint Entity.id, -1 none / not an agent */
    float pos[3];
    float rot[4];           /* w x y z */
    float scale[3];
    float vel[6];           /* linear, angular */
    float ext_force[3];
    float ext_torque[3];
} hsref_body;

typedef struct {
    int32_t entity_id;
    int32_t e1, e2;         /* Entity.id of the two bodies */
    float attach1[4], attach2[4];
    float separation;
    float r1[3], r2[3];
} hsref_joint;

typedef struct {
    float bounds[24];       /* minX[4] minY[4] minZ[4] maxX[4] maxY[4] maxZ[4] */
    int32_t children[4];
    int32_t parent;
} hsref_node;

typedef struct {
    /* Sim bookkeeping (src/sim.hpp:326-362) */
    int32_t cur_episode_step;
    int32_t cur_world_episode;
    uint32_t episode_key[2];        /* curEpisodeRNDCounter */
    uint32_t rng_key[2];
    uint32_t rng_count;
    int32_t num_hiders, num_seekers, num_active_agents;
    int32_t num_obstacles, num_boxes, num_ramps;
    float hider_team_reward;
    int32_t seekers_first;
    int32_t running_scores[2];
    int32_t obstacles[HSREF_MAX_BODIES];   /* Entity.id in obstacles[] order */
    int32_t boxes[9];
    int32_t ramps[2];
    int32_t hiders[3];
    int32_t seekers[3];
    int32_t agent_iface_sim[HSREF_MAX_AGENTS];  /* SimEntity.e.id per interface slot */
    float box_sizes[9][3];

    /* rigid bodies in ECS iteration order: DynamicObject rows, then DynAgent rows */
    int32_t num_bodies;
    hsref_body bodies[HSREF_MAX_BODIES];

    int32_t num_joints;
    hsref_joint joints[HSREF_MAX_JOINTS];

    /* BVH */
    int32_t bvh_num_leaves;
    int32_t bvh_num_nodes;
    int32_t bvh_force_rebuild;
    hsref_node nodes[HSREF_MAX_NODES];
    float leaf_aabbs[HSREF_MAX_BODIES][6];
    uint32_t leaf_parents[HSREF_MAX_BODIES];
    int32_t leaf_entities[HSREF_MAX_BODIES];
    int32_t leaf_obj_ids[HSREF_MAX_BODIES];
} hsref_world;

/* Object table dump (ObjectManager after processRigidBodyAssets). */
#define HSREF_NUM_OBJECTS 8
#define HSREF_MAX_HULL_VERTS 8
#define HSREF_MAX_HULL_FACES 6
#define HSREF_MAX_HULL_HEDGES 24

typedef struct {
    int32_t prim_type;      /* 1 sphere, 2 hull, 4 plane */
    int32_t num_verts, num_faces, num_hedges;
    float prim_aabb[6];
    float obj_aabb[6];
    float inv_mass;
    float inv_inertia[3];
    float mu_s, mu_d;
    float verts[HSREF_MAX_HULL_VERTS][3];
    float planes[HSREF_MAX_HULL_FACES][4];
    uint32_t face_base_hedge[HSREF_MAX_HULL_FACES];
    uint32_t hedges[HSREF_MAX_HULL_HEDGES][3];   /* next, rootVertex, face */
} hsref_object;

/* Compact per-world record for bulk parity runs: the same layout as hs_compact_world
 * (include/hideseek/debug_dump.h), restated here so that nothing under oracle/ depends on the
 * product tree; tests/compact.py asserts the two sizes agree.  Body arrays are indexed by ECS
 * iteration order.  cand / contact_* / lidar_hit are filled by hsref_step_probe. */
#define HSREF_C_BODIES 52
#define HSREF_C_NODES 69
#define HSREF_C_CAND 256
#define HSREF_C_CONTACTS 96

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
    int32_t num_cand, num_contacts;
    uint8_t obj[HSREF_C_BODIES], response[HSREF_C_BODIES], owner[HSREF_C_BODIES], leaf[HSREF_C_BODIES];
    uint8_t eid_rank[HSREF_C_BODIES];
    int8_t joint_other[HSREF_MAX_AGENTS];
    int8_t pad0[2];
    float joint_sep[HSREF_MAX_AGENTS];
    float pose[HSREF_C_BODIES][7];
    float scale[HSREF_C_BODIES][3];
    float vel[HSREF_C_BODIES][6];
    float force[HSREF_C_BODIES][6];
    float node_bounds[HSREF_C_NODES][24];
    int32_t node_children[HSREF_C_NODES][4];
    int32_t node_parent[HSREF_C_NODES];
    float leaf_aabb[HSREF_C_BODIES][6];
    uint32_t leaf_parent[HSREF_C_BODIES];
    uint8_t cand[HSREF_C_CAND][2];
    uint8_t contact_ref[HSREF_C_CONTACTS], contact_alt[HSREF_C_CONTACTS], contact_npts[HSREF_C_CONTACTS];
    float contact_pts[HSREF_C_CONTACTS][4][4];
    float contact_normal[HSREF_C_CONTACTS][3];
    float contact_lambda[HSREF_C_CONTACTS];
    int8_t lidar_hit[HSREF_MAX_AGENTS][30];
} hsref_compact;

#endif
