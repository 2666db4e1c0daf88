/* hideseek B200 — C ABI of the CUDA batch simulator.
 *
 * This is the boundary the reference's `Manager` (src/mgr.hpp:14-99) is
 * re-pointed at: where the reference holds a Madrona executor
 * (CPU `TaskGraphExecutor` / GPU `MWCudaExecutor`, src/mgr.cpp:100-437) and asks
 * it for `getExported(slot)` pointers (src/mgr.cpp:824-842), the drop-in holds an
 * `hs_sim*` and calls the functions below.  Plain pointers and sizes only; no
 * torch / nanobind types.  All functions return 0 on success, non-zero on
 * failure with a message available from hs_last_error() (the reference aborts
 * via FATAL/REQ_CUDA instead; the Manager layer turns errors back into aborts).
 *
 * One hs_sim owns one shard of worlds on one CUDA device.  Worlds are
 * independent, so multi-GPU runs create one hs_sim per device with
 * `world_base` set to the shard's first global world index (the episode RNG
 * is keyed by global world index, src/sim.cpp:105-114).
 */
#ifndef HIDESEEK_C_API_H
#define HIDESEEK_C_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hs_sim hs_sim;

/* Mirrors Manager::Config (src/mgr.hpp:16-32); renderer fields are absent
 * because the graphics path is unavailable on this box. */
typedef struct hs_config {
    int32_t gpu_id;
    int32_t num_worlds;       /* worlds on this shard */
    int32_t world_base;       /* global index of this shard's world 0 */
    uint32_t sim_flags;       /* SimFlags bitmask, src/sim_flags.hpp:7-13 */
    uint32_t rand_seed;
    int32_t min_hiders, max_hiders;
    int32_t min_seekers, max_seekers;
    int32_t num_pbt_policies;
} hs_config;

/* ExportID numbering is the reference's (src/sim.hpp:45-68). */
enum hs_export_id {
    HS_EXPORT_RESET = 0, HS_EXPORT_PREP_COUNTER, HS_EXPORT_ACTION, HS_EXPORT_SELF_OBS,
    HS_EXPORT_SELF_TYPE, HS_EXPORT_SELF_MASK, HS_EXPORT_AGENT_OBS, HS_EXPORT_BOX_OBS,
    HS_EXPORT_RAMP_OBS, HS_EXPORT_AGENT_VIS, HS_EXPORT_BOX_VIS, HS_EXPORT_RAMP_VIS,
    HS_EXPORT_LIDAR, HS_EXPORT_SEED, HS_EXPORT_REWARD, HS_EXPORT_DONE,
    HS_EXPORT_GLOBAL_POS, HS_EXPORT_AGENT_POLICY, HS_EXPORT_EPISODE_RESULT,
    HS_EXPORT_CKPT_CTRL, HS_EXPORT_CKPT, HS_NUM_EXPORTS,
    /* not ExportIDs in the reference (they come from its render manager,
     * src/mgr.cpp:1241-1263); valid after hs_enable_views */
    HS_EXPORT_VIEW_DEPTH = HS_NUM_EXPORTS, HS_EXPORT_VIEW_RGB
};

/* TensorElementType subset (M/include/madrona/py/utils.hpp) */
enum hs_dtype { HS_U8 = 0, HS_I32 = 1, HS_F32 = 2 };

/* Manager::Manager / ~Manager (src/mgr.cpp:844-859) */
int hs_create(const hs_config *cfg, hs_sim **out);
void hs_destroy(hs_sim *sim);

/* Manager::init (src/mgr.cpp:861-881): runs the Init task graph
 * (reset every world, build BVHs, first observations).  `stream` is a
 * cudaStream_t (0 = legacy default stream); the call is asynchronous. */
int hs_init(hs_sim *sim, void *stream);

/* Manager::step (src/mgr.cpp:883-903): one 1/30 s tick of every world on
 * `stream`, asynchronous.  hs_sync waits for it. */
int hs_step(hs_sim *sim, void *stream);
int hs_sync(hs_sim *sim, void *stream);

/* Manager::*Tensor() getters (src/mgr.cpp:1062-1331): borrowed device pointer,
 * stable until hs_destroy.  dims has room for 4 entries. */
int hs_export(hs_sim *sim, int export_id, void **dev_ptr, int *dtype, int *ndim, int64_t *dims);

/* Manager::triggerReset / setAction (src/mgr.cpp:1265-1310): host-side pokes,
 * synchronous H2D writes like the reference's cudaMemcpy. */
int hs_trigger_reset(hs_sim *sim, int world, int level);
int hs_set_action(hs_sim *sim, int agent_row, int x, int y, int r, int g, int l);

/* Host-buffer convenience used by the end-to-end path: copy a full action
 * tensor [num_worlds*A,5] i32 from (pinned) host memory, step, and copy
 * reward/done back.  Every copy is issued on `stream`. */
int hs_step_host(hs_sim *sim, void *stream, const int32_t *actions_host,
                 float *reward_host, int32_t *done_host);

/* Manager::saveCheckpoint / loadCheckpoint / loadCheckpoints (src/mgr.cpp:905-985) and the
 * SaveCheckpoints / LoadCheckpoints task graphs (src/sim.cpp:956-1137, 1315-1344).
 * hs_save_checkpoint / hs_load_checkpoint set CheckpointControl.trigger of one world and run the
 * corresponding graph synchronously; the *_checkpoints variants run the graph over every
 * world whose trigger is already set in the exported CheckpointControl tensor (the JAX path,
 * src/mgr.cpp:400-437), asynchronously on `stream`.  The Checkpoint tensor keeps the
 * reference's byte layout (src/sim.hpp:283-313). */
int hs_save_checkpoint(hs_sim *sim, int world);
int hs_load_checkpoint(hs_sim *sim, int world);
int hs_save_checkpoints(hs_sim *sim, void *stream);
int hs_load_checkpoints(hs_sim *sim, void *stream);

/* The JAX / XLA path (Manager::gpuJAXInit / gpuJAXStep / gpuJAXSaveCheckpoints /
 * gpuJAXLoadCheckpoints, src/mgr.cpp:365-437, 987-1060).  `buffers` is the flat list of DEVICE
 * pointers of an XLA GPU custom call after the two entries Madrona's wrapper skips:
 *   init:  out 11 observations in JAXIOObservations order (src/mgr.cpp:168-201):
 *          prep_counter, self_data, self_type, self_mask, lidar, agent_data, box_data,
 *          ramp_data, vis_agents_mask, vis_boxes_mask, vis_ramps_mask      (waits for the stream)
 *   step:  in actions, resets, policy_assignments; out 11 observations, rewards, dones,
 *          episode_results                                                  (fully asynchronous)
 *   save:  in checkpoint_control; out checkpoint_data
 *   load:  in checkpoint_control, checkpoint_data; out 11 observations
 * Copies are device-to-device on `stream`; shapes/dtypes are those of hs_export. */
int hs_jax_init(hs_sim *sim, void *stream, void **buffers);
int hs_jax_step(hs_sim *sim, void *stream, void **buffers);
int hs_jax_save_checkpoints(hs_sim *sim, void *stream, void **buffers);
int hs_jax_load_checkpoints(hs_sim *sim, void *stream, void **buffers);

/* The entry points an XLA GPU custom-call registration is given (JAXInterface::gpuEntryFn,
 * external/madrona/include/madrona/py/bindings.inl:107-119): `opaque` carries the hs_sim*,
 * buffers[0] is the CPU-backend slot and buffers[1] JAX's ordering token (both skipped).
 * They abort on failure, as the reference's REQ_CUDA does. */
void hs_xla_gpu_init(void *stream, void **buffers, const char *opaque, size_t opaque_len);
void hs_xla_gpu_step(void *stream, void **buffers, const char *opaque, size_t opaque_len);
void hs_xla_gpu_save_checkpoints(void *stream, void **buffers, const char *opaque, size_t opaque_len);
void hs_xla_gpu_load_checkpoints(void *stream, void **buffers, const char *opaque, size_t opaque_len);

/* Agent camera views = the batch renderer's outputs (Config::enableBatchRenderer,
 * batchRenderViewWidth/Height, src/mgr.hpp:27-29; RenderManager::batchRender after init and
 * every step, src/mgr.cpp:876-880, 898-902; depthTensor / rgbTensor, src/mgr.cpp:1241-1263).
 * The Vulkan rasteriser is unavailable here; this is a CUDA ray caster with the arithmetic of
 * Madrona's software ray caster (M/src/mw/device/bvh_raycast.cpp) — see csrc/hs_views.cuh.
 *   hs_enable_views: call once between hs_create and hs_init.  `data_dir` = the reference's
 *     data/ directory (render OBJ meshes and the three PNG textures, src/mgr.cpp:590-665);
 *     NULL/"" = $HIDESEEK_DATA_DIR, else built-in meshes without textures.  Allocates [N*A,H,W,1] f32 depth and [N*A,H,W,4] u8 RGBA.
 *   hs_render_views: renders every agent slot's view of the current state on `stream`
 *     (asynchronous); the Manager calls it after hs_init / hs_step like the reference.
 *   hs_export(HS_EXPORT_VIEW_DEPTH / HS_EXPORT_VIEW_RGB) returns the two tensors.
 *   hs_build_view_mesh_table: parity/diagnostic, host only — the render-mesh table
 *     (hs::ViewMeshTable, csrc/hs_views.cuh) hs_enable_views builds for `data_dir`. */
int hs_enable_views(hs_sim *sim, int width, int height, const char *data_dir);
int hs_render_views(hs_sim *sim, void *stream);
int hs_view_mesh_table_bytes(void);
int hs_build_view_mesh_table(const char *data_dir, void *out, int out_bytes);
int hs_view_tex_table_bytes(void);       /* hs::ViewTexTable: per-triangle UVs, material -> texture */
int hs_build_view_tex_table(const char *data_dir, void *out, int out_bytes);
/* host only: the PNG reader used for the view textures (8-bit RGB / RGBA, non-interlaced) */
int hs_decode_png(const char *path, int *width, int *height, uint8_t *out, int64_t cap);

/* host only: the collision asset table (hs::AssetTable, csrc/hs_state.cuh — hulls as half-edge
 * meshes with face planes, AABBs, inverse mass / inertia, friction for the 8 SimObjects) that
 * hs_create derives from the `*_collision.obj` files of `data_dir` (the reference's data/
 * directory; hs_create itself reads $HIDESEEK_DATA_DIR), or from the built-in hull descriptions when
 * data_dir is NULL / "".  Replaces loadPhysicsObjects + processRigidBodyAssets
 * (src/mgr.cpp:441-588, M/src/physics/physics_assets.cpp:1268). */
int hs_asset_table_bytes(void);
int hs_dump_assets(const char *data_dir, void *out, int out_bytes);

/* Diagnostics / parity: flat per-world internal state (layout in
 * hideseek/debug_dump.h), number of kernels launched by the last hs_step,
 * capacity-overflow counter. */
int hs_dump_world(hs_sim *sim, int world, void *out, int out_bytes);
int hs_world_dump_bytes(void);
/* Bulk parity dump: `count` hs_compact_world records (debug_dump.h) of worlds [world0, world0 +
 * count) into host memory.  hs_set_debug(flags): bit 0 makes the ray stage record which body
 * every lidar ray hit (reported in the compact record); takes effect from the next hs_step. */
int hs_compact_dump_bytes(void);
int hs_dump_compact(hs_sim *sim, int world0, int count, void *out, int64_t out_bytes);
int hs_set_debug(hs_sim *sim, unsigned flags);
/* Allocation-free variant of hs_dump_compact.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments and same records as
 * hs_dump_compact, but the device-side staging scratch is a static device array of the library
 * (hs_dump_compact_noalloc_max() = 64 records, 1.5 MB held by every GPU that loads the library)
 * instead of a cudaMalloc'ed buffer: the call makes no cudaMalloc / cudaFree / allocating library
 * call.  count above the maximum is rejected with an error.  All calls share the array (it is
 * cleared as far as a call uses it): calls that overlap on the GPU race — serialise them. */
int hs_dump_compact_noalloc_max(void);
int hs_dump_compact_noalloc(hs_sim *sim, int world0, int count, void *out, int64_t out_bytes);
int hs_last_step_launches(hs_sim *sim);
int hs_overflow_count(hs_sim *sim, uint32_t *out);

/* Measurement only (bench.py roofline): one step launched kernel by kernel with a
 * CUDA event after each launch on `stream`; per-launch milliseconds in ms_out. */
int hs_step_profiled(hs_sim *sim, void *stream, float *ms_out, int cap, int *n_out);
const char *hs_profiled_kernel_name(int i);
/* totals over the shard: live bodies, active agents, candidates, contacts */
int hs_world_stats(hs_sim *sim, int64_t *out4);

const char *hs_last_error(void);

#ifdef __cplusplus
}
#endif

#endif
