// GPUHideSeek::Manager — host-side owner of one batch simulator, API-compatible with the
// reference's Manager (src/mgr.hpp:14-99): same Config fields, init()/step(), the same 21
// tensor getters, triggerReset / setAction.  Where the reference's Impl holds a Madrona
// executor (src/mgr.cpp:100-437), this one holds an `hs_sim*` from the C ABI
// (include/hideseek/c_api.h) — hand-written sm_100a kernels, no CPU execution mode.
//
// Differences, all forced by the environment (DESIGN.md "Boundary"):
//   * execMode must be ExecMode::CUDA: there is no CPU backend and no fallback.
//   * enableBatchRenderer = true turns on a CUDA ray caster for the per-agent views
//     (depthTensor()/rgbTensor(), csrc/hs_views.cuh) instead of the Vulkan rasteriser;
//     extRenderAPI/extRenderDev and getRenderManager() abort (graphics APIs are unavailable
//     on this box).  Config::renderDataDir (extension): the reference's data/ directory with
//     the render OBJ meshes; nullptr = $HIDESEEK_DATA_DIR, else built-in meshes.
//   * Config::worldBase (extension, default 0): global index of this Manager's world 0, used
//     to key the episode RNG when worlds are sharded across GPUs (one Manager per GPU).
#pragma once

#include <cstdint>

#include "madrona_py.hpp"
#include "sim_flags.hpp"

namespace madrona::render { struct APIBackend; struct GPUDevice; class RenderManager; }

// the CUDA runtime's own typedef, restated so this header needs no CUDA include
typedef struct CUstream_st *cudaStream_t;

namespace GPUHideSeek {

class Manager {
public:
    using Tensor = madrona::py::Tensor;      // non-owning device tensor descriptor
    using CountT = madrona::CountT;

    struct Config {      // same members, types and order as the reference's Manager::Config
        madrona::ExecMode execMode;  int gpuID;  uint32_t numWorlds;  SimFlags simFlags;  uint32_t randSeed;
        uint32_t minHiders, maxHiders, minSeekers, maxSeekers, numPBTPolicies;
        bool enableBatchRenderer;
        uint32_t batchRenderViewWidth = 64, batchRenderViewHeight = 64;
        madrona::render::APIBackend *extRenderAPI = nullptr;  madrona::render::GPUDevice *extRenderDev = nullptr;
        uint32_t worldBase = 0;      // extension: first global world index of this shard
        const char *renderDataDir = nullptr;     // extension: render OBJ directory for the agent views
    };


    Manager(const Config &cfg);
    ~Manager();
    Manager(const Manager &) = delete;
    Manager &operator=(const Manager &) = delete;

    void init();
    void step();     // synchronous, like the reference's graph launch + stream sync

    void saveCheckpoint(CountT world_idx);
    void loadCheckpoint(CountT world_idx);
    void loadCheckpoints();

    // XLA custom-call bodies (src/mgr.hpp:44-50, src/mgr.cpp:365-437): `buffers` = device
    // pointers in the order documented at hs_jax_* in include/hideseek/c_api.h; everything is
    // enqueued on `strm`, only gpuJAXInit waits for it.
    void gpuJAXInit(cudaStream_t strm, void **buffers);
    void gpuJAXStep(cudaStream_t strm, void **buffers);
    void gpuJAXSaveCheckpoints(cudaStream_t strm, void **buffers);
    void gpuJAXLoadCheckpoints(cudaStream_t strm, void **buffers);

    Tensor resetTensor() const;
    Tensor doneTensor() const;
    Tensor prepCounterTensor() const;
    Tensor actionTensor() const;
    Tensor rewardTensor() const;
    Tensor selfDataTensor() const;
    Tensor selfTypeTensor() const;
    Tensor selfMaskTensor() const;
    Tensor agentDataTensor() const;
    Tensor boxDataTensor() const;
    Tensor rampDataTensor() const;
    Tensor visibleAgentsMaskTensor() const;
    Tensor visibleBoxesMaskTensor() const;
    Tensor visibleRampsMaskTensor() const;
    Tensor globalPositionsTensor() const;
    Tensor lidarTensor() const;
    Tensor seedTensor() const;

    Tensor checkpointControlTensor() const;
    Tensor checkpointTensor() const;

    Tensor depthTensor() const;
    Tensor rgbTensor() const;

    void triggerReset(CountT world_idx, CountT level_idx);
    void setAction(CountT agent_idx, int32_t x, int32_t y, int32_t r, bool g, bool l);

    madrona::render::RenderManager &getRenderManager();

    Tensor policyAssignmentsTensor() const;
    Tensor episodeResultTensor() const;
    Tensor simControlTensor() const;       // empty, as in the reference

    madrona::py::TrainInterface trainInterface() const;

    // ---- extensions (not in the reference) -------------------------------------------------
    // Asynchronous step on a caller stream (cudaStream_t as void*) + explicit wait; the
    // reference's JAX entry points are asynchronous in the same way (src/mgr.cpp:379-398).
    void stepAsync(void *stream);
    void sync(void *stream);
    // Host-buffer step: H2D actions, step, D2H reward+done, all on `stream`.
    void stepHost(void *stream, const int32_t *actions, float *reward, int32_t *done);
    void *handle() const;            // the underlying hs_sim*
    uint32_t numWorlds() const;
    uint32_t agentsPerWorld() const;
    int gpuID() const;

private:
    struct Impl;
    Impl *impl_;
};

}  // namespace GPUHideSeek
