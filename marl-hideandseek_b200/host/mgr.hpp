// GPUHideSeek::Manager — host-side owner of one batch simulator, API-compatible with the
// reference's Manager (src/mgr.hpp:14-99): same Config fields, init()/step(), the same 21
// tensor getters, triggerReset / setAction.  Where the reference's Impl holds a Madrona
// executor (src/mgr.cpp:100-437), this one holds an `hs_sim*` from the C ABI
// (include/hideseek/c_api.h) — hand-written sm_100a kernels, no CPU execution mode.
//
// Differences, all forced by the environment (DESIGN.md "Boundary"):
//   * execMode must be ExecMode::CUDA: there is no CPU backend and no fallback.
//   * enableBatchRenderer must be false; depthTensor()/rgbTensor()/getRenderManager() abort
//     (graphics APIs are unavailable on this box).
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
    struct Config {
        madrona::ExecMode execMode;
        int gpuID;
        uint32_t numWorlds;
        SimFlags simFlags;
        uint32_t randSeed;
        uint32_t minHiders;
        uint32_t maxHiders;
        uint32_t minSeekers;
        uint32_t maxSeekers;
        uint32_t numPBTPolicies;
        bool enableBatchRenderer;
        uint32_t batchRenderViewWidth = 64;
        uint32_t batchRenderViewHeight = 64;
        madrona::render::APIBackend *extRenderAPI = nullptr;
        madrona::render::GPUDevice *extRenderDev = nullptr;
        uint32_t worldBase = 0;      // extension: first global world index of this shard
    };

    Manager(const Config &cfg);
    ~Manager();
    Manager(const Manager &) = delete;
    Manager &operator=(const Manager &) = delete;

    void init();
    void step();     // synchronous, like the reference's graph launch + stream sync

    void saveCheckpoint(madrona::CountT world_idx);
    void loadCheckpoint(madrona::CountT world_idx);
    void loadCheckpoints();

    // XLA custom-call bodies (src/mgr.hpp:44-50, src/mgr.cpp:365-437): `buffers` = device
    // pointers in the order documented at hs_jax_* in include/hideseek/c_api.h; everything is
    // enqueued on `strm`, only gpuJAXInit waits for it.
    void gpuJAXInit(cudaStream_t strm, void **buffers);
    void gpuJAXStep(cudaStream_t strm, void **buffers);
    void gpuJAXSaveCheckpoints(cudaStream_t strm, void **buffers);
    void gpuJAXLoadCheckpoints(cudaStream_t strm, void **buffers);

    madrona::py::Tensor resetTensor() const;
    madrona::py::Tensor doneTensor() const;
    madrona::py::Tensor prepCounterTensor() const;
    madrona::py::Tensor actionTensor() const;
    madrona::py::Tensor rewardTensor() const;
    madrona::py::Tensor selfDataTensor() const;
    madrona::py::Tensor selfTypeTensor() const;
    madrona::py::Tensor selfMaskTensor() const;
    madrona::py::Tensor agentDataTensor() const;
    madrona::py::Tensor boxDataTensor() const;
    madrona::py::Tensor rampDataTensor() const;
    madrona::py::Tensor visibleAgentsMaskTensor() const;
    madrona::py::Tensor visibleBoxesMaskTensor() const;
    madrona::py::Tensor visibleRampsMaskTensor() const;
    madrona::py::Tensor globalPositionsTensor() const;
    madrona::py::Tensor lidarTensor() const;
    madrona::py::Tensor seedTensor() const;

    madrona::py::Tensor checkpointControlTensor() const;
    madrona::py::Tensor checkpointTensor() const;

    madrona::py::Tensor depthTensor() const;
    madrona::py::Tensor rgbTensor() const;

    void triggerReset(madrona::CountT world_idx, madrona::CountT level_idx);
    void setAction(madrona::CountT agent_idx, int32_t x, int32_t y, int32_t r, bool g, bool l);

    madrona::render::RenderManager &getRenderManager();

    madrona::py::Tensor policyAssignmentsTensor() const;
    madrona::py::Tensor episodeResultTensor() const;
    madrona::py::Tensor simControlTensor() const;       // empty, as in the reference

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
