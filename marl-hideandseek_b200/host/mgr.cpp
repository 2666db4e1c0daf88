// GPUHideSeek::Manager over the C ABI.  Error convention follows the reference: any failure
// is fatal (the reference's FATAL / REQ_CUDA abort the process, M/include/madrona/crash.hpp);
// the nanobind module of the reference is built -fno-exceptions, so scripts never see a
// Python exception from the simulator.  Set HIDESEEK_THROW=1 to get std::runtime_error
// instead (the pybind11 module below turns it into a Python RuntimeError), which is what
// the tests use to check the failure paths.
#include "mgr.hpp"

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include <hideseek/c_api.h>

namespace GPUHideSeek {

using madrona::py::Tensor;
using madrona::py::TensorElementType;

namespace {

[[noreturn]] void fatal(const std::string &msg)
{
    const char *t = getenv("HIDESEEK_THROW");
    if (t && t[0] == '1') throw std::runtime_error(msg);
    fprintf(stderr, "gpu_hideseek FATAL: %s\n", msg.c_str());
    fflush(stderr);
    abort();
}

inline void req(int rc, const char *what)
{
    if (rc != 0) fatal(std::string(what) + ": " + hs_last_error());
}

}  // namespace

struct Manager::Impl {
    Config cfg;
    hs_sim *sim = nullptr;
};

Manager::Manager(const Config &cfg) : impl_(new Impl{cfg, nullptr})
{
    if (cfg.execMode != madrona::ExecMode::CUDA)
        fatal("ExecMode::CPU requested: this build has only the CUDA (sm_100a) backend and no CPU fallback");
    if (cfg.extRenderAPI || cfg.extRenderDev)
        fatal("external render device requested: graphics APIs are unavailable on this box");
    hs_config c{};
    c.gpu_id = cfg.gpuID;
    c.num_worlds = (int32_t)cfg.numWorlds;
    c.world_base = (int32_t)cfg.worldBase;
    c.sim_flags = (uint32_t)cfg.simFlags;
    c.rand_seed = cfg.randSeed;
    c.min_hiders = (int32_t)cfg.minHiders;
    c.max_hiders = (int32_t)cfg.maxHiders;
    c.min_seekers = (int32_t)cfg.minSeekers;
    c.max_seekers = (int32_t)cfg.maxSeekers;
    c.num_pbt_policies = (int32_t)cfg.numPBTPolicies;
    if (hs_create(&c, &impl_->sim) != 0) {
        std::string msg = std::string("hs_create: ") + hs_last_error();
        delete impl_;
        impl_ = nullptr;
        fatal(msg);
    }
    // initRenderManager (src/mgr.cpp:54-88): the batch renderer's per-agent views are ray-cast
    // in CUDA (hs_enable_views), not rasterised through a graphics API
    if (cfg.enableBatchRenderer)
        req(hs_enable_views(impl_->sim, (int)cfg.batchRenderViewWidth, (int)cfg.batchRenderViewHeight,
                            cfg.renderDataDir), "hs_enable_views");
}

Manager::~Manager()
{
    if (!impl_) return;
    hs_destroy(impl_->sim);
    delete impl_;
}

// src/mgr.cpp:861-881
void Manager::init()
{
    req(hs_init(impl_->sim, nullptr), "hs_init");
    if (impl_->cfg.enableBatchRenderer) req(hs_render_views(impl_->sim, nullptr), "hs_render_views");   // :878-880
    req(hs_sync(impl_->sim, nullptr), "hs_sync");
}

// src/mgr.cpp:883-903: the legacy default stream orders the step after any prior
// default-stream writes of the caller (PyTorch's action writes) and the sync makes the
// outputs complete on return.
void Manager::step()
{
    req(hs_step(impl_->sim, nullptr), "hs_step");
    if (impl_->cfg.enableBatchRenderer) req(hs_render_views(impl_->sim, nullptr), "hs_render_views");   // :900-902
    req(hs_sync(impl_->sim, nullptr), "hs_sync");
}

void Manager::stepAsync(void *stream)
{
    req(hs_step(impl_->sim, stream), "hs_step");
    if (impl_->cfg.enableBatchRenderer) req(hs_render_views(impl_->sim, stream), "hs_render_views");
}
void Manager::sync(void *stream) { req(hs_sync(impl_->sim, stream), "hs_sync"); }

void Manager::stepHost(void *stream, const int32_t *actions, float *reward, int32_t *done)
{
    req(hs_step_host(impl_->sim, stream, actions, reward, done), "hs_step_host");
}

void Manager::saveCheckpoint(madrona::CountT world_idx)
{
    req(hs_save_checkpoint(impl_->sim, world_idx), "hs_save_checkpoint");
}

void Manager::loadCheckpoint(madrona::CountT world_idx)
{
    req(hs_load_checkpoint(impl_->sim, world_idx), "hs_load_checkpoint");
    if (impl_->cfg.enableBatchRenderer) {       // src/mgr.cpp:960-962
        req(hs_render_views(impl_->sim, nullptr), "hs_render_views");
        req(hs_sync(impl_->sim, nullptr), "hs_sync");
    }
}

void Manager::loadCheckpoints()
{
    req(hs_load_checkpoints(impl_->sim, nullptr), "hs_load_checkpoints");
    if (impl_->cfg.enableBatchRenderer) req(hs_render_views(impl_->sim, nullptr), "hs_render_views");   // :982-984
}

static Tensor exported(const hs_sim *sim, int gpu, int id)
{
    void *ptr = nullptr;
    int dtype = 0, ndim = 0;
    int64_t dims[4] = {0, 0, 0, 0};
    req(hs_export(const_cast<hs_sim *>(sim), id, &ptr, &dtype, &ndim, dims), "hs_export");
    TensorElementType t = dtype == HS_U8 ? TensorElementType::UInt8
                        : dtype == HS_I32 ? TensorElementType::Int32 : TensorElementType::Float32;
    return Tensor(ptr, t, dims, ndim, gpu);
}

#define HS_GETTER(name, id) \
    Tensor Manager::name() const { return exported(impl_->sim, impl_->cfg.gpuID, id); }

HS_GETTER(resetTensor, HS_EXPORT_RESET)
HS_GETTER(doneTensor, HS_EXPORT_DONE)
HS_GETTER(prepCounterTensor, HS_EXPORT_PREP_COUNTER)
HS_GETTER(actionTensor, HS_EXPORT_ACTION)
HS_GETTER(rewardTensor, HS_EXPORT_REWARD)
HS_GETTER(selfDataTensor, HS_EXPORT_SELF_OBS)
HS_GETTER(selfTypeTensor, HS_EXPORT_SELF_TYPE)
HS_GETTER(selfMaskTensor, HS_EXPORT_SELF_MASK)
HS_GETTER(agentDataTensor, HS_EXPORT_AGENT_OBS)
HS_GETTER(boxDataTensor, HS_EXPORT_BOX_OBS)
HS_GETTER(rampDataTensor, HS_EXPORT_RAMP_OBS)
HS_GETTER(visibleAgentsMaskTensor, HS_EXPORT_AGENT_VIS)
HS_GETTER(visibleBoxesMaskTensor, HS_EXPORT_BOX_VIS)
HS_GETTER(visibleRampsMaskTensor, HS_EXPORT_RAMP_VIS)
HS_GETTER(globalPositionsTensor, HS_EXPORT_GLOBAL_POS)
HS_GETTER(lidarTensor, HS_EXPORT_LIDAR)
HS_GETTER(seedTensor, HS_EXPORT_SEED)
HS_GETTER(checkpointControlTensor, HS_EXPORT_CKPT_CTRL)
HS_GETTER(checkpointTensor, HS_EXPORT_CKPT)
HS_GETTER(policyAssignmentsTensor, HS_EXPORT_AGENT_POLICY)
HS_GETTER(episodeResultTensor, HS_EXPORT_EPISODE_RESULT)
#undef HS_GETTER

// src/mgr.cpp:1241-1263: renderer outputs
Tensor Manager::depthTensor() const
{
    if (!impl_->cfg.enableBatchRenderer) fatal("depthTensor: enableBatchRenderer is false");
    return exported(impl_->sim, impl_->cfg.gpuID, HS_EXPORT_VIEW_DEPTH);
}
Tensor Manager::rgbTensor() const
{
    if (!impl_->cfg.enableBatchRenderer) fatal("rgbTensor: enableBatchRenderer is false");
    return exported(impl_->sim, impl_->cfg.gpuID, HS_EXPORT_VIEW_RGB);
}
madrona::render::RenderManager &Manager::getRenderManager()
{
    fatal("getRenderManager: batch renderer unavailable on this box");
}

// src/mgr.cpp:1265-1310
void Manager::triggerReset(madrona::CountT world_idx, madrona::CountT level_idx)
{
    req(hs_trigger_reset(impl_->sim, world_idx, level_idx), "hs_trigger_reset");
}

void Manager::setAction(madrona::CountT agent_idx, int32_t x, int32_t y, int32_t r, bool g, bool l)
{
    req(hs_set_action(impl_->sim, agent_idx, x, y, r, g ? 1 : 0, l ? 1 : 0), "hs_set_action");
}

// src/mgr.cpp:987-1060 -> CUDAImpl::gpuStream* (src/mgr.cpp:365-437)
void Manager::gpuJAXInit(cudaStream_t strm, void **buffers)
{
    req(hs_jax_init(impl_->sim, strm, buffers), "hs_jax_init");
}

void Manager::gpuJAXStep(cudaStream_t strm, void **buffers)
{
    req(hs_jax_step(impl_->sim, strm, buffers), "hs_jax_step");
}

void Manager::gpuJAXSaveCheckpoints(cudaStream_t strm, void **buffers)
{
    req(hs_jax_save_checkpoints(impl_->sim, strm, buffers), "hs_jax_save_checkpoints");
}

void Manager::gpuJAXLoadCheckpoints(cudaStream_t strm, void **buffers)
{
    req(hs_jax_load_checkpoints(impl_->sim, strm, buffers), "hs_jax_load_checkpoints");
}

// src/mgr.cpp:1333-1336: no simulator control tensor
Tensor Manager::simControlTensor() const
{
    const int64_t dims[1] = {0};
    return Tensor(nullptr, TensorElementType::Int32, dims, 1, -1);
}

// src/mgr.cpp:1338-1375: same names, same order
madrona::py::TrainInterface Manager::trainInterface() const
{
    using namespace madrona::py;
    return TrainInterface(
        TrainStepInputInterface{
            {{"actions", actionTensor()}},
            resetTensor(),
            simControlTensor(),
            {{"policy_assignments", policyAssignmentsTensor()}},
        },
        TrainStepOutputInterface{
            {
                {"prep_counter", prepCounterTensor()},
                {"self_data", selfDataTensor()},
                {"self_type", selfTypeTensor()},
                {"self_mask", selfMaskTensor()},
                {"self_lidar", lidarTensor()},
                {"agent_data", agentDataTensor()},
                {"box_data", boxDataTensor()},
                {"ramp_data", rampDataTensor()},
                {"vis_agents_mask", visibleAgentsMaskTensor()},
                {"vis_boxes_mask", visibleBoxesMaskTensor()},
                {"vis_ramps_mask", visibleRampsMaskTensor()},
            },
            rewardTensor(),
            doneTensor(),
            {},
            {{"episode_results", episodeResultTensor()}},
        },
        TrainCheckpointingInterface{checkpointTensor()});
}

void *Manager::handle() const { return impl_->sim; }
uint32_t Manager::numWorlds() const { return impl_->cfg.numWorlds; }
uint32_t Manager::agentsPerWorld() const { return impl_->cfg.maxHiders + impl_->cfg.maxSeekers; }
int Manager::gpuID() const { return impl_->cfg.gpuID; }

}  // namespace GPUHideSeek
