// Python module `gpu_hideseek._native` — the drop-in for the reference's nanobind module
// (src/bindings.cpp:20-121 + the `madrona` submodule of M/src/python/bindings.cpp:320-378):
// same class name, constructor keywords, method names and `madrona.Tensor.to_torch()`.
// Written against pybind11 (nanobind is not installed in this image); tensors are exported
// zero-copy through the DLPack protocol exactly like the reference's nb::ndarray path
// (M/src/python/bindings.cpp:52-68): no owner, the pointer stays valid for the lifetime of
// the simulator object, which the returned Python tensor keeps alive.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstring>
#include <stdexcept>
#include <string>

#include "mgr.hpp"

namespace py = pybind11;
using GPUHideSeek::Manager;
using GPUHideSeek::SimFlags;
using madrona::py::Tensor;
using madrona::py::TensorElementType;

namespace {

// ---- DLPack ABI (dlpack.h v0.8, public C structs) -----------------------------------------
struct DLDevice { int32_t device_type; int32_t device_id; };
struct DLDataType { uint8_t code; uint8_t bits; uint16_t lanes; };
struct DLTensor {
    void *data; DLDevice device; int32_t ndim; DLDataType dtype;
    int64_t *shape; int64_t *strides; uint64_t byte_offset;
};
struct DLManagedTensor { DLTensor dl_tensor; void *manager_ctx; void (*deleter)(DLManagedTensor *); };
constexpr int32_t kDLCPU = 1, kDLCUDA = 2;
constexpr uint8_t kDLInt = 0, kDLUInt = 1, kDLFloat = 2;

struct ManagedCtx { int64_t shape[Tensor::maxDimensions]; PyObject *owner; };

void dlDeleter(DLManagedTensor *t)
{
    ManagedCtx *ctx = (ManagedCtx *)t->manager_ctx;
    if (ctx->owner) {
        py::gil_scoped_acquire gil;
        Py_DECREF(ctx->owner);
    }
    delete ctx;
    delete t;
}

void capsuleDestructor(PyObject *cap)
{
    // an unconsumed capsule still owns the managed tensor
    if (PyCapsule_IsValid(cap, "dltensor")) {
        DLManagedTensor *t = (DLManagedTensor *)PyCapsule_GetPointer(cap, "dltensor");
        if (t && t->deleter) t->deleter(t);
    }
}

DLDataType dlType(TensorElementType t)
{
    switch (t) {
    case TensorElementType::UInt8: return {kDLUInt, 8, 1};
    case TensorElementType::Int8: return {kDLInt, 8, 1};
    case TensorElementType::Int16: return {kDLInt, 16, 1};
    case TensorElementType::Int32: return {kDLInt, 32, 1};
    case TensorElementType::Int64: return {kDLInt, 64, 1};
    case TensorElementType::Float16: return {kDLFloat, 16, 1};
    case TensorElementType::Float32: return {kDLFloat, 32, 1};
    }
    throw std::runtime_error("unknown tensor element type");
}

// madrona.Tensor as seen from Python: the descriptor + a reference that keeps the
// simulator (owner of the device memory) alive.
struct PyTensor {
    Tensor t;
    py::object owner;

    py::capsule dlpack() const
    {
        auto *m = new DLManagedTensor();
        auto *ctx = new ManagedCtx();
        for (int64_t i = 0; i < t.numDims(); i++) ctx->shape[i] = t.dims()[i];
        ctx->owner = owner.ptr();
        Py_XINCREF(ctx->owner);
        m->dl_tensor.data = t.devicePtr();
        m->dl_tensor.device = t.isOnGPU() ? DLDevice{kDLCUDA, t.gpuID()} : DLDevice{kDLCPU, 0};
        m->dl_tensor.ndim = (int32_t)t.numDims();
        m->dl_tensor.dtype = dlType(t.type());
        m->dl_tensor.shape = ctx->shape;
        m->dl_tensor.strides = nullptr;          // compact row-major
        m->dl_tensor.byte_offset = 0;
        m->manager_ctx = ctx;
        m->deleter = dlDeleter;
        return py::reinterpret_steal<py::capsule>(PyCapsule_New(m, "dltensor", capsuleDestructor));
    }

    py::object toTorch() const
    {
        py::object from_dlpack = py::module_::import("torch.utils.dlpack").attr("from_dlpack");
        return from_dlpack(dlpack());
    }

    py::object toJax() const
    {
        py::object jdl = py::module_::import("jax.dlpack");     // raises ImportError without jax
        return jdl.attr("from_dlpack")(py::cast(*this));
    }

    py::tuple shape() const
    {
        py::tuple s(t.numDims());
        for (int64_t i = 0; i < t.numDims(); i++) s[i] = t.dims()[i];
        return s;
    }
};

PyTensor wrap(py::object self, Tensor t) { return PyTensor{t, std::move(self)}; }

}  // namespace

PYBIND11_MODULE(_native, m)
{
    m.doc() = "gpu_hideseek (B200): hide-and-seek batch simulator, CUDA sm_100a backend";

    // ---- `madrona` submodule (M/src/python/bindings.cpp:320-378) ---------------------------
    py::module_ mad = m.def_submodule("madrona");
    py::enum_<madrona::ExecMode>(mad, "ExecMode")
        .value("CPU", madrona::ExecMode::CPU)
        .value("CUDA", madrona::ExecMode::CUDA);
    py::class_<PyTensor>(mad, "Tensor")
        .def("to_torch", &PyTensor::toTorch)
        .def("to_jax", &PyTensor::toJax)
        .def("__dlpack__", [](const PyTensor &t, py::kwargs) { return t.dlpack(); })
        .def("__dlpack_device__", [](const PyTensor &t) {
            return py::make_tuple(t.t.isOnGPU() ? kDLCUDA : kDLCPU, t.t.isOnGPU() ? t.t.gpuID() : 0);
        })
        .def_property_readonly("shape", &PyTensor::shape)
        .def_property_readonly("dtype", [](const PyTensor &t) -> std::string {
            // numpy / jax.numpy dtype name (JAXModule::getDType, M/src/python/bindings.cpp:107-135)
            using ET = madrona::py::TensorElementType;
            switch (t.t.type()) {
            case ET::UInt8: return "uint8";
            case ET::Int8: return "int8";
            case ET::Int16: return "int16";
            case ET::Int32: return "int32";
            case ET::Int64: return "int64";
            case ET::Float16: return "float16";
            case ET::Float32: return "float32";
            }
            return "unknown";
        })
        .def("device_ptr", [](const PyTensor &t) { return (uintptr_t)t.t.devicePtr(); })
        .def("gpu_id", [](const PyTensor &t) { return t.t.gpuID(); });

    // arithmetic enum like nb::is_arithmetic()
    py::enum_<SimFlags>(m, "SimFlags", py::arithmetic())
        .value("Default", SimFlags::Default)
        .value("UseFixedWorld", SimFlags::UseFixedWorld)
        .value("IgnoreEpisodeLength", SimFlags::IgnoreEpisodeLength)
        .value("RandomFlipTeams", SimFlags::RandomFlipTeams)
        .value("ZeroAgentVelocity", SimFlags::ZeroAgentVelocity);

#define HS_TENSOR(pyname, method) \
    .def(pyname, [](py::object self) { return wrap(self, self.cast<Manager &>().method()); })

    py::class_<Manager>(m, "HideAndSeekSimulator")
        .def(py::init([](madrona::ExecMode exec_mode, int64_t gpu_id, int64_t num_worlds,
                         int64_t sim_flags, uint32_t rand_seed, uint32_t min_hiders,
                         uint32_t max_hiders, uint32_t min_seekers, uint32_t max_seekers,
                         uint32_t num_pbt_policies, bool enable_batch_renderer,
                         int64_t batch_render_width, int64_t batch_render_height,
                         int64_t world_base, const std::string &render_data_dir) {
                 Manager::Config cfg{};
                 cfg.execMode = exec_mode;
                 cfg.gpuID = (int)gpu_id;
                 cfg.numWorlds = (uint32_t)num_worlds;
                 cfg.simFlags = (SimFlags)sim_flags;
                 cfg.randSeed = rand_seed;
                 cfg.minHiders = min_hiders;
                 cfg.maxHiders = max_hiders;
                 cfg.minSeekers = min_seekers;
                 cfg.maxSeekers = max_seekers;
                 cfg.numPBTPolicies = num_pbt_policies;
                 cfg.enableBatchRenderer = enable_batch_renderer;
                 cfg.batchRenderViewWidth = (uint32_t)batch_render_width;
                 cfg.batchRenderViewHeight = (uint32_t)batch_render_height;
                 cfg.worldBase = (uint32_t)world_base;
                 cfg.renderDataDir = render_data_dir.empty() ? nullptr : render_data_dir.c_str();
                 return new Manager(cfg);
             }),
             py::arg("exec_mode"), py::arg("gpu_id"), py::arg("num_worlds"), py::arg("sim_flags"),
             py::arg("rand_seed"), py::arg("min_hiders"), py::arg("max_hiders"),
             py::arg("min_seekers"), py::arg("max_seekers"), py::arg("num_pbt_policies"),
             py::arg("enable_batch_renderer") = false, py::arg("batch_render_width") = 64,
             py::arg("batch_render_height") = 64, py::arg("world_base") = 0,
             py::arg("render_data_dir") = std::string())
        .def("init", &Manager::init)
        .def("step", &Manager::step, py::call_guard<py::gil_scoped_release>())
        HS_TENSOR("reset_tensor", resetTensor)
        HS_TENSOR("done_tensor", doneTensor)
        HS_TENSOR("prep_counter_tensor", prepCounterTensor)
        HS_TENSOR("action_tensor", actionTensor)
        HS_TENSOR("reward_tensor", rewardTensor)
        HS_TENSOR("self_data_tensor", selfDataTensor)
        HS_TENSOR("self_type_tensor", selfTypeTensor)
        HS_TENSOR("self_mask_tensor", selfMaskTensor)
        HS_TENSOR("agent_data_tensor", agentDataTensor)
        HS_TENSOR("box_data_tensor", boxDataTensor)
        HS_TENSOR("ramp_data_tensor", rampDataTensor)
        HS_TENSOR("visible_agents_mask_tensor", visibleAgentsMaskTensor)
        HS_TENSOR("visible_boxes_mask_tensor", visibleBoxesMaskTensor)
        HS_TENSOR("visible_ramps_mask_tensor", visibleRampsMaskTensor)
        HS_TENSOR("global_positions_tensor", globalPositionsTensor)
        HS_TENSOR("depth_tensor", depthTensor)
        HS_TENSOR("rgb_tensor", rgbTensor)
        HS_TENSOR("lidar_tensor", lidarTensor)
        HS_TENSOR("seed_tensor", seedTensor)
        HS_TENSOR("ckpt_ctrl_tensor", checkpointControlTensor)
        HS_TENSOR("ckpt_tensor", checkpointTensor)
        HS_TENSOR("policy_assignments_tensor", policyAssignmentsTensor)
        HS_TENSOR("episode_result_tensor", episodeResultTensor)
        .def("trigger_reset", &Manager::triggerReset, py::arg("world"), py::arg("level") = 1)
        .def("set_action", &Manager::setAction, py::arg("agent_row"), py::arg("x"), py::arg("y"),
             py::arg("r"), py::arg("g") = false, py::arg("l") = false)
        .def("save_checkpoint", &Manager::saveCheckpoint)
        .def("load_checkpoint", &Manager::loadCheckpoint)
        .def("load_checkpoints", &Manager::loadCheckpoints)
        // `jax(xla_gpu)` (src/bindings.cpp:118, JAXInterface::buildEntry) is provided by the Python
        // layer: gpu_hideseek.HideAndSeekSimulator.jax -> gpu_hideseek/jax_register.py, which
        // registers the hs_xla_gpu_* entry points of the C ABI as XLA custom calls.
        // TrainInterface (src/mgr.cpp:1338-1375) as nested dicts of named tensors, in the
        // reference's order (Python dicts keep insertion order)
        .def("train_interface", [](py::object self) {
            Manager &mgr = self.cast<Manager &>();
            madrona::py::TrainInterface ti = mgr.trainInterface();
            auto named = [&](const std::vector<madrona::py::NamedTensor> &v) {
                py::dict d;
                for (const auto &nt : v) d[nt.name] = wrap(self, nt.tensor);
                return d;
            };
            py::dict in, out, ck, r;
            in["actions"] = named(ti.stepInputs().actions);
            in["resets"] = wrap(self, ti.stepInputs().resets);
            in["pbt"] = named(ti.stepInputs().pbt);
            out["observations"] = named(ti.stepOutputs().observations);
            out["rewards"] = wrap(self, ti.stepOutputs().rewards);
            out["dones"] = wrap(self, ti.stepOutputs().dones);
            out["pbt"] = named(ti.stepOutputs().pbt);
            ck["checkpoint_data"] = wrap(self, ti.checkpointing().checkpointData);
            r["step_inputs"] = in; r["step_outputs"] = out; r["checkpointing"] = ck;
            return r;
        })
        // the bodies of the XLA GPU custom calls (src/mgr.hpp:44-50); `buffers` = device pointers
#define HS_JAX(pyname, method)                                                                    \
        .def(pyname, [](Manager &mgr, uintptr_t stream, std::vector<uintptr_t> buffers) {        \
                 std::vector<void *> b(buffers.size());                                          \
                 for (size_t i = 0; i < buffers.size(); i++) b[i] = (void *)buffers[i];          \
                 mgr.method((cudaStream_t)stream, b.data());                                     \
             }, py::arg("stream"), py::arg("buffers"))
        HS_JAX("gpu_jax_init", gpuJAXInit)
        HS_JAX("gpu_jax_step", gpuJAXStep)
        HS_JAX("gpu_jax_save_checkpoints", gpuJAXSaveCheckpoints)
        HS_JAX("gpu_jax_load_checkpoints", gpuJAXLoadCheckpoints)
#undef HS_JAX
        // ---- extensions ----------------------------------------------------------------
        .def("step_async", [](Manager &mgr, uintptr_t stream) { mgr.stepAsync((void *)stream); },
             py::arg("stream") = 0)
        .def("sync", [](Manager &mgr, uintptr_t stream) { mgr.sync((void *)stream); },
             py::arg("stream") = 0, py::call_guard<py::gil_scoped_release>())
        .def("step_host", [](Manager &mgr, uintptr_t actions, uintptr_t reward, uintptr_t done,
                             uintptr_t stream) {
                 mgr.stepHost((void *)stream, (const int32_t *)actions, (float *)reward, (int32_t *)done);
             }, py::arg("actions_ptr"), py::arg("reward_ptr"), py::arg("done_ptr"), py::arg("stream") = 0)
        .def("handle", [](Manager &mgr) { return (uintptr_t)mgr.handle(); })
        .def_property_readonly("num_worlds", &Manager::numWorlds)
        .def_property_readonly("agents_per_world", &Manager::agentsPerWorld)
        .def_property_readonly("gpu_id", &Manager::gpuID);
#undef HS_TENSOR
}
