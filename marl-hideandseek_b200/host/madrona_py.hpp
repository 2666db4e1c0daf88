// Minimal restatement of the two Madrona host types the Manager API exposes:
// `madrona::ExecMode` (M/include/madrona/exec_mode.hpp) and the non-owning
// `madrona::py::Tensor` descriptor (M/include/madrona/py/utils.hpp:60-141).
// Same names, accessors and element-type enumerators, so code written against
// the reference's Manager compiles against this one.
#pragma once

#include <array>
#include <cstdint>
#include <initializer_list>
#include <vector>

namespace madrona {

using CountT = int32_t;

enum class ExecMode : uint32_t { CPU, CUDA };

namespace py {

enum class TensorElementType { UInt8, Int8, Int16, Int32, Int64, Float16, Float32 };

class Tensor final {
public:
    static inline constexpr int64_t maxDimensions = 16;

    Tensor(void *dev_ptr, TensorElementType type, const int64_t *dims, int64_t num_dims, int gpu_id)
        : dev_ptr_(dev_ptr), type_(type), gpu_id_(gpu_id), num_dimensions_(num_dims), dimensions_{}
    {
        for (int64_t i = 0; i < num_dims && i < maxDimensions; i++) dimensions_[i] = dims[i];
    }

    inline void *devicePtr() const { return dev_ptr_; }
    inline TensorElementType type() const { return type_; }
    inline bool isOnGPU() const { return gpu_id_ != -1; }
    inline int gpuID() const { return gpu_id_; }
    inline int64_t numDims() const { return num_dimensions_; }
    inline const int64_t *dims() const { return dimensions_.data(); }
    inline int64_t numBytesPerItem() const
    {
        switch (type_) {
        case TensorElementType::UInt8: case TensorElementType::Int8: return 1;
        case TensorElementType::Int16: case TensorElementType::Float16: return 2;
        case TensorElementType::Int32: case TensorElementType::Float32: return 4;
        case TensorElementType::Int64: return 8;
        }
        return 0;
    }
    inline int64_t numItems() const
    {
        int64_t n = 1;
        for (int64_t i = 0; i < num_dimensions_; i++) n *= dimensions_[i];
        return n;
    }

private:
    void *dev_ptr_;
    TensorElementType type_;
    int gpu_id_;
    int64_t num_dimensions_;
    std::array<int64_t, maxDimensions> dimensions_;
};

// The training-loop view of a simulator (M/include/madrona/py/utils.hpp:143-195): which tensors
// are step inputs, which are outputs, and the checkpoint blob.  Value types here (std::vector
// where the reference holds Span + pimpl); same member names.
struct NamedTensor {
    const char *name;
    Tensor tensor;
};

struct TrainStepInputInterface {
    std::vector<NamedTensor> actions;
    Tensor resets;
    Tensor simCtrl;
    std::vector<NamedTensor> pbt;
};

struct TrainStepOutputInterface {
    std::vector<NamedTensor> observations;
    Tensor rewards;
    Tensor dones;
    std::vector<NamedTensor> stats;
    std::vector<NamedTensor> pbt;
};

struct TrainCheckpointingInterface {
    Tensor checkpointData;
};

class TrainInterface {
public:
    TrainInterface(TrainStepInputInterface in, TrainStepOutputInterface out, TrainCheckpointingInterface ckpt)
        : in_(std::move(in)), out_(std::move(out)), ckpt_(ckpt) {}
    const TrainStepInputInterface &stepInputs() const { return in_; }
    const TrainStepOutputInterface &stepOutputs() const { return out_; }
    const TrainCheckpointingInterface &checkpointing() const { return ckpt_; }

private:
    TrainStepInputInterface in_;
    TrainStepOutputInterface out_;
    TrainCheckpointingInterface ckpt_;
};

}  // namespace py
}  // namespace madrona
