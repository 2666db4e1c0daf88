// Headless C++ driver with the reference's CLI (src/headless.cpp:12-120):
//   headless CUDA <num_worlds> <num_steps> [--rand-actions]
// Builds a Manager (3 hiders + 2 seekers, seed 5 like the reference driver), steps it and
// prints the FPS line.  ExecMode "CPU" is rejected: this build has no CPU backend.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include <cuda_runtime_api.h>

#include "mgr.hpp"

using namespace GPUHideSeek;

int main(int argc, char *argv[])
{
    if (argc < 4) {
        fprintf(stderr, "%s TYPE NUM_WORLDS NUM_STEPS [--rand-actions]\n", argv[0]);
        return 1;
    }
    if (strcmp(argv[1], "CUDA") != 0) {
        fprintf(stderr, "only TYPE=CUDA exists in this build (no CPU backend)\n");
        return 1;
    }
    uint32_t num_worlds = (uint32_t)atoi(argv[2]);
    uint32_t num_steps = (uint32_t)atoi(argv[3]);
    bool rand_actions = argc >= 5 && strcmp(argv[4], "--rand-actions") == 0;

    Manager::Config cfg{};
    cfg.execMode = madrona::ExecMode::CUDA;
    cfg.gpuID = 0;
    cfg.numWorlds = num_worlds;
    cfg.simFlags = SimFlags::Default;
    cfg.randSeed = 5;
    cfg.minHiders = 3; cfg.maxHiders = 3;
    cfg.minSeekers = 2; cfg.maxSeekers = 2;
    cfg.numPBTPolicies = 0;
    cfg.enableBatchRenderer = false;
    Manager mgr(cfg);
    mgr.init();

    const uint32_t A = mgr.agentsPerWorld();
    const size_t rows = size_t(num_worlds) * A;
    std::mt19937 rng(5);
    std::uniform_int_distribution<int32_t> act(0, 10), bit(0, 1);
    std::vector<int32_t> host(rows * 5);
    madrona::py::Tensor action = mgr.actionTensor();

    auto start = std::chrono::steady_clock::now();
    for (uint32_t i = 0; i < num_steps; i++) {
        if (rand_actions) {
            for (size_t r = 0; r < rows; r++) {
                host[r * 5 + 0] = act(rng); host[r * 5 + 1] = act(rng); host[r * 5 + 2] = act(rng);
                host[r * 5 + 3] = bit(rng); host[r * 5 + 4] = bit(rng);
            }
            cudaMemcpy(action.devicePtr(), host.data(), host.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
        }
        mgr.step();
    }
    auto end = std::chrono::steady_clock::now();
    double secs = std::chrono::duration<double>(end - start).count();
    printf("FPS %f\n", double(num_steps) * double(num_worlds) / secs);

    std::vector<float> reward(rows);
    cudaMemcpy(reward.data(), mgr.rewardTensor().devicePtr(), rows * 4, cudaMemcpyDeviceToHost);
    double sum = 0;
    for (float v : reward) sum += v;
    printf("reward checksum %f\n", sum);
    return 0;
}
