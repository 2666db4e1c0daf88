// Windowless record / replay driver: the replay loop of the reference's viewer
// (src/viewer.cpp:12-27 readReplayLog, :186-216 replayStep, :224-245 printObs) without the window.
//   replay [NUM_WORLDS] [--record path] [--replay path] [--steps N] [--print-obs] [--dump-views dir]
// A log is what scripts/jax_infer.py --record-log writes and viewer --replay reads: for every
// step, the Checkpoint records (src/sim.hpp:283-313, 1392 bytes) of all worlds, back to back.
//   --record : run N steps with seeded random actions, saving every world's checkpoint after
//              each step (CheckpointControl trigger + SaveCheckpoints graph) and appending it.
//   --replay : for each logged step set every world's trigger, copy the step's checkpoints into
//              the Checkpoint tensor and call Manager::loadCheckpoints(), like replayStep().
//   --print-obs   : self data / prep counter / reward of world 0 after every step (printObs).
//   --dump-views  : with the CUDA agent-view ray caster on, write world 0's agent views of every
//                   step as dir/step%04d_agent%d.pgm (depth, 8-bit, 0 = miss) and .ppm (colour) —
//                   the headless stand-in for the viewer's window.
// Manager configuration as the viewer's: 3 hiders + 3 seekers, seed 5, ZeroAgentVelocity,
// IgnoreEpisodeLength unless replaying (src/viewer.cpp:123-160).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include <cuda_runtime_api.h>
#include <hideseek/c_api.h>

#include "mgr.hpp"

using namespace GPUHideSeek;

namespace {

constexpr size_t kCheckpointBytes = 1392;

void usage(const char *argv0)
{
    fprintf(stderr, "%s [NUM_WORLDS] [--record path] [--replay path] [--steps N] [--print-obs] [--dump-views dir]\n", argv0);
    exit(EXIT_FAILURE);
}

template <typename T>
std::vector<T> download(const madrona::py::Tensor &t, size_t count)
{
    std::vector<T> h(count);
    cudaMemcpy(h.data(), t.devicePtr(), count * sizeof(T), cudaMemcpyDeviceToHost);
    return h;
}

void writeImage(const std::string &path, const char *magic, int w, int h, const std::vector<uint8_t> &px)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(EXIT_FAILURE); }
    fprintf(f, "%s\n%d %d\n255\n", magic, w, h);
    fwrite(px.data(), 1, px.size(), f);
    fclose(f);
}

}  // namespace

int main(int argc, char *argv[])
{
    uint32_t num_worlds = 1, num_steps = 0;
    bool num_worlds_set = false, print_obs = false;
    const char *record_path = nullptr, *replay_path = nullptr, *dump_dir = nullptr;
    for (int i = 1; i < argc; i++) {
        char *arg = argv[i];
        auto value = [&]() -> char * { if (++i == argc) usage(argv[0]); return argv[i]; };
        if (arg[0] == '-' && arg[1] == '-') {
            arg += 2;
            if (!strcmp("record", arg)) record_path = value();
            else if (!strcmp("replay", arg)) replay_path = value();
            else if (!strcmp("steps", arg)) num_steps = (uint32_t)atoi(value());
            else if (!strcmp("dump-views", arg)) dump_dir = value();
            else if (!strcmp("print-obs", arg)) print_obs = true;
            else usage(argv[0]);
        } else {
            if (num_worlds_set) usage(argv[0]);
            num_worlds_set = true;
            num_worlds = (uint32_t)atoi(arg);
        }
    }
    if ((record_path != nullptr) == (replay_path != nullptr) || num_worlds == 0) usage(argv[0]);

    // readReplayLog
    std::vector<uint8_t> log;
    if (replay_path) {
        FILE *f = fopen(replay_path, "rb");
        if (!f) { fprintf(stderr, "cannot read %s\n", replay_path); return EXIT_FAILURE; }
        fseek(f, 0, SEEK_END);
        long size = ftell(f);
        fseek(f, 0, SEEK_SET);
        log.resize(size_t(size) / kCheckpointBytes * kCheckpointBytes);
        if (fread(log.data(), 1, log.size(), f) != log.size()) { fprintf(stderr, "short read\n"); return EXIT_FAILURE; }
        fclose(f);
        num_steps = uint32_t(log.size() / kCheckpointBytes / num_worlds);
    }

    const int view = 64;
    Manager::Config cfg{};
    cfg.execMode = madrona::ExecMode::CUDA;
    cfg.gpuID = 0;
    cfg.numWorlds = num_worlds;
    cfg.simFlags = SimFlags::ZeroAgentVelocity;
    if (!replay_path) cfg.simFlags = cfg.simFlags | SimFlags::IgnoreEpisodeLength;
    cfg.randSeed = 5;
    cfg.minHiders = 3; cfg.maxHiders = 3;
    cfg.minSeekers = 3; cfg.maxSeekers = 3;
    cfg.numPBTPolicies = 1;
    cfg.enableBatchRenderer = dump_dir != nullptr;
    cfg.batchRenderViewWidth = view; cfg.batchRenderViewHeight = view;
    Manager mgr(cfg);
    mgr.init();

    const uint32_t A = mgr.agentsPerWorld();
    const size_t rows = size_t(num_worlds) * A;
    std::vector<int32_t> trigger_all(num_worlds, 1);
    std::vector<int32_t> actions(rows * 5);
    std::vector<uint8_t> ckpts(size_t(num_worlds) * kCheckpointBytes);
    std::mt19937 rng(5);
    std::uniform_int_distribution<int32_t> act(0, 10), bit(0, 9);
    FILE *rec = nullptr;
    if (record_path && !(rec = fopen(record_path, "wb"))) { fprintf(stderr, "cannot write %s\n", record_path); return EXIT_FAILURE; }

    for (uint32_t step = 0; step < num_steps; step++) {
        if (replay_path) {
            cudaMemcpy(mgr.checkpointControlTensor().devicePtr(), trigger_all.data(), num_worlds * sizeof(int32_t),
                       cudaMemcpyHostToDevice);
            cudaMemcpy(mgr.checkpointTensor().devicePtr(), &log[size_t(step) * num_worlds * kCheckpointBytes],
                       ckpts.size(), cudaMemcpyHostToDevice);
            mgr.loadCheckpoints();
            cudaDeviceSynchronize();
        } else {
            for (size_t r = 0; r < rows; r++) {
                actions[r * 5 + 0] = act(rng); actions[r * 5 + 1] = act(rng); actions[r * 5 + 2] = act(rng);
                actions[r * 5 + 3] = bit(rng) == 0; actions[r * 5 + 4] = bit(rng) == 0;
            }
            cudaMemcpy(mgr.actionTensor().devicePtr(), actions.data(), actions.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
            mgr.step();
            // every world's checkpoint of this step: triggers + the SaveCheckpoints graph
            cudaMemcpy(mgr.checkpointControlTensor().devicePtr(), trigger_all.data(), num_worlds * sizeof(int32_t),
                       cudaMemcpyHostToDevice);
            if (hs_save_checkpoints((hs_sim *)mgr.handle(), nullptr) != 0 || hs_sync((hs_sim *)mgr.handle(), nullptr) != 0) {
                fprintf(stderr, "save checkpoints: %s\n", hs_last_error());
                return EXIT_FAILURE;
            }
            cudaMemcpy(ckpts.data(), mgr.checkpointTensor().devicePtr(), ckpts.size(), cudaMemcpyDeviceToHost);
            fwrite(ckpts.data(), 1, ckpts.size(), rec);
        }

        // one line per step that must agree between the recording run and its replay
        std::vector<float> self = download<float>(mgr.selfDataTensor(), rows * 13);
        std::vector<int32_t> prep = download<int32_t>(mgr.prepCounterTensor(), rows);
        double sum = 0;
        for (size_t r = 0; r < rows; r++) for (int c = 0; c < 6; c++) sum += std::fabs((double)self[r * 13 + c]);
        printf("step %u prep %d pose_checksum %.6f\n", step, prep[0], sum);
        if (print_obs) {
            std::vector<float> reward = download<float>(mgr.rewardTensor(), rows);
            printf("Self Data\n");
            for (uint32_t a = 0; a < A; a++) {
                for (int c = 0; c < 13; c++) printf(" %9.4f", self[a * 13 + c]);
                printf("\n");
            }
            printf("Prep Counter\n");
            for (uint32_t a = 0; a < A; a++) printf(" %d", prep[a]);
            printf("\nReward\n");
            for (uint32_t a = 0; a < A; a++) printf(" %g", reward[a]);
            printf("\n\n");
        }
        if (dump_dir) {
            const size_t px = size_t(view) * view;
            std::vector<float> depth = download<float>(mgr.depthTensor(), A * px);
            std::vector<uint8_t> rgba = download<uint8_t>(mgr.rgbTensor(), A * px * 4);
            for (uint32_t a = 0; a < A; a++) {
                std::vector<uint8_t> g(px), c(px * 3);
                for (size_t p = 0; p < px; p++) {
                    float d = depth[a * px + p];
                    g[p] = d <= 0.f ? 0 : (uint8_t)(255.f - fminf(d, 50.f) * (254.f / 50.f));
                    for (int k = 0; k < 3; k++) c[p * 3 + k] = rgba[(a * px + p) * 4 + k];
                }
                char name[64];
                snprintf(name, sizeof(name), "/step%04u_agent%u", step, a);
                writeImage(std::string(dump_dir) + name + ".pgm", "P5", view, view, g);
                writeImage(std::string(dump_dir) + name + ".ppm", "P6", view, view, c);
            }
        }
    }
    if (rec) fclose(rec);
    return 0;
}
