// Translation unit of the agent-view ray caster (hs_views.cuh).  Separate from hs_sim.cu because
// the simulation kernels are compiled -fmad=false to reproduce the reference CPU backend's fp32
// results, while the ray caster has no such constraint and runs with FMA contraction.
#include "hs_views.cuh"

namespace hs {

cudaError_t launchAgentViews(const DeviceState &S, const ViewParams &P, unsigned views, cudaStream_t stream)
{
    k_agent_views<<<views, kViewThreads, 0, stream>>>(S, P);
    return cudaGetLastError();
}

}  // namespace hs
