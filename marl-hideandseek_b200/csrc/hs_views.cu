// Translation unit of the agent-view ray caster (hs_views.cuh).  Separate from hs_sim.cu because
// the simulation kernels are compiled -fmad=false to reproduce the reference CPU backend's fp32
// results, while the ray caster has no such constraint and runs with FMA contraction.
#include "hs_views.cuh"

namespace hs {

cudaError_t launchAgentViews(const DeviceState &S, const ViewParams &P, unsigned views, cudaStream_t stream)
{
    // persistent CTAs: kViewCtasPerSm per SM, each with the mesh table in dynamic shared memory
    static int ctas = 0;
    if (ctas == 0) {
        int dev = 0, sms = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaFuncSetAttribute(k_agent_views, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sizeof(ViewMeshTable));
        if (e != cudaSuccess) return e;
        ctas = sms * kViewCtasPerSm;
    }
    const unsigned grid = views < unsigned(ctas) ? views : unsigned(ctas);
    k_agent_views<<<grid, kViewThreads, sizeof(ViewMeshTable), stream>>>(S, P, int(views));
    return cudaGetLastError();
}

}  // namespace hs
