// Translation unit of the agent-view ray caster (hs_views.cuh).  Separate from hs_sim.cu because
// the simulation kernels are compiled -fmad=false to reproduce the reference CPU backend's fp32
// results, while the ray caster has no such constraint and runs with FMA contraction.
#include "hs_views.cuh"

namespace hs {

cudaError_t launchAgentViews(const DeviceState &S, const ViewParams &P, unsigned views, cudaStream_t stream)
{
    // persistent CTAs: kViewCtasPerSm per SM, each with the mesh table in dynamic shared memory
    // (the shared-memory opt-in is a per-device function attribute: cache it per device)
    static int ctasOfDevice[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (ctasOfDevice[dev] == 0) {
        int sms = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaFuncSetAttribute(k_agent_views, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sizeof(ViewMeshTable));
        if (e != cudaSuccess) return e;
        ctasOfDevice[dev] = sms * kViewCtasPerSm;
    }
    const int ctas = ctasOfDevice[dev];
    const unsigned grid = views < unsigned(ctas) ? views : unsigned(ctas);
    k_agent_views<<<grid, kViewThreads, sizeof(ViewMeshTable), stream>>>(S, P, int(views));
    return cudaGetLastError();
}

}  // namespace hs
