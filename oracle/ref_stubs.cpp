// TEST INFRASTRUCTURE ONLY — link-time stubs for reference symbols that the
// headless CPU oracle never calls: the Vulkan/Metal render manager
// (M/src/render/*, graphics API — unavailable on this box) and the image
// importer used only for render textures.  src/mgr.cpp references them but
// skips every call when enableBatchRenderer == false && extRenderDev == nullptr
// (src/mgr.cpp:36-84, 896-902).  Any stub that is reached aborts loudly.
#include <cstdio>
#include <cstdlib>

#include <madrona/importer.hpp>
#include <madrona/render/api.hpp>
#include <madrona/render/render_mgr.hpp>

#define HS_STUB_ABORT() do { \
    fprintf(stderr, "oracle stub reached: %s\n", __PRETTY_FUNCTION__); \
    abort(); } while (0)

namespace madrona::render {

struct RenderContext {};

RenderManager::RenderManager(APIBackend *, GPUDevice *, const Config &) { HS_STUB_ABORT(); }
RenderManager::RenderManager(RenderManager &&) = default;
RenderManager::~RenderManager() = default;
CountT RenderManager::loadObjects(Span<const imp::SourceObject>,
                                  Span<const imp::SourceMaterial>,
                                  Span<const imp::SourceTexture>, bool) { HS_STUB_ABORT(); }
void RenderManager::configureLighting(Span<const LightConfig>) { HS_STUB_ABORT(); }
const RenderECSBridge * RenderManager::bridge() const { HS_STUB_ABORT(); }
void RenderManager::readECS() { HS_STUB_ABORT(); }
void RenderManager::batchRender() { HS_STUB_ABORT(); }
const uint8_t * RenderManager::batchRendererRGBOut() const { HS_STUB_ABORT(); }
const float * RenderManager::batchRendererDepthOut() const { HS_STUB_ABORT(); }

GPUHandle::GPUHandle(APIBackendSelect backend, GPUDevice *dev) : backend_(backend), dev_(dev) {}
GPUHandle::~GPUHandle() {}
APILibHandle::APILibHandle(APIBackendSelect backend, APILib *lib) : backend_(backend), lib_(lib) {}
APILibHandle::~APILibHandle() {}
APILibHandle APIManager::loadDefaultLib(APIBackendSelect) { HS_STUB_ABORT(); }
APIManager::APIManager(APILib *, const Config &, APIBackendSelect) { HS_STUB_ABORT(); }
APIManager::APIManager(APIManager &&) = default;
APIManager::~APIManager() {}
GPUHandle APIManager::initGPU(CountT) { HS_STUB_ABORT(); }

}

namespace madrona::imp {

struct ImageImporter::Impl {};

ImageImporter::ImageImporter() : impl_(nullptr) {}
ImageImporter::ImageImporter(ImageImporter &&) = default;
ImageImporter::~ImageImporter() = default;
Span<SourceTexture> ImageImporter::importImages(StackAlloc &, Span<const char * const>) { HS_STUB_ABORT(); }
void ImageImporter::deallocImportedImages(Span<SourceTexture>) { HS_STUB_ABORT(); }

}

// ---------------------------------------------------------------------------
// Deterministic sinf/cosf/sincosf.  The shared library is
// linked with -Bsymbolic-functions so every sinf/cosf/sincosf call inside the
// reference objects binds to these definitions instead of libm's; no reference
// source is edited.  The deviation from glibc is <= 1 ulp and is the ONLY
// arithmetic in the oracle that is not the reference's own.
#include "../include/hideseek/detmath.h"

extern "C" {
float sinf(float x) noexcept { return hs_det_sinf(x); }
float cosf(float x) noexcept { return hs_det_cosf(x); }
void sincosf(float x, float *s, float *c) noexcept { hs_det_sincosf(x, s, c); }
}
