// pth_capi.cpp — plain-C entry points over the controller side of the host mirror (PathTracer_Main and the CUDA_* shims) for the
// Python binding, the tests and bench.py. The scene side (importers, BVH builder) is pth_scene_capi.cpp / libpathtracer_scene.so.
// Not part of the device boundary (that is include/pathtracer_cuda.h).
#include <chrono>
#include <cstring>
#include <stdexcept>
#include <string>

#include "PathTracer.h"
#include "PathTracer_FileImporter.h"
#include "PathTracer_ProceduralImporter.h"
#include "pathtracer_cuda.h"
#include "pth_scene.h"

using namespace PathTracerNS;

static thread_local std::string g_err;
PTH_API const char *pth_host_last_error() { return g_err.c_str(); }

// The whole reference pipeline through PathTracer_Main (importer -> BVH_Create -> CUDA_* shims -> statistics),
// i.e. what Alone/main.cpp does, with the results copied out. Any output pointer may be NULL.
PTH_API int pth_main(const char *scene, uint32_t seed, uint32_t width, uint32_t height, uint32_t numImagesToRender, int saveRenderedImages,
                     int loadSky, int sampler, uint32_t rayMaxDepth, int mode, uint32_t readBackEvery, int device, int nDevices, const char *exportFolder,
                     float *imageColor, float *imageRayNb, uint32_t *rayDepths, uint32_t *rayBbx, uint32_t *rayTri)
{
    CUDA_Options opt;
    opt.device = device;
    opt.nDevices = nDevices;
    opt.mode = mode;
    opt.readBackEvery = readBackEvery ? readBackEvery : 1;
    CUDA_SetOptions(opt);
    if (exportFolder) PathTracerDialog::exportFolderPath = exportFolder;
    PathTracer_SetQuiet(true);
    PathTracer_KeepResults(true);
    PathTracer_SetImporter(new PathTracerProceduralImporter(scene, seed));
    bool ok = PathTracer_Main(width, height, numImagesToRender, saveRenderedImages != 0, loadSky != 0, false, (Sampler)sampler, rayMaxDepth, false, false);
    if (ok)
    {
        GlobalVars &g = PathTracer_GetGlobalVars();
        const size_t n = (size_t)g.imageWidth * g.imageHeight;
        if (imageColor) std::memcpy(imageColor, g.imageColor, n * 16);
        if (imageRayNb) std::memcpy(imageRayNb, g.imageRayNb, n * 4);
        if (rayDepths) std::memcpy(rayDepths, g.rayDepths, (g.rayMaxDepth + 1) * 4);
        if (rayBbx) std::memcpy(rayBbx, g.rayIntersectedBBx, MAX_INTERSETCION_NUMBER * 4);
        if (rayTri) std::memcpy(rayTri, g.rayIntersectedTri, MAX_INTERSETCION_NUMBER * 4);
        PathTracer_ReleaseResults();
    }
    PathTracer_KeepResults(false);
    return ok ? 0 : 1;
}

// The drop-in DEFAULT, timed: what a maintainer gets from INTEGRATION.md's three-line change with CUDA_Options{} untouched —
// PARITY arithmetic, read-back of imageColor + imageRayNb and the callback after EVERY iteration, as the reference's loop does
// (Controleur/PathTracer_OpenCL.cpp:76-107) — on a scene that is already imported and has its BVH. times = {setup + upload,
// path tracing, read-back, display} in milliseconds as CUDA_RunKernel reports them. Returns 0 on success.
PTH_API int pth_run_default(pth_scene *s, uint32_t numImagesToRender, uint32_t rayMaxDepth, double times[4], uint32_t *rayDepths)
{
    GlobalVars &g = s->gv;
    const size_t n = (size_t)g.imageWidth * g.imageHeight;
    RGBAColor *color = new RGBAColor[n];
    float *nb = new float[n];
    uint *depths = new uint[rayMaxDepth + 1](), *bbx = new uint[MAX_INTERSETCION_NUMBER](), *tri = new uint[MAX_INTERSETCION_NUMBER]();
    g.imageColor = color; g.imageRayNb = nb; g.rayDepths = depths; g.rayIntersectedBBx = bbx; g.rayIntersectedTri = tri;
    g.rayMaxDepth = rayMaxDepth;
    int rc = 0;
    try
    {
        CUDA_SetOptions(CUDA_Options{});
        using clock = std::chrono::steady_clock;
        const clock::time_point t0 = clock::now();
        CUDA_SetupContext(g, JITTERED);
        CUDA_InitializeMemory(g);
        times[0] = std::chrono::duration<double, std::milli>(clock::now() - t0).count();
        CUDA_RunKernel(g, nullptr, numImagesToRender, &times[1], &times[2], &times[3]);
        if (rayDepths) std::memcpy(rayDepths, depths, sizeof(uint) * (rayMaxDepth + 1));
    }
    catch (std::exception const &e)
    {
        g_err = e.what();
        rc = 1;
    }
    delete[] color; delete[] nb; delete[] depths; delete[] bbx; delete[] tri;
    g.imageColor = nullptr; g.imageRayNb = nullptr; g.rayDepths = nullptr; g.rayIntersectedBBx = nullptr; g.rayIntersectedTri = nullptr;
    return rc;
}
