// PathTracer_CUDA.cpp — see PathTracer_CUDA.h. Restates the call sequence of reference
// Controleur/PathTracer_OpenCL.cpp:56-198 (setup, buffer creation, launch loop with per-iteration
// read-back and callback, statistics read-back, release) over the PathTracer_CUDA C ABI.
#include "PathTracer_CUDA.h"

#include <chrono>
#include <stdexcept>
#include <string>

#include "pathtracer_cuda.h"

namespace PathTracerNS
{
namespace
{
ptc_ctx *g_ctx = nullptr;  // the reference keeps its cl_context / cl_mem objects in file statics too (:19-42)
CUDA_Options g_options;

void check(int rc, const char *what)
{
    if (rc == PTC_OK) return;
    std::string msg = std::string(what) + " : " + ptc_last_error(g_ctx);
    if (g_ctx)
    {
        ptc_destroy(g_ctx);
        g_ctx = nullptr;
    }
    throw std::runtime_error(msg);
}
inline double nowMs()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}
}  // namespace

void CUDA_SetOptions(const CUDA_Options &options) { g_options = options; }
const CUDA_Options &CUDA_GetOptions() { return g_options; }

void CUDA_SetupContext(GlobalVars &globalVars, Sampler sampler)
{
    if (g_ctx)
    {
        ptc_destroy(g_ctx);
        g_ctx = nullptr;
    }
    ptc_config cfg{};
    cfg.struct_size = sizeof(cfg);
    cfg.device = g_options.device;
    cfg.image_width = globalVars.imageWidth;
    cfg.image_height = globalVars.imageHeight;
    cfg.ray_max_depth = globalVars.rayMaxDepth;
    cfg.sampler = (uint32_t)sampler;
    cfg.mode = (uint32_t)g_options.mode;
    cfg.iterations_per_batch = g_options.iterationsPerBatch;
    cfg.flags = g_options.flags;
    if (globalVars.superSampling) cfg.flags &= ~(uint32_t)PTC_FLAG_NO_VARIANCE;  // the criterion reads imageV
    globalVars.sampler = sampler;
    int nDev = g_options.nDevices;
    if (nDev < 0) nDev = ptc_device_count();
    if (nDev > 16) nDev = 16;
    if (nDev > 1)
    {
        int32_t devs[16];
        for (int i = 0; i < nDev; i++) devs[i] = g_options.devices[i];
        if (globalVars.superSampling) throw std::runtime_error("CUDA_SetupContext : superSampling reads the running variance image, which several devices cannot share");
        check(ptc_create_multi(&cfg, devs, (uint32_t)nDev, &g_ctx), "CUDA_SetupContext");
    }
    else
    {
        if (nDev == 1) cfg.device = g_options.devices[0];
        check(ptc_create(&cfg, &g_ctx), "CUDA_SetupContext");
    }
    // reference Controleur/PathTracer_OpenCL.cpp:311 `-D SUPER_SAMPLING`. Its table file Kernel/X2inv.cl is empty (the option
    // cannot compile there); the library's own 95 % chi-square quantiles stand in, 4096 entries
    if (globalVars.superSampling) check(ptc_set_supersampling(g_ctx, nullptr, 4096), "CUDA_SetupContext");
}

void CUDA_InitializeMemory(GlobalVars &globalVars)
{
    if (!g_ctx) throw std::runtime_error("CUDA_InitializeMemory : CUDA_SetupContext was not called");
    ptc_scene_view v{};
    v.struct_size = sizeof(v);
    v.bvh_size = globalVars.bvhSize;
    v.triangulation_size = globalVars.triangulationSize;
    v.lights_size = globalVars.lightsSize;
    v.materiaux_size = globalVars.materiauxSize;
    v.textures_size = globalVars.texturesSize;
    v.textures_data_size = globalVars.texturesDataSize;
    v.bvh = globalVars.bvh;
    v.triangulation = globalVars.triangulation;
    v.lights = globalVars.lights;
    v.materiaux = globalVars.materiaux;
    v.textures = globalVars.textures;
    v.textures_data = globalVars.texturesData;
    v.sky = &globalVars.sky;
    v.camera_position = globalVars.cameraPosition;
    v.camera_direction = globalVars.cameraDirection;
    v.camera_right = globalVars.cameraRight;
    v.camera_up = globalVars.cameraUp;
    check(ptc_upload_scene(g_ctx, &v), "CUDA_InitializeMemory");
}

void CUDA_RunKernel(GlobalVars &globalVars, bool (*UpdateWindowFunc)(void), uint numImagesToRender, double *pathTracingTime,
                    double *memoryTime, double *displayTime)
{
    if (!g_ctx) throw std::runtime_error("CUDA_RunKernel : CUDA_SetupContext was not called");
    *pathTracingTime = 0;
    *memoryTime = 0;
    *displayTime = 0;
    const uint step = g_options.readBackEvery ? g_options.readBackEvery : 1;
    uint imageId = 0;
    while (imageId < numImagesToRender)
    {
        const uint n = (numImagesToRender - imageId) < step ? (numImagesToRender - imageId) : step;
        double t0 = nowMs();
        check(ptc_render(g_ctx, g_options.firstIteration + imageId, n), "CUDA_RunKernel");
        check(ptc_synchronize(g_ctx), "CUDA_RunKernel");
        *pathTracingTime += nowMs() - t0;

        t0 = nowMs();
        check(ptc_read_image(g_ctx, (float *)globalVars.imageColor, globalVars.imageRayNb), "CUDA_RunKernel");
        *memoryTime += nowMs() - t0;

        t0 = nowMs();
        if (UpdateWindowFunc) (*UpdateWindowFunc)();
        *displayTime += nowMs() - t0;
        imageId += n;
    }
    check(ptc_read_stats(g_ctx, globalVars.rayDepths, globalVars.rayIntersectedBBx, globalVars.rayIntersectedTri), "CUDA_RunKernel");
    // the reference releases every device object at the end of OpenCL_RunKernel (:120-139)
    ptc_destroy(g_ctx);
    g_ctx = nullptr;
}
}  // namespace PathTracerNS
