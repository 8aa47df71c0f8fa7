// PathTracer.cpp — the controller behind PathTracer.h.
//
// What is fixed by the drop-in claim: the entry points of reference Controleur/PathTracer.h:26-45 and the
// ORDER in which PathTracer_Main drives the pipeline — import, BVH_Create, the two scene limits, optional export, the three
// device calls, statistics, clear (reference Controleur/PathTracer.cpp:25-110) — with CUDA_SetupContext / CUDA_InitializeMemory /
// CUDA_RunKernel in the place of the OpenCL_* calls. Everything else here is this repository's own: phases are timed by one
// stopwatch, host buffers come from one allocator, and the end-of-run report is a table computed from the work summary
// (ptc-style totals derived from the three histograms the device layer returns), not a transcript of the reference's printer.
#include "PathTracer.h"

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <iostream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "PathTracer_FileImporter.h"
#include "pathtracer_cuda.h"

namespace PathTracerNS
{
GlobalVars globalVars;  // the one scene / image block the front-ends and the device layer share (reference Controleur/PathTracer.cpp:13)

namespace
{
bool g_keepResults = false, g_quiet = false, g_live = false;
std::string g_sceneFolder = ".";

struct Stopwatch
{
    using clock = std::chrono::steady_clock;
    clock::time_point t0 = clock::now();
    double lapMs()
    {
        const clock::time_point t1 = clock::now();
        const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        t0 = t1;
        return ms;
    }
};

std::string text(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
std::string text(const char *fmt, ...)
{
    char buf[256];
    va_list ap;
    va_start(ap, fmt);
    std::vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return buf;
}

void say(const std::string &line)
{
    if (!g_quiet) std::cout << line << std::endl;
}

template <typename T>
T *zeroed(size_t n, const T &zero = T())
{
    T *p = new T[n];
    std::fill(p, p + n, zero);
    return p;
}

template <typename T>
void drop(T *&p)
{
    delete[] p;
    p = nullptr;
}

// The limits the kernel's fixed-size arrays impose (reference Controleur/PathTracer.cpp:54-65: traversal stack BVH_MAX_DEPTH,
// LIGHTS_SIZE < MAX_LIGHT_SIZE); the device layer refuses the same scenes with PTC_ERR_LIMIT.
void requireSceneLimits(const GlobalVars &g)
{
    if (g.bvhMaxDepth >= BVH_MAX_DEPTH)
        throw std::runtime_error(text("BVH is too deep : it has a size of %u and should be under %u", g.bvhMaxDepth, (unsigned)BVH_MAX_DEPTH));
    if (g.lightsSize >= MAX_LIGHT_SIZE)
        throw std::runtime_error(text("The scene has too many lights : it has %u and should be under %u", g.lightsSize, (unsigned)MAX_LIGHT_SIZE));
}

// Totals of a finished run, from what the device layer hands back: rayDepths[d] = paths that ended after d bounces,
// rayIntersectedBBx/Tri[k] = paths that tested k boxes / triangles. "Charged" rays follow the reference's accounting
// (Controleur/PathTracer.cpp:238-243): a path that ended at depth d is charged (1 + lights) rays for each of its d + 1 segments.
struct WorkSummary
{
    unsigned long long paths = 0, chargedRays = 0, chargedHits = 0, boxTests = 0, triangleTests = 0;
};

WorkSummary summarise(const GlobalVars &g)
{
    WorkSummary w;
    const unsigned long long perSegment = 1ull + g.lightsSize;
    for (uint d = 0; d < g.rayMaxDepth; d++)
    {
        const unsigned long long n = g.rayDepths[d];
        w.paths += n;
        w.chargedRays += perSegment * (d + 1) * n;
        w.chargedHits += perSegment * d * n;
    }
    for (uint k = 0; k < MAX_INTERSETCION_NUMBER; k++)
    {
        w.boxTests += (unsigned long long)k * g.rayIntersectedBBx[k];
        w.triangleTests += (unsigned long long)k * g.rayIntersectedTri[k];
    }
    return w;
}

double ratio(double a, double b) { return b != 0 ? a / b : 0.0; }
}  // namespace

void PathTracer_SetQuiet(bool quiet) { g_quiet = quiet; }
void PathTracer_KeepResults(bool keep) { g_keepResults = keep; }
GlobalVars &PathTracer_GetGlobalVars() { return globalVars; }
void PathTracer_SetSceneFolder(const char *folder) { g_sceneFolder = folder ? folder : "."; }
void PathTracer_PrintSection(const char *section) { say(std::string("\n**** ") + section + " ****"); }

void PathTracer_SetImporter(PathTracerImporter *importer)
{
    PathTracer_ReleaseResults();
    globalVars.importer = importer;
}

bool PathTracer_Main(uint image_width, uint image_height, uint numImagesToRender, bool saveRenderedImages, bool loadSky,
                     bool exportScene, Sampler sampler, uint rayMaxDepth, bool printLogInfos, bool superSampling)
{
    try
    {
        g_live = true;
        Stopwatch watch;

        PathTracer_PrintSection("LOADING SCENE");
        PathTracer_Initialize(image_width, image_height, saveRenderedImages, loadSky, rayMaxDepth, printLogInfos, superSampling);
        const double loadMs = watch.lapMs();

        PathTracer_PrintSection("BUILDING BVH");
        // with PTC_FLAG_GPU_BVH (FAST) the device builds its own wide BVH from the triangle list inside CUDA_InitializeMemory:
        // the single-threaded host builder does not run at all
        if (!(CUDA_GetOptions().flags & PTC_FLAG_GPU_BVH)) BVH_Create(globalVars);
        const double bvhMs = watch.lapMs();
        requireSceneLimits(globalVars);

        if (exportScene)  // the scene as imported, triangles in BVH order, for a later run from files
        {
            PathTracer_PrintSection("EXPORTING SCENE");
            PathTracer_Export();
        }

        PathTracer_PrintSection("SETTING CUDA CONTEXT");
        watch.lapMs();
        CUDA_SetupContext(globalVars, sampler);
        CUDA_InitializeMemory(globalVars);
        const double setupMs = watch.lapMs();

        PathTracer_PrintSection("PATH TRACING");
        double traceMs = 0, memoryMs = 0, displayMs = 0;
        CUDA_RunKernel(globalVars, &PathTracer_UpdateWindow, numImagesToRender, &traceMs, &memoryMs, &displayMs);

        PathTracer_PrintSection("STATISTICS");
        PathTracer_ComputeStatistics(numImagesToRender, loadMs, bvhMs, setupMs, traceMs, memoryMs, displayMs);

        if (!g_keepResults) PathTracer_Clear();
    }
    catch (std::exception const &e)
    {
        // the device layer reports through exceptions, as OpenCL_ErrorHandling does: log, release, tell the front-end
        std::cerr << "ERROR : " << e.what() << std::endl;
        try { PathTracer_Clear(); }
        catch (std::exception const &again) { std::cerr << "ERROR 2 : " << again.what() << std::endl; }
        return false;
    }
    return true;
}

void PathTracer_Initialize(uint image_width, uint image_height, bool saveRenderedImages, bool loadSky, uint rayMaxDepth,
                           bool printLogInfos, bool superSampling)
{
    if (!globalVars.importer) throw std::runtime_error("PathTracer_SetImporter was not called");
    globalVars.printLogInfos = printLogInfos;
    globalVars.superSampling = superSampling;
    globalVars.importer->Initialize(globalVars);
    globalVars.importer->Import(image_width, image_height, loadSky);
    globalVars.rayMaxDepth = rayMaxDepth;
    PathTracer_InitializeImage();
    PathTracer_InitializeWindow(saveRenderedImages);
    globalVars.window->TestPaintWindow();
}

// host accumulators and histograms, all zero (the device layer overwrites them at every read-back)
void PathTracer_InitializeImage()
{
    GlobalVars &g = globalVars;
    if (g.imageWidth == 0 || g.imageHeight == 0)
        throw std::runtime_error(text("Bad image dimension : width x height : %u x %u", g.imageWidth, g.imageHeight));
    const size_t pixels = (size_t)g.imageWidth * g.imageHeight;
    g.imageColor = zeroed<RGBAColor>(pixels, RGBAColor(0, 0, 0, 0));
    g.imageRayNb = zeroed<float>(pixels);
    g.rayDepths = zeroed<uint>(g.rayMaxDepth + 1);
    g.rayIntersectedBBx = zeroed<uint>(MAX_INTERSETCION_NUMBER);
    g.rayIntersectedTri = zeroed<uint>(MAX_INTERSETCION_NUMBER);
}

void PathTracer_InitializeWindow(bool saveRenderedImages)
{
    PathTracerDialog *w = new PathTracerDialog();
    w->SetWidth(globalVars.imageWidth);
    w->SetHeight(globalVars.imageHeight);
    w->Create();
    w->saveRenderedImages = saveRenderedImages;
    w->SetNumberOfTriangles(globalVars.triangulationSize);
    w->SetMaxDepthOfBVH(globalVars.bvhMaxDepth);
    w->SetSizeOfBVH(globalVars.bvhSize);
    globalVars.window = w;
}

void PathTracer_Clear()
{
    GlobalVars &g = globalVars;
    delete g.importer;
    delete g.window;
    drop(g.imageColor); drop(g.imageRayNb); drop(g.rayDepths); drop(g.rayIntersectedBBx); drop(g.rayIntersectedTri);
    drop(g.bvh); drop(g.triangulation); drop(g.lights); drop(g.materiaux); drop(g.textures); drop(g.texturesData);
    g = GlobalVars();
    g_live = false;
}

void PathTracer_ReleaseResults()
{
    if (g_live) PathTracer_Clear();
}

bool PathTracer_UpdateWindow() { return globalVars.window->PaintWindow(globalVars.imageColor, globalVars.imageRayNb); }

// End-of-run report (the reference prints a similar one, Controleur/PathTracer.cpp:230-290, in clock() ticks): one table of
// (label, value) rows per block, values from summarise() and the phase times in milliseconds.
void PathTracer_ComputeStatistics(uint numImageToRender, double loadingTime, double bvhBuildingTime, double openclSettingTime,
                                  double pathTracingTime, double memoryTime, double displayTime)
{
    const GlobalVars &g = globalVars;
    const WorkSummary w = summarise(g);
    typedef std::vector<std::pair<std::string, std::string>> Rows;
    auto block = [](const char *title, const Rows &rows) {
        say(title);
        size_t width = 0;
        for (const auto &r : rows) width = std::max(width, r.first.size());
        for (const auto &r : rows) say("\t" + r.first + std::string(width - r.first.size(), ' ') + " : " + r.second);
        say("");
    };
    block("Scene", {{"image", text("%u x %u", g.imageWidth, g.imageHeight)},
                    {"triangles", text("%u", g.triangulationSize)},
                    {"lights", text("%u", g.lightsSize)},
                    {"BVH nodes", text("%u", g.bvhSize)},
                    {"BVH depth", text("%u", g.bvhMaxDepth)}});
    Rows work = {{"paths", text("%llu", w.paths)},
                 {"rays charged (1 + lights per segment)", text("%llu", w.chargedRays)},
                 {"box tests per charged ray", text("%.3f", ratio((double)w.boxTests, (double)w.chargedRays))},
                 {"triangle tests per charged ray", text("%.3f", ratio((double)w.triangleTests, (double)w.chargedRays))},
                 {"hits per triangle test", text("%.2f %%", 100.0 * ratio((double)w.chargedHits, (double)w.triangleTests))}};
    for (uint d = 0; d < g.rayMaxDepth; d++)
        work.push_back({text("paths ending after %u bounces", d), text("%.3f %%", 100.0 * ratio((double)g.rayDepths[d], (double)w.paths))});
    block("Work", work);
    block("Time (ms)", {{"import", text("%.1f", loadingTime)},
                        {"BVH_Create", text("%.1f", bvhBuildingTime)},
                        {"device set-up + scene upload", text("%.1f", openclSettingTime)},
                        {"path tracing", text("%.1f", pathTracingTime)},
                        {"read-back", text("%.1f", memoryTime)},
                        {"display", text("%.1f", displayTime)},
                        {"per picture", text("%.3f", ratio(pathTracingTime, numImageToRender))},
                        {"per 1000 paths", text("%.5f", 1000.0 * ratio(pathTracingTime, (double)w.paths))},
                        {"per 1000 charged rays", text("%.5f", 1000.0 * ratio(pathTracingTime, (double)w.chargedRays))}});
}

void PathTracer_PrintBVHCharacteristics()
{
    const BVHCharacteristics c = BVH_GetCharacteristics(globalVars.bvh);
    say(text("BVH : nodes %u, leafs %u, depth %u..%u, leaf size %u..%u", c.nNodes, c.nLeafs, c.minDepth, c.maxDepth, c.minLeafSize, c.maxLeafSize));
}

// the .pth triple of the scene as it is now (PathTracer_FileImporter.h); the folder is settable, the reference's is a build-time path
void PathTracer_Export()
{
    PathTracerFileImporter exporter(g_sceneFolder);
    exporter.Initialize(globalVars);
    exporter.Export();
}
}  // namespace PathTracerNS
