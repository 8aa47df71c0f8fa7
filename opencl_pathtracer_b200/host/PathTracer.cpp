// PathTracer.cpp — see PathTracer.h.
#include "PathTracer.h"
#include "PathTracer_FileImporter.h"
#include "pathtracer_cuda.h"

#include <chrono>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace PathTracerNS
{
GlobalVars globalVars;  // reference Controleur/PathTracer.cpp:13
static bool g_keepResults = false, g_quiet = false, g_cleared = true;

#define CONSOLE_LOG \
    if (!g_quiet) std::cout
#define ENDL std::endl

static inline double nowMs()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

void PathTracer_SetQuiet(bool quiet) { g_quiet = quiet; }
void PathTracer_KeepResults(bool keep) { g_keepResults = keep; }
GlobalVars &PathTracer_GetGlobalVars() { return globalVars; }

void PathTracer_SetImporter(PathTracerImporter *importer)
{
    if (!g_cleared) PathTracer_ReleaseResults();
    globalVars.importer = importer;
}

bool PathTracer_Main(uint image_width, uint image_height, uint numImagesToRender, bool saveRenderedImages, bool loadSky,
                     bool exportScene, Sampler sampler, uint rayMaxDepth, bool printLogInfos, bool superSampling)
{
    try
    {
        double start, loadingTime, bvhBuildingTime, openclSettingTime, pathTracingTime, memoryTime, displayTime;
        g_cleared = false;

        PathTracer_PrintSection("LOADING SCENE");
        start = nowMs();
        PathTracer_Initialize(image_width, image_height, saveRenderedImages, loadSky, rayMaxDepth, printLogInfos, superSampling);
        loadingTime = nowMs() - start;

        PathTracer_PrintSection("BUILDING BVH");
        start = nowMs();
        // CUDA_Options.flags & PTC_FLAG_GPU_BVH (FAST mode): the device builds its own wide BVH from the triangle list inside
        // CUDA_InitializeMemory, so the host builder of the reference (:50, single-threaded) does not run at all
        if (!(CUDA_GetOptions().flags & PTC_FLAG_GPU_BVH)) BVH_Create(globalVars);
        bvhBuildingTime = nowMs() - start;

        if (globalVars.bvhMaxDepth >= BVH_MAX_DEPTH)
        {
            std::ostringstream m;
            m << "BVH is too deep : it has a size of " << globalVars.bvhMaxDepth << " and should be under " << BVH_MAX_DEPTH;
            throw std::runtime_error(m.str());
        }
        if (globalVars.lightsSize >= MAX_LIGHT_SIZE)
        {
            std::ostringstream m;
            m << "The scene has too many lights : it has " << globalVars.lightsSize << " and should be under " << MAX_LIGHT_SIZE;
            throw std::runtime_error(m.str());
        }
        if (exportScene)
        {
            // reference :67-71: dump the imported scene (after the BVH build has ordered the triangles) for a later run
            PathTracer_PrintSection("EXPORTING SCENE");
            PathTracer_Export();
        }

        PathTracer_PrintSection("SETTING CUDA CONTEXT");
        start = nowMs();
        CUDA_SetupContext(globalVars, sampler);
        CUDA_InitializeMemory(globalVars);
        openclSettingTime = nowMs() - start;

        PathTracer_PrintSection("PATH TRACING");
        CUDA_RunKernel(globalVars, &PathTracer_UpdateWindow, numImagesToRender, &pathTracingTime, &memoryTime, &displayTime);

        PathTracer_PrintSection("STATISTICS");
        PathTracer_ComputeStatistics(numImagesToRender, loadingTime, bvhBuildingTime, openclSettingTime, pathTracingTime, memoryTime, displayTime);

        if (!g_keepResults) PathTracer_Clear();
    }
    catch (std::exception const &e)
    {
        std::cerr << "ERROR : " << e.what() << std::endl;
        try { PathTracer_Clear(); }
        catch (std::exception const &e2) { std::cerr << "ERROR 2 : " << e2.what() << std::endl; }
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

void PathTracer_InitializeImage()
{
    if (!(globalVars.imageWidth > 0 && globalVars.imageHeight > 0))
    {
        std::ostringstream m;
        m << "Bad image dimension : width x height : " << globalVars.imageWidth << " x " << globalVars.imageHeight;
        throw std::runtime_error(m.str());
    }
    const size_t n = (size_t)globalVars.imageWidth * globalVars.imageHeight;
    globalVars.imageColor = new RGBAColor[n];
    globalVars.imageRayNb = new float[n];
    globalVars.rayDepths = new uint[globalVars.rayMaxDepth + 1];
    globalVars.rayIntersectedBBx = new uint[MAX_INTERSETCION_NUMBER];
    globalVars.rayIntersectedTri = new uint[MAX_INTERSETCION_NUMBER];
    for (uint x = 0; x < globalVars.rayMaxDepth + 1; x++) globalVars.rayDepths[x] = 0;
    for (uint x = 0; x < MAX_INTERSETCION_NUMBER; x++) globalVars.rayIntersectedBBx[x] = globalVars.rayIntersectedTri[x] = 0;
    for (size_t x = 0; x < n; x++)
    {
        globalVars.imageColor[x] = RGBAColor(0, 0, 0, 0);
        globalVars.imageRayNb[x] = 0;
    }
}

void PathTracer_InitializeWindow(bool saveRenderedImages)
{
    globalVars.window = new PathTracerDialog();
    globalVars.window->SetWidth(globalVars.imageWidth);
    globalVars.window->SetHeight(globalVars.imageHeight);
    globalVars.window->Create();
    globalVars.window->saveRenderedImages = saveRenderedImages;
    globalVars.window->SetNumberOfTriangles(globalVars.triangulationSize);
    globalVars.window->SetMaxDepthOfBVH(globalVars.bvhMaxDepth);
    globalVars.window->SetSizeOfBVH(globalVars.bvhSize);
}

void PathTracer_Clear()
{
    delete globalVars.importer;
    delete globalVars.window;
    delete[] globalVars.imageColor;
    delete[] globalVars.imageRayNb;
    delete[] globalVars.rayDepths;
    delete[] globalVars.rayIntersectedBBx;
    delete[] globalVars.rayIntersectedTri;
    delete[] globalVars.bvh;
    delete[] globalVars.triangulation;
    delete[] globalVars.lights;
    delete[] globalVars.materiaux;
    delete[] globalVars.textures;
    delete[] globalVars.texturesData;
    globalVars = GlobalVars();
    g_cleared = true;
}

void PathTracer_ReleaseResults()
{
    if (!g_cleared) PathTracer_Clear();
}

bool PathTracer_UpdateWindow() { return globalVars.window->PaintWindow(globalVars.imageColor, globalVars.imageRayNb); }

// reference Controleur/PathTracer.cpp:230-290 (times are milliseconds here; the reference prints clock() ticks)
void PathTracer_ComputeStatistics(uint numImageToRender, double loadingTime, double bvhBuildingTime, double openclSettingTime,
                                  double pathTracingTime, double memoryTime, double displayTime)
{
    unsigned long numberOfShotRays = 0, numberOfProcessedRays = 0, numberOfTestedBBx = 0, numberOfTestedTri = 0, numberOfIntersectedTri = 0;
    for (uint i = 0; i < globalVars.rayMaxDepth; i++)
    {
        numberOfShotRays += globalVars.rayDepths[i];
        numberOfProcessedRays += (unsigned long)(1 + globalVars.lightsSize) * (i + 1) * globalVars.rayDepths[i];
        numberOfIntersectedTri += (unsigned long)(1 + globalVars.lightsSize) * i * globalVars.rayDepths[i];
    }
    for (uint i = 0; i < MAX_INTERSETCION_NUMBER; i++)
    {
        numberOfTestedBBx += (unsigned long)i * globalVars.rayIntersectedBBx[i];
        numberOfTestedTri += (unsigned long)i * globalVars.rayIntersectedTri[i];
    }
    CONSOLE_LOG << "Scene information" << ENDL;
    CONSOLE_LOG << "\timage size          : " << globalVars.imageWidth << " x " << globalVars.imageHeight << ENDL;
    CONSOLE_LOG << "\tnumber of triangles : " << globalVars.triangulationSize << ENDL;
    CONSOLE_LOG << "\tnumber of lights    : " << globalVars.lightsSize << ENDL;
    CONSOLE_LOG << "\tsize of the bvh     : " << globalVars.bvhSize << ENDL;
    CONSOLE_LOG << "\tdepth of bvh        : " << globalVars.bvhMaxDepth << ENDL << ENDL;
    CONSOLE_LOG << "Path tracing information" << ENDL;
    CONSOLE_LOG << "\tNumber of shot rays                    : " << numberOfShotRays << ENDL;
    CONSOLE_LOG << "\tNumber of processed rays               : " << numberOfProcessedRays << ENDL;
    CONSOLE_LOG << "\tAverage number of tested BBx per ray   : " << ((float)numberOfTestedBBx) / ((float)numberOfProcessedRays) << ENDL;
    CONSOLE_LOG << "\tAverage number of tested Tri per ray   : " << ((float)numberOfTestedTri) / ((float)numberOfProcessedRays) << ENDL;
    CONSOLE_LOG << "\tPercentage of successful intersection : " << ((float)numberOfIntersectedTri * 100.) / ((float)numberOfTestedTri) << " % " << ENDL;
    CONSOLE_LOG << "\tDepth histogram in %" << ENDL;
    for (uint i = 0; i < globalVars.rayMaxDepth; i++)
        CONSOLE_LOG << "\t\t" << i << " : " << globalVars.rayDepths[i] * 100.0 / numberOfShotRays << " %" << ENDL;
    CONSOLE_LOG << ENDL << "Time information" << ENDL;
    CONSOLE_LOG << "\tLoading time        : " << loadingTime << " ms" << ENDL;
    CONSOLE_LOG << "\tBVH building time   : " << bvhBuildingTime << " ms" << ENDL;
    CONSOLE_LOG << "\tCUDA setting time   : " << openclSettingTime << " ms" << ENDL;
    CONSOLE_LOG << "\tPath Tracing time   : " << pathTracingTime << " ms" << ENDL;
    CONSOLE_LOG << "\tMemory time         : " << memoryTime << " ms" << ENDL;
    CONSOLE_LOG << "\tDisplay time        : " << displayTime << " ms" << ENDL << ENDL;
    CONSOLE_LOG << "Time statistics" << ENDL;
    CONSOLE_LOG << "\tAverage time per picture                  : " << pathTracingTime / numImageToRender << " ms" << ENDL;
    CONSOLE_LOG << "\tAverage time per 1000 full rays           : " << 1000. * pathTracingTime / numberOfShotRays << " ms" << ENDL;
    CONSOLE_LOG << "\tAverage time per 1000 (ray + shadow rays) : " << 1000. * pathTracingTime / numberOfProcessedRays << " ms" << ENDL << ENDL;
}

void PathTracer_PrintBVHCharacteristics()
{
    BVHCharacteristics c = BVH_GetCharacteristics(globalVars.bvh);
    CONSOLE_LOG << "BVH : nodes " << c.nNodes << ", leafs " << c.nLeafs << ", depth " << c.minDepth << ".." << c.maxDepth << ", leaf size "
                << c.minLeafSize << ".." << c.maxLeafSize << ENDL;
}

// reference Controleur/PathTracer.cpp:202-207 (the folder is settable instead of a build-time Windows path)
static std::string g_sceneFolder = ".";
void PathTracer_SetSceneFolder(const char *folder) { g_sceneFolder = folder ? folder : "."; }
void PathTracer_Export()
{
    PathTracerFileImporter fileExporter(g_sceneFolder);
    fileExporter.Initialize(globalVars);
    fileExporter.Export();
}

void PathTracer_PrintSection(const char *section) { CONSOLE_LOG << ENDL << "**** " << section << " ****" << ENDL; }
}  // namespace PathTracerNS
