// PathTracer.h — the controller, with the interface of reference Controleur/PathTracer.h:26-45.
//
// PathTracer_SetImporter + PathTracer_Main are what the reference's three front-ends call
// (Alone/main.cpp:26-27, Maya/RayTracer.cpp:120-121, LumenRT). The orchestration restates
// Controleur/PathTracer.cpp:25-110; the only substantive change is that the three OpenCL_* calls are the
// CUDA_* shims of PathTracer_CUDA.h.
#ifndef PATHTRACER_B200_PATHTRACER_H
#define PATHTRACER_B200_PATHTRACER_H

#include "PathTracer_BVH.h"
#include "PathTracer_CUDA.h"
#include "PathTracer_Dialog.h"
#include "PathTracer_Importer.h"

namespace PathTracerNS
{
void PathTracer_SetImporter(PathTracerImporter *importer);
bool PathTracer_Main(uint image_width, uint image_height, uint numImagesToRender, bool saveRenderedImages, bool loadSky,
                     bool exportScene, Sampler sampler, uint rayDepth, bool printLogInfos, bool superSampling);
void PathTracer_Initialize(uint image_width, uint image_height, bool saveRenderedImages, bool loadSky, uint rayMaxDepth,
                           bool printLogInfos, bool superSampling);
void PathTracer_InitializeImage();
void PathTracer_InitializeWindow(bool saveRenderedImages);
void PathTracer_Clear();
bool PathTracer_UpdateWindow();
void PathTracer_ComputeStatistics(uint numImageToRender, double loadingTime, double bvhBuildingTime, double openclSettingTime,
                                  double pathTracingTime, double memoryTime, double displayTime);
void PathTracer_PrintBVHCharacteristics();
void PathTracer_PrintSection(const char *section);
void PathTracer_Export();                            // reference :202-207: sizes.pth / pointers.pth / textureData.pth (PathTracer_FileImporter.h)
void PathTracer_SetSceneFolder(const char *folder);  // where PathTracer_Export writes (default "."; the reference hard-codes a folder)

// New: keep the results of the last PathTracer_Main alive for callers that want the buffers (tests, the
// Python binding). When set, PathTracer_Clear() is deferred until PathTracer_ReleaseResults().
void PathTracer_KeepResults(bool keep);
GlobalVars &PathTracer_GetGlobalVars();
void PathTracer_ReleaseResults();
void PathTracer_SetQuiet(bool quiet);
}  // namespace PathTracerNS

#endif
