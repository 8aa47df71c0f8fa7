// PathTracer_Importer.h — scene import interface with the reference's shape.
//
// Same abstract interface as reference Controleur/PathTracer_Importer.h:14-51: an importer is handed
// aliases into GlobalVars by Initialize() and fills them in Import(width, height, loadSky). Front-ends
// install one with PathTracer_SetImporter() before calling PathTracer_Main(). The reference ships a Maya,
// a LumenRT and a raw-dump file importer (SDK / MSVC-ABI bound, out of scope — SURVEY.md §2); this build
// ships a procedural one (PathTracer_ProceduralImporter.h) that follows the Maya importer's conventions.
#ifndef PATHTRACER_B200_IMPORTER_H
#define PATHTRACER_B200_IMPORTER_H

#include "PathTracer_Types.h"

namespace PathTracerNS
{
class PathTracerImporter
{
public:
    PathTracerImporter() {}
    virtual ~PathTracerImporter() {}

    // reference Controleur/PathTracer_Importer.cpp:32-55
    void Initialize(GlobalVars &g)
    {
        ptr__global__cameraDirection = &g.cameraDirection;
        ptr__global__cameraRight = &g.cameraRight;
        ptr__global__cameraUp = &g.cameraUp;
        ptr__global__cameraPosition = &g.cameraPosition;
        ptr__global__triangulation = &g.triangulation;
        ptr__global__lights = &g.lights;
        ptr__global__materiaux = &g.materiaux;
        ptr__global__textures = &g.textures;
        ptr__global__texturesData = &g.texturesData;
        ptr__global__triangulationSize = &g.triangulationSize;
        ptr__global__lightsSize = &g.lightsSize;
        ptr__global__materiauxSize = &g.materiauxSize;
        ptr__global__texturesSize = &g.texturesSize;
        ptr__global__texturesDataSize = &g.texturesDataSize;
        ptr__global__imageWidth = &g.imageWidth;
        ptr__global__imageHeight = &g.imageHeight;
        ptr__global__imageSize = &g.imageSize;
        ptr__global__sky = &g.sky;
    }

    virtual void Import(uint image_width, uint image_height, bool loadSky) = 0;

    Float4 *ptr__global__cameraDirection = nullptr;
    Float4 *ptr__global__cameraRight = nullptr;
    Float4 *ptr__global__cameraUp = nullptr;
    Float4 *ptr__global__cameraPosition = nullptr;

    Triangle **ptr__global__triangulation = nullptr;
    Light **ptr__global__lights = nullptr;
    Material **ptr__global__materiaux = nullptr;
    Texture **ptr__global__textures = nullptr;
    Uchar4 **ptr__global__texturesData = nullptr;
    uint *ptr__global__triangulationSize = nullptr;
    uint *ptr__global__lightsSize = nullptr;
    uint *ptr__global__materiauxSize = nullptr;
    uint *ptr__global__texturesSize = nullptr;
    uint *ptr__global__texturesDataSize = nullptr;

    uint *ptr__global__imageWidth = nullptr;
    uint *ptr__global__imageHeight = nullptr;
    uint *ptr__global__imageSize = nullptr;

    Sky *ptr__global__sky = nullptr;
};
}  // namespace PathTracerNS

#endif
