// pth_scene.h — the scene handle shared by the pth_* C entry points (pth_scene_capi.cpp, pth_capi.cpp).
#pragma once

#include <cstdint>
#include <string>

#include "PathTracer_Types.h"

#define PTH_API extern "C" __attribute__((visibility("default")))

struct pth_scene
{
    PathTracerNS::GlobalVars gv;
    std::string err;
};

struct pth_scene_info
{
    uint32_t bvhSize, bvhMaxDepth, triangulationSize, lightsSize, materiauxSize, texturesSize, texturesDataSize, imageWidth, imageHeight;
    uint32_t reserved;
    const void *bvh, *triangulation, *lights, *materiaux, *textures, *texturesData, *sky;
    float camera[16];  // position, direction, right, up
};
