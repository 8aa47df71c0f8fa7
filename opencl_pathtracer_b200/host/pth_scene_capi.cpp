// pth_scene_capi.cpp — plain-C entry points over the scene side of the host mirror (procedural / file importer, BVH builder)
// for the Python binding, the tests and bench.py. This translation unit and libpathtracer_scene.so it is linked into have NO
// dependency on the CUDA library: bench.py's reference arm generates its scene here without mapping any device code.
#include <cstring>
#include <stdexcept>
#include <string>

#include "PathTracer_BVH.h"
#include "PathTracer_FileImporter.h"
#include "PathTracer_ProceduralImporter.h"
#include "pathtracer_cuda.h"  // ptc_scene_view (a plain struct; no symbol of the library is referenced)
#include "pth_scene.h"

using namespace PathTracerNS;

static thread_local std::string g_err;
PTH_API const char *pth_last_error() { return g_err.c_str(); }

// Runs the procedural importer (no BVH yet). Returns NULL on failure.
PTH_API pth_scene *pth_scene_create(const char *name, uint32_t seed, uint32_t width, uint32_t height, int loadSky)
{
    pth_scene *s = new pth_scene();
    try
    {
        PathTracerProceduralImporter imp(name, seed);
        imp.Initialize(s->gv);
        imp.Import(width, height, loadSky != 0);
    }
    catch (std::exception const &e)
    {
        g_err = e.what();
        delete s;
        return nullptr;
    }
    return s;
}

// PathTracerFileImporter::Import: a scene from sizes.pth / pointers.pth / textureData.pth in `folder` (no BVH yet).
PTH_API pth_scene *pth_scene_import(const char *folder, uint32_t width, uint32_t height, int loadSky)
{
    pth_scene *s = new pth_scene();
    try
    {
        PathTracerFileImporter imp(folder);
        imp.Initialize(s->gv);
        imp.Import(width, height, loadSky != 0);
    }
    catch (std::exception const &e)
    {
        g_err = e.what();
        GlobalVars &g = s->gv;
        delete[] g.triangulation; delete[] g.lights; delete[] g.materiaux; delete[] g.textures; delete[] g.texturesData;
        delete s;
        return nullptr;
    }
    return s;
}

// PathTracerFileImporter::Export of the scene as it is now (after pth_scene_build_bvh: triangles in BVH order).
PTH_API int pth_scene_export(pth_scene *s, const char *folder)
{
    try
    {
        PathTracerFileImporter exp(folder);
        exp.Initialize(s->gv);
        exp.Export();
    }
    catch (std::exception const &e)
    {
        g_err = e.what();
        return 1;
    }
    return 0;
}

// BVH_Create of this repository (host/PathTracer_BVH.cpp). Permutes the triangles.
PTH_API int pth_scene_build_bvh(pth_scene *s)
{
    try
    {
        delete[] s->gv.bvh;
        s->gv.bvh = nullptr;
        BVH_Create(s->gv);
    }
    catch (std::exception const &e)
    {
        g_err = e.what();
        return 1;
    }
    return 0;
}

// Installs an externally built BVH (e.g. the reference builder's output from oracle/_ref) — nodes are copied.
PTH_API int pth_scene_set_bvh(pth_scene *s, const void *nodes, uint32_t bvhSize, uint32_t bvhMaxDepth)
{
    delete[] s->gv.bvh;
    s->gv.bvh = new Node[bvhSize ? bvhSize : 1];
    std::memcpy((void *)s->gv.bvh, nodes, sizeof(Node) * (size_t)bvhSize);
    s->gv.bvhSize = bvhSize;
    s->gv.bvhMaxDepth = bvhMaxDepth;
    return 0;
}

PTH_API void pth_scene_get_info(pth_scene *s, pth_scene_info *o)
{
    const GlobalVars &g = s->gv;
    o->bvhSize = g.bvhSize; o->bvhMaxDepth = g.bvhMaxDepth; o->triangulationSize = g.triangulationSize; o->lightsSize = g.lightsSize;
    o->materiauxSize = g.materiauxSize; o->texturesSize = g.texturesSize; o->texturesDataSize = g.texturesDataSize;
    o->imageWidth = g.imageWidth; o->imageHeight = g.imageHeight; o->reserved = 0;
    o->bvh = g.bvh; o->triangulation = g.triangulation; o->lights = g.lights; o->materiaux = g.materiaux; o->textures = g.textures;
    o->texturesData = g.texturesData; o->sky = &g.sky;
    std::memcpy(o->camera + 0, &g.cameraPosition, 16);
    std::memcpy(o->camera + 4, &g.cameraDirection, 16);
    std::memcpy(o->camera + 8, &g.cameraRight, 16);
    std::memcpy(o->camera + 12, &g.cameraUp, 16);
}

PTH_API void pth_scene_get_view(pth_scene *s, ptc_scene_view *v)
{
    const GlobalVars &g = s->gv;
    std::memset(v, 0, sizeof(*v));
    v->struct_size = sizeof(*v);
    v->bvh_size = g.bvhSize; v->triangulation_size = g.triangulationSize; v->lights_size = g.lightsSize; v->materiaux_size = g.materiauxSize;
    v->textures_size = g.texturesSize; v->textures_data_size = g.texturesDataSize;
    v->bvh = g.bvh; v->triangulation = g.triangulation; v->lights = g.lights; v->materiaux = g.materiaux; v->textures = g.textures;
    v->textures_data = g.texturesData; v->sky = &g.sky;
    v->camera_position = g.cameraPosition; v->camera_direction = g.cameraDirection; v->camera_right = g.cameraRight; v->camera_up = g.cameraUp;
}

PTH_API void pth_scene_free(pth_scene *s)
{
    if (!s) return;
    GlobalVars &g = s->gv;
    delete[] g.bvh; delete[] g.triangulation; delete[] g.lights; delete[] g.materiaux; delete[] g.textures; delete[] g.texturesData;
    delete s;
}
